/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * Plain-C CPU restatement of the reference tile rasteriser (the ashawkey
 * "depth + alpha-mask" variant of diff-gaussian-rasterization that
 * HandSomeWay/myGauHuman vendors).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may call into this file.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/).
 *
 * Floating-point contract: the integer outputs of the reference (radii,
 * tiles_touched, keys, tile ranges) depend on the exact fp32 operation
 * sequence that nvcc 12.9 + ptxas emit for the reference kernels (which
 * mul/add pairs were fused into FMA).  That sequence was read off the
 * reference's sm_100 SASS (DESIGN.md "bit-exactness") and is reproduced here
 * with explicit fmaf(); this file MUST be compiled with -ffp-contract=off so
 * that gcc adds no fusions of its own.  Division, reciprocal and sqrt are
 * IEEE round-to-nearest on both sides.  expf() is NOT bit-identical between
 * glibc and CUDA libdevice, so blended images agree to ~1e-6 and n_contrib can
 * differ on rare threshold pixels (tests bound that rate); the bit-exact pin
 * for those is the golden fixture produced by the reference CUDA build.
 *
 * Parity status: pinned by tests/golden/raster_*.npz (generated on a B200 by
 * the unmodified reference CUDA sources, see oracle/gen_golden_raster.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16
#define BLOCK_SIZE (TILE * TILE)

/* Spherical-harmonics constants: cuda_rasterizer/auxiliary.h:22-39 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

/* CUDA cvt.rzi.s32.f32 saturates and maps NaN to 0; a C cast is UB there. */
static inline int f2i_rz(float f)
{
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (-2147483647 - 1);
    return (int)f;
}

/* m0*x + m4*y + m8*z + m12 as nvcc emits it (auxiliary.h:58-77):
 * t = y*m4; t = fma(x,m0,t); t = fma(z,m8,t); t = m12 + t                */
static inline float affine_row(float x, float y, float z, float m0, float m4, float m8, float m12)
{
    float t = y * m4;
    t = fmaf(x, m0, t);
    t = fmaf(z, m8, t);
    return m12 + t;
}

/* a0*b0 + a1*b1 + a2*b2 as nvcc emits a GLM mat3 product element:
 * t = a1*b1; t = fma(a0,b0,t); t = fma(a2,b2,t)                           */
static inline float dot3_glm(float a0, float b0, float a1, float b1, float a2, float b2)
{
    float t = a1 * b1;
    t = fmaf(a0, b0, t);
    t = fmaf(a2, b2, t);
    return t;
}

/* auxiliary.h:41-44 -- evaluated in fp64: (v+1.0) then fma(.,S,-1.0) then *0.5 */
static inline float ndc2pix(float v, int S)
{
    double d = (double)v + 1.0;
    d = fma(d, (double)S, -1.0);
    d = d * 0.5;
    return (float)d;
}

/* auxiliary.h:45-55 */
static inline void get_rect(float px, float py, int max_radius, unsigned gx, unsigned gy,
                            unsigned* minx, unsigned* miny, unsigned* maxx, unsigned* maxy)
{
    float rf = (float)max_radius;
    int a;
    a = f2i_rz((px - rf) * 0.0625f); if (a < 0) a = 0; *minx = ((unsigned)a < gx) ? (unsigned)a : gx;
    a = f2i_rz((py - rf) * 0.0625f); if (a < 0) a = 0; *miny = ((unsigned)a < gy) ? (unsigned)a : gy;
    a = f2i_rz((((px + rf) + 16.0f) + -1.0f) * 0.0625f); if (a < 0) a = 0; *maxx = ((unsigned)a < gx) ? (unsigned)a : gx;
    a = f2i_rz((((py + rf) + 16.0f) + -1.0f) * 0.0625f); if (a < 0) a = 0; *maxy = ((unsigned)a < gy) ? (unsigned)a : gy;
}

/* forward.cu:118-152 computeCov3D (quaternion NOT normalised, forward.cu:127).
 * FMA placement read from the reference SASS.                              */
static void cov3d_from_scale_rot(const float* s3, float mod, const float* q4, float* cov6)
{
    float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
    float sx = s3[0] * mod, sy = s3[1] * mod, sz = s3[2] * mod;
    float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    float xz_p_ry = fmaf(r, y, xz);
    float xz_m_ry = fmaf(-r, y, xz);
    float yz_m_rx = fmaf(y, z, -rx);
    float yz_p_rx = fmaf(y, z, rx);
    float xy_m_rz = fmaf(x, y, -rz);
    float xy_p_rz = fmaf(x, y, rz);
    float xx_p_yy = fmaf(x, x, yy);
    float xx_p_zz = fmaf(x, x, zz);
    float yy_p_zz = yy + zz;
    /* R columns (GLM column-major constructor, forward.cu:134-138) */
    float R00 = 1.0f - (yy_p_zz + yy_p_zz), R01 = xy_m_rz + xy_m_rz, R02 = xz_p_ry + xz_p_ry;
    float R10 = xy_p_rz + xy_p_rz, R11 = 1.0f - (xx_p_zz + xx_p_zz), R12 = yz_m_rx + yz_m_rx;
    float R20 = xz_m_ry + xz_m_ry, R21 = yz_p_rx + yz_p_rx, R22 = 1.0f - (xx_p_yy + xx_p_yy);
    /* M = S * R  ->  M[c][k] = s_k * R[c][k] */
    float M00 = sx * R00, M01 = sy * R01, M02 = sz * R02;
    float M10 = sx * R10, M11 = sy * R11, M12 = sz * R12;
    float M20 = sx * R20, M21 = sy * R21, M22 = sz * R22;
    /* Sigma = M^T * M  ->  Sigma[j][i] = sum_k M[i][k] * M[j][k] */
    cov6[0] = dot3_glm(M00, M00, M01, M01, M02, M02);
    cov6[1] = dot3_glm(M10, M00, M11, M01, M12, M02);
    cov6[2] = dot3_glm(M20, M00, M21, M01, M22, M02);
    cov6[3] = dot3_glm(M10, M10, M11, M11, M12, M12);
    cov6[4] = dot3_glm(M20, M10, M21, M11, M22, M12);
    cov6[5] = dot3_glm(M20, M20, M21, M21, M22, M22);
}

/* forward.cu:20-71 computeColorFromSH (fp32, source order; tolerance-checked) */
static void sh_to_rgb(int deg, int M, const float* pos, const float* campos, const float* sh /*[M][3]*/,
                      float* rgb, unsigned char* clamped)
{
    float dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float x = dx / len, y = dy / len, z = dz / len;
    (void)M;
    for (int c = 0; c < 3; c++) {
        float res = SH_C0 * sh[0 * 3 + c];
        if (deg > 0) {
            res = res - SH_C1 * y * sh[1 * 3 + c] + SH_C1 * z * sh[2 * 3 + c] - SH_C1 * x * sh[3 * 3 + c];
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SH_C2[0] * xy * sh[4 * 3 + c] + SH_C2[1] * yz * sh[5 * 3 + c] +
                      SH_C2[2] * (2.0f * zz - xx - yy) * sh[6 * 3 + c] + SH_C2[3] * xz * sh[7 * 3 + c] +
                      SH_C2[4] * (xx - yy) * sh[8 * 3 + c];
                if (deg > 2) {
                    res = res + SH_C3[0] * y * (3.0f * xx - yy) * sh[9 * 3 + c] + SH_C3[1] * xy * z * sh[10 * 3 + c] +
                          SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11 * 3 + c] +
                          SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12 * 3 + c] +
                          SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13 * 3 + c] +
                          SH_C3[5] * z * (xx - yy) * sh[14 * 3 + c] + SH_C3[6] * x * (xx - 3.0f * yy) * sh[15 * 3 + c];
                }
            }
        }
        res += 0.5f;
        clamped[c] = (res < 0);
        rgb[c] = res < 0.0f ? 0.0f : res;
    }
}

/* The intermediate T = W*J of forward.cu:74-113 / backward.cu:144-199. */
typedef struct {
    float T00, T01, T02, T10, T11, T12; /* T[c][r], third column is zero */
    float tx, ty, tz;                   /* clamped view-space mean */
    float txtz, tytz, limx, limy;
} cov2d_ctx;

static void cov2d_T(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy, const float* V, cov2d_ctx* c)
{
    float tx = affine_row(mean[0], mean[1], mean[2], V[0], V[4], V[8], V[12]);
    float ty = affine_row(mean[0], mean[1], mean[2], V[1], V[5], V[9], V[13]);
    float tz = affine_row(mean[0], mean[1], mean[2], V[2], V[6], V[10], V[14]);
    float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    float txtz = tx / tz, tytz = ty / tz;
    float clx = fminf(limx, fmaxf(-limx, txtz));
    float cly = fminf(limy, fmaxf(-limy, tytz));
    float J00 = fx / tz;
    float J02 = (fx * (clx * -tz)) / (tz * tz);
    float J11 = fy / tz;
    float J12 = (fy * (cly * -tz)) / (tz * tz);
    c->T00 = fmaf(V[2], J02, V[0] * J00);
    c->T01 = fmaf(V[6], J02, V[4] * J00);
    c->T02 = fmaf(V[10], J02, V[8] * J00);
    c->T10 = fmaf(V[2], J12, J11 * V[1]);
    c->T11 = fmaf(V[6], J12, J11 * V[5]);
    c->T12 = fmaf(V[10], J12, J11 * V[9]);
    c->tx = clx * tz; c->ty = cly * tz; c->tz = tz;
    c->txtz = txtz; c->tytz = tytz; c->limx = limx; c->limy = limy;
}

/* forward.cu:74-113 computeCov2D; returns (cov00+0.3, cov01, cov11+0.3) */
static void cov2d(const cov2d_ctx* c, const float* v /*cov3D[6]*/, float* out3)
{
    float A00 = dot3_glm(c->T00, v[0], c->T01, v[1], c->T02, v[2]);
    float A01 = dot3_glm(c->T10, v[0], c->T11, v[1], c->T12, v[2]);
    float A10 = dot3_glm(c->T00, v[1], c->T01, v[3], c->T02, v[4]);
    float A11 = dot3_glm(c->T10, v[1], c->T11, v[3], c->T12, v[4]);
    float A20 = dot3_glm(c->T00, v[2], c->T01, v[4], c->T02, v[5]);
    float A21 = dot3_glm(c->T10, v[2], c->T11, v[4], c->T12, v[5]);
    float c00 = dot3_glm(c->T00, A00, c->T01, A10, c->T02, A20);
    float c01 = dot3_glm(c->T00, A01, c->T01, A11, c->T02, A21);
    float c11 = dot3_glm(c->T10, A01, c->T11, A11, c->T12, A21);
    out3[0] = c00 + 0.3f;
    out3[1] = c01;
    out3[2] = c11 + 0.3f;
}

/* rasterizer_impl.cu:35-50 getHigherMsb */
uint32_t ora_higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* rasterizer_impl.cu:54-66 checkFrustum / auxiliary.h:139-164 in_frustum */
void ora_mark_visible(int P, const float* means3D, const float* V, const float* Pm, unsigned char* present)
{
    (void)Pm;
    for (int i = 0; i < P; i++) {
        float z = affine_row(means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2], V[2], V[6], V[10], V[14]);
        present[i] = !(z <= 0.2f);
    }
}

/*
 * forward.cu:156-256 preprocessCUDA + rasterizer_impl.cu:279 InclusiveSum.
 * All per-Gaussian arrays are caller-allocated with P entries (rows).
 * Arrays that the reference leaves untouched for culled Gaussians are left
 * untouched here too (callers zero them first).  Returns num_rendered.
 */
int ora_preprocess(int P, int D, int M, int W, int H, const float* means3D, const float* shs,
                   const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                   const float* rotations, const float* cov3D_precomp, const float* V, const float* Pm,
                   const float* campos, float tan_fovx, float tan_fovy,
                   int* radii, float* depths, float* means2D, float* cov3Ds, float* conic_opacity, float* rgb,
                   unsigned char* clamped, uint32_t* tiles_touched, uint32_t* point_offsets)
{
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
    const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        const float x = means3D[3 * idx], y = means3D[3 * idx + 1], z = means3D[3 * idx + 2];
        float pvz = affine_row(x, y, z, V[2], V[6], V[10], V[14]);
        if (pvz <= 0.2f) continue;
        float hx = affine_row(x, y, z, Pm[0], Pm[4], Pm[8], Pm[12]);
        float hy = affine_row(x, y, z, Pm[1], Pm[5], Pm[9], Pm[13]);
        float hw = affine_row(x, y, z, Pm[3], Pm[7], Pm[11], Pm[15]);
        float p_w = 1.0f / (hw + 0.0000001f);
        float projx = hx * p_w, projy = hy * p_w;
        const float* c3;
        if (cov3D_precomp) {
            c3 = cov3D_precomp + 6 * idx;
        } else {
            cov3d_from_scale_rot(scales + 3 * idx, scale_modifier, rotations + 4 * idx, cov3Ds + 6 * idx);
            c3 = cov3Ds + 6 * idx;
        }
        cov2d_ctx ctx;
        float cov[3];
        cov2d_T(means3D + 3 * idx, focal_x, focal_y, tan_fovx, tan_fovy, V, &ctx);
        cov2d(&ctx, c3, cov);
        float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
        if (det == 0.0f) continue;
        float det_inv = 1.0f / det;
        float con0 = cov[2] * det_inv, con1 = det_inv * (-cov[1]), con2 = cov[0] * det_inv;
        float mid = (cov[0] + cov[2]) * 0.5f;
        float sq = sqrtf(fmaxf(fmaf(mid, mid, -det), 0.1f));
        float lambda1 = mid + sq, lambda2 = mid - sq;
        float my_radius = ceilf(sqrtf(fmaxf(lambda1, lambda2)) * 3.0f);
        float px = ndc2pix(projx, W), py = ndc2pix(projy, H);
        unsigned minx, miny, maxx, maxy;
        get_rect(px, py, f2i_rz(my_radius), gx, gy, &minx, &miny, &maxx, &maxy);
        if ((maxx - minx) * (maxy - miny) == 0) continue;
        if (!colors_precomp)
            sh_to_rgb(D, M, means3D + 3 * idx, campos, shs + (size_t)idx * M * 3, rgb + 3 * idx, clamped + 3 * idx);
        depths[idx] = pvz;
        radii[idx] = f2i_rz(my_radius);
        means2D[2 * idx] = px; means2D[2 * idx + 1] = py;
        conic_opacity[4 * idx] = con0; conic_opacity[4 * idx + 1] = con1; conic_opacity[4 * idx + 2] = con2;
        conic_opacity[4 * idx + 3] = opacities[idx];
        tiles_touched[idx] = (maxy - miny) * (maxx - minx);
    }
    uint32_t acc = 0;
    for (int i = 0; i < P; i++) { acc += tiles_touched[i]; point_offsets[i] = acc; }
    return (int)acc;
}

/* Stable LSD radix sort of (key,val) pairs on the low `bits` bits
 * (what cub::DeviceRadixSort::SortPairs(..., 0, bits) guarantees,
 * rasterizer_impl.cu:305-310).                                             */
static void stable_sort_pairs(uint64_t* k_in, uint32_t* v_in, uint64_t* k_out, uint32_t* v_out, int N, int bits)
{
    uint64_t *ka = k_in, *kb = k_out; uint32_t *va = v_in, *vb = v_out;
    size_t* cnt = (size_t*)malloc(sizeof(size_t) * 65537);
    int passes = 0;
    for (int shift = 0; shift < bits; shift += 16) {
        int nb = bits - shift < 16 ? bits - shift : 16;
        uint64_t mask = ((uint64_t)1 << nb) - 1;
        memset(cnt, 0, sizeof(size_t) * 65537);
        for (int i = 0; i < N; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int i = 0; i < 65536; i++) cnt[i + 1] += cnt[i];
        for (int i = 0; i < N; i++) { size_t d = cnt[(ka[i] >> shift) & mask]++; kb[d] = ka[i]; vb[d] = va[i]; }
        uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
        passes++;
    }
    if (ka != k_out) { memcpy(k_out, ka, sizeof(uint64_t) * N); memcpy(v_out, va, sizeof(uint32_t) * N); }
    free(cnt);
    (void)passes;
}

/*
 * rasterizer_impl.cu:70-111 duplicateWithKeys, :305 SortPairs,
 * :116-138 identifyTileRanges (+ memset :312).
 * keys_unsorted/vals_unsorted/keys/vals hold N entries, ranges 2*tiles.
 */
void ora_bin(int P, int W, int H, int N, const int* radii, const float* means2D, const float* depths,
             const uint32_t* point_offsets, uint64_t* keys_unsorted, uint32_t* vals_unsorted, uint64_t* keys,
             uint32_t* vals, uint32_t* ranges)
{
    const unsigned gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int idx = 0; idx < P; idx++) {
        if (radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : point_offsets[idx - 1];
        unsigned minx, miny, maxx, maxy;
        get_rect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gx, gy, &minx, &miny, &maxx, &maxy);
        uint32_t dbits; memcpy(&dbits, &depths[idx], 4);
        for (unsigned ty = miny; ty < maxy; ty++)
            for (unsigned tx = minx; tx < maxx; tx++) {
                uint64_t key = (uint64_t)(ty * gx + tx);
                key <<= 32; key |= dbits;
                keys_unsorted[off] = key; vals_unsorted[off] = (uint32_t)idx; off++;
            }
    }
    int bit = (int)ora_higher_msb(gx * gy);
    if (N > 0) {
        /* the sort clobbers its input buffer's role, so sort copies */
        uint64_t* kt = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)N);
        uint32_t* vt = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)N);
        memcpy(kt, keys_unsorted, sizeof(uint64_t) * (size_t)N);
        memcpy(vt, vals_unsorted, sizeof(uint32_t) * (size_t)N);
        stable_sort_pairs(kt, vt, keys, vals, N, 32 + bit);
        free(kt); free(vt);
    }
    memset(ranges, 0, sizeof(uint32_t) * 2 * gx * gy);
    for (int i = 0; i < N; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == N - 1) ranges[2 * cur + 1] = (uint32_t)N;
    }
}

/* forward.cu:261-383 renderCUDA (per pixel; tiles in parallel) */
void ora_blend_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                       const float* features, const float* depths, const float* conic_opacity, const float* bg,
                       float* out_color, float* out_depth, float* out_alpha, float* final_T, uint32_t* n_contrib)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx0 = (tile % gx) * TILE, ty0 = (tile / gx) * TILE;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int py = ty0; py < ty0 + TILE && py < H; py++)
            for (int px = tx0; px < tx0 + TILE && px < W; px++) {
                const float pxf = (float)px, pyf = (float)py;
                float T = 1.0f, C0 = 0, C1 = 0, C2 = 0, wsum = 0, Dacc = 0;
                uint32_t contributor = 0, last = 0;
                for (uint32_t k = r0; k < r1; k++) {
                    const uint32_t id = point_list[k];
                    contributor++;
                    float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                    const float* co = conic_opacity + 4 * id;
                    float s = fmaf(dx, dx * co[0], dy * (dy * co[2]));
                    float power = fmaf(s, -0.5f, -(dy * (dx * co[1])));
                    if (power > 0.0f) continue;
                    float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    float test_T = T * (1.0f - alpha);
                    if (test_T < 0.0001f) break; /* done = true */
                    C0 = fmaf(T, alpha * features[3 * id], C0);
                    C1 = fmaf(T, alpha * features[3 * id + 1], C1);
                    C2 = fmaf(T, alpha * features[3 * id + 2], C2);
                    wsum = fmaf(T, alpha, wsum);
                    Dacc = fmaf(T, alpha * depths[id], Dacc);
                    T = test_T;
                    last = contributor;
                }
                const size_t pix = (size_t)py * W + px;
                final_T[pix] = T; n_contrib[pix] = last;
                out_color[0 * (size_t)H * W + pix] = fmaf(T, bg[0], C0);
                out_color[1 * (size_t)H * W + pix] = fmaf(T, bg[1], C1);
                out_color[2 * (size_t)H * W + pix] = fmaf(T, bg[2], C2);
                out_alpha[pix] = wsum;
                out_depth[pix] = Dacc;
            }
    }
}

/*
 * backward.cu:399-587 renderCUDA (bwd).  Per-pixel math in fp32 exactly as
 * the reference; the cross-pixel sums that the reference does with fp32
 * global atomics (order-nondeterministic) are accumulated here in fp64.
 * acc9 is [P][9] doubles: dcolor(3), dmean2D(2), dconic(x,y,w), dopacity.
 */
void ora_blend_backward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                        const float* means2D, const float* conic_opacity, const float* colors, const float* depths,
                        const float* final_Ts, const uint32_t* n_contrib, const float* dL_dpixels,
                        const float* dL_ddepths, const float* dL_dalphas, double* acc9)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx0 = (tile % gx) * TILE, ty0 = (tile / gx) * TILE;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        const int todo = (int)(r1 - r0);
        for (int py = ty0; py < ty0 + TILE && py < H; py++)
            for (int px = tx0; px < tx0 + TILE && px < W; px++) {
                const size_t pix = (size_t)py * W + px;
                const float pxf = (float)px, pyf = (float)py;
                const float T_final = final_Ts[pix];
                float T = T_final;
                uint32_t contributor = (uint32_t)todo;
                const int last_contributor = (int)n_contrib[pix];
                float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0};
                float dpix[3] = {dL_dpixels[pix], dL_dpixels[(size_t)H * W + pix], dL_dpixels[2 * (size_t)H * W + pix]};
                float dL_ddepth = dL_ddepths[pix], dL_dalpha = dL_dalphas[pix];
                float accum_depth_rec = 0, accum_alpha_rec = 0, last_alpha = 0, last_depth = 0;
                for (int k = 0; k < todo; k++) {
                    contributor--;
                    if ((int)contributor >= last_contributor) continue;
                    const uint32_t id = point_list[r1 - 1 - k];
                    const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                    const float* co = conic_opacity + 4 * id;
                    float s = fmaf(dx, dx * co[0], dy * (dy * co[2]));
                    float power = fmaf(s, -0.5f, -(dy * (dx * co[1])));
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.0f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dopa = 0.0f;
                    double* a = acc9 + 9 * (size_t)id;
                    for (int ch = 0; ch < 3; ch++) {
                        const float c = colors[3 * id + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.0f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dopa += (c - accum_rec[ch]) * dpix[ch];
                        double g = (double)(dchannel_dcolor * dpix[ch]);
#pragma omp atomic
                        a[ch] += g;
                    }
                    const float c_d = depths[id];
                    accum_depth_rec = last_alpha * last_depth + (1.0f - last_alpha) * accum_depth_rec;
                    last_depth = c_d;
                    dL_dopa += (c_d - accum_depth_rec) * dL_ddepth;
                    accum_alpha_rec = last_alpha + (1.0f - last_alpha) * accum_alpha_rec;
                    dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
                    dL_dopa *= T;
                    last_alpha = alpha;
                    float bg_dot = 0;
                    for (int i = 0; i < 3; i++) bg_dot += bg[i] * dpix[i];
                    dL_dopa += (-T_final / (1.0f - alpha)) * bg_dot;
                    const float dL_dG = co[3] * dL_dopa;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
                    double g3 = (double)(dL_dG * dG_ddelx * ddelx_dx), g4 = (double)(dL_dG * dG_ddely * ddely_dy);
                    double g5 = (double)(-0.5f * gdx * dx * dL_dG), g6 = (double)(-0.5f * gdx * dy * dL_dG);
                    double g7 = (double)(-0.5f * gdy * dy * dL_dG), g8 = (double)(G * dL_dopa);
#pragma omp atomic
                    a[3] += g3;
#pragma omp atomic
                    a[4] += g4;
#pragma omp atomic
                    a[5] += g5;
#pragma omp atomic
                    a[6] += g6;
#pragma omp atomic
                    a[7] += g7;
#pragma omp atomic
                    a[8] += g8;
                }
            }
    }
}

/* backward.cu:20-139 computeColorFromSH (bwd): writes dL_dsh[M][3], adds to dL_dmean[3] */
static void sh_backward(int deg, int M, const float* pos, const float* campos, const float* sh,
                        const unsigned char* clamped, const float* dL_dcolor, float* dL_dmean, float* dL_dsh)
{
    float dox = pos[0] - campos[0], doy = pos[1] - campos[1], doz = pos[2] - campos[2];
    float len = sqrtf(dox * dox + doy * doy + doz * doz);
    float x = dox / len, y = doy / len, z = doz / len;
    float dRGB[3];
    for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0.0f : 1.0f);
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
    (void)M;
#define SH(i, c) sh[(i) * 3 + (c)]
#define DSH(i, v) for (int c = 0; c < 3; c++) dL_dsh[(i) * 3 + c] = (v) * dRGB[c]
    DSH(0, SH_C0);
    if (deg > 0) {
        DSH(1, -SH_C1 * y); DSH(2, SH_C1 * z); DSH(3, -SH_C1 * x);
        for (int c = 0; c < 3; c++) { dRGBdx[c] = -SH_C1 * SH(3, c); dRGBdy[c] = -SH_C1 * SH(1, c); dRGBdz[c] = SH_C1 * SH(2, c); }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            DSH(4, SH_C2[0] * xy); DSH(5, SH_C2[1] * yz); DSH(6, SH_C2[2] * (2.f * zz - xx - yy));
            DSH(7, SH_C2[3] * xz); DSH(8, SH_C2[4] * (xx - yy));
            for (int c = 0; c < 3; c++) {
                dRGBdx[c] += SH_C2[0] * y * SH(4, c) + SH_C2[2] * 2.f * -x * SH(6, c) + SH_C2[3] * z * SH(7, c) + SH_C2[4] * 2.f * x * SH(8, c);
                dRGBdy[c] += SH_C2[0] * x * SH(4, c) + SH_C2[1] * z * SH(5, c) + SH_C2[2] * 2.f * -y * SH(6, c) + SH_C2[4] * 2.f * -y * SH(8, c);
                dRGBdz[c] += SH_C2[1] * y * SH(5, c) + SH_C2[2] * 2.f * 2.f * z * SH(6, c) + SH_C2[3] * x * SH(7, c);
            }
            if (deg > 2) {
                DSH(9, SH_C3[0] * y * (3.f * xx - yy)); DSH(10, SH_C3[1] * xy * z);
                DSH(11, SH_C3[2] * y * (4.f * zz - xx - yy)); DSH(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                DSH(13, SH_C3[4] * x * (4.f * zz - xx - yy)); DSH(14, SH_C3[5] * z * (xx - yy));
                DSH(15, SH_C3[6] * x * (xx - 3.f * yy));
                for (int c = 0; c < 3; c++) {
                    dRGBdx[c] += (SH_C3[0] * SH(9, c) * 3.f * 2.f * xy + SH_C3[1] * SH(10, c) * yz + SH_C3[2] * SH(11, c) * -2.f * xy +
                                  SH_C3[3] * SH(12, c) * -3.f * 2.f * xz + SH_C3[4] * SH(13, c) * (-3.f * xx + 4.f * zz - yy) +
                                  SH_C3[5] * SH(14, c) * 2.f * xz + SH_C3[6] * SH(15, c) * 3.f * (xx - yy));
                    dRGBdy[c] += (SH_C3[0] * SH(9, c) * 3.f * (xx - yy) + SH_C3[1] * SH(10, c) * xz +
                                  SH_C3[2] * SH(11, c) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12, c) * -3.f * 2.f * yz +
                                  SH_C3[4] * SH(13, c) * -2.f * xy + SH_C3[5] * SH(14, c) * -2.f * yz + SH_C3[6] * SH(15, c) * -3.f * 2.f * xy);
                    dRGBdz[c] += (SH_C3[1] * SH(10, c) * xy + SH_C3[2] * SH(11, c) * 4.f * 2.f * yz +
                                  SH_C3[3] * SH(12, c) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13, c) * 4.f * 2.f * xz +
                                  SH_C3[5] * SH(14, c) * (xx - yy));
                }
            }
        }
    }
#undef SH
#undef DSH
    float ddx = 0, ddy = 0, ddz = 0;
    for (int c = 0; c < 3; c++) { ddx += dRGBdx[c] * dRGB[c]; ddy += dRGBdy[c] * dRGB[c]; ddz += dRGBdz[c] * dRGB[c]; }
    /* auxiliary.h:108-119 dnormvdv */
    float sum2 = dox * dox + doy * doy + doz * doz;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dL_dmean[0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
    dL_dmean[1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
    dL_dmean[2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
}

/* backward.cu:278-341 computeCov3D (bwd) */
static void cov3d_backward(const float* s3, float mod, const float* q4, const float* dL_dcov3D, float* dL_dscale,
                           float* dL_drot)
{
    float r = q4[0], x = q4[1], y = q4[2], z = q4[3];
    /* R[c][r] as GLM columns */
    float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                     {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                     {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float s[3] = {mod * s3[0], mod * s3[1], mod * s3[2]};
    float Mm[3][3];
    for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) Mm[c][k] = s[k] * R[c][k];
    const float* d = dL_dcov3D;
    float dS[3][3] = {{d[0], 0.5f * d[1], 0.5f * d[2]}, {0.5f * d[1], d[3], 0.5f * d[4]}, {0.5f * d[2], 0.5f * d[4], d[5]}};
    /* dL_dM = 2 * M * dL_dSigma (GLM column-major product) */
    float dM[3][3];
    for (int j = 0; j < 3; j++) for (int i = 0; i < 3; i++)
        dM[j][i] = 2.0f * (Mm[0][i] * dS[j][0] + Mm[1][i] * dS[j][1] + Mm[2][i] * dS[j][2]);
    /* Rt = transpose(R), dMt = transpose(dM) : Xt[c][k] = X[k][c] */
    float dMt[3][3], Rt[3][3];
    for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) { dMt[c][k] = dM[k][c]; Rt[c][k] = R[k][c]; }
    for (int c = 0; c < 3; c++) dL_dscale[c] = Rt[c][0] * dMt[c][0] + Rt[c][1] * dMt[c][1] + Rt[c][2] * dMt[c][2];
    for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) dMt[c][k] *= s[c];
    dL_drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    dL_drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
    dL_drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
    dL_drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

/*
 * backward.cu:144-274 computeCov2DCUDA + :346-396 preprocessCUDA (bwd).
 * dL_dmean2D [P][3], dL_dconic [P][4] (x,y,.,w), dL_dcolor [P][3] in;
 * dL_dmean3D [P][3], dL_dcov3D [P][6], dL_dsh [P][M][3], dL_dscale [P][3],
 * dL_drot [P][4] out (callers pre-zero them, as rasterize_points.cu:159-167).
 */
void ora_preprocess_backward(int P, int D, int M, int W, int H, const float* means3D, const int* radii,
                             const float* shs, const unsigned char* clamped, const float* scales,
                             const float* rotations, float scale_modifier, const float* cov3Ds, const float* V,
                             const float* Pm, float tan_fovx, float tan_fovy, const float* campos,
                             const float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D, float* dL_dcolor,
                             float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot)
{
    const float h_y = H / (2.0f * tan_fovy);
    const float h_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float* cov3D = cov3Ds + 6 * idx;
        const float* mean = means3D + 3 * idx;
        float dcx = dL_dconic[4 * idx], dcy = dL_dconic[4 * idx + 1], dcz = dL_dconic[4 * idx + 3];
        cov2d_ctx c; float cv[3];
        cov2d_T(mean, h_x, h_y, tan_fovx, tan_fovy, V, &c);
        cov2d(&c, cov3D, cv);
        const float x_grad_mul = (c.txtz < -c.limx || c.txtz > c.limx) ? 0.f : 1.f;
        const float y_grad_mul = (c.tytz < -c.limy || c.tytz > c.limy) ? 0.f : 1.f;
        float a = cv[0], b = cv[1], cc = cv[2];
        float denom = a * cc - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = dL_dcov3D + 6 * idx;
        /* T[c][r]: T0r = (T00,T01,T02), T1r = (T10,T11,T12) */
        const float T00 = c.T00, T01 = c.T01, T02 = c.T02, T10 = c.T10, T11 = c.T11, T12 = c.T12;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-cc * cc * dcx + 2 * b * cc * dcy + (denom - a * cc) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * cc) * dcx);
            dL_db = denom2inv * 2 * (b * cc * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc);
            dcov[3] = (T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc);
            dcov[5] = (T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc);
            dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
            dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
            dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
        } else {
            for (int i = 0; i < 6; i++) dcov[i] = 0;
        }
        const float V00 = cov3D[0], V01 = cov3D[1], V02 = cov3D[2], V11 = cov3D[3], V12 = cov3D[4], V22 = cov3D[5];
        float dL_dT00 = 2 * (T00 * V00 + T01 * V01 + T02 * V02) * dL_da + (T10 * V00 + T11 * V01 + T12 * V02) * dL_db;
        float dL_dT01 = 2 * (T00 * V01 + T01 * V11 + T02 * V12) * dL_da + (T10 * V01 + T11 * V11 + T12 * V12) * dL_db;
        float dL_dT02 = 2 * (T00 * V02 + T01 * V12 + T02 * V22) * dL_da + (T10 * V02 + T11 * V12 + T12 * V22) * dL_db;
        float dL_dT10 = 2 * (T10 * V00 + T11 * V01 + T12 * V02) * dL_dc + (T00 * V00 + T01 * V01 + T02 * V02) * dL_db;
        float dL_dT11 = 2 * (T10 * V01 + T11 * V11 + T12 * V12) * dL_dc + (T00 * V01 + T01 * V11 + T02 * V12) * dL_db;
        float dL_dT12 = 2 * (T10 * V02 + T11 * V12 + T12 * V22) * dL_dc + (T00 * V02 + T01 * V12 + T02 * V22) * dL_db;
        /* W[c][r] = V[4r + c]  (W columns are viewmatrix rows 0..2) */
        float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
        float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
        float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
        float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
        float tz = 1.f / c.tz, tz2 = tz * tz, tz3 = tz2 * tz;
        float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
        float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
        float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * c.tx) * tz3 * dL_dJ02 + (2 * h_y * c.ty) * tz3 * dL_dJ12;
        /* transformVec4x3Transpose, auxiliary.h:89-97; assignment, backward.cu:273 */
        float dm[3] = {V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz, V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz,
                       V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz};
        /* backward.cu:363-387 : 2D mean -> 3D mean */
        const float mx = mean[0], my = mean[1], mz = mean[2];
        float m_hw = Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15];
        float m_w = 1.0f / (m_hw + 0.0000001f);
        float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
        float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
        const float g2x = dL_dmean2D[3 * idx], g2y = dL_dmean2D[3 * idx + 1];
        dm[0] += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
        dm[1] += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
        dm[2] += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
        if (shs)
            sh_backward(D, M, mean, campos, shs + (size_t)idx * M * 3, clamped + 3 * idx, dL_dcolor + 3 * idx, dm,
                        dL_dsh + (size_t)idx * M * 3);
        dL_dmean3D[3 * idx] = dm[0]; dL_dmean3D[3 * idx + 1] = dm[1]; dL_dmean3D[3 * idx + 2] = dm[2];
        if (scales)
            cov3d_backward(scales + 3 * idx, scale_modifier, rotations + 4 * idx, dcov, dL_dscale + 3 * idx, dL_drot + 4 * idx);
    }
}
