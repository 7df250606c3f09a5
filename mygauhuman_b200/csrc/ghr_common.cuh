// Shared device helpers for libghr (sm_100a).
//
// Bit-exactness note: radii / tiles_touched / keys depend on the exact fp32
// operation sequence of the reference kernels *as compiled* (which mul/add
// pairs nvcc+ptxas fused into FFMA).  That sequence was read off the
// reference's sm_100 SASS; here it is spelled with __fmaf_rn/__fmul_rn/
// __fadd_rn/__fsub_rn/__frcp_rn/__fdiv_rn/__fsqrt_rn, which the compiler never
// contracts or reorders, so the result does not depend on our own compiler's
// fusion choices.  oracle/raster_oracle.c spells the same sequence with fmaf.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "../../include/ghr.h"

#define GHR_BLOCK 256

namespace ghr {

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t lanemask_lt()
{
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// m0*x + m4*y + m8*z + m12 as the reference compiles it
// (auxiliary.h:58-77): t = y*m4; t = fma(x,m0,t); t = fma(z,m8,t); t = m12 + t
__device__ __forceinline__ float affine_row(float x, float y, float z, float m0, float m4, float m8, float m12)
{
    float t = __fmul_rn(y, m4);
    t = __fmaf_rn(x, m0, t);
    t = __fmaf_rn(z, m8, t);
    return __fadd_rn(m12, t);
}

// a0*b0 + a1*b1 + a2*b2 as the reference compiles a GLM mat3 product element
__device__ __forceinline__ float dot3_glm(float a0, float b0, float a1, float b1, float a2, float b2)
{
    float t = __fmul_rn(a1, b1);
    t = __fmaf_rn(a0, b0, t);
    t = __fmaf_rn(a2, b2, t);
    return t;
}

// auxiliary.h:41-44, evaluated in fp64 with one DFMA
__device__ __forceinline__ float ndc2pix(float v, int S)
{
    double d = __dadd_rn((double)v, 1.0);
    d = __fma_rn(d, (double)S, -1.0);
    d = __dmul_rn(d, 0.5);
    return (float)d;
}

// auxiliary.h:45-55
__device__ __forceinline__ void get_rect(float px, float py, int max_radius, uint32_t gx, uint32_t gy, uint32_t& minx,
                                         uint32_t& miny, uint32_t& maxx, uint32_t& maxy)
{
    const float rf = (float)max_radius;
    minx = min(gx, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fsub_rn(px, rf), 0.0625f))));
    miny = min(gy, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fsub_rn(py, rf), 0.0625f))));
    maxx = min(gx, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, rf), 16.0f), -1.0f), 0.0625f))));
    maxy = min(gy, (uint32_t)max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, rf), 16.0f), -1.0f), 0.0625f))));
}

// T = W*J of computeCov2D (forward.cu:74-113), shared by forward and backward.
struct Cov2DCtx {
    float T00, T01, T02, T10, T11, T12;  // T[c][r]; third column is zero
    float tx, ty, tz;                    // clamped view-space mean
    float txtz, tytz, limx, limy;
};

__device__ __forceinline__ void cov2d_T(float mx, float my, float mz, float fx, float fy, float tan_fovx,
                                        float tan_fovy, const float* __restrict__ V, Cov2DCtx& c)
{
    const float tx = affine_row(mx, my, mz, V[0], V[4], V[8], V[12]);
    const float ty = affine_row(mx, my, mz, V[1], V[5], V[9], V[13]);
    const float tz = affine_row(mx, my, mz, V[2], V[6], V[10], V[14]);
    const float limx = __fmul_rn(1.3f, tan_fovx), limy = __fmul_rn(1.3f, tan_fovy);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    const float clx = fminf(limx, fmaxf(-limx, txtz));
    const float cly = fminf(limy, fmaxf(-limy, tytz));
    const float tz2 = __fmul_rn(tz, tz);
    const float J00 = __fdiv_rn(fx, tz);
    const float J02 = __fdiv_rn(__fmul_rn(fx, __fmul_rn(clx, -tz)), tz2);
    const float J11 = __fdiv_rn(fy, tz);
    const float J12 = __fdiv_rn(__fmul_rn(fy, __fmul_rn(cly, -tz)), tz2);
    c.T00 = __fmaf_rn(V[2], J02, __fmul_rn(V[0], J00));
    c.T01 = __fmaf_rn(V[6], J02, __fmul_rn(V[4], J00));
    c.T02 = __fmaf_rn(V[10], J02, __fmul_rn(V[8], J00));
    c.T10 = __fmaf_rn(V[2], J12, __fmul_rn(J11, V[1]));
    c.T11 = __fmaf_rn(V[6], J12, __fmul_rn(J11, V[5]));
    c.T12 = __fmaf_rn(V[10], J12, __fmul_rn(J11, V[9]));
    c.tx = __fmul_rn(clx, tz);
    c.ty = __fmul_rn(cly, tz);
    c.tz = tz;
    c.txtz = txtz; c.tytz = tytz; c.limx = limx; c.limy = limy;
}

// cov = T^T * Vrk^T * T, + 0.3 low-pass (forward.cu:107-112)
__device__ __forceinline__ void cov2d(const Cov2DCtx& c, const float* v, float& c00, float& c01, float& c11)
{
    const float A00 = dot3_glm(c.T00, v[0], c.T01, v[1], c.T02, v[2]);
    const float A01 = dot3_glm(c.T10, v[0], c.T11, v[1], c.T12, v[2]);
    const float A10 = dot3_glm(c.T00, v[1], c.T01, v[3], c.T02, v[4]);
    const float A11 = dot3_glm(c.T10, v[1], c.T11, v[3], c.T12, v[4]);
    const float A20 = dot3_glm(c.T00, v[2], c.T01, v[4], c.T02, v[5]);
    const float A21 = dot3_glm(c.T10, v[2], c.T11, v[4], c.T12, v[5]);
    c00 = __fadd_rn(dot3_glm(c.T00, A00, c.T01, A10, c.T02, A20), 0.3f);
    c01 = dot3_glm(c.T00, A01, c.T01, A11, c.T02, A21);
    c11 = __fadd_rn(dot3_glm(c.T10, A01, c.T11, A11, c.T12, A21), 0.3f);
}

// computeCov3D (forward.cu:118-152), FMA placement from the reference SASS.
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float r, float x,
                                                     float y, float z, float* cov6)
{
    const float sx = __fmul_rn(s0, mod), sy = __fmul_rn(s1, mod), sz = __fmul_rn(s2, mod);
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
    const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
    const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
    const float xx_p_yy = __fmaf_rn(x, x, yy), xx_p_zz = __fmaf_rn(x, x, zz);
    const float yy_p_zz = __fadd_rn(yy, zz);
    const float R00 = __fsub_rn(1.0f, __fadd_rn(yy_p_zz, yy_p_zz)), R01 = __fadd_rn(xy_m_rz, xy_m_rz),
                R02 = __fadd_rn(xz_p_ry, xz_p_ry);
    const float R10 = __fadd_rn(xy_p_rz, xy_p_rz), R11 = __fsub_rn(1.0f, __fadd_rn(xx_p_zz, xx_p_zz)),
                R12 = __fadd_rn(yz_m_rx, yz_m_rx);
    const float R20 = __fadd_rn(xz_m_ry, xz_m_ry), R21 = __fadd_rn(yz_p_rx, yz_p_rx),
                R22 = __fsub_rn(1.0f, __fadd_rn(xx_p_yy, xx_p_yy));
    const float M00 = __fmul_rn(sx, R00), M01 = __fmul_rn(sy, R01), M02 = __fmul_rn(sz, R02);
    const float M10 = __fmul_rn(sx, R10), M11 = __fmul_rn(sy, R11), M12 = __fmul_rn(sz, R12);
    const float M20 = __fmul_rn(sx, R20), M21 = __fmul_rn(sy, R21), M22 = __fmul_rn(sz, R22);
    cov6[0] = dot3_glm(M00, M00, M01, M01, M02, M02);
    cov6[1] = dot3_glm(M10, M00, M11, M01, M12, M02);
    cov6[2] = dot3_glm(M20, M00, M21, M01, M22, M02);
    cov6[3] = dot3_glm(M10, M10, M11, M11, M12, M12);
    cov6[4] = dot3_glm(M20, M10, M21, M11, M22, M12);
    cov6[5] = dot3_glm(M20, M20, M21, M21, M22, M22);
}

// SH constants, auxiliary.h:22-39
#define GHR_SH_C0 0.28209479177387814f
#define GHR_SH_C1 0.4886025119029199f
__device__ __constant__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                                 -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                                 -0.5900435899266435f};

// forward.cu:20-71 computeColorFromSH (fp32; tolerance-checked, not bit-pinned)
__device__ __forceinline__ void sh_to_rgb(int deg, const float* __restrict__ sh, float dx, float dy, float dz,
                                          float* rgb, uint32_t& clamp_mask)
{
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    clamp_mask = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float res = GHR_SH_C0 * sh[c];
        if (deg > 0) {
            res = res - GHR_SH_C1 * y * sh[3 + c] + GHR_SH_C1 * z * sh[6 + c] - GHR_SH_C1 * x * sh[9 + c];
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + kSH_C2[0] * xy * sh[12 + c] + kSH_C2[1] * yz * sh[15 + c] +
                      kSH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] + kSH_C2[3] * xz * sh[21 + c] +
                      kSH_C2[4] * (xx - yy) * sh[24 + c];
                if (deg > 2) {
                    res = res + kSH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + kSH_C3[1] * xy * z * sh[30 + c] +
                          kSH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
                          kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
                          kSH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + kSH_C3[5] * z * (xx - yy) * sh[42 + c] +
                          kSH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
                }
            }
        }
        res += 0.5f;
        if (res < 0.0f) clamp_mask |= 1u << c;
        rgb[c] = fmaxf(res, 0.0f);
    }
}

// backward.cu:20-139
__device__ __forceinline__ void sh_backward(int deg, int M, const float* __restrict__ sh, float dox, float doy,
                                            float doz, uint32_t clamp_mask, const float* dL_dcolor, float* dm,
                                            float* __restrict__ dL_dsh, int accumulate = 0)
{
    const float len = sqrtf(dox * dox + doy * doy + doz * doz);
    const float x = dox / len, y = doy / len, z = doz / len;
    float dRGB[3];
#pragma unroll
    for (int c = 0; c < 3; c++) dRGB[c] = (clamp_mask >> c) & 1u ? 0.0f : dL_dcolor[c];
    float ddx = 0.f, ddy = 0.f, ddz = 0.f;
    float coef[16];
#pragma unroll
    for (int k = 0; k < 16; k++) coef[k] = 0.0f;
    coef[0] = GHR_SH_C0;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    if (deg > 0) {
        coef[1] = -GHR_SH_C1 * y; coef[2] = GHR_SH_C1 * z; coef[3] = -GHR_SH_C1 * x;
        if (deg > 1) {
            coef[4] = kSH_C2[0] * xy; coef[5] = kSH_C2[1] * yz; coef[6] = kSH_C2[2] * (2.f * zz - xx - yy);
            coef[7] = kSH_C2[3] * xz; coef[8] = kSH_C2[4] * (xx - yy);
            if (deg > 2) {
                coef[9] = kSH_C3[0] * y * (3.f * xx - yy); coef[10] = kSH_C3[1] * xy * z;
                coef[11] = kSH_C3[2] * y * (4.f * zz - xx - yy);
                coef[12] = kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                coef[13] = kSH_C3[4] * x * (4.f * zz - xx - yy); coef[14] = kSH_C3[5] * z * (xx - yy);
                coef[15] = kSH_C3[6] * x * (xx - 3.f * yy);
            }
        }
    }
    const int ncoef = (deg + 1) * (deg + 1);
    for (int k = 0; k < M; k++) {
        const float cf = k < ncoef && k < 16 ? coef[k] : 0.0f;
        if (accumulate == 2) {  // several streams add into the same buffer (multi-lane batches): L2 reductions
            atomicAdd(dL_dsh + 3 * k, cf * dRGB[0]); atomicAdd(dL_dsh + 3 * k + 1, cf * dRGB[1]);
            atomicAdd(dL_dsh + 3 * k + 2, cf * dRGB[2]);
        } else if (accumulate) {
            dL_dsh[3 * k] += cf * dRGB[0]; dL_dsh[3 * k + 1] += cf * dRGB[1]; dL_dsh[3 * k + 2] += cf * dRGB[2];
        } else {
            dL_dsh[3 * k] = cf * dRGB[0]; dL_dsh[3 * k + 1] = cf * dRGB[1]; dL_dsh[3 * k + 2] = cf * dRGB[2];
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float gx = 0.f, gy = 0.f, gz = 0.f;
#define SHV(i) sh[(i) * 3 + c]
        if (deg > 0) {
            gx = -GHR_SH_C1 * SHV(3); gy = -GHR_SH_C1 * SHV(1); gz = GHR_SH_C1 * SHV(2);
            if (deg > 1) {
                gx += kSH_C2[0] * y * SHV(4) + kSH_C2[2] * 2.f * -x * SHV(6) + kSH_C2[3] * z * SHV(7) + kSH_C2[4] * 2.f * x * SHV(8);
                gy += kSH_C2[0] * x * SHV(4) + kSH_C2[1] * z * SHV(5) + kSH_C2[2] * 2.f * -y * SHV(6) + kSH_C2[4] * 2.f * -y * SHV(8);
                gz += kSH_C2[1] * y * SHV(5) + kSH_C2[2] * 2.f * 2.f * z * SHV(6) + kSH_C2[3] * x * SHV(7);
                if (deg > 2) {
                    gx += (kSH_C3[0] * SHV(9) * 3.f * 2.f * xy + kSH_C3[1] * SHV(10) * yz + kSH_C3[2] * SHV(11) * -2.f * xy +
                           kSH_C3[3] * SHV(12) * -3.f * 2.f * xz + kSH_C3[4] * SHV(13) * (-3.f * xx + 4.f * zz - yy) +
                           kSH_C3[5] * SHV(14) * 2.f * xz + kSH_C3[6] * SHV(15) * 3.f * (xx - yy));
                    gy += (kSH_C3[0] * SHV(9) * 3.f * (xx - yy) + kSH_C3[1] * SHV(10) * xz +
                           kSH_C3[2] * SHV(11) * (-3.f * yy + 4.f * zz - xx) + kSH_C3[3] * SHV(12) * -3.f * 2.f * yz +
                           kSH_C3[4] * SHV(13) * -2.f * xy + kSH_C3[5] * SHV(14) * -2.f * yz + kSH_C3[6] * SHV(15) * -3.f * 2.f * xy);
                    gz += (kSH_C3[1] * SHV(10) * xy + kSH_C3[2] * SHV(11) * 4.f * 2.f * yz +
                           kSH_C3[3] * SHV(12) * 3.f * (2.f * zz - xx - yy) + kSH_C3[4] * SHV(13) * 4.f * 2.f * xz +
                           kSH_C3[5] * SHV(14) * (xx - yy));
                }
            }
        }
#undef SHV
        ddx += gx * dRGB[c]; ddy += gy * dRGB[c]; ddz += gz * dRGB[c];
    }
    // auxiliary.h:108-119 dnormvdv
    const float sum2 = dox * dox + doy * doy + doz * doz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dm[0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
    dm[1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
    dm[2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
}

// power / alpha of one (pixel, Gaussian) pair exactly as renderCUDA compiles
// (forward.cu:336-347): s = fma(dx, dx*a, dy*(dy*c)); power = fma(s,-0.5,-(dy*(dx*b)))
__device__ __forceinline__ float pair_power(float dx, float dy, float ca, float cb, float cc)
{
    const float s = __fmaf_rn(dx, __fmul_rn(dx, ca), __fmul_rn(dy, __fmul_rn(dy, cc)));
    return __fmaf_rn(s, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, cb)));
}


// ---- conservative screen-space extent of a splat (blend culling) ----------
// A (pixel, Gaussian) pair contributes only if alpha = min(0.99, o*exp(power))
// >= 1/255 (forward.cu:343-347), i.e. only inside the ellipse
// q(d) = a dx^2 + 2b dx dy + c dy^2 <= 2 ln(255 o).  Its axis-aligned bounding
// box has half-extents sqrt(2t * c/det), sqrt(2t * a/det) with det = ac - b^2.
// We store both, rounded UP to fp16 with a safety margin that swallows the fp32
// rounding of power/expf in the blend, packed in one 32-bit word.  A warp skips
// a Gaussian whose box misses its pixel block; the skip is conservative, so
// final_T / n_contrib stay bit-exact.  -1 = never contributes, +inf = unknown
// (NaN / indefinite conic): never skipped.
__device__ __forceinline__ uint32_t pack_cull_extent(float con_a, float con_b, float con_c, float opacity)
{
    const float t = 2.0f * (logf(255.0f * opacity) + 0.05f);
    float ex, ey;
    if (t > 0.0f) {
        const double det = (double)con_a * (double)con_c - (double)con_b * (double)con_b;
        if (det > 0.0) {
            ex = (float)sqrt((double)t * (double)con_c / det) * 1.01f + 0.02f;
            ey = (float)sqrt((double)t * (double)con_a / det) * 1.01f + 0.02f;
            if (!(ex >= 0.0f)) ex = __int_as_float(0x7f800000);
            if (!(ey >= 0.0f)) ey = __int_as_float(0x7f800000);
        } else {
            ex = ey = __int_as_float(0x7f800000);
        }
    } else if (t <= 0.0f) {
        ex = ey = -1.0f;  // o <= exp(-0.05)/255: alpha < 1/255 everywhere
    } else {
        ex = ey = __int_as_float(0x7f800000);  // NaN opacity: fminf(0.99, NaN) = 0.99 contributes
    }
    const uint32_t hx = __half_as_ushort(__float2half_ru(ex)), hy = __half_as_ushort(__float2half_ru(ey));
    return hx | (hy << 16);
}
__device__ __forceinline__ void unpack_cull_extent(uint32_t w, float& ex, float& ey)
{
    ex = __half2float(__ushort_as_half((unsigned short)(w & 0xFFFFu)));
    ey = __half2float(__ushort_as_half((unsigned short)(w >> 16)));
}

// Exact ellipse-vs-pixel-block refinement of the bounding-box test: the minimum over the block of
//   q(u,v) = a u^2 + 2 b u v + c v^2,   (u,v) = centre - pixel,
// (the reference's power is -q/2) is compared with the alpha >= 1/255 level 2 ln(255 o).  A convex quadratic
// whose unconstrained minimum (the centre) lies outside the block attains its block minimum on one of the four edges,
// where it is a clamped 1-D parabola.  Conservative by 0.05 in ln(alpha) plus 1e-5 of the term magnitudes (fp32
// rounding of the reference's own power), and anything indefinite / NaN is kept, so nothing the reference would blend
// into some pixel of the block is dropped: final_T / n_contrib / images stay bit-exact.
struct CullQuad {
    float a, b, c, tau, nb_c, nb_a;
    bool definite;  // false: indefinite / NaN conic, never culled by the exact test
};
__device__ __forceinline__ CullQuad cull_quad(float a, float b, float c, float opacity)
{
    CullQuad q;
    q.a = a; q.b = b; q.c = c;
    q.definite = (a > 0.0f) && (c > 0.0f) && (a * c > b * b);
    q.tau = 2.0f * (__logf(255.0f * opacity) + 0.05f);
    q.nb_c = __fdividef(-b, c);
    q.nb_a = __fdividef(-b, a);
    return q;
}
__device__ __forceinline__ bool cull_keep_exact(const CullQuad& g, float cx, float cy, float rx0, float rx1, float ry0,
                                                float ry1)
{
    const float u0 = cx - rx1, u1 = cx - rx0, v0 = cy - ry1, v1 = cy - ry0;  // u in [u0,u1], v in [v0,v1]
    if (u0 <= 0.0f && u1 >= 0.0f && v0 <= 0.0f && v1 >= 0.0f) return true;   // centre inside the block
    if (!g.definite) return true;
    const float a = g.a, b2 = 2.0f * g.b, c = g.c;
    float qmin;
    {
        const float v = fminf(fmaxf(g.nb_c * u0, v0), v1);
        qmin = fmaf(a * u0, u0, fmaf(b2 * u0, v, c * v * v));
    }
    {
        const float v = fminf(fmaxf(g.nb_c * u1, v0), v1);
        qmin = fminf(qmin, fmaf(a * u1, u1, fmaf(b2 * u1, v, c * v * v)));
    }
    {
        const float u = fminf(fmaxf(g.nb_a * v0, u0), u1);
        qmin = fminf(qmin, fmaf(c * v0, v0, fmaf(b2 * v0, u, a * u * u)));
    }
    {
        const float u = fminf(fmaxf(g.nb_a * v1, u0), u1);
        qmin = fminf(qmin, fmaf(c * v1, v1, fmaf(b2 * v1, u, a * u * u)));
    }
    const float mu = fmaxf(fabsf(u0), fabsf(u1)), mv = fmaxf(fabsf(v0), fabsf(v1));
    const float mag = fmaf(a * mu, mu, fmaf(fabsf(b2) * mu, mv, c * mv * mv));
    return !(qmin - 1e-5f * mag > g.tau);
}

// ---- decoupled look-back status words -------------------------------------
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

}  // namespace ghr
