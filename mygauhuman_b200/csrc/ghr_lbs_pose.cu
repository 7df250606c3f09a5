// Pose-level part of the LBS deformation (24 joints, a few hundred floats):
//   ghr_lbs_prepare         fold J_regressor into template / shape directions
//   ghr_lbs_pose            Rodrigues + rest joints + kinematic chain + pose feature
//   ghr_lbs_vertex_offsets  per-vertex blend-shape displacement (GEMV over posedirs)
//   ghr_lbs_pose_backward   gradient of the chain / pose feature w.r.t. correct_Rs
//   ghr_smpl_lbs_vertices   smplx lbs() over the model's own vertices
// Reference: scene/gaussian_model.py:894-980, :805-844; smplx/lbs.py:156-405.
// The reference runs ~60 ATen micro-kernels for this; here the whole
// pose stage is ONE single-warp kernel and the 17 MB posedirs GEMV is one
// HBM-streaming kernel (algorithmic bytes 4*(V*3*207 + V*3*NB)).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

constexpr int kJ = GHR_LBS_JOINTS;

// ---------------------------------------------------------------------------
// prepare: joint_template[j][c] = sum_v Jreg[j][v] vt[v][c];  joint_shapedirs[j][c][b] likewise
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lbs_prepare_kernel(int V, int NB, const float* __restrict__ vt,
                                                          const float* __restrict__ sd, const float* __restrict__ Jreg,
                                                          float* __restrict__ jt, float* __restrict__ jsd)
{
    // one block per (joint, component c, slot b) with b == NB meaning the template
    const int j = blockIdx.x / (3 * (NB + 1));
    const int rem = blockIdx.x % (3 * (NB + 1));
    const int c = rem / (NB + 1), b = rem % (NB + 1);
    double acc = 0.0;  // one-time, fp64 accumulation: more accurate than the reference's fp32 matmul
    for (int v = threadIdx.x; v < V; v += 256) {
        const float w = Jreg[(size_t)j * V + v];
        if (w != 0.0f) acc += (double)w * (double)(b == NB ? vt[3 * v + c] : sd[((size_t)v * 3 + c) * NB + b]);
    }
    __shared__ double s[256];
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (b == NB) jt[j * 3 + c] = (float)s[0];
        else jsd[(j * 3 + c) * NB + b] = (float)s[0];
    }
}

// ---------------------------------------------------------------------------
// pose: one warp
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mat3_mul(const float* A, const float* B, float* C)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}

__device__ __forceinline__ void mat3_inverse(const float* m, float* inv)
{
    const float c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    const float det = m[0] * c00 + m[1] * c01 + m[2] * c02;
    const float id = 1.0f / det;
    inv[0] = c00 * id; inv[1] = (m[2] * m[7] - m[1] * m[8]) * id; inv[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    inv[3] = c01 * id; inv[4] = (m[0] * m[8] - m[2] * m[6]) * id; inv[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    inv[6] = c02 * id; inv[7] = (m[1] * m[6] - m[0] * m[7]) * id; inv[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// Rodrigues (gaussian_model.py:894-912): angle = |r + 1e-8|, k = r/angle;  K = [0 -kz ky; kz 0 -kx; -ky kx 0];
// R = I + s K + c1 K K
__device__ __forceinline__ void rodrigues(float rx, float ry, float rz, float* R)
{
    const float ax = rx + 1e-8f, ay = ry + 1e-8f, az = rz + 1e-8f;
    const float angle = sqrtf(ax * ax + ay * ay + az * az);
    const float kx = rx / angle, ky = ry / angle, kz = rz / angle;
    const float s = sinf(angle), c1 = 1.0f - cosf(angle);
    R[0] = 1.0f + c1 * (-kz * kz - ky * ky); R[1] = s * -kz + c1 * (kx * ky); R[2] = s * ky + c1 * (kx * kz);
    R[3] = s * kz + c1 * (kx * ky); R[4] = 1.0f + c1 * (-kz * kz - kx * kx); R[5] = s * -kx + c1 * (ky * kz);
    R[6] = s * -ky + c1 * (kx * kz); R[7] = s * kx + c1 * (ky * kz); R[8] = 1.0f + c1 * (-ky * ky - kx * kx);
}

// rotation of joint j >= 1 as the pose features see it: Rodrigues(poses[j]) (* correct_Rs[j-1])
__device__ __forceinline__ void joint_rotation(const float* poses, const float* correct_Rs, int j, float* R)
{
    rodrigues(poses[3 * j], poses[3 * j + 1], poses[3 * j + 2], R);
    if (correct_Rs && j > 0) {
        float C[9], RC[9];
#pragma unroll
        for (int k = 0; k < 9; k++) C[k] = correct_Rs[9 * (j - 1) + k];
        mat3_mul(R, C, RC);
#pragma unroll
        for (int k = 0; k < 9; k++) R[k] = RC[k];
    }
}

struct PoseScratch {
    float rot[kJ][9];
    float joint[kJ][3];
    float G[kJ][12];
};

// The pose stage of one view, run by ONE warp (j = lane): rotations, pose features, rest joints, kinematic chain.
__device__ __forceinline__ void pose_chain(int j, int NB, const float* __restrict__ jt, const float* __restrict__ jsd,
                                           const int32_t* __restrict__ parents, const float* __restrict__ poses,
                                           const float* __restrict__ shapes, const float* __restrict__ correct_Rs,
                                           const float* __restrict__ Rg, const float* __restrict__ Th,
                                           float* __restrict__ out, PoseScratch& sc)
{
    if (j < kJ) {
        float R[9];
        rodrigues(poses[3 * j], poses[3 * j + 1], poses[3 * j + 2], R);
#pragma unroll
        for (int k = 0; k < 9; k++) out[GHR_POSE_ROD + 9 * j + k] = R[k];
        if (correct_Rs && j > 0) {
            float C[9], RC[9];
#pragma unroll
            for (int k = 0; k < 9; k++) C[k] = correct_Rs[9 * (j - 1) + k];
            mat3_mul(R, C, RC);
#pragma unroll
            for (int k = 0; k < 9; k++) R[k] = RC[k];
        }
#pragma unroll
        for (int k = 0; k < 9; k++) { sc.rot[j][k] = R[k]; out[GHR_POSE_ROT + 9 * j + k] = R[k]; }
        if (j > 0) {
#pragma unroll
            for (int k = 0; k < 9; k++) out[GHR_POSE_PF + 9 * (j - 1) + k] = R[k] - ((k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f);
        }
        // rest joints: J_regressor (v_template + shapedirs beta), folded (gaussian_model.py:951-970)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float v = jt[3 * j + c];
            for (int b = 0; b < NB; b++) v += jsd[(3 * j + c) * NB + b] * shapes[b];
            sc.joint[j][c] = v;
            out[GHR_POSE_JOINTS + 3 * j + c] = v;
        }
    }
    if (j < 9) { out[GHR_POSE_RG + j] = Rg[j]; }
    if (j < 3) { out[GHR_POSE_TH + j] = Th[j]; }
    __syncwarp();
    if (j == 0) {
        float inv[9], R[9];
#pragma unroll
        for (int k = 0; k < 9; k++) R[k] = Rg[k];
        mat3_inverse(R, inv);
#pragma unroll
        for (int k = 0; k < 9; k++) out[GHR_POSE_RGINV + k] = inv[k];
    }
    // kinematic chain (gaussian_model.py:914-944): G_0 = [R_0 | j_0]; G_i = G_parent [R_i | j_i - j_parent].
    // 24 dependent steps; lanes 0..11 each own one element (r,c) of the 3x4 matrix, so a step is ~4 FMAs deep.
    const int r = (j < 12) ? j / 4 : 0, c = (j < 12) ? j % 4 : 0;
    const int my_parent = (j > 0 && j < kJ) ? parents[j] : 0;  // one global load up front, not one per chain step
    for (int i = 0; i < kJ; i++) {
        float gval = 0.0f;
        const int p = __shfl_sync(0xffffffffu, my_parent, i);
        if (j < 12) {
            if (i == 0) {
                gval = c < 3 ? sc.rot[0][3 * r + c] : sc.joint[0][r];
            } else {
                const float g0 = sc.G[p][4 * r], g1 = sc.G[p][4 * r + 1], g2 = sc.G[p][4 * r + 2], g3 = sc.G[p][4 * r + 3];
                if (c < 3) {
                    gval = g0 * sc.rot[i][c] + g1 * sc.rot[i][3 + c] + g2 * sc.rot[i][6 + c];
                } else {
                    gval = g0 * (sc.joint[i][0] - sc.joint[p][0]) + g1 * (sc.joint[i][1] - sc.joint[p][1]) +
                           g2 * (sc.joint[i][2] - sc.joint[p][2]) + g3;
                }
            }
            sc.G[i][j] = gval;
            out[GHR_POSE_G + 12 * i + j] = gval;
        }
        __syncwarp();
        if (j < 12) {
            // A_i = G_i with translation - G_i^{3x3} j_i (:939-942)
            float a = gval;
            if (c == 3)
                a = gval - (sc.G[i][4 * r] * sc.joint[i][0] + sc.G[i][4 * r + 1] * sc.joint[i][1] + sc.G[i][4 * r + 2] * sc.joint[i][2]);
            out[GHR_POSE_A + 12 * i + j] = a;
        }
    }
}

__global__ void __launch_bounds__(32) lbs_pose_kernel(int NB, const float* __restrict__ jt, const float* __restrict__ jsd,
                                                      const int32_t* __restrict__ parents, const float* __restrict__ poses,
                                                      const float* __restrict__ shapes, const float* __restrict__ correct_Rs,
                                                      const float* __restrict__ Rg, const float* __restrict__ Th,
                                                      float* __restrict__ out)
{
    __shared__ PoseScratch sc;
    pose_chain(threadIdx.x, NB, jt, jsd, parents, poses, shapes, correct_Rs, Rg, Th, out, sc);
}

// ---------------------------------------------------------------------------
// vertex offsets: one warp per row r of posedirs [V*3, 207]
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lbs_vertex_offsets_kernel(int rows, int NB, const float* __restrict__ posedirs,
                                                                 const float* __restrict__ shapedirs,
                                                                 const float* __restrict__ pose_big,
                                                                 const float* __restrict__ pose_tgt,
                                                                 const float* __restrict__ shapes, float* __restrict__ off)
{
    __shared__ float s_dpf[GHR_LBS_POSE_DIM];
    for (int k = threadIdx.x; k < GHR_LBS_POSE_DIM; k += 256)
        s_dpf[k] = pose_tgt[GHR_POSE_PF + k] - (pose_big ? pose_big[GHR_POSE_PF + k] : 0.0f);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float* pr = posedirs + (size_t)row * GHR_LBS_POSE_DIM;
    float acc = 0.0f;
    for (int k = lane; k < GHR_LBS_POSE_DIM; k += 32) acc = fmaf(pr[k], s_dpf[k], acc);
    if (shapes && lane < NB) acc = fmaf(shapedirs[(size_t)row * NB + lane], shapes[lane], acc);
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) off[row] = acc;
}

// ---------------------------------------------------------------------------
// The pose stage of up to GHR_POSE_VIEWS_MAX views in ONE launch (the views of a multi-view batch): the first n blocks
// run one view's pose chain each (warp 0), the others are vertex-offset blocks that read every row of the 17 MB pose
// basis ONCE for all the views (the per-view kernel reads it once per view).  An offset block derives the pose
// features of its views itself (23 Rodrigues per view, the pose chain's own code), so it does not wait for the pose
// blocks; per view the arithmetic, and therefore the result, is that of lbs_pose_kernel / lbs_vertex_offsets_kernel.
// ---------------------------------------------------------------------------
struct PoseViewsArgs {
    int n, NB, rows;
    const float *jt, *jsd, *posedirs, *shapedirs, *pose_big;
    const int32_t* parents;
    ghr_pose_view v[GHR_POSE_VIEWS_MAX];
};

__global__ void __launch_bounds__(256) lbs_pose_views_kernel(const PoseViewsArgs a)
{
    __shared__ PoseScratch sc;
    __shared__ float s_dpf[GHR_POSE_VIEWS_MAX][GHR_LBS_POSE_DIM];
    if ((int)blockIdx.x < a.n) {
        if (threadIdx.x < 32) {
            const ghr_pose_view& v = a.v[blockIdx.x];
            pose_chain(threadIdx.x, a.NB, a.jt, a.jsd, a.parents, v.poses, v.shapes, v.correct_Rs, v.R, v.Th, v.pose_out, sc);
        }
        return;
    }
    for (int t = threadIdx.x; t < a.n * (kJ - 1); t += 256) {
        const int vi = t / (kJ - 1), j = 1 + t % (kJ - 1);
        float R[9];
        joint_rotation(a.v[vi].poses, a.v[vi].correct_Rs, j, R);
#pragma unroll
        for (int k = 0; k < 9; k++) {
            const float pf = R[k] - ((k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f);
            s_dpf[vi][9 * (j - 1) + k] = pf - (a.pose_big ? a.pose_big[GHR_POSE_PF + 9 * (j - 1) + k] : 0.0f);
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x - a.n) * 8 + (threadIdx.x >> 5);
    if (row >= a.rows) return;
    const float* pr = a.posedirs + (size_t)row * GHR_LBS_POSE_DIM;
    float acc[GHR_POSE_VIEWS_MAX];
#pragma unroll
    for (int vi = 0; vi < GHR_POSE_VIEWS_MAX; vi++) acc[vi] = 0.0f;
    for (int k = lane; k < GHR_LBS_POSE_DIM; k += 32) {
        const float pv = pr[k];
#pragma unroll
        for (int vi = 0; vi < GHR_POSE_VIEWS_MAX; vi++)
            if (vi < a.n) acc[vi] = fmaf(pv, s_dpf[vi][k], acc[vi]);
    }
    const float sd = (lane < a.NB) ? a.shapedirs[(size_t)row * a.NB + lane] : 0.0f;
#pragma unroll
    for (int vi = 0; vi < GHR_POSE_VIEWS_MAX; vi++) {
        if (vi >= a.n) break;
        float x = acc[vi];
        if (a.v[vi].shapes && lane < a.NB) x = fmaf(sd, a.v[vi].shapes[lane], x);
#pragma unroll
        for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) a.v[vi].offsets_out[row] = x;
    }
}

// dpf[k] = sum_r posedirs[r][k] * doff[r]  (transposed GEMV; 207 outputs)
__global__ void __launch_bounds__(256) lbs_vertex_offsets_backward_kernel(int rows, const float* __restrict__ posedirs,
                                                                          const float* __restrict__ doff,
                                                                          float* __restrict__ dpf /*[207], zeroed*/)
{
    // each block handles a slab of rows; thread k < 207 accumulates column k (coalesced across k)
    const int k = threadIdx.x;
    const int rows_per_block = (rows + gridDim.x - 1) / gridDim.x;
    const int r0 = blockIdx.x * rows_per_block, r1 = min(rows, r0 + rows_per_block);
    if (k >= GHR_LBS_POSE_DIM) return;
    float acc = 0.0f;
    for (int r = r0; r < r1; r++) {
        const float d = doff[r];
        if (d != 0.0f) acc = fmaf(posedirs[(size_t)r * GHR_LBS_POSE_DIM + k], d, acc);
    }
    atomicAdd(&dpf[k], acc);
}

// ---------------------------------------------------------------------------
// pose backward: dA[24][12], dpf[207] -> d correct_Rs[23][9]   (one thread; ~2k flops)
// ---------------------------------------------------------------------------
__global__ void lbs_pose_backward_kernel(const int32_t* __restrict__ parents, const float* __restrict__ pose,
                                         const float* __restrict__ dA, const float* __restrict__ dpf,
                                         float* __restrict__ dC)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float dG3[kJ][9], dg[kJ][3];
    for (int i = 0; i < kJ; i++) {
        const float* j = pose + GHR_POSE_JOINTS + 3 * i;
        for (int r = 0; r < 3; r++) {
            const float da = dA[12 * i + 4 * r + 3];
            dg[i][r] = da;
            for (int c = 0; c < 3; c++) dG3[i][3 * r + c] = dA[12 * i + 4 * r + c] - da * j[c];
        }
    }
    for (int i = kJ - 1; i >= 0; i--) {
        float dR[9];
        if (i == 0) {
            for (int k = 0; k < 9; k++) dR[k] = dG3[0][k];
        } else {
            const int p = parents[i];
            const float* Gp = pose + GHR_POSE_G + 12 * p;
            const float* Ri = pose + GHR_POSE_ROT + 9 * i;
            const float* ji = pose + GHR_POSE_JOINTS + 3 * i;
            const float* jp = pose + GHR_POSE_JOINTS + 3 * p;
            const float rel[3] = {ji[0] - jp[0], ji[1] - jp[1], ji[2] - jp[2]};
            // G_i3 = Gp3 R_i ; g_i = Gp3 rel + g_p
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++) {
                    // dGp3[r][c] += sum_k dG3_i[r][k] R_i[c][k] + dg_i[r] rel[c]
                    dG3[p][3 * r + c] += dG3[i][3 * r] * Ri[3 * c] + dG3[i][3 * r + 1] * Ri[3 * c + 1] +
                                         dG3[i][3 * r + 2] * Ri[3 * c + 2] + dg[i][r] * rel[c];
                    // dR_i[r][c] = sum_k Gp3[k][r] dG3_i[k][c]
                    dR[3 * r + c] = Gp[r] * dG3[i][c] + Gp[4 + r] * dG3[i][3 + c] + Gp[8 + r] * dG3[i][6 + c];
                }
            for (int r = 0; r < 3; r++) dg[p][r] += dg[i][r];
            // pose feature = rot[i] - I for i >= 1
            if (dpf)
                for (int k = 0; k < 9; k++) dR[k] += dpf[9 * (i - 1) + k];
            // rot_i = Rod_i C_i  ->  dC_i = Rod_i^T dR_i
            const float* Rod = pose + GHR_POSE_ROD + 9 * i;
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++)
                    dC[9 * (i - 1) + 3 * r + c] = Rod[r] * dR[c] + Rod[3 + r] * dR[3 + c] + Rod[6 + r] * dR[6 + c];
        }
    }
}

// ---------------------------------------------------------------------------
// smplx lbs() over the template vertices
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) smpl_lbs_vertices_kernel(int V, int NB, const float* __restrict__ vt,
                                                                const float* __restrict__ weights,
                                                                const float* __restrict__ offsets,
                                                                const float* __restrict__ pose, float* __restrict__ verts,
                                                                float* __restrict__ T_out)
{
    __shared__ float s_A[kJ * 12];
    for (int k = threadIdx.x; k < kJ * 12; k += 128) s_A[k] = pose[GHR_POSE_A + k];
    __syncthreads();
    const int v = blockIdx.x * 128 + threadIdx.x;
    if (v >= V) return;
    float T[12];
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = 0.0f;
    for (int j = 0; j < kJ; j++) {
        const float w = weights[(size_t)v * kJ + j];
#pragma unroll
        for (int k = 0; k < 12; k++) T[k] = fmaf(w, s_A[12 * j + k], T[k]);
    }
    const float x = vt[3 * v] + offsets[3 * v], y = vt[3 * v + 1] + offsets[3 * v + 1], z = vt[3 * v + 2] + offsets[3 * v + 2];
#pragma unroll
    for (int r = 0; r < 3; r++) verts[3 * v + r] = T[4 * r] * x + T[4 * r + 1] * y + T[4 * r + 2] * z + T[4 * r + 3];
    if (T_out)
#pragma unroll
        for (int k = 0; k < 12; k++) T_out[(size_t)v * 12 + k] = T[k];
}

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
cudaError_t launch_lbs_prepare(int V, int NB, const float* vt, const float* sd, const float* Jreg, float* jt, float* jsd,
                               cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_JOINTS, s);
    count_launch();
    lbs_prepare_kernel<<<kJ * 3 * (NB + 1), 256, 0, s>>>(V, NB, vt, sd, Jreg, jt, jsd);
    return cudaGetLastError();
}

cudaError_t launch_lbs_pose(const ghr_lbs_model& m, const float* poses, const float* shapes, const float* correct_Rs,
                            const float* R, const float* Th, float* out, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_JOINTS, s);
    count_launch();
    lbs_pose_kernel<<<1, 32, 0, s>>>(m.NB, m.joint_template, m.joint_shapedirs, m.parents, poses, shapes, correct_Rs, R,
                                     Th, out);
    return cudaGetLastError();
}

cudaError_t launch_lbs_vertex_offsets(const ghr_lbs_model& m, const float* pose_big, const float* pose_tgt,
                                      const float* shapes, float* off, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_JOINTS, s);
    count_launch();
    const int rows = m.V * 3;
    lbs_vertex_offsets_kernel<<<(rows + 7) / 8, 256, 0, s>>>(rows, m.NB, m.posedirs, m.shapedirs, pose_big, pose_tgt,
                                                             shapes, off);
    return cudaGetLastError();
}

cudaError_t launch_lbs_pose_views(const ghr_lbs_model& m, int n, const ghr_pose_view* views, const float* pose_big,
                                  cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_JOINTS, s);
    count_launch();
    PoseViewsArgs a;
    a.n = n; a.NB = m.NB; a.rows = m.V * 3;
    a.jt = m.joint_template; a.jsd = m.joint_shapedirs; a.posedirs = m.posedirs; a.shapedirs = m.shapedirs;
    a.pose_big = pose_big; a.parents = m.parents;
    for (int i = 0; i < GHR_POSE_VIEWS_MAX; i++) a.v[i] = views[i < n ? i : 0];
    lbs_pose_views_kernel<<<n + (a.rows + 7) / 8, 256, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_lbs_pose_backward(const ghr_lbs_model& m, const float* pose_tgt, const float* dA, const float* doff,
                                     float* dC, float* dpf_scratch, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_BWD, s);
    const float* dpf = nullptr;
    if (doff) {
        cudaError_t e = cudaMemsetAsync(dpf_scratch, 0, sizeof(float) * GHR_LBS_POSE_DIM, s);
        if (e != cudaSuccess) return e;
        count_launch();
        lbs_vertex_offsets_backward_kernel<<<sm_count(), 224, 0, s>>>(m.V * 3, m.posedirs, doff, dpf_scratch);
        dpf = dpf_scratch;
    }
    count_launch();
    lbs_pose_backward_kernel<<<1, 32, 0, s>>>(m.parents, pose_tgt, dA, dpf, dC);
    return cudaGetLastError();
}

cudaError_t launch_smpl_lbs_vertices(const ghr_lbs_model& m, const float* vt, const float* pose, const float* offsets,
                                     float* verts, float* T_out, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_FWD, s);
    count_launch();
    smpl_lbs_vertices_kernel<<<(m.V + 127) / 128, 128, 0, s>>>(m.V, m.NB, vt, m.weights, offsets, pose, verts, T_out);
    return cudaGetLastError();
}

}  // namespace ghr

// ---------------------------------------------------------------------------
// BodyPoseRefiner (reference nets/mlp_delta_body_pose.py:5-41,164-193): Linear(69, W) ReLU [Linear(W, W) ReLU]* Linear(W, 69)
// on ONE pose vector per view, then the module's own Rodrigues (theta = sqrt(1e-5 + |r|^2)) -> correct_Rs [23,3,3].
// ~35 kFLOP: the reference spends eight launches on it; here it is one block -- a warp per output row, lanes striding the
// input (coalesced weight reads), shuffle reduction -- so the whole pose stage of a view stays a couple of launches.
// ---------------------------------------------------------------------------
namespace ghr {

struct RefinerArgs {
    int width, n_layers;              // n_layers linear layers in all (>= 2)
    const float* w[GHR_REFINER_MAX_LAYERS];
    const float* b[GHR_REFINER_MAX_LAYERS];
    const float* pose;                // [69]
    float* Rs;                        // [23*9]
};

__global__ void __launch_bounds__(256) pose_refiner_kernel(const RefinerArgs a)
{
    __shared__ float s_x[2][GHR_REFINER_MAX_WIDTH];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t < 69) s_x[0][t] = a.pose[t];
    __syncthreads();
    int cur = 0, n_in = 69;
    for (int l = 0; l < a.n_layers; l++) {
        const int n_out = l == a.n_layers - 1 ? 69 : a.width;
        const float* W = a.w[l];
        for (int r = warp; r < n_out; r += 8) {
            float acc = 0.0f;
            for (int k = lane; k < n_in; k += 32) acc = fmaf(W[(size_t)r * n_in + k], s_x[cur][k], acc);
#pragma unroll
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) {
                acc += a.b[l][r];
                s_x[cur ^ 1][r] = l == a.n_layers - 1 ? acc : fmaxf(acc, 0.0f);
            }
        }
        __syncthreads();
        cur ^= 1;
        n_in = n_out;
    }
    if (t < 23) {
        const float rx = s_x[cur][3 * t], ry = s_x[cur][3 * t + 1], rz = s_x[cur][3 * t + 2];
        const float theta = sqrtf(1e-5f + rx * rx + ry * ry + rz * rz);
        const float x = rx / theta, y = ry / theta, z = rz / theta;
        const float c = cosf(theta), s = sinf(theta);
        float* R = a.Rs + 9 * t;
        R[0] = x * x + (1.f - x * x) * c; R[1] = x * y * (1.f - c) - z * s; R[2] = x * z * (1.f - c) + y * s;
        R[3] = x * y * (1.f - c) + z * s; R[4] = y * y + (1.f - y * y) * c; R[5] = y * z * (1.f - c) - x * s;
        R[6] = x * z * (1.f - c) - y * s; R[7] = y * z * (1.f - c) + x * s; R[8] = z * z + (1.f - z * z) * c;
    }
}

cudaError_t launch_pose_refiner(int width, int n_layers, const float* const* w, const float* const* b, const float* pose,
                                float* Rs, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_JOINTS, s);
    count_launch();
    RefinerArgs a;
    a.width = width; a.n_layers = n_layers; a.pose = pose; a.Rs = Rs;
    for (int i = 0; i < GHR_REFINER_MAX_LAYERS; i++) { a.w[i] = i < n_layers ? w[i] : nullptr; a.b[i] = i < n_layers ? b[i] : nullptr; }
    pose_refiner_kernel<<<1, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace ghr
