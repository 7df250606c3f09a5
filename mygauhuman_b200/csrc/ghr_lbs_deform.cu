// Fused per-Gaussian LBS deformation + world covariance + view-dependent colour,
// forward and backward (one thread per Gaussian).
//
// Reference: scene/gaussian_model.py:768-872 coarse_deform_c2source,
// :35-42 build_covariance_from_scaling_rotation (+ utils/general_utils.py:60-117),
// gaussian_renderer/__init__.py:124-125,191-195 (view directions, eval_sh, clamp),
// utils/sh_utils.py:135-190.  The reference runs ~60 ATen kernels forward and
// ~150 backward for this (two batched 3x3 torch.inverse among them); here it is
// one streaming kernel each way.  The 24 joint matrices of both poses live in
// shared memory; the skinning-weight row of a Gaussian's nearest vertex is a
// 96-byte gather served by L2 (6,890 rows = 661 KB).
//
// Roofline: HBM stream.  Algorithmic bytes per Gaussian, forward:
//   read  xyz 12 + normal 12 + vid 4 + weight row 96 (+96 offsets) + scale 12 + quat 16 + SH 12*M
//   write world 12 + transform 36 + normal 12 + cov 24 + colour 12 + clamp 1
//   = 453 B at M=16 without offsets.  Backward re-reads the inputs, reads the five
//   incoming gradients (96 B) and writes 12+12+12+16+12*M (+96) B.
#include <stdio.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"

#include "ghr_deform.cuh"

namespace ghr {


__global__ void __launch_bounds__(128) lbs_deform_forward_kernel(const ghr_deform_params p, const DeformOut o)
{
    __shared__ DeformShared s;
    load_shared(p, s);
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= p.P) return;
    const int vid = p.vert_ids[idx];
    DeformState d;
    float xw[3];
    deform_forward_core(p, s, idx, vid, d, xw);
#pragma unroll
    for (int c = 0; c < 3; c++) o.world_pts[3 * (size_t)idx + c] = xw[c];
#pragma unroll
    for (int k = 0; k < 9; k++) o.transforms[9 * (size_t)idx + k] = d.Tw[k];
    if (o.smpl_pts)
#pragma unroll
        for (int c = 0; c < 3; c++) o.smpl_pts[3 * (size_t)idx + c] = d.x3[c];
    if (o.bweights)
#pragma unroll
        for (int j = 0; j < kJ; j++) o.bweights[kJ * (size_t)idx + j] = d.bw[j];
    if (o.world_normals && p.normals) {
        const float n[3] = {p.normals[3 * idx], p.normals[3 * idx + 1], p.normals[3 * idx + 2]};
        float n3[3], nw[3];
        mat3_vec(d.M, n, n3);  // At3 (Rinv n)
        mat3_tvec(s.Rginv, n3, nw);
#pragma unroll
        for (int c = 0; c < 3; c++) o.world_normals[3 * (size_t)idx + c] = nw[c];
    }
    if (o.translation) {
        // translation of the per-Gaussian affine map x_world = transforms x + translation (:796-799,815,824,844,859,869)
        const float nt[3] = {-d.Ab[3], -d.Ab[7], -d.Ab[11]};
        float t1[3], t3[3], tw[3];
        mat3_vec(d.Rinv, nt, t1);
#pragma unroll
        for (int c = 0; c < 3; c++) t1[c] += p.vertex_offsets[3 * (size_t)vid + c];
        const float At3[9] = {d.At[0], d.At[1], d.At[2], d.At[4], d.At[5], d.At[6], d.At[8], d.At[9], d.At[10]};
        mat3_vec(At3, t1, t3);
        t3[0] += d.At[3]; t3[1] += d.At[7]; t3[2] += d.At[11];
        mat3_tvec(s.Rginv, t3, tw);
#pragma unroll
        for (int c = 0; c < 3; c++) o.translation[3 * (size_t)idx + c] = tw[c] + s.Th[c];
    }
    if (p.opacity_logits && p.opacity_out)  // get_opacity (gaussian_model.py:174): sigmoid
        p.opacity_out[idx] = 1.0f / (1.0f + expf(-p.opacity_logits[idx]));
    if (o.cov3D && p.scales) {
        DeformCov k;
        float c3[6];
        deform_cov_forward(p, idx, d.Tw, k, c3);
        float* co = o.cov3D + 6 * (size_t)idx;
#pragma unroll
        for (int i = 0; i < 6; i++) co[i] = c3[i];
    }
    if (o.colors && p.shs) {
        float sh[48];
        load_sh_row(p.shs, idx, p.M, sh);
        float rgb[3];
        uint32_t cm;
        sh_to_rgb(p.sh_degree, sh, xw[0] - s.campos[0], xw[1] - s.campos[1], xw[2] - s.campos[2], rgb, cm);
#pragma unroll
        for (int c = 0; c < 3; c++) o.colors[3 * (size_t)idx + c] = rgb[c];
        if (o.clamped) o.clamped[idx] = (uint8_t)cm;
    }
}


__global__ void __launch_bounds__(128) lbs_deform_backward_kernel(const ghr_deform_params p, const DeformGrads g)
{
    __shared__ DeformShared s;
    __shared__ float s_dA[kJ * 12];
    load_shared(p, s);
    if (g.d_A_target) {
        for (int k = threadIdx.x; k < kJ * 12; k += 128) s_dA[k] = 0.0f;
        __syncthreads();
    }
    const int idx = blockIdx.x * 128 + threadIdx.x;
    float dAt[12];
    float bw_keep[kJ];
#pragma unroll
    for (int k = 0; k < 12; k++) dAt[k] = 0.0f;
#pragma unroll
    for (int j = 0; j < kJ; j++) bw_keep[j] = 0.0f;
    // per-Gaussian gradients are stored, or added to what the buffers hold (multi-view gradient accumulation)
    // (2 = with atomics: concurrent views on several streams add into one buffer, MultiLaneRenderStep)
    if (idx < p.P) {
        const int vid = p.vert_ids[idx];
        DeformState d;
        float xw[3];
        deform_forward_core(p, s, idx, vid, d, xw);
        DeformUpstream up;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            up.gxw[c] = g.g_world ? g.g_world[3 * idx + c] : 0.0f;
            up.gcol[c] = g.g_colors ? g.g_colors[3 * idx + c] : 0.0f;
            up.gn[c] = g.g_normals ? g.g_normals[3 * idx + c] : 0.0f;
        }
#pragma unroll
        for (int k = 0; k < 9; k++) up.dTw[k] = g.g_transforms ? g.g_transforms[9 * (size_t)idx + k] : 0.0f;
#pragma unroll
        for (int k = 0; k < 6; k++) up.gcov[k] = g.g_cov ? g.g_cov[6 * (size_t)idx + k] : 0.0f;
        up.has_cov = g.g_cov != nullptr; up.has_col = g.g_colors != nullptr; up.has_n = g.g_normals != nullptr;
        up.gop = p.dL_dopacity ? p.dL_dopacity[idx] : 0.0f;
        up.clamp_mask = g.clamped ? g.clamped[idx] : 0u;
        deform_backward_gaussian(p, s, g, idx, vid, d, xw, up, dAt, bw_keep);
    }
    if (g.d_A_target) reduce_pose_grads(s_dA, g.d_A_target, bw_keep, dAt);
}

cudaError_t launch_lbs_deform_forward(const ghr_deform_params& p, const DeformOut& o, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_FWD, s);
    count_launch();
    lbs_deform_forward_kernel<<<(p.P + 127) / 128, 128, 0, s>>>(p, o);
    return cudaGetLastError();
}

cudaError_t launch_lbs_deform_backward(const ghr_deform_params& p, const DeformGrads& g, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_LBS_BWD, s);
    count_launch();
    lbs_deform_backward_kernel<<<(p.P + 127) / 128, 128, 0, s>>>(p, g);
    return cudaGetLastError();
}

// ---- stand-alone operators: the same math as the fused kernels above, for callers that use
// GaussianModel.get_covariance (scene/gaussian_model.py:35-42,208-209) or eval_sh + clamp
// (gaussian_renderer/__init__.py:124-125,191-195; utils/sh_utils.py:135-190) on their own. ----------------------------

// Sigma' = (T L)(T L)^T, L = R(q/|q|) diag(mod * s), packed [xx,xy,xz,yy,yz,zz]; T = identity when transforms == NULL
__global__ void __launch_bounds__(128) covariance_forward_kernel(int P, const float* __restrict__ scaling,
                                                                 const float* __restrict__ rotation, float modifier,
                                                                 const float* __restrict__ transforms,
                                                                 float* __restrict__ cov)
{
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= P) return;
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotation) + idx);
    float Rq[9], inv_norm;
    quat_rot(q.x, q.y, q.z, q.w, Rq, inv_norm);
    const float sc[3] = {modifier * scaling[3 * idx], modifier * scaling[3 * idx + 1], modifier * scaling[3 * idx + 2]};
    float L[9], TL[9];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) L[3 * r + c] = Rq[3 * r + c] * sc[c];
    if (transforms) {
        float T[9];
#pragma unroll
        for (int k = 0; k < 9; k++) T[k] = transforms[9 * (size_t)idx + k];
        mat3_mul_d(T, L, TL);
    } else {
#pragma unroll
        for (int k = 0; k < 9; k++) TL[k] = L[k];
    }
    float* co = cov + 6 * (size_t)idx;
    co[0] = TL[0] * TL[0] + TL[1] * TL[1] + TL[2] * TL[2];
    co[1] = TL[0] * TL[3] + TL[1] * TL[4] + TL[2] * TL[5];
    co[2] = TL[0] * TL[6] + TL[1] * TL[7] + TL[2] * TL[8];
    co[3] = TL[3] * TL[3] + TL[4] * TL[4] + TL[5] * TL[5];
    co[4] = TL[3] * TL[6] + TL[4] * TL[7] + TL[5] * TL[8];
    co[5] = TL[6] * TL[6] + TL[7] * TL[7] + TL[8] * TL[8];
}

__global__ void __launch_bounds__(128) covariance_backward_kernel(int P, const float* __restrict__ scaling,
                                                                  const float* __restrict__ rotation, float modifier,
                                                                  const float* __restrict__ transforms,
                                                                  const float* __restrict__ g_cov,
                                                                  float* __restrict__ d_scaling,
                                                                  float* __restrict__ d_rotation,
                                                                  float* __restrict__ d_transforms)
{
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= P) return;
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotation) + idx);
    float Rq[9], inv_norm;
    quat_rot(q.x, q.y, q.z, q.w, Rq, inv_norm);
    const float sc[3] = {modifier * scaling[3 * idx], modifier * scaling[3 * idx + 1], modifier * scaling[3 * idx + 2]};
    float L[9], TL[9], T[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) L[3 * r + c] = Rq[3 * r + c] * sc[c];
    if (transforms) {
#pragma unroll
        for (int k = 0; k < 9; k++) T[k] = transforms[9 * (size_t)idx + k];
    }
    mat3_mul_d(T, L, TL);
    const float* gc = g_cov + 6 * (size_t)idx;
    // dS' + dS'^T for the upper-triangular packing [xx,xy,xz,yy,yz,zz] (strip_symmetric, general_utils.py:64-72)
    const float S2[9] = {2.f * gc[0], gc[1], gc[2], gc[1], 2.f * gc[3], gc[4], gc[2], gc[4], 2.f * gc[5]};
    float dTL[9], dL[9], t[9];
    mat3_mul_d(S2, TL, dTL);
    if (d_transforms) {
        mat3_mult_d(dTL, L, t);  // dT = dTL L^T
#pragma unroll
        for (int k = 0; k < 9; k++) d_transforms[9 * (size_t)idx + k] = t[k];
    }
    mat3_tmul_d(T, dTL, dL);  // dL = T^T dTL
    float dsc[3], dR[9];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        dsc[c] = (Rq[c] * dL[c] + Rq[3 + c] * dL[3 + c] + Rq[6 + c] * dL[6 + c]) * modifier;
        dR[c] = dL[c] * sc[c]; dR[3 + c] = dL[3 + c] * sc[c]; dR[6 + c] = dL[6 + c] * sc[c];
    }
    const float r = q.x * inv_norm, x = q.y * inv_norm, y = q.z * inv_norm, z = q.w * inv_norm;
    const float dr_ = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
    const float dx_ = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - r * dR[5] + z * dR[6] + r * dR[7]) - 4.f * x * (dR[4] + dR[8]);
    const float dy_ = 2.f * (x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7]) - 4.f * y * (dR[0] + dR[8]);
    const float dz_ = 2.f * (-r * dR[1] + x * dR[2] + r * dR[3] + y * dR[5] + x * dR[6] + y * dR[7]) - 4.f * z * (dR[0] + dR[4]);
    const float dot = r * dr_ + x * dx_ + y * dy_ + z * dz_;  // q^ = q/|q|: dq = (dq^ - q^ (q^ . dq^)) / |q|
    d_scaling[3 * idx] = dsc[0]; d_scaling[3 * idx + 1] = dsc[1]; d_scaling[3 * idx + 2] = dsc[2];
    reinterpret_cast<float4*>(d_rotation)[idx] = make_float4((dr_ - r * dot) * inv_norm, (dx_ - x * dot) * inv_norm,
                                                             (dy_ - y * dot) * inv_norm, (dz_ - z * dot) * inv_norm);
}

// colors = clamp_min(eval_sh(deg, shs, (means - campos)/|.|) + 0.5, 0); clamped[i] bit c = channel c was clamped
__global__ void __launch_bounds__(128) sh_colors_forward_kernel(int P, int deg, int M, const float* __restrict__ shs,
                                                                const float* __restrict__ means,
                                                                const float* __restrict__ campos,
                                                                float* __restrict__ colors, uint8_t* __restrict__ clamped)
{
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= P) return;
    float sh[48];
    const int nf = min(M * 3, 48);
    const float* sp = shs + (size_t)idx * M * 3;
#pragma unroll
    for (int k = 0; k < 48; k++)
        if (k < nf) sh[k] = __ldg(sp + k);
    float rgb[3];
    uint32_t cm;
    sh_to_rgb(deg, sh, means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1], means[3 * idx + 2] - campos[2], rgb, cm);
    colors[3 * (size_t)idx] = rgb[0]; colors[3 * (size_t)idx + 1] = rgb[1]; colors[3 * (size_t)idx + 2] = rgb[2];
    if (clamped) clamped[idx] = (uint8_t)cm;
}

__global__ void __launch_bounds__(128) sh_colors_backward_kernel(int P, int deg, int M, const float* __restrict__ shs,
                                                                 const float* __restrict__ means,
                                                                 const float* __restrict__ campos,
                                                                 const uint8_t* __restrict__ clamped,
                                                                 const float* __restrict__ g_colors,
                                                                 float* __restrict__ d_shs, float* __restrict__ d_means)
{
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= P) return;
    const float gc[3] = {g_colors[3 * idx], g_colors[3 * idx + 1], g_colors[3 * idx + 2]};
    float dm[3] = {0.f, 0.f, 0.f};
    sh_backward(deg, M, shs + (size_t)idx * M * 3, means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1],
                means[3 * idx + 2] - campos[2], clamped ? clamped[idx] : 0u, gc, dm, d_shs + (size_t)idx * M * 3, 0);
    if (d_means) { d_means[3 * idx] = dm[0]; d_means[3 * idx + 1] = dm[1]; d_means[3 * idx + 2] = dm[2]; }
}

}  // namespace ghr

using namespace ghr;

extern "C" {

static int ops_status(const char* what)
{
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return GHR_OK;
    fprintf(stderr, "[libghr] CUDA error in %s: %s\n", what, cudaGetErrorString(e));
    return GHR_ERR_CUDA;
}

int ghr_covariance_forward(int32_t P, const float* scaling, const float* rotation, float scaling_modifier,
                           const float* transforms, float* cov3D, void* stream)
{
    if (P < 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!scaling || !rotation || !cov3D || (reinterpret_cast<uintptr_t>(rotation) & 15)) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_LBS_FWD, s);
    count_launch();
    covariance_forward_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, scaling, rotation, scaling_modifier, transforms, cov3D);
    return ops_status("covariance_forward");
}

int ghr_covariance_backward(int32_t P, const float* scaling, const float* rotation, float scaling_modifier,
                            const float* transforms, const float* dL_dcov3D, float* dL_dscaling, float* dL_drotation,
                            float* dL_dtransforms, void* stream)
{
    if (P < 0) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!scaling || !rotation || !dL_dcov3D || !dL_dscaling || !dL_drotation ||
        (reinterpret_cast<uintptr_t>(rotation) & 15) || (reinterpret_cast<uintptr_t>(dL_drotation) & 15))
        return GHR_ERR_BAD_ARG;
    if (dL_dtransforms && !transforms) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_LBS_BWD, s);
    count_launch();
    covariance_backward_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, scaling, rotation, scaling_modifier, transforms, dL_dcov3D,
                                                               dL_dscaling, dL_drotation, dL_dtransforms);
    return ops_status("covariance_backward");
}

int ghr_sh_colors_forward(int32_t P, int32_t sh_degree, int32_t M, const float* shs, const float* means3D,
                          const float* campos, float* colors, uint8_t* clamped, void* stream)
{
    if (P < 0 || sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1) || M > 16) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!shs || !means3D || !campos || !colors) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_LBS_FWD, s);
    count_launch();
    sh_colors_forward_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, sh_degree, M, shs, means3D, campos, colors, clamped);
    return ops_status("sh_colors_forward");
}

int ghr_sh_colors_backward(int32_t P, int32_t sh_degree, int32_t M, const float* shs, const float* means3D,
                           const float* campos, const uint8_t* clamped, const float* dL_dcolors, float* dL_dshs,
                           float* dL_dmeans3D, void* stream)
{
    if (P < 0 || sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1) || M > 16) return GHR_ERR_BAD_ARG;
    if (P == 0) return GHR_OK;
    if (!shs || !means3D || !campos || !dL_dcolors || !dL_dshs) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_LBS_BWD, s);
    count_launch();
    sh_colors_backward_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, sh_degree, M, shs, means3D, campos, clamped, dL_dcolors,
                                                              dL_dshs, dL_dmeans3D);
    return ops_status("sh_colors_backward");
}

}  // extern "C"
