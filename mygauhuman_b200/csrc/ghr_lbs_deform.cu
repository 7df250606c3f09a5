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

namespace ghr {

constexpr int kJ = GHR_LBS_JOINTS;

struct DeformState {
    float bw[kJ];
    float Ab[12], At[12];  // blended 3x4 transforms (row-major), big pose / target pose
    float Rinv[9];         // inverse of Ab's 3x3
    float xc[3];           // x - t_big
    float x2[3];           // T-pose point after blend-shape offsets
    float x3[3];           // target-pose (SMPL space) point
    float M[9];            // At3 * Rinv
    float Tw[9];           // Rg * M
};

__device__ __forceinline__ void mat3_mul_d(const float* A, const float* B, float* C)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
// C = A^T * B
__device__ __forceinline__ void mat3_tmul_d(const float* A, const float* B, float* C)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) C[3 * r + c] = A[r] * B[c] + A[3 + r] * B[3 + c] + A[6 + r] * B[6 + c];
}
// C = A * B^T
__device__ __forceinline__ void mat3_mult_d(const float* A, const float* B, float* C)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++)
            C[3 * r + c] = A[3 * r] * B[3 * c] + A[3 * r + 1] * B[3 * c + 1] + A[3 * r + 2] * B[3 * c + 2];
}
__device__ __forceinline__ void mat3_vec(const float* A, const float* v, float* o)
{
#pragma unroll
    for (int r = 0; r < 3; r++) o[r] = A[3 * r] * v[0] + A[3 * r + 1] * v[1] + A[3 * r + 2] * v[2];
}
__device__ __forceinline__ void mat3_tvec(const float* A, const float* v, float* o)
{
#pragma unroll
    for (int r = 0; r < 3; r++) o[r] = A[r] * v[0] + A[3 + r] * v[1] + A[6 + r] * v[2];
}
__device__ __forceinline__ void mat3_inv_d(const float* m, float* inv)
{
    const float c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    const float id = 1.0f / (m[0] * c00 + m[1] * c01 + m[2] * c02);
    inv[0] = c00 * id; inv[1] = (m[2] * m[7] - m[1] * m[8]) * id; inv[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    inv[3] = c01 * id; inv[4] = (m[0] * m[8] - m[2] * m[6]) * id; inv[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    inv[6] = c02 * id; inv[7] = (m[1] * m[6] - m[0] * m[7]) * id; inv[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// Shared per-block constants: A of both poses, global R, R^-1, Th, camera position.
struct DeformShared {
    float A_big[kJ * 12];
    float A_tgt[kJ * 12];
    float Rg[9], Rginv[9], Th[3], campos[3];
};

__device__ __forceinline__ void load_shared(const ghr_deform_params& p, DeformShared& s)
{
    for (int k = threadIdx.x; k < kJ * 12; k += blockDim.x) {
        s.A_big[k] = p.pose_big[GHR_POSE_A + k];
        s.A_tgt[k] = p.pose_target[GHR_POSE_A + k];
    }
    if (threadIdx.x < 9) { s.Rg[threadIdx.x] = p.pose_target[GHR_POSE_RG + threadIdx.x]; s.Rginv[threadIdx.x] = p.pose_target[GHR_POSE_RGINV + threadIdx.x]; }
    if (threadIdx.x < 3) { s.Th[threadIdx.x] = p.pose_target[GHR_POSE_TH + threadIdx.x]; s.campos[threadIdx.x] = p.campos ? p.campos[threadIdx.x] : 0.0f; }
    __syncthreads();
}

// forward of the skinning part for Gaussian idx (gaussian_model.py:775-870)
__device__ __forceinline__ void deform_forward_core(const ghr_deform_params& p, const DeformShared& s, int idx, int vid,
                                                    DeformState& d, float* xw)
{
    // skinning weights of the nearest vertex, optionally softmax(log(w + 1e-9) + offset) (:776-781)
    const float4* wr = reinterpret_cast<const float4*>(p.weights + (size_t)vid * kJ);
#pragma unroll
    for (int k = 0; k < 6; k++) {
        const float4 w = __ldg(wr + k);
        d.bw[4 * k] = w.x; d.bw[4 * k + 1] = w.y; d.bw[4 * k + 2] = w.z; d.bw[4 * k + 3] = w.w;
    }
    if (p.lbs_offsets) {
        float mx = -3.4e38f;
#pragma unroll
        for (int j = 0; j < kJ; j++) {
            d.bw[j] = logf(d.bw[j] + 1e-9f) + p.lbs_offsets[(size_t)idx * p.lbs_stride_p + (size_t)j * p.lbs_stride_j];
            mx = fmaxf(mx, d.bw[j]);
        }
        float sum = 0.0f;
#pragma unroll
        for (int j = 0; j < kJ; j++) { d.bw[j] = expf(d.bw[j] - mx); sum += d.bw[j]; }
        const float inv = 1.0f / sum;
#pragma unroll
        for (int j = 0; j < kJ; j++) d.bw[j] *= inv;
    }
#pragma unroll
    for (int k = 0; k < 12; k++) { d.Ab[k] = 0.0f; d.At[k] = 0.0f; }
#pragma unroll
    for (int j = 0; j < kJ; j++) {
        const float w = d.bw[j];
#pragma unroll
        for (int k = 0; k < 12; k++) {
            d.Ab[k] = fmaf(w, s.A_big[12 * j + k], d.Ab[k]);
            d.At[k] = fmaf(w, s.A_tgt[12 * j + k], d.At[k]);
        }
    }
    const float Ab3[9] = {d.Ab[0], d.Ab[1], d.Ab[2], d.Ab[4], d.Ab[5], d.Ab[6], d.Ab[8], d.Ab[9], d.Ab[10]};
    const float At3[9] = {d.At[0], d.At[1], d.At[2], d.At[4], d.At[5], d.At[6], d.At[8], d.At[9], d.At[10]};
    mat3_inv_d(Ab3, d.Rinv);
    d.xc[0] = p.xyz[3 * idx] - d.Ab[3]; d.xc[1] = p.xyz[3 * idx + 1] - d.Ab[7]; d.xc[2] = p.xyz[3 * idx + 2] - d.Ab[11];
    float x1[3];
    mat3_vec(d.Rinv, d.xc, x1);
#pragma unroll
    for (int c = 0; c < 3; c++) d.x2[c] = x1[c] + p.vertex_offsets[3 * (size_t)vid + c];
    mat3_vec(At3, d.x2, d.x3);
    d.x3[0] += d.At[3]; d.x3[1] += d.At[7]; d.x3[2] += d.At[11];
    // world = smpl @ R^-1 + Th  (row vector times matrix, :864)
    mat3_tvec(s.Rginv, d.x3, xw);
    xw[0] += s.Th[0]; xw[1] += s.Th[1]; xw[2] += s.Th[2];
    mat3_mul_d(At3, d.Rinv, d.M);
    mat3_mul_d(s.Rg, d.M, d.Tw);
}

// utils/general_utils.py:76-99 build_rotation (normalises q)
__device__ __forceinline__ void quat_rot(float qr, float qx, float qy, float qz, float* R, float& inv_norm)
{
    inv_norm = 1.0f / sqrtf(qr * qr + qx * qx + qy * qy + qz * qz);
    const float r = qr * inv_norm, x = qx * inv_norm, y = qy * inv_norm, z = qz * inv_norm;
    R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
    R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
    R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
}


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
        // L = R(q) diag(mod*s); Sigma = L L^T; Sigma' = Tw Sigma Tw^T (gaussian_model.py:35-42)
        const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
        float Rq[9], inv_norm;
        quat_rot(q.x, q.y, q.z, q.w, Rq, inv_norm);
        float sa[3] = {p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]};
        if (p.scales_are_log) { sa[0] = expf(sa[0]); sa[1] = expf(sa[1]); sa[2] = expf(sa[2]); }
        const float sc[3] = {p.scale_modifier * sa[0], p.scale_modifier * sa[1], p.scale_modifier * sa[2]};
        float L[9], TL[9];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 3; c++) L[3 * r + c] = Rq[3 * r + c] * sc[c];
        mat3_mul_d(d.Tw, L, TL);  // Sigma' = (Tw L)(Tw L)^T
        float* co = o.cov3D + 6 * (size_t)idx;
        co[0] = TL[0] * TL[0] + TL[1] * TL[1] + TL[2] * TL[2];
        co[1] = TL[0] * TL[3] + TL[1] * TL[4] + TL[2] * TL[5];
        co[2] = TL[0] * TL[6] + TL[1] * TL[7] + TL[2] * TL[8];
        co[3] = TL[3] * TL[3] + TL[4] * TL[4] + TL[5] * TL[5];
        co[4] = TL[3] * TL[6] + TL[4] * TL[7] + TL[5] * TL[8];
        co[5] = TL[6] * TL[6] + TL[7] * TL[7] + TL[8] * TL[8];
    }
    if (o.colors && p.shs) {
        float sh[48];
        const int nf = min(p.M * 3, 48);
        const float* sp = p.shs + (size_t)idx * p.M * 3;
        if ((nf & 3) == 0 && ((reinterpret_cast<uintptr_t>(sp) & 15) == 0)) {
            const float4* sp4 = reinterpret_cast<const float4*>(sp);
#pragma unroll
            for (int k = 0; k < 12; k++)
                if (4 * k < nf) {
                    const float4 v = __ldg(sp4 + k);
                    sh[4 * k] = v.x; sh[4 * k + 1] = v.y; sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
                }
        } else {
#pragma unroll
            for (int k = 0; k < 48; k++)
                if (k < nf) sh[k] = __ldg(sp + k);
        }
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
    const bool active = idx < p.P;
    float dAt[12];
    float bw_keep[kJ];
#pragma unroll
    for (int k = 0; k < 12; k++) dAt[k] = 0.0f;
#pragma unroll
    for (int j = 0; j < kJ; j++) bw_keep[j] = 0.0f;
    // per-Gaussian gradients are stored, or added to what the buffers hold (multi-view gradient accumulation)
    // (2 = with atomics: concurrent views on several streams add into one buffer, MultiLaneRenderStep)
    const int acc = p.accumulate_grads;
    auto put = [acc](float* dst, float v) {
        if (acc == 2) atomicAdd(dst, v);
        else *dst = acc ? *dst + v : v;
    };
    if (active) {
        const int vid = p.vert_ids[idx];
        DeformState d;
        float xw[3];
        deform_forward_core(p, s, idx, vid, d, xw);
        const float At3[9] = {d.At[0], d.At[1], d.At[2], d.At[4], d.At[5], d.At[6], d.At[8], d.At[9], d.At[10]};

        float gxw[3] = {0.f, 0.f, 0.f};
        if (g.g_world) { gxw[0] = g.g_world[3 * idx]; gxw[1] = g.g_world[3 * idx + 1]; gxw[2] = g.g_world[3 * idx + 2]; }
        float dTw[9];
#pragma unroll
        for (int k = 0; k < 9; k++) dTw[k] = g.g_transforms ? g.g_transforms[9 * (size_t)idx + k] : 0.0f;

        // ---- colour: clamp_min(eval_sh + 0.5, 0) with dirs from the deformed mean ----
        if (p.shs && g.d_shs) {
            float* dsh = g.d_shs + (size_t)idx * p.M * 3;
            if (g.g_colors) {
                const float gc[3] = {g.g_colors[3 * idx], g.g_colors[3 * idx + 1], g.g_colors[3 * idx + 2]};
                sh_backward(p.sh_degree, p.M, p.shs + (size_t)idx * p.M * 3, xw[0] - s.campos[0], xw[1] - s.campos[1],
                            xw[2] - s.campos[2], g.clamped ? g.clamped[idx] : 0u, gc, gxw, dsh, acc);
            } else if (!acc) {
                for (int k = 0; k < p.M * 3; k++) dsh[k] = 0.0f;
            }
        }

        if (p.opacity_logits && p.dL_dopacity_logits) {
            const float sg = 1.0f / (1.0f + expf(-p.opacity_logits[idx]));
            put(p.dL_dopacity_logits + idx, p.dL_dopacity ? p.dL_dopacity[idx] * sg * (1.0f - sg) : 0.0f);
        }
        // ---- covariance: Sigma' = (Tw L)(Tw L)^T, L = R(q^) diag(mod*s) ----
        if (p.scales && g.d_scales) {
            float dsc[3] = {0.f, 0.f, 0.f}, dq[4] = {0.f, 0.f, 0.f, 0.f};
            if (g.g_cov) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
                float Rq[9], inv_norm;
                quat_rot(q.x, q.y, q.z, q.w, Rq, inv_norm);
                float sa[3] = {p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]};
                if (p.scales_are_log) { sa[0] = expf(sa[0]); sa[1] = expf(sa[1]); sa[2] = expf(sa[2]); }
                const float sc[3] = {p.scale_modifier * sa[0], p.scale_modifier * sa[1], p.scale_modifier * sa[2]};
                float L[9], TL[9];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++) L[3 * r + c] = Rq[3 * r + c] * sc[c];
                mat3_mul_d(d.Tw, L, TL);
                const float* gc = g.g_cov + 6 * (size_t)idx;
                // dS' + dS'^T for the upper-triangular packing [xx,xy,xz,yy,yz,zz] (strip_symmetric)
                const float S2[9] = {2.f * gc[0], gc[1], gc[2], gc[1], 2.f * gc[3], gc[4], gc[2], gc[4], 2.f * gc[5]};
                float dTL[9], dL[9], t[9];
                mat3_mul_d(S2, TL, dTL);
                mat3_mult_d(dTL, L, t);  // dTw += dTL L^T
#pragma unroll
                for (int k = 0; k < 9; k++) dTw[k] += t[k];
                mat3_tmul_d(d.Tw, dTL, dL);  // dL = Tw^T dTL
                float dR[9];
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    dsc[c] = Rq[c] * dL[c] + Rq[3 + c] * dL[3 + c] + Rq[6 + c] * dL[6 + c];
                    dR[c] = dL[c] * sc[c]; dR[3 + c] = dL[3 + c] * sc[c]; dR[6 + c] = dL[6 + c] * sc[c];
                }
                const float r = q.x * inv_norm, x = q.y * inv_norm, y = q.z * inv_norm, z = q.w * inv_norm;
                const float dr_ = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
                const float dx_ = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - r * dR[5] + z * dR[6] + r * dR[7]) -
                                  4.f * x * (dR[4] + dR[8]);
                const float dy_ = 2.f * (x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7]) -
                                  4.f * y * (dR[0] + dR[8]);
                const float dz_ = 2.f * (-r * dR[1] + x * dR[2] + r * dR[3] + y * dR[5] + x * dR[6] + y * dR[7]) -
                                  4.f * z * (dR[0] + dR[4]);
                // q^ = q/|q|: dq = (dq^ - q^ (q^ . dq^)) / |q|
                const float dot = r * dr_ + x * dx_ + y * dy_ + z * dz_;
                dq[0] = (dr_ - r * dot) * inv_norm; dq[1] = (dx_ - x * dot) * inv_norm;
                dq[2] = (dy_ - y * dot) * inv_norm; dq[3] = (dz_ - z * dot) * inv_norm;
#pragma unroll
                for (int c = 0; c < 3; c++) dsc[c] *= p.scale_modifier;
                if (p.scales_are_log) { dsc[0] *= sa[0]; dsc[1] *= sa[1]; dsc[2] *= sa[2]; }  // d exp(l) / dl = exp(l)
            }
            put(g.d_scales + 3 * idx, dsc[0]); put(g.d_scales + 3 * idx + 1, dsc[1]); put(g.d_scales + 3 * idx + 2, dsc[2]);
            if (g.d_rot) {
                float4* o4 = reinterpret_cast<float4*>(g.d_rot) + idx;
                float4 v4 = make_float4(dq[0], dq[1], dq[2], dq[3]);
                if (acc == 2) {
                    atomicAdd(&o4->x, v4.x); atomicAdd(&o4->y, v4.y); atomicAdd(&o4->z, v4.z); atomicAdd(&o4->w, v4.w);
                } else {
                    if (acc) { const float4 old = *o4; v4.x += old.x; v4.y += old.y; v4.z += old.z; v4.w += old.w; }
                    *o4 = v4;
                }
            }
        }

        // ---- world <- smpl <- T pose <- big pose ----
        float dM[9];
        mat3_tmul_d(s.Rg, dTw, dM);  // Tw = Rg M
        if (p.normals && (g.g_normals || g.d_normals)) {
            float dn[3] = {0.f, 0.f, 0.f};
            if (g.g_normals) {
                const float gn[3] = {g.g_normals[3 * idx], g.g_normals[3 * idx + 1], g.g_normals[3 * idx + 2]};
                const float n[3] = {p.normals[3 * idx], p.normals[3 * idx + 1], p.normals[3 * idx + 2]};
                float dn3[3];
                mat3_vec(s.Rginv, gn, dn3);  // nw = Rginv^T n3
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++) dM[3 * r + c] += dn3[r] * n[c];
                mat3_tvec(d.M, dn3, dn);
            }
            if (g.d_normals) { put(g.d_normals + 3 * idx, dn[0]); put(g.d_normals + 3 * idx + 1, dn[1]); put(g.d_normals + 3 * idx + 2, dn[2]); }
        }
        float dx3[3], dx2[3];
        mat3_vec(s.Rginv, gxw, dx3);  // xw = Rginv^T x3 + Th
        mat3_tvec(At3, dx3, dx2);     // x3 = At3 x2 + tt
        float dAt3[9], dRinv[9], t9[9];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 3; c++) dAt3[3 * r + c] = dx3[r] * d.x2[c];
        mat3_mult_d(dM, d.Rinv, t9);  // M = At3 Rinv: dAt3 += dM Rinv^T
#pragma unroll
        for (int k = 0; k < 9; k++) dAt3[k] += t9[k];
        mat3_tmul_d(At3, dM, dRinv);  // dRinv = At3^T dM
        if (g.d_vertex_offsets) {
#pragma unroll
            for (int c = 0; c < 3; c++) atomicAdd(g.d_vertex_offsets + 3 * (size_t)vid + c, dx2[c]);
        }
        float dxc[3];
        mat3_tvec(d.Rinv, dx2, dxc);  // x1 = Rinv xc
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 3; c++) dRinv[3 * r + c] += dx2[r] * d.xc[c];
        put(g.d_xyz + 3 * idx, dxc[0]); put(g.d_xyz + 3 * idx + 1, dxc[1]); put(g.d_xyz + 3 * idx + 2, dxc[2]);
        // Rinv = Ab3^-1: dAb3 = -Rinv^T dRinv Rinv^T
        float dAb3[9];
        mat3_tmul_d(d.Rinv, dRinv, t9);
        mat3_mult_d(t9, d.Rinv, dAb3);
        float dAb[12];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            dAb[4 * r] = -dAb3[3 * r]; dAb[4 * r + 1] = -dAb3[3 * r + 1]; dAb[4 * r + 2] = -dAb3[3 * r + 2];
            dAb[4 * r + 3] = -dxc[r];
            dAt[4 * r] = dAt3[3 * r]; dAt[4 * r + 1] = dAt3[3 * r + 1]; dAt[4 * r + 2] = dAt3[3 * r + 2];
            dAt[4 * r + 3] = dx3[r];
        }
#pragma unroll
        for (int j = 0; j < kJ; j++) bw_keep[j] = d.bw[j];
        if (p.lbs_offsets && g.d_offsets) {
            // bw = softmax(logits): dlogit_j = bw_j (dbw_j - sum_k bw_k dbw_k)
            float dbw[kJ];
            float dot = 0.0f;
#pragma unroll
            for (int j = 0; j < kJ; j++) {
                float v = 0.0f;
#pragma unroll
                for (int k = 0; k < 12; k++) v += dAt[k] * s.A_tgt[12 * j + k] + dAb[k] * s.A_big[12 * j + k];
                dbw[j] = v;
                dot += d.bw[j] * v;
            }
#pragma unroll
            for (int j = 0; j < kJ; j++) put(g.d_offsets + kJ * (size_t)idx + j, d.bw[j] * (dbw[j] - dot));
        }
    }
    // ---- dL/dA_target[j] = sum over Gaussians of bw_j dAt  (pose branch) ----
    if (g.d_A_target) {
        const int lane = threadIdx.x & 31;
        for (int j = 0; j < kJ; j++) {
            const float w = bw_keep[j];
            if (!__any_sync(0xffffffffu, w != 0.0f)) continue;
#pragma unroll
            for (int k = 0; k < 12; k++) {
                float v = w * dAt[k];
#pragma unroll
                for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) atomicAdd(&s_dA[12 * j + k], v);
            }
        }
        __syncthreads();
        for (int k = threadIdx.x; k < kJ * 12; k += 128) {
            const float v = s_dA[k];
            if (v != 0.0f) atomicAdd(g.d_A_target + k, v);
        }
    }
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
