// Device-side building blocks of the per-Gaussian LBS deformation, shared by the stand-alone LBS kernels
// (ghr_lbs_deform.cu) and the fused LBS + projection kernels (ghr_fused.cu).
// Reference: scene/gaussian_model.py:768-872 coarse_deform_c2source, :35-42 build_covariance_from_scaling_rotation.
#pragma once
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

// Sigma' = (Tw L)(Tw L)^T with L = R(q/|q|) diag(mod * s)  (gaussian_model.py:35-42), packed [xx,xy,xz,yy,yz,zz].
struct DeformCov {
    float4 q;
    float Rq[9], inv_norm;
    float sa[3], sc[3];  // activated scales, scales * modifier
    float L[9], TL[9];
};
__device__ __forceinline__ void deform_cov_forward(const ghr_deform_params& p, int idx, const float* Tw, DeformCov& k,
                                                   float* c3)
{
    k.q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
    quat_rot(k.q.x, k.q.y, k.q.z, k.q.w, k.Rq, k.inv_norm);
    k.sa[0] = p.scales[3 * idx]; k.sa[1] = p.scales[3 * idx + 1]; k.sa[2] = p.scales[3 * idx + 2];
    if (p.scales_are_log) { k.sa[0] = expf(k.sa[0]); k.sa[1] = expf(k.sa[1]); k.sa[2] = expf(k.sa[2]); }
#pragma unroll
    for (int c = 0; c < 3; c++) k.sc[c] = p.scale_modifier * k.sa[c];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) k.L[3 * r + c] = k.Rq[3 * r + c] * k.sc[c];
    mat3_mul_d(Tw, k.L, k.TL);
    const float* TL = k.TL;
    c3[0] = TL[0] * TL[0] + TL[1] * TL[1] + TL[2] * TL[2];
    c3[1] = TL[0] * TL[3] + TL[1] * TL[4] + TL[2] * TL[5];
    c3[2] = TL[0] * TL[6] + TL[1] * TL[7] + TL[2] * TL[8];
    c3[3] = TL[3] * TL[3] + TL[4] * TL[4] + TL[5] * TL[5];
    c3[4] = TL[3] * TL[6] + TL[4] * TL[7] + TL[5] * TL[8];
    c3[5] = TL[6] * TL[6] + TL[7] * TL[7] + TL[8] * TL[8];
}

// loads the SH row of Gaussian idx (M coefficients x 3 channels, up to 48 floats)
__device__ __forceinline__ void load_sh_row(const float* shs, int idx, int M, float* sh)
{
    const int nf = min(M * 3, 48);
    const float* sp = shs + (size_t)idx * M * 3;
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
}

// Gradients arriving at the outputs of the deformation of ONE Gaussian (zero where nothing arrives).
struct DeformUpstream {
    float gxw[3];   // d/d world point
    float dTw[9];   // d/d transforms
    float gcov[6];  // d/d packed world covariance
    float gcol[3];  // d/d colour
    float gn[3];    // d/d world normal
    float gop;      // d/d activated opacity
    bool has_cov, has_col, has_n;
    uint32_t clamp_mask;
};

// Backward of the deformation of Gaussian idx, given its recomputed forward state (d, xw).  Writes / adds (mode
// p.accumulate_grads) every per-Gaussian parameter gradient of `g` and returns the pose-branch terms: dAt[12] (gradient
// of the blended target transform) and bw_keep[24] (the skinning weights), which the caller reduces over the block.
__device__ __forceinline__ void deform_backward_gaussian(const ghr_deform_params& p, const DeformShared& s,
                                                         const DeformGrads& g, int idx, int vid, const DeformState& d,
                                                         const float* xw, DeformUpstream& up, float* dAt, float* bw_keep)
{
    const int acc = p.accumulate_grads;
    auto put = [acc](float* dst, float v) {
        if (acc == 2) atomicAdd(dst, v);
        else *dst = acc ? *dst + v : v;
    };
    const float At3[9] = {d.At[0], d.At[1], d.At[2], d.At[4], d.At[5], d.At[6], d.At[8], d.At[9], d.At[10]};
    float* gxw = up.gxw;
    float* dTw = up.dTw;

    // ---- colour: clamp_min(eval_sh + 0.5, 0) with dirs from the deformed mean ----
    if (p.shs && g.d_shs) {
        float* dsh = g.d_shs + (size_t)idx * p.M * 3;
        if (up.has_col) {
            sh_backward(p.sh_degree, p.M, p.shs + (size_t)idx * p.M * 3, xw[0] - s.campos[0], xw[1] - s.campos[1],
                        xw[2] - s.campos[2], up.clamp_mask, up.gcol, gxw, dsh, acc);
        } else if (!acc) {
            for (int k = 0; k < p.M * 3; k++) dsh[k] = 0.0f;
        }
    }

    if (p.opacity_logits && p.dL_dopacity_logits) {
        const float sg = 1.0f / (1.0f + expf(-p.opacity_logits[idx]));
        put(p.dL_dopacity_logits + idx, up.gop * sg * (1.0f - sg));
    }
    // ---- covariance: Sigma' = (Tw L)(Tw L)^T, L = R(q^) diag(mod*s) ----
    if (p.scales && g.d_scales) {
        float dsc[3] = {0.f, 0.f, 0.f}, dq[4] = {0.f, 0.f, 0.f, 0.f};
        if (up.has_cov) {
            DeformCov k;
            float c3[6];
            deform_cov_forward(p, idx, d.Tw, k, c3);
            const float* gc = up.gcov;
            // dS' + dS'^T for the upper-triangular packing [xx,xy,xz,yy,yz,zz] (strip_symmetric)
            const float S2[9] = {2.f * gc[0], gc[1], gc[2], gc[1], 2.f * gc[3], gc[4], gc[2], gc[4], 2.f * gc[5]};
            float dTL[9], dL[9], t[9];
            mat3_mul_d(S2, k.TL, dTL);
            mat3_mult_d(dTL, k.L, t);  // dTw += dTL L^T
#pragma unroll
            for (int i = 0; i < 9; i++) dTw[i] += t[i];
            mat3_tmul_d(d.Tw, dTL, dL);  // dL = Tw^T dTL
            float dR[9];
#pragma unroll
            for (int c = 0; c < 3; c++) {
                dsc[c] = k.Rq[c] * dL[c] + k.Rq[3 + c] * dL[3 + c] + k.Rq[6 + c] * dL[6 + c];
                dR[c] = dL[c] * k.sc[c]; dR[3 + c] = dL[3 + c] * k.sc[c]; dR[6 + c] = dL[6 + c] * k.sc[c];
            }
            const float r = k.q.x * k.inv_norm, x = k.q.y * k.inv_norm, y = k.q.z * k.inv_norm, z = k.q.w * k.inv_norm;
            const float dr_ = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
            const float dx_ = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - r * dR[5] + z * dR[6] + r * dR[7]) -
                              4.f * x * (dR[4] + dR[8]);
            const float dy_ = 2.f * (x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7]) -
                              4.f * y * (dR[0] + dR[8]);
            const float dz_ = 2.f * (-r * dR[1] + x * dR[2] + r * dR[3] + y * dR[5] + x * dR[6] + y * dR[7]) -
                              4.f * z * (dR[0] + dR[4]);
            // q^ = q/|q|: dq = (dq^ - q^ (q^ . dq^)) / |q|
            const float dot = r * dr_ + x * dx_ + y * dy_ + z * dz_;
            dq[0] = (dr_ - r * dot) * k.inv_norm; dq[1] = (dx_ - x * dot) * k.inv_norm;
            dq[2] = (dy_ - y * dot) * k.inv_norm; dq[3] = (dz_ - z * dot) * k.inv_norm;
#pragma unroll
            for (int c = 0; c < 3; c++) dsc[c] *= p.scale_modifier;
            if (p.scales_are_log) { dsc[0] *= k.sa[0]; dsc[1] *= k.sa[1]; dsc[2] *= k.sa[2]; }  // d exp(l) / dl = exp(l)
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
    if (p.normals && (up.has_n || g.d_normals)) {
        float dn[3] = {0.f, 0.f, 0.f};
        if (up.has_n) {
            const float n[3] = {p.normals[3 * idx], p.normals[3 * idx + 1], p.normals[3 * idx + 2]};
            float dn3[3];
            mat3_vec(s.Rginv, up.gn, dn3);  // nw = Rginv^T n3
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

// dL/dA_target[j] = sum over the block's Gaussians of bw_j dAt (pose branch): warp shuffles -> shared -> one atomic per
// entry and block.  Called by every thread of the block (s_dA zeroed beforehand).
__device__ __forceinline__ void reduce_pose_grads(float* s_dA, float* d_A_target, const float* bw_keep, const float* dAt)
{
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
    for (int k = threadIdx.x; k < kJ * 12; k += blockDim.x) {
        const float v = s_dA[k];
        if (v != 0.0f) atomicAdd(d_A_target + k, v);
    }
}

}  // namespace ghr
