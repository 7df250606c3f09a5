// Per-Gaussian backward: reference backward.cu:144-274 computeCov2DCUDA and
// :346-396 preprocessCUDA (with :20-139 SH backward and :278-341 cov3D
// backward) fused into one kernel.  It consumes the packed accumulator written
// by the blend backward and writes every output gradient for all P Gaussians
// (zeros for culled ones), so callers never memset 75 floats per Gaussian as
// rasterize_points.cu:159-167 does.
// Roofline: HBM stream, ~132*P bytes (precomputed path) / ~580*P (SH + scale/rot).
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

struct PreBwdArgs {
    int P, D, M, W, H;
    float scale_modifier, tan_fovx, tan_fovy, h_x, h_y;
    const float *means3D, *shs, *scales, *rotations, *cov3D;  // cov3D: precomputed or geom copy
    const float *view, *proj, *campos;
    const int32_t* radii;
    const uint8_t* clamped;
    const float* accum;  // [P][12]
    float *dL_dmean2D, *dL_dcolor, *dL_dopacity, *dL_dmean3D, *dL_dcov3D, *dL_dsh, *dL_dscale, *dL_drot;
};

// backward.cu:278-341
__device__ __forceinline__ void cov3d_backward(float s0, float s1, float s2, float mod, float r, float x, float y,
                                               float z, const float* d, float* dL_dscale, float* dL_drot)
{
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float s[3] = {mod * s0, mod * s1, mod * s2};
    float Mm[3][3];
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) Mm[c][k] = s[k] * R[c][k];
    const float dS[3][3] = {{d[0], 0.5f * d[1], 0.5f * d[2]}, {0.5f * d[1], d[3], 0.5f * d[4]}, {0.5f * d[2], 0.5f * d[4], d[5]}};
    float dMt[3][3];  // dMt[c][k] = dL_dM[k][c]
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++)
            dMt[i][j] = 2.0f * (Mm[0][i] * dS[j][0] + Mm[1][i] * dS[j][1] + Mm[2][i] * dS[j][2]);
#pragma unroll
    for (int c = 0; c < 3; c++) dL_dscale[c] = R[0][c] * dMt[c][0] + R[1][c] * dMt[c][1] + R[2][c] * dMt[c][2];
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) dMt[c][k] *= s[c];
    dL_drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    dL_drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
    dL_drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
    dL_drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

__global__ void __launch_bounds__(GHR_BLOCK) preprocess_backward_kernel(const PreBwdArgs a)
{
    __shared__ float s_cam[36];
    const int tid = threadIdx.x;
    if (tid < 16) s_cam[tid] = a.view[tid];
    else if (tid < 32) s_cam[tid] = a.proj[tid - 16];
    else if (tid < 35) s_cam[tid] = a.campos[tid - 32];
    __syncthreads();
    const int idx = blockIdx.x * GHR_BLOCK + tid;
    if (idx >= a.P) return;
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    float dm[3] = {0, 0, 0}, dcov[6] = {0, 0, 0, 0, 0, 0}, dcol[3] = {0, 0, 0}, dsc[3] = {0, 0, 0}, drt[4] = {0, 0, 0, 0};
    float g2x = 0.f, g2y = 0.f, dop = 0.f;
    const bool vis = a.radii[idx] > 0;
    if (vis) {
        const float4* ap = reinterpret_cast<const float4*>(a.accum + 12 * (size_t)idx);
        const float4 a0 = ap[0], a1 = ap[1], a2 = ap[2];
        dcol[0] = a0.x; dcol[1] = a0.y; dcol[2] = a0.z; g2x = a0.w; g2y = a1.x;
        const float dcx = a1.y, dcy = a1.z, dcz = a1.w;
        dop = a2.x;
        const float mx = a.means3D[3 * idx], my = a.means3D[3 * idx + 1], mz = a.means3D[3 * idx + 2];
        float c3[6];
        {
            const float2* cp = reinterpret_cast<const float2*>(a.cov3D + 6 * (size_t)idx);
            const float2 c01 = cp[0], c23 = cp[1], c45 = cp[2];
            c3[0] = c01.x; c3[1] = c01.y; c3[2] = c23.x; c3[3] = c23.y; c3[4] = c45.x; c3[5] = c45.y;
        }
        Cov2DCtx c;
        cov2d_T(mx, my, mz, a.h_x, a.h_y, a.tan_fovx, a.tan_fovy, V, c);
        float ca, cb, cc;
        cov2d(c, c3, ca, cb, cc);
        const float x_grad_mul = (c.txtz < -c.limx || c.txtz > c.limx) ? 0.f : 1.f;
        const float y_grad_mul = (c.tytz < -c.limy || c.tytz > c.limy) ? 0.f : 1.f;
        const float denom = ca * cc - cb * cb;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        const float T00 = c.T00, T01 = c.T01, T02 = c.T02, T10 = c.T10, T11 = c.T11, T12 = c.T12;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
            dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
            dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
            dcov[0] = (T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc);
            dcov[3] = (T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc);
            dcov[5] = (T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc);
            dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
            dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
            dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
        }
        const float V00 = c3[0], V01 = c3[1], V02 = c3[2], V11 = c3[3], V12 = c3[4], V22 = c3[5];
        const float dL_dT00 = 2 * (T00 * V00 + T01 * V01 + T02 * V02) * dL_da + (T10 * V00 + T11 * V01 + T12 * V02) * dL_db;
        const float dL_dT01 = 2 * (T00 * V01 + T01 * V11 + T02 * V12) * dL_da + (T10 * V01 + T11 * V11 + T12 * V12) * dL_db;
        const float dL_dT02 = 2 * (T00 * V02 + T01 * V12 + T02 * V22) * dL_da + (T10 * V02 + T11 * V12 + T12 * V22) * dL_db;
        const float dL_dT10 = 2 * (T10 * V00 + T11 * V01 + T12 * V02) * dL_dc + (T00 * V00 + T01 * V01 + T02 * V02) * dL_db;
        const float dL_dT11 = 2 * (T10 * V01 + T11 * V11 + T12 * V12) * dL_dc + (T00 * V01 + T01 * V11 + T02 * V12) * dL_db;
        const float dL_dT12 = 2 * (T10 * V02 + T11 * V12 + T12 * V22) * dL_dc + (T00 * V02 + T01 * V12 + T02 * V22) * dL_db;
        const float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
        const float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
        const float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
        const float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
        const float tz = 1.f / c.tz, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -a.h_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -a.h_y * tz2 * dL_dJ12;
        const float dL_dtz = -a.h_x * tz2 * dL_dJ00 - a.h_y * tz2 * dL_dJ11 + (2 * a.h_x * c.tx) * tz3 * dL_dJ02 +
                             (2 * a.h_y * c.ty) * tz3 * dL_dJ12;
        dm[0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
        dm[1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
        dm[2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;
        // backward.cu:363-387
        const float m_hw = Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15];
        const float m_w = 1.0f / (m_hw + 0.0000001f);
        const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
        const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
        dm[0] += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
        dm[1] += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
        dm[2] += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
        if (a.shs)
            sh_backward(a.D, a.M, a.shs + (size_t)idx * a.M * 3, mx - s_cam[32], my - s_cam[33], mz - s_cam[34],
                        a.clamped[idx], dcol, dm, a.dL_dsh + (size_t)idx * a.M * 3);
        if (a.scales) {
            const float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
            cov3d_backward(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.scale_modifier, q.x, q.y,
                           q.z, q.w, dcov, dsc, drt);
        }
    } else if (a.shs) {
        float* o = a.dL_dsh + (size_t)idx * a.M * 3;
        for (int k = 0; k < a.M * 3; k++) o[k] = 0.0f;
    }
    a.dL_dmean2D[3 * idx] = g2x; a.dL_dmean2D[3 * idx + 1] = g2y; a.dL_dmean2D[3 * idx + 2] = 0.0f;
    a.dL_dcolor[3 * idx] = dcol[0]; a.dL_dcolor[3 * idx + 1] = dcol[1]; a.dL_dcolor[3 * idx + 2] = dcol[2];
    a.dL_dopacity[idx] = dop;
    a.dL_dmean3D[3 * idx] = dm[0]; a.dL_dmean3D[3 * idx + 1] = dm[1]; a.dL_dmean3D[3 * idx + 2] = dm[2];
#pragma unroll
    for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    if (a.dL_dscale) { a.dL_dscale[3 * idx] = dsc[0]; a.dL_dscale[3 * idx + 1] = dsc[1]; a.dL_dscale[3 * idx + 2] = dsc[2]; }
    if (a.dL_drot) reinterpret_cast<float4*>(a.dL_drot)[idx] = make_float4(drt[0], drt[1], drt[2], drt[3]);
}

cudaError_t launch_preprocess_backward(const ghr_raster_params& p, const int32_t* radii, const char* geom,
                                       const ghr_layout& L, const float* accum, float* dL_dmean2D, float* dL_dcolor,
                                       float* dL_dopacity, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                                       float* dL_dscale, float* dL_drot, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_PREPROCESS_BWD, s);
    count_launch();
    PreBwdArgs a;
    a.P = p.P; a.D = p.D; a.M = p.M; a.W = p.W; a.H = p.H;
    a.scale_modifier = p.scale_modifier; a.tan_fovx = p.tan_fovx; a.tan_fovy = p.tan_fovy;
    a.h_y = p.H / (2.0f * p.tan_fovy);
    a.h_x = p.W / (2.0f * p.tan_fovx);
    a.means3D = p.means3D; a.shs = p.shs; a.scales = p.scales; a.rotations = p.rotations;
    a.cov3D = p.cov3D_precomp ? p.cov3D_precomp : reinterpret_cast<const float*>(geom + L.g_cov3D);
    a.view = p.viewmatrix; a.proj = p.projmatrix; a.campos = p.campos;
    a.radii = radii;
    a.clamped = reinterpret_cast<const uint8_t*>(geom + L.g_clamped);
    a.accum = accum;
    a.dL_dmean2D = dL_dmean2D; a.dL_dcolor = dL_dcolor; a.dL_dopacity = dL_dopacity; a.dL_dmean3D = dL_dmean3D;
    a.dL_dcov3D = dL_dcov3D; a.dL_dsh = dL_dsh; a.dL_dscale = dL_dscale; a.dL_drot = dL_drot;
    preprocess_backward_kernel<<<(p.P + GHR_BLOCK - 1) / GHR_BLOCK, GHR_BLOCK, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace ghr
