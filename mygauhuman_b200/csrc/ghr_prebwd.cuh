// Per-Gaussian backward of the projection (reference backward.cu:144-274 computeCov2DCUDA, :346-396 preprocessCUDA up to
// the mean), shared by preprocess_backward_kernel (ghr_preprocess_bwd.cu) and the fused projection + LBS backward
// (ghr_fused.cu).  Input: the packed accumulator row the blend backward wrote for this Gaussian.
#pragma once
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

struct PreBwdOut {
    float dm[3];    // dL/d mean3D (without the SH view-direction term)
    float dcov[6];  // dL/d packed cov3D
    float dcol[3];  // dL/d colour
    float g2x, g2y; // dL/d mean2D (screen space, what densification reads)
    float dop;      // dL/d opacity
};

__device__ __forceinline__ void preprocess_backward_core(const float* V, const float* Pm, float h_x, float h_y,
                                                         float tan_fovx, float tan_fovy, float mx, float my, float mz,
                                                         const float* c3, const float* accum_row, PreBwdOut& o)
{
    float* dm = o.dm;
    float* dcov = o.dcov;
    float* dcol = o.dcol;
#pragma unroll
    for (int k = 0; k < 6; k++) dcov[k] = 0.0f;
    const float4* ap = reinterpret_cast<const float4*>(accum_row);
    const float4 a0 = ap[0], a1 = ap[1], a2 = ap[2];
    dcol[0] = a0.x; dcol[1] = a0.y; dcol[2] = a0.z; o.g2x = a0.w; o.g2y = a1.x;
    const float g2x = o.g2x, g2y = o.g2y;
    const float dcx = a1.y, dcy = a1.z, dcz = a1.w;
    o.dop = a2.x;
    Cov2DCtx c;
    cov2d_T(mx, my, mz, h_x, h_y, tan_fovx, tan_fovy, V, c);
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
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * c.tx) * tz3 * dL_dJ02 +
                         (2 * h_y * c.ty) * tz3 * dL_dJ12;
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
}

}  // namespace ghr
