// Per-Gaussian backward: reference backward.cu:144-274 computeCov2DCUDA and
// :346-396 preprocessCUDA (with :20-139 SH backward and :278-341 cov3D
// backward) fused into one kernel.  It consumes the packed accumulator written
// by the blend backward and writes every output gradient for all P Gaussians
// (zeros for culled ones), so callers never memset 75 floats per Gaussian as
// rasterize_points.cu:159-167 does.
// Roofline: HBM stream, ~132*P bytes (precomputed path) / ~580*P (SH + scale/rot).
#include "ghr_prebwd.cuh"

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
        const float mx = a.means3D[3 * idx], my = a.means3D[3 * idx + 1], mz = a.means3D[3 * idx + 2];
        float c3[6];
        {
            const float2* cp = reinterpret_cast<const float2*>(a.cov3D + 6 * (size_t)idx);
            const float2 c01 = cp[0], c23 = cp[1], c45 = cp[2];
            c3[0] = c01.x; c3[1] = c01.y; c3[2] = c23.x; c3[3] = c23.y; c3[4] = c45.x; c3[5] = c45.y;
        }
        PreBwdOut pb;
        preprocess_backward_core(V, Pm, a.h_x, a.h_y, a.tan_fovx, a.tan_fovy, mx, my, mz, c3, a.accum + 12 * (size_t)idx, pb);
#pragma unroll
        for (int k = 0; k < 3; k++) { dm[k] = pb.dm[k]; dcol[k] = pb.dcol[k]; }
#pragma unroll
        for (int k = 0; k < 6; k++) dcov[k] = pb.dcov[k];
        g2x = pb.g2x; g2y = pb.g2y; dop = pb.dop;
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
