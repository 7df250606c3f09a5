// The per-Gaussian chain of a view as ONE kernel each way ("fused LBS + projection + SH"):
//
//   forward   LBS deformation + world covariance T Sigma T^T + sigmoid / exp activations  (ghr_lbs_deform.cu)
//             -> projection / EWA covariance / radius / tile rect, SH -> RGB for the survivors, tiles_touched scan
//                (ghr_preprocess.cu)
//   backward  projection backward (ghr_preprocess_bwd.cu) -> SH / covariance / skinning backward (ghr_lbs_deform.cu)
//
// Reference: gaussian_renderer/__init__.py:94-195 (coarse_deform_c2source, get_covariance, eval_sh, rasteriser call),
// forward.cu:156-256 preprocessCUDA, backward.cu:144-396, scene/gaussian_model.py:768-872.
//
// The deformed mean, the world covariance, the colour and the opacity of a Gaussian never exist in HBM: the forward
// hands them to the projection in registers, the backward recomputes them from the canonical parameters (which it has to
// read anyway for the skinning backward).  Against the two-kernel chain that removes, per Gaussian and view,
// 12 + 24 + 12 + 4 B written and read back forward, the same read again backward, and 12 + 24 + 12 + 4 B of
// intermediate gradients written and read back: 260 B of ~1.1 KB.  The device functions are the ones the stand-alone
// kernels use (ghr_deform.cuh, ghr_project.cuh, ghr_prebwd.cuh): the projection's pinned fp32 sequence is untouched, so
// radii / tile rects / sort keys are those of the two-kernel chain bit for bit (tests/test_fused_gpu.py).
#include "ghr_deform.cuh"
#include "ghr_prebwd.cuh"
#include "ghr_project.cuh"

namespace ghr {

struct FusedFwdArgs {
    ghr_deform_params dp;
    PreArgs a;
    // optional copies of the intermediate values for callers that want them (tests, normal maps): NULL = not written
    float *world_pts, *world_normals, *cov3D, *colors, *opacity;
};

__global__ void __launch_bounds__(GHR_BLOCK) deform_project_kernel(const FusedFwdArgs f)
{
    __shared__ DeformShared s;
    __shared__ float s_cam[32];  // view[16] proj[16]
    __shared__ uint32_t s_ticket;
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_excl;
    const ghr_deform_params& p = f.dp;
    const PreArgs& a = f.a;

    const int tid = threadIdx.x;
    if (tid < 16) s_cam[tid] = a.view[tid];
    else if (tid < 32) s_cam[tid] = a.proj[tid - 16];
    if (tid == 0) s_ticket = atomicAdd(&a.hdr->scan_ticket, 1u);
    load_shared(p, s);  // ends with __syncthreads()
    const int tile = (int)s_ticket;
    const int idx = tile * GHR_BLOCK + tid;
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    uint32_t tiles = 0;
    uint32_t dbits_hi = 0u, dbits_lo_inv = 0u;
    if (idx < p.P) {
        // the SH row is needed ~600 instructions from here and nearly every Gaussian of a body is visible: pull it into
        // L2 now (no register is tied up)
        if (p.shs) {
            const char* sp = reinterpret_cast<const char*>(p.shs + (size_t)idx * p.M * 3);
            const int nb = p.M * 12;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sp));
            if (nb > 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(sp + 128));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sp + nb - 4));
        }
        const int vid = p.vert_ids[idx];
        const float opac = p.opacity_logits ? 1.0f / (1.0f + expf(-p.opacity_logits[idx])) : __ldg(a.opacities + idx);
        DeformState d;
        float xw[3];
        deform_forward_core(p, s, idx, vid, d, xw);
        if (f.world_pts) {
#pragma unroll
            for (int c = 0; c < 3; c++) f.world_pts[3 * (size_t)idx + c] = xw[c];
        }
        if (f.opacity) f.opacity[idx] = opac;
        if (f.world_normals && p.normals) {
            const float n[3] = {p.normals[3 * idx], p.normals[3 * idx + 1], p.normals[3 * idx + 2]};
            float n3[3], nw[3];
            mat3_vec(d.M, n, n3);
            mat3_tvec(s.Rginv, n3, nw);
#pragma unroll
            for (int c = 0; c < 3; c++) f.world_normals[3 * (size_t)idx + c] = nw[c];
        }
        auto hint = [](float, float) {};
        auto cov3 = [&](float* c3) {
            DeformCov k;
            deform_cov_forward(p, idx, d.Tw, k, c3);
            if (f.cov3D) {
#pragma unroll
                for (int i = 0; i < 6; i++) f.cov3D[6 * (size_t)idx + i] = c3[i];
            }
        };
        bool coloured = false;
        auto rgbf = [&](float* rgb) {
            float sh[48];
            load_sh_row(p.shs, idx, p.M, sh);
            uint32_t cm;
            sh_to_rgb(p.sh_degree, sh, xw[0] - s.campos[0], xw[1] - s.campos[1], xw[2] - s.campos[2], rgb, cm);
            a.clamped[idx] = (uint8_t)cm;
            if (f.colors) {
#pragma unroll
                for (int c = 0; c < 3; c++) f.colors[3 * (size_t)idx + c] = rgb[c];
            }
            coloured = true;
        };
        int radius;
        bool near_culled;
        float depth = 0.0f;
        tiles = project_core(a, (size_t)idx, xw[0], xw[1], xw[2], opac, V, Pm, radius, near_culled, hint, cov3, rgbf, &depth);
        if (near_culled && a.prefiltered) atomicMax(&a.hdr->error, (uint32_t)(-GHR_ERR_PREFILTER));
        a.radii[idx] = radius;
        a.tiles_touched[idx] = tiles;
        if (tiles) {
            dbits_hi = __float_as_uint(depth);
            dbits_lo_inv = ~dbits_hi;
        }
        if (!coloured && f.colors) {
#pragma unroll
            for (int c = 0; c < 3; c++) f.colors[3 * (size_t)idx + c] = 0.0f;
        }
        if (near_culled && f.cov3D) {
#pragma unroll
            for (int i = 0; i < 6; i++) f.cov3D[6 * (size_t)idx + i] = 0.0f;
        }
    }
    finish_preprocess_block(a, tile, idx, tiles, dbits_hi, dbits_lo_inv, s_warp, s_excl);
}

struct FusedBwdArgs {
    ghr_deform_params dp;
    DeformGrads g;
    const float *view, *proj;
    float h_x, h_y, tan_fovx, tan_fovy;
    const int32_t* radii;
    const float* accum;      // [P][12] from the blend backward
    const uint8_t* clamped;  // geom
    float* dL_dmean2D;       // [P,3]
    // densification statistics of the training loop (train.py:401-404, gaussian_model.py:764-766), fused: NULL = skipped
    float *max_radii2D, *xyz_gradient_accum, *denom;
};

__global__ void __launch_bounds__(128) project_deform_backward_kernel(const FusedBwdArgs f)
{
    __shared__ DeformShared s;
    __shared__ float s_dA[kJ * 12];
    __shared__ float s_cam[32];
    const ghr_deform_params& p = f.dp;
    const DeformGrads& g = f.g;
    if (threadIdx.x < 16) s_cam[threadIdx.x] = f.view[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = f.proj[threadIdx.x - 16];
    if (g.d_A_target)
        for (int k = threadIdx.x; k < kJ * 12; k += 128) s_dA[k] = 0.0f;
    load_shared(p, s);
    const int idx = blockIdx.x * 128 + threadIdx.x;
    float dAt[12];
    float bw_keep[kJ];
#pragma unroll
    for (int k = 0; k < 12; k++) dAt[k] = 0.0f;
#pragma unroll
    for (int j = 0; j < kJ; j++) bw_keep[j] = 0.0f;
    if (idx < p.P) {
        const bool vis = f.radii[idx] > 0;
        float g2x = 0.0f, g2y = 0.0f;
        // a culled Gaussian has no gradient: nothing to add in the accumulating modes, zeros to store otherwise
        if (vis || p.accumulate_grads == 0) {
            const int vid = p.vert_ids[idx];
            DeformState d;
            float xw[3];
            deform_forward_core(p, s, idx, vid, d, xw);
            DeformUpstream up;
#pragma unroll
            for (int c = 0; c < 3; c++) { up.gxw[c] = 0.0f; up.gcol[c] = 0.0f; up.gn[c] = 0.0f; }
#pragma unroll
            for (int k = 0; k < 9; k++) up.dTw[k] = 0.0f;
#pragma unroll
            for (int k = 0; k < 6; k++) up.gcov[k] = 0.0f;
            up.has_cov = false; up.has_col = false; up.has_n = false; up.gop = 0.0f; up.clamp_mask = 0u;
            if (vis) {
                DeformCov k;
                float c3[6];
                deform_cov_forward(p, idx, d.Tw, k, c3);
                PreBwdOut pb;
                preprocess_backward_core(s_cam, s_cam + 16, f.h_x, f.h_y, f.tan_fovx, f.tan_fovy, xw[0], xw[1], xw[2], c3,
                                         f.accum + 12 * (size_t)idx, pb);
#pragma unroll
                for (int c = 0; c < 3; c++) { up.gxw[c] = pb.dm[c]; up.gcol[c] = pb.dcol[c]; }
#pragma unroll
                for (int i = 0; i < 6; i++) up.gcov[i] = pb.dcov[i];
                up.has_cov = true; up.has_col = true; up.gop = pb.dop;
                up.clamp_mask = f.clamped[idx];
                g2x = pb.g2x; g2y = pb.g2y;
            }
            deform_backward_gaussian(p, s, g, idx, vid, d, xw, up, dAt, bw_keep);
        }
        if (vis) {  // visibility_filter = radii > 0; atomics: the views of a batch may run concurrently
            if (f.max_radii2D)  // max of non-negative floats = max of their bit patterns
                atomicMax(reinterpret_cast<int*>(f.max_radii2D + idx), __float_as_int((float)f.radii[idx]));
            if (f.xyz_gradient_accum) atomicAdd(f.xyz_gradient_accum + idx, sqrtf(g2x * g2x + g2y * g2y));
            if (f.denom) atomicAdd(f.denom + idx, 1.0f);
        }
        if (f.dL_dmean2D) {
            f.dL_dmean2D[3 * (size_t)idx] = g2x; f.dL_dmean2D[3 * (size_t)idx + 1] = g2y; f.dL_dmean2D[3 * (size_t)idx + 2] = 0.0f;
        }
    }
    if (g.d_A_target) reduce_pose_grads(s_dA, g.d_A_target, bw_keep, dAt);
}

cudaError_t launch_deform_project(const ghr_deform_params& dp, const ghr_raster_params& p, int32_t* radii, char* geom,
                                  const ghr_layout& L, float* world_pts, float* world_normals, float* cov3D, float* colors,
                                  float* opacity, int32_t* n_host, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_DEFORM_PROJECT, s);
    FusedFwdArgs f;
    f.dp = dp;
    f.dp.campos = p.campos;
    PreArgs& a = f.a;
    a.P = dp.P; a.D = dp.sh_degree; a.M = dp.M; a.W = p.W; a.H = p.H;
    a.scale_modifier = dp.scale_modifier; a.tan_fovx = p.tan_fovx; a.tan_fovy = p.tan_fovy;
    a.focal_y = p.H / (2.0f * p.tan_fovy);
    a.focal_x = p.W / (2.0f * p.tan_fovx);
    a.prefiltered = p.prefiltered;
    a.means3D = nullptr; a.shs = dp.shs; a.colors_precomp = nullptr; a.opacities = p.opacities;
    a.scales = nullptr; a.rotations = nullptr; a.cov3D_precomp = nullptr;
    a.view = p.viewmatrix; a.proj = p.projmatrix; a.campos = p.campos;
    a.radii = radii;
    a.rec = reinterpret_cast<float4*>(geom + L.g_rec);
    a.tiles_touched = reinterpret_cast<uint32_t*>(geom + L.g_tiles_touched);
    a.point_offsets = reinterpret_cast<uint32_t*>(geom + L.g_point_offsets);
    a.cov3D_out = nullptr;
    a.clamped = reinterpret_cast<uint8_t*>(geom + L.g_clamped);
    a.scan_status = reinterpret_cast<uint64_t*>(geom + L.g_scan_status);
    a.hdr = reinterpret_cast<ghr_geom_header*>(geom + L.g_header);
    a.n_host = n_host;
    a.nblocks = (dp.P + GHR_BLOCK - 1) / GHR_BLOCK;
    a.skip_rgb = 0;
    a.split_scan = 0;
    f.world_pts = world_pts; f.world_normals = world_normals; f.cov3D = cov3D; f.colors = colors; f.opacity = opacity;
    cudaError_t e = cudaMemsetAsync(geom + L.g_header, 0, L.g_rec - L.g_header, s);
    if (e != cudaSuccess) return e;
    deform_project_kernel<<<a.nblocks, GHR_BLOCK, 0, s>>>(f);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_project_deform_backward(const ghr_deform_params& dp, const ghr_raster_params& p, const int32_t* radii,
                                           const char* geom, const ghr_layout& L, const float* accum,
                                           const DeformGrads& g, float* dL_dmean2D, float* max_radii2D,
                                           float* xyz_gradient_accum, float* denom, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_PROJECT_DEFORM_BWD, s);
    count_launch();
    FusedBwdArgs f;
    f.dp = dp;
    f.dp.campos = p.campos;
    f.g = g;
    f.view = p.viewmatrix; f.proj = p.projmatrix;
    f.h_y = p.H / (2.0f * p.tan_fovy);
    f.h_x = p.W / (2.0f * p.tan_fovx);
    f.tan_fovx = p.tan_fovx; f.tan_fovy = p.tan_fovy;
    f.radii = radii;
    f.accum = accum;
    f.clamped = reinterpret_cast<const uint8_t*>(geom + L.g_clamped);
    f.dL_dmean2D = dL_dmean2D;
    f.max_radii2D = max_radii2D; f.xyz_gradient_accum = xyz_gradient_accum; f.denom = denom;
    project_deform_backward_kernel<<<(dp.P + 127) / 128, 128, 0, s>>>(f);
    return cudaGetLastError();
}

}  // namespace ghr
