// Stage 1: per-Gaussian projection / EWA covariance / radius / tile rect /
// SH->RGB (reference forward.cu:156-256 preprocessCUDA) fused with the
// prefix sum over tiles_touched (reference rasterizer_impl.cu:279,
// cub::DeviceScan::InclusiveSum) as a single-pass decoupled look-back scan,
// so num_rendered is known when this one kernel retires.  Also
// checkFrustum (rasterizer_impl.cu:54-66).
//
// Roofline: HBM stream.  Algorithmic bytes per Gaussian:
// 48*P + 28*V (precomputed colour + cov3D) or 244*P + 67*V (SH deg3 + scale/rot);
// we additionally write the packed 48-B blend record per visible Gaussian.
#include "ghr_project.cuh"

// Resident blocks per SM the register allocation aims at (tuning macro).  Measured at 3 M Gaussians / 1080p on B200:
// 3 (79 registers) 320-334 us, 4 (64 registers, 160 B spilled) 330 us, 5 (48 registers) 378 us -- occupancy is not what
// holds this kernel at 2.4 TB/s.
#ifndef GHR_PRE_MIN_BLOCKS
#define GHR_PRE_MIN_BLOCKS 3
#endif

namespace ghr {

// Above this many Gaussians the tiles_touched scan is split off the projection kernel (GHR_SPLIT_SCAN_P overrides)
static const int kSplitScanP = [] { const char* e = getenv("GHR_SPLIT_SCAN_P"); return e ? atoi(e) : (1 << 20); }();

__global__ void __launch_bounds__(GHR_BLOCK, GHR_PRE_MIN_BLOCKS) preprocess_kernel(const PreArgs a)
{
    __shared__ float s_cam[36];  // view[16] proj[16] campos[3]
    __shared__ uint32_t s_ticket;
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_excl;

    const int tid = threadIdx.x;
    if (tid < 16) s_cam[tid] = a.view[tid];
    else if (tid < 32) s_cam[tid] = a.proj[tid - 16];
    else if (tid < 35) s_cam[tid] = a.campos[tid - 32];
    if (tid == 0) s_ticket = a.split_scan ? blockIdx.x : atomicAdd(&a.hdr->scan_ticket, 1u);  // no chain, no ticket
    __syncthreads();
    const int tile = (int)s_ticket;
    const int idx = tile * GHR_BLOCK + tid;
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    uint32_t tiles = 0;
    uint32_t dbits_hi = 0u, dbits_lo_inv = 0u;  // depth bit pattern of a visible Gaussian (hi) and its complement (lo)
    if (idx < a.P) {
        int radius;
        bool near_culled;
        float depth = 0.0f;
        tiles = project_gaussian(a, idx, (size_t)idx, V, Pm, s_cam + 32, radius, near_culled, &depth);
        if (near_culled && a.prefiltered) atomicMax(&a.hdr->error, (uint32_t)(-GHR_ERR_PREFILTER));
        a.radii[idx] = radius;
        a.tiles_touched[idx] = tiles;
        if (tiles) {
            dbits_hi = __float_as_uint(depth);  // the depth in the record just written
            dbits_lo_inv = ~dbits_hi;
        }
    }
    finish_preprocess_block(a, tile, idx, tiles, dbits_hi, dbits_lo_inv, s_warp, s_excl);
}

// (Measured and not kept: the split-scan projection in blocks of 128 threads with one barrier -- smaller blocks should
// waste fewer warp slots waiting for their slowest warp -- 265 us vs 254 us at 3 M Gaussians.)
// Large P (split_scan): exclusive scan of the per-block totals (one CTA; 3 M Gaussians = 11 719 totals), num_rendered, then
// the block prefixes are added to the block-local inclusive scans.  The fused single-pass look-back is what a small frame
// wants (one launch); at 3 M Gaussians every block held its SM slot while its warp 0 waited for 32 predecessors.
__global__ void __launch_bounds__(1024) scan_block_totals_kernel(uint64_t* status, int nblocks, ghr_geom_header* hdr, int32_t* n_host)
{
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        const int i = base + t;
        const uint32_t x = i < nblocks ? (uint32_t)status[i] : 0u;
        uint32_t v = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += y;
        }
        if (lane == 31) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const uint32_t incl = s_carry + (warp ? s_warp[warp - 1] : 0u) + v;
        if (i < nblocks) status[i] = incl - x;  // exclusive prefix of block i
        __syncthreads();
        if (t == 1023) s_carry = incl;
        __syncthreads();
    }
    if (t == 0) {
        hdr->num_rendered = s_carry;
        if (n_host) *n_host = (int32_t)s_carry;
    }
}

__global__ void __launch_bounds__(GHR_BLOCK) add_block_prefix_kernel(int P, const uint64_t* __restrict__ status,
                                                                    uint32_t* __restrict__ point_offsets)
{
    const int idx = blockIdx.x * GHR_BLOCK + threadIdx.x;
    if (idx < P && blockIdx.x) point_offsets[idx] += (uint32_t)status[blockIdx.x];
}

__global__ void __launch_bounds__(GHR_BLOCK) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                                 const float* __restrict__ V, uint8_t* present)
{
    const int idx = blockIdx.x * GHR_BLOCK + threadIdx.x;
    if (idx >= P) return;
    const float z = affine_row(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2], V[2], V[6], V[10], V[14]);
    present[idx] = !(z <= 0.2f);
}

cudaError_t launch_preprocess(const ghr_raster_params& p, int32_t* radii, char* geom, const ghr_layout& L,
                              int32_t* n_host, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_PREPROCESS, s);
    PreArgs a;
    a.P = p.P; a.D = p.D; a.M = p.M; a.W = p.W; a.H = p.H;
    a.scale_modifier = p.scale_modifier; a.tan_fovx = p.tan_fovx; a.tan_fovy = p.tan_fovy;
    a.focal_y = p.H / (2.0f * p.tan_fovy);  // rasterizer_impl.cu:224-225
    a.focal_x = p.W / (2.0f * p.tan_fovx);
    a.prefiltered = p.prefiltered;
    a.means3D = p.means3D; a.shs = p.shs; a.colors_precomp = p.colors_precomp; a.opacities = p.opacities;
    a.scales = p.scales; a.rotations = p.rotations; a.cov3D_precomp = p.cov3D_precomp;
    a.view = p.viewmatrix; a.proj = p.projmatrix; a.campos = p.campos;
    a.radii = radii;
    a.rec = reinterpret_cast<float4*>(geom + L.g_rec);
    a.tiles_touched = reinterpret_cast<uint32_t*>(geom + L.g_tiles_touched);
    a.point_offsets = reinterpret_cast<uint32_t*>(geom + L.g_point_offsets);
    a.cov3D_out = reinterpret_cast<float*>(geom + L.g_cov3D);
    a.clamped = reinterpret_cast<uint8_t*>(geom + L.g_clamped);
    a.scan_status = reinterpret_cast<uint64_t*>(geom + L.g_scan_status);
    a.hdr = reinterpret_cast<ghr_geom_header*>(geom + L.g_header);
    a.n_host = n_host;
    a.nblocks = (p.P + GHR_BLOCK - 1) / GHR_BLOCK;
    a.skip_rgb = 0;
    a.split_scan = p.P > kSplitScanP;
    // header + look-back words are contiguous: one memset
    cudaError_t e = cudaMemsetAsync(geom + L.g_header, 0, L.g_rec - L.g_header, s);
    if (e != cudaSuccess) return e;
    preprocess_kernel<<<a.nblocks, GHR_BLOCK, 0, s>>>(a);
    count_launch();
    if (a.split_scan) {
        scan_block_totals_kernel<<<1, 1024, 0, s>>>(a.scan_status, a.nblocks, a.hdr, n_host);
        add_block_prefix_kernel<<<a.nblocks, GHR_BLOCK, 0, s>>>(p.P, a.scan_status, a.point_offsets);
        count_launch(2);
    }
    return cudaGetLastError();
}

cudaError_t launch_mark_visible(int P, const float* means3D, const float* V, uint8_t* present, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_MISC, s);
    count_launch();
    mark_visible_kernel<<<(P + GHR_BLOCK - 1) / GHR_BLOCK, GHR_BLOCK, 0, s>>>(P, means3D, V, present);
    return cudaGetLastError();
}

}  // namespace ghr
