/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * Thin C-ABI wrapper (ours) around the UNMODIFIED reference CUDA sources,
 * which are compiled from where they lie under /root/reference by
 * oracle/build_ref.sh into oracle/_ref/libref_gauhuman.so:
 *   submodules/diff-gaussian-rasterization/cuda_rasterizer/{rasterizer_impl,forward,backward}.cu
 *   submodules/simple-knn/simple_knn.cu
 * It replaces the torch binding (rasterize_points.cu / spatial.cu / ext.cpp)
 * so the reference kernels can be driven through ctypes on the GPU box, where
 * /root/reference does not exist.  It plays two roles there: the live
 * bit-exact oracle for the -m gpu parity tests and the generator of
 * tests/golden/, and the "reference CUDA build on the same box" timing arm.
 */
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cub/cub.cuh>

#include "rasterizer.h"
#include "rasterizer_impl.h"
#include "simple_knn.h"

namespace {
struct Buf {
    char* p = nullptr;
    size_t cap = 0;
    size_t used = 0;
    char* get(size_t n)
    {
        used = n;
        if (n > cap) {
            if (p) cudaFree(p);
            cap = n + n / 4 + 256;
            cudaMalloc(&p, cap);
        }
        return p;
    }
};
struct Handle {
    Buf geom, binning, img;
    int P = 0, N = 0, W = 0, H = 0;
};
}  // namespace

extern "C" {

void* ref_create() { return new Handle(); }

void ref_destroy(void* h)
{
    Handle* s = (Handle*)h;
    if (s->geom.p) cudaFree(s->geom.p);
    if (s->binning.p) cudaFree(s->binning.p);
    if (s->img.p) cudaFree(s->img.p);
    delete s;
}

/* rasterize_points.cu:36-120 without the torch allocations: outputs are
 * caller-owned device buffers, pre-zeroed by the caller like torch::full(0). */
int ref_forward(void* h, int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                const float* campos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color,
                float* out_depth, float* out_alpha, int* radii, int debug)
{
    Handle* s = (Handle*)h;
    s->P = P; s->W = W; s->H = H;
    std::function<char*(size_t)> g = [s](size_t n) { return s->geom.get(n); };
    std::function<char*(size_t)> b = [s](size_t n) { return s->binning.get(n); };
    std::function<char*(size_t)> i = [s](size_t n) { return s->img.get(n); };
    int n = 0;
    try {
        n = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
                                                scales, scale_modifier, rotations, cov3D_precomp, view, proj, campos,
                                                tan_fovx, tan_fovy, prefiltered != 0, out_color, out_depth, out_alpha,
                                                radii, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_forward: %s\n", e.what());
        return -1;
    }
    s->N = n;
    return n;
}

/* rasterize_points.cu:122-207; grads are caller-owned, pre-zeroed. */
int ref_backward(void* h, int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D,
                 const float* shs, const float* colors_precomp, const float* alphas, const float* scales,
                 float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* view,
                 const float* proj, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                 const float* dL_dpix, const float* dL_ddepths, const float* dL_dalphas, float* dL_dmean2D,
                 float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot, int debug)
{
    Handle* s = (Handle*)h;
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, alphas, scales,
                                             scale_modifier, rotations, cov3D_precomp, view, proj, campos, tan_fovx,
                                             tan_fovy, radii, s->geom.p, s->binning.p, s->img.p, dL_dpix, dL_ddepths,
                                             dL_dalphas, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D,
                                             dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_backward: %s\n", e.what());
        return -1;
    }
    return 0;
}

void ref_mark_visible(int P, float* means3D, float* view, float* proj, bool* present)
{
    CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
}

/* Device pointers of the reference's internal state, located with the
 * reference's own fromChunk (rasterizer_impl.cu:155-194).  Order:
 * 0 depths 1 clamped 2 internal_radii 3 means2D 4 cov3D 5 conic_opacity 6 rgb
 * 7 point_offsets 8 tiles_touched | 9 point_list 10 point_list_unsorted
 * 11 point_list_keys 12 point_list_keys_unsorted | 13 accum_alpha
 * 14 n_contrib 15 ranges                                                   */
void ref_state_ptrs(void* h, void** out16)
{
    Handle* s = (Handle*)h;
    char* c = s->geom.p;
    CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c, s->P);
    out16[0] = g.depths; out16[1] = g.clamped; out16[2] = g.internal_radii; out16[3] = g.means2D;
    out16[4] = g.cov3D; out16[5] = g.conic_opacity; out16[6] = g.rgb; out16[7] = g.point_offsets;
    out16[8] = g.tiles_touched;
    c = s->binning.p;
    if (c) {
        CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c, s->N);
        out16[9] = b.point_list; out16[10] = b.point_list_unsorted; out16[11] = b.point_list_keys;
        out16[12] = b.point_list_keys_unsorted;
    } else {
        out16[9] = out16[10] = out16[11] = out16[12] = nullptr;
    }
    c = s->img.p;
    CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(c, (size_t)s->W * s->H);
    out16[13] = im.accum_alpha; out16[14] = im.n_contrib; out16[15] = im.ranges;
}

size_t ref_state_bytes(void* h, int which)
{
    Handle* s = (Handle*)h;
    return which == 0 ? s->geom.used : which == 1 ? s->binning.used : s->img.used;
}

int ref_d2h(void* dst, const void* src, size_t bytes)
{
    return (int)cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost);
}

int ref_sync() { return (int)cudaDeviceSynchronize(); }

/* spatial.cu:15-26 distCUDA2 without torch: out[P] caller-owned, zeroed. */
void ref_dist_cuda2(int P, const float* points, float* out) { SimpleKNN::knn(P, (float3*)points, out); }

/* The reference's sort call in isolation, for the Mkeys/s comparison:
 * cub::DeviceRadixSort::SortPairs(uint64 keys, uint32 vals, N, 0, bits) exactly
 * as rasterizer_impl.cu:305-310 issues it (legacy default stream, CUB from
 * the toolkit).  temp == nullptr queries the temp size.                    */
size_t ref_sort_pairs(void* temp, size_t temp_bytes, int N, int bits, const uint64_t* keys_in, uint64_t* keys_out,
                      const uint32_t* vals_in, uint32_t* vals_out)
{
    cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, N, 0, bits);
    return temp_bytes;
}

}  // extern "C"
