// PLY checkpoint rows: the reference's GaussianModel.save_ply / load_ply
// (scene/gaussian_model.py:310-407) move nine parameter tensors to the host one by one, concatenate them with numpy and
// build the structured array through list(map(tuple, ...)).  Here the interleaved rows the file holds are produced /
// consumed on the device in one pass (one D2H / H2D copy of P x (18 + 3M) floats):
//   x y z | nx ny nz | ar ag ab | roughness | f_dc_0..2 | f_rest_* (CHANNEL-major: features_rest.transpose(1,2),
//   :330) | opacity | scale_0..2 | rot_0..3                    (construct_list_of_attributes, :309-326)
// Pure HBM streams: 2 x 4 x (18 + 3M) bytes per Gaussian.
#include <stdio.h>
#include "ghr_internal.h"

namespace ghr {
namespace {

struct PlyFields {
    float *xyz, *normal, *albedo, *roughness, *f_dc, *f_rest, *opacity, *scaling, *rotation;
    int P, M;          // M = SH coefficients per channel including the DC term
    int cols;          // floats per file row
    int col_of[80];    // file column of canonical column c (identity when packing)
};

// canonical column c of Gaussian p -> address of the parameter element it holds (NULL: absent optional field)
__device__ __forceinline__ float* ply_field(const PlyFields& f, int p, int c)
{
    const int R = f.M - 1;  // rest coefficients per channel
    if (c < 3) return f.xyz + 3 * (size_t)p + c;
    if (c < 6) return f.normal ? f.normal + 3 * (size_t)p + (c - 3) : nullptr;
    if (c < 9) return f.albedo ? f.albedo + 3 * (size_t)p + (c - 6) : nullptr;
    if (c < 10) return f.roughness ? f.roughness + p : nullptr;
    if (c < 13) return f.f_dc + 3 * (size_t)p + (c - 10);
    c -= 13;
    if (c < 3 * R) {  // file order: channel-major [3][R]; tensor order: [R][3]
        const int ch = c / R, k = c - ch * R;
        return f.f_rest + ((size_t)p * R + k) * 3 + ch;
    }
    c -= 3 * R;
    if (c < 1) return f.opacity + p;
    if (c < 4) return f.scaling + 3 * (size_t)p + (c - 1);
    return f.rotation + 4 * (size_t)p + (c - 4);
}

template <bool kPack>
__global__ void __launch_bounds__(256) ply_rows_kernel(const __grid_constant__ PlyFields f, float* __restrict__ rows)
{
    const int canon = 18 + 3 * f.M;  // 10 + 3 + 3(M-1) + 1 + 3 + 4
    const int64_t total = (int64_t)f.P * canon;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
        const int p = (int)(i / canon), c = (int)(i - (int64_t)p * canon);
        float* src = ply_field(f, p, c);
        float* cell = rows + (size_t)p * f.cols + f.col_of[c];
        if (kPack) *cell = src ? *src : 0.0f;
        else if (src) *src = *cell;
    }
}

int launch(bool pack, const PlyFields& f, float* rows, cudaStream_t s)
{
    const int64_t total = (int64_t)f.P * (18 + 3 * f.M);
    int64_t grid = (total + 255) / 256;
    if (grid > sm_count() * 32) grid = sm_count() * 32;
    StageTimer timer(GHR_STAGE_MISC, s);
    if (pack) ply_rows_kernel<true><<<(unsigned)grid, 256, 0, s>>>(f, rows);
    else ply_rows_kernel<false><<<(unsigned)grid, 256, 0, s>>>(f, rows);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[libghr] CUDA error in ply rows: %s\n", cudaGetErrorString(e));
        return GHR_ERR_CUDA;
    }
    return GHR_OK;
}

int fill(PlyFields& f, int32_t P, int32_t M, const float* xyz, const float* normal, const float* albedo,
         const float* roughness, const float* f_dc, const float* f_rest, const float* opacity, const float* scaling,
         const float* rotation, int32_t row_floats, const int32_t* columns)
{
    if (P < 0 || M < 1 || M > 16) return GHR_ERR_BAD_ARG;
    const int canon = 18 + 3 * M;
    if (row_floats < canon) return GHR_ERR_BAD_ARG;
    if (P > 0 && (!xyz || !f_dc || (M > 1 && !f_rest) || !opacity || !scaling || !rotation)) return GHR_ERR_BAD_ARG;
    auto nc = [](const float* p) { return const_cast<float*>(p); };
    f.xyz = nc(xyz); f.normal = nc(normal); f.albedo = nc(albedo); f.roughness = nc(roughness); f.f_dc = nc(f_dc);
    f.f_rest = nc(f_rest); f.opacity = nc(opacity); f.scaling = nc(scaling); f.rotation = nc(rotation);
    f.P = P; f.M = M; f.cols = row_floats;
    for (int c = 0; c < canon; c++) {
        f.col_of[c] = columns ? columns[c] : c;
        if (f.col_of[c] < 0 || f.col_of[c] >= row_floats) return GHR_ERR_BAD_ARG;
    }
    return GHR_OK;
}

}  // namespace
}  // namespace ghr

using namespace ghr;

extern "C" {

int ghr_ply_row_floats(int32_t M) { return (M < 1 || M > 16) ? 0 : 18 + 3 * M; }

int ghr_ply_pack_rows(int32_t P, int32_t M, const float* xyz, const float* normal, const float* albedo,
                      const float* roughness, const float* f_dc, const float* f_rest, const float* opacity,
                      const float* scaling, const float* rotation, float* rows, void* stream)
{
    PlyFields f;
    int r = fill(f, P, M, xyz, normal, albedo, roughness, f_dc, f_rest, opacity, scaling, rotation, 18 + 3 * (M < 1 ? 1 : M),
                 nullptr);
    if (r != GHR_OK) return r;
    if (P == 0) return GHR_OK;
    if (!rows) return GHR_ERR_BAD_ARG;
    return launch(true, f, rows, (cudaStream_t)stream);
}

int ghr_ply_unpack_rows(int32_t P, int32_t M, const float* rows, int32_t row_floats, const int32_t* columns, float* xyz,
                        float* normal, float* albedo, float* roughness, float* f_dc, float* f_rest, float* opacity,
                        float* scaling, float* rotation, void* stream)
{
    PlyFields f;
    int r = fill(f, P, M, xyz, normal, albedo, roughness, f_dc, f_rest, opacity, scaling, rotation, row_floats, columns);
    if (r != GHR_OK) return r;
    if (P == 0) return GHR_OK;
    if (!rows) return GHR_ERR_BAD_ARG;
    return launch(false, f, const_cast<float*>(rows), (cudaStream_t)stream);
}

}  // extern "C"
