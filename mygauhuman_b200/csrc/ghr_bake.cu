// Cube-map -> direction lookup of the occlusion baking: the part of reference baking.py::bake_set after
// the cube-face rasterisation (:286-309).  There the six 32x32 alpha maps of a grid cell are resampled onto the 16x32
// lat-long direction grid with nvdiffrast's dr.texture(filter_mode="nearest", boundary_mode="cube") (:290-297), every
// Gaussian of the cell receives 1 - alpha of its cell (:305), and the result is masked by the hemisphere of the
// Gaussian's normal (:231,307).  nvdiffrast is an un-vendored third party; its cube addressing is the OpenGL one
// (major axis selects the face +x,-x,+y,-y,+z,-z; (sc, tc) per face as in the OpenGL 4.6 specification, table 8.19),
// restated here and in oracle/cubemap_oracle.py.  No graphics API is involved: it is a gather.
//
// HBM stream: reads n_maps * 6 * res^2 * ch floats once (L2-resident: 24 KB per cell), writes n_maps * n_dirs * ch;
// the fused occlusion kernel writes P * n_dirs floats (2 KB per Gaussian at 16x32) and reads 16 bytes per Gaussian.
#include <math.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

// direction -> texel index inside one cube map laid out [6][res][res]; -1 for a direction without a major axis (NaN / 0)
__device__ __forceinline__ int cube_texel_nearest(float x, float y, float z, int res)
{
    const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
    int face;
    float c, u, v;
    if (az > fmaxf(ax, ay)) { face = 4; c = z; u = x; v = y; }
    else if (ay > ax) { face = 2; c = y; u = x; v = z; }
    else { face = 0; c = x; u = z; v = y; }
    if (c < 0.0f) face += 1;
    const float m = __frcp_rz(fabsf(c)) * 0.5f;
    // sign of the two in-face coordinates per face: u is mirrored on +x and -z, v on every face but +y
    const float mu = (face == 0 || face == 5) ? -m : m;
    const float mv = (face != 2) ? -m : m;
    float s = fmaf(u, mu, 0.5f), t = fmaf(v, mv, 0.5f);
    if (!isfinite(s) || !isfinite(t)) return -1;
    s = fminf(fmaxf(s, 0.0f), 1.0f);
    t = fminf(fmaxf(t, 0.0f), 1.0f);
    const int iu = min(max((int)floorf(s * (float)res), 0), res - 1);
    const int iv = min(max((int)floorf(t * (float)res), 0), res - 1);
    return (face * res + iv) * res + iu;
}

// out[m][d][c] = tex[m][texel(d)][c]   (tex [n_maps][6][res][res][ch], dirs [n_dirs][3])
__global__ void __launch_bounds__(256) cubemap_sample_kernel(int n_maps, int res, int ch, const float* __restrict__ tex,
                                                             int n_dirs, const float* __restrict__ dirs,
                                                             float* __restrict__ out)
{
    const int64_t total = (int64_t)n_maps * n_dirs;
    const int64_t map_floats = (int64_t)6 * res * res * ch;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int d = (int)(i % n_dirs);
        const int64_t m = i / n_dirs;
        const int tx = cube_texel_nearest(dirs[3 * d], dirs[3 * d + 1], dirs[3 * d + 2], res);
        for (int c = 0; c < ch; c++) out[i * ch + c] = tx < 0 ? 0.0f : tex[m * map_floats + (int64_t)tx * ch + c];
    }
}

// occlusion[p][d] = dot(dirs[d], normal[p]) > 0 ? 1 - alpha_cube[cell_of[p]][texel(d)] : 0      (bake_set :231,305-307)
// One CTA per Gaussian group; the per-direction texel table is built once per CTA in shared memory.
__global__ void __launch_bounds__(256) bake_occlusion_kernel(int P, int n_dirs, int n_cells, int res,
                                                             const float* __restrict__ cubes,  // [n_cells][6][res][res]
                                                             const int64_t* __restrict__ cell_of,
                                                             const float* __restrict__ normals,
                                                             const float* __restrict__ dirs, float* __restrict__ out)
{
    extern __shared__ int s_texel[];  // [n_dirs]
    float* s_dir = reinterpret_cast<float*>(s_texel + n_dirs);  // [n_dirs][3]
    for (int d = threadIdx.x; d < n_dirs; d += blockDim.x) {
        const float x = dirs[3 * d], y = dirs[3 * d + 1], z = dirs[3 * d + 2];
        s_texel[d] = cube_texel_nearest(x, y, z, res);
        s_dir[3 * d] = x; s_dir[3 * d + 1] = y; s_dir[3 * d + 2] = z;
    }
    __syncthreads();
    const int64_t cube_floats = (int64_t)6 * res * res;
    for (int p = blockIdx.x; p < P; p += gridDim.x) {
        const int64_t cell = cell_of[p];
        const float nx = normals[3 * (size_t)p], ny = normals[3 * (size_t)p + 1], nz = normals[3 * (size_t)p + 2];
        const float* cube = cubes + cell * cube_floats;
        const bool ok = cell >= 0 && cell < n_cells;
        for (int d = threadIdx.x; d < n_dirs; d += blockDim.x) {
            // the reference multiplies and sums in this order: (dirs * normal).sum(-1) > 0
            const float dot = s_dir[3 * d] * nx + s_dir[3 * d + 1] * ny + s_dir[3 * d + 2] * nz;
            const int tx = s_texel[d];
            const float a = (ok && tx >= 0) ? cube[tx] : 0.0f;
            out[(size_t)p * n_dirs + d] = dot > 0.0f ? 1.0f - a : 0.0f;
        }
    }
}

}  // namespace ghr

using namespace ghr;

extern "C" {

int ghr_cubemap_sample_nearest(int32_t n_maps, int32_t res, int32_t channels, const float* tex, int32_t n_dirs,
                               const float* dirs, float* out, void* stream)
{
    if (n_maps < 0 || res <= 0 || channels <= 0 || n_dirs < 0) return GHR_ERR_BAD_ARG;
    if (n_maps == 0 || n_dirs == 0) return GHR_OK;
    if (!tex || !dirs || !out) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_MISC, s);
    const int64_t total = (int64_t)n_maps * n_dirs;
    int64_t grid = (total + 255) / 256;
    if (grid > (int64_t)sm_count() * 16) grid = (int64_t)sm_count() * 16;
    cubemap_sample_kernel<<<(unsigned)grid, 256, 0, s>>>(n_maps, res, channels, tex, n_dirs, dirs, out);
    count_launch();
    return cudaGetLastError() == cudaSuccess ? GHR_OK : GHR_ERR_CUDA;
}

int ghr_bake_occlusion(int32_t P, int32_t n_dirs, int32_t n_cells, int32_t res, const float* cubemaps,
                       const int64_t* cell_of, const float* normals, const float* dirs, float* out, void* stream)
{
    if (P < 0 || n_dirs < 0 || n_cells < 0 || res <= 0 || n_dirs > 8192) return GHR_ERR_BAD_ARG;
    if (P == 0 || n_dirs == 0) return GHR_OK;
    if (!cubemaps || !cell_of || !normals || !dirs || !out) return GHR_ERR_BAD_ARG;
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer(GHR_STAGE_MISC, s);
    int grid = P < sm_count() * 8 ? P : sm_count() * 8;
    const size_t smem = (size_t)n_dirs * (sizeof(int) + 3 * sizeof(float));
    cudaError_t e = cudaFuncSetAttribute(bake_occlusion_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return GHR_ERR_CUDA;
    bake_occlusion_kernel<<<grid, 256, smem, s>>>(P, n_dirs, n_cells, res, cubemaps, cell_of, normals, dirs, out);
    count_launch();
    return cudaGetLastError() == cudaSuccess ? GHR_OK : GHR_ERR_CUDA;
}

}  // extern "C"
