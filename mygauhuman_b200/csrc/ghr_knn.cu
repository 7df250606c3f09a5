// distCUDA2 (reference submodules/simple-knn/simple_knn.cu:185-221 via
// spatial.cu:15-26): mean of the three smallest squared distances from each
// point to the other points (self excluded by index).
//
// The reference Morton-sorts the points and prunes 1024-point boxes; the
// result is the exact 3-NN value either way.  Here: the points are binned into
// a uniform grid whose cell size is chosen from the AABB and P (about 2 points
// per cell), cells are counted/scanned/filled on the device, and each point
// searches rings of cells around its own cell until the 3rd-best distance is
// provably final.  Small inputs (P <= kBruteMax = 4096) use a shared-memory
// tiled all-pairs kernel instead, which needs no auxiliary structure; measured
// on B200 the grid wins from a few thousand points on (6,890 points: grid
// faster than the 0.41 ms all-pairs pass; 50,000: 3.5 ms all-pairs).
#include <float.h>
#include "ghr_common.cuh"
#include "ghr_internal.h"

namespace ghr {

constexpr int kBruteMax = 1 << 12;
constexpr int kKnnTile = 1024;

__device__ __forceinline__ void kbest3(float d, float& b0, float& b1, float& b2)
{
    // updateKBest<3>, simple_knn.cu:131-145
    if (b0 > d) { const float t = b0; b0 = d; d = t; }
    if (b1 > d) { const float t = b1; b1 = d; d = t; }
    if (b2 > d) { b2 = d; }
}

__device__ __forceinline__ float dist2(float ax, float ay, float az, float bx, float by, float bz)
{
    const float dx = __fsub_rn(bx, ax), dy = __fsub_rn(by, ay), dz = __fsub_rn(bz, az);
    float t = __fmul_rn(dy, dy);
    t = __fmaf_rn(dx, dx, t);
    return __fmaf_rn(dz, dz, t);
}

__global__ void __launch_bounds__(256) knn_brute_kernel(int P, const float* __restrict__ pts, float* __restrict__ out)
{
    __shared__ float s_p[kKnnTile * 3];
    const int i = blockIdx.x * 256 + threadIdx.x;
    float px = 0.f, py = 0.f, pz = 0.f;
    if (i < P) { px = pts[3 * i]; py = pts[3 * i + 1]; pz = pts[3 * i + 2]; }
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    for (int base = 0; base < P; base += kKnnTile) {
        const int n = min(kKnnTile, P - base);
        __syncthreads();
        for (int k = threadIdx.x; k < n * 3; k += 256) s_p[k] = pts[3 * (size_t)base + k];
        __syncthreads();
        if (i < P) {
            for (int j = 0; j < n; j++) {
                if (base + j == i) continue;
                kbest3(dist2(px, py, pz, s_p[3 * j], s_p[3 * j + 1], s_p[3 * j + 2]), b0, b1, b2);
            }
        }
    }
    if (i < P) out[i] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

// ---- uniform-grid path for large P -----------------------------------------
struct KnnGrid {
    float minx, miny, minz, inv_cell, cell;
    int nx, ny, nz;
};

__global__ void knn_aabb_kernel(int P, const float* __restrict__ pts, float* aabb /*min3,max3 as ordered ints*/)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
        for (int c = 0; c < 3; c++) { const float v = pts[3 * i + c]; mn[c] = fminf(mn[c], v); mx[c] = fmaxf(mx[c], v); }
    for (int c = 0; c < 3; c++) {
        for (int o = 16; o; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        // order-preserving float<->int mapping for atomicMin/Max on signed ints
        for (int c = 0; c < 3; c++) {
            int a = __float_as_int(mn[c]); a = a >= 0 ? a : a ^ 0x7FFFFFFF;
            int b = __float_as_int(mx[c]); b = b >= 0 ? b : b ^ 0x7FFFFFFF;
            atomicMin(reinterpret_cast<int*>(aabb) + c, a);
            atomicMax(reinterpret_cast<int*>(aabb) + 3 + c, b);
        }
    }
}

__device__ __forceinline__ float ordered_to_float(int a) { return __int_as_float(a >= 0 ? a : a ^ 0x7FFFFFFF); }

__global__ void knn_grid_setup_kernel(int P, const float* aabb, KnnGrid* g, int max_cells)
{
    const int* ia = reinterpret_cast<const int*>(aabb);
    const float mnx = ordered_to_float(ia[0]), mny = ordered_to_float(ia[1]), mnz = ordered_to_float(ia[2]);
    const float ex = fmaxf(ordered_to_float(ia[3]) - mnx, 1e-20f), ey = fmaxf(ordered_to_float(ia[4]) - mny, 1e-20f),
                ez = fmaxf(ordered_to_float(ia[5]) - mnz, 1e-20f);
    // ~2 points per cell on average over the occupied volume
    float cell = cbrtf(ex * ey * ez * 2.0f / (float)P);
    cell = fmaxf(cell, fmaxf(ex, fmaxf(ey, ez)) / 1024.0f);
    int nx, ny, nz;
    while (true) {
        nx = (int)(ex / cell) + 1; ny = (int)(ey / cell) + 1; nz = (int)(ez / cell) + 1;
        if ((long long)nx * ny * nz <= max_cells) break;
        cell *= 1.26f;
    }
    g->minx = mnx; g->miny = mny; g->minz = mnz; g->cell = cell; g->inv_cell = 1.0f / cell;
    g->nx = nx; g->ny = ny; g->nz = nz;
}

__device__ __forceinline__ int3 cell_of(const KnnGrid& g, float x, float y, float z)
{
    int cx = (int)((x - g.minx) * g.inv_cell), cy = (int)((y - g.miny) * g.inv_cell), cz = (int)((z - g.minz) * g.inv_cell);
    cx = min(max(cx, 0), g.nx - 1); cy = min(max(cy, 0), g.ny - 1); cz = min(max(cz, 0), g.nz - 1);
    return make_int3(cx, cy, cz);
}

__global__ void knn_count_kernel(int P, const float* __restrict__ pts, const KnnGrid* gp, uint32_t* counts)
{
    const KnnGrid g = *gp;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int3 c = cell_of(g, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
    atomicAdd(&counts[(c.z * g.ny + c.y) * g.nx + c.x], 1u);
}

// single-block exclusive scan over the cells (cells <= max_cells, a few M)
__global__ void __launch_bounds__(1024) knn_scan_kernel(const KnnGrid* gp, const uint32_t* counts, uint32_t* starts)
{
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    const int n = gp->nx * gp->ny * gp->nz;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t c = i < n ? counts[i] : 0;
        uint32_t v = c;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, v, o); if ((threadIdx.x & 31) >= o) v += t; }
        if ((threadIdx.x & 31) == 31) s_w[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = s_w[threadIdx.x];
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, w, o); if (threadIdx.x >= o) w += t; }
            s_w[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t wp = (threadIdx.x >> 5) ? s_w[(threadIdx.x >> 5) - 1] : 0;
        const uint32_t carry = s_carry;
        if (i < n) starts[i] = carry + wp + v - c;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + wp + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) starts[n] = s_carry;
}

__global__ void knn_fill_kernel(int P, const float* __restrict__ pts, const KnnGrid* gp, const uint32_t* starts,
                                uint32_t* cursor, float4* sorted /*xyz + index*/)
{
    const KnnGrid g = *gp;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
    const int3 c = cell_of(g, x, y, z);
    const int cid = (c.z * g.ny + c.y) * g.nx + c.x;
    const uint32_t slot = starts[cid] + atomicAdd(&cursor[cid], 1u);
    sorted[slot] = make_float4(x, y, z, __int_as_float(i));
}

__global__ void __launch_bounds__(128) knn_grid_search_kernel(int P, const KnnGrid* gp, const uint32_t* __restrict__ starts,
                                                              const float4* __restrict__ sorted, float* __restrict__ out)
{
    const KnnGrid g = *gp;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;  // iterate in cell order: neighbours share cache lines
    if (s >= P) return;
    const float4 me = sorted[s];
    const int self = __float_as_int(me.w);
    const int3 c = cell_of(g, me.x, me.y, me.z);
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    const int maxr = max(g.nx, max(g.ny, g.nz));
    for (int r = 0; r <= maxr; r++) {
        // every point outside the (2r-1)^3 block around the cell is at least this far away
        if (r > 0) {
            const float lo = fminf(fminf(me.x - (g.minx + (c.x - r + 1) * g.cell), (g.minx + (c.x + r) * g.cell) - me.x),
                                   fminf(fminf(me.y - (g.miny + (c.y - r + 1) * g.cell), (g.miny + (c.y + r) * g.cell) - me.y),
                                         fminf(me.z - (g.minz + (c.z - r + 1) * g.cell), (g.minz + (c.z + r) * g.cell) - me.z)));
            // conservative against the rounding of cell assignment / plane coordinates
            const float eps = 4e-7f * (fabsf(me.x) + fabsf(me.y) + fabsf(me.z) + g.cell * (float)maxr);
            const float bound = fmaxf(lo - eps, 0.0f) * 0.999f;
            if (bound * bound > b2) break;
        }
        const int z0 = max(c.z - r, 0), z1 = min(c.z + r, g.nz - 1);
        const int y0 = max(c.y - r, 0), y1 = min(c.y + r, g.ny - 1);
        const int x0 = max(c.x - r, 0), x1 = min(c.x + r, g.nx - 1);
        for (int zz = z0; zz <= z1; zz++)
            for (int yy = y0; yy <= y1; yy++) {
                const bool shell_row = (abs(zz - c.z) == r) || (abs(yy - c.y) == r);
                const int row = (zz * g.ny + yy) * g.nx;
                if (shell_row) {
                    const uint32_t a = starts[row + x0], b = starts[row + x1 + 1];
                    for (uint32_t k = a; k < b; k++) {
                        const float4 q = sorted[k];
                        if (__float_as_int(q.w) == self) continue;
                        kbest3(dist2(me.x, me.y, me.z, q.x, q.y, q.z), b0, b1, b2);
                    }
                } else {
                    // interior row of the shell: only the two end cells are new
                    for (int e = 0; e < 2; e++) {
                        const int xx = e ? c.x + r : c.x - r;
                        if (xx < 0 || xx >= g.nx || (e && r == 0)) continue;
                        const uint32_t a = starts[row + xx], b = starts[row + xx + 1];
                        for (uint32_t k = a; k < b; k++) {
                            const float4 q = sorted[k];
                            if (__float_as_int(q.w) == self) continue;
                            kbest3(dist2(me.x, me.y, me.z, q.x, q.y, q.z), b0, b1, b2);
                        }
                    }
                }
            }
    }
    out[self] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

// ---- nearest model vertex for arbitrary query points (LBS, knn_cuda KNN(k=1)) ----
// Exact: rings of cells are searched until the best distance is provably final;
// lowest vertex index wins ties.  Faces of the searched block that lie on or
// beyond the grid boundary hide no unexamined cell and are ignored, so queries
// outside the vertex AABB terminate as quickly as interior ones.
__global__ void __launch_bounds__(128) nearest_vertex_kernel(int P, const float* __restrict__ query, const KnnGrid* gp,
                                                             const uint32_t* __restrict__ starts,
                                                             const float4* __restrict__ sorted, int32_t* __restrict__ vid)
{
    const KnnGrid g = *gp;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float qx = query[3 * i], qy = query[3 * i + 1], qz = query[3 * i + 2];
    const int3 c = cell_of(g, qx, qy, qz);
    float best = FLT_MAX;
    int best_id = 0x7FFFFFFF;
    const int maxr = max(g.nx, max(g.ny, g.nz));
    const float inf = __int_as_float(0x7f800000);
    for (int r = 0; r <= maxr; r++) {
        if (r > 0) {
            float lo = inf;
            if (c.x - r + 1 > 0) lo = fminf(lo, qx - (g.minx + (c.x - r + 1) * g.cell));
            if (c.x + r < g.nx) lo = fminf(lo, (g.minx + (c.x + r) * g.cell) - qx);
            if (c.y - r + 1 > 0) lo = fminf(lo, qy - (g.miny + (c.y - r + 1) * g.cell));
            if (c.y + r < g.ny) lo = fminf(lo, (g.miny + (c.y + r) * g.cell) - qy);
            if (c.z - r + 1 > 0) lo = fminf(lo, qz - (g.minz + (c.z - r + 1) * g.cell));
            if (c.z + r < g.nz) lo = fminf(lo, (g.minz + (c.z + r) * g.cell) - qz);
            if (lo == inf) break;  // every cell has been examined
            const float eps = 4e-7f * (fabsf(qx) + fabsf(qy) + fabsf(qz) + fabsf(g.minx) + fabsf(g.miny) + fabsf(g.minz) +
                                       g.cell * (float)maxr);
            const float bound = fmaxf(lo - eps, 0.0f) * 0.999f;
            if (bound * bound > best) break;
        }
        const int z0 = max(c.z - r, 0), z1 = min(c.z + r, g.nz - 1);
        const int y0 = max(c.y - r, 0), y1 = min(c.y + r, g.ny - 1);
        const int x0 = max(c.x - r, 0), x1 = min(c.x + r, g.nx - 1);
        for (int zz = z0; zz <= z1; zz++)
            for (int yy = y0; yy <= y1; yy++) {
                const bool shell_row = (abs(zz - c.z) == r) || (abs(yy - c.y) == r);
                const int row = (zz * g.ny + yy) * g.nx;
                for (int e = 0; e < 2; e++) {
                    uint32_t a, b;
                    if (shell_row) {
                        if (e) break;
                        a = starts[row + x0]; b = starts[row + x1 + 1];
                    } else {
                        const int xx = e ? c.x + r : c.x - r;
                        if (xx < 0 || xx >= g.nx || (e && r == 0)) continue;
                        a = starts[row + xx]; b = starts[row + xx + 1];
                    }
                    for (uint32_t k = a; k < b; k++) {
                        const float4 v = sorted[k];
                        const float d = dist2(qx, qy, qz, v.x, v.y, v.z);
                        const int id = __float_as_int(v.w);
                        if (d < best || (d == best && id < best_id)) { best = d; best_id = id; }
                    }
                }
            }
    }
    vid[i] = best_id;
}

// Variant with the whole grid (sorted vertices + cell starts) resident in shared memory (6,890 vertices =
// 110 KB + <= 8192 cells = 32 KB: fits B200's 227 KB).  Eight lanes answer one query: the (2r+1)^2 cell rows
// of the searched block are dealt round-robin to the eight lanes, each walks its rows' vertices, then a
// 3-level arg-min butterfly inside the group.  Compared with one thread per query through L2 this removes
// most of the divergence (7/32 active lanes measured there) and turns ~30 dependent L2 round trips per
// query into shared-memory reads; compared with a whole warp per query it keeps 4 queries in flight per warp.
constexpr int kVertexGridCells = 8192;

__global__ void __launch_bounds__(1024, 1) nearest_vertex_smem_kernel(int P, int V, const float* __restrict__ query,
                                                                      const KnnGrid* gp,
                                                                      const uint32_t* __restrict__ starts_g,
                                                                      const float4* __restrict__ sorted_g,
                                                                      int32_t* __restrict__ vid)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* s_pts = reinterpret_cast<float4*>(smem_raw);
    uint32_t* s_starts = reinterpret_cast<uint32_t*>(smem_raw + sizeof(float4) * (size_t)V);
    const KnnGrid g = *gp;
    const int ncell = g.nx * g.ny * g.nz;
    for (int k = threadIdx.x; k < V; k += blockDim.x) s_pts[k] = sorted_g[k];
    for (int k = threadIdx.x; k <= ncell; k += blockDim.x) s_starts[k] = starts_g[k];
    __syncthreads();
    const int sub = threadIdx.x & 7;                                   // lane inside the 8-lane group
    const int group = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;    // global group id
    const int ngroups = (gridDim.x * blockDim.x) >> 3;
    const int maxr = max(g.nx, max(g.ny, g.nz));
    const float inf = __int_as_float(0x7f800000);
    // all 32 lanes of a warp run the same number of outer iterations (shuffles below are full-warp)
    const int iters = (P + ngroups - 1) / ngroups;
    for (int it = 0; it < iters; it++) {
        const int qi = it * ngroups + group;
        const bool live = qi < P;
        const float qx = live ? query[3 * qi] : g.minx, qy = live ? query[3 * qi + 1] : g.miny,
                    qz = live ? query[3 * qi + 2] : g.minz;
        const int3 c = cell_of(g, qx, qy, qz);
        const float eps = 4e-7f * (fabsf(qx) + fabsf(qy) + fabsf(qz) + fabsf(g.minx) + fabsf(g.miny) + fabsf(g.minz) +
                                   g.cell * (float)maxr);
        float best = FLT_MAX;
        int best_id = 0x7FFFFFFF;
        auto scan = [&](uint32_t a, uint32_t b, uint32_t first, uint32_t step) {
            for (uint32_t k = a + first; k < b; k += step) {
                const float4 v = s_pts[k];
                const float d = dist2(qx, qy, qz, v.x, v.y, v.z);
                const int id = __float_as_int(v.w);
                if (d < best || (d == best && id < best_id)) { best = d; best_id = id; }
            }
        };
        auto group_min = [&]() {
#pragma unroll
            for (int o = 4; o; o >>= 1) {
                const float od = __shfl_xor_sync(0xffffffffu, best, o);
                const int oid = __shfl_xor_sync(0xffffffffu, best_id, o);
                if (od < best || (od == best && oid < best_id)) { best = od; best_id = oid; }
            }
        };
        // phase A: the query's own cell, its vertices dealt to the 8 lanes -> a tight bound for pruning
        if (live) {
            const int cid = (c.z * g.ny + c.y) * g.nx + c.x;
            scan(s_starts[cid], s_starts[cid + 1], sub, 8);
        }
        group_min();
        // phase B: shells r = 1, 2, ...; a cell is skipped when its box is provably farther than the best so far
        bool done = !live;
        for (int r = 1; r <= maxr + 1; r++) {
            if (!done && r == 1) {
                // 3x3x3 shell: per axis, a neighbour slab is needed only if its face is closer than the bound
                // from phase A (group-uniform).  The few needed cells are walked one after the other with the
                // eight lanes splitting each cell's vertices.
                const float bnd = best;
                const float fxm = fmaxf(qx - (g.minx + c.x * g.cell) - eps, 0.0f), fxp = fmaxf((g.minx + (c.x + 1) * g.cell) - qx - eps, 0.0f);
                const float fym = fmaxf(qy - (g.miny + c.y * g.cell) - eps, 0.0f), fyp = fmaxf((g.miny + (c.y + 1) * g.cell) - qy - eps, 0.0f);
                const float fzm = fmaxf(qz - (g.minz + c.z * g.cell) - eps, 0.0f), fzp = fmaxf((g.minz + (c.z + 1) * g.cell) - qz - eps, 0.0f);
                const float xm2 = fxm * fxm * 0.998f, xp2 = fxp * fxp * 0.998f, ym2 = fym * fym * 0.998f,
                            yp2 = fyp * fyp * 0.998f, zm2 = fzm * fzm * 0.998f, zp2 = fzp * fzp * 0.998f;
                const int ox0 = (c.x > 0 && xm2 <= bnd) ? -1 : 0, ox1 = (c.x + 1 < g.nx && xp2 <= bnd) ? 1 : 0;
                const int oy0 = (c.y > 0 && ym2 <= bnd) ? -1 : 0, oy1 = (c.y + 1 < g.ny && yp2 <= bnd) ? 1 : 0;
                const int oz0 = (c.z > 0 && zm2 <= bnd) ? -1 : 0, oz1 = (c.z + 1 < g.nz && zp2 <= bnd) ? 1 : 0;
                for (int oz = oz0; oz <= oz1; oz++)
                    for (int oy = oy0; oy <= oy1; oy++)
                        for (int ox = ox0; ox <= ox1; ox++) {
                            if ((ox | oy | oz) == 0) continue;  // own cell: phase A
                            const float lb = (ox < 0 ? xm2 : (ox > 0 ? xp2 : 0.0f)) + (oy < 0 ? ym2 : (oy > 0 ? yp2 : 0.0f)) +
                                             (oz < 0 ? zm2 : (oz > 0 ? zp2 : 0.0f));
                            if (lb > bnd) continue;
                            const int cid = ((c.z + oz) * g.ny + (c.y + oy)) * g.nx + (c.x + ox);
                            scan(s_starts[cid], s_starts[cid + 1], sub, 8);
                        }
            } else if (!done) {
                const int z0 = max(c.z - r, 0), z1 = min(c.z + r, g.nz - 1);
                const int y0 = max(c.y - r, 0), y1 = min(c.y + r, g.ny - 1);
                const int x0 = max(c.x - r, 0), x1 = min(c.x + r, g.nx - 1);
                const int bx = x1 - x0 + 1, by = y1 - y0 + 1, ncells = bx * by * (z1 - z0 + 1);
                for (int t = sub; t < ncells; t += 8) {
                    const int xx = x0 + t % bx, yy = y0 + (t / bx) % by, zz = z0 + t / (bx * by);
                    const int ax = abs(xx - c.x), ay = abs(yy - c.y), az = abs(zz - c.z);
                    if (max(ax, max(ay, az)) != r) continue;  // inner cells were examined by earlier shells
                    // distance from the query to the cell's box (0 along an axis where the cell index matches)
                    const float ddx = xx < c.x ? qx - (g.minx + (xx + 1) * g.cell) : (xx > c.x ? (g.minx + xx * g.cell) - qx : 0.0f);
                    const float ddy = yy < c.y ? qy - (g.miny + (yy + 1) * g.cell) : (yy > c.y ? (g.miny + yy * g.cell) - qy : 0.0f);
                    const float ddz = zz < c.z ? qz - (g.minz + (zz + 1) * g.cell) : (zz > c.z ? (g.minz + zz * g.cell) - qz : 0.0f);
                    const float ex = fmaxf(ddx - eps, 0.0f), ey = fmaxf(ddy - eps, 0.0f), ez = fmaxf(ddz - eps, 0.0f);
                    if ((ex * ex + ey * ey + ez * ez) * 0.998f > best) continue;
                    const int cid = (zz * g.ny + yy) * g.nx + xx;
                    scan(s_starts[cid], s_starts[cid + 1], 0, 1);
                }
            }
            group_min();
            if (!done) {
                // unexamined cells lie beyond the faces of this block that are inside the grid
                float lo = inf;
                if (c.x - r > 0) lo = fminf(lo, qx - (g.minx + (c.x - r) * g.cell));
                if (c.x + r + 1 < g.nx) lo = fminf(lo, (g.minx + (c.x + r + 1) * g.cell) - qx);
                if (c.y - r > 0) lo = fminf(lo, qy - (g.miny + (c.y - r) * g.cell));
                if (c.y + r + 1 < g.ny) lo = fminf(lo, (g.miny + (c.y + r + 1) * g.cell) - qy);
                if (c.z - r > 0) lo = fminf(lo, qz - (g.minz + (c.z - r) * g.cell));
                if (c.z + r + 1 < g.nz) lo = fminf(lo, (g.minz + (c.z + r + 1) * g.cell) - qz);
                const float bound = fmaxf(lo - eps, 0.0f) * 0.999f;
                if (lo == inf || bound * bound > best) done = true;
            }
            if (__all_sync(0xffffffffu, done)) break;
        }
        if (live && sub == 0) vid[qi] = best_id;
    }
}

constexpr int kMaxCells = 1 << 22;

size_t knn_workspace_bytes(int P)
{
    if (P < 0) return 0;
    if (P <= kBruteMax) return 256;
    // aabb(32) + grid(64) | counts[cells] | starts[cells+1] | sorted float4[P]
    return 256 + align_up(sizeof(uint32_t) * (size_t)kMaxCells) + align_up(sizeof(uint32_t) * ((size_t)kMaxCells + 1)) +
           align_up(sizeof(float4) * (size_t)P);
}

cudaError_t launch_knn(int P, const float* points, float* out, char* ws, size_t ws_bytes, cudaStream_t s)
{
    if (P <= 0) return cudaSuccess;
    StageTimer timer(GHR_STAGE_KNN, s);
    if (P <= kBruteMax) {
        count_launch();
        knn_brute_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, out);
        return cudaGetLastError();
    }
    (void)ws_bytes;
    count_launch(6);
    float* aabb = reinterpret_cast<float*>(ws);
    KnnGrid* grid = reinterpret_cast<KnnGrid*>(ws + 64);
    uint32_t* counts = reinterpret_cast<uint32_t*>(ws + 256);
    uint32_t* starts = reinterpret_cast<uint32_t*>(ws + 256 + align_up(sizeof(uint32_t) * (size_t)kMaxCells));
    float4* sorted = reinterpret_cast<float4*>(reinterpret_cast<char*>(starts) + align_up(sizeof(uint32_t) * ((size_t)kMaxCells + 1)));
    // aabb init: min = +max ordered int, max = -max
    const int init[6] = {0x7F7FFFFF, 0x7F7FFFFF, 0x7F7FFFFF, (int)0x80800000, (int)0x80800000, (int)0x80800000};
    cudaError_t e = cudaMemcpyAsync(aabb, init, sizeof(init), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
    knn_aabb_kernel<<<sm_count() * 4, 256, 0, s>>>(P, points, aabb);
    knn_grid_setup_kernel<<<1, 1, 0, s>>>(P, aabb, grid, kMaxCells);
    e = cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (size_t)kMaxCells, s);
    if (e != cudaSuccess) return e;
    knn_count_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, grid, counts);
    knn_scan_kernel<<<1, 1024, 0, s>>>(grid, counts, starts);
    e = cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (size_t)kMaxCells, s);
    if (e != cudaSuccess) return e;
    knn_fill_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, grid, starts, counts, sorted);
    knn_grid_search_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, grid, starts, sorted, out);
    return cudaGetLastError();
}


// grid workspace: aabb(32 B @0) + KnnGrid(@64) | counts[cells] | starts[cells+1] | sorted float4[V]
static size_t grid_ws_bytes(int n)
{
    return 256 + align_up(sizeof(uint32_t) * (size_t)kMaxCells) + align_up(sizeof(uint32_t) * ((size_t)kMaxCells + 1)) +
           align_up(sizeof(float4) * (size_t)n);
}
size_t vertex_grid_workspace_bytes(int V) { return V < 0 ? 0 : grid_ws_bytes(V); }
static size_t vertex_grid_smem_bytes(int V) { return sizeof(float4) * (size_t)V + sizeof(uint32_t) * (kVertexGridCells + 1); }
static bool vertex_grid_fits_smem(int V) { return vertex_grid_smem_bytes(V) <= 200 * 1024; }

static cudaError_t build_grid(int P, const float* points, char* ws, int max_cells, cudaStream_t s)
{
    float* aabb = reinterpret_cast<float*>(ws);
    KnnGrid* grid = reinterpret_cast<KnnGrid*>(ws + 64);
    uint32_t* counts = reinterpret_cast<uint32_t*>(ws + 256);
    uint32_t* starts = reinterpret_cast<uint32_t*>(ws + 256 + align_up(sizeof(uint32_t) * (size_t)kMaxCells));
    float4* sorted = reinterpret_cast<float4*>(reinterpret_cast<char*>(starts) + align_up(sizeof(uint32_t) * ((size_t)kMaxCells + 1)));
    const int init[6] = {0x7F7FFFFF, 0x7F7FFFFF, 0x7F7FFFFF, (int)0x80800000, (int)0x80800000, (int)0x80800000};
    cudaError_t e = cudaMemcpyAsync(aabb, init, sizeof(init), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
    knn_aabb_kernel<<<sm_count() * 4, 256, 0, s>>>(P, points, aabb);
    knn_grid_setup_kernel<<<1, 1, 0, s>>>(P, aabb, grid, max_cells);
    e = cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (size_t)max_cells, s);
    if (e != cudaSuccess) return e;
    knn_count_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, grid, counts);
    knn_scan_kernel<<<1, 1024, 0, s>>>(grid, counts, starts);
    e = cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (size_t)max_cells, s);
    if (e != cudaSuccess) return e;
    knn_fill_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, points, grid, starts, counts, sorted);
    count_launch(5);
    return cudaGetLastError();
}

cudaError_t launch_vertex_grid_build(int V, const float* verts, char* ws, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_KNN, s);
    return build_grid(V, verts, ws, vertex_grid_fits_smem(V) ? kVertexGridCells : kMaxCells, s);
}

cudaError_t launch_nearest_vertex(int P, const float* query, int V, const char* ws, int32_t* vid, cudaStream_t s)
{
    StageTimer timer(GHR_STAGE_KNN, s);
    const KnnGrid* grid = reinterpret_cast<const KnnGrid*>(ws + 64);
    const uint32_t* starts = reinterpret_cast<const uint32_t*>(ws + 256 + align_up(sizeof(uint32_t) * (size_t)kMaxCells));
    const float4* sorted = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(starts) + align_up(sizeof(uint32_t) * ((size_t)kMaxCells + 1)));
    count_launch();
    if (vertex_grid_fits_smem(V)) {
        static unsigned long long attr_set = 0ull;
        const size_t smem = vertex_grid_smem_bytes(V);
        if (first_use_on_current_device(attr_set)) {
            cudaError_t e = cudaFuncSetAttribute(nearest_vertex_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) return e;
        }
        const int need = (P + 127) / 128;  // 128 eight-lane groups per 1024-thread block
        const int blocks = need < sm_count() ? need : sm_count();
        nearest_vertex_smem_kernel<<<blocks, 1024, smem, s>>>(P, V, query, grid, starts, sorted, vid);
    } else {
        nearest_vertex_kernel<<<(P + 127) / 128, 128, 0, s>>>(P, query, grid, starts, sorted, vid);
    }
    return cudaGetLastError();
}

}  // namespace ghr
