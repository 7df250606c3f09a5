/*
 * ghr.h -- C ABI of the B200-native GauHuman render hot path (libghr.so).
 *
 * Plain pointers and sizes only: no torch types, no C++ types, never throws.
 * Every entry point returns 0 (GHR_OK) or a negative ghr_status; all work is
 * enqueued on the caller's cudaStream_t (passed as void*), nothing is
 * allocated behind the caller's back, no global state is kept between calls
 * (two documented exceptions: the process-wide switches ghr_set_shared_gpu /
 * ghr_profile_enable, and the static device scratch of ghr_sort_pairs_u64_static).
 *
 * Each entry point names the reference interface it replaces (paths relative
 * to /root/reference/).  `dgr/` = submodules/diff-gaussian-rasterization/.
 */
#ifndef GHR_H_INCLUDED
#define GHR_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GHR_VERSION 1
#define GHR_TILE 16 /* dgr/cuda_rasterizer/config.h:16-17 BLOCK_X/BLOCK_Y */

typedef enum ghr_status {
    GHR_OK = 0,
    GHR_ERR_BAD_ARG = -1,   /* null/misaligned pointer, negative size, bad shape */
    GHR_ERR_BUFFER = -2,    /* a state/workspace buffer is smaller than required */
    GHR_ERR_CUDA = -3,      /* a CUDA runtime call or (debug) a kernel failed    */
    GHR_ERR_CAPACITY = -4,  /* num_rendered exceeded the binning capacity         */
    GHR_ERR_PREFILTER = -5  /* prefiltered=1 but a Gaussian was culled            */
} ghr_status;

/*
 * Per-call rasteriser configuration + inputs.  Mirrors the 19 positional
 * arguments of _C.rasterize_gaussians (dgr/rasterize_points.h:19-39,
 * dgr/rasterize_points.cu:36-120) and the fields of
 * GaussianRasterizationSettings (dgr/diff_gaussian_rasterization/__init__.py:160-172).
 * Absent optional inputs are NULL (the reference passes 0-element tensors and
 * tests ptr==nullptr: forward.cu:205,241).  All pointers are device pointers
 * to contiguous fp32 unless noted.  Matrices are the reference's transposed
 * (column-major-as-read) 4x4s, dgr/cuda_rasterizer/auxiliary.h:58-77.
 */
typedef struct ghr_raster_params {
    int32_t P;                /* number of Gaussians                              */
    int32_t D;                /* active SH degree (0..3)                          */
    int32_t M;                /* SH coefficients per channel in `shs` (0 if none) */
    int32_t W, H;             /* image size in pixels                             */
    float scale_modifier;
    float tan_fovx, tan_fovy;
    int32_t prefiltered;
    int32_t debug;            /* 1: synchronise + check after every stage         */
    const float* background;  /* [3]                                              */
    const float* means3D;     /* [P,3]                                            */
    const float* shs;         /* [P,M,3] or NULL                                  */
    const float* colors_precomp; /* [P,3] or NULL                                 */
    const float* opacities;   /* [P]                                              */
    const float* scales;      /* [P,3] or NULL                                    */
    const float* rotations;   /* [P,4] or NULL (not normalised, forward.cu:127)   */
    const float* cov3D_precomp; /* [P,6] or NULL                                  */
    const float* viewmatrix;  /* [16]                                             */
    const float* projmatrix;  /* [16]                                             */
    const float* campos;      /* [3]                                              */
} ghr_raster_params;

/*
 * Byte offsets of the arrays inside the three opaque state buffers (the
 * reference's geomBuffer / binningBuffer / imgBuffer, carved by fromChunk,
 * dgr/cuda_rasterizer/rasterizer_impl.cu:155-194).  Layout is ours; exposed so
 * parity tests can read tiles_touched / point_offsets / keys / ranges /
 * n_contrib.  Sorted keys/ids live in buffer A or B depending on the device
 * word `sorted_sel` (0 = A, 1 = B).
 */
typedef struct ghr_layout {
    /* geom buffer */
    size_t geom_bytes;
    size_t g_header;        /* ghr_geom_header                                    */
    size_t g_rec;           /* float4[P][3]: {px,py,depth,_} {conic xyz,opacity} {rgb,_} */
    size_t g_tiles_touched; /* uint32[P]                                          */
    size_t g_point_offsets; /* uint32[P] inclusive scan                           */
    size_t g_cov3D;         /* float[P][6] (only when computed from scale/rot)    */
    size_t g_clamped;       /* uint8[P] bit c set = channel c clamped             */
    size_t g_scan_status;   /* uint64[ceil(P/256)+1] look-back words              */
    /* img buffer */
    size_t img_bytes;
    size_t i_final_T;       /* float[W*H]                                         */
    size_t i_n_contrib;     /* uint32[W*H]                                        */
    size_t i_ranges;        /* uint2[tiles]                                       */
    /* binning buffer */
    size_t binning_bytes;
    size_t b_header;        /* ghr_sort_header                                    */
    size_t b_keys_a;        /* uint64[cap]  (unsorted keys are emitted here)      */
    size_t b_keys_b;        /* uint64[cap]                                        */
    size_t b_vals_a;        /* uint32[cap]                                        */
    size_t b_vals_b;        /* uint32[cap]                                        */
    size_t b_status;        /* uint32[passes][tiles_per_pass][256] look-back      */
    int64_t capacity;       /* pairs the binning buffer holds                     */
    int32_t sort_bits;      /* 32 + getHigherMsb(tiles), rasterizer_impl.cu:302   */
    int32_t sort_passes;
    size_t i_order;         /* uint32[tiles]: tile ids by decreasing list length (blend work queue) */
} ghr_layout;

/* First bytes of the geom buffer (device memory). */
typedef struct ghr_geom_header {
    uint32_t num_rendered;  /* N = sum of tiles_touched (rasterizer_impl.cu:283)  */
    uint32_t error;         /* 0 or -(ghr_status) raised on the device            */
    uint32_t scan_ticket;   /* work counter of the fused preprocess+scan kernel   */
    uint32_t blend_fwd_ticket; /* work-queue heads of the two blend kernels        */
    uint32_t blend_bwd_ticket;
    uint32_t depth_hi;      /* max over the visible Gaussians of the depth's fp32 bit pattern (0 if none)          */
    uint32_t depth_lo_inv;  /* ~min of the same: the sort squeezes the constant high bits of (depth - min) out    */
    uint32_t reserved[57];
} ghr_geom_header;

int ghr_version(void);
const char* ghr_error_string(int status);

/* Sizes of the three state buffers.  `capacity` = max (Gaussian,tile) pairs
 * the binning buffer can hold (the reference sizes it exactly from a blocking
 * D2H read of num_rendered, rasterizer_impl.cu:283-287).                     */
int ghr_state_layout(int32_t P, int64_t capacity, int32_t W, int32_t H, ghr_layout* out);
/* The capacity (a multiple of 64) whose binning buffer is exactly `binning_bytes` long, or -1 if there is none.  The
 * reference's backward takes only the buffer and num_rendered (dgr/rasterize_points.cu:122-147); here the buffer may have
 * been sized from a guess larger than num_rendered, and its array offsets depend on that capacity, so the backward entry
 * points recover it from the size (binning_bytes is strictly increasing in the capacity) instead of trusting a side
 * table, and reject a buffer whose size does not match the capacity they are given (GHR_ERR_BUFFER). */
int64_t ghr_binning_capacity(int32_t P, int32_t W, int32_t H, size_t binning_bytes);

/*
 * Stage 1 of Rasterizer::forward (dgr/cuda_rasterizer/rasterizer_impl.cu:198-283):
 * preprocessCUDA (forward.cu:156-256) fused with the InclusiveSum
 * (rasterizer_impl.cu:279).  Writes radii[P] (int32), the packed geometry
 * records, tiles_touched, point_offsets and num_rendered (into the geom header
 * and, if non-NULL, into *num_rendered_host, which must be pinned/mapped host
 * memory the device can write).
 */
int ghr_forward_preprocess(const ghr_raster_params* prm, int32_t* radii, void* geom, size_t geom_bytes,
                           int32_t* num_rendered_host, void* stream);

/*
 * Stage 2 of Rasterizer::forward (rasterizer_impl.cu:285-341): duplicateWithKeys
 * (:70-111), the depth|tile radix sort (:305-310, here a hand-written onesweep),
 * identifyTileRanges (:116-138) and renderCUDA (forward.cu:261-383).
 * num_rendered is read on the device; if it exceeds `capacity` nothing is
 * rendered and GHR_ERR_CAPACITY is latched in the geom header.
 * out_color [3,H,W], out_depth [H,W], out_alpha [H,W].
 */
int ghr_forward_render(const ghr_raster_params* prm, void* geom, size_t geom_bytes, void* binning,
                       size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                       float* out_depth, float* out_alpha, void* stream);

/*
 * Rasterizer::backward (rasterizer_impl.cu:345-446): renderCUDA bwd
 * (backward.cu:399-587), computeCov2DCUDA (:144-274) and preprocessCUDA bwd
 * (:346-396) for _C.rasterize_gaussians_backward (dgr/rasterize_points.cu:122-207).
 * `accum` is a float[P*12] workspace the caller zeroes (the packed per-Gaussian
 * accumulator of the blend backward: dL_dcolor 0-2, dL_dmean2D 3-4, dL_dconic
 * 5-7, dL_dopacity 8).  Every output is fully written for all P Gaussians
 * (the reference instead zero-fills 75 floats per Gaussian first,
 * rasterize_points.cu:159-167).  dL_dmeans2D [P,3], dL_dcolors [P,3],
 * dL_dopacity [P], dL_dmeans3D [P,3], dL_dcov3D [P,6]; dL_dsh [P,M,3],
 * dL_dscales [P,3], dL_drotations [P,4] may be NULL when the matching input is
 * absent.
 */
int ghr_backward(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                 const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                 const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha, float* accum,
                 float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscales, float* dL_drotations, void* stream);

/* ---- multi-attribute rendering ------------------------------------------------------------------------------------
 * The reference's render() (gaussian_renderer/__init__.py:203-272) calls the whole rasteriser seven times per view on the
 * same Gaussians, changing only colors_precomp (image, normal, world normal, albedo, occlusion, roughness, axis).  Here
 * projection, key emission, sort, tile ranges and block masks are done once (ghr_forward_preprocess + ghr_forward_render
 * with the first attribute) and every further attribute only re-runs the blend on the same lists:
 *
 *   ghr_forward_reblend   writes `colors` ([P,3], used as given, like colors_precomp) into the geometry records and
 *                         blends again; out_depth / out_alpha receive the same values as in the first pass.
 *   ghr_backward_blend    blend backward of ONE attribute (its colours, its image gradient).  It ADDS into `accum`
 *                         ([P,12] floats, zeroed by the caller before the first attribute): slots 3-8 (mean2D, conic,
 *                         opacity) sum over the attributes; slots 0-2 hold this attribute's dL_dcolors, which the
 *                         caller copies out and clears before the next attribute.
 *   ghr_backward_finish   the per-Gaussian backward (reference backward.cu:144-396) from the summed record.
 * dL_dout_depth / dL_dout_alpha may be NULL (= zero) in ghr_backward_blend. */
int ghr_forward_reblend(const ghr_raster_params* prm, const float* colors, void* geom, size_t geom_bytes,
                        const void* binning, size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes,
                        float* out_color, float* out_depth, float* out_alpha, void* stream);
int ghr_backward_blend(const ghr_raster_params* prm, const float* colors, void* geom, size_t geom_bytes,
                       const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                       const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
                       const float* zeros_plane /* [H*W] zeros, used when a gradient plane is NULL */, float* accum,
                       void* stream);
int ghr_backward_finish(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                        int64_t capacity, const float* accum, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                        float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
                        void* stream);

/* Fused form of the above, for 2, 4 or 6 extra attributes at a time: ONE blend kernel walks the tile lists once and
 * accumulates the main colour, depth, alpha AND the attribute channels (forward), and one backward kernel produces the
 * gradients of all of them.  Attribute colours travel packed, ghr_attribute_stride(n) floats per Gaussian:
 *
 *   ghr_attribute_stride       floats per Gaussian of a packed row for n attributes; 0 = n is not fused (use reblend)
 *   ghr_attributes_fused       1 if the fused kernels serve a frame of this binning capacity, 0 = use reblend (frames
 *                              with millions of (Gaussian, tile) pairs run the throughput kernel family, which has no
 *                              fused variant)
 *   ghr_pack_attributes        colors[k] ([P,3], k < n) -> packed [P][stride]
 *   ghr_forward_render_multi   ghr_forward_render + out_images [3n][H][W] (attribute k = planes 3k..3k+2); every image
 *                              equals a separate single-attribute pass bit for bit
 *   ghr_backward_multi         ghr_backward with dL_dout_images [3n][H][W]; packed_grad [P][stride] receives the colour
 *                              gradients of the attributes (zeroed here)
 *   ghr_unpack_attribute_grads packed_grad -> dL_dcolors[k] ([P,3]; NULL entries are skipped)
 * `colors` / `dL_dcolors` are HOST arrays of n device pointers. */
/* The sorted list of Gaussian ids in the binning buffer (reference: point_list) is uint32 per pair.  In frames with more
 * than 2^22 pairs and fewer than 2^24 Gaussians the upper 8 bits of each value carry the mask of the tile's 8x4-pixel
 * blocks the Gaussian can touch (written by the key emission, read by the blend kernels): AND the values with this
 * mask to get the ids (0xFFFFFFFF when nothing is packed). */
uint32_t ghr_point_list_id_mask(int32_t P, int64_t capacity);

int ghr_attribute_stride(int n_attr);
int ghr_attributes_fused(int64_t capacity);
int ghr_pack_attributes(int32_t P, int n_attr, const float* const* colors, float* packed, void* stream);
int ghr_unpack_attribute_grads(int32_t P, int n_attr, const float* packed_grad, float* const* dL_dcolors, void* stream);
int ghr_forward_render_multi(const ghr_raster_params* prm, void* geom, size_t geom_bytes, void* binning,
                             size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                             float* out_depth, float* out_alpha, int n_attr, const float* packed, float* out_images,
                             void* stream);
int ghr_backward_multi(const ghr_raster_params* prm, const int32_t* radii, const void* geom, size_t geom_bytes,
                       const void* binning, size_t binning_bytes, int64_t capacity, const void* img, size_t img_bytes,
                       const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha, int n_attr,
                       const float* packed, const float* dL_dout_images, float* packed_grad, float* accum,
                       float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                       float* dL_dsh, float* dL_dscales, float* dL_drotations, void* stream);

/* Rasterizer::markVisible / checkFrustum (rasterizer_impl.cu:54-66,141-153)
 * for _C.mark_visible (dgr/rasterize_points.cu:209-228).  present: uint8[P]. */
int ghr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/*
 * The sort of rasterizer_impl.cu:305-310 in isolation (bench hook, Mkeys/s):
 * stable LSD radix sort of N (uint64 key, uint32 value) pairs on key bits
 * [0,bits).  keys_a/vals_a hold the input; the result is in A or B as
 * reported through *sorted_sel_dev (device uint32).  workspace from
 * ghr_sort_workspace_bytes.
 */
size_t ghr_sort_workspace_bytes(int64_t N);
int ghr_sort_pairs_u64(int64_t N, int32_t bits, uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a,
                       uint32_t* vals_b, void* workspace, size_t workspace_bytes, uint32_t* sorted_sel_dev,
                       void* stream);

/*
 * ghr_sort_pairs_u64 for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO
 * COUNTERPART IN THE REFERENCE.
 * Same kernels, arguments and result as ghr_sort_pairs_u64 minus the
 * workspace: the histogram / look-back scratch is a file-scope static
 * __device__ array inside libghr.so (ghr_noalloc.cu, ~4.2 MB, held by every
 * GPU that loads the library until exit).  Limit: N <=
 * GHR_STATIC_SORT_MAX_KEYS, beyond it GHR_ERR_CAPACITY and a message on
 * stderr.  Calls no cudaMalloc / cudaFree / allocating routine.  All calls
 * share the one array (cleared on `stream` at the start of each call): calls
 * that overlap on the GPU race -- serialise them.
 */
#define GHR_STATIC_SORT_MAX_KEYS (1 << 20)
int64_t ghr_sort_static_max_keys(void);
int ghr_sort_pairs_u64_static(int64_t N, int32_t bits, uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a,
                              uint32_t* vals_b, uint32_t* sorted_sel_dev, void* stream);

/*
 * distCUDA2 (submodules/simple-knn/spatial.cu:15-26 -> SimpleKNN::knn,
 * simple_knn.cu:185-221): out[i] = mean of the 3 smallest squared distances
 * from point i to the other points.
 */
size_t ghr_knn_workspace_bytes(int32_t P);
int ghr_dist2_knn3(int32_t P, const float* points, float* out, void* workspace, size_t workspace_bytes,
                   void* stream);

/* ======================================================================== *
 *  LBS deformation of SMPL-initialised Gaussians
 *  (reference scene/gaussian_model.py:768-872 coarse_deform_c2source,
 *   :894-980 batch_rodrigues_torch / get_rigid_transformation_torch /
 *   get_transform_params_torch, :35-42 build_covariance_from_scaling_rotation,
 *   utils/sh_utils.py:135-190 eval_sh, smplx/lbs.py:156-405)
 * ======================================================================== */
#define GHR_LBS_JOINTS 24
#define GHR_LBS_POSE_DIM 207 /* (24-1)*9 pose-feature entries */

/* Body model in device memory (keys of SMPL_to_tensor, gaussian_model.py:880-892). */
typedef struct ghr_lbs_model {
    int32_t V;                    /* vertices (6890)                                    */
    int32_t NB;                   /* shape coefficients used (10)                       */
    const float* weights;         /* [V,24]   skinning weights                          */
    const float* posedirs;        /* [V*3,207]                                          */
    const float* shapedirs;       /* [V,3,NB]                                           */
    const float* joint_template;  /* [24,3]    J_regressor . v_template   (ghr_lbs_prepare) */
    const float* joint_shapedirs; /* [24,3,NB] J_regressor . shapedirs    (ghr_lbs_prepare) */
    const int32_t* parents;       /* [24] kintree_table[0]; parents[0] ignored           */
} ghr_lbs_model;

/* Layout (float offsets) of a "pose buffer", the output of ghr_lbs_pose. */
#define GHR_POSE_A 0        /* [24][12] rows 0-2 of A_j (rigid transform relative to the rest pose) */
#define GHR_POSE_ROT 288    /* [24][9]  joint rotations (after correct_Rs)               */
#define GHR_POSE_PF 504     /* [207]    pose feature (rot[1:] - I)                       */
#define GHR_POSE_JOINTS 720 /* [24][3]  rest joints                                      */
#define GHR_POSE_RGINV 800  /* [9]      inverse of the global rotation params['R']       */
#define GHR_POSE_RG 816     /* [9]      params['R']                                      */
#define GHR_POSE_TH 832     /* [3]      params['Th']                                     */
#define GHR_POSE_G 848      /* [24][12] world chain G_j (before removing the rest joint) */
#define GHR_POSE_ROD 1136   /* [24][9]  Rodrigues rotations before correct_Rs            */
#define GHR_POSE_FLOATS 1536

/* One-time model preparation: because joints = J_regressor (v_template + shapedirs beta)
 * is linear in beta (gaussian_model.py:951-970), the regressor is folded into the template
 * and shape directions once.  J_regressor [24,V], v_template [V,3]. */
int ghr_lbs_prepare(int32_t V, int32_t NB, const float* v_template, const float* shapedirs, const float* J_regressor,
                    float* joint_template, float* joint_shapedirs, void* stream);

/* get_transform_params_torch (gaussian_model.py:947-980): Rodrigues (:894-912), optional
 * correct_Rs right-multiplication (:833-836,964-967), rest joints, the 24-joint chain
 * (:914-944), pose feature (:811,838) and inverse of the global R (:863).
 * poses [72], shapes [NB], correct_Rs [23,3,3] or NULL, R [3,3], Th [3] (all device). */
int ghr_lbs_pose(const ghr_lbs_model* model, const float* poses, const float* shapes, const float* correct_Rs,
                 const float* R, const float* Th, float* pose_out, void* stream);

/* Per-vertex blend-shape displacement applied between the two skinning steps
 * (gaussian_model.py:805-844): off[v] = shapedirs[v].beta + posedirs[v].(pf_target - pf_big). */
int ghr_lbs_vertex_offsets(const ghr_lbs_model* model, const float* pose_big, const float* pose_target,
                           const float* shapes, float* offsets /*[V,3]*/, void* stream);

/* Nearest SMPL vertex (the reference's knn_cuda KNN(k=1), gaussian_model.py:775).
 * build: bins t_vertices [V,3] into a uniform grid inside `workspace`;
 * query: exact nearest vertex (lowest index on ties) for P points -> vert_ids [P] int32; V as built. */
/* The pose stage (ghr_lbs_pose + ghr_lbs_vertex_offsets) of the n_views <= GHR_POSE_VIEWS_MAX views of a multi-view
 * batch in ONE launch: the 17 MB pose basis (posedirs) is read once for the batch instead of once per view.  `views` is a
 * HOST array; per view the results equal those of the two per-view entry points bit for bit.  Reference: the same lines
 * as ghr_lbs_pose / ghr_lbs_vertex_offsets (the reference runs them per iteration, batch size 1). */
#define GHR_POSE_VIEWS_MAX 8
typedef struct ghr_pose_view {
    const float* poses;       /* [72]                                    */
    const float* shapes;      /* [NB]                                    */
    const float* correct_Rs;  /* [23,3,3] or NULL                        */
    const float* R;           /* [9]                                     */
    const float* Th;          /* [3]                                     */
    float* pose_out;          /* [GHR_POSE_FLOATS] pose buffer           */
    float* offsets_out;       /* [V,3] vertex offsets against pose_big   */
} ghr_pose_view;
int ghr_lbs_pose_views(const ghr_lbs_model* model, int32_t n_views, const ghr_pose_view* views, const float* pose_big,
                       void* stream);

size_t ghr_vertex_grid_workspace_bytes(int32_t V);
int ghr_vertex_grid_build(int32_t V, const float* t_vertices, void* workspace, size_t workspace_bytes, void* stream);
int ghr_nearest_vertex(int32_t P, const float* query, int32_t V, const void* workspace, int32_t* vert_ids,
                       void* stream);

/* Fused per-Gaussian deformation: skinning-weight blend (optionally
 * softmax(log(w+1e-9)+offset), :776-781), big pose -> T pose (:785-799), blend-shape
 * offsets, T pose -> target pose (:847-860), SMPL -> world (:863-870); then, if
 * requested, the world-space covariance T Sigma T^T (gaussian_model.py:35-42,
 * general_utils.py:60-117) and the view-dependent colour clamp_min(eval_sh+0.5,0)
 * (gaussian_renderer/__init__.py:124-125,191-195).  NULL outputs are skipped. */
typedef struct ghr_deform_params {
    int32_t P;
    int32_t V;
    int32_t sh_degree;           /* active degree 0..3                                  */
    int32_t M;                   /* SH coefficients per channel in shs (0: no colours)  */
    float scale_modifier;
    const float* xyz;            /* [P,3] canonical (big-pose) means                    */
    const float* normals;        /* [P,3] or NULL                                       */
    const int32_t* vert_ids;     /* [P]                                                 */
    const float* weights;        /* [V,24] model skinning weights                       */
    const float* lbs_offsets;    /* logit offsets, element (p,j) at p*stride_p + j*stride_j, or NULL */
    int64_t lbs_stride_p, lbs_stride_j;
    const float* vertex_offsets; /* [V,3] from ghr_lbs_vertex_offsets                   */
    const float* pose_big;       /* pose buffer of the big pose                         */
    const float* pose_target;    /* pose buffer of the target pose                      */
    const float* scales;         /* [P,3] activated scales or NULL (no covariance)      */
    const float* rotations;      /* [P,4] raw quaternions (normalised inside)           */
    const float* shs;            /* [P,M,3] or NULL                                     */
    const float* campos;         /* [3]                                                 */
    /* Optional fused activations (reference get_scaling = exp, get_opacity = sigmoid, gaussian_model.py:152-175).
     * All zero / NULL = plain behaviour. */
    int32_t scales_are_log;      /* 1: `scales` holds log-scales; exp() is applied inside and dL_dscales is returned
                                    with respect to the log-scales                                                  */
    const float* opacity_logits; /* [P] or NULL                                                                     */
    float* opacity_out;          /* forward: [P] sigmoid(opacity_logits)                                            */
    const float* dL_dopacity;    /* backward: [P] incoming gradient of opacity_out, or NULL (= zero)                */
    float* dL_dopacity_logits;   /* backward: [P] written when opacity_logits is given                              */
    int32_t accumulate_grads;    /* backward: 1 = the per-Gaussian outputs (dL_dxyz, dL_dnormals, dL_dlbs_offsets,
                                    dL_dscales, dL_drotations, dL_dshs, dL_dopacity_logits) are ADDED to the buffers
                                    (gradient accumulation over a batch of views) instead of stored; 2 = added with
                                    atomics (views of the batch running concurrently on several streams)          */
} ghr_deform_params;

int ghr_lbs_deform_forward(const ghr_deform_params* prm, float* smpl_pts /*[P,3]|NULL*/, float* world_pts /*[P,3]*/,
                           float* bweights /*[P,24]|NULL*/, float* transforms /*[P,3,3]*/,
                           float* translation /*[P,3]|NULL*/, float* world_normals /*[P,3]|NULL*/,
                           float* cov3D /*[P,6]|NULL*/, float* colors /*[P,3]|NULL*/, uint8_t* clamped /*[P]|NULL*/,
                           void* stream);

/* Backward of ghr_lbs_deform_forward.  Incoming gradients may be NULL (= zero).
 * dL_dA_target [24*12] and dL_dvertex_offsets [V*3] are ACCUMULATED (caller zeroes
 * them) and may be NULL when the pose branch needs no gradient. */
int ghr_lbs_deform_backward(const ghr_deform_params* prm, const uint8_t* clamped, const float* dL_dworld_pts,
                            const float* dL_dtransforms, const float* dL_dworld_normals, const float* dL_dcov3D,
                            const float* dL_dcolors, float* dL_dxyz, float* dL_dnormals,
                            float* dL_dlbs_offsets /*[P,24]*/, float* dL_dscales, float* dL_drotations,
                            float* dL_dshs, float* dL_dA_target, float* dL_dvertex_offsets, void* stream);

/*
 * The per-Gaussian chain of a view as one kernel each way ("fused LBS + projection + SH"): what render() does between
 * reading the canonical Gaussians and calling the rasteriser (gaussian_renderer/__init__.py:94-195:
 * coarse_deform_c2source, get_covariance, eval_sh + clamp, get_opacity) fused with the rasteriser's own per-Gaussian
 * stage (forward.cu:156-256 preprocessCUDA; backward.cu:144-396).  The deformed mean, world covariance, colour and
 * opacity stay in registers; results are those of ghr_lbs_deform_forward -> ghr_forward_preprocess (radii, tile rects and
 * sort keys bit for bit) and of ghr_backward -> ghr_lbs_deform_backward.
 *
 *   ghr_forward_deform_preprocess   dp: scales, rotations and shs are required; opacity from dp->opacity_logits
 *                                   (sigmoid) or view->opacities.  view: P (= dp->P), W, H, tan_fov*, background,
 *                                   viewmatrix, projmatrix, campos, prefiltered, debug -- its per-Gaussian pointers are
 *                                   not read.  Writes radii and the geometry state like ghr_forward_preprocess.
 *                                   world_pts [P,3], world_normals [P,3], cov3D [P,6], colors [P,3] (zero for culled
 *                                   Gaussians), opacity [P]: optional copies of the intermediate values, NULL = skipped.
 *   ghr_forward_render_deformed     ghr_forward_render for such a view (no per-Gaussian input pointer is checked).
 *   ghr_backward_deformed           blend backward + the fused per-Gaussian backward.  Outputs as
 *                                   ghr_lbs_deform_backward (same accumulate_grads modes, dL_dopacity_logits through
 *                                   dp), plus dL_dmeans2D [P,3] (screen-space gradient, always stored; NULL = skipped).
 *                                   Culled Gaussians get zeros (nothing is added in the accumulating modes).
 *                                   max_radii2D / xyz_gradient_accum / denom: the densification statistics of the
 *                                   training loop (ghr_densify_stats: train.py:401-404, scene/gaussian_model.py:764-766)
 *                                   updated in the same kernel for the visible Gaussians, with atomics (views of a batch
 *                                   may run concurrently); NULL = skipped.
 */
int ghr_forward_deform_preprocess(const ghr_deform_params* dp, const ghr_raster_params* view, int32_t* radii, void* geom,
                                  size_t geom_bytes, float* world_pts, float* world_normals, float* cov3D, float* colors,
                                  float* opacity, int32_t* num_rendered_host, void* stream);
int ghr_forward_render_deformed(const ghr_raster_params* view, void* geom, size_t geom_bytes, void* binning,
                                size_t binning_bytes, int64_t capacity, void* img, size_t img_bytes, float* out_color,
                                float* out_depth, float* out_alpha, void* stream);
int ghr_backward_deformed(const ghr_deform_params* dp, const ghr_raster_params* view, const int32_t* radii,
                          const void* geom, size_t geom_bytes, const void* binning, size_t binning_bytes, int64_t capacity,
                          const void* img, size_t img_bytes, const float* dL_dout_color, const float* dL_dout_depth,
                          const float* dL_dout_alpha, float* accum /* [P,12] zeroed */, float* dL_dmeans2D,
                          float* dL_dxyz, float* dL_dlbs_offsets /* [P,24]|NULL */, float* dL_dscales,
                          float* dL_drotations, float* dL_dshs, float* dL_dA_target /* [24*12]|NULL, accumulated */,
                          float* dL_dvertex_offsets /* [V*3]|NULL, accumulated */,
                          float* max_radii2D, float* xyz_gradient_accum, float* denom /* each [P]|NULL, see below */,
                          void* stream);

/* BodyPoseRefiner (nets/mlp_delta_body_pose.py:5-41 + RodriguesModule :164-193; called per view from
 * gaussian_renderer/__init__.py:100-102): Linear(69, width) ReLU [Linear(width, width) ReLU]* Linear(width, 69) on the 69
 * pose values of a view, then Rodrigues with theta = sqrt(1e-5 + |r|^2) -> correct_Rs [23,3,3] (the `correct_Rs` input of
 * ghr_lbs_pose / ghr_lbs_pose_views).  n_layers = number of Linear layers (2..GHR_REFINER_MAX_LAYERS), weights[i]
 * row-major [out,in], fp32 throughout.  One launch. */
#define GHR_REFINER_MAX_LAYERS 6
#define GHR_REFINER_MAX_WIDTH 256
int ghr_pose_refiner_forward(int32_t width, int32_t n_layers, const float* const* weights, const float* const* biases,
                             const float* pose69, float* correct_Rs, void* stream);

/*
 * LBSOffsetDecoder (nets/mlp_delta_weight_lbs.py:5-33; called per view from gaussian_renderer/__init__.py:105): positional
 * encoding (10 octaves, 63 values) -> Conv1d 63->128, 128->128, 128->128, [63+128]->128 (ReLU each) -> Conv1d 128->24, for
 * every Gaussian; its output is the `lbs_offsets` input of the deform entry points.  One kernel on the tcgen05 tensor
 * cores (TF32 operands, fp32 accumulation in TMEM -- the precision of the reference's own Conv1d layers, cuDNN allows
 * TF32 by default), weights streamed by cp.async.bulk from a prepacked blob.
 *   ghr_lbs_mlp_pack     weights[5] = the Conv1d weights [out,in(,1)] of bw_linears[0..3] and bw_fc ([128,63] [128,128]
 *                        [128,128] [128,191] [24,128]), biases[5] likewise -> blob (ghr_lbs_mlp_blob_bytes(), 16-byte
 *                        aligned).  Repack after every optimiser step that touches the network.
 *   ghr_lbs_mlp_forward  xyz [P,3] -> logits [P,24] (row-major, 16-byte aligned: `lbs_offsets` with strides 24, 1).
 */
size_t ghr_lbs_mlp_blob_bytes(void);
int ghr_lbs_mlp_pack(const float* const* weights, const float* const* biases, void* blob, void* stream);
int ghr_lbs_mlp_forward(int32_t P, const float* xyz, const void* blob, float* logits, void* stream);

/* Pose-branch backward: dL/dA_target [24,12] and dL/dvertex_offsets [V,3] (or NULL) -> dL/dcorrect_Rs
 * [23,3,3] (the only learnable input of the pose branch, nets/mlp_delta_body_pose.py).
 * dL_dcorrect_Rs must hold 23*9 + 207 floats: the tail is scratch for the pose-feature gradient. */
int ghr_lbs_pose_backward(const ghr_lbs_model* model, const float* pose_target, const float* correct_Rs,
                          const float* dL_dA_target, const float* dL_dvertex_offsets, float* dL_dcorrect_Rs,
                          void* stream);

/* smplx/lbs.py:156-252 lbs() for batch size 1 over the model's own V vertices
 * (BASELINE.json config 1): v = T (v_template + shapedirs beta + posedirs pf), T = W A. */
int ghr_smpl_lbs_vertices(const ghr_lbs_model* model, const float* v_template, const float* pose_buf,
                          const float* vertex_offsets /*[V,3] from ghr_lbs_vertex_offsets(pose_big=NULL)*/,
                          float* verts /*[V,3]*/, float* T_out /*[V,12]|NULL*/, void* stream);

/*
 * Stand-alone forms of two operators that ghr_lbs_deform_forward/_backward also compute fused,
 * for callers that use them on their own.
 *  - covariance: GaussianModel.get_covariance(scaling_modifier, transforms) (scene/gaussian_model.py:35-42,208-209;
 *    utils/general_utils.py:60-117): q^ = q/|q|, L = R(q^) diag(mod*s), Sigma = L L^T, Sigma' = T Sigma T^T, packed
 *    [xx,xy,xz,yy,yz,zz].  scaling [P,3] is the ACTIVATED scaling (get_scaling), rotation [P,4] the raw quaternion,
 *    transforms [P,3,3] or NULL (identity).  backward writes dL_dscaling [P,3], dL_drotation [P,4] and, if non-NULL,
 *    dL_dtransforms [P,3,3].
 *  - sh_colors: clamp_min(eval_sh(deg, shs, (means3D - campos)/|.|) + 0.5, 0) (gaussian_renderer/__init__.py:124-125,
 *    191-195; utils/sh_utils.py:135-190).  shs [P,M,3] (get_features layout), clamped [P] uint8 (bit c = channel c was
 *    clamped; forward output, backward input, may be NULL).  backward writes dL_dshs [P,M,3] and, if non-NULL,
 *    dL_dmeans3D [P,3] (through the view direction).
 */
int ghr_covariance_forward(int32_t P, const float* scaling, const float* rotation, float scaling_modifier,
                           const float* transforms, float* cov3D, void* stream);
int ghr_covariance_backward(int32_t P, const float* scaling, const float* rotation, float scaling_modifier,
                            const float* transforms, const float* dL_dcov3D, float* dL_dscaling, float* dL_drotation,
                            float* dL_dtransforms, void* stream);
int ghr_sh_colors_forward(int32_t P, int32_t sh_degree, int32_t M, const float* shs, const float* means3D,
                          const float* campos, float* colors, uint8_t* clamped, void* stream);
int ghr_sh_colors_backward(int32_t P, int32_t sh_degree, int32_t M, const float* shs, const float* means3D,
                           const float* campos, const uint8_t* clamped, const float* dL_dcolors, float* dL_dshs,
                           float* dL_dmeans3D, void* stream);

/*
 * View-batched forward rasterisation: n_views small views of the SAME Gaussians in one set of
 * launches -- what baking.py::bake_set (:136-309) does with one _C.rasterize_gaussians call per cube-map face
 * (six 32x32 faces per occupied grid cell, thousands of calls).  prm carries the shared inputs (its viewmatrix /
 * projmatrix / campos are ignored); cameras is a device array of n_views rows of 36 floats
 * [viewmatrix 16 | projmatrix 16 | campos 3 | pad 1] in the reference's transposed convention.  If labels [P] and
 * skip_labels [n_views] are given, view v leaves out every Gaussian g with labels[g] == skip_labels[v] (the
 * reference gathers means3D[pc_grid_indices != grid_id] per cell, baking.py:237-246; the images are the same).
 * Outputs: out_alpha [n_views,1,H,W]; optional out_color [n_views,3,H,W], out_depth [n_views,1,H,W] (needs
 * out_color), radii [n_views,P].  Per view they equal ghr_forward_preprocess + ghr_forward_render on that view.
 * The workspace (ghr_views_workspace_bytes, 256-byte aligned) begins with a ghr_geom_header: after the stream has
 * drained, num_rendered is the total number of (view, Gaussian, tile) pairs and error is -GHR_ERR_CAPACITY if that
 * exceeded `capacity` (outputs are then not written: call again with a larger capacity).
 */
size_t ghr_views_workspace_bytes(int32_t P, int32_t n_views, int32_t W, int32_t H, int64_t capacity);
int ghr_forward_views(const ghr_raster_params* prm, int32_t n_views, const float* cameras, const int32_t* labels,
                      const int32_t* skip_labels, void* workspace, size_t workspace_bytes, int64_t capacity,
                      float* out_color, float* out_depth, float* out_alpha, int32_t* radii, void* stream);

/*
 * Cube-map -> direction lookup of the occlusion baking (reference baking.py:286-309).
 *
 * ghr_cubemap_sample_nearest: what bake_set asks of nvdiffrast (un-vendored third party),
 *   dr.texture(tex[1,6,res,res,C], uv = dirs[1,H,W,3], filter_mode="nearest", boundary_mode="cube") (:290-297),
 * for n_maps cube maps at once: out[m][d][c] = tex[m][face(d)][v(d)][u(d)][c].  Cube addressing is the OpenGL one
 * (major axis -> face +x,-x,+y,-y,+z,-z; in-face coordinates per the OpenGL 4.6 specification, table 8.19), nearest
 * texel = floor(coordinate * res) clamped to [0, res-1]; a direction without a major axis (NaN, zero) reads 0.
 * tex [n_maps,6,res,res,channels], dirs [n_dirs,3] (need not be normalised), out [n_maps,n_dirs,channels].
 *
 * ghr_bake_occlusion: the rest of bake_set fused with that lookup (:231,305-307) for single-channel alpha cube maps:
 *   occlusion[p][d] = dot(dirs[d], normals[p]) > 0 ? 1 - cubemaps[cell_of[p]][texel(d)] : 0
 * cubemaps [n_cells,6,res,res] (the opacity maps ghr_forward_views renders per grid cell), cell_of [P] int64
 * (pc_grid_indices of pc_to_grid, :104-134), normals [P,3], dirs [n_dirs,3], out [P,n_dirs]  (n_dirs <= 8192).
 */
int ghr_cubemap_sample_nearest(int32_t n_maps, int32_t res, int32_t channels, const float* tex, int32_t n_dirs,
                               const float* dirs, float* out, void* stream);
int ghr_bake_occlusion(int32_t P, int32_t n_dirs, int32_t n_cells, int32_t res, const float* cubemaps,
                       const int64_t* cell_of, const float* normals, const float* dirs, float* out, void* stream);

/*
 * Optimiser side of the training iteration (no native counterpart in the reference, which runs
 * these as chains of ATen kernels).
 *
 * ghr_adam_step: torch.optim.Adam(l, lr=0.0, eps=1e-15).step() of scene/gaussian_model.py:283 / train.py:418
 * (no weight decay, no amsgrad, maximize=False) for up to GHR_ADAM_MAX_GROUPS parameter tensors in ONE launch:
 *   exp_avg += (1-beta1)(grad - exp_avg); exp_avg_sq = beta2 exp_avg_sq + (1-beta2) grad^2;
 *   param  -= lr/(1-beta1^step) * exp_avg / (sqrt(exp_avg_sq)/sqrt(1-beta2^step) + eps)
 * `step` is the group's 1-based step count after this step (torch keeps state['step'] per parameter; tensors
 * re-created by densification carry theirs along, gaussian_model.py:423-483).  The groups table is host memory, the
 * pointers in it device memory, fp32.  zero_grads != 0 also clears the gradients in
 * the same pass (optimizer.zero_grad, train.py:419).  28 (32) bytes of HBM traffic per parameter.
 */
#define GHR_ADAM_MAX_GROUPS 16
typedef struct ghr_adam_group {
    float* param;
    float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t count;  /* floats */
    double lr;      /* param_group['lr'] of this step (update_learning_rate, gaussian_model.py:290-295) */
    int32_t step;   /* >= 1 */
    int32_t pad;
} ghr_adam_group;
int ghr_adam_step(int n_groups, const ghr_adam_group* groups, double beta1, double beta2, double eps, int zero_grads,
                  void* stream); /* betas / eps as the Python doubles torch holds: 1 - beta is rounded to fp32 once */

/*
 * View-sharded data parallelism, optimiser step fused with its collective (no counterpart in the reference, which is
 * single-GPU): reduce-scatter of the ranks' flat gradient buckets + Adam on the owned slice + all-gather of the new
 * parameters, in ONE kernel over NVLink peer memory.  peer_grads[r] / peer_params[r] (host arrays of `world` device
 * pointers, r = rank order, entry `rank` = this process's own buffers) are the ranks' flat gradient buckets and flat
 * parameter replicas, mapped into this process (CUDA IPC; mygauhuman_b200/optim.py ShardedFusedAdam does the exchange).
 * This rank owns floats [shard_lo, shard_hi) (multiples of 4) and holds the moments of that slice only
 * (exp_avg_shard / exp_avg_sq_shard, shard_hi - shard_lo floats, 16-byte aligned).  groups: the parameter groups as
 * (offset, count) ranges of the flat layout (offsets multiples of 4) with their lr and 1-based step.  The caller
 * brackets the call with cross-rank barriers on the stream (all buckets complete before; no bucket reset or parameter
 * read before every rank is done).  Element arithmetic = ghr_adam_step on the rank-ordered sum of the gradients.
 */
#define GHR_MAX_PEERS 8
typedef struct ghr_shard_group {
    int64_t offset;
    int64_t count;
    double lr;
    int32_t step;
    int32_t pad;
} ghr_shard_group;
int ghr_adam_step_sharded(int world, int rank, const float* const* peer_grads, float* const* peer_params, float* exp_avg_shard,
                          float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi, int n_groups,
                          const ghr_shard_group* groups, double beta1, double beta2, double eps, void* stream);
/* Same step with the two cross-rank barriers INSIDE the kernel (no NCCL call at all): peer_flags[r] = rank r's flag buffer
 * (ghr_sharded_flag_words() zero-initialised uint32 words, mapped like the other buffers), epoch = 1, 2, 3, ... (the step
 * count, the same on every rank).  At its start the kernel publishes "rank arrived" to every peer with a system-scope
 * release store and waits for all peers; at its end the last CTA publishes "rank done" and waits for all peers, so
 * kernel completion means every replica is written and every bucket has been read.  Every rank must make the call
 * (also one whose slice is empty). */
int ghr_adam_step_sharded_synced(int world, int rank, const float* const* peer_grads, float* const* peer_params,
                                 float* exp_avg_shard, float* exp_avg_sq_shard, int64_t shard_lo, int64_t shard_hi,
                                 int n_groups, const ghr_shard_group* groups, double beta1, double beta2, double eps,
                                 uint32_t* const* peer_flags, uint32_t epoch, void* stream);
int ghr_sharded_flag_words(void);
/* The in-kernel waits are BOUNDED (GHR_SHARDED_TIMEOUT_MS, default 20 000 ms): a rank whose peer never arrives (its host
 * raised before the launch, it skipped a step, its epoch differs) gives up, skips its update, latches the reason in word
 * ghr_sharded_error_word() of ITS flag buffer (1 = timed out waiting for the buckets, 2 = timed out waiting for the
 * replicas, 3 = aborted by a peer) and raises the abort word of every peer, so that the whole node stops at once.  The
 * host reads the error word at its next synchronisation point (ShardedFusedAdam.check()).  ghr_sharded_abort raises the
 * abort word of every rank from the host side (enqueued on `stream`): call it when a step cannot be launched. */
int ghr_sharded_error_word(void);
int ghr_sharded_abort(int world, uint32_t* const* peer_flags, void* stream);
/* Lets kernels on the CURRENT device dereference memory of `peer_device` (cudaDeviceEnablePeerAccess; already enabled
 * is not an error).  Needed once per peer before ghr_adam_step_sharded. */
int ghr_enable_peer_access(int peer_device);
/* Maps another process's allocation (64-byte cudaIpcMemHandle_t, e.g. the one torch's storage._share_cuda_() returns)
 * into THIS process under the CURRENT device, with peer access enabled as needed; *base_out is the allocation's base.
 * One open per handle and process.  ghr_ipc_close unmaps. */
int ghr_ipc_open(const void* handle, void** base_out);
int ghr_ipc_close(void* base);

/* train.py:401-404 + GaussianModel.add_densification_stats (scene/gaussian_model.py:764-766) with
 * visibility_filter = radii > 0, fused: max_radii2D[i] = max(max_radii2D[i], radii[i]);
 * xyz_gradient_accum[i] += ||dL_dmean2D[i, :2]||; denom[i] += 1.  Any of the three outputs may be NULL.
 * grad_stride = floats per row of dL_dmean2D (3 for the [P,3] viewspace tensor). */
int ghr_densify_stats(int32_t P, const int32_t* radii, const float* dL_dmean2D, int32_t grad_stride, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, void* stream);

/*
 * PLY checkpoint rows: the interleaved float rows of the file GaussianModel.save_ply writes and
 * load_ply reads (scene/gaussian_model.py:309-407), packed / unpacked on the device in one pass.  Row =
 * x y z | nx ny nz | ar ag ab | roughness | f_dc_0..2 | f_rest_0..3(M-1)-1 (channel-major, :330) | opacity |
 * scale_0..2 | rot_0..3 = ghr_ply_row_floats(M) = 18 + 3M floats; M = SH coefficients per channel incl. DC (16).
 * Tensors: xyz [P,3], normal [P,3], albedo [P,3], roughness [P,1], f_dc [P,1,3], f_rest [P,M-1,3], opacity [P,1],
 * scaling [P,3], rotation [P,4] (the reference's raw parameters); normal / albedo / roughness may be NULL (written
 * as zeros, skipped on unpack).  unpack: rows may hold row_floats >= 18+3M columns in any order; columns[c] (host
 * array of 18+3M ints, NULL = identity) is the file column of canonical column c.
 */
int ghr_ply_row_floats(int32_t M);
int ghr_ply_pack_rows(int32_t P, int32_t M, const float* xyz, const float* normal, const float* albedo,
                      const float* roughness, const float* f_dc, const float* f_rest, const float* opacity,
                      const float* scaling, const float* rotation, float* rows, void* stream);
int ghr_ply_unpack_rows(int32_t P, int32_t M, const float* rows, int32_t row_floats, const int32_t* columns, float* xyz,
                        float* normal, float* albedo, float* roughness, float* f_dc, float* f_rest, float* opacity,
                        float* scaling, float* rotation, void* stream);

/*
 * Measurement hooks (no reference counterpart; the reference has no profiling
 * facility beyond wall-clock prints, train.py:209).  Process-wide, off by
 * default, the only state the library keeps between calls.
 *  - ghr_launch_count: kernels this library has launched since it was loaded.
 *  - ghr_profile_enable(1): every entry point brackets each stage with CUDA
 *    events on the caller's stream.  ghr_profile_read synchronises those
 *    events, returns the summed device milliseconds and the span count per
 *    stage (arrays of GHR_STAGE_COUNT) and clears the record.
 */
enum ghr_stage {
    GHR_STAGE_PREPROCESS = 0,   /* fused preprocess + scan                       */
    GHR_STAGE_EMIT_KEYS = 1,
    GHR_STAGE_SORT = 2,         /* histogram + plan + onesweep passes            */
    GHR_STAGE_TILE_RANGES = 3,
    GHR_STAGE_BLEND_FWD = 4,
    GHR_STAGE_BLEND_BWD = 5,
    GHR_STAGE_PREPROCESS_BWD = 6,
    GHR_STAGE_KNN = 7,
    GHR_STAGE_LBS_JOINTS = 8,
    GHR_STAGE_LBS_FWD = 9,
    GHR_STAGE_LBS_BWD = 10,
    GHR_STAGE_MISC = 11,
    GHR_STAGE_OPTIMIZER = 12,   /* fused Adam, densification statistics           */
    GHR_STAGE_DEFORM_PROJECT = 13,     /* fused LBS + projection forward (ghr_forward_deform_preprocess) */
    GHR_STAGE_PROJECT_DEFORM_BWD = 14, /* fused projection + LBS backward (ghr_backward_deformed)        */
    GHR_STAGE_COUNT = 15
};
uint64_t ghr_launch_count(void);
/* Scheduling hint (process-wide, default 0): 1 = several independent views are in flight on different streams
 * (mygauhuman_b200.graphed.MultiLaneRenderStep), so per-view kernels should leave room for each other: the blend
 * kernels then stage half as many entries per warp and two of them fit an SM together.  Results do not change. */
int ghr_set_shared_gpu(int on);
int ghr_shared_gpu(void); /* the current setting */
int ghr_profile_enable(int on);
int ghr_profile_read(double* ms_out, int64_t* count_out);

#ifdef __cplusplus
}
#endif
#endif /* GHR_H_INCLUDED */
