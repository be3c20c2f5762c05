/*
 * hugs_b200.h -- C ABI of the B200-native HUGS splat hot path.
 *
 * One shared library (libhugs_b200.so), plain C linkage, raw device pointers and sizes only:
 * no torch / pybind / C++ types cross this boundary.  Every entry point names the reference
 * interface it replaces (paths relative to the apple/ml-hugs tree):
 *
 *   DGR = submodules/diff-gaussian-rasterization
 *   CR  = submodules/diff-gaussian-rasterization/cuda_rasterizer
 *
 * Ownership: all device memory is caller-owned (torch-owned in the Python host).  The library
 * never allocates or frees device memory; growable workspaces are obtained through caller
 * supplied resize callbacks, exactly like the reference's std::function<char*(size_t)>
 * arguments (CR/rasterizer.h:31-34).  All work is enqueued on the caller's stream; the only
 * host synchronisation is the 4-byte read-back of the number of (Gaussian, tile) pairs that
 * the reference also performs (CR/rasterizer_impl.cu:281).
 *
 * Return value: 0 on success, a positive cudaError_t, or a negative HUGS_B200_E* code.
 * hugs_b200_last_error() returns a thread-local, human-readable message for the last failure.
 */
#ifndef HUGS_B200_H
#define HUGS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define HUGS_B200_ABI_VERSION 10

#define HUGS_B200_EINVAL   (-1)  /* bad argument (null pointer, negative size, ...)          */
#define HUGS_B200_ENOMEM   (-2)  /* a resize callback returned NULL                          */
#define HUGS_B200_EUNSUP   (-3)  /* unsupported configuration (e.g. SH degree > 3)           */

/* Opaque CUDA stream handle (cudaStream_t). NULL = legacy default stream. */
/* The `debug` argument of the rasterizer entry points is a flag word (0 / 1 keep the reference's meaning):
 *   HUGS_B200_FLAG_DEBUG    synchronise after every stage and attribute asynchronous faults to it
 *                           (CHECK_CUDA, cuda_rasterizer/auxiliary.h:166-173);
 *   HUGS_B200_FLAG_CLAMP01  the renderer's clamp(image, 0, 1) (hugs/renderer/gs_renderer.py:153) in the blend
 *                           kernels: the forward call stores the clamped image (and the layer's, if any) plus one
 *                           mask byte per pixel in the image workspace; a backward call made with the same flag
 *                           takes dL/d(clamped image) and gives the clamped channels no gradient. */
#define HUGS_B200_FLAG_DEBUG   1
#define HUGS_B200_FLAG_CLAMP01 2
#define HUGS_B200_FLAG_DEFER_SEG1 4   /* backward of a two-segment call: see hugs_b200_scene_backward_views */
#define HUGS_B200_FLAG_SEG1_PREPROCESSED 8   /* forward of a two-segment call: see hugs_b200_preprocess_scene_views */

typedef void* hugs_b200_stream_t;

/* Grow-callback: must return a device pointer to at least `nbytes` bytes (256-B aligned), valid
 * until the matching backward call has finished.  Replaces the std::function<char*(size_t)>
 * geometryBuffer / binningBuffer / imageBuffer arguments of
 * CudaRasterizer::Rasterizer::forward (CR/rasterizer.h:31-34) and the resizeFunctional lambdas
 * of DGR/rasterize_points.cu:27-33. */
typedef char* (*hugs_b200_resize_fn)(void* user, size_t nbytes);

int         hugs_b200_abi_version(void);
const char* hugs_b200_last_error(void);

/* ---------------------------------------------------------------------------------------------
 * Rasterizer forward.  Replaces CudaRasterizer::Rasterizer::forward (CR/rasterizer.h:31-57,
 * CR/rasterizer_impl.cu:198-336), i.e. what DGR/rasterize_points.cu:35-115
 * (RasterizeGaussiansCUDA) calls.
 *
 *   preprocess (CR/forward.cu:155-256) -> inclusive scan (CR/rasterizer_impl.cu:277)
 *   -> read R (..:281) -> duplicateWithKeys (..:70-111) -> radix sort on 32+bit bits (..:300-308)
 *   -> identifyTileRanges (..:116-138) -> per-tile blend (CR/forward.cu:261-374)
 *
 * Argument meaning is the reference's: all tensors are float32 device arrays, matrices are the
 * row-major-transposed 4x4 the reference expects (CR/auxiliary.h:58-77), shs is [P,M,3] AoS,
 * rotations are (r,x,y,z), any of shs / colors_precomp / scales / rotations / cov3D_precomp may be
 * NULL subject to the reference's exactly-one-of rule.  out_color is [3,H,W], radii is int32[P].
 * *num_rendered receives R.  debug != 0 synchronises and checks after every stage
 * (CR/auxiliary.h:166-173).
 * CUDA-graph capture (no counterpart in the reference, whose forward blocks on a D2H copy of R): when `stream`
 * is being captured, the call is enqueued for the pair capacity remembered from earlier uncaptured calls with
 * the same (P, width, height) on this host thread, never waits, copies nothing to the host, and reports that
 * capacity in *num_rendered (an upper bound of R; every kernel reads the real count on the device).  Without
 * such an earlier call it returns HUGS_B200_EUNSUP.  The resize callbacks are called during the capture: what
 * they return must stay valid for every replay.
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_rasterize_forward(
    hugs_b200_resize_fn geometry_buffer, void* geometry_user,
    hugs_b200_resize_fn binning_buffer,  void* binning_user,
    hugs_b200_resize_fn image_buffer,    void* image_user,
    int P, int D, int M,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* opacities,
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* cam_pos,
    float tan_fovx, float tan_fovy,
    int prefiltered,
    float* out_color,
    int* radii,
    int debug,
    hugs_b200_stream_t stream,
    int* num_rendered);

/* ---------------------------------------------------------------------------------------------
 * Rasterizer backward.  Replaces CudaRasterizer::Rasterizer::backward (CR/rasterizer.h:59-84,
 * CR/rasterizer_impl.cu:340-434), i.e. what DGR/rasterize_points.cu:117-196 calls:
 *   per-tile reverse blend (CR/backward.cu:399-557) -> cov2D backward (..:144-274)
 *   -> projection / SH / cov3D backward (..:346-396, 20-139, 278-341).
 *
 * geom/binning/image buffers are the ones forward filled.  Outputs (all float32, caller
 * allocated, NOT required to be zeroed -- the library clears what it accumulates into):
 *   dL_dmean2D [P,3] (z column zero), dL_dconic [P,4] (x,y,-,w as CR/backward.cu:549-551),
 *   dL_dopacity [P], dL_dcolor [P,3], dL_dmean3D [P,3], dL_dcov3D [P,6], dL_dsh [P,M,3],
 *   dL_dscale [P,3], dL_drot [P,4].
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_rasterize_backward(
    int P, int D, int M, int R,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* cam_pos,
    float tan_fovx, float tan_fovy,
    const int* radii,
    char* geom_buffer,
    char* binning_buffer,
    size_t binning_bytes,   /* size of the binning workspace as the forward resize callback was asked for */
    char* image_buffer,
    const float* dL_dpix,
    float* dL_dmean2D,
    float* dL_dconic,
    float* dL_dopacity,
    float* dL_dcolor,
    float* dL_dmean3D,
    float* dL_dcov3D,
    float* dL_dsh,
    float* dL_dscale,
    float* dL_drot,
    int debug,
    hugs_b200_stream_t stream);

/* The same backward pass for view-batched training (gradient accumulation over the views a rank
 * renders before the data-parallel allreduce; the reference renders one view per optimiser step and
 * leaves accumulation to torch's AccumulateGrad: read old + read new + write, for all P rows).
 * dL_dsh_accumulate == 0: identical to hugs_b200_rasterize_backward (every row of dL_dsh written,
 * zero rows for culled Gaussians).  dL_dsh_accumulate != 0: dL_dsh must hold valid data; the SH
 * gradient of every VISIBLE Gaussian is ADDED to its row, rows of culled Gaussians are not touched.
 * All other outputs are overwritten as before. */
int hugs_b200_rasterize_backward_acc(
    int P, int D, int M, int R,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* cam_pos,
    float tan_fovx, float tan_fovy,
    const int* radii,
    char* geom_buffer,
    char* binning_buffer,
    size_t binning_bytes,
    char* image_buffer,
    const float* dL_dpix,
    float* dL_dmean2D,
    float* dL_dconic,
    float* dL_dopacity,
    float* dL_dcolor,
    float* dL_dmean3D,
    float* dL_dcov3D,
    float* dL_dsh,
    float* dL_dscale,
    float* dL_drot,
    int debug,
    hugs_b200_stream_t stream,
    int dL_dsh_accumulate);

/* Replaces CudaRasterizer::Rasterizer::markVisible (CR/rasterizer.h:23-29,
 * CR/rasterizer_impl.cu:141-153; kernel checkFrustum :54-66).  present is uint8/bool [P]. */
int hugs_b200_mark_visible(int P, const float* means3D, const float* viewmatrix,
                           const float* projmatrix, uint8_t* present, hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Workspace introspection (used by the parity tests to read sort keys, the sorted Gaussian
 * list, tile ranges and per-pixel contributor counts out of the opaque byte buffers; the
 * reference's layout is CR/rasterizer_impl.cu:155-194, ours is in DESIGN.md).
 * Each call fills byte offsets from the buffer base; returns the number of entries written.
 * ------------------------------------------------------------------------------------------- */
typedef struct hugs_b200_geom_layout {
    size_t depths;         /* f32  [P]   view-space z                                   */
    size_t means2D;        /* f32x2[P]   pixel-space centre                             */
    size_t conic_opacity;  /* f32x4[P]   (conic.x, conic.y, conic.z, opacity)           */
    size_t rgb;            /* f32x4[P]   (r,g,b,unused) from SH                         */
    size_t clamped;        /* u8   [P]   bit c set <=> channel c clamped at 0           */
    size_t tiles_touched;  /* u32  [P]                                                  */
    size_t point_offsets;  /* u32  [P]   inclusive prefix sum of tiles_touched, in depth-sorted order */
    size_t scan_temp;      /* decoupled look-back state                                 */
    size_t sorted_depths;  /* u32  [P]   depth bits ascending (culled Gaussians 0xFFFFFFFF, last) */
    size_t perm;           /* u32  [P]   Gaussian index at each depth-sorted position   */
    size_t total;          /* bytes required                                            */
} hugs_b200_geom_layout;

typedef struct hugs_b200_image_layout {
    size_t accum_alpha;    /* f32  [H*W] final transmittance T                          */
    size_t n_contrib;      /* u32  [H*W] last contributing list position (1-based)      */
    size_t ranges;         /* u32x2[tiles] [first,last) into the sorted list            */
    size_t total;
} hugs_b200_image_layout;

typedef struct hugs_b200_binning_layout {
    size_t point_list_keys;           /* u64[R] sorted keys  (tile<<32 | depth bits); two-level modes: after
                                         hugs_b200_materialize_sorted_keys                                  */
    size_t point_list;                /* u32[R] sorted Gaussian indices                */
    size_t point_list_keys_unsorted;  /* u64[R] scratch; two-level modes: tile id per list position (u16, or
                                         u32 beyond 65536 tiles) once the forward call returns              */
    size_t point_list_unsorted;       /* u32[R]                                        */
    size_t sort_temp;                 /* histograms + look-back state                  */
    size_t hit_words;                 /* u32[(R+31)/32 + tiles + 1][8]: per 32 list positions of a tile and per
                                         warp of the blend CTA, which Gaussians the forward pass blended     */
    size_t total;
} hugs_b200_binning_layout;

int hugs_b200_geom_layout_of(int P, hugs_b200_geom_layout* out);
/* In the two-level binning modes nothing downstream reads the sorted 64-bit keys, so the forward pass
 * does not write them (it keeps the sorted Gaussian list and a tile id per list position).  This
 * fills point_list_keys[i] = tile(i) << 32 | depth bits of Gaussian point_list[i] for i < R -- the
 * array cub::DeviceRadixSort::SortPairs leaves in the reference's binningState.point_list_keys
 * (CR/rasterizer_impl.cu:303-308) -- from the buffers of the forward call (parity tests, debugging).
 * No-op in HUGS_B200_BINNING_REFERENCE mode, where the keys are the sort's own output. */
int hugs_b200_materialize_sorted_keys(int P, int R, int width, int height, char* geom_buffer,
                                      char* binning_buffer, size_t binning_bytes, hugs_b200_stream_t stream);
/* The binning workspace is laid out for a CAPACITY of pairs that may exceed the R of the call that
 * filled it (see hugs_b200_set_binning_mode); this recovers the capacity from the buffer size so that
 * hugs_b200_binning_layout_of(capacity, width, height) describes the buffer.  point_list always sits at
 * offset 0. */
size_t hugs_b200_binning_capacity_of(size_t nbytes, int width, int height);
int hugs_b200_image_layout_of(int width, int height, hugs_b200_image_layout* out);
int hugs_b200_binning_layout_of(int R, int width, int height, hugs_b200_binning_layout* out);

/* ---------------------------------------------------------------------------------------------
 * How the forward pass bins (Gaussian, tile) pairs.  All four produce bit-identical keys, lists,
 * ranges and images; they differ in structure and in when the host learns R.
 *   ASYNC      (default) depth argsort of the P Gaussians; the depth-ordered Gaussians are split
 *              stably by super-tile (8x4 tiles; one radix pass over ~1.5 P entries), then every
 *              super-tile list is split into its 32 tile lists by transposing 32-bit coverage masks
 *              (csrc/binning2.cu): 4 bytes written per pair, no per-pair sort.  The workspace is sized
 *              from the pair counts of earlier calls with the same (P, W, H), the whole stage is enqueued
 *              before R reaches the host, R is checked before returning and an overflow is redone with
 *              the exact size.
 *   SYNC       same kernels, but the host waits for R before sizing the workspace (what
 *              cuda_rasterizer/rasterizer_impl.cu:281 does).
 *   REFERENCE  the reference's structure: pairs emitted in Gaussian-index order, one sort over all
 *              32+bit key bits (rasterizer_impl.cu:289-308), host waits for R.
 *   TILESORT   round-1 structure, kept for cross-checks: depth argsort, pairs emitted in depth order,
 *              stable radix sort of the R pairs on the tile bits only; sized like ASYNC.
 * Returns the previous mode, or HUGS_B200_EINVAL.  Process-wide.
 * ------------------------------------------------------------------------------------------- */
#define HUGS_B200_BINNING_ASYNC     0
#define HUGS_B200_BINNING_SYNC      1
#define HUGS_B200_BINNING_REFERENCE 2
#define HUGS_B200_BINNING_TILESORT  3
int hugs_b200_set_binning_mode(int mode);

/* Launch priority of the two blend kernels (render forward / backward).  They fill every register of every SM
 * for hundreds of microseconds; a caller that runs several views on several HIGH-priority streams
 * (cudaStreamCreateWithPriority) can have the blend kernels launched at a lower priority than those streams'
 * (enable = 1, priority = e.g. the least priority of cudaDeviceGetStreamPriorityRange), so that the other views'
 * short latency-bound kernels take the CTA slots a blend kernel frees instead of queueing behind its grid.
 * enable = 0 (the default): plain launches at the stream's priority.  Process-wide; returns the previous
 * enable flag.  No counterpart in the reference (it runs one view on one stream). */
int hugs_b200_set_blend_priority(int enable, int priority);

/* The pair counts the last uncaptured hugs_b200_rasterize_forward* call of this host thread read back, once the
 * call has returned: (Gaussian, tile) pairs (= *num_rendered) and the (Gaussian, super-tile) entries of the
 * two-level binning's first level.  Measurement only (bench.py sizes its per-stage byte counts with them). */
int hugs_b200_last_pair_counts(int* pairs, int* level1_entries);

/* ---------------------------------------------------------------------------------------------
 * Stand-alone stages, exported so each can be parity-tested and timed on its own.
 * ------------------------------------------------------------------------------------------- */

/* Number of key bits the reference sorts on: 32 + getHigherMsb(tiles) (CR/rasterizer_impl.cu:35-50,300). */
int hugs_b200_sort_bits(int width, int height);

/* Bytes of temp storage hugs_b200_sort_pairs needs for n pairs. */
size_t hugs_b200_sort_temp_bytes(int n);

/* Stable LSD radix sort of (u64 key, u32 value) pairs on key bits [0, end_bit): hand-written
 * onesweep (one histogram pass + ceil(end_bit/8) decoupled-look-back scatter passes).  Replaces
 * cub::DeviceRadixSort::SortPairs at CR/rasterizer_impl.cu:303-308.  keys_in/vals_in are
 * clobbered (used as the ping-pong buffer); the result is always in keys_out/vals_out. */
int hugs_b200_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                         int n, int end_bit, void* temp, size_t temp_bytes, hugs_b200_stream_t stream);

/* The same sort restricted to key bits [begin_bit, end_bit), split into equal digits of at most 8
 * bits.  This is the second level of the rasterizer's binning: pairs emitted in (depth, index) order
 * are sorted stably on the tile bits [32, 32+bit) only, which yields the order of the reference's
 * single sort on [0, 32+bit) (CR/rasterizer_impl.cu:300-308). */
int hugs_b200_sort_pairs_range(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                               int n, int begin_bit, int end_bit, void* temp, size_t temp_bytes,
                               hugs_b200_stream_t stream);

/* Bytes of temp storage hugs_b200_inclusive_scan needs for n items. */
size_t hugs_b200_scan_temp_bytes(int n);

/* Single-pass decoupled-look-back inclusive prefix sum of u32.  Replaces
 * cub::DeviceScan::InclusiveSum at CR/rasterizer_impl.cu:277. */
int hugs_b200_inclusive_scan(const uint32_t* in, uint32_t* out, int n, void* temp, size_t temp_bytes,
                             hugs_b200_stream_t stream);

/* The same scan without any temp storage from the caller and without any allocation: the look-back state
 * lives in a file-scope device array of the library (16.3 KB per GPU, held until exit).  No counterpart in
 * the reference: added for callers that cannot call cudaMalloc (CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  n > HUGS_B200_SCAN_STATIC_MAX_ITEMS is
 * rejected (HUGS_B200_EUNSUP, message in hugs_b200_last_error).  in/out: device pointers, 16-byte aligned.
 * Every call shares the array: calls that overlap on the GPU (different streams or host threads) race --
 * serialise them.  Same result as hugs_b200_inclusive_scan, bit for bit. */
#define HUGS_B200_SCAN_STATIC_MAX_ITEMS (1 << 24)
int hugs_b200_inclusive_scan_static_max_items(void);
int hugs_b200_inclusive_scan_static(const uint32_t* in, uint32_t* out, int n, hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * SMPL linear-blend-skinning deformation of the human Gaussians (fused).  Replaces the PyTorch
 * chain  lbs_extra (hugs/models/modules/lbs.py:19-73, with disable_posedirs=True: v_posed=v_shaped)
 *  + the deformation tail hugs/models/hugs_trimlp.py:510-518
 *  + matrix_to_quaternion (hugs/utils/rotations.py:105-156):
 *
 *   T_i   = sum_j w_ij A_fj            (A: [F,J,4,4] row-major joint transforms, w: [N,J])
 *   xyz'  = (T_i [xyz_i;1])[:3] * smpl_scale + transl_f
 *   R'_i  = T_i[:3,:3] R_i ;  q_i = matrix_to_quaternion(R'_i)   (real first, NOT normalised)
 *   s'_i  = s_i * smpl_scale
 *
 * F pose frames are processed in one launch (F=1 for training; F=300 for the animation config).
 * Inputs: A [F,J,16], xyz [N,3], weights [N,J], rotmat [N,9] row-major, scales [N,3] (may be
 * NULL), transl [F,3] (may be NULL => 0), smpl_scale scalar.
 * Outputs: out_xyz [F,N,3], out_rotq [F,N,4], out_scales [F,N,3] (if scales), out_T [F,N,16]
 * (optional, NULL to skip; the reference materialises it).
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_lbs_forward(int F, int N, int J,
                          const float* A, const float* xyz, const float* weights,
                          const float* rotmat, const float* scales,
                          const float* transl, float smpl_scale,
                          float* out_xyz, float* out_rotq, float* out_scales, float* out_T,
                          hugs_b200_stream_t stream);

/* Backward of hugs_b200_lbs_forward (the reference relies on torch autograd through the chain
 * above).  Upstream grads: g_xyz [F,N,3], g_rotq [F,N,4], g_scales [F,N,3] (may be NULL).
 * Outputs (caller allocated; the library zero-fills what it accumulates into):
 *   dA [F,J,16], d_xyz [N,3], d_weights [N,J], d_rotmat [N,9], d_scales [N,3] (if scales),
 *   d_transl [F,3] (may be NULL), d_smpl_scale [1] (may be NULL). */
int hugs_b200_lbs_backward(int F, int N, int J,
                           const float* A, const float* xyz, const float* weights,
                           const float* rotmat, const float* scales,
                           const float* transl, float smpl_scale,
                           const float* g_xyz, const float* g_rotq, const float* g_scales,
                           float* dA, float* d_xyz, float* d_weights, float* d_rotmat,
                           float* d_scales, float* d_transl, float* d_smpl_scale,
                           hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Image-space losses, the step right after the rasterizer in the reference's training loop
 * (SURVEY.md 8(f) N3).  Replaces l1_loss + ssim/_ssim (hugs/losses/utils.py:52-55, 63-103: window 11,
 * sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2) as called by hugs/losses/loss.py:87-107, and
 * their autograd backward.  Images are float32 [C,H,W] (a batch is folded into C: the reference's
 * conv2d is depthwise).
 *   forward : sums[0] = sum |img1 - img2|, sums[1] = sum of the SSIM map (double[2], device; divide
 *             by C*H*W for the means the reference returns); dmap float32 [3][C][H][W] receives the
 *             per-pixel derivative planes the backward needs (NULL: forward only).
 *   backward: dL_dimg1 [C,H,W] = g[0] * d(l1 mean)/d img1 + g[1] * d(ssim mean)/d img1, with
 *             g = float32[2] on the device (the autograd gradients of the two scalars).
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_l1_ssim_forward(int C, int H, int W, const float* img1, const float* img2,
                              float* dmap, double* sums, hugs_b200_stream_t stream);
int hugs_b200_l1_ssim_backward(int C, int H, int W, const float* img1, const float* img2,
                               const float* dmap, const float* g, float* dL_dimg1,
                               hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Mean SQUARED distance of every point to its 3 nearest neighbours (exact).  Replaces simple-knn's
 * distCUDA2 (submodules/simple-knn/simple_knn.cu:185-221, spatial.cu), which hugs/models/scene.py:181
 * uses to initialise the scene Gaussians' scales.  points float32 [P,3],
 * mean_dist2 float32 [P]; temp: hugs_b200_knn_temp_bytes(P) bytes of device scratch.  Fewer than 4
 * points: the missing neighbours count as FLT_MAX, as in the reference (huge or inf results).
 * ------------------------------------------------------------------------------------------- */
size_t hugs_b200_knn_temp_bytes(int P);
int hugs_b200_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* temp, size_t temp_bytes,
                             hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Two input segments: the renderer's human_scene mode concatenates the human and
 * the scene Gaussians with five torch.cat calls per render (hugs/renderer/gs_renderer.py:33-37), a
 * copy of 236 B per Gaussian.  These entry points take the two sets of arrays as they are: Gaussians
 * [0, P0) come from segment 0 (the human), [P0, P0+P1) from segment 1 (the scene).  SH path with
 * scales / rotations only (what that renderer passes).  Outputs, workspaces and gradients are those
 * of hugs_b200_rasterize_forward / _backward on the concatenation, bit for bit: radii and every
 * dL_d* array have P0 + P1 rows, rows [0, P0) belonging to segment 0.
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_rasterize_forward_seg2(
    hugs_b200_resize_fn geometry_buffer, void* geometry_user,
    hugs_b200_resize_fn binning_buffer,  void* binning_user,
    hugs_b200_resize_fn image_buffer,    void* image_user,
    int D, int M, const float* background, int width, int height,
    int P0, const float* means3D0, const float* shs0, const float* opacities0, const float* scales0,
    const float* rotations0,
    int P1, const float* means3D1, const float* shs1, const float* opacities1, const float* scales1,
    const float* rotations1,
    float scale_modifier, const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy, int prefiltered, float* out_color, int* radii, int debug,
    hugs_b200_stream_t stream, int* num_rendered);
int hugs_b200_rasterize_backward_seg2(
    int D, int M, int R, const float* background, int width, int height,
    int P0, const float* means3D0, const float* shs0, const float* scales0, const float* rotations0,
    int P1, const float* means3D1, const float* shs1, const float* scales1, const float* rotations1,
    float scale_modifier, const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy, const int* radii, char* geom_buffer, char* binning_buffer,
    size_t binning_bytes, char* image_buffer, const float* dL_dpix,
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
    float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, int debug,
    hugs_b200_stream_t stream);


/* ---- a second image of the first n_layer Gaussians, from the same sorted lists -----------------------
 * Replaces the second, human-only render() of hugs/renderer/gs_renderer.py:68-82 (consumed by
 * hugs/trainer/gs_trainer.py:255-271 and the human-separate loss): instead of a second complete
 * rasterizer pass over the human Gaussians, the joint pass's tile lists are filtered (stable) to the
 * Gaussians with index < n_layer and blended against `background`; contributor counts, transmittance and
 * image are those of CudaRasterizer::Rasterizer::forward (cuda_rasterizer/rasterizer_impl.cu:198-336) on
 * the first n_layer Gaussians alone.  Call after hugs_b200_rasterize_forward[_seg2] on the same stream with
 * that call's three workspaces; `layer_buffer` (caller-owned, hugs_b200_layer_bytes) keeps what the
 * backward needs.  colors_precomp as in the forward call (NULL with SHs). */
size_t hugs_b200_layer_bytes(int n_layer, size_t binning_bytes, int width, int height);
/* byte offsets inside the layer buffer, for tests: filtered list u32[capacity], hit masks, final T f32[W*H],
 * contributor counts u32[W*H], tile ranges u32x2[tiles], packed gradient records f32[n_layer][8] */
int hugs_b200_layer_layout_of(int n_layer, size_t binning_bytes, int width, int height, size_t* offsets6);
int hugs_b200_layer_forward(int P, int n_layer, const float* background, int width, int height,
                            const float* colors_precomp, char* geom_buffer, char* binning_buffer,
                            size_t binning_bytes, char* image_buffer, char* layer_buffer, size_t layer_bytes,
                            float* out_color, int debug, hugs_b200_stream_t stream);
/* hugs_b200_rasterize_backward_acc with the layer's image gradient consumed in the same pass
 * (CudaRasterizer::Rasterizer::backward, rasterizer_impl.cu:340-434, run once instead of twice): every
 * gradient is the sum over both images, except dL_dmean2D, which stays the main image's (the reference's
 * second pass has its own, unread, screen-space tensor).  dL_dpix_layer == NULL: plain backward. */
int hugs_b200_rasterize_backward_layer(int P, int D, int M, int R, const float* background, int width, int height,
                                       const float* means3D, const float* shs, const float* colors_precomp,
                                       const float* scales, float scale_modifier, const float* rotations,
                                       const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                       const float* cam_pos, float tan_fovx, float tan_fovy, const int* radii,
                                       char* geom_buffer, char* binning_buffer, size_t binning_bytes,
                                       char* image_buffer, const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                                       float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                                       float* dL_dsh, float* dL_dscale, float* dL_drot, int debug,
                                       hugs_b200_stream_t stream, int dL_dsh_accumulate, int n_layer,
                                       const float* background_layer, char* layer_buffer, size_t layer_bytes,
                                       const float* dL_dpix_layer);
int hugs_b200_rasterize_backward_seg2_layer(int D, int M, int R, const float* background, int width, int height,
                                            int P0, const float* means3D0, const float* shs0, const float* scales0,
                                            const float* rotations0, int P1, const float* means3D1,
                                            const float* shs1, const float* scales1, const float* rotations1,
                                            float scale_modifier, const float* viewmatrix, const float* projmatrix,
                                            const float* cam_pos, float tan_fovx, float tan_fovy, const int* radii,
                                            char* geom_buffer, char* binning_buffer, size_t binning_bytes,
                                            char* image_buffer, const float* dL_dpix, float* dL_dmean2D,
                                            float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                                            float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                            int debug, hugs_b200_stream_t stream, int n_layer,
                                            const float* background_layer, char* layer_buffer, size_t layer_bytes,
                                            const float* dL_dpix_layer);


/* ---- gradient sinks (view-batched training) ------------------------------------------------------------
 * The reference renders one view per optimiser step (hugs/trainer/gs_trainer.py:226-290); torch's
 * AccumulateGrad adds every further view's gradients to .grad with one elementwise kernel per parameter.
 * Here the per-Gaussian backward can deliver a segment's rows of a gradient straight into the parameter's
 * accumulation buffer: entry [s] != NULL means "rows of segment s (0: [0,P0), 1: [P0,P0+P1)) of this
 * gradient are ADDED, for visible Gaussians only, to this [P_s, ...] buffer (16-byte aligned) and the
 * segment's rows of the joint output array are left unwritten".  The caller zeroes the buffers at the
 * start of a step; several views may add concurrently from different streams. */
typedef struct hugs_b200_grad_sinks {
    float* means3D[2];    /* [P_s,3]   */
    float* sh[2];         /* [P_s,M,3] */
    float* opacity[2];    /* [P_s,1]   */
    float* scales[2];     /* [P_s,3]   */
    float* rotations[2];  /* [P_s,4]   */
} hugs_b200_grad_sinks;
/* hugs_b200_rasterize_backward_seg2_layer + optional sinks (NULL: none); dL_dpix_layer == NULL: no layer. */
int hugs_b200_rasterize_backward_seg2_ex(int D, int M, int R, const float* background, int width, int height,
                                         int P0, const float* means3D0, const float* shs0, const float* scales0,
                                         const float* rotations0, int P1, const float* means3D1, const float* shs1,
                                         const float* scales1, const float* rotations1, float scale_modifier,
                                         const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                                         float tan_fovx, float tan_fovy, const int* radii, char* geom_buffer,
                                         char* binning_buffer, size_t binning_bytes, char* image_buffer,
                                         const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                                         float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                                         float* dL_dsh, float* dL_dscale, float* dL_drot, int debug,
                                         hugs_b200_stream_t stream, int n_layer, const float* background_layer,
                                         char* layer_buffer, size_t layer_bytes, const float* dL_dpix_layer,
                                         const hugs_b200_grad_sinks* sinks);

/* The scene segment's per-Gaussian backward for ALL views of a view-batched training step in one launch (no
 * counterpart in the reference, which trains on one view per step).  A backward call made with
 * HUGS_B200_FLAG_DEFER_SEG1 in `debug` (two segments, all five sinks of segment 1 set) runs the blend backward
 * and segment 0's per-Gaussian backward only and leaves segment 1's share in the view's geometry workspace
 * (packed records, clamp bits, opacity scratch).  When the last view of the step has done so, this call reads
 * every scene Gaussian's parameters and SH row ONCE, walks the views, and adds the summed gradients to the sinks
 * once -- ~110 B per Gaussian and view instead of 744.  Per view: its camera, its radii [P0+P1] and its geometry
 * workspace (unchanged since the view's backward call).  The deferring backward call still writes segment 1's
 * rows of dL_dmean2D (the screen-space gradient it reports); its dL_dconic, dL_dcolor and dL_dcov3D rows are
 * not produced.  Must be ordered after every view's backward call on the device (same stream, or an event). */
typedef struct hugs_b200_scene_view {
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float tan_fovx, tan_fovy;
    int* radii;          /* [P0+P1] of the view */
    char* geom_buffer;   /* the view's geometry workspace */
} hugs_b200_scene_view;
int hugs_b200_scene_backward_views(int n_views, const hugs_b200_scene_view* views, int P0, int P1, int D, int M,
                                   int width, int height, const float* means3D1, const float* shs1,
                                   const float* scales1, float scale_modifier, const float* rotations1,
                                   const hugs_b200_grad_sinks* sinks, hugs_b200_stream_t stream);
/* The forward counterpart: the scene segment's preprocess (CR/forward.cu:155-256) for ALL views of a step in one
 * launch -- parameters and SH row read once per Gaussian, the 3D covariance built once, the cameras walked; per
 * (Gaussian, camera) the arithmetic of the per-view kernel, operation for operation.  For every view it clears
 * the pair counts of the view's geometry workspace (geom_buffer: hugs_b200_geom_layout_of(P0 + P1).total bytes,
 * owned by the caller), fills segment 1's rows of it and of radii [P0+P1], and adds segment 1's share of the
 * counts.  The view's forward call is then made with HUGS_B200_FLAG_SEG1_PREPROCESSED and a geometry callback
 * that returns that same buffer: it preprocesses segment 0 only and carries on from there. */
int hugs_b200_preprocess_scene_views(int n_views, const hugs_b200_scene_view* views, int P0, int P1, int D, int M,
                                     int width, int height, const float* means3D1, const float* shs1,
                                     const float* opacities1, const float* scales1, float scale_modifier,
                                     const float* rotations1, int prefiltered, hugs_b200_stream_t stream);

/* ---- fused decoder MLPs on the tensor cores (tcgen05 / TMEM, 3xTF32) -----------------------------------
 * Replaces the torch.nn.Linear / GELU chains of hugs/models/modules/decoders.py:24-110 (AppearanceDecoder
 * :24-44, DeformationDecoder :47-84, GeometryDecoder :87-110) as hugs/models/hugs_trimlp.py:205-238 runs them
 * on the [N, 96] tri-plane features.  A chain is a sequence of Linear layers: every layer but the last is
 * followed by `act` and feeds the next; the last layer's columns are written to up to three outputs, each
 * with its own activation (the decoders' heads concatenated into one layer).  Up to three chains read the
 * same input in one launch.  Weights are handed over in the packed form hugs_b200_mlp_pack produces from a
 * Linear's [n_true, K] weight and [n_true] bias (n_pad: n_true rounded up to 16, hidden layers to 32).
 * z_out (optional): the layer's pre-activations [rows, N], kept for the backward pass. */
#define HUGS_B200_MLP_ACT_NONE    0
#define HUGS_B200_MLP_ACT_GELU    1   /* exact, erf: torch.nn.GELU() */
#define HUGS_B200_MLP_ACT_SIGMOID 2
typedef struct hugs_b200_mlp_out {
    float* ptr;             /* out[row * ld + (col - col0)] for col in [col0, col0 + ncols) */
    int ld, col0, ncols, act;
} hugs_b200_mlp_out;
typedef struct hugs_b200_mlp_layer {
    int K, N, act;          /* N: padded width */
    const float* packed;
    float* z_out;
} hugs_b200_mlp_layer;
typedef struct hugs_b200_mlp_chain {
    int n_layers, n_outs;
    hugs_b200_mlp_layer layer[4];
    hugs_b200_mlp_out out[3];
} hugs_b200_mlp_chain;
size_t hugs_b200_mlp_packed_floats(int K, int n_pad);
int hugs_b200_mlp_pack(const float* W, const float* bias, int n_true, int n_pad, int K, float* packed,
                       hugs_b200_stream_t stream);
int hugs_b200_mlp_chains_forward(int rows, const float* x, int ldx, int Kx, int n_chains,
                                 const hugs_b200_mlp_chain* chains, hugs_b200_stream_t stream);
/* Backward of the same chains (what autograd runs through the reference's Linear / GELU modules): from the
 * gradients of the head outputs and the pre-activations the forward call saved, per layer
 * dW^T [K, N] and db [N] are ADDED to caller-zeroed buffers (N: the layer's width rounded up to 32; the
 * caller transposes / trims) and every chain writes its share of dL/dx [rows, Kx].  packed_t: the layer's
 * weight through hugs_b200_mlp_pack_t (W^T, split for 3xTF32). */
typedef struct hugs_b200_mlp_bwd_out {
    const float* dptr;      /* NULL: this output has no gradient */
    int ld, col0, ncols, act;
} hugs_b200_mlp_bwd_out;
typedef struct hugs_b200_mlp_bwd_layer {
    int K, N;
    const float* packed_t;
    const float* z_prev;    /* [rows, K] pre-activations of the layer below; NULL for the first layer */
    float* dWt;             /* [K, N] */
    float* db;              /* [N] */
} hugs_b200_mlp_bwd_layer;
typedef struct hugs_b200_mlp_bwd_chain {
    int n_layers, n_outs;
    hugs_b200_mlp_bwd_layer layer[4];
    hugs_b200_mlp_bwd_out out[3];
    const float* z_head;    /* [rows, N_head] */
    float* dx;              /* [rows, Kx] or NULL */
} hugs_b200_mlp_bwd_chain;
int hugs_b200_mlp_pack_t(const float* W, int n_true, int k_true, int n_pad, float* packed, hugs_b200_stream_t stream);
int hugs_b200_mlp_chains_backward(int rows, const float* x, int ldx, int Kx, int n_chains,
                                  const hugs_b200_mlp_bwd_chain* chains, hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Densification statistics after a backward pass, one launch, no host sync.
 * For every i < n with radii[i] > 0 (the reference's visibility_filter, gs_renderer.py:156):
 *   max_radii2D[i]        = max(max_radii2D[i], radii[i])          gs_trainer.py:406-411, :429-435
 *   xyz_gradient_accum[i] += ||grad_means2D[i, 0:2]||              scene.py:460-461
 *   denom[i]              += 1                                     scene.py:462
 * radii int32 [n]; grad_means2D float32 rows of grad_stride floats (3 for the rasterizer's
 * dL_dmeans2D); the three statistics float32 [n] (the reference keeps them as float tensors, [n] or
 * [n,1]).  Any of the three statistics may be NULL (skipped).
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_densification_stats(int n, const int* radii, const float* grad_means2D, int grad_stride,
                                  float* xyz_gradient_accum, float* denom, float* max_radii2D,
                                  hugs_b200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Tri-plane feature sampling.  Replaces TriPlane.forward,
 * hugs/models/modules/triplane.py:25-39: g = ((xyz - center)/scale + 0.5)*2 - 1, then three bilinear
 * grid samples (zero padding, align_corners=True) of plane_xy at (g.x,g.y), plane_xz at (g.x,g.z),
 * plane_yz at (g.y,g.z) -- the first coordinate indexes the LAST plane axis, as F.grid_sample does --
 * concatenated to features [N, 3C].
 * Planes are logically [1,C,resA,resB] (xy: resX,resY; xz: resX,resZ; yz: resY,resZ) and are passed
 * CHANNELS-LAST, i.e. as dense [resA][resB][C] float32 (torch.channels_last of the same tensor).
 * out_of_range (int32 on the device, may be NULL) is OR-ed with 1 when a normalised coordinate leaves
 * [-eps, 1+eps] -- the condition the reference asserts on (triplane.py:28).
 * backward: dL_dplane_* (same layout) are ACCUMULATED into (zero them first; NULL skips the planes);
 * dL_dxyz [N,3] is written (NULL skips it).
 * ------------------------------------------------------------------------------------------- */
int hugs_b200_triplane_forward(int N, int C, int resX, int resY, int resZ, const float* plane_xy,
                               const float* plane_xz, const float* plane_yz, const float* xyz, float center,
                               float scale, float eps, float* features, int* out_of_range,
                               hugs_b200_stream_t stream);
int hugs_b200_triplane_backward(int N, int C, int resX, int resY, int resZ, const float* plane_xy,
                                const float* plane_xz, const float* plane_yz, const float* xyz, float center,
                                float scale, const float* dL_dfeatures, float* dL_dplane_xy, float* dL_dplane_xz,
                                float* dL_dplane_yz, float* dL_dxyz, hugs_b200_stream_t stream);

/* Number of kernel launches issued by this library (all host threads) since the last reset
 * (bench.py reports it as gpu_launches). */
long long hugs_b200_launch_count(int reset);

/* ---------------------------------------------------------------------------------------------
 * Per-stage device timing (measurement aid; the reference has no counterpart -- SURVEY.md 5.1).
 * While enabled, each stage of the calls above is bracketed by two CUDA events recorded on the
 * caller's stream.  hugs_b200_stage_timing_read waits for the recorded events, adds them to the
 * per-stage totals and copies the totals out (ms_sum / calls: HUGS_B200_NUM_STAGES entries each).
 * Stage order: preprocess, inclusive_scan, duplicate_with_keys, sort_pairs, identify_tile_ranges,
 * render_forward, render_backward, preprocess_backward, lbs_forward, lbs_backward, tile_split.
 * ------------------------------------------------------------------------------------------- */
#define HUGS_B200_NUM_STAGES 11
int         hugs_b200_stage_timing_enable(int on); /* returns the previous setting */
int         hugs_b200_stage_timing_read(double* ms_sum, long long* calls, int reset);
const char* hugs_b200_stage_name(int stage);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* HUGS_B200_H */
