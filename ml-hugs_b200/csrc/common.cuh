// Shared declarations for the B200 (sm_100a) splat hot path.  Internal header: nothing here
// crosses the C ABI (include/hugs_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define HB_TILE_X 16
#define HB_TILE_Y 16
#define HB_TILE_PIX (HB_TILE_X * HB_TILE_Y)
#define HB_ALIGN 256
#define HUGS_B200_EUNSUP_INTERNAL (-3)

namespace hb {

// ---- error plumbing (thread-local message, see api.cu) -----------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define HB_CUDA_OK(expr)                                                                      \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            hb::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,                \
                          cudaGetErrorString(_e));                                            \
            return (int)_e;                                                                   \
        }                                                                                     \
    } while (0)

// After a kernel launch: always pick up launch-configuration errors; in debug mode also
// synchronise so that asynchronous faults are attributed to the stage that caused them
// (same contract as CHECK_CUDA in the reference, cuda_rasterizer/auxiliary.h:166-173).
#define HB_LAUNCH_OK(stage, stream, debug)                                                    \
    do {                                                                                      \
        hb::count_launch();                                                                   \
        cudaError_t _e = cudaGetLastError();                                                  \
        if (_e == cudaSuccess && (debug)) _e = cudaStreamSynchronize(stream);                 \
        if (_e != cudaSuccess) {                                                              \
            hb::set_error("stage '%s' failed: %s", stage, cudaGetErrorString(_e));            \
            return (int)_e;                                                                   \
        }                                                                                     \
    } while (0)

static inline size_t align_up(size_t x, size_t a = HB_ALIGN) { return (x + a - 1) / a * a; }
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- workspace views --------------------------------------------------------------------
struct GeomView {
    float* depths;
    float2* means2D;
    float4* conic_opacity;
    float4* rgb;
    uint8_t* clamped;
    uint32_t* tiles_touched;
    uint32_t* point_offsets;  // inclusive scan, in depth-sorted order, of st_touched (tiles_touched in TILESORT mode)
    char* scan_temp;
    size_t scan_temp_bytes;
    uint32_t* sorted_depths;  // depth bits, ascending (culled Gaussians: 0xFFFFFFFF, last)
    uint32_t* perm;           // Gaussian index at each depth-sorted position
    uint32_t* dsort_keys_tmp;
    uint32_t* dsort_vals_tmp;
    char* dsort_temp;
    size_t dsort_temp_bytes;
    float* grad_opac;         // [P] backward scratch: dL/dopacity when it is bound for a gradient sink
    float* grad_pack;         // [P][8] backward scratch: dL/d(colour 3, mean2D 2, conic 3), see render_backward.cu
    uint2* rects;             // tile rectangle of every Gaussian (x0 | x1 << 16, y0 | y1 << 16)
    uint32_t* st_touched;     // super-tiles its rectangle overlaps
    uint32_t* pair_counts;    // [0] R = sum of tiles_touched, [1] R1 = sum of st_touched
    size_t total;
};
struct ImageView {
    float* accum_alpha;
    uint32_t* n_contrib;
    uint2* ranges;
    uint8_t* clamp_mask;   // [W*H] bit c: channel c of the pixel was clamped to [0, 1] (written only when asked for)
    size_t total;
};
struct LayerView {   // second image of the first n Gaussians (layer.cu)
    uint32_t* list;       // [capacity] filtered tile lists, each at its joint offset
    uint32_t* hit_words;  // same shape as BinningView::hit_words
    float* accum_alpha;   // [W*H]
    uint32_t* n_contrib;  // [W*H]
    uint2* ranges;        // [tiles]
    float* grad_pack;     // [n][8] the layer's share of the packed blend gradients
    uint8_t* clamp_mask;  // [W*H] as ImageView::clamp_mask, for the layer's image
    size_t total;
};
struct BinningView {
    uint64_t* keys;
    uint32_t* list;
    uint64_t* keys_unsorted;
    uint32_t* list_unsorted;
    char* sort_temp;
    size_t sort_temp_bytes;
    char* l2_temp;        // level-2 binning scratch (binning2.cu: carve_l2_temp)
    size_t l2_temp_bytes;
    uint32_t* hit_words;  // [(R+31)/32 + tiles + 1][8 warps]: what each warp of the forward blend actually blended
    size_t total;
};
// Super-tiles: 8 x 4 tiles.
static inline int super_grid_x(int grid_x) { return (grid_x + 7) / 8; }
static inline int super_grid_y(int grid_y) { return (grid_y + 3) / 4; }
struct L2Temp {
    uint32_t chunk_cap;
    uint32_t* counts;          // [chunk_cap][32]
    uint2* chunks;             // [chunk_cap] (super-tile | entries << 16, first list position)
    uint2* st_ranges;          // [n_st] list range of every super-tile
    uint32_t* st_chunk_start;  // [n_st + 1]
    uint32_t* tile_total;      // [tiles]
    uint32_t* tile_start;      // [tiles]
    uint32_t* n_chunks;        // [1]
    size_t total;
};
L2Temp carve_l2_temp(char* base, int cap, int n_st, int num_tiles);
int launch_l2_split(const L2Temp& t, const uint32_t* st_list, const uint2* rects, int n_st, int gx_s, int grid_x,
                    int grid_y, int cap1, int cap, uint32_t* list, uint2* ranges, cudaStream_t stream);
int launch_keys_from_ranges(int num_tiles, const uint32_t* tile_start, const uint32_t* tile_total,
                            const uint32_t* list, const uint32_t* depth_bits, uint64_t* keys, cudaStream_t stream);
GeomView carve_geom(char* base, int P);
ImageView carve_image(char* base, int W, int H);
BinningView carve_binning(char* base, int R, int grid_x, int grid_y);
LayerView carve_layer(char* base, int R, int n_layer, int W, int H);

constexpr int kSortMaxPasses = 8;
struct SortPassPlan {  // pass p sorts on `bits[p]` key bits starting at bit `shift[p]`
    int num_passes;
    int shift[kSortMaxPasses];
    int bits[kSortMaxPasses];
};

// ---- stage launchers (one per .cu file) ----------------------------------------------------
struct Camera {
    const float* view;   // 16 floats, device
    const float* proj;   // 16 floats, device
    const float* campos; // 3 floats, device
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y;
};

int launch_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* opacities, const float* shs,
                      const float* cov3D_precomp, const float* colors_precomp, const Camera& cam,
                      int* radii, const GeomView& g, bool prefiltered, cudaStream_t stream,
                      struct SegBatch* batch = nullptr);
int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj, uint8_t* present,
                        cudaStream_t stream);

size_t scan_temp_bytes(int n);
int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, void* temp, size_t temp_bytes,
                          cudaStream_t stream, const uint32_t* gather = nullptr);

int launch_duplicate(int P, const float2* means2D, const float* depths, const uint32_t* offsets,
                     const int* radii, uint64_t* keys, uint32_t* vals, int grid_x, int grid_y,
                     cudaStream_t stream);
int launch_duplicate_sorted(int P, const uint32_t* perm, const uint32_t* offsets, const float2* means2D,
                            const int* radii, uint64_t* pairs, int grid_x, int grid_y, int cap,
                            const SortPassPlan& plan, uint32_t* hist, cudaStream_t stream);
int launch_duplicate_super(int P, const uint32_t* perm, const uint32_t* offsets, const uint2* rects, uint64_t* pairs,
                           int gx_s, int gy_s, int cap1, const SortPassPlan& plan, uint32_t* hist,
                           cudaStream_t stream);
int launch_tile_ranges(int R, const uint64_t* keys, uint2* ranges, int num_tiles, cudaStream_t stream,
                       const uint32_t* n_dev = nullptr);
// the same from the per-position tile ids the two-level sort leaves behind (u16 or u32)
int launch_tile_ranges_ids(int R, const void* tile_ids, bool tile16, uint2* ranges, int num_tiles,
                           cudaStream_t stream, const uint32_t* n_dev);
// keys[i] = tile_ids[i] << 32 | depth_bits[list[i]]: the sorted 64-bit keys, on demand
int launch_materialize_keys(int R, const void* tile_ids, bool tile16, const uint32_t* list,
                            const uint32_t* depth_bits, uint64_t* keys, cudaStream_t stream);

// ---- radix sort (radix_sort.cu) ------------------------------------------------------------------
size_t sort_temp_bytes(int n);
SortPassPlan tile_pass_plan(int tile_bits);  // passes over key bits [32, 32+tile_bits)
int sort_temp_clear(void* temp, size_t temp_bytes, int n_cap, cudaStream_t stream);
uint32_t* sort_temp_histogram(void* temp);   // [kSortMaxPasses][256] digit counts, filled by the duplicate kernel
int launch_depth_argsort(const uint32_t* keys0, uint32_t* kB, uint32_t* vB, uint32_t* kC, uint32_t* vC, int P,
                         void* temp, size_t temp_bytes, cudaStream_t stream);
int launch_tile_sort(uint64_t* a, uint64_t* b, void* tiles_out, uint32_t* list_out, bool tile16, int n_cap,
                     const uint32_t* n_dev, const SortPassPlan& plan, void* temp, size_t temp_bytes,
                     cudaStream_t stream);
int launch_sort_pairs_pingpong(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                               int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream,
                               bool* result_in_input, int begin_bit = 0);
int launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                      int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream, int begin_bit = 0);

int launch_render_forward(const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x, int grid_y,
                          const float2* means2D, const float* colors, int color_stride,
                          const float4* conic_opacity, float* final_T, uint32_t* n_contrib,
                          const float* bg, float* out_color, uint32_t* hit_words, cudaStream_t stream, uint8_t* clamp_mask = nullptr);
int launch_render_backward(const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x, int grid_y,
                           const float* bg, const float2* means2D, const float4* conic_opacity,
                           const float* colors, int color_stride, const float* final_T,
                           const uint32_t* n_contrib, const uint32_t* hit_words, const float* dL_dpix,
                           float* grad_pack /*[P,8], zeroed*/, float* dL_dopacity /*[P], zeroed*/,
                           cudaStream_t stream, const uint8_t* clamp_mask = nullptr);
// Two-segment calls: the per-segment launches of a per-Gaussian kernel are collected here and issued as ONE
// launch (flush_*): the human segment alone (110 k rows) is too small a grid to fill the machine.
struct SegBatch {
    int n = 0;
    alignas(16) unsigned char args[2][768];   // the kernel's argument record (private to its .cu file)
    int blocks[2] = {0, 0};
    size_t smem[2] = {0, 0};
};
int flush_preprocess(SegBatch& b, cudaStream_t stream);
int flush_preprocess_backward(SegBatch& b, cudaStream_t stream);

// Optional inputs of the per-Gaussian backward: a second image's packed records (layer.cu) and gradient sinks.
constexpr int kAccMean3D = 1, kAccScale = 2, kAccRot = 4, kAccOpacity = 8;
struct PreBwdExtra {
    const float* grad_pack2 = nullptr;  // [n2][8], rows [0, n2)
    int n2 = 0;
    int acc_mask = 0;                   // kAcc* bits: that output is ADDED for visible rows, culled rows untouched
    const float* opac_scratch = nullptr;   // [P]: when set, dL/dopacity is taken from here and written / added to dL_dopacity
    float* dL_dopacity = nullptr;
};
int launch_preprocess_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs,
                               const uint8_t* clamped, const float* scales, const float* rotations,
                               float scale_modifier, const float* cov3D_precomp, const Camera& cam,
                               const float* grad_pack, float* dL_dmean2D, float* dL_dconic, float* dL_dmean3D,
                               float* dL_dcolor, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                               float* dL_drot, bool sh_accumulate, cudaStream_t stream,
                               const struct PreBwdExtra* extra = nullptr, SegBatch* batch = nullptr);
// The scene segment's preprocess for all views of a step in one launch (preprocess.cu).
struct PreViewHost {
    Camera cam;
    int* radii;     // scene rows of the view's radii
    GeomView g;     // the view's geometry workspace, per-Gaussian arrays moved to the scene rows
};
int launch_preprocess_views(int P, int D, int M, int n_views, const PreViewHost* views, const float* means3D,
                            const float* scales, float scale_modifier, const float* rotations, const float* opacities,
                            const float* shs, bool prefiltered, cudaStream_t stream);
// The scene segment's per-Gaussian backward for all views of a step in one launch (preprocess_backward.cu).
struct SceneViewHost {
    Camera cam;
    const int* radii;            // scene rows of the view's arrays
    const float* grad_pack;
    const uint8_t* clamped;
    const float* opac_scratch;
};
int launch_unpack_mean2d(int P, const int* radii, const float* grad_pack, float* dL_dmean2D, cudaStream_t stream);
int launch_scene_backward_views(int P, int D, int M, int n_views, const SceneViewHost* views, const float* means3D,
                                const float* shs, const float* scales, const float* rotations, float scale_modifier,
                                float* g_mean, float* g_sh, float* g_scale, float* g_rot, float* g_opac,
                                cudaStream_t stream);
int launch_layer_filter(int num_tiles, const uint2* ranges, const uint32_t* list, int n_layer, uint32_t* list_out,
                        uint2* ranges_out, cudaStream_t stream);

// ---- fused decoder MLPs on tcgen05 (mlp_tc.cu) -----------------------------------------------------------
#define HB_MLP_ACT_NONE 0
#define HB_MLP_ACT_GELU 1
#define HB_MLP_ACT_SIGMOID 2
#define HB_MLP_MAX_CHAINS 3
#define HB_MLP_MAX_LAYERS 4
#define HB_MLP_MAX_OUTS 3
struct MlpOut {          // columns [col0, col0 + ncols) of a chain's last layer -> out[row * ld + (col - col0)]
    float* ptr;
    int ld, col0, ncols, act;
};
struct MlpLayerArgs {
    int K, N;            // N padded to a multiple of 16 (hidden layers: of 32)
    int act;             // activation between this layer and the next (hidden layers)
    const float* packed; // mlp_pack_kernel's output for this layer
    float* z_out;        // optional [rows x N] pre-activations (bias included)
};
struct MlpChainArgs {
    int n_layers, n_outs;
    MlpLayerArgs layer[HB_MLP_MAX_LAYERS];
    MlpOut out[HB_MLP_MAX_OUTS];
};
struct MlpArgs {
    int rows, Kx, ldx, n_chains;
    const float* x;      // [rows x ldx], the first Kx columns are the input features
    MlpChainArgs chain[HB_MLP_MAX_CHAINS];
};
size_t mlp_packed_floats(int K, int n_pad);
int launch_mlp_pack(const float* W, const float* bias, int n_true, int n_pad, int K, float* packed, cudaStream_t stream);
int launch_mlp_chains(const MlpArgs& a, cudaStream_t stream);
struct MlpBwdOut {       // gradient of one head output: dL/dout[row * ld + (col - col0)]
    const float* dptr;   // null: no gradient for this output
    int ld, col0, ncols, act;
};
struct MlpBwdLayerArgs {
    int K, N;               // N padded to a multiple of 32
    const float* packed_t;  // launch_mlp_pack_t of the layer's weight
    const float* z_prev;    // [rows x K] pre-activations of the layer below (null for the first layer)
    float* dWt;             // [K x N] += dW^T (zeroed by the caller)
    float* db;              // [N] += db
};
struct MlpBwdChainArgs {
    int n_layers, n_outs;
    MlpBwdLayerArgs layer[HB_MLP_MAX_LAYERS];
    MlpBwdOut out[HB_MLP_MAX_OUTS];
    const float* z_head;    // [rows x N_head] pre-activations of the last layer
    float* dx;              // [rows x Kx] this chain's share of dL/dx (or null)
};
struct MlpBwdArgs {
    int rows, Kx, ldx, n_chains;
    const float* x;
    MlpBwdChainArgs chain[HB_MLP_MAX_CHAINS];
};
int launch_mlp_pack_t(const float* W, int n_true, int k_true, int n_pad, float* packed, cudaStream_t stream);
int launch_mlp_chains_backward(const MlpBwdArgs& a, cudaStream_t stream);

int launch_lbs_forward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                       const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                       float* out_xyz, float* out_rotq, float* out_scales, float* out_T, cudaStream_t stream);
int launch_lbs_backward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                        const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                        const float* g_xyz, const float* g_rotq, const float* g_scales, float* dA,
                        float* d_xyz, float* d_weights, float* d_rotmat, float* d_scales, float* d_transl,
                        float* d_smpl_scale, cudaStream_t stream);

// ---- image-space losses (image_loss.cu) -------------------------------------------------------
int launch_l1_ssim_forward(int C, int H, int W, const float* img1, const float* img2, float* dmap, double* sums,
                           cudaStream_t stream);
int launch_l1_ssim_backward(int C, int H, int W, const float* img1, const float* img2, const float* dmap,
                            const float* g, float* dL_dimg1, cudaStream_t stream);

// ---- 3-nearest-neighbour mean squared distance (knn.cu) -----------------------------------------
size_t knn_temp_bytes(int P);
int launch_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* temp, size_t temp_bytes,
                          cudaStream_t stream);

// ---- renderer post-ops (post_ops.cu) ------------------------------------------------------------
int launch_densification_stats(int n, const int* radii, const float* grad, int grad_stride, float* grad_accum,
                               float* denom, float* max_radii, cudaStream_t stream);

// ---- tri-plane sampling (triplane.cu) ------------------------------------------------------------
int launch_triplane_forward(int N, int C, int resX, int resY, int resZ, const float* plane_xy, const float* plane_xz,
                            const float* plane_yz, const float* xyz, float center, float scale, float eps,
                            float* out, int* flag, cudaStream_t stream);
int launch_triplane_backward(int N, int C, int resX, int resY, int resZ, const float* plane_xy, const float* plane_xz,
                             const float* plane_yz, const float* xyz, float center, float scale, const float* gout,
                             float* g_xy, float* g_xz, float* g_yz, float* gxyz, cudaStream_t stream);

// ---- device helpers ------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ float4 ldg_f4(const float4* p) { return __ldg(p); }
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// Spherical-harmonics constants (same values as cuda_rasterizer/auxiliary.h:22-39 and
// hugs/utils/spherical_harmonics.py:30-47).
#define HB_SH_C0 0.28209479177387814f
#define HB_SH_C1 0.4886025119029199f
#define HB_SH_C2_0 1.0925484305920792f
#define HB_SH_C2_1 (-1.0925484305920792f)
#define HB_SH_C2_2 0.31539156525252005f
#define HB_SH_C2_3 (-1.0925484305920792f)
#define HB_SH_C2_4 0.5462742152960396f
#define HB_SH_C3_0 (-0.5900435899266435f)
#define HB_SH_C3_1 2.890611442640554f
#define HB_SH_C3_2 (-0.4570457994644658f)
#define HB_SH_C3_3 0.3731763325901154f
#define HB_SH_C3_4 (-0.4570457994644658f)
#define HB_SH_C3_5 1.445305721320277f
#define HB_SH_C3_6 (-0.5900435899266435f)
#endif

}  // namespace hb
