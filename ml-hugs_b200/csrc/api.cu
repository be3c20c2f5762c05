// C ABI of libhugs_b200.so (declared in include/hugs_b200.h) and the forward / backward stage
// orchestration that replaces CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (cuda_rasterizer/rasterizer_impl.cu:141-153,198-336,340-434).
#include "../../include/hugs_b200.h"
#include "common.cuh"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

namespace hb {

static thread_local char g_error[512] = "";
// Process-wide: torch.autograd runs the backward call on its own worker thread.
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- optional per-stage device timing ------------------------------------------------------
// When enabled (hugs_b200_stage_timing_enable), every stage is bracketed by a pair of CUDA events
// recorded on the stream the stage is launched on; hugs_b200_stage_timing_read drains the pairs.
// bench.py uses this for the per-stage roofline numbers.  Off by default: no events, no cost.
enum Stage {
    ST_PREPROCESS = 0,
    ST_SCAN,
    ST_DUPLICATE,
    ST_SORT,
    ST_RANGES,
    ST_RENDER_FWD,
    ST_RENDER_BWD,
    ST_PREPROCESS_BWD,
    ST_LBS_FWD,
    ST_LBS_BWD,
    ST_TILE_SPLIT,
    ST_COUNT
};
static_assert(ST_COUNT == HUGS_B200_NUM_STAGES, "stage table out of sync with the header");
static const char* const kStageNames[ST_COUNT] = {
    "preprocess", "inclusive_scan", "duplicate_with_keys", "sort_pairs", "identify_tile_ranges",
    "render_forward", "render_backward", "preprocess_backward", "lbs_forward", "lbs_backward",
    "tile_split"};

struct StageTimer {
    std::mutex mu;
    std::atomic<int> enabled{0};
    struct Pair {
        int stage, dev;
        cudaEvent_t a, b;
    };
    std::vector<Pair> pending;
    // CUDA events belong to the device that was current when they were created: one pool per device
    static constexpr int kMaxDevices = 64;
    std::vector<cudaEvent_t> free_events[kMaxDevices];
    double ms[ST_COUNT] = {0};
    long long calls[ST_COUNT] = {0};

    static int device_slot() {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -1;
        return dev;
    }
    cudaEvent_t get(int dev) {
        if (dev < 0) return nullptr;
        if (!free_events[dev].empty()) {
            cudaEvent_t e = free_events[dev].back();
            free_events[dev].pop_back();
            return e;
        }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
};
static StageTimer g_timer;

struct StageScope {
    int stage;
    cudaStream_t stream;
    cudaEvent_t a = nullptr;
    int dev = -1;
    StageScope(int st, cudaStream_t s) : stage(st), stream(s) {
        if (g_timer.enabled.load(std::memory_order_relaxed)) {
            std::lock_guard<std::mutex> lk(g_timer.mu);
            dev = StageTimer::device_slot();
            a = g_timer.get(dev);
            if (a) cudaEventRecord(a, stream);
        }
    }
    ~StageScope() {
        if (a) {
            std::lock_guard<std::mutex> lk(g_timer.mu);
            cudaEvent_t b = g_timer.get(dev);
            if (b) {
                cudaEventRecord(b, stream);
                g_timer.pending.push_back({stage, dev, a, b});
            } else {
                g_timer.free_events[dev].push_back(a);
            }
        }
    }
};

// ---- workspace carving ---------------------------------------------------------------------
// All slices are 256-byte aligned relative to the buffer base (the base itself comes from the
// caller's allocator: torch CUDA allocations are 512-byte aligned).
template <typename T>
static T* take(char*& p, size_t count) {
    T* r = reinterpret_cast<T*>(p);
    p += align_up(sizeof(T) * count);
    return r;
}

GeomView carve_geom(char* base, int P) {
    GeomView g;
    char* p = base;
    const size_t n = (size_t)(P > 0 ? P : 0);
    g.depths = take<float>(p, n);
    g.means2D = take<float2>(p, n);
    g.conic_opacity = take<float4>(p, n);
    g.rgb = take<float4>(p, n);
    g.clamped = take<uint8_t>(p, n);
    g.tiles_touched = take<uint32_t>(p, n);
    g.point_offsets = take<uint32_t>(p, n);
    g.scan_temp_bytes = scan_temp_bytes(P);
    g.scan_temp = take<char>(p, g.scan_temp_bytes);
    g.sorted_depths = take<uint32_t>(p, n);
    g.perm = take<uint32_t>(p, n);
    g.dsort_keys_tmp = take<uint32_t>(p, n);
    g.dsort_vals_tmp = take<uint32_t>(p, n);
    g.dsort_temp_bytes = sort_temp_bytes(P);
    g.dsort_temp = take<char>(p, g.dsort_temp_bytes);
    // appended after every exported offset
    g.grad_pack = take<float>(p, 8 * n);
    g.grad_opac = take<float>(p, n);   // right behind grad_pack: one fill clears both
    g.rects = take<uint2>(p, n);
    g.st_touched = take<uint32_t>(p, n);
    g.pair_counts = take<uint32_t>(p, 2);
    g.total = (size_t)(p - base);
    return g;
}

ImageView carve_image(char* base, int W, int H) {
    ImageView v;
    char* p = base;
    const size_t n = (size_t)W * (size_t)H;
    const size_t tiles = (size_t)cdiv(W, HB_TILE_X) * (size_t)cdiv(H, HB_TILE_Y);
    v.accum_alpha = take<float>(p, n);
    v.n_contrib = take<uint32_t>(p, n);
    v.ranges = take<uint2>(p, tiles);
    v.clamp_mask = take<uint8_t>(p, n);   // appended after every exported offset
    v.total = (size_t)(p - base);
    return v;
}

BinningView carve_binning(char* base, int R, int grid_x, int grid_y) {
    BinningView b;
    char* p = base;
    const size_t n = (size_t)(R > 0 ? R : 0);
    const int num_tiles = grid_x * grid_y;
    b.list = take<uint32_t>(p, n);  // first: the backward pass finds it without knowing the capacity
    b.keys = take<uint64_t>(p, n);
    b.keys_unsorted = take<uint64_t>(p, n);
    b.list_unsorted = take<uint32_t>(p, n);
    b.sort_temp_bytes = sort_temp_bytes(R);
    b.sort_temp = take<char>(p, b.sort_temp_bytes);
    b.l2_temp_bytes = carve_l2_temp(nullptr, R > 0 ? R : 0, super_grid_x(grid_x) * super_grid_y(grid_y), num_tiles).total;
    b.l2_temp = take<char>(p, b.l2_temp_bytes);
    // last, so that no other offset depends on the tile count
    b.hit_words = take<uint32_t>(p, ((n + 31) / 32 + (size_t)(num_tiles > 0 ? num_tiles : 0) + 1) * 8);
    b.total = (size_t)(p - base);
    return b;
}

LayerView carve_layer(char* base, int R, int n_layer, int W, int H) {
    LayerView v;
    char* p = base;
    const size_t n = (size_t)(R > 0 ? R : 0);
    const size_t pix = (size_t)W * (size_t)H;
    const size_t tiles = (size_t)cdiv(W, HB_TILE_X) * (size_t)cdiv(H, HB_TILE_Y);
    v.list = take<uint32_t>(p, n);
    v.hit_words = take<uint32_t>(p, ((n + 31) / 32 + tiles + 1) * 8);
    v.accum_alpha = take<float>(p, pix);
    v.n_contrib = take<uint32_t>(p, pix);
    v.ranges = take<uint2>(p, tiles);
    v.grad_pack = take<float>(p, 8 * (size_t)(n_layer > 0 ? n_layer : 0));
    v.clamp_mask = take<uint8_t>(p, pix);
    v.total = (size_t)(p - base);
    return v;
}

// Helper function to find the next-highest bit of the MSB (rasterizer_impl.cu:35-50): the
// reference sorts on 32 + this many key bits.
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step;
        else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// Pinned 4-byte landing pad for the number of rendered pairs (one per host thread).
static int* pinned_counter() {
    static thread_local int* p = nullptr;
    if (p == nullptr) {
        if (cudaHostAlloc(reinterpret_cast<void**>(&p), 64, cudaHostAllocDefault) != cudaSuccess) p = nullptr;
    }
    return p;
}

// Event that marks "the pair count has landed in the pinned counter": one per host thread and device
// (an event can only be recorded on a stream of the device it was created on).
static cudaEvent_t count_event() {
    static thread_local cudaEvent_t ev[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (ev[dev] == nullptr) {
        if (cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming) != cudaSuccess) ev[dev] = nullptr;
    }
    return ev[dev];
}

static std::atomic<int> g_binning_mode{HUGS_B200_BINNING_ASYNC};

// launch priority of the blend kernels: unset (plain launches) unless hugs_b200_set_blend_priority says otherwise
static std::atomic<int> g_blend_prio_set{0};
static std::atomic<int> g_blend_prio{0};
bool blend_priority(int* priority) {
    *priority = g_blend_prio.load(std::memory_order_relaxed);
    return g_blend_prio_set.load(std::memory_order_relaxed) != 0;
}

// Pair-count history per (P, W, H): capacity = recent maximum + 25 % slack, decaying slowly.
struct CapacityHint {
    int P, W, H, cap, cap1;   // cap: (Gaussian, tile) pairs; cap1: level-1 (Gaussian, super-tile) entries
    unsigned long long stamp;
};
static thread_local CapacityHint g_hints[8];
static thread_local unsigned long long g_hint_clock = 0;

static int capacity_hint(int P, int W, int H, int* cap1) {
    for (CapacityHint& h : g_hints)
        if (h.cap > 0 && h.P == P && h.W == W && h.H == H) {
            h.stamp = ++g_hint_clock;
            *cap1 = h.cap1;
            return h.cap;
        }
    return 0;
}

static int with_slack(int R) {
    const long long want_ll = (long long)R + R / 4 + 4096;
    return want_ll > 0x7fff0000LL ? 0x7fff0000 : (int)want_ll;
}

static void update_capacity_hint(int P, int W, int H, int R, int R1) {
    const int want = with_slack(R), want1 = with_slack(R1);
    CapacityHint* slot = nullptr;
    for (CapacityHint& h : g_hints)
        if (h.cap > 0 && h.P == P && h.W == W && h.H == H) slot = &h;
    if (slot == nullptr) {
        slot = &g_hints[0];
        for (CapacityHint& h : g_hints)
            if (h.stamp < slot->stamp) slot = &h;
        slot->cap = 0;
        slot->cap1 = 0;
    }
    const int decayed = (int)(slot->cap * 0.98), decayed1 = (int)(slot->cap1 * 0.98);
    slot->P = P;
    slot->W = W;
    slot->H = H;
    slot->cap = want > decayed ? want : decayed;
    slot->cap1 = want1 > decayed1 ? want1 : decayed1;
    slot->stamp = ++g_hint_clock;
}

static Camera make_camera(const float* view, const float* proj, const float* campos, int W, int H, float tan_fovx,
                          float tan_fovy) {
    Camera c;
    c.view = view;
    c.proj = proj;
    c.campos = campos;
    c.W = W;
    c.H = H;
    c.tan_fovx = tan_fovx;
    c.tan_fovy = tan_fovy;
    // rasterizer_impl.cu:222-223
    c.focal_y = H / (2.0f * tan_fovy);
    c.focal_x = W / (2.0f * tan_fovx);
    c.grid_x = cdiv(W, HB_TILE_X);
    c.grid_y = cdiv(H, HB_TILE_Y);
    return c;
}

#define HB_STAGE(call, stage_id)                                                              \
    do {                                                                                      \
        const char* name = kStageNames[stage_id];                                             \
        int _rc;                                                                              \
        {                                                                                     \
            StageScope _scope(stage_id, stream);                                              \
            _rc = (call);                                                                     \
        }                                                                                     \
        if (_rc) return _rc;                                                                  \
        if (debug) {                                                                          \
            cudaError_t _e = cudaStreamSynchronize(stream);                                   \
            if (_e != cudaSuccess) {                                                          \
                set_error("stage '%s' failed: %s", name, cudaGetErrorString(_e));             \
                return (int)_e;                                                               \
            }                                                                                 \
        }                                                                                     \
    } while (0)

}  // namespace hb

using namespace hb;

extern "C" {

int hugs_b200_abi_version(void) { return HUGS_B200_ABI_VERSION; }
const char* hugs_b200_last_error(void) { return g_error; }

long long hugs_b200_launch_count(int reset) {
    return reset ? g_launches.exchange(0) : g_launches.load();
}

int hugs_b200_set_binning_mode(int mode) {
    if (mode < HUGS_B200_BINNING_ASYNC || mode > HUGS_B200_BINNING_TILESORT) return HUGS_B200_EINVAL;
    return g_binning_mode.exchange(mode);
}

static int tiles_of(int width, int height) { return cdiv(width, HB_TILE_X) * cdiv(height, HB_TILE_Y); }
static BinningView carve_binning_wh(char* base, int R, int width, int height) {
    return carve_binning(base, R, cdiv(width, HB_TILE_X), cdiv(height, HB_TILE_Y));
}

size_t hugs_b200_binning_capacity_of(size_t nbytes, int width, int height) {
    // largest capacity whose layout fits in nbytes (the layout size is monotone in the capacity)
    if (width <= 0 || height <= 0) return 0;
    size_t lo = 0, hi = nbytes / 24 + 1;
    while (lo < hi) {
        const size_t mid = lo + (hi - lo + 1) / 2;
        if (mid <= 0x7fffffffULL && carve_binning_wh(nullptr, (int)mid, width, height).total <= nbytes) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

int hugs_b200_stage_timing_enable(int on) { return g_timer.enabled.exchange(on ? 1 : 0); }

const char* hugs_b200_stage_name(int stage) {
    return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : nullptr;
}

int hugs_b200_stage_timing_read(double* ms_sum, long long* calls, int reset) {
    std::lock_guard<std::mutex> lk(g_timer.mu);
    for (const StageTimer::Pair& p : g_timer.pending) {
        float ms = 0.f;
        if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
            g_timer.ms[p.stage] += ms;
            g_timer.calls[p.stage] += 1;
        }
        g_timer.free_events[p.dev].push_back(p.a);
        g_timer.free_events[p.dev].push_back(p.b);
    }
    g_timer.pending.clear();
    for (int i = 0; i < ST_COUNT; i++) {
        if (ms_sum) ms_sum[i] = g_timer.ms[i];
        if (calls) calls[i] = g_timer.calls[i];
        if (reset) {
            g_timer.ms[i] = 0;
            g_timer.calls[i] = 0;
        }
    }
    return ST_COUNT;
}

int hugs_b200_geom_layout_of(int P, hugs_b200_geom_layout* out) {
    if (!out || P < 0) return HUGS_B200_EINVAL;
    const GeomView g = carve_geom(nullptr, P);
    out->depths = (size_t)reinterpret_cast<char*>(g.depths);
    out->means2D = (size_t)reinterpret_cast<char*>(g.means2D);
    out->conic_opacity = (size_t)reinterpret_cast<char*>(g.conic_opacity);
    out->rgb = (size_t)reinterpret_cast<char*>(g.rgb);
    out->clamped = (size_t)reinterpret_cast<char*>(g.clamped);
    out->tiles_touched = (size_t)reinterpret_cast<char*>(g.tiles_touched);
    out->point_offsets = (size_t)reinterpret_cast<char*>(g.point_offsets);
    out->scan_temp = (size_t)g.scan_temp;
    out->sorted_depths = (size_t)reinterpret_cast<char*>(g.sorted_depths);
    out->perm = (size_t)reinterpret_cast<char*>(g.perm);
    out->total = g.total;
    return 11;
}

int hugs_b200_image_layout_of(int width, int height, hugs_b200_image_layout* out) {
    if (!out || width <= 0 || height <= 0) return HUGS_B200_EINVAL;
    const ImageView v = carve_image(nullptr, width, height);
    out->accum_alpha = (size_t)reinterpret_cast<char*>(v.accum_alpha);
    out->n_contrib = (size_t)reinterpret_cast<char*>(v.n_contrib);
    out->ranges = (size_t)reinterpret_cast<char*>(v.ranges);
    out->total = v.total;
    return 4;
}

int hugs_b200_binning_layout_of(int R, int width, int height, hugs_b200_binning_layout* out) {
    if (!out || R < 0 || width <= 0 || height <= 0) return HUGS_B200_EINVAL;
    const BinningView b = carve_binning_wh(nullptr, R, width, height);
    out->point_list_keys = (size_t)reinterpret_cast<char*>(b.keys);
    out->point_list = (size_t)reinterpret_cast<char*>(b.list);
    out->point_list_keys_unsorted = (size_t)reinterpret_cast<char*>(b.keys_unsorted);
    out->point_list_unsorted = (size_t)reinterpret_cast<char*>(b.list_unsorted);
    out->sort_temp = (size_t)b.sort_temp;
    out->hit_words = (size_t)reinterpret_cast<char*>(b.hit_words);
    out->total = b.total;
    return 7;
}

int hugs_b200_sort_bits(int width, int height) {
    const uint32_t tiles = (uint32_t)(cdiv(width, HB_TILE_X) * cdiv(height, HB_TILE_Y));
    return 32 + (int)higher_msb(tiles);
}

size_t hugs_b200_sort_temp_bytes(int n) { return sort_temp_bytes(n); }
size_t hugs_b200_scan_temp_bytes(int n) { return scan_temp_bytes(n); }

int hugs_b200_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                         int end_bit, void* temp, size_t temp_bytes, hugs_b200_stream_t stream) {
    if (n < 0 || end_bit < 0 || end_bit > 64) {
        set_error("sort_pairs: bad n / end_bit");
        return HUGS_B200_EINVAL;
    }
    if (n > 0 && (!keys_in || !vals_in || !keys_out || !vals_out)) {
        set_error("sort_pairs: null buffer");
        return HUGS_B200_EINVAL;
    }
    return launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, n, end_bit, temp, temp_bytes,
                             static_cast<cudaStream_t>(stream));
}

int hugs_b200_sort_pairs_range(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                               int begin_bit, int end_bit, void* temp, size_t temp_bytes,
                               hugs_b200_stream_t stream) {
    if (n < 0 || begin_bit < 0 || end_bit < begin_bit || end_bit > 64) {
        set_error("sort_pairs_range: bad n / bit range");
        return HUGS_B200_EINVAL;
    }
    if (n > 0 && (!keys_in || !vals_in || !keys_out || !vals_out)) {
        set_error("sort_pairs_range: null buffer");
        return HUGS_B200_EINVAL;
    }
    if (end_bit == begin_bit && n > 0) {
        HB_CUDA_OK(cudaMemcpyAsync(keys_out, keys_in, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToDevice,
                                   static_cast<cudaStream_t>(stream)));
        HB_CUDA_OK(cudaMemcpyAsync(vals_out, vals_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice,
                                   static_cast<cudaStream_t>(stream)));
        return 0;
    }
    return launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, n, end_bit, temp, temp_bytes,
                             static_cast<cudaStream_t>(stream), begin_bit);
}

int hugs_b200_inclusive_scan(const uint32_t* in, uint32_t* out, int n, void* temp, size_t temp_bytes,
                             hugs_b200_stream_t stream) {
    if (n < 0 || (n > 0 && (!in || !out))) {
        set_error("inclusive_scan: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_inclusive_scan(in, out, n, temp, temp_bytes, static_cast<cudaStream_t>(stream));
}

int hugs_b200_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                           uint8_t* present, hugs_b200_stream_t stream) {
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) {
        set_error("mark_visible: bad arguments");
        return HUGS_B200_EINVAL;
    }
    if (P == 0) return 0;
    return launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, static_cast<cudaStream_t>(stream));
}

// Second input segment of the *_seg2 entry points: rows [P0, P) of the per-Gaussian inputs come from a
// second set of arrays.  Every per-Gaussian OUTPUT of the preprocess kernels is a plain array indexed by
// the Gaussian, so a segment is just one more launch of the same kernel with the output pointers moved to
// the segment's first row; everything downstream (scan, binning, sort, blend) sees one Gaussian range.
struct SecondSegment {
    int P0;
    const float* means3D;
    const float* shs;
    const float* opacities;
    const float* scales;
    const float* rotations;
};
static thread_local const SecondSegment* tl_seg2 = nullptr;   // set only for the duration of a *_seg2 call

// Second image of the first `n` Gaussians (hugs_b200_layer_forward): its blend backward runs between the
// main blend backward and the per-Gaussian backward of the *_layer entry points.
struct LayerBackward {
    int n;
    const float* background;
    char* buffer;
    size_t bytes;
    const float* dL_dpix;
};
static thread_local const LayerBackward* tl_layer = nullptr;   // set only for the duration of a *_layer call
static thread_local const hugs_b200_grad_sinks* tl_sinks = nullptr;   // likewise, for *_seg2_ex

static GeomView geom_rows_from(const GeomView& g, int row0) {
    GeomView o = g;
    o.depths += row0;
    o.means2D += row0;
    o.conic_opacity += row0;
    o.rgb += row0;
    o.clamped += row0;
    o.tiles_touched += row0;
    o.rects += row0;
    o.st_touched += row0;
    return o;
}

int hugs_b200_rasterize_forward(hugs_b200_resize_fn geometry_buffer, void* geometry_user,
                                hugs_b200_resize_fn binning_buffer, void* binning_user,
                                hugs_b200_resize_fn image_buffer, void* image_user, int P, int D, int M,
                                const float* background, int width, int height, const float* means3D,
                                const float* shs, const float* colors_precomp, const float* opacities,
                                const float* scales, float scale_modifier, const float* rotations,
                                const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered,
                                float* out_color, int* radii, int debug, hugs_b200_stream_t stream_,
                                int* num_rendered) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool clamp01 = (debug & HUGS_B200_FLAG_CLAMP01) != 0;   // `debug` doubles as a flag word
    // segment 1's rows of the geometry workspace (and its share of the pair counts) are already there:
    // hugs_b200_preprocess_scene_views
    const bool seg1_done = (debug & HUGS_B200_FLAG_SEG1_PREPROCESSED) != 0;
    debug &= HUGS_B200_FLAG_DEBUG;
    if (num_rendered) *num_rendered = 0;
    if (!geometry_buffer || !binning_buffer || !image_buffer || !num_rendered) {
        set_error("rasterize_forward: null callback / output");
        return HUGS_B200_EINVAL;
    }
    if (P < 0 || width <= 0 || height <= 0 || !background || !out_color || !viewmatrix || !projmatrix || !cam_pos) {
        set_error("rasterize_forward: bad size or null camera / output pointer");
        return HUGS_B200_EINVAL;
    }
    if (P > 0 && (!means3D || !opacities || !radii)) {
        set_error("rasterize_forward: null means3D / opacities / radii");
        return HUGS_B200_EINVAL;
    }
    if (P > 0 && colors_precomp == nullptr) {
        if (shs == nullptr) {
            set_error("rasterize_forward: provide exactly one of SHs or precomputed colors");
            return HUGS_B200_EINVAL;
        }
        if (D < 0 || D > 3 || (D + 1) * (D + 1) > M) {
            set_error("rasterize_forward: SH degree %d needs 0 <= D <= 3 and M >= (D+1)^2 (M=%d)", D, M);
            return HUGS_B200_EUNSUP;
        }
    }
    if (P > 0 && cov3D_precomp == nullptr && (scales == nullptr || rotations == nullptr)) {
        set_error("rasterize_forward: provide scale/rotation pair or precomputed 3D covariance");
        return HUGS_B200_EINVAL;
    }
    const Camera cam = make_camera(viewmatrix, projmatrix, cam_pos, width, height, tan_fovx, tan_fovy);
    const int num_tiles = cam.grid_x * cam.grid_y;

    char* img_base = image_buffer(image_user, carve_image(nullptr, width, height).total);
    if (!img_base) {
        set_error("rasterize_forward: image buffer callback returned NULL");
        return HUGS_B200_ENOMEM;
    }
    const ImageView img = carve_image(img_base, width, height);

    int R = 0;
    GeomView geom;
    BinningView bin;
    memset(&geom, 0, sizeof(geom));
    memset(&bin, 0, sizeof(bin));
    const int tile_bits = (int)higher_msb((uint32_t)num_tiles);
    const int mode = g_binning_mode.load(std::memory_order_relaxed);
    // Both two-level structures start from a depth argsort of the Gaussians and need tile grids of at most
    // 1023 x 1023 (packed rectangles); otherwise fall back to the reference's structure.
    //   multisplit (ASYNC / SYNC, the default): super-tile split + mask transposition (binning2.cu)
    //   tilesort   (TILESORT, round 1): radix sort of the R pairs on their tile bits
    const bool fits = cam.grid_x <= 1023 && cam.grid_y <= 1023;
    const bool multisplit = fits && (mode == HUGS_B200_BINNING_ASYNC || mode == HUGS_B200_BINNING_SYNC);
    const bool tilesort = fits && mode == HUGS_B200_BINNING_TILESORT;
    const bool two_level = tilesort;
    const bool async = (mode == HUGS_B200_BINNING_ASYNC || mode == HUGS_B200_BINNING_TILESORT) && !debug;
    const bool tile16 = num_tiles <= 65536;
    const int gx_s = super_grid_x(cam.grid_x), gy_s = super_grid_y(cam.grid_y), n_st = gx_s * gy_s;
    uint32_t* R_dev = nullptr;   // tilesort: pair count; multisplit: level-1 entry count
    int* pinned = pinned_counter();
    cudaEvent_t ev = count_event();
    bool capturing = false;   // CUDA-graph capture in progress on the caller's stream
    {
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &st) == cudaSuccess) capturing = (st == cudaStreamCaptureStatusActive);
    }
    if (P > 0) {
        if (!pinned || !ev) {
            set_error("rasterize_forward: could not create the pinned counter / event");
            return HUGS_B200_ENOMEM;
        }
        char* geom_base = geometry_buffer(geometry_user, carve_geom(nullptr, P).total);
        if (!geom_base) {
            set_error("rasterize_forward: geometry buffer callback returned NULL");
            return HUGS_B200_ENOMEM;
        }
        geom = carve_geom(geom_base, P);
        if (seg1_done && !tl_seg2) {
            set_error("rasterize_forward: HUGS_B200_FLAG_SEG1_PREPROCESSED needs a two-segment call");
            return HUGS_B200_EINVAL;
        }
        if (!seg1_done) HB_CUDA_OK(cudaMemsetAsync(geom.pair_counts, 0, 2 * sizeof(uint32_t), stream));

        if (const SecondSegment* s2 = tl_seg2) {
            // both segments in ONE launch (SegBatch): the human segment alone would be a 431-block grid
            auto both = [&]() -> int {
                SegBatch batch;
                int rc = 0;
                if (s2->P0 > 0)
                    rc = launch_preprocess(s2->P0, D, M, means3D, scales, scale_modifier, rotations, opacities, shs,
                                           nullptr, nullptr, cam, radii, geom, prefiltered != 0, stream, &batch);
                if (rc == 0 && P - s2->P0 > 0 && !seg1_done)
                    rc = launch_preprocess(P - s2->P0, D, M, s2->means3D, s2->scales, scale_modifier, s2->rotations,
                                           s2->opacities, s2->shs, nullptr, nullptr, cam, radii + s2->P0,
                                           geom_rows_from(geom, s2->P0), prefiltered != 0, stream, &batch);
                return rc ? rc : flush_preprocess(batch, stream);
            };
            HB_STAGE(both(), ST_PREPROCESS);
        } else
        HB_STAGE(launch_preprocess(P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs,
                                   (scales && rotations) ? nullptr : cov3D_precomp, colors_precomp, cam, radii, geom,
                                   prefiltered != 0, stream),
                 ST_PREPROCESS);
        // The number of (Gaussian, tile) pairs must reach the host (rasterizer_impl.cu:281).  The preprocess
        // kernel has summed it (and the level-1 entry count): async copy into pinned memory on the caller's
        // stream + an event, instead of a blocking cudaMemcpy after the scan.
        if (!capturing) {
            HB_CUDA_OK(cudaMemcpyAsync(pinned, geom.pair_counts, 2 * sizeof(int), cudaMemcpyDeviceToHost, stream));
            HB_CUDA_OK(cudaEventRecord(ev, stream));
        }
        if (multisplit || tilesort) {
            HB_STAGE(launch_depth_argsort(reinterpret_cast<const uint32_t*>(geom.depths), geom.dsort_keys_tmp,
                                          geom.dsort_vals_tmp, geom.sorted_depths, geom.perm, P, geom.dsort_temp,
                                          geom.dsort_temp_bytes, stream),
                     ST_SORT);
        }
        HB_STAGE(launch_inclusive_scan(multisplit ? geom.st_touched : geom.tiles_touched, geom.point_offsets, P,
                                       geom.scan_temp, geom.scan_temp_bytes, stream,
                                       (multisplit || tilesort) ? geom.perm : nullptr),
                 ST_SCAN);
        R_dev = geom.point_offsets + (P - 1);
    }
    int R1 = 0;
    bool R_known = (P == 0);
    auto wait_R = [&]() -> int {
        HB_CUDA_OK(cudaEventSynchronize(ev));
        R = pinned[0];
        R1 = pinned[1];
        R_known = true;
        if (R < 0) {
            set_error("rasterize_forward: more than 2^31 (Gaussian, tile) pairs");
            return HUGS_B200_EUNSUP;
        }
        return 0;
    };
    // Workspace capacity: in the default mode it comes from the pair counts of earlier calls with the
    // same (P, W, H), so everything below is enqueued before the count has arrived; the count is
    // checked before returning and an overflow is redone with the exact size.
    int cap = 0, cap1 = 0;
    if (P > 0 && (multisplit || tilesort) && async) cap = capacity_hint(P, width, height, &cap1);
    // Under CUDA-graph capture the host cannot wait for the count at all: the call is enqueued for the
    // remembered capacity and reports it as num_rendered (an upper bound; every kernel reads the real count on
    // the device, lists beyond the capacity are truncated).  One uncaptured call with the same (P, W, H) must
    // have gone before, so that a capacity is known.
    if (capturing) {
        if (cap <= 0 && P > 0) {
            set_error("rasterize_forward: stream capture needs a known pair capacity -- run the same (P, W, H) once "
                      "outside the capture first (binning mode ASYNC)");
            return HUGS_B200_EUNSUP;
        }
        R = cap;
        R1 = cap1;
        R_known = true;
    }
    if (cap <= 0 && !R_known) {
        const int rc = wait_R();
        if (rc) return rc;
        cap = R;
        cap1 = R1;
    }
    const float* feature_ptr = colors_precomp != nullptr ? colors_precomp : reinterpret_cast<const float*>(geom.rgb);
    const int stride = colors_precomp != nullptr ? 3 : 4;
    if (P == 0) feature_ptr = nullptr;
    for (;;) {
        char* bin_base = binning_buffer(binning_user, carve_binning(nullptr, cap, cam.grid_x, cam.grid_y).total);
        if (!bin_base) {
            set_error("rasterize_forward: binning buffer callback returned NULL");
            return HUGS_B200_ENOMEM;
        }
        bin = carve_binning(bin_base, cap, cam.grid_x, cam.grid_y);
        if (cap1 > cap) cap1 = cap;   // a level-1 entry holds at least one pair; the buffers are sized by cap
        const uint32_t* n_dev = nullptr;
        if (cap > 0 && multisplit) {
            // Level 1: (super-tile << 32 | Gaussian) entries in depth order, stably sorted on the super-tile
            // bits (one 8-bit pass at 1080p).  The entries ping-pong between the two 64-bit regions; the last
            // pass leaves the Gaussian list in `list_unsorted` and the super-tile of every position (u16) in
            // the then dead `keys_unsorted` region.
            const SortPassPlan plan = tile_pass_plan((int)higher_msb((uint32_t)n_st));
            uint64_t* const buf_a = bin.keys_unsorted;
            uint64_t* const buf_b = bin.keys;
            uint64_t* first_buf = (plan.num_passes % 2 == 0) ? buf_a : buf_b;
            const L2Temp l2 = carve_l2_temp(bin.l2_temp, cap, n_st, num_tiles);
            auto emit = [&]() -> int {
                const int rc = sort_temp_clear(bin.sort_temp, bin.sort_temp_bytes, cap1, stream);
                if (rc) return rc;
                return launch_duplicate_super(P, geom.perm, geom.point_offsets, geom.rects, first_buf, gx_s, gy_s,
                                              cap1, plan, sort_temp_histogram(bin.sort_temp), stream);
            };
            HB_STAGE(emit(), ST_DUPLICATE);
            HB_STAGE(launch_tile_sort(buf_a, buf_b, bin.keys_unsorted, bin.list_unsorted, true, cap1, R_dev, plan,
                                      bin.sort_temp, bin.sort_temp_bytes, stream),
                     ST_SORT);
            // Level 2: super-tile lists -> tile lists and ranges.
            auto split = [&]() -> int {
                const int rc = launch_tile_ranges_ids(cap1, bin.keys_unsorted, true, l2.st_ranges, n_st, stream, R_dev);
                if (rc) return rc;
                return launch_l2_split(l2, bin.list_unsorted, geom.rects, n_st, gx_s, cam.grid_x, cam.grid_y, cap1, cap,
                                       bin.list, img.ranges, stream);
            };
            HB_STAGE(split(), ST_TILE_SPLIT);
        } else if (cap > 0 && two_level) {
            n_dev = R_dev;
            const SortPassPlan plan = tile_pass_plan(tile_bits);
            // Packed (tile << 32 | index) pairs ping-pong between the two 64-bit regions; the last pass
            // reads from `keys` and leaves the sorted list plus the tile id of every list position
            // (in the then dead `keys_unsorted` region).  The sorted 64-bit keys themselves are not
            // needed by anything downstream; hugs_b200_materialize_sorted_keys writes them on demand.
            uint64_t* const buf_a = bin.keys_unsorted;
            uint64_t* const buf_b = bin.keys;
            uint64_t* first_buf = (plan.num_passes % 2 == 0) ? buf_a : buf_b;
            auto emit = [&]() -> int {
                const int rc = sort_temp_clear(bin.sort_temp, bin.sort_temp_bytes, cap, stream);
                if (rc) return rc;
                return launch_duplicate_sorted(P, geom.perm, geom.point_offsets, geom.means2D, radii, first_buf,
                                               cam.grid_x, cam.grid_y, cap, plan, sort_temp_histogram(bin.sort_temp),
                                               stream);
            };
            HB_STAGE(emit(), ST_DUPLICATE);
            HB_STAGE(launch_tile_sort(buf_a, buf_b, bin.keys_unsorted, bin.list, tile16, cap, n_dev, plan,
                                      bin.sort_temp, bin.sort_temp_bytes, stream),
                     ST_SORT);
        } else if (cap > 0) {
            // Reference structure: pairs emitted in Gaussian-index order, one sort on all 32+bit key bits.
            const int end_bit = 32 + tile_bits;
            const int passes = cdiv(end_bit, 8);
            uint64_t* k_src = (passes % 2 == 0) ? bin.keys : bin.keys_unsorted;
            uint32_t* v_src = (passes % 2 == 0) ? bin.list : bin.list_unsorted;
            uint64_t* k_dst = (passes % 2 == 0) ? bin.keys_unsorted : bin.keys;
            uint32_t* v_dst = (passes % 2 == 0) ? bin.list_unsorted : bin.list;
            HB_STAGE(launch_duplicate(P, geom.means2D, geom.depths, geom.point_offsets, radii, k_src, v_src, cam.grid_x,
                                      cam.grid_y, stream),
                     ST_DUPLICATE);
            bool in_input = false;
            HB_STAGE(launch_sort_pairs_pingpong(k_src, v_src, k_dst, v_dst, cap, end_bit, bin.sort_temp,
                                                bin.sort_temp_bytes, stream, &in_input),
                     ST_SORT);
            const bool landed_in_keys = in_input ? (k_src == bin.keys) : (k_dst == bin.keys);
            if (!landed_in_keys) {
                HB_CUDA_OK(cudaMemcpyAsync(bin.keys, bin.keys_unsorted, sizeof(uint64_t) * (size_t)cap,
                                           cudaMemcpyDeviceToDevice, stream));
                HB_CUDA_OK(cudaMemcpyAsync(bin.list, bin.list_unsorted, sizeof(uint32_t) * (size_t)cap,
                                           cudaMemcpyDeviceToDevice, stream));
            }
        }
        if (multisplit) {
            if (cap <= 0) HB_CUDA_OK(cudaMemsetAsync(img.ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream));
        } else if (cap > 0 && two_level)
            HB_STAGE(launch_tile_ranges_ids(cap, bin.keys_unsorted, tile16, img.ranges, num_tiles, stream, n_dev),
                     ST_RANGES);
        else
            HB_STAGE(launch_tile_ranges(cap, bin.keys, img.ranges, num_tiles, stream, n_dev), ST_RANGES);
        HB_STAGE(launch_render_forward(img.ranges, bin.list, width, height, cam.grid_x, cam.grid_y,
                                       P > 0 ? geom.means2D : nullptr, feature_ptr, stride,
                                       P > 0 ? geom.conic_opacity : nullptr, img.accum_alpha, img.n_contrib,
                                       background, out_color, bin.hit_words, stream,
                                       clamp01 ? img.clamp_mask : nullptr),
                 ST_RENDER_FWD);
        if (!R_known) {
            const int rc = wait_R();
            if (rc) return rc;
        }
        if (R <= cap && (!multisplit || R1 <= cap1)) break;
        cap = R;  // the hint was too small: redo the binning with the exact size
        cap1 = R1;
    }
    if (P > 0 && !capturing) update_capacity_hint(P, width, height, R, R1);
    *num_rendered = R;
    return 0;
}

int hugs_b200_last_pair_counts(int* pairs, int* level1_entries) {
    // what the last uncaptured forward call of this host thread read back: (Gaussian, tile) pairs and
    // (Gaussian, super-tile) entries of the two-level binning
    int* pinned = pinned_counter();
    if (!pinned) return HUGS_B200_ENOMEM;
    if (pairs) *pairs = pinned[0];
    if (level1_entries) *level1_entries = pinned[1];
    return 0;
}

int hugs_b200_set_blend_priority(int enable, int priority) {
    g_blend_prio.store(priority);
    return g_blend_prio_set.exchange(enable ? 1 : 0);
}

int hugs_b200_materialize_sorted_keys(int P, int R, int width, int height, char* geom_buffer, char* binning_buffer,
                                      size_t binning_bytes, hugs_b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P < 0 || R < 0 || width <= 0 || height <= 0 || (R > 0 && (!geom_buffer || !binning_buffer))) {
        set_error("materialize_sorted_keys: bad arguments");
        return HUGS_B200_EINVAL;
    }
    const int grid_x = cdiv(width, HB_TILE_X), grid_y = cdiv(height, HB_TILE_Y);
    const int mode = g_binning_mode.load(std::memory_order_relaxed);
    const bool fits = grid_x <= 1023 && grid_y <= 1023;
    const bool two_level = fits && mode != HUGS_B200_BINNING_REFERENCE;
    const bool multisplit = fits && (mode == HUGS_B200_BINNING_ASYNC || mode == HUGS_B200_BINNING_SYNC);
    if (!two_level || R == 0) return 0;  // the reference-structure sort leaves the keys in place
    const size_t cap = hugs_b200_binning_capacity_of(binning_bytes, width, height);
    if ((size_t)R > cap) {
        set_error("materialize_sorted_keys: binning buffer of %zu bytes cannot hold %d pairs", binning_bytes, R);
        return HUGS_B200_EINVAL;
    }
    const GeomView geom = carve_geom(geom_buffer, P);
    const BinningView bin = carve_binning(binning_buffer, (int)cap, grid_x, grid_y);
    if (multisplit) {
        const L2Temp l2 = carve_l2_temp(bin.l2_temp, (int)cap, super_grid_x(grid_x) * super_grid_y(grid_y), grid_x * grid_y);
        return launch_keys_from_ranges(grid_x * grid_y, l2.tile_start, l2.tile_total, bin.list,
                                       reinterpret_cast<const uint32_t*>(geom.depths), bin.keys, stream);
    }
    return launch_materialize_keys(R, bin.keys_unsorted, grid_x * grid_y <= 65536, bin.list,
                                   reinterpret_cast<const uint32_t*>(geom.depths), bin.keys, stream);
}

int hugs_b200_rasterize_backward_acc(int P, int D, int M, int R, const float* background, int width, int height,
                                 const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* scales, float scale_modifier, const float* rotations,
                                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                 const float* cam_pos, float tan_fovx, float tan_fovy, const int* radii,
                                 char* geom_buffer, char* binning_buffer, size_t binning_bytes, char* image_buffer,
                                 const float* dL_dpix,
                                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                                 float* dL_drot, int debug, hugs_b200_stream_t stream_, int dL_dsh_accumulate) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool clamp01 = (debug & HUGS_B200_FLAG_CLAMP01) != 0;   // the forward call was made with the same flag
    const bool defer1 = (debug & HUGS_B200_FLAG_DEFER_SEG1) != 0;  // segment 1 is left to hugs_b200_scene_backward_views
    debug &= HUGS_B200_FLAG_DEBUG;
    if (P < 0 || R < 0 || width <= 0 || height <= 0) {
        set_error("rasterize_backward: bad sizes");
        return HUGS_B200_EINVAL;
    }
    if (P == 0) return 0;
    if (!geom_buffer || !binning_buffer || !image_buffer || !dL_dpix || !radii || !means3D || !background ||
        !viewmatrix || !projmatrix || !cam_pos || !dL_dmean2D || !dL_dconic || !dL_dopacity || !dL_dcolor ||
        !dL_dmean3D || !dL_dcov3D) {
        set_error("rasterize_backward: null pointer argument");
        return HUGS_B200_EINVAL;
    }
    if (colors_precomp == nullptr && (shs == nullptr || dL_dsh == nullptr)) {
        set_error("rasterize_backward: SH path needs shs and dL_dsh");
        return HUGS_B200_EINVAL;
    }
    const Camera cam = make_camera(viewmatrix, projmatrix, cam_pos, width, height, tan_fovx, tan_fovy);
    const GeomView geom = carve_geom(geom_buffer, P);
    // The workspace was laid out for a capacity >= R (hugs_b200_set_binning_mode); recover it from the size.
    const size_t cap = hugs_b200_binning_capacity_of(binning_bytes, width, height);
    if ((size_t)R > cap) {
        set_error("rasterize_backward: binning buffer of %zu bytes cannot hold %d pairs", binning_bytes, R);
        return HUGS_B200_EINVAL;
    }
    const BinningView bin = carve_binning(binning_buffer, (int)cap, cam.grid_x, cam.grid_y);
    const ImageView img = carve_image(image_buffer, width, height);

    const float* color_ptr = colors_precomp != nullptr ? colors_precomp : reinterpret_cast<const float*>(geom.rgb);
    const int stride = colors_precomp != nullptr ? 3 : 4;
    // The blend backward accumulates with reductions into the packed per-Gaussian records (geometry
    // workspace) and into dL_dopacity, so both are cleared first (part of the stage's cost); the
    // per-Gaussian kernel below unpacks the records into dL_dmean2D / dL_dconic / dL_dcolor.
    // Gradient sinks (two-segment calls only): a segment's rows of a gradient are ADDED to the caller's
    // accumulation buffer by the per-Gaussian kernel.  dL/dopacity is produced by the blend backward, so with
    // an opacity sink it goes through a scratch array of the geometry workspace first.
    const hugs_b200_grad_sinks* sinks = tl_seg2 ? tl_sinks : nullptr;
    const bool opac_scratch = sinks && (sinks->opacity[0] || sinks->opacity[1]);
    if (defer1 && !(tl_seg2 && sinks && sinks->means3D[1] && sinks->sh[1] && sinks->opacity[1] && sinks->scales[1] &&
                    sinks->rotations[1])) {
        set_error("rasterize_backward: HUGS_B200_FLAG_DEFER_SEG1 needs a two-segment call with all five sinks of "
                  "segment 1");
        return HUGS_B200_EINVAL;
    }
    float* const opac_target = opac_scratch ? geom.grad_opac : dL_dopacity;
    auto blend_backward = [&]() -> int {
        if (opac_scratch) {   // grad_pack and grad_opac are neighbours: one fill
            const size_t span = (size_t)(reinterpret_cast<char*>(geom.grad_opac + P) - reinterpret_cast<char*>(geom.grad_pack));
            HB_CUDA_OK(cudaMemsetAsync(geom.grad_pack, 0, span, stream));
        } else {
            HB_CUDA_OK(cudaMemsetAsync(geom.grad_pack, 0, sizeof(float) * 8 * (size_t)P, stream));
            HB_CUDA_OK(cudaMemsetAsync(dL_dopacity, 0, sizeof(float) * (size_t)P, stream));
        }
        if (R <= 0) return 0;
        return launch_render_backward(img.ranges, bin.list, width, height, cam.grid_x, cam.grid_y, background,
                                      geom.means2D, geom.conic_opacity, color_ptr, stride, img.accum_alpha,
                                      img.n_contrib, bin.hit_words, dL_dpix, geom.grad_pack, opac_target, stream,
                                      clamp01 ? img.clamp_mask : nullptr);
    };
    HB_STAGE(blend_backward(), ST_RENDER_BWD);
    // The layer's blend backward: same kernel on the filtered lists, gradients into the layer's own packed
    // records (summed with the main ones by the per-Gaussian kernel) and into the shared dL_dopacity.
    const float* pack2 = nullptr;
    int n2 = 0;
    if (const LayerBackward* L = tl_layer) {
        if (L->n < 0 || L->n > P || !L->buffer || !L->dL_dpix || !L->background) {
            set_error("rasterize_backward_layer: bad layer arguments");
            return HUGS_B200_EINVAL;
        }
        if (L->bytes < carve_layer(nullptr, (int)cap, L->n, width, height).total) {
            set_error("rasterize_backward_layer: layer buffer of %zu bytes is too small", L->bytes);
            return HUGS_B200_EINVAL;
        }
        const LayerView lay = carve_layer(L->buffer, (int)cap, L->n, width, height);
        auto layer_backward = [&]() -> int {
            HB_CUDA_OK(cudaMemsetAsync(lay.grad_pack, 0, sizeof(float) * 8 * (size_t)L->n, stream));
            if (R <= 0 || L->n == 0) return 0;
            return launch_render_backward(lay.ranges, lay.list, width, height, cam.grid_x, cam.grid_y, L->background,
                                          geom.means2D, geom.conic_opacity, color_ptr, stride, lay.accum_alpha,
                                          lay.n_contrib, lay.hit_words, L->dL_dpix, lay.grad_pack, opac_target,
                                          stream, clamp01 ? lay.clamp_mask : nullptr);
        };
        HB_STAGE(layer_backward(), ST_RENDER_BWD);
        pack2 = lay.grad_pack;
        n2 = L->n;
    }
    if (const SecondSegment* s2 = tl_seg2) {
        const size_t r0 = (size_t)s2->P0;
        SegBatch batch;   // both segments in one launch
        auto both = [&]() -> int {
        for (int seg = 0; seg < 2; seg++) {
            const int n = seg == 0 ? s2->P0 : P - s2->P0;
            if (n <= 0) continue;
            if (defer1 && seg == 1) {   // only what the caller is owed at once: the view's screen-space gradient
                const int rc = launch_unpack_mean2d(n, radii + r0, geom.grad_pack + 8 * r0, dL_dmean2D + 3 * r0, stream);
                if (rc) return rc;
                continue;
            }
            const size_t o = seg == 0 ? 0 : r0;   // first row of the segment in the joint arrays
            PreBwdExtra ex;
            ex.grad_pack2 = pack2 ? pack2 + 8 * o : nullptr;
            ex.n2 = seg == 0 ? (n2 < s2->P0 ? n2 : s2->P0) : (n2 > s2->P0 ? n2 - s2->P0 : 0);
            float* o_mean = dL_dmean3D + 3 * o;
            float* o_scale = dL_dscale + 3 * o;
            float* o_rot = dL_drot + 4 * o;
            float* o_sh = dL_dsh + (size_t)M * 3 * o;
            bool sh_acc = dL_dsh_accumulate != 0;
            if (opac_scratch) {
                ex.opac_scratch = geom.grad_opac + o;
                ex.dL_dopacity = dL_dopacity + o;
            }
            if (sinks) {
                if (sinks->means3D[seg]) { o_mean = sinks->means3D[seg]; ex.acc_mask |= kAccMean3D; }
                if (sinks->scales[seg]) { o_scale = sinks->scales[seg]; ex.acc_mask |= kAccScale; }
                if (sinks->rotations[seg]) { o_rot = sinks->rotations[seg]; ex.acc_mask |= kAccRot; }
                if (sinks->opacity[seg]) { ex.dL_dopacity = sinks->opacity[seg]; ex.acc_mask |= kAccOpacity; }
                if (sinks->sh[seg]) { o_sh = sinks->sh[seg]; sh_acc = true; }
            }
            const int rc = launch_preprocess_backward(
                n, D, M, seg == 0 ? means3D : s2->means3D, radii + o, seg == 0 ? shs : s2->shs, geom.clamped + o,
                seg == 0 ? scales : s2->scales, seg == 0 ? rotations : s2->rotations, scale_modifier, nullptr, cam,
                geom.grad_pack + 8 * o, dL_dmean2D + 3 * o, dL_dconic + 4 * o, o_mean, dL_dcolor + 3 * o,
                dL_dcov3D + 6 * o, o_sh, o_scale, o_rot, sh_acc, stream, &ex, &batch);
            if (rc) return rc;
        }
        return flush_preprocess_backward(batch, stream);
        };
        HB_STAGE(both(), ST_PREPROCESS_BWD);
        return 0;
    }
    PreBwdExtra ex1;
    ex1.grad_pack2 = pack2;
    ex1.n2 = n2;
    HB_STAGE(launch_preprocess_backward(P, D, M, means3D, radii, colors_precomp ? nullptr : shs, geom.clamped, scales,
                                        rotations, scale_modifier, cov3D_precomp, cam, geom.grad_pack, dL_dmean2D,
                                        dL_dconic, dL_dmean3D, dL_dcolor, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
                                        dL_dsh_accumulate != 0, stream, &ex1),
             ST_PREPROCESS_BWD);
    return 0;
}

size_t hugs_b200_layer_bytes(int n_layer, size_t binning_bytes, int width, int height) {
    if (n_layer < 0 || width <= 0 || height <= 0) return 0;
    const size_t cap = hugs_b200_binning_capacity_of(binning_bytes, width, height);
    return carve_layer(nullptr, (int)cap, n_layer, width, height).total;
}

int hugs_b200_layer_layout_of(int n_layer, size_t binning_bytes, int width, int height, size_t* offsets6) {
    if (!offsets6 || n_layer < 0 || width <= 0 || height <= 0) return HUGS_B200_EINVAL;
    const size_t cap = hugs_b200_binning_capacity_of(binning_bytes, width, height);
    const LayerView v = carve_layer(nullptr, (int)cap, n_layer, width, height);
    offsets6[0] = (size_t)reinterpret_cast<char*>(v.list);
    offsets6[1] = (size_t)reinterpret_cast<char*>(v.hit_words);
    offsets6[2] = (size_t)reinterpret_cast<char*>(v.accum_alpha);
    offsets6[3] = (size_t)reinterpret_cast<char*>(v.n_contrib);
    offsets6[4] = (size_t)reinterpret_cast<char*>(v.ranges);
    offsets6[5] = (size_t)reinterpret_cast<char*>(v.grad_pack);
    return 6;
}

int hugs_b200_layer_forward(int P, int n_layer, const float* background, int width, int height,
                            const float* colors_precomp, char* geom_buffer, char* binning_buffer,
                            size_t binning_bytes, char* image_buffer, char* layer_buffer, size_t layer_bytes,
                            float* out_color, int debug, hugs_b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool clamp01 = (debug & HUGS_B200_FLAG_CLAMP01) != 0;
    debug &= HUGS_B200_FLAG_DEBUG;
    if (P <= 0 || n_layer < 0 || n_layer > P || width <= 0 || height <= 0 || !background || !out_color || !geom_buffer ||
        !binning_buffer || !image_buffer || !layer_buffer) {
        set_error("layer_forward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    const int grid_x = cdiv(width, HB_TILE_X), grid_y = cdiv(height, HB_TILE_Y);
    const size_t cap = hugs_b200_binning_capacity_of(binning_bytes, width, height);
    if (layer_bytes < carve_layer(nullptr, (int)cap, n_layer, width, height).total) {
        set_error("layer_forward: layer buffer of %zu bytes is too small (hugs_b200_layer_bytes)", layer_bytes);
        return HUGS_B200_EINVAL;
    }
    const GeomView geom = carve_geom(geom_buffer, P);
    const BinningView bin = carve_binning(binning_buffer, (int)cap, grid_x, grid_y);
    const ImageView img = carve_image(image_buffer, width, height);
    const LayerView lay = carve_layer(layer_buffer, (int)cap, n_layer, width, height);
    const float* feature_ptr = colors_precomp != nullptr ? colors_precomp : reinterpret_cast<const float*>(geom.rgb);
    const int stride = colors_precomp != nullptr ? 3 : 4;
    HB_STAGE(launch_layer_filter(grid_x * grid_y, img.ranges, bin.list, n_layer, lay.list, lay.ranges, stream),
             ST_RANGES);
    HB_STAGE(launch_render_forward(lay.ranges, lay.list, width, height, grid_x, grid_y, geom.means2D, feature_ptr, stride,
                                   geom.conic_opacity, lay.accum_alpha, lay.n_contrib, background, out_color,
                                   lay.hit_words, stream, clamp01 ? lay.clamp_mask : nullptr),
             ST_RENDER_FWD);
    return 0;
}

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
                                       const float* dL_dpix_layer) {
    const LayerBackward L{n_layer, background_layer, layer_buffer, layer_bytes, dL_dpix_layer};
    tl_layer = dL_dpix_layer ? &L : nullptr;   // no gradient for the layer image: plain backward
    const int rc = hugs_b200_rasterize_backward_acc(
        P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
        cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer,
        binning_bytes, image_buffer, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D,
        dL_dsh, dL_dscale, dL_drot, debug, stream, dL_dsh_accumulate);
    tl_layer = nullptr;
    return rc;
}

int hugs_b200_rasterize_backward(int P, int D, int M, int R, const float* background, int width, int height,
                                 const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* scales, float scale_modifier, const float* rotations,
                                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                 const float* cam_pos, float tan_fovx, float tan_fovy, const int* radii,
                                 char* geom_buffer, char* binning_buffer, size_t binning_bytes, char* image_buffer,
                                 const float* dL_dpix,
                                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                                 float* dL_drot, int debug, hugs_b200_stream_t stream) {
    return hugs_b200_rasterize_backward_acc(P, D, M, R, background, width, height, means3D, shs, colors_precomp, scales,
                                            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos,
                                            tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, binning_bytes,
                                            image_buffer, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                                            dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug, stream, 0);
}

int hugs_b200_lbs_forward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                          const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                          float* out_xyz, float* out_rotq, float* out_scales, float* out_T,
                          hugs_b200_stream_t stream) {
    if (N > 0 && (!A || !xyz || !weights || !rotmat || !out_xyz || !out_rotq)) {
        set_error("lbs_forward: null pointer argument");
        return HUGS_B200_EINVAL;
    }
    StageScope scope(ST_LBS_FWD, static_cast<cudaStream_t>(stream));
    return launch_lbs_forward(F, N, J, A, xyz, weights, rotmat, scales, transl, smpl_scale, out_xyz, out_rotq,
                              out_scales, out_T, static_cast<cudaStream_t>(stream));
}

int hugs_b200_lbs_backward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                           const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                           const float* g_xyz, const float* g_rotq, const float* g_scales, float* dA, float* d_xyz,
                           float* d_weights, float* d_rotmat, float* d_scales, float* d_transl, float* d_smpl_scale,
                           hugs_b200_stream_t stream) {
    if (!dA || (N > 0 && (!A || !xyz || !weights || !rotmat || !g_xyz || !g_rotq || !d_xyz || !d_weights || !d_rotmat))) {
        set_error("lbs_backward: null pointer argument");
        return HUGS_B200_EINVAL;
    }
    StageScope scope(ST_LBS_BWD, static_cast<cudaStream_t>(stream));
    return launch_lbs_backward(F, N, J, A, xyz, weights, rotmat, scales, transl, smpl_scale, g_xyz, g_rotq, g_scales,
                               dA, d_xyz, d_weights, d_rotmat, d_scales, d_transl, d_smpl_scale,
                               static_cast<cudaStream_t>(stream));
}

int hugs_b200_l1_ssim_forward(int C, int H, int W, const float* img1, const float* img2, float* dmap, double* sums,
                              hugs_b200_stream_t stream) {
    if (C < 0 || H < 0 || W < 0 || !sums || ((size_t)C * H * W > 0 && (!img1 || !img2))) {
        set_error("l1_ssim_forward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_l1_ssim_forward(C, H, W, img1, img2, dmap, sums, static_cast<cudaStream_t>(stream));
}

int hugs_b200_l1_ssim_backward(int C, int H, int W, const float* img1, const float* img2, const float* dmap,
                               const float* g, float* dL_dimg1, hugs_b200_stream_t stream) {
    if (C < 0 || H < 0 || W < 0 || ((size_t)C * H * W > 0 && (!img1 || !img2 || !dmap || !g || !dL_dimg1))) {
        set_error("l1_ssim_backward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_l1_ssim_backward(C, H, W, img1, img2, dmap, g, dL_dimg1, static_cast<cudaStream_t>(stream));
}

size_t hugs_b200_knn_temp_bytes(int P) { return knn_temp_bytes(P); }

int hugs_b200_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* temp, size_t temp_bytes,
                             hugs_b200_stream_t stream) {
    if (P < 0 || (P > 0 && (!points || !mean_dist2))) {
        set_error("knn_mean_dist2: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_knn_mean_dist2(P, points, mean_dist2, temp, temp_bytes, static_cast<cudaStream_t>(stream));
}

static int check_seg2(const char* what, int P0, int P1, const float* const* a, const float* const* b) {
    if (P0 < 0 || P1 < 0 || (long long)P0 + P1 > 0x7fffffffll) {
        set_error("%s: bad segment sizes %d + %d", what, P0, P1);
        return HUGS_B200_EINVAL;
    }
    for (int i = 0; i < 5; i++)
        if ((P0 > 0 && !a[i]) || (P1 > 0 && !b[i])) {
            set_error("%s: a segment lacks means3D / shs / opacities / scales / rotations", what);
            return HUGS_B200_EINVAL;
        }
    return 0;
}

int hugs_b200_rasterize_forward_seg2(hugs_b200_resize_fn geometry_buffer, void* geometry_user,
                                     hugs_b200_resize_fn binning_buffer, void* binning_user,
                                     hugs_b200_resize_fn image_buffer, void* image_user, int D, int M,
                                     const float* background, int width, int height, int P0, const float* means3D0,
                                     const float* shs0, const float* opacities0, const float* scales0,
                                     const float* rotations0, int P1, const float* means3D1, const float* shs1,
                                     const float* opacities1, const float* scales1, const float* rotations1,
                                     float scale_modifier, const float* viewmatrix, const float* projmatrix,
                                     const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered,
                                     float* out_color, int* radii, int debug, hugs_b200_stream_t stream,
                                     int* num_rendered) {
    const float* a[5] = {means3D0, shs0, opacities0, scales0, rotations0};
    const float* b[5] = {means3D1, shs1, opacities1, scales1, rotations1};
    if (const int rc = check_seg2("rasterize_forward_seg2", P0, P1, a, b)) return rc;
    const bool first_empty = P0 == 0;   // the argument checks below look at the first segment's pointers
    const SecondSegment s2{P0, means3D1, shs1, opacities1, scales1, rotations1};
    tl_seg2 = &s2;
    const int rc = hugs_b200_rasterize_forward(
        geometry_buffer, geometry_user, binning_buffer, binning_user, image_buffer, image_user, P0 + P1, D, M,
        background, width, height, first_empty ? means3D1 : means3D0, first_empty ? shs1 : shs0, nullptr,
        first_empty ? opacities1 : opacities0, first_empty ? scales1 : scales0, scale_modifier,
        first_empty ? rotations1 : rotations0, nullptr, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
        prefiltered, out_color, radii, debug, stream, num_rendered);
    tl_seg2 = nullptr;
    return rc;
}

int hugs_b200_rasterize_backward_seg2(int D, int M, int R, const float* background, int width, int height, int P0,
                                      const float* means3D0, const float* shs0, const float* scales0,
                                      const float* rotations0, int P1, const float* means3D1, const float* shs1,
                                      const float* scales1, const float* rotations1, float scale_modifier,
                                      const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                                      float tan_fovx, float tan_fovy, const int* radii, char* geom_buffer,
                                      char* binning_buffer, size_t binning_bytes, char* image_buffer,
                                      const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                                      float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                                      float* dL_dscale, float* dL_drot, int debug, hugs_b200_stream_t stream) {
    const float* a[5] = {means3D0, shs0, means3D0, scales0, rotations0};
    const float* b[5] = {means3D1, shs1, means3D1, scales1, rotations1};
    if (const int rc = check_seg2("rasterize_backward_seg2", P0, P1, a, b)) return rc;
    const bool first_empty = P0 == 0;
    const SecondSegment s2{P0, means3D1, shs1, nullptr, scales1, rotations1};
    tl_seg2 = &s2;
    const int rc = hugs_b200_rasterize_backward_acc(
        P0 + P1, D, M, R, background, width, height, first_empty ? means3D1 : means3D0, first_empty ? shs1 : shs0,
        nullptr, first_empty ? scales1 : scales0, scale_modifier, first_empty ? rotations1 : rotations0, nullptr,
        viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, binning_bytes,
        image_buffer, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
        dL_drot, debug, stream, 0);
    tl_seg2 = nullptr;
    return rc;
}

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
                                            const float* dL_dpix_layer) {
    const LayerBackward L{n_layer, background_layer, layer_buffer, layer_bytes, dL_dpix_layer};
    tl_layer = dL_dpix_layer ? &L : nullptr;
    const int rc = hugs_b200_rasterize_backward_seg2(
        D, M, R, background, width, height, P0, means3D0, shs0, scales0, rotations0, P1, means3D1, shs1, scales1,
        rotations1, scale_modifier, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, radii, geom_buffer,
        binning_buffer, binning_bytes, image_buffer, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D,
        dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug, stream);
    tl_layer = nullptr;
    return rc;
}

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
                                         const hugs_b200_grad_sinks* sinks) {
    tl_sinks = sinks;
    const int rc = hugs_b200_rasterize_backward_seg2_layer(
        D, M, R, background, width, height, P0, means3D0, shs0, scales0, rotations0, P1, means3D1, shs1, scales1,
        rotations1, scale_modifier, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, radii, geom_buffer,
        binning_buffer, binning_bytes, image_buffer, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D,
        dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug, stream, n_layer, background_layer, layer_buffer, layer_bytes,
        dL_dpix_layer);
    tl_sinks = nullptr;
    return rc;
}

int hugs_b200_preprocess_scene_views(int n_views, const hugs_b200_scene_view* views, int P0, int P1, int D, int M,
                                     int width, int height, const float* means3D1, const float* shs1,
                                     const float* opacities1, const float* scales1, float scale_modifier,
                                     const float* rotations1, int prefiltered, hugs_b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n_views < 0 || P0 < 0 || P1 < 0 || width <= 0 || height <= 0 || (n_views > 0 && !views)) {
        set_error("preprocess_scene_views: bad arguments");
        return HUGS_B200_EINVAL;
    }
    if (n_views == 0) return 0;
    if (P1 > 0 && (!means3D1 || !shs1 || !opacities1 || !scales1 || !rotations1)) {
        set_error("preprocess_scene_views: null pointer (segment 1 needs its five input arrays)");
        return HUGS_B200_EINVAL;
    }
    const int P = P0 + P1;
    std::vector<PreViewHost> host((size_t)n_views);
    for (int v = 0; v < n_views; v++) {
        const hugs_b200_scene_view& in = views[v];
        if (!in.viewmatrix || !in.projmatrix || !in.cam_pos || !in.radii || !in.geom_buffer) {
            set_error("preprocess_scene_views: view %d: null pointer", v);
            return HUGS_B200_EINVAL;
        }
        const GeomView geom = carve_geom(in.geom_buffer, P);
        // the counts start here; the view's forward call (HUGS_B200_FLAG_SEG1_PREPROCESSED) adds segment 0's share
        HB_CUDA_OK(cudaMemsetAsync(geom.pair_counts, 0, 2 * sizeof(uint32_t), stream));
        PreViewHost& h = host[(size_t)v];
        h.cam = make_camera(in.viewmatrix, in.projmatrix, in.cam_pos, width, height, in.tan_fovx, in.tan_fovy);
        h.radii = const_cast<int*>(in.radii) + P0;
        h.g = geom_rows_from(geom, P0);
    }
    if (P1 == 0) return 0;
    StageScope scope(ST_PREPROCESS, stream);
    return launch_preprocess_views(P1, D, M, n_views, host.data(), means3D1, scales1, scale_modifier, rotations1,
                                   opacities1, shs1, prefiltered != 0, stream);
}

int hugs_b200_scene_backward_views(int n_views, const hugs_b200_scene_view* views, int P0, int P1, int D, int M,
                                   int width, int height, const float* means3D1, const float* shs1,
                                   const float* scales1, float scale_modifier, const float* rotations1,
                                   const hugs_b200_grad_sinks* sinks, hugs_b200_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n_views < 0 || P0 < 0 || P1 < 0 || width <= 0 || height <= 0 || (n_views > 0 && !views)) {
        set_error("scene_backward_views: bad arguments");
        return HUGS_B200_EINVAL;
    }
    if (n_views == 0 || P1 == 0) return 0;
    if (!means3D1 || !shs1 || !scales1 || !rotations1 || !sinks || !sinks->means3D[1] || !sinks->sh[1] ||
        !sinks->opacity[1] || !sinks->scales[1] || !sinks->rotations[1]) {
        set_error("scene_backward_views: null pointer (segment 1 needs its four input arrays and all five sinks)");
        return HUGS_B200_EINVAL;
    }
    const int P = P0 + P1;
    std::vector<SceneViewHost> host((size_t)n_views);
    for (int v = 0; v < n_views; v++) {
        const hugs_b200_scene_view& in = views[v];
        if (!in.viewmatrix || !in.projmatrix || !in.cam_pos || !in.radii || !in.geom_buffer) {
            set_error("scene_backward_views: view %d: null pointer", v);
            return HUGS_B200_EINVAL;
        }
        const GeomView geom = carve_geom(in.geom_buffer, P);
        SceneViewHost& h = host[(size_t)v];
        h.cam = make_camera(in.viewmatrix, in.projmatrix, in.cam_pos, width, height, in.tan_fovx, in.tan_fovy);
        h.radii = in.radii + P0;
        h.grad_pack = geom.grad_pack + 8 * (size_t)P0;
        h.clamped = geom.clamped + P0;
        h.opac_scratch = geom.grad_opac + P0;
    }
    StageScope scope(ST_PREPROCESS_BWD, stream);
    return launch_scene_backward_views(P1, D, M, n_views, host.data(), means3D1, shs1, scales1, rotations1,
                                       scale_modifier, sinks->means3D[1], sinks->sh[1], sinks->scales[1],
                                       sinks->rotations[1], sinks->opacity[1], stream);
}

size_t hugs_b200_mlp_packed_floats(int K, int n_pad) {
    return (K > 0 && n_pad > 0) ? mlp_packed_floats(K, n_pad) : 0;
}

int hugs_b200_mlp_pack(const float* W, const float* bias, int n_true, int n_pad, int K, float* packed,
                       hugs_b200_stream_t stream) {
    if (!W || !packed || n_true <= 0) {
        set_error("mlp_pack: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_mlp_pack(W, bias, n_true, n_pad, K, packed, static_cast<cudaStream_t>(stream));
}

int hugs_b200_mlp_chains_forward(int rows, const float* x, int ldx, int Kx, int n_chains,
                                 const hugs_b200_mlp_chain* chains, hugs_b200_stream_t stream) {
    if (rows < 0 || n_chains < 1 || n_chains > HB_MLP_MAX_CHAINS || !chains || (rows > 0 && !x)) {
        set_error("mlp_chains_forward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    MlpArgs a;
    memset(&a, 0, sizeof(a));
    a.rows = rows;
    a.x = x;
    a.ldx = ldx;
    a.Kx = Kx;
    a.n_chains = n_chains;
    for (int c = 0; c < n_chains; c++) {
        const hugs_b200_mlp_chain& src = chains[c];
        if (src.n_layers < 1 || src.n_layers > HB_MLP_MAX_LAYERS || src.n_outs < 0 || src.n_outs > HB_MLP_MAX_OUTS) {
            set_error("mlp_chains_forward: chain %d: bad layer / output count", c);
            return HUGS_B200_EINVAL;
        }
        MlpChainArgs& dst = a.chain[c];
        dst.n_layers = src.n_layers;
        dst.n_outs = src.n_outs;
        for (int l = 0; l < src.n_layers; l++)
            dst.layer[l] = MlpLayerArgs{src.layer[l].K, src.layer[l].N, src.layer[l].act, src.layer[l].packed,
                                        src.layer[l].z_out};
        for (int o = 0; o < src.n_outs; o++) {
            if (!src.out[o].ptr || src.out[o].ncols <= 0 || src.out[o].col0 < 0) {
                set_error("mlp_chains_forward: chain %d output %d is malformed", c, o);
                return HUGS_B200_EINVAL;
            }
            dst.out[o] = MlpOut{src.out[o].ptr, src.out[o].ld, src.out[o].col0, src.out[o].ncols, src.out[o].act};
        }
    }
    return launch_mlp_chains(a, static_cast<cudaStream_t>(stream));
}

int hugs_b200_mlp_pack_t(const float* W, int n_true, int k_true, int n_pad, float* packed, hugs_b200_stream_t stream) {
    if (!W || !packed || n_true <= 0) {
        set_error("mlp_pack_t: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_mlp_pack_t(W, n_true, k_true, n_pad, packed, static_cast<cudaStream_t>(stream));
}

int hugs_b200_mlp_chains_backward(int rows, const float* x, int ldx, int Kx, int n_chains,
                                  const hugs_b200_mlp_bwd_chain* chains, hugs_b200_stream_t stream) {
    if (rows < 0 || n_chains < 1 || n_chains > HB_MLP_MAX_CHAINS || !chains || (rows > 0 && !x)) {
        set_error("mlp_chains_backward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    MlpBwdArgs a;
    memset(&a, 0, sizeof(a));
    a.rows = rows;
    a.x = x;
    a.ldx = ldx;
    a.Kx = Kx;
    a.n_chains = n_chains;
    for (int c = 0; c < n_chains; c++) {
        const hugs_b200_mlp_bwd_chain& src = chains[c];
        if (src.n_layers < 1 || src.n_layers > HB_MLP_MAX_LAYERS || src.n_outs < 0 || src.n_outs > HB_MLP_MAX_OUTS) {
            set_error("mlp_chains_backward: chain %d: bad layer / output count", c);
            return HUGS_B200_EINVAL;
        }
        MlpBwdChainArgs& dst = a.chain[c];
        dst.n_layers = src.n_layers;
        dst.n_outs = src.n_outs;
        dst.z_head = src.z_head;
        dst.dx = src.dx;
        for (int l = 0; l < src.n_layers; l++)
            dst.layer[l] = MlpBwdLayerArgs{src.layer[l].K, src.layer[l].N, src.layer[l].packed_t, src.layer[l].z_prev,
                                           src.layer[l].dWt, src.layer[l].db};
        for (int o = 0; o < src.n_outs; o++)
            dst.out[o] = MlpBwdOut{src.out[o].dptr, src.out[o].ld, src.out[o].col0, src.out[o].ncols, src.out[o].act};
    }
    return launch_mlp_chains_backward(a, static_cast<cudaStream_t>(stream));
}

int hugs_b200_densification_stats(int n, const int* radii, const float* grad_means2D, int grad_stride,
                                  float* xyz_gradient_accum, float* denom, float* max_radii2D,
                                  hugs_b200_stream_t stream) {
    if (n < 0 || (n > 0 && (!radii || (xyz_gradient_accum && (!grad_means2D || grad_stride < 2))))) {
        set_error("densification_stats: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_densification_stats(n, radii, grad_means2D, grad_stride, xyz_gradient_accum, denom, max_radii2D,
                                      static_cast<cudaStream_t>(stream));
}

int hugs_b200_triplane_forward(int N, int C, int resX, int resY, int resZ, const float* plane_xy,
                               const float* plane_xz, const float* plane_yz, const float* xyz, float center,
                               float scale, float eps, float* features, int* out_of_range,
                               hugs_b200_stream_t stream) {
    if (N < 0 || C <= 0 || resX <= 0 || resY <= 0 || resZ <= 0 || scale == 0.f ||
        (N > 0 && (!plane_xy || !plane_xz || !plane_yz || !xyz || !features))) {
        set_error("triplane_forward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_triplane_forward(N, C, resX, resY, resZ, plane_xy, plane_xz, plane_yz, xyz, center, scale, eps,
                                   features, out_of_range, static_cast<cudaStream_t>(stream));
}

int hugs_b200_triplane_backward(int N, int C, int resX, int resY, int resZ, const float* plane_xy,
                                const float* plane_xz, const float* plane_yz, const float* xyz, float center,
                                float scale, const float* dL_dfeatures, float* dL_dplane_xy, float* dL_dplane_xz,
                                float* dL_dplane_yz, float* dL_dxyz, hugs_b200_stream_t stream) {
    if (N < 0 || C <= 0 || resX <= 0 || resY <= 0 || resZ <= 0 || scale == 0.f ||
        (N > 0 && (!plane_xy || !plane_xz || !plane_yz || !xyz || !dL_dfeatures))) {
        set_error("triplane_backward: bad arguments");
        return HUGS_B200_EINVAL;
    }
    return launch_triplane_backward(N, C, resX, resY, resZ, plane_xy, plane_xz, plane_yz, xyz, center, scale,
                                    dL_dfeatures, dL_dplane_xy, dL_dplane_xz, dL_dplane_yz, dL_dxyz,
                                    static_cast<cudaStream_t>(stream));
}

}  // extern "C"
