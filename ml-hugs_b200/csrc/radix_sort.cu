// Hand-written onesweep-style LSD radix sort for sm_100a (u32 or u64 keys, u32 payload).
// Replaces cub::DeviceRadixSort::SortPairs at cuda_rasterizer/rasterizer_impl.cu:303-308
// (CUB is used only as a correctness oracle in tests/).
//
// Building blocks:
//   * radix_histogram_kernel   one read of the keys builds the digit histogram of every pass;
//   * radix_offsets_kernel     histogram -> exclusive bin offsets (one CTA per pass);
//   * radix_onesweep_kernel    one scatter pass.  Each CTA (512 threads x 8 pairs) owns a 4096-pair
//     tile (atomic ticket order), ranks its keys stably with per-bit ballots + per-warp counters,
//     obtains the tile's per-digit global offset by a WINDOWED decoupled look-back over 64-bit
//     (flag|count) words, stages the tile digit-sorted in shared memory and writes it out in
//     coalesced runs.
// Stability: a tile is loaded warp-striped (warp w owns 256 consecutive pairs, item k of lane l
// is pair 32k+l), items are ranked in k order and lanes in lane order, warps in warp order, tiles
// in ticket == index order.
//
// The pair count may live in device memory (`n_dev`): the grid is then sized for a capacity and
// CTAs beyond the real count retire immediately.  That is what lets the rasterizer enqueue the
// whole binning stage before the number of (Gaussian, tile) pairs has reached the host.
//
// How the rasterizer uses it (api.cu): the reference sorts R pairs on 32+bit key bits
// (tile | depth) in one go.  The same order -- tile ascending, then depth bits ascending, ties in
// ascending Gaussian index -- is produced here by (1) sorting the P Gaussians once by depth bits
// (u32 keys, 4 passes over P items), (2) emitting the pairs in that order and (3) a stable sort
// of the R pairs on the tile bits only (2 passes at 1080p instead of 6).  See DESIGN.md.
#include "common.cuh"

#include <cstring>

namespace hb {

namespace {
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 512;
constexpr int kSortItems = 8;
constexpr int kMinSortTile = kSortThreads * kSortItems;  // 4096 pairs per CTA: sizes the look-back array
constexpr int kHistThreads = 512;
static_assert(kSortMaxPasses == 8, "header constant");

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- 1. histograms of all passes in one read of the keys --------------------------------------
// Pass p looks at `bits[p]` key bits starting at `shift[p]`.
using PassPlan = SortPassPlan;

template <typename KeyT>
__global__ void __launch_bounds__(kHistThreads)
    radix_histogram_kernel(const KeyT* __restrict__ keys, int n, const PassPlan plan, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[kSortMaxPasses * kRadix];
    for (int i = threadIdx.x; i < plan.num_passes * kRadix; i += kHistThreads) s_hist[i] = 0;
    __syncthreads();
    const int stride = gridDim.x * kHistThreads;
    for (int i = blockIdx.x * kHistThreads + threadIdx.x; i < n; i += stride) {
        const KeyT k = __ldg(keys + i);
#pragma unroll
        for (int p = 0; p < kSortMaxPasses; p++) {
            if (p < plan.num_passes) {
                const uint32_t d = (uint32_t)(k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u);
                atomicAdd(&s_hist[p * kRadix + d], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.num_passes * kRadix; i += kHistThreads) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// ---- 2. histogram -> exclusive offsets (one CTA per pass) --------------------------------------
__global__ void __launch_bounds__(kRadix) radix_offsets_kernel(uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_warp[kRadix / 32];
    uint32_t* h = hist + blockIdx.x * kRadix;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t c = h[threadIdx.x];
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (unsigned)d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (unsigned w = 0; w < warp; w++) base += s_warp[w];
    h[threadIdx.x] = base + incl - c;
}

// ---- 3. one scatter pass ---------------------------------------------------------------------
// The kernel is bound by instruction issue and shared-memory latency, not by HBM (ncu: 93 M warp
// instructions per pass over 12.25 M pairs at 1.7 IPC/SM), so the three payload modes below exist to
// execute fewer instructions per pair, and the CTA shape (512 threads x 8 pairs) to keep 32 warps
// per SM resident with every global load of the tile in flight before its first use.
//   kPairs     (key, u32 payload) pairs; vals_in == nullptr means "the payload of item i is i"
//              (first pass of an argsort).  Generic stable pair sort.
//   kKeysOnly  the payload rides in the low 32 bits of a u64 key (the sorted digit lies above them):
//              one load / stage / store per pair instead of two.
//   kSplit     last pass of a kKeysOnly sort: emits the low word (payload) to vals_out and the high
//              word to a u16 / u32 array at keys_out.  The rasterizer sorts (tile << 32 | Gaussian
//              index) and gets the sorted Gaussian list plus the tile id of every list position
//              (6 B per pair written instead of 12).
enum SortMode { kPairs = 0, kKeysOnly = 1, kSplit = 2 };

template <typename KeyT, int THREADS, int ITEMS, int MODE>
struct SortSmem {
    KeyT keys[THREADS * ITEMS];                              // 32 KB (u64) / 16 KB (u32)
    uint32_t vals[MODE == kPairs ? THREADS * ITEMS : 1];     // 16 KB
    uint32_t warp_hist[THREADS / 32][kRadix];                // 16 KB
    uint32_t local_start[kRadix];  // first staged slot of each digit
    uint32_t count[kRadix];        // pairs of each digit in this tile
    uint32_t bin_off[kRadix];      // global index of staged slot 0 of each digit (mod 2^32)
    uint32_t scan_warp[kRadix / 32];
    unsigned tile;
};

// Decoupled look-back with a WINDOW.  The textbook walk (one thread per digit, one predecessor per
// dependent L2 round trip) makes the time for an inclusive prefix to reach a tile grow like
// sqrt(2 x tiles in flight) round trips; with ~300-450 tiles in flight on 148 SMs that chain cost
// more than the HBM traffic of the pass.  Here the `T` threads of a digit fetch WIN x T predecessors
// per round trip, all loads in flight at once, and combine with shuffles inside the digit's lane
// group: a narrow first round (the nearest predecessors usually suffice), wide ones after it.
template <int WIN>
__device__ __forceinline__ bool lookback_round(const unsigned long long* __restrict__ lookback, int t0, int d,
                                               int sub, int T, unsigned gmask, unsigned flag_agg, unsigned flag_inc,
                                               uint32_t& prefix) {
    const int kBig = 1 << 30;
    unsigned long long w[WIN];
#pragma unroll
    for (int k = 0; k < WIN; k++) {
        const int t = t0 - sub - T * k;
        w[k] = (t >= 0) ? ld_relaxed_u64(lookback + (size_t)t * kRadix + d) : ((unsigned long long)flag_inc << 32);
    }
#pragma unroll
    for (int k = 0; k < WIN; k++) {
        const int t = t0 - sub - T * k;
        if (t >= 0)
            while ((unsigned)(w[k] >> 32) < flag_agg) w[k] = ld_relaxed_u64(lookback + (size_t)t * kRadix + d);
    }
    // distance (0-based) of the nearest predecessor that already holds an inclusive prefix
    int best = kBig;
#pragma unroll
    for (int k = WIN - 1; k >= 0; k--)
        if ((unsigned)(w[k] >> 32) == flag_inc) best = sub + T * k;
    for (int o = T >> 1; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(gmask, best, o));
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < WIN; k++)
        if (sub + T * k <= best) sum += (uint32_t)w[k];
    for (int o = T >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(gmask, sum, o);
    prefix += sum;
    return best != kBig;
}

__device__ __forceinline__ uint32_t lookback_prefix(const unsigned long long* __restrict__ lookback, int tile, int d,
                                                    int sub, int T, unsigned gmask, unsigned flag_agg,
                                                    unsigned flag_inc) {
    uint32_t prefix = 0;
    int t0 = tile - 1;
    if (t0 < 0) return 0;
    if (lookback_round<2>(lookback, t0, d, sub, T, gmask, flag_agg, flag_inc, prefix)) return prefix;
    for (t0 -= 2 * T; t0 >= 0; t0 -= 8 * T)
        if (lookback_round<8>(lookback, t0, d, sub, T, gmask, flag_agg, flag_inc, prefix)) break;
    return prefix;
}

// Stable warp-local rank of one item.  Lanes holding the same digit are found with one ballot per
// digit bit (constant cost); match.any takes time proportional to the number of distinct digits in
// the warp (profiles/r01_b_sort_passes.txt).  The leader bumps the warp's counter with one
// shared-memory atomic (returns the old value).
template <bool FULL, int BITS>
__device__ __forceinline__ uint32_t warp_rank(uint32_t d, bool valid, int digit_bits, uint32_t* warp_hist,
                                              unsigned lane, unsigned lanemask_lt) {
    const int nbits = BITS ? BITS : digit_bits;  // BITS != 0: known at compile time, the loop below is straight-line
    unsigned peers = FULL ? 0xffffffffu : __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < kRadixBits; b++) {
        if (b < nbits) {
            const bool bit = (d >> b) & 1u;
            const unsigned m = __ballot_sync(0xffffffffu, FULL ? bit : (valid && bit));
            peers &= bit ? m : ~m;
        }
    }
    // every lane takes part in the shuffle (a data-dependent member mask costs a match + reduction)
    const int leader = (__ffs(peers) - 1) & 31;
    uint32_t old = 0;
    if ((FULL || valid) && (int)lane == leader) old = atomicAdd(warp_hist + d, (uint32_t)__popc(peers));
    old = __shfl_sync(0xffffffffu, old, leader);
    __syncwarp();
    return (FULL || valid) ? old + __popc(peers & lanemask_lt) : 0u;
}

template <typename KeyT, int THREADS, int ITEMS, int MINB, int MODE, int BITS>
__global__ void __launch_bounds__(THREADS, MINB)
    radix_onesweep_kernel(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                          KeyT* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n_host,
                          const uint32_t* __restrict__ n_dev, int shift, uint32_t digit_mask, int pass,
                          const uint32_t* __restrict__ pass_offsets, unsigned long long* lookback,
                          unsigned* tickets, int tile16) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int kSortThreads = THREADS, kSortItems = ITEMS, kSortWarps = THREADS / 32, kSortTile = THREADS * ITEMS;
    static_assert(THREADS >= kRadix && THREADS % kRadix == 0, "digit phases use the first kRadix threads");
    static_assert(MODE == kPairs || sizeof(KeyT) == 8, "payload-in-key modes need 64-bit keys");
    using Smem = SortSmem<KeyT, THREADS, ITEMS, MODE>;
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    const unsigned lanemask_lt = (1u << lane) - 1u;
    const int digit_bits = BITS ? BITS : __popc(digit_mask);
    if (BITS) digit_mask = (1u << BITS) - 1u;

    // n_host is the capacity the grid was sized for when the real count lives on the device
    int n = n_host;
    if (n_dev != nullptr) n = (int)min((uint32_t)n_host, __ldg(n_dev));

    if (tid == 0) s.tile = atomicAdd(tickets + pass, 1u);
    for (int i = tid; i < kSortWarps * kRadix; i += kSortThreads) (&s.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int tile = (int)s.tile;
    const long long tile_base_ll = (long long)tile * kSortTile;
    if (tile_base_ll >= (long long)n) return;  // beyond the real count: nothing to do, nobody waits on us
    const int tile_base = (int)tile_base_ll;
    const int warp_base = tile_base + (int)warp * (32 * kSortItems);
    const bool full_tile = tile_base + kSortTile <= n;

    // Load keys (and payloads), warp-striped; every load of the tile is in flight before the first use.
    KeyT key[kSortItems];
    uint32_t val[MODE == kPairs ? kSortItems : 1];
    if (full_tile) {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) key[k] = __ldg(keys_in + warp_base + k * 32 + lane);
        if (MODE == kPairs) {
#pragma unroll
            for (int k = 0; k < kSortItems; k++) {
                const int i = warp_base + k * 32 + (int)lane;
                val[MODE == kPairs ? k : 0] = (vals_in != nullptr) ? __ldg(vals_in + i) : (uint32_t)i;
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) {
            const int i = warp_base + k * 32 + (int)lane;
            key[k] = (i < n) ? __ldg(keys_in + i) : (KeyT)~(KeyT)0;
            if (MODE == kPairs) val[MODE == kPairs ? k : 0] = (i < n && vals_in != nullptr) ? __ldg(vals_in + i) : (uint32_t)i;
        }
    }

    uint32_t dig[kSortItems], rank[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; k++) dig[k] = (uint32_t)(key[k] >> shift) & digit_mask;
    if (full_tile) {
#pragma unroll
        for (int k = 0; k < kSortItems; k++)
            rank[k] = warp_rank<true, BITS>(dig[k], true, digit_bits, s.warp_hist[warp], lane, lanemask_lt);
    } else {
#pragma unroll
        for (int k = 0; k < kSortItems; k++)
            rank[k] = warp_rank<false, BITS>(dig[k], warp_base + k * 32 + (int)lane < n, digit_bits, s.warp_hist[warp], lane,
                                       lanemask_lt);
    }
    __syncthreads();

    const unsigned flag_agg = (unsigned)(2 * pass + 1), flag_inc = (unsigned)(2 * pass + 2);
    // Per digit (thread tid == digit): exclusive scan over warps, tile count, publish the aggregate.
    uint32_t tile_count = 0, incl = 0;
    if (tid < kRadix) {
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t t = s.warp_hist[w][tid];
            s.warp_hist[w][tid] = sum;
            sum += t;
        }
        tile_count = sum;
        st_relaxed_u64(lookback + (size_t)tile * kRadix + tid,
                       ((unsigned long long)(tile == 0 ? flag_inc : flag_agg) << 32) | tile_count);
        s.count[tid] = tile_count;
        // exclusive scan of tile_count over digits -> local_start
        incl = tile_count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (unsigned)o) incl += t;
        }
        if (lane == 31) s.scan_warp[warp] = incl;
    }
    __syncthreads();
    if (tid < kRadix) {
        uint32_t base = 0;
        for (unsigned w = 0; w < warp; w++) base += s.scan_warp[w];
        s.local_start[tid] = base + incl - tile_count;
    }
    __syncthreads();

    // Windowed decoupled look-back: T = threads per digit (a power of two, at most one warp).
    {
        const int bins = 1 << digit_bits;
        const int T = min(kSortThreads / bins, 32);
        const int d = tid / T, sub = tid & (T - 1);
        if (d < bins) {
            const unsigned gmask = (T == 32) ? 0xffffffffu : (((1u << T) - 1u) << (lane & ~(unsigned)(T - 1)));
            const uint32_t prefix = lookback_prefix(lookback, tile, d, sub, T, gmask, flag_agg, flag_inc);
            if (sub == 0) {
                const uint32_t cnt = s.count[d];
                if (tile != 0)
                    st_relaxed_u64(lookback + (size_t)tile * kRadix + d,
                                   ((unsigned long long)flag_inc << 32) | (unsigned long long)(prefix + cnt));
                s.bin_off[d] = __ldg(pass_offsets + d) + prefix - s.local_start[d];
            }
        }
    }

    // Stage the tile digit-sorted in shared memory (local_start / warp_hist were final before the
    // barrier above; bin_off is not read until after the next one).
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const int i = warp_base + k * 32 + (int)lane;
        if (full_tile || i < n) {
            const uint32_t pos = s.local_start[dig[k]] + s.warp_hist[warp][dig[k]] + rank[k];
            s.keys[pos] = key[k];
            if (MODE == kPairs) s.vals[pos] = val[MODE == kPairs ? k : 0];
        }
    }
    __syncthreads();

    // Coalesced write-out: consecutive staged slots of one digit are consecutive in global memory.
    // Shared-memory reads are batched ahead of the stores (the compiler cannot move a load above a
    // store to a possibly aliasing pointer on its own).
    const int count = full_tile ? kSortTile : (n - tile_base);
    KeyT ko[kSortItems];
    uint32_t go[kSortItems], vo[MODE == kPairs ? kSortItems : 1];
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const int i = k * kSortThreads + tid;
        if (i < count) {
            ko[k] = s.keys[i];
            if (MODE == kPairs) vo[MODE == kPairs ? k : 0] = s.vals[i];
            go[k] = s.bin_off[(uint32_t)(ko[k] >> shift) & digit_mask] + (uint32_t)i;
        }
    }
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        if (k * kSortThreads + tid < count) {
            if (MODE == kPairs) {
                keys_out[go[k]] = ko[k];
                vals_out[go[k]] = vo[MODE == kPairs ? k : 0];
            } else if (MODE == kKeysOnly) {
                keys_out[go[k]] = ko[k];
            } else {
                vals_out[go[k]] = (uint32_t)ko[k];
                const uint32_t hi = (uint32_t)((unsigned long long)ko[k] >> 32);
                if (tile16) reinterpret_cast<uint16_t*>(keys_out)[go[k]] = (uint16_t)hi;
                else reinterpret_cast<uint32_t*>(keys_out)[go[k]] = hi;
            }
        }
    }
}

struct SortTemp {
    uint32_t* hist;                // [kSortMaxPasses][kRadix]
    unsigned* tickets;             // [kSortMaxPasses]
    unsigned long long* lookback;  // [tiles][kRadix]
    size_t bytes;
};

SortTemp carve_sort_temp(void* temp, int n) {
    SortTemp t;
    char* p = reinterpret_cast<char*>(temp);
    t.hist = reinterpret_cast<uint32_t*>(p);
    p += align_up(sizeof(uint32_t) * kSortMaxPasses * kRadix);
    t.tickets = reinterpret_cast<unsigned*>(p);
    p += align_up(sizeof(unsigned) * kSortMaxPasses);
    t.lookback = reinterpret_cast<unsigned long long*>(p);
    const int tiles = cdiv(n > 0 ? n : 1, kMinSortTile);
    p += align_up(sizeof(unsigned long long) * (size_t)tiles * kRadix);
    t.bytes = (size_t)(p - reinterpret_cast<char*>(temp));
    return t;
}

// One scatter pass.  CTA shape 512 x 8 (256x8, 256x16 and 512x4 measured within 3 % of it and are gone);
// the common digit widths get a kernel with the width known at compile time.
template <typename KeyT, int MODE, int BITS>
int launch_pass_bits(const KeyT* kin, const uint32_t* vin, KeyT* kout, uint32_t* vout, int n_cap,
                     const uint32_t* n_dev, int shift, uint32_t mask, int pass, const SortTemp& t, int tile16,
                     cudaStream_t stream) {
    constexpr int THREADS = kSortThreads, ITEMS = kSortItems;
    using Smem = SortSmem<KeyT, THREADS, ITEMS, MODE>;
    auto kernel = radix_onesweep_kernel<KeyT, THREADS, ITEMS, 2, MODE, BITS>;
    {  // opt in to > 48 KB of dynamic shared memory: per instantiation and per device
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
            HB_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                            cudaSharedmemCarveoutMaxShared));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    const int tiles = cdiv(n_cap, THREADS * ITEMS);
    kernel<<<tiles, THREADS, sizeof(Smem), stream>>>(kin, vin, kout, vout, n_cap, n_dev, shift, mask, pass,
                                                     t.hist + pass * kRadix, t.lookback, t.tickets, tile16);
    HB_LAUNCH_OK("radix_onesweep", stream, false);
    return 0;
}

template <typename KeyT, int MODE>
int launch_pass(const KeyT* kin, const uint32_t* vin, KeyT* kout, uint32_t* vout, int n_cap, const uint32_t* n_dev,
                int shift, uint32_t mask, int pass, const SortTemp& t, cudaStream_t stream, int tile16 = 0) {
#define HB_BITS(B) \
    launch_pass_bits<KeyT, MODE, B>(kin, vin, kout, vout, n_cap, n_dev, shift, mask, pass, t, tile16, stream)
    switch (__builtin_popcount(mask)) {
        case 8: return HB_BITS(8);
        case 7: return HB_BITS(7);
        case 6: return HB_BITS(6);
        default: return HB_BITS(0);
    }
#undef HB_BITS
}

// Run the scatter passes of `plan` (histogram already in t.hist, raw counts).  Pairs ping-pong
// between (k0,v0) and (k1,v1) starting from (k0,v0); v0 may be nullptr for "payload = index".
template <typename KeyT>
int run_passes(KeyT* k0, uint32_t* v0, KeyT* k1, uint32_t* v1, int n_cap, const uint32_t* n_dev,
               const SortPassPlan& plan, const SortTemp& t, cudaStream_t stream, bool* result_in_first) {
    radix_offsets_kernel<<<plan.num_passes, kRadix, 0, stream>>>(t.hist);
    HB_LAUNCH_OK("radix_offsets", stream, false);
    KeyT* kbuf[2] = {k0, k1};
    uint32_t* vbuf[2] = {v0, v1};
    int cur = 0;
    for (int p = 0; p < plan.num_passes; p++) {
        const uint32_t mask = (1u << plan.bits[p]) - 1u;
        const int rc = launch_pass<KeyT, kPairs>(kbuf[cur], vbuf[cur], kbuf[cur ^ 1], vbuf[cur ^ 1], n_cap, n_dev,
                                                 plan.shift[p], mask, p, t, stream);
        if (rc) return rc;
        cur ^= 1;
    }
    const bool in_first = (cur == 0);
    *result_in_first = in_first;
    return 0;
}
}  // namespace

size_t sort_temp_bytes(int n) { return carve_sort_temp(nullptr, n).bytes; }

SortPassPlan tile_pass_plan(int tile_bits) {
    SortPassPlan sp;
    memset(&sp, 0, sizeof(sp));
    if (tile_bits <= 0) tile_bits = 1;  // always at least one pass: the last pass is what unpacks the pairs
    const int n = cdiv(tile_bits, kRadixBits);
    const int base = tile_bits / n, extra = tile_bits % n;
    int shift = 32;
    for (int p = 0; p < n; p++) {
        sp.shift[p] = shift;
        sp.bits[p] = base + (p < extra ? 1 : 0);
        shift += sp.bits[p];
    }
    sp.num_passes = n;
    return sp;
}

int sort_temp_clear(void* temp, size_t temp_bytes, int n_cap, cudaStream_t stream) {
    const SortTemp t = carve_sort_temp(temp, n_cap);
    if (temp == nullptr || temp_bytes < t.bytes) {
        set_error("sort: temp storage too small (%zu < %zu)", temp_bytes, t.bytes);
        return -1;
    }
    HB_CUDA_OK(cudaMemsetAsync(temp, 0, t.bytes, stream));
    return 0;
}

uint32_t* sort_temp_histogram(void* temp) { return carve_sort_temp(temp, 1).hist; }

// Stable argsort of the P depth keys (u32).  keys0 is read only; the sorted keys land in kC and the
// permutation (original indices) in vC.  kB/vB are scratch.
int launch_depth_argsort(const uint32_t* keys0, uint32_t* kB, uint32_t* vB, uint32_t* kC, uint32_t* vC, int P,
                         void* temp, size_t temp_bytes, cudaStream_t stream) {
    if (P <= 0) return 0;
    int rc = sort_temp_clear(temp, temp_bytes, P, stream);
    if (rc) return rc;
    const SortTemp t = carve_sort_temp(temp, P);
    SortPassPlan sp;
    memset(&sp, 0, sizeof(sp));
    sp.num_passes = 4;
    for (int p = 0; p < 4; p++) {
        sp.shift[p] = 8 * p;
        sp.bits[p] = 8;
    }
    int hist_blocks = cdiv(P, kHistThreads * 8);
    if (hist_blocks > 148 * 2) hist_blocks = 148 * 2;
    radix_histogram_kernel<uint32_t><<<hist_blocks, kHistThreads, 0, stream>>>(keys0, P, sp, t.hist);
    HB_LAUNCH_OK("radix_histogram", stream, false);
    radix_offsets_kernel<<<4, kRadix, 0, stream>>>(t.hist);
    HB_LAUNCH_OK("radix_offsets", stream, false);
    // keys0 -> B -> C -> B -> C
    const uint32_t* kin[4] = {keys0, kB, kC, kB};
    const uint32_t* vin[4] = {nullptr, vB, vC, vB};
    uint32_t* kout[4] = {kB, kC, kB, kC};
    uint32_t* vout[4] = {vB, vC, vB, vC};
    for (int p = 0; p < 4; p++) {
        rc = launch_pass<uint32_t, kPairs>(kin[p], vin[p], kout[p], vout[p], P, nullptr, 8 * p, 255u, p, t, stream);
        if (rc) return rc;
    }
    return 0;
}

// Stable sort of the emitted pairs on the tile bits only.  The pairs arrive packed as
// (tile << 32 | Gaussian index) in depth order; the digit histograms were accumulated by the
// duplicate kernel into sort_temp_histogram(temp) (cleared by sort_temp_clear before it).  All
// passes but the last move 8 B per pair; the last one reads from `b` and writes list_out = index
// and tiles_out = tile id (u16 when tile16, else u32; tiles_out may alias `a`, which is dead by
// then).  With an odd number of passes the packed pairs must start in `b`, with an even number in `a`.
int launch_tile_sort(uint64_t* a, uint64_t* b, void* tiles_out, uint32_t* list_out, bool tile16, int n_cap,
                     const uint32_t* n_dev, const SortPassPlan& plan, void* temp, size_t temp_bytes,
                     cudaStream_t stream) {
    if (n_cap <= 0 || plan.num_passes <= 0) return 0;
    const SortTemp t = carve_sort_temp(temp, n_cap);
    if (temp == nullptr || temp_bytes < t.bytes) {
        set_error("tile_sort: temp storage too small (%zu < %zu)", temp_bytes, t.bytes);
        return -1;
    }
    radix_offsets_kernel<<<plan.num_passes, kRadix, 0, stream>>>(t.hist);
    HB_LAUNCH_OK("radix_offsets", stream, false);
    uint64_t* buf[2] = {a, b};
    int cur = (plan.num_passes % 2 == 0) ? 0 : 1;
    for (int p = 0; p < plan.num_passes; p++) {
        const uint32_t mask = (1u << plan.bits[p]) - 1u;
        int rc;
        if (p + 1 < plan.num_passes)
            rc = launch_pass<uint64_t, kKeysOnly>(buf[cur], nullptr, buf[cur ^ 1], nullptr, n_cap, n_dev, plan.shift[p],
                                                  mask, p, t, stream);
        else
            rc = launch_pass<uint64_t, kSplit>(buf[cur], nullptr, reinterpret_cast<uint64_t*>(tiles_out), list_out, n_cap,
                                               n_dev, plan.shift[p], mask, p, t, stream, tile16 ? 1 : 0);
        if (rc) return rc;
        cur ^= 1;
    }
    return 0;
}

// Generic entry (C ABI hugs_b200_sort_pairs): sorts on key bits [0, end_bit) with 8-bit digits.
// The pairs ping-pong between (keys_in, vals_in) and (keys_out, vals_out); *result_in_input tells
// the caller where they ended up.
int launch_sort_pairs_pingpong(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                               int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream,
                               bool* result_in_input, int begin_bit) {
    *result_in_input = true;
    if (end_bit > 64) end_bit = 64;
    if (begin_bit < 0) begin_bit = 0;
    if (n <= 0 || end_bit <= begin_bit) return 0;
    int rc = sort_temp_clear(temp, temp_bytes, n, stream);
    if (rc) return rc;
    const SortTemp t = carve_sort_temp(temp, n);
    SortPassPlan sp;
    memset(&sp, 0, sizeof(sp));
    if (begin_bit == 0) {
        sp.num_passes = cdiv(end_bit, kRadixBits);
        for (int p = 0; p < sp.num_passes; p++) {
            sp.shift[p] = p * kRadixBits;
            sp.bits[p] = (end_bit - sp.shift[p]) < kRadixBits ? (end_bit - sp.shift[p]) : kRadixBits;
        }
    } else {
        // a bit range is split into equal digits (what the rasterizer does for the tile bits)
        const int nbits = end_bit - begin_bit, np = cdiv(nbits, kRadixBits);
        int shift = begin_bit;
        for (int p = 0; p < np; p++) {
            sp.shift[p] = shift;
            sp.bits[p] = nbits / np + (p < nbits % np ? 1 : 0);
            shift += sp.bits[p];
        }
        sp.num_passes = np;
    }
    int hist_blocks = cdiv(n, kHistThreads * 8);
    if (hist_blocks > 148 * 4) hist_blocks = 148 * 4;
    radix_histogram_kernel<uint64_t><<<hist_blocks, kHistThreads, 0, stream>>>(keys_in, n, sp, t.hist);
    HB_LAUNCH_OK("radix_histogram", stream, false);
    return run_passes<uint64_t>(keys_in, vals_in, keys_out, vals_out, n, nullptr, sp, t, stream, result_in_input);
}

int launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                      int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream, int begin_bit) {
    bool in_input = false;
    const int rc = launch_sort_pairs_pingpong(keys_in, vals_in, keys_out, vals_out, n, end_bit, temp, temp_bytes,
                                              stream, &in_input, begin_bit);
    if (rc) return rc;
    if (in_input && n > 0) {
        HB_CUDA_OK(cudaMemcpyAsync(keys_out, keys_in, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream));
        HB_CUDA_OK(cudaMemcpyAsync(vals_out, vals_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream));
    }
    return 0;
}

}  // namespace hb
