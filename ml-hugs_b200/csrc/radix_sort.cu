// Hand-written onesweep-style LSD radix sort of (u64 key, u32 value) pairs for sm_100a.
// Replaces cub::DeviceRadixSort::SortPairs at cuda_rasterizer/rasterizer_impl.cu:303-308
// (CUB is used only as a correctness oracle in tests/).
//
// Structure (8-bit digits):
//   1. one histogram kernel reads the keys once and builds the digit histogram of every pass;
//   2. a tiny kernel turns each histogram into exclusive bin offsets;
//   3. one scatter kernel per pass.  Each CTA owns a 4096-pair tile (atomic ticket order),
//      ranks its keys stably with warp-level match_any + per-warp counters, obtains the
//      tile's per-digit global offset by decoupled look-back over 64-bit (flag|count) words,
//      stages the tile digit-sorted in shared memory and writes it out in coalesced runs.
// Stability: a tile is loaded warp-striped (warp w owns 512 consecutive pairs, item k of lane l
// is pair 32k+l), items are ranked in k order and lanes in lane order, warps in warp order,
// tiles in ticket == index order.
// Algorithmic traffic: 8 B/pair (histogram) + 24 B/pair/pass.
#include "common.cuh"

namespace hb {

namespace {
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 pairs
constexpr int kMaxPasses = 8;
constexpr int kHistThreads = 512;
static_assert(kSortThreads == kRadix, "one thread per digit in the look-back phase");

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- 1. histograms of all passes in one read of the keys --------------------------------------
__global__ void __launch_bounds__(kHistThreads) radix_histogram_kernel(const uint64_t* __restrict__ keys, int n,
                                                                       int num_passes, int end_bit,
                                                                       uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[kMaxPasses * kRadix];
    for (int i = threadIdx.x; i < num_passes * kRadix; i += kHistThreads) s_hist[i] = 0;
    __syncthreads();
    const int stride = gridDim.x * kHistThreads;
    // Bits at and above end_bit are not part of the sort: mask them so the last digit matches
    // what the scatter pass extracts.
    const uint64_t keep = (end_bit >= 64) ? ~0ull : ((1ull << end_bit) - 1ull);
    const int n2 = n >> 1;  // 128-bit loads, two keys each
    for (int i = blockIdx.x * kHistThreads + threadIdx.x; i < n2; i += stride) {
        const ulonglong2 kk = __ldg(reinterpret_cast<const ulonglong2*>(keys) + i);
        const uint64_t k0 = kk.x & keep, k1 = kk.y & keep;
        for (int p = 0; p < num_passes; p++) {
            atomicAdd(&s_hist[p * kRadix + (int)((k0 >> (p * kRadixBits)) & (kRadix - 1))], 1u);
            atomicAdd(&s_hist[p * kRadix + (int)((k1 >> (p * kRadixBits)) & (kRadix - 1))], 1u);
        }
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        const uint64_t k0 = keys[n - 1] & keep;
        for (int p = 0; p < num_passes; p++)
            atomicAdd(&s_hist[p * kRadix + (int)((k0 >> (p * kRadixBits)) & (kRadix - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < num_passes * kRadix; i += kHistThreads) {
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
struct SortSmem {
    uint64_t keys[kSortTile];               // 32 KB
    uint32_t vals[kSortTile];               // 16 KB
    uint32_t warp_hist[kSortWarps][kRadix]; //  8 KB
    uint32_t local_start[kRadix];           //  1 KB  first staged slot of each digit
    int64_t bin_base[kRadix];               //  2 KB  global index of staged slot 0 of each digit
    uint32_t scan_warp[kSortWarps];
    unsigned tile;
};

__global__ void __launch_bounds__(kSortThreads)
    radix_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                          uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n, int shift,
                          uint32_t digit_mask, int pass, const uint32_t* __restrict__ pass_offsets,
                          unsigned long long* lookback, unsigned* tickets) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem& s = *reinterpret_cast<SortSmem*>(smem_raw);
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    const unsigned lanemask_lt = (1u << lane) - 1u;

    if (tid == 0) s.tile = atomicAdd(tickets + pass, 1u);
#pragma unroll
    for (int w = 0; w < kSortWarps; w++) s.warp_hist[w][tid] = 0;
    __syncthreads();
    const int tile = (int)s.tile;
    const int tile_base = tile * kSortTile;
    const int warp_base = tile_base + (int)warp * (32 * kSortItems);
    const bool full_tile = tile_base + kSortTile <= n;

    // Load keys, warp-striped.
    uint64_t key[kSortItems];
    if (full_tile) {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) key[k] = __ldg(keys_in + warp_base + k * 32 + lane);
    } else {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) {
            const int i = warp_base + k * 32 + (int)lane;
            key[k] = (i < n) ? __ldg(keys_in + i) : ~0ull;
        }
    }

    // Stable warp-local ranks.
    uint32_t rank[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const bool valid = full_tile || (warp_base + k * 32 + (int)lane < n);
        const unsigned active = __ballot_sync(0xffffffffu, valid);
        rank[k] = 0;
        if (valid) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & digit_mask;
            const unsigned peers = __match_any_sync(active, d);
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if ((int)lane == leader) {
                old = s.warp_hist[warp][d];
                s.warp_hist[warp][d] = old + __popc(peers);
            }
            old = __shfl_sync(peers, old, leader);
            rank[k] = old + __popc(peers & lanemask_lt);
        }
        __syncwarp();
    }
    __syncthreads();

    // Per digit (thread tid == digit): exclusive scan over warps, tile count, look-back.
    {
        const int d = tid;
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t t = s.warp_hist[w][d];
            s.warp_hist[w][d] = sum;
            sum += t;
        }
        const uint32_t tile_count = sum;
        const unsigned long long flag_agg = (unsigned long long)(2 * pass + 1) << 32;
        const unsigned long long flag_inc = (unsigned long long)(2 * pass + 2) << 32;
        unsigned long long* my = lookback + (size_t)tile * kRadix + d;
        st_relaxed_u64(my, (tile == 0 ? flag_inc : flag_agg) | tile_count);

        // block exclusive scan of tile_count over digits -> local_start
        uint32_t incl = tile_count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (unsigned)o) incl += t;
        }
        if (lane == 31) s.scan_warp[warp] = incl;
        __syncthreads();
        uint32_t base = 0;
        for (unsigned w = 0; w < warp; w++) base += s.scan_warp[w];
        const uint32_t local_start = base + incl - tile_count;
        s.local_start[d] = local_start;

        // decoupled look-back: sum of this digit's counts over all previous tiles
        uint32_t prefix = 0;
        for (int t = tile - 1; t >= 0; --t) {
            const unsigned long long* p = lookback + (size_t)t * kRadix + d;
            unsigned long long w;
            do {
                w = ld_relaxed_u64(p);
            } while ((w >> 32) < (flag_agg >> 32));
            prefix += (uint32_t)w;
            if ((w >> 32) == (flag_inc >> 32)) break;
        }
        if (tile != 0) st_relaxed_u64(my, flag_inc | (unsigned long long)(prefix + tile_count));
        s.bin_base[d] = (int64_t)__ldg(pass_offsets + d) + (int64_t)prefix - (int64_t)local_start;
    }
    __syncthreads();

    // Stage the tile digit-sorted in shared memory.
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const int i = warp_base + k * 32 + (int)lane;
        if (full_tile || i < n) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & digit_mask;
            const uint32_t pos = s.local_start[d] + s.warp_hist[warp][d] + rank[k];
            s.keys[pos] = key[k];
            s.vals[pos] = __ldg(vals_in + i);
        }
    }
    __syncthreads();

    // Coalesced write-out: consecutive staged slots of one digit are consecutive in global memory.
    const int count = full_tile ? kSortTile : (n - tile_base);
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const int i = k * kSortThreads + tid;
        if (i < count) {
            const uint64_t kk = s.keys[i];
            const uint32_t d = (uint32_t)(kk >> shift) & digit_mask;
            const int64_t g = s.bin_base[d] + i;
            keys_out[g] = kk;
            vals_out[g] = s.vals[i];
        }
    }
}

struct SortTemp {
    uint32_t* hist;        // [kMaxPasses][kRadix]
    unsigned* tickets;     // [kMaxPasses]
    unsigned long long* lookback;  // [tiles][kRadix]
    size_t bytes;
};

SortTemp carve_sort_temp(void* temp, int n) {
    SortTemp t;
    char* p = reinterpret_cast<char*>(temp);
    t.hist = reinterpret_cast<uint32_t*>(p);
    p += align_up(sizeof(uint32_t) * kMaxPasses * kRadix);
    t.tickets = reinterpret_cast<unsigned*>(p);
    p += align_up(sizeof(unsigned) * kMaxPasses);
    t.lookback = reinterpret_cast<unsigned long long*>(p);
    const int tiles = cdiv(n > 0 ? n : 1, kSortTile);
    p += align_up(sizeof(unsigned long long) * (size_t)tiles * kRadix);
    t.bytes = (size_t)(p - reinterpret_cast<char*>(temp));
    return t;
}
}  // namespace

size_t sort_temp_bytes(int n) { return carve_sort_temp(nullptr, n).bytes; }

// Sorts on key bits [0, end_bit).  The pairs ping-pong between (keys_in, vals_in) and
// (keys_out, vals_out); *result_in_input tells the caller where they ended up.
int launch_sort_pairs_pingpong(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                               int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream,
                               bool* result_in_input) {
    *result_in_input = true;
    if (n <= 0 || end_bit <= 0) return 0;
    if (end_bit > 64) end_bit = 64;
    const int num_passes = cdiv(end_bit, kRadixBits);
    const SortTemp t = carve_sort_temp(temp, n);
    if (temp == nullptr || temp_bytes < t.bytes) {
        set_error("sort_pairs: temp storage too small (%zu < %zu)", temp_bytes, t.bytes);
        return -1;
    }
    if ((reinterpret_cast<uintptr_t>(keys_in) & 15u) || (reinterpret_cast<uintptr_t>(keys_out) & 15u)) {
        set_error("sort_pairs: key buffers must be 16-byte aligned");
        return -1;
    }
    HB_CUDA_OK(cudaFuncSetAttribute(radix_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)sizeof(SortSmem)));
    HB_CUDA_OK(cudaMemsetAsync(temp, 0, t.bytes, stream));
    int hist_blocks = cdiv(n, kHistThreads * 2 * 8);
    if (hist_blocks > 148 * 4) hist_blocks = 148 * 4;
    if (hist_blocks < 1) hist_blocks = 1;
    radix_histogram_kernel<<<hist_blocks, kHistThreads, 0, stream>>>(keys_in, n, num_passes, end_bit, t.hist);
    HB_LAUNCH_OK("radix_histogram", stream, false);
    radix_offsets_kernel<<<num_passes, kRadix, 0, stream>>>(t.hist);
    HB_LAUNCH_OK("radix_offsets", stream, false);
    const int tiles = cdiv(n, kSortTile);
    uint64_t* kin = keys_in;
    uint32_t* vin = vals_in;
    uint64_t* kout = keys_out;
    uint32_t* vout = vals_out;
    for (int p = 0; p < num_passes; p++) {
        const int shift = p * kRadixBits;
        const int bits = (end_bit - shift) < kRadixBits ? (end_bit - shift) : kRadixBits;
        const uint32_t mask = (1u << bits) - 1u;
        radix_onesweep_kernel<<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(
            kin, vin, kout, vout, n, shift, mask, p, t.hist + p * kRadix, t.lookback, t.tickets);
        HB_LAUNCH_OK("radix_onesweep", stream, false);
        uint64_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    *result_in_input = (kin == keys_in);
    return 0;
}

int launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, int n,
                      int end_bit, void* temp, size_t temp_bytes, cudaStream_t stream) {
    bool in_input = false;
    const int rc = launch_sort_pairs_pingpong(keys_in, vals_in, keys_out, vals_out, n, end_bit, temp, temp_bytes,
                                              stream, &in_input);
    if (rc) return rc;
    if (in_input && n > 0) {
        HB_CUDA_OK(cudaMemcpyAsync(keys_out, keys_in, sizeof(uint64_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream));
        HB_CUDA_OK(cudaMemcpyAsync(vals_out, vals_in, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream));
    }
    return 0;
}

}  // namespace hb
