// Single-pass inclusive prefix sum of u32 with decoupled look-back (chained scan), sm_100a.
// Replaces cub::DeviceScan::InclusiveSum at cuda_rasterizer/rasterizer_impl.cu:277.
//
// One CTA per 8192-item tile (the look-back walks its predecessors 32 at a time, one L2 round trip per
// window: with 2048-item tiles the last of 543 tiles needed up to 17 dependent round trips -- the stage took
// 22 us for 9 MB; 136 tiles need at most 5), tiles handed out by an atomic ticket so that every predecessor
// of a running tile has already started (forward progress for the look-back spin).  Per-tile
// state is one 64-bit word (flag << 32 | value) so flag and value are read atomically.
// Algorithmic traffic: 4 B read + 4 B written per item.
#include "../../include/hugs_b200.h"
#include "common.cuh"

#include <atomic>

namespace hb {

namespace {
constexpr int kScanThreads = 512;
constexpr int kScanItems = 16;
constexpr int kScanTile = kScanThreads * kScanItems;
constexpr unsigned long long kFlagAggregate = 1ull << 32;
constexpr unsigned long long kFlagInclusive = 2ull << 32;

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// `gather` (optional): scan in[gather[i]] instead of in[i] -- the rasterizer scans the per-Gaussian tile
// counts in depth-sorted order without materialising the permuted array.
__global__ void __launch_bounds__(kScanThreads) inclusive_scan_kernel(const uint32_t* __restrict__ in,
                                                                      const uint32_t* __restrict__ gather,
                                                                      uint32_t* __restrict__ out, int n,
                                                                      unsigned long long* state,
                                                                      unsigned* ticket) {
    __shared__ unsigned s_tile;
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_prefix;
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int base = tile * kScanTile + tid * kScanItems;

    uint32_t v[kScanItems];
    const bool full = base + kScanItems <= n;
    if (gather != nullptr) {
        uint32_t g[kScanItems];
        if (full) {
#pragma unroll
            for (int q = 0; q < kScanItems / 4; q++) {
                const uint4 a = __ldg(reinterpret_cast<const uint4*>(gather + base) + q);
                g[4 * q + 0] = a.x; g[4 * q + 1] = a.y; g[4 * q + 2] = a.z; g[4 * q + 3] = a.w;
            }
#pragma unroll
            for (int i = 0; i < kScanItems; i++) v[i] = __ldg(in + g[i]);
        } else {
#pragma unroll
            for (int i = 0; i < kScanItems; i++) v[i] = (base + i < n) ? __ldg(in + __ldg(gather + base + i)) : 0u;
        }
    } else if (full) {
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(in + base) + q);
            v[4 * q + 0] = a.x; v[4 * q + 1] = a.y; v[4 * q + 2] = a.z; v[4 * q + 3] = a.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++) v[i] = (base + i < n) ? __ldg(in + base + i) : 0u;
    }
#pragma unroll
    for (int i = 1; i < kScanItems; i++) v[i] += v[i - 1];
    uint32_t incl = v[kScanItems - 1];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (unsigned)d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_excl = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; w++) {
        const uint32_t t = s_warp[w];
        if ((unsigned)w < warp) warp_excl += t;
        tile_total += t;
    }
    const uint32_t thread_excl = warp_excl + incl - v[kScanItems - 1];

    if (warp == 0) {
        if (lane == 0)
            st_relaxed_u64(state + tile, (tile == 0 ? kFlagInclusive : kFlagAggregate) | tile_total);
        uint32_t prefix = 0;
        int idx = tile - 1;
        while (idx >= 0) {
            const int j = idx - (int)lane;
            unsigned long long w;
            do {
                w = (j >= 0) ? ld_relaxed_u64(state + j) : kFlagInclusive;
            } while (__any_sync(0xffffffffu, (w >> 32) == 0));
            const unsigned incl_mask = __ballot_sync(0xffffffffu, (w >> 32) == 2);
            uint32_t val = (uint32_t)w;
            if (incl_mask) {
                const unsigned first = __ffs(incl_mask) - 1;
                if (lane > first) val = 0;
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
            prefix += val;
            if (incl_mask) break;
            idx -= 32;
        }
        if (lane == 0) {
            if (tile != 0) st_relaxed_u64(state + tile, kFlagInclusive | (unsigned long long)(prefix + tile_total));
            s_prefix = prefix;
        }
    }
    __syncthreads();
    const uint32_t add = s_prefix + thread_excl;
    if (full) {
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            uint4 a;
            a.x = v[4 * q + 0] + add; a.y = v[4 * q + 1] + add; a.z = v[4 * q + 2] + add; a.w = v[4 * q + 3] + add;
            reinterpret_cast<uint4*>(out + base)[q] = a;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++)
            if (base + i < n) out[base + i] = v[i] + add;
    }
}
}  // namespace

size_t scan_temp_bytes(int n) {
    const int tiles = cdiv(n > 0 ? n : 1, kScanTile);
    return align_up(sizeof(unsigned long long) * (size_t)tiles + 256);
}

int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, void* temp, size_t temp_bytes,
                          cudaStream_t stream, const uint32_t* gather) {
    if (n <= 0) return 0;
    const size_t need = scan_temp_bytes(n);
    if (temp == nullptr || temp_bytes < need) {
        set_error("inclusive_scan: temp storage too small (%zu < %zu)", temp_bytes, need);
        return -1;
    }
    if ((reinterpret_cast<uintptr_t>(in) & 15u) || (reinterpret_cast<uintptr_t>(out) & 15u) ||
        (reinterpret_cast<uintptr_t>(gather) & 15u)) {
        set_error("inclusive_scan: in/out must be 16-byte aligned");
        return -1;
    }
    const int tiles = cdiv(n, kScanTile);
    HB_CUDA_OK(cudaMemsetAsync(temp, 0, need, stream));
    unsigned* ticket = reinterpret_cast<unsigned*>(temp);
    unsigned long long* state = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(temp) + 256);
    inclusive_scan_kernel<<<tiles, kScanThreads, 0, stream>>>(in, gather, out, n, state, ticket);
    HB_LAUNCH_OK("inclusive_scan", stream, false);
    return 0;
}

}  // namespace hb

// ---------------------------------------------------------------------------------------------------------
// hugs_b200_inclusive_scan_static: the same scan with its look-back state in a file-scope device array, for
// callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.  Nothing below allocates.
// Every GPU that loads the library holds the array (16.3 KB) until exit.  All calls share it: calls that
// overlap on the GPU (two streams, two host threads) race on it -- serialise them.
namespace hb {
namespace {
constexpr int kStaticScanMaxItems = HUGS_B200_SCAN_STATIC_MAX_ITEMS;
// what scan_temp_bytes(kStaticScanMaxItems) asks for: ticket block + one word per tile, rounded up
constexpr size_t kStaticScanBytes =
    (256 + sizeof(unsigned long long) * (size_t)(kStaticScanMaxItems / kScanTile) + HB_ALIGN - 1) / HB_ALIGN * HB_ALIGN;
static_assert(kStaticScanMaxItems % kScanTile == 0, "whole tiles");

static __device__ __align__(256) unsigned char g_scan_scratch[kStaticScanBytes];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_scan_scratch_ptr[kMaxDevices];   // device address of g_scan_scratch, per device

// In host code `g_scan_scratch` is only a handle for the runtime: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
int static_scan_scratch(void** out) {
    int dev = -1;
    HB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("inclusive_scan_static: device ordinal %d out of range", dev);
        return HUGS_B200_EUNSUP;
    }
    void* p = g_scan_scratch_ptr[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        HB_CUDA_OK(cudaGetSymbolAddress(&p, g_scan_scratch));
        if (p == nullptr) {
            set_error("inclusive_scan_static: cudaGetSymbolAddress returned a null address");
            return HUGS_B200_EUNSUP;
        }
        g_scan_scratch_ptr[dev].store(p, std::memory_order_release);   // (same value from every thread)
    }
    *out = p;
    return 0;
}
}  // namespace
}  // namespace hb

extern "C" int hugs_b200_inclusive_scan_static_max_items(void) { return hb::kStaticScanMaxItems; }

extern "C" int hugs_b200_inclusive_scan_static(const uint32_t* in, uint32_t* out, int n, hugs_b200_stream_t stream) {
    if (n < 0 || (n > 0 && (!in || !out))) {
        hb::set_error("inclusive_scan_static: bad arguments");
        return HUGS_B200_EINVAL;
    }
    if (n > hb::kStaticScanMaxItems) {
        hb::set_error("inclusive_scan_static: %d items exceed the static scratch's maximum of %d "
                      "(use hugs_b200_inclusive_scan with caller-provided temp storage)", n, hb::kStaticScanMaxItems);
        return HUGS_B200_EUNSUP;
    }
    if (n == 0) return 0;
    void* scratch = nullptr;
    const int rc = hb::static_scan_scratch(&scratch);
    if (rc) return rc;
    // launch_inclusive_scan clears the state it uses (ticket + one word per tile, cudaMemsetAsync on `stream`):
    // the array keeps the last call's contents otherwise
    return hb::launch_inclusive_scan(in, out, n, scratch, hb::kStaticScanBytes, static_cast<cudaStream_t>(stream),
                                     nullptr);
}
