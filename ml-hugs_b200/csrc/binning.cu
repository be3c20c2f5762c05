// Key duplication and tile-range identification for sm_100a.
//
// duplicate: replaces duplicateWithKeys (cuda_rasterizer/rasterizer_impl.cu:70-111).  For every
// visible Gaussian and every tile of its rectangle it emits key = (tile_id << 32) | depth_bits,
// value = Gaussian index, at offsets[idx-1] + (y-y0)*w + (x-x0): the same slot and the same
// (y outer, x inner) emission order as the reference, so that a stable sort reproduces its
// point_list bit for bit.  Instead of one thread looping over its own rectangle (imbalanced and
// uncoalesced), a CTA owns 256 consecutive Gaussians and all its threads walk the CTA's
// contiguous output range together: output slot -> owning Gaussian by binary search over the
// CTA's 257 offsets in shared memory, so stores are fully coalesced and perfectly balanced.
// Algorithmic traffic: 24 B per Gaussian + 12 B per emitted pair.
//
// tile_ranges: replaces identifyTileRanges (rasterizer_impl.cu:116-138) + the cudaMemset at :310.
#include "common.cuh"

namespace hb {

namespace {
constexpr int kDupThreads = 256;

__global__ void __launch_bounds__(kDupThreads)
    duplicate_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                     const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                     uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, unsigned grid_x, unsigned grid_y) {
    __shared__ uint32_t s_off[kDupThreads + 1];  // exclusive start offset of each Gaussian, [256] = CTA end
    __shared__ uint32_t s_depth[kDupThreads];
    __shared__ uint32_t s_rect[kDupThreads];  // min_x | min_y << 10 | width << 20 (grid <= 1023 tiles per axis)
    const int tid = threadIdx.x;
    const int first = blockIdx.x * kDupThreads;
    const int idx = first + tid;

    uint32_t end_off = 0;
    if (idx < P) {
        end_off = __ldg(offsets + idx);
        const int r = __ldg(radii + idx);
        uint32_t rect = 0, dbits = 0;
        if (r > 0) {
            // getRect (auxiliary.h:46-56), same float sequence as preprocess.
            const float2 p = __ldg(means2D + idx);
            const float rf = (float)r;
            const unsigned min_x = min(grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.x, -rf), 0.0625f)));
            const unsigned min_y = min(grid_y, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.y, -rf), 0.0625f)));
            const unsigned max_x = min(
                grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.x, rf), 16.f), -1.f), 0.0625f)));
            rect = min_x | (min_y << 10) | ((max_x - min_x) << 20);
            dbits = __float_as_uint(__ldg(depths + idx));
        }
        s_rect[tid] = rect;
        s_depth[tid] = dbits;
    }
    // s_off[t] = inclusive offset of Gaussian first+t-1 (== exclusive start of first+t)
    if (tid == 0) s_off[0] = (first == 0) ? 0u : __ldg(offsets + first - 1);
    const int in_block = min(kDupThreads, P - first);
    if (tid < in_block) s_off[tid + 1] = end_off;
    __syncthreads();

    const uint32_t begin = s_off[0];
    const uint32_t end = s_off[in_block];
    for (uint32_t o = begin + tid; o < end; o += kDupThreads) {
        // largest g with s_off[g] <= o  (Gaussians with zero tiles have s_off[g] == s_off[g+1])
        int lo = 0, hi = in_block;  // invariant: s_off[lo] <= o < s_off[hi]
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (s_off[mid] <= o) lo = mid; else hi = mid;
        }
        const uint32_t rect = s_rect[lo];
        const uint32_t k = o - s_off[lo];
        const uint32_t w = rect >> 20;
        const uint32_t y = ((rect >> 10) & 1023u) + k / w;
        const uint32_t x = (rect & 1023u) + k % w;
        const uint64_t key = ((uint64_t)(y * grid_x + x) << 32) | (uint64_t)s_depth[lo];
        keys[o] = key;
        vals[o] = (uint32_t)(first + lo);
    }
}

// Depth-sorted variant used by the rasterizer: CTA b owns the Gaussians at depth-sorted positions
// [256 b, 256 b + 256) (`perm` = argsort of the depth bits, `offsets` the inclusive scan of
// tiles_touched in that order).  Pairs therefore leave this kernel ordered by (depth bits, Gaussian
// index), and a stable sort on the tile bits alone reproduces the reference's order on the full
// 32+bit key.  Because the depth is implied by the position, a pair is emitted as ONE 64-bit word
// (tile << 32 | Gaussian index): 8 B instead of the reference's 12 B (8 B key + 4 B value); the
// last sort pass turns it back into (tile << 32 | depth bits, index).  While emitting, the CTA also
// counts the tile digits of the following sort passes (`plan`) into `hist`, so the sort needs no
// histogram pass of its own.  Emission is clamped to `cap` slots: the workspace is sized from the
// previous frames' pair count before the real count has reached the host (api.cu); an overflow is
// detected there and redone.
// kSuper: the units are super-tiles (8x4 tiles, binning2.cu) instead of tiles; the rectangle comes from the
// packed tile rectangle the preprocess kernel left (`rects`), grid_x is the super-tile grid width.
template <bool kSuper>
__global__ void __launch_bounds__(kDupThreads)
    duplicate_sorted_kernel(int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ offsets,
                            const float2* __restrict__ means2D, const int* __restrict__ radii,
                            const uint2* __restrict__ rects,
                            uint64_t* __restrict__ pairs, unsigned grid_x, unsigned grid_y, uint32_t cap,
                            const SortPassPlan plan, uint32_t* __restrict__ hist) {
    // Each lane loads ONE depth-sorted Gaussian (slot range + tile rectangle); the slots are then
    // written by groups of 8 lanes, every group walking the Gaussians its own 8 lanes hold, so four
    // Gaussians are emitted per warp step and a typical 5..16-tile rectangle takes one or two steps.
    // No slot-to-owner search, no CTA barrier in the emission loop: slot k of a Gaussian is tile
    // (min_y + k / w, min_x + k % w), and a rectangle's slots are contiguous in the output.
    // (History: per-slot binary search over the CTA's offsets, then a per-chunk max-scan of head
    // markers -- 250 instructions per 32 slots, 72 % issue-bound; this form needs about a third.)
    __shared__ uint32_t s_hist[3][256];
    __shared__ uint4 s_huge[kDupThreads];   // (start, len, rect, index) of rectangles the whole CTA emits
    __shared__ int s_nhuge;
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31u;
    const unsigned full = 0xffffffffu;
    for (int i = tid; i < plan.num_passes * 256; i += kDupThreads) s_hist[i >> 8][i & 255] = 0;
    if (tid == 0) s_nhuge = 0;
    __syncthreads();
    // Warp w of CTA b takes the 32 depth-sorted positions of chunk w * gridDim.x + b: neighbouring chunks
    // (the nearest Gaussians come first and have by far the largest rectangles) land on different CTAs
    // and SMs instead of piling up in the first few CTAs.
    const int pos = (int)(((unsigned)(tid >> 5) * gridDim.x + blockIdx.x) * 32u + lane);
    uint32_t end_off = 0, idx = 0;
    if (pos < P) end_off = __ldg(offsets + pos);
    uint32_t start = __shfl_up_sync(full, end_off, 1);
    if (lane == 0) start = (pos > 0 && pos < P) ? __ldg(offsets + pos - 1) : 0u;
    uint32_t len = (pos < P) ? end_off - start : 0u;
    // the workspace may be smaller than this frame needs (api.cu redoes the frame): never write past it
    len = (start < cap) ? min(len, cap - start) : 0u;
    uint32_t rect = 0;
    if (len != 0u) {
        idx = __ldg(perm + pos);
        if (kSuper) {
            const uint2 r = __ldg(rects + idx);
            const unsigned sx0 = (r.x & 0xffffu) >> 3, sx1 = ((r.x >> 16) + 7u) >> 3, sy0 = (r.y & 0xffffu) >> 2;
            rect = sx0 | (sy0 << 10) | ((sx1 - sx0) << 20);
        } else {
            // getRect (auxiliary.h:46-56), same float sequence as preprocess.
            const float2 p = __ldg(means2D + idx);
            const float rf = (float)__ldg(radii + idx);
            const unsigned min_x = min(grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.x, -rf), 0.0625f)));
            const unsigned min_y = min(grid_y, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.y, -rf), 0.0625f)));
            const unsigned max_x = min(
                grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.x, rf), 16.f), -1.f), 0.0625f)));
            rect = min_x | (min_y << 10) | ((max_x - min_x) << 20);
        }
    }
    // digit layout of the sort passes relative to the tile id
    const int sh0 = plan.shift[0] - 32, sh1 = plan.shift[1] - 32, sh2 = plan.shift[2] - 32;
    const uint32_t mk0 = (1u << plan.bits[0]) - 1u, mk1 = (1u << plan.bits[1]) - 1u, mk2 = (1u << plan.bits[2]) - 1u;
    // emits slots first, first + step, ... (< g_len) of one rectangle; `iters` is warp-uniform
    auto emit = [&](uint32_t first, uint32_t step, unsigned iters, uint32_t g_len, uint32_t g_start, uint32_t g_rect,
                    uint32_t g_index) {
        const uint32_t w = max(g_rect >> 20, 1u);
        const uint32_t base_tile = ((g_rect >> 10) & 1023u) * grid_x + (g_rect & 1023u);
        const float rcp = __frcp_rn((float)w);
        uint64_t* out = pairs + g_start;
        const uint64_t low = (uint64_t)g_index;
#pragma unroll 1
        for (unsigned i = 0; i < iters; ++i) {
            const uint32_t k = first + step * i;
            const bool valid = k < g_len;
            const unsigned active = __ballot_sync(full, valid);
            if (valid) {
                // k / w through the float reciprocal: k < 2^20 and w <= 1023, so the estimate is q or q - 1
                uint32_t q = (uint32_t)__fmul_rz((float)k, rcp);
                uint32_t r = k - q * w;
                if (r >= w) { q++; r -= w; }
                const uint32_t tile = base_tile + q * grid_x + r;
                out[k] = ((uint64_t)tile << 32) | low;
                // Digit counts for the sort passes.  Neighbouring slots are neighbouring tiles: their low
                // digits differ (plain shared-memory atomics), their higher digits mostly coincide
                // (lanes with equal digits are counted with one atomic by the lowest of them).
                atomicAdd(&s_hist[0][(tile >> sh0) & mk0], 1u);
                if (plan.num_passes > 1) {
                    const uint32_t d = (tile >> sh1) & mk1;
                    const unsigned same = __match_any_sync(active, d);
                    if ((int)lane == __ffs(same) - 1) atomicAdd(&s_hist[1][d], (uint32_t)__popc(same));
                }
                if (plan.num_passes > 2) {
                    const uint32_t d = (tile >> sh2) & mk2;
                    const unsigned same = __match_any_sync(active, d);
                    if ((int)lane == __ffs(same) - 1) atomicAdd(&s_hist[2][d], (uint32_t)__popc(same));
                }
            }
        }
    };
    // rectangles of more than kHuge tiles (close-ups: up to the whole tile grid) are queued for the whole
    // CTA; a warp left alone with 32 of them (a quarter of a million slots) was a 0.2 ms tail
    constexpr uint32_t kWide = 32, kHuge = 512;
    if (len > kHuge) s_huge[atomicAdd(&s_nhuge, 1)] = make_uint4(start, len, rect, idx);
    // rectangles of more than kWide tiles: the whole warp, one rectangle at a time
    unsigned wide = __ballot_sync(full, len > kWide && len <= kHuge);
    while (wide) {
        const int src = __ffs(wide) - 1;
        wide &= wide - 1u;
        const uint32_t g_len = __shfl_sync(full, len, src);
        emit(lane, 32u, (g_len + 31u) >> 5, g_len, __shfl_sync(full, start, src), __shfl_sync(full, rect, src),
             __shfl_sync(full, idx, src));
    }
    // the rest: groups of 8 lanes, each walking the rectangles its own 8 lanes hold
    const uint32_t small_len = (len > kWide) ? 0u : len;
    const unsigned sub = lane & 7u, group_base = lane & 24u;
#pragma unroll 1
    for (unsigned j = 0; j < 8; ++j) {
        const int src = (int)(group_base | j);
        const uint32_t g_len = __shfl_sync(full, small_len, src);
        const unsigned iters = __reduce_max_sync(full, (g_len + 7u) >> 3);      // warp-uniform trip count
        if (iters == 0u) continue;
        emit(sub, 8u, iters, g_len, __shfl_sync(full, start, src), __shfl_sync(full, rect, src),
             __shfl_sync(full, idx, src));
    }
    __syncthreads();
    for (int e = 0, n = s_nhuge; e < n; ++e) {
        const uint4 h = s_huge[e];
        emit((uint32_t)tid, (uint32_t)kDupThreads, (h.y + kDupThreads - 1) / kDupThreads, h.y, h.x, h.z, h.w);
    }
    __syncthreads();
    for (int i = tid; i < plan.num_passes * 256; i += kDupThreads) {
        const uint32_t c = s_hist[i >> 8][i & 255];
        if (c) atomicAdd(hist + i, c);
    }
}

// Generic fallback for tile grids wider than 1023 tiles per axis (images beyond 16K pixels):
// one thread per Gaussian, the reference's loop.
__global__ void __launch_bounds__(kDupThreads)
    duplicate_serial_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                            const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                            uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, unsigned grid_x,
                            unsigned grid_y) {
    const int idx = blockIdx.x * kDupThreads + threadIdx.x;
    if (idx >= P) return;
    const int r = __ldg(radii + idx);
    if (r <= 0) return;
    uint32_t off = (idx == 0) ? 0u : __ldg(offsets + idx - 1);
    const float2 p = __ldg(means2D + idx);
    const float rf = (float)r;
    const unsigned min_x = min(grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.x, -rf), 0.0625f)));
    const unsigned min_y = min(grid_y, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(p.y, -rf), 0.0625f)));
    const unsigned max_x =
        min(grid_x, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.x, rf), 16.f), -1.f), 0.0625f)));
    const unsigned max_y =
        min(grid_y, (unsigned)max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(p.y, rf), 16.f), -1.f), 0.0625f)));
    const uint64_t dbits = __float_as_uint(__ldg(depths + idx));
    for (unsigned y = min_y; y < max_y; y++)
        for (unsigned x = min_x; x < max_x; x++) {
            keys[off] = ((uint64_t)(y * grid_x + x) << 32) | dbits;
            vals[off] = (uint32_t)idx;
            off++;
        }
}

// ranges[t] = [first, last) of tile t in the sorted list; tiles without pairs keep (0,0).
// Two keys per thread, 128-bit loads; the neighbour comes from the previous thread's pair or one
// extra (cached) load.
__global__ void __launch_bounds__(256)
    tile_ranges_kernel(int R, const uint32_t* __restrict__ n_dev, const uint64_t* __restrict__ keys,
                       uint2* __restrict__ ranges) {
    if (n_dev != nullptr) R = (int)min((uint32_t)R, __ldg(n_dev));  // R was the capacity
    const int i0 = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i0 >= R) return;
    uint64_t k0, k1 = 0;
    const bool has1 = i0 + 1 < R;
    if (has1) {
        const ulonglong2 kk = __ldg(reinterpret_cast<const ulonglong2*>(keys + i0));
        k0 = kk.x;
        k1 = kk.y;
    } else {
        k0 = __ldg(keys + i0);
    }
    const uint32_t t0 = (uint32_t)(k0 >> 32);
    if (i0 == 0) {
        ranges[t0].x = 0;
    } else {
        const uint32_t tp = (uint32_t)(__ldg(keys + i0 - 1) >> 32);
        if (tp != t0) {
            ranges[tp].y = i0;
            ranges[t0].x = i0;
        }
    }
    if (has1) {
        const uint32_t t1 = (uint32_t)(k1 >> 32);
        if (t1 != t0) {
            ranges[t0].y = i0 + 1;
            ranges[t1].x = i0 + 1;
        }
        if (i0 + 1 == R - 1) ranges[t1].y = R;
    } else {
        ranges[t0].y = R;  // i0 == R-1
    }
}
// Tile ranges from the per-position tile ids (u16 / u32) that the two-level sort leaves behind: same
// boundaries as identifyTileRanges finds in the high words of the sorted keys.  8 positions per thread.
template <typename T>
__global__ void __launch_bounds__(256)
    tile_ranges_ids_kernel(int R, const uint32_t* __restrict__ n_dev, const T* __restrict__ ids,
                           uint2* __restrict__ ranges) {
    if (n_dev != nullptr) R = (int)min((uint32_t)R, __ldg(n_dev));  // R was the capacity
    constexpr int kPer = 8;
    const int i0 = kPer * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i0 >= R) return;
    uint32_t t[kPer];
    if (i0 + kPer <= R) {
        if (sizeof(T) == 2) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(ids + i0));
            t[0] = v.x & 0xffffu; t[1] = v.x >> 16; t[2] = v.y & 0xffffu; t[3] = v.y >> 16;
            t[4] = v.z & 0xffffu; t[5] = v.z >> 16; t[6] = v.w & 0xffffu; t[7] = v.w >> 16;
        } else {
            const uint4 v0 = __ldg(reinterpret_cast<const uint4*>(ids + i0));
            const uint4 v1 = __ldg(reinterpret_cast<const uint4*>(ids + i0) + 1);
            t[0] = v0.x; t[1] = v0.y; t[2] = v0.z; t[3] = v0.w;
            t[4] = v1.x; t[5] = v1.y; t[6] = v1.z; t[7] = v1.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < kPer; k++) t[k] = (i0 + k < R) ? (uint32_t)__ldg(ids + i0 + k) : 0u;
    }
    uint32_t prev = (i0 == 0) ? 0xffffffffu : (uint32_t)__ldg(ids + i0 - 1);
#pragma unroll
    for (int k = 0; k < kPer; k++) {
        const int i = i0 + k;
        if (i < R) {
            if (t[k] != prev) {
                if (i != 0) ranges[prev].y = (uint32_t)i;
                ranges[t[k]].x = (uint32_t)i;
            }
            if (i == R - 1) ranges[t[k]].y = (uint32_t)R;
            prev = t[k];
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
    materialize_keys_kernel(int R, const T* __restrict__ ids, const uint32_t* __restrict__ list,
                            const uint32_t* __restrict__ depth_bits, uint64_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    keys[i] = ((uint64_t)__ldg(ids + i) << 32) | (uint64_t)__ldg(depth_bits + __ldg(list + i));
}
}  // namespace

int launch_tile_ranges_ids(int R, const void* tile_ids, bool tile16, uint2* ranges, int num_tiles,
                           cudaStream_t stream, const uint32_t* n_dev) {
    HB_CUDA_OK(cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream));
    if (R <= 0) return 0;
    const int blocks = cdiv(cdiv(R, 8), 256);
    if (tile16)
        tile_ranges_ids_kernel<uint16_t><<<blocks, 256, 0, stream>>>(R, n_dev, static_cast<const uint16_t*>(tile_ids), ranges);
    else
        tile_ranges_ids_kernel<uint32_t><<<blocks, 256, 0, stream>>>(R, n_dev, static_cast<const uint32_t*>(tile_ids), ranges);
    HB_LAUNCH_OK("identify_tile_ranges", stream, false);
    return 0;
}

int launch_materialize_keys(int R, const void* tile_ids, bool tile16, const uint32_t* list,
                            const uint32_t* depth_bits, uint64_t* keys, cudaStream_t stream) {
    if (R <= 0) return 0;
    if (tile16)
        materialize_keys_kernel<uint16_t><<<cdiv(R, 256), 256, 0, stream>>>(R, static_cast<const uint16_t*>(tile_ids), list, depth_bits, keys);
    else
        materialize_keys_kernel<uint32_t><<<cdiv(R, 256), 256, 0, stream>>>(R, static_cast<const uint32_t*>(tile_ids), list, depth_bits, keys);
    HB_LAUNCH_OK("materialize_keys", stream, false);
    return 0;
}


int launch_duplicate(int P, const float2* means2D, const float* depths, const uint32_t* offsets, const int* radii,
                     uint64_t* keys, uint32_t* vals, int grid_x, int grid_y, cudaStream_t stream) {
    if (P <= 0) return 0;
    if (grid_x <= 1023 && grid_y <= 1023) {
        duplicate_kernel<<<cdiv(P, kDupThreads), kDupThreads, 0, stream>>>(P, means2D, depths, offsets, radii, keys,
                                                                           vals, (unsigned)grid_x, (unsigned)grid_y);
    } else {
        duplicate_serial_kernel<<<cdiv(P, kDupThreads), kDupThreads, 0, stream>>>(
            P, means2D, depths, offsets, radii, keys, vals, (unsigned)grid_x, (unsigned)grid_y);
    }
    HB_LAUNCH_OK("duplicate_with_keys", stream, false);
    return 0;
}

int launch_duplicate_sorted(int P, const uint32_t* perm, const uint32_t* offsets, const float2* means2D,
                            const int* radii, uint64_t* pairs, int grid_x, int grid_y, int cap,
                            const SortPassPlan& plan, uint32_t* hist, cudaStream_t stream) {
    if (P <= 0) return 0;
    if (grid_x > 1023 || grid_y > 1023 || plan.num_passes > 3) {
        set_error("duplicate_sorted: tile grid %dx%d too large", grid_x, grid_y);
        return HUGS_B200_EUNSUP_INTERNAL;
    }
    duplicate_sorted_kernel<false><<<cdiv(P, kDupThreads), kDupThreads, 0, stream>>>(
        P, perm, offsets, means2D, radii, nullptr, pairs, (unsigned)grid_x, (unsigned)grid_y, (uint32_t)cap, plan,
        hist);
    HB_LAUNCH_OK("duplicate_with_keys", stream, false);
    return 0;
}

// Level 1 of the two-level binning (binning2.cu): (super-tile << 32 | Gaussian index) entries in depth order.
// `offsets` is the inclusive scan of the per-Gaussian super-tile counts in depth order, gx_s x gy_s the
// super-tile grid.
int launch_duplicate_super(int P, const uint32_t* perm, const uint32_t* offsets, const uint2* rects, uint64_t* pairs,
                           int gx_s, int gy_s, int cap1, const SortPassPlan& plan, uint32_t* hist,
                           cudaStream_t stream) {
    if (P <= 0) return 0;
    if (gx_s > 1023 || gy_s > 1023 || plan.num_passes > 3) {
        set_error("duplicate_super: super-tile grid %dx%d too large", gx_s, gy_s);
        return HUGS_B200_EUNSUP_INTERNAL;
    }
    duplicate_sorted_kernel<true><<<cdiv(P, kDupThreads), kDupThreads, 0, stream>>>(
        P, perm, offsets, nullptr, nullptr, rects, pairs, (unsigned)gx_s, (unsigned)gy_s, (uint32_t)cap1, plan, hist);
    HB_LAUNCH_OK("duplicate_super", stream, false);
    return 0;
}

// R is the pair count, or -- when n_dev is given -- the capacity the launch is sized for while the
// real count is read from device memory.
int launch_tile_ranges(int R, const uint64_t* keys, uint2* ranges, int num_tiles, cudaStream_t stream,
                       const uint32_t* n_dev) {
    HB_CUDA_OK(cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream));
    if (R <= 0) return 0;
    tile_ranges_kernel<<<cdiv(cdiv(R, 2), 256), 256, 0, stream>>>(R, n_dev, keys, ranges);
    HB_LAUNCH_OK("identify_tile_ranges", stream, false);
    return 0;
}

}  // namespace hb
