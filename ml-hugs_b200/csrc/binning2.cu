// Two-level tile binning, level 2: super-tile lists -> per-tile lists, for sm_100a.
// Together with level 1 (binning.cu / radix_sort.cu at super-tile granularity) this replaces
// duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges
// (cuda_rasterizer/rasterizer_impl.cu:70-138, 289-310) and produces the same point_list and ranges
// bit for bit.
//
// What the reference sorts is R (Gaussian, tile) pairs on tile | depth.  After the depth argsort of the
// P Gaussians (radix_sort.cu) the pairs can be PRODUCED in depth order, so all that remains is a stable
// split of depth-ordered pairs by tile.  Round 1 did that split on the R pairs themselves (duplicate +
// two radix passes over R = 12 M pairs: 0.47 ms, latency- and issue-bound).  Here it is done in two
// levels, and the expensive level never touches a per-pair intermediate:
//   level 1  the depth-ordered Gaussians are split by SUPER-TILE (8x4 tiles = 128x64 pixels): R1 ~ 1.5 P
//            entries instead of R ~ 11 P, one 8-bit onesweep pass at 1080p (255 super-tiles);
//   level 2  a super-tile has 32 tiles: one 32-bit mask per list entry says which of them the Gaussian's
//            rectangle covers.  A warp transposes the masks of 32 consecutive entries (5 shuffle stages):
//            lane j then holds WHICH entries cover tile j, already in depth order, so counting is a
//            popcount and the stable rank of an entry within tile j is its bit position -- no sort, no
//            atomics, no keys.  A chunk of 256 entries is one CTA; chunk counts are scanned per tile
//            (that scan IS identifyTileRanges), then every chunk stages its 32 tile runs in shared memory
//            and writes each as one contiguous piece: 4 bytes written per pair, nothing else.
// Kernels: chunk table -> count -> per-tile scan over chunks -> scan over tiles (ranges, R) -> emit.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kL2Chunk = 256;  // list entries per CTA
constexpr int kL2Warps = kL2Chunk / 32;
constexpr unsigned kFull = 0xffffffffu;

// Which of the 32 tiles of super-tile (sx, sy) the tile rectangle [x0,x1) x [y0,y1) covers:
// bit 8 ly + lx is tile (8 sx + lx, 4 sy + ly).
__device__ __forceinline__ uint32_t super_tile_mask(uint2 rect, unsigned sx, unsigned sy) {
    const int x0 = (int)(rect.x & 0xffffu) - (int)(8u * sx), x1 = (int)(rect.x >> 16) - (int)(8u * sx);
    const int y0 = (int)(rect.y & 0xffffu) - (int)(4u * sy), y1 = (int)(rect.y >> 16) - (int)(4u * sy);
    const int lx0 = max(x0, 0), lx1 = min(x1, 8), ly0 = max(y0, 0), ly1 = min(y1, 4);
    if (lx1 <= lx0 || ly1 <= ly0) return 0u;
    const uint32_t row = ((1u << (lx1 - lx0)) - 1u) << lx0;          // 8 bits
    const uint32_t rows = ((1u << (ly1 - ly0)) - 1u) << ly0;         // 4 bits
    const uint32_t spread = (rows & 1u) | ((rows & 2u) << 7) | ((rows & 4u) << 14) | ((rows & 8u) << 21);
    return row * spread;  // no carries: row < 256
}

// 32 x 32 bit-matrix transpose across a warp: in, lane i holds row i; out, lane j holds column j
// (bit i = row i's bit j).
__device__ __forceinline__ uint32_t transpose32(uint32_t x, unsigned lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        // bits whose index has bit s clear
        const uint32_t lo = (s == 16) ? 0x0000ffffu : (s == 8) ? 0x00ff00ffu : (s == 4) ? 0x0f0f0f0fu
                          : (s == 2) ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(kFull, x, s);
        x = (lane & (unsigned)s) ? ((x & ~lo) | ((y & ~lo) >> s)) : ((x & lo) | ((y & lo) << s));
    }
    return x;
}

// ---- chunk table ---------------------------------------------------------------------------------------
// One CTA.  st_chunk_start[s] = number of chunks of the super-tiles before s; chunks[c] = (s | n << 16,
// first list position).  A super-tile's list is cut into chunks of 256 entries.
__global__ void __launch_bounds__(1024)
    l2_chunk_table_kernel(const uint2* __restrict__ st_ranges, int n_st, uint32_t* __restrict__ st_chunk_start,
                          uint2* __restrict__ chunks, uint32_t* __restrict__ n_chunks_out, uint32_t chunk_cap) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n_st; base += 1024) {
        const int s = base + tid;
        uint32_t nc = 0;
        if (s < n_st) {
            const uint2 r = st_ranges[s];
            nc = (r.y - r.x + (kL2Chunk - 1)) / kL2Chunk;
        }
        uint32_t incl = nc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= (unsigned)d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(kFull, wi, d);
                if (lane >= (unsigned)d) wi += t;
            }
            s_warp[lane] = wi - w;  // exclusive prefix of the warp totals
        }
        __syncthreads();
        const uint32_t excl = s_carry + s_warp[warp] + incl - nc;
        if (s < n_st) st_chunk_start[s] = excl;
        __syncthreads();
        if (tid == 1023) s_carry = excl + nc;
        __syncthreads();
    }
    const uint32_t total = s_carry;
    const uint32_t n_chunks = min(total, chunk_cap);
    if (tid == 0) {
        st_chunk_start[n_st] = total;
        *n_chunks_out = n_chunks;
    }
    __syncthreads();  // st_chunk_start (global, written by this CTA) is read back below
    for (uint32_t c = tid; c < n_chunks; c += 1024) {
        int lo = 0, hi = n_st;  // largest s with st_chunk_start[s] <= c
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (st_chunk_start[mid] <= c) lo = mid; else hi = mid;
        }
        const uint2 r = st_ranges[lo];
        const uint32_t begin = r.x + (c - st_chunk_start[lo]) * kL2Chunk;
        const uint32_t n = min((uint32_t)kL2Chunk, r.y - begin);
        chunks[c] = make_uint2((uint32_t)lo | (n << 16), begin);
    }
}

// Column masks of this warp's 32 entries of the chunk: lane j gets the entries that cover tile j.
__device__ __forceinline__ uint32_t chunk_columns(const uint2 desc, const uint32_t* __restrict__ st_list,
                                                  const uint2* __restrict__ rects, unsigned gx_s, uint32_t& id) {
    const unsigned s = desc.x & 0xffffu, n = desc.x >> 16;
    const unsigned sx = s % gx_s, sy = s / gx_s;
    uint32_t mask = 0;
    id = 0;
    if (threadIdx.x < n) {
        id = __ldg(st_list + desc.y + threadIdx.x);
        mask = super_tile_mask(__ldg(rects + id), sx, sy);
    }
    return transpose32(mask, threadIdx.x & 31u);
}

// ---- count: counts[c][j] = pairs of chunk c in tile j of its super-tile ----------------------------------
__global__ void __launch_bounds__(kL2Chunk)
    l2_count_kernel(const uint2* __restrict__ chunks, const uint32_t* __restrict__ n_chunks,
                    const uint32_t* __restrict__ st_list, const uint2* __restrict__ rects, unsigned gx_s,
                    uint32_t* __restrict__ counts) {
    __shared__ uint32_t s_cnt[kL2Warps][32];
    const unsigned c = blockIdx.x;
    if (c >= __ldg(n_chunks)) return;
    uint32_t id;
    const uint32_t col = chunk_columns(__ldg(chunks + c), st_list, rects, gx_s, id);
    s_cnt[threadIdx.x >> 5][threadIdx.x & 31u] = __popc(col);
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < kL2Warps; w++) sum += s_cnt[w][threadIdx.x];
        counts[(size_t)c * 32 + threadIdx.x] = sum;
    }
}

__device__ __forceinline__ int tile_of(unsigned s, unsigned j, unsigned gx_s, unsigned grid_x, unsigned grid_y) {
    const unsigned tx = 8u * (s % gx_s) + (j & 7u), ty = 4u * (s / gx_s) + (j >> 3);
    return (tx < grid_x && ty < grid_y) ? (int)(ty * grid_x + tx) : -1;
}

// ---- per tile: exclusive scan of its counts over the chunks of its super-tile ------------------------------
// Thread (s, j).  counts[c][j] becomes the number of pairs of tile j in the chunks before c; the total goes
// to tile_total.
__global__ void __launch_bounds__(256)
    l2_scan_chunks_kernel(const uint32_t* __restrict__ st_chunk_start, int n_st, const uint32_t* __restrict__ n_chunks,
                          uint32_t* __restrict__ counts, unsigned gx_s, unsigned grid_x, unsigned grid_y,
                          uint32_t* __restrict__ tile_total) {
    const unsigned g = blockIdx.x * 256u + threadIdx.x;
    const unsigned s = g >> 5, j = g & 31u;
    if (s >= (unsigned)n_st) return;
    const uint32_t limit = __ldg(n_chunks);
    const uint32_t c0 = min(__ldg(st_chunk_start + s), limit), c1 = min(__ldg(st_chunk_start + s + 1), limit);
    uint32_t run = 0;
    uint32_t c = c0;
    for (; c + 4 <= c1; c += 4) {
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = counts[(size_t)(c + k) * 32 + j];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            counts[(size_t)(c + k) * 32 + j] = run;
            run += v[k];
        }
    }
    for (; c < c1; c++) {
        const uint32_t v = counts[(size_t)c * 32 + j];
        counts[(size_t)c * 32 + j] = run;
        run += v;
    }
    const int t = tile_of(s, j, gx_s, grid_x, grid_y);
    if (t >= 0) tile_total[t] = run;
}

// ---- over tiles: exclusive scan of the totals = identifyTileRanges ------------------------------------------
// One CTA.  ranges[t] = [start, start + total), (0,0) for a tile without pairs (the reference's memset,
// rasterizer_impl.cu:310), clamped to the list capacity (an overflowing frame is redone by the host; until
// then nothing may point past the workspace).
__global__ void __launch_bounds__(1024)
    l2_scan_tiles_kernel(const uint32_t* __restrict__ tile_total, int num_tiles, uint32_t cap,
                         uint32_t* __restrict__ tile_start, uint2* __restrict__ ranges) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x;
    const unsigned lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    constexpr int kPer = 8;
    for (int base = 0; base < num_tiles; base += 1024 * kPer) {
        const int t0 = base + tid * kPer;
        uint32_t v[kPer], sum = 0;
#pragma unroll
        for (int k = 0; k < kPer; k++) {
            v[k] = (t0 + k < num_tiles) ? tile_total[t0 + k] : 0u;
            sum += v[k];
        }
        uint32_t incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= (unsigned)d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(kFull, wi, d);
                if (lane >= (unsigned)d) wi += t;
            }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        uint32_t run = s_carry + s_warp[warp] + incl - sum;
#pragma unroll
        for (int k = 0; k < kPer; k++) {
            if (t0 + k < num_tiles) {
                tile_start[t0 + k] = run;
                const uint32_t a = min(run, cap), b = min(run + v[k], cap);
                ranges[t0 + k] = (v[k] != 0u) ? make_uint2(a, b) : make_uint2(0u, 0u);
            }
            run += v[k];
        }
        __syncthreads();
        if (tid == 1023) s_carry = run;
        __syncthreads();
    }
}

// ---- emit: every chunk writes its 32 tile runs ---------------------------------------------------------------
__global__ void __launch_bounds__(kL2Chunk)
    l2_emit_kernel(const uint2* __restrict__ chunks, const uint32_t* __restrict__ n_chunks,
                   const uint32_t* __restrict__ st_list, const uint2* __restrict__ rects, unsigned gx_s,
                   unsigned grid_x, unsigned grid_y, const uint32_t* __restrict__ counts,
                   const uint32_t* __restrict__ tile_start, uint32_t* __restrict__ list, uint32_t cap) {
    __shared__ uint32_t s_stage[kL2Chunk * 32];   // the chunk's pairs, tile run after tile run
    __shared__ uint32_t s_cnt[kL2Warps][32];
    __shared__ uint32_t s_run_off[32], s_total[32], s_gbase[32];
    const unsigned c = blockIdx.x;
    if (c >= __ldg(n_chunks)) return;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint2 desc = __ldg(chunks + c);
    uint32_t id;
    uint32_t col = chunk_columns(desc, st_list, rects, gx_s, id);
    const uint32_t mine = __popc(col);
    s_cnt[warp][lane] = mine;
    __syncthreads();
    if (warp == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < kL2Warps; w++) total += s_cnt[w][lane];
        uint32_t incl = total;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= (unsigned)d) incl += t;
        }
        s_run_off[lane] = incl - total;
        s_total[lane] = total;
        const int t = tile_of(desc.x & 0xffffu, lane, gx_s, grid_x, grid_y);
        s_gbase[lane] = (t >= 0 && total != 0u) ? __ldg(tile_start + t) + __ldg(counts + (size_t)c * 32 + lane) : 0u;
    }
    __syncthreads();
    // lane j of warp w appends, in list order, the entries of its 32 that cover tile j
    uint32_t at = s_run_off[lane];
    for (unsigned w = 0; w < warp; w++) at += s_cnt[w][lane];
    const unsigned iters = __reduce_max_sync(kFull, mine);
    for (unsigned k = 0; k < iters; k++) {
        const int i = __ffs(col) - 1;              // -1 when this lane has run out
        const uint32_t v = __shfl_sync(kFull, id, i & 31);
        if (i >= 0) {
            s_stage[at + k] = v;
            col &= col - 1u;
        }
    }
    __syncthreads();
    // tile runs out, each a contiguous piece of the final list
    for (unsigned j = warp; j < 32u; j += kL2Warps) {
        const uint32_t total = s_total[j], off = s_run_off[j], gbase = s_gbase[j];
        for (uint32_t k = lane; k < total; k += 32u)
            if (gbase + k < cap) list[gbase + k] = s_stage[off + k];
    }
}

// keys[i] = tile << 32 | depth bits of list[i], for every list position of every tile (parity tests only)
__global__ void __launch_bounds__(256)
    keys_from_ranges_kernel(const uint32_t* __restrict__ tile_start, const uint32_t* __restrict__ tile_total,
                            const uint32_t* __restrict__ list, const uint32_t* __restrict__ depth_bits,
                            uint64_t* __restrict__ keys) {
    const uint32_t begin = __ldg(tile_start + blockIdx.x), end = begin + __ldg(tile_total + blockIdx.x);
    for (uint32_t i = begin + threadIdx.x; i < end; i += 256u)
        keys[i] = ((uint64_t)blockIdx.x << 32) | (uint64_t)__ldg(depth_bits + __ldg(list + i));
}
}  // namespace

L2Temp carve_l2_temp(char* base, int cap, int n_st, int num_tiles) {
    L2Temp t;
    char* p = base;
    const size_t chunk_cap = (size_t)(cap > 0 ? cap : 0) / kL2Chunk + (size_t)n_st + 1;
    auto take = [&](size_t bytes) {
        char* r = p;
        p += align_up(bytes);
        return r;
    };
    t.chunk_cap = (uint32_t)chunk_cap;
    t.counts = reinterpret_cast<uint32_t*>(take(chunk_cap * 32 * sizeof(uint32_t)));
    t.chunks = reinterpret_cast<uint2*>(take(chunk_cap * sizeof(uint2)));
    t.st_ranges = reinterpret_cast<uint2*>(take((size_t)n_st * sizeof(uint2)));
    t.st_chunk_start = reinterpret_cast<uint32_t*>(take(((size_t)n_st + 1) * sizeof(uint32_t)));
    t.tile_total = reinterpret_cast<uint32_t*>(take((size_t)num_tiles * sizeof(uint32_t)));
    t.tile_start = reinterpret_cast<uint32_t*>(take((size_t)num_tiles * sizeof(uint32_t)));
    t.n_chunks = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t)));
    t.total = (size_t)(p - base);
    return t;
}

// Level 2 of the binning.  `cap1` bounds the level-1 entries this launch is sized for (the real count lives
// on the device and is what st_ranges describe); `cap` is the capacity of `list`.
int launch_l2_split(const L2Temp& t, const uint32_t* st_list, const uint2* rects, int n_st, int gx_s, int grid_x,
                    int grid_y, int cap1, int cap, uint32_t* list, uint2* ranges, cudaStream_t stream) {
    const int num_tiles = grid_x * grid_y;
    const uint32_t chunk_limit = (uint32_t)min((long long)t.chunk_cap, (long long)cap1 / kL2Chunk + n_st + 1);
    l2_chunk_table_kernel<<<1, 1024, 0, stream>>>(t.st_ranges, n_st, t.st_chunk_start, t.chunks, t.n_chunks,
                                                  chunk_limit);
    HB_LAUNCH_OK("l2_chunk_table", stream, false);
    l2_count_kernel<<<chunk_limit, kL2Chunk, 0, stream>>>(t.chunks, t.n_chunks, st_list, rects, (unsigned)gx_s,
                                                          t.counts);
    HB_LAUNCH_OK("l2_count", stream, false);
    HB_CUDA_OK(cudaMemsetAsync(t.tile_total, 0, sizeof(uint32_t) * (size_t)num_tiles, stream));
    l2_scan_chunks_kernel<<<cdiv(n_st * 32, 256), 256, 0, stream>>>(t.st_chunk_start, n_st, t.n_chunks, t.counts,
                                                                    (unsigned)gx_s, (unsigned)grid_x,
                                                                    (unsigned)grid_y, t.tile_total);
    HB_LAUNCH_OK("l2_scan_chunks", stream, false);
    l2_scan_tiles_kernel<<<1, 1024, 0, stream>>>(t.tile_total, num_tiles, (uint32_t)cap, t.tile_start, ranges);
    HB_LAUNCH_OK("l2_scan_tiles", stream, false);
    l2_emit_kernel<<<chunk_limit, kL2Chunk, 0, stream>>>(t.chunks, t.n_chunks, st_list, rects, (unsigned)gx_s,
                                                         (unsigned)grid_x, (unsigned)grid_y, t.counts, t.tile_start,
                                                         list, (uint32_t)cap);
    HB_LAUNCH_OK("l2_emit", stream, false);
    return 0;
}

int launch_keys_from_ranges(int num_tiles, const uint32_t* tile_start, const uint32_t* tile_total,
                            const uint32_t* list, const uint32_t* depth_bits, uint64_t* keys, cudaStream_t stream) {
    if (num_tiles <= 0) return 0;
    keys_from_ranges_kernel<<<num_tiles, 256, 0, stream>>>(tile_start, tile_total, list, depth_bits, keys);
    HB_LAUNCH_OK("keys_from_ranges", stream, false);
    return 0;
}

}  // namespace hb
