// A second image of the FIRST n Gaussians from the joint pass's sorted lists.
//
// hugs/renderer/gs_renderer.py:68-82 renders the human Gaussians a second time, alone, whenever the
// trainer asks for the separate human image (hugs/trainer/gs_trainer.py:255-271): a complete second
// rasterizer pass (preprocess, scan, duplicate, sort, ranges, blend; and all of it again backwards)
// over Gaussians whose projection, conic, colour, depth order and tile lists the joint pass has just
// computed.  The human-only tile lists are the joint lists with the scene entries removed: the sort
// key is (tile, depth, index) in both passes, so a STABLE filter of a tile's joint list on
// `index < n` is, entry for entry, the list the second pass would have sorted.  This kernel is that
// filter; the blend kernels then run on the filtered lists unchanged (contributor counts of the
// layer are bit-identical to the reference's second pass).
//
// The filtered list of a tile is written at the tile's own offset of a second list buffer, so the
// layer's ranges are (joint begin, joint begin + kept) and no scan over tiles is needed; the hit
// masks of the layer (indexed by range begin and tile, render_forward.cu) stay disjoint for the
// same reason.  4 B read per joint pair, 4 B written per kept pair.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kFilterThreads = 256;
constexpr int kFilterWarps = kFilterThreads / 32;
constexpr int kFilterRounds = 4;   // list entries per thread and iteration (independent loads)
static_assert(kFilterWarps * kFilterRounds == 32, "one warp scans the (round, warp) counts");

__global__ void __launch_bounds__(kFilterThreads)
    layer_filter_kernel(int num_tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
                        uint32_t n_layer, uint32_t* __restrict__ list_out, uint2* __restrict__ ranges_out) {
    __shared__ uint32_t s_cnt[32];
    const int tile = (int)blockIdx.x;
    if (tile >= num_tiles) return;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint2 range = __ldg(ranges + tile);
    const uint32_t n = range.y - range.x;
    const uint32_t* const src = list + range.x;
    uint32_t out = range.x;   // next free slot of the filtered list (CTA-uniform)
    for (uint32_t base = 0; base < n; base += kFilterThreads * kFilterRounds) {
        uint32_t id[kFilterRounds];
        unsigned bal[kFilterRounds];
#pragma unroll
        for (int k = 0; k < kFilterRounds; k++) {
            const uint32_t pos = base + (uint32_t)k * kFilterThreads + threadIdx.x;
            id[k] = pos < n ? __ldg(src + pos) : 0xffffffffu;
        }
#pragma unroll
        for (int k = 0; k < kFilterRounds; k++) {
            bal[k] = __ballot_sync(0xffffffffu, id[k] < n_layer);
            if (lane == 0) s_cnt[k * kFilterWarps + warp] = (uint32_t)__popc(bal[k]);
        }
        __syncthreads();
        // list order within the iteration is (round, warp, lane): every warp scans the 32 counts itself
        const uint32_t mine = s_cnt[lane];
        uint32_t incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if ((int)lane >= d) incl += up;
        }
        const uint32_t excl = incl - mine;
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
        for (int k = 0; k < kFilterRounds; k++) {
            const uint32_t first = __shfl_sync(0xffffffffu, excl, k * kFilterWarps + (int)warp);
            if (id[k] < n_layer) list_out[out + first + (uint32_t)__popc(bal[k] & ((1u << lane) - 1u))] = id[k];
        }
        out += total;
        __syncthreads();   // s_cnt is rewritten by the next iteration
    }
    if (threadIdx.x == 0) ranges_out[tile] = make_uint2(range.x, out);
}
}  // namespace

int launch_layer_filter(int num_tiles, const uint2* ranges, const uint32_t* list, int n_layer, uint32_t* list_out,
                        uint2* ranges_out, cudaStream_t stream) {
    if (num_tiles <= 0) return 0;
    layer_filter_kernel<<<num_tiles, kFilterThreads, 0, stream>>>(num_tiles, ranges, list, (uint32_t)n_layer, list_out,
                                                                  ranges_out);
    HB_LAUNCH_OK("layer_filter", stream, false);
    return 0;
}

}  // namespace hb
