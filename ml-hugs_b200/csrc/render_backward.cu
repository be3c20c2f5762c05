// Per-tile back-to-front blend backward for sm_100a.
// Replaces renderCUDA<3> backward (cuda_rasterizer/backward.cu:399-557).
//
// The reference issues 9 global float atomicAdds per (pixel, Gaussian) hit, all 256 pixels of a
// tile hammering the same 9 addresses in lock step.  This kernel is issue-bound (not HBM-bound),
// so everything below is about executing fewer warp instructions per blended Gaussian:
//   * a warp owns an 8x4 sub-tile and walks ONLY the Gaussians it blended in the forward pass: the
//     forward blend leaves one 32-bit mask per (32 list positions, warp) (render_forward.cu), so
//     there is no cull test and no evaluation that ends in "no pixel of this warp was hit";
//     positions behind the warp's own largest contributor count are never visited
//     (backward.cu:486-488 skips them pixel by pixel);
//   * the 9 partial gradients of a Gaussian are reduced over the warp's 32 pixels with a halving
//     butterfly (12 SHFL instead of 45) and STORED -- not added -- to the warp's private slot of
//     the staged batch: each warp meets each Gaussian once per batch, so the 8 warps of a tile
//     need neither shared-memory atomics nor a zero fill; the flush sums the slots whose mask bit
//     is set and issues one set of 9 global reductions per (tile, Gaussian) -- 256x fewer global
//     atomics than the reference;
//   * dL/dmean2D and dL/dconic are accumulated in a factored form (sums of X, Y, X dx, X dy, Y dy
//     with X = dL/dG G dx, Y = dL/dG G dy) and multiplied by the conic once per (tile, Gaussian);
//   * gradients need 1e-3, not bit-exactness: exp and the reciprocal of (1 - alpha) use the
//     special-function unit (ex2.approx / rcp.approx) instead of the IEEE sequences the forward
//     pass keeps for its contributor counts.
// Batches are aligned to multiples of 128 list positions so that the forward masks line up.
// Per-hit arithmetic follows backward.cu:490-554.
#include "render_common.cuh"

namespace hb {

namespace {

constexpr int kBwdBatch = 128;                 // Gaussians staged per round
constexpr int kBwdChunks = kBwdBatch / 32;     // forward-mask words per warp and round
constexpr int kBwdWarps = kBlendThreads / 32;  // 8

// Sum 9 per-lane values over the warp and store the totals to the 9 floats at shared address `dst`.
__device__ __forceinline__ void warp_reduce9_store(const float (&v)[9], uint32_t dst, unsigned lane) {
    const bool b4 = lane & 16u, b3 = lane & 8u, b2 = lane & 4u, b1 = lane & 2u;
    float w[4], p8 = v[8];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float send = b4 ? v[i] : v[i + 4];
        const float keep = b4 ? v[i + 4] : v[i];
        w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    p8 += __shfl_xor_sync(0xffffffffu, p8, 16);
    float x[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float send = b3 ? w[i] : w[i + 2];
        const float keep = b3 ? w[i + 2] : w[i];
        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    p8 += __shfl_xor_sync(0xffffffffu, p8, 8);
    float y;
    {
        const float send = b2 ? x[0] : x[1];
        const float keep = b2 ? x[1] : x[0];
        y = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    p8 += __shfl_xor_sync(0xffffffffu, p8, 4);
    // y: value (lane >> 2) summed over the 8 lanes that share (lane & 3); p8: value 8, likewise.
    // Lanes with bit 1 clear finish y, lanes with bit 1 set finish p8.
    float r;
    {
        const float send = b1 ? y : p8;
        const float keep = b1 ? p8 : y;
        r = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    // lanes 0,4,..,28 hold values 0..7, lane 2 holds value 8: one predicated store
    const unsigned slot = (lane == 2u) ? 8u : (lane >> 2);
    if ((lane & 3u) == 0u || lane == 2u)
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(dst + 4u * slot), "f"(r) : "memory");
}

__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(kBlendThreads, 5)
    render_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                           int grid_x, const float* __restrict__ bg, const float2* __restrict__ means2D,
                           const float4* __restrict__ conic_opacity, const float* __restrict__ colors,
                           int color_stride, const float* __restrict__ final_T,
                           const uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ hit_words,
                           const float* __restrict__ dL_dpix, float* __restrict__ dL_dmean2D,
                           float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
                           float* __restrict__ dL_dcolor) {
    __shared__ uint32_t s_id[kBwdBatch];
    __shared__ __align__(16) unsigned char s_stage[kBwdBatch * kStageBytes];
    __shared__ float s_part[kBwdWarps][kBwdBatch][9];
    __shared__ uint32_t s_hit[kBwdChunks][kBwdWarps];    // forward masks of this round
    __shared__ uint32_t s_touch[kBwdWarps][kBwdChunks];  // what each warp actually wrote to s_part
    __shared__ uint32_t s_max[kBwdWarps];
    const uint32_t s_base = (uint32_t)__cvta_generic_to_shared(s_stage);

    const unsigned tile_x = blockIdx.x, tile_y = blockIdx.y;
    const unsigned tile_id = tile_y * grid_x + tile_x;
    const PixelMap pm = pixel_map(tile_x, tile_y);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t part_base = (uint32_t)__cvta_generic_to_shared(&s_part[warp][0][0]);
    const bool inside = pm.px < (unsigned)W && pm.py < (unsigned)H;
    const float pixfx = (float)pm.px, pixfy = (float)pm.py;
    const size_t pix_id = (size_t)W * pm.py + pm.px;
    const size_t plane = (size_t)W * H;

    const uint2 range = __ldg(ranges + tile_id);

    // In the forward, we stored the final value for T, the product of all (1 - alpha) factors,
    // and the (1-based) list position of the last contributing Gaussian of each pixel.
    const float T_final = inside ? __ldg(final_T + pix_id) : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? __ldg(n_contrib + pix_id) : 0u;

    // Largest contributor count of this warp / of the tile: nothing behind it was blended.
    const uint32_t warp_max = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) s_max[warp] = warp_max;
    __syncthreads();
    uint32_t max_contrib = 0;
#pragma unroll
    for (int w = 0; w < kBwdWarps; w++) max_contrib = max(max_contrib, s_max[w]);
    const int todo = (int)min(max_contrib, range.y - range.x);
    if (todo == 0) return;

    float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f;
    if (inside) {
        dLp0 = __ldg(dL_dpix + pix_id);
        dLp1 = __ldg(dL_dpix + plane + pix_id);
        dLp2 = __ldg(dL_dpix + 2 * plane + pix_id);
    }
    // alpha also influences how much of the background colour is added (backward.cu:531-534)
    const float bg_term = -T_final * (__ldg(bg + 0) * dLp0 + __ldg(bg + 1) * dLp1 + __ldg(bg + 2) * dLp2);

    float accum0 = 0.f, accum1 = 0.f, accum2 = 0.f;
    float last_alpha = 0.f, last_c0 = 0.f, last_c1 = 0.f, last_c2 = 0.f;

    // Gradient of pixel coordinate w.r.t. normalized screen-space viewport coordinates (-1 to 1).
    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;
    const uint32_t* const hit_base = hit_words + ((size_t)(range.x >> 5) + tile_id) * kBwdWarps;

    // The list entries of the next round are fetched while this one is processed (first of the two
    // dependent round trips of the gather).
    const int pb0 = ((todo - 1) / kBwdBatch) * kBwdBatch;
    uint32_t next_id = 0;
    if ((int)threadIdx.x < todo - pb0) next_id = __ldg(point_list + range.x + (uint32_t)(pb0 + (int)threadIdx.x));

    for (int pb = pb0; pb >= 0; pb -= kBwdBatch) {
        __syncthreads();  // the previous round's flush has read s_part / s_touch / s_id / s_stage
        const int nb = min(kBwdBatch, todo - pb);
        // Stage list positions [pb, pb + nb): slot j holds position pb + j.
        if ((int)threadIdx.x < nb) {
            s_id[threadIdx.x] = next_id;
            store_gaussian(s_base + threadIdx.x * kStageBytes,
                           load_gaussian(next_id, means2D, conic_opacity, colors, color_stride));
        } else if (threadIdx.x >= kBwdBatch && threadIdx.x < kBwdBatch + kBwdChunks * kBwdWarps) {
            const int q = ((int)threadIdx.x - kBwdBatch) / kBwdWarps, w = ((int)threadIdx.x - kBwdBatch) % kBwdWarps;
            uint32_t m = 0;
            if (pb + 32 * q < todo) m = __ldg(hit_base + (size_t)((pb >> 5) + q) * kBwdWarps + w);
            s_hit[q][w] = m;
        }
        if (pb >= kBwdBatch && threadIdx.x < kBwdBatch)  // the round after this one is always a full batch
            next_id = __ldg(point_list + range.x + (uint32_t)(pb - kBwdBatch + (int)threadIdx.x));
        __syncthreads();

        // This warp's masks, clipped to positions below its own largest contributor count (words
        // beyond it were never written by the forward pass).
        uint32_t mq = 0;
        if (lane < (unsigned)kBwdChunks) {
            const int p0 = pb + 32 * (int)lane;
            if (p0 < (int)warp_max) {
                mq = s_hit[lane][warp];
                const int rem = (int)warp_max - p0;
                if (rem < 32) mq &= (1u << rem) - 1u;
            }
            s_touch[warp][lane] = mq;
        }
#pragma unroll 1
        for (int q = kBwdChunks - 1; q >= 0; q--) {
            unsigned m = __shfl_sync(0xffffffffu, mq, q);
            while (m) {
                const int bit = 31 - __clz(m);
                m ^= 1u << bit;
                const int e = 32 * q + bit;
                const uint32_t addr = s_base + (uint32_t)e * kStageBytes;
                const float4 g0 = lds_f4(addr);       // x, y, A, B
                const float4 g1 = lds_f4(addr + 16);  // C, opacity, r, g
                const float cb = lds_f1(addr + 32);   // b
                // Straight-line: nearly every lane is hit by a Gaussian its warp blended (measured: 28
                // of 32), so the state updates are predicated instead of branched around.
                const float dx = g0.x - pixfx;
                const float dy = g0.y - pixfy;
                const float power = -0.5f * (g0.z * dx * dx + g1.x * dy * dy) - g0.w * dx * dy;
                // clamp: a lane outside the footprint must not make inf.  ex2 / rcp are the bare
                // special-function instructions (flush-to-zero forms: no denormal fix-up code; 1 - alpha
                // >= 0.01 and a G that underflows belongs to a lane that is not hit).
                const float G = ex2_ftz(fminf(power, 0.f) * 1.4426950408889634f);
                const float alpha = fminf(0.99f, g1.y * G);
                // Skip if this one is behind the last contributor of this pixel, or was skipped by it
                // in the forward pass (backward.cu:486-501).
                const bool hit = ((uint32_t)(pb + e) < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                const float rcp = rcp_ftz(1.f - alpha);
                if (hit) {
                    T *= rcp;
                    accum0 = fmaf(last_alpha, last_c0 - accum0, accum0);
                    accum1 = fmaf(last_alpha, last_c1 - accum1, accum1);
                    accum2 = fmaf(last_alpha, last_c2 - accum2, accum2);
                    last_c0 = g1.z;
                    last_c1 = g1.w;
                    last_c2 = cb;
                    last_alpha = alpha;
                }
                // Propagate gradients to per-Gaussian colours and keep gradients w.r.t. alpha
                // (blending factor for a Gaussian/pixel pair); both vanish for a lane that is not hit.
                const float dchannel_dcolor = hit ? alpha * T : 0.f;
                float dL_dalpha = (g1.z - accum0) * dLp0 + (g1.w - accum1) * dLp1 + (cb - accum2) * dLp2;
                dL_dalpha = hit ? fmaf(dL_dalpha, T, bg_term * rcp) : 0.f;
                const float dL_dG = g1.y * dL_dalpha;
                const float X = dL_dG * (G * dx);
                const float Y = dL_dG * (G * dy);
                float g[9];
                g[0] = dchannel_dcolor * dLp0;
                g[1] = dchannel_dcolor * dLp1;
                g[2] = dchannel_dcolor * dLp2;
                g[3] = X;
                g[4] = Y;
                g[5] = X * dx;
                g[6] = X * dy;
                g[7] = Y * dy;
                g[8] = G * dL_dalpha;
                warp_reduce9_store(g, part_base + (uint32_t)e * 36u, lane);
            }
        }
        __syncthreads();

        // One set of global reductions per (tile, Gaussian) pair that some warp blended.
        if ((int)threadIdx.x < nb) {
            const int e = (int)threadIdx.x;
            float S[9];
#pragma unroll
            for (int v = 0; v < 9; v++) S[v] = 0.f;
            bool any = false;
#pragma unroll
            for (int w = 0; w < kBwdWarps; w++) {
                if ((s_touch[w][e >> 5] >> (e & 31)) & 1u) {
                    any = true;
#pragma unroll
                    for (int v = 0; v < 9; v++) S[v] += s_part[w][e][v];
                }
            }
            if (any) {
                const size_t id = s_id[e];
                const float4 r0 = lds_f4(s_base + (uint32_t)e * kStageBytes);       // x, y, A, B
                const float cC = lds_f1(s_base + (uint32_t)e * kStageBytes + 16);   // C
                const float4 co = make_float4(r0.z, r0.w, cC, 0.f);
                atomicAdd(dL_dcolor + id * 3 + 0, S[0]);
                atomicAdd(dL_dcolor + id * 3 + 1, S[1]);
                atomicAdd(dL_dcolor + id * 3 + 2, S[2]);
                // dG/ddelx = -G dx A - G dy B, dG/ddely = -G dy C - G dx B (backward.cu:520-521)
                atomicAdd(dL_dmean2D + id * 3 + 0, -(co.x * S[3] + co.y * S[4]) * ddelx_dx);
                atomicAdd(dL_dmean2D + id * 3 + 1, -(co.z * S[4] + co.y * S[3]) * ddely_dy);
                atomicAdd(dL_dconic + id * 4 + 0, -0.5f * S[5]);
                atomicAdd(dL_dconic + id * 4 + 1, -0.5f * S[6]);
                atomicAdd(dL_dconic + id * 4 + 3, -0.5f * S[7]);
                atomicAdd(dL_dopacity + id, S[8]);
            }
        }
    }
}
}  // namespace

int launch_render_backward(const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x, int grid_y,
                           const float* bg, const float2* means2D, const float4* conic_opacity, const float* colors,
                           int color_stride, const float* final_T, const uint32_t* n_contrib,
                           const uint32_t* hit_words, const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                           cudaStream_t stream) {
    const dim3 grid((unsigned)grid_x, (unsigned)grid_y, 1);
    render_backward_kernel<<<grid, kBlendThreads, 0, stream>>>(ranges, point_list, W, H, grid_x, bg, means2D,
                                                               conic_opacity, colors, color_stride, final_T, n_contrib,
                                                               hit_words, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor);
    HB_LAUNCH_OK("render_backward", stream, false);
    return 0;
}

}  // namespace hb
