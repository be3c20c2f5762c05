// Per-tile back-to-front blend backward for sm_100a.
// Replaces renderCUDA<3> backward (cuda_rasterizer/backward.cu:399-557).
//
// The reference issues 9 global float atomicAdds per (pixel, Gaussian) hit, all 256 pixels of a
// tile hammering the same 9 addresses in lock step.  Here the 9 partial gradients
// (colour 3, mean2D 2, conic 3, opacity 1) are
//   1. reduced across the 32 pixels of a warp with shuffles (a halving butterfly: 14 SHFL
//      instead of 45),
//   2. accumulated across the 8 warps of the tile in shared memory (per-Gaussian slots of the
//      staged batch, conflict-free [slot][9] layout),
//   3. flushed once per (tile, Gaussian) pair with 9 global reductions -- 256x fewer global
//      atomics than the reference.
// Traversal starts at the tile's largest contributor count instead of the end of the tile's list
// (entries past it were never blended by any pixel, backward.cu:486-488 skips them one by one),
// and uses the same per-warp 8x4 sub-tile reject test as the forward pass.
// Per-hit arithmetic follows backward.cu:490-554.
#include "render_common.cuh"

namespace hb {

namespace {

// Sum 9 per-lane values over the warp.  On return every lane of quad q (lanes 4q..4q+3) holds the
// total of v[q] in `r8` (q = 0..7) and every lane holds the total of v[8] in `r1`.
__device__ __forceinline__ void warp_reduce9(const float (&v)[9], float& r8, float& r1) {
    const unsigned lane = threadIdx.x & 31u;
    const bool b4 = lane & 16u, b3 = lane & 8u, b2 = lane & 4u;
    float w[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float send = b4 ? v[i] : v[i + 4];
        const float keep = b4 ? v[i + 4] : v[i];
        w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float x[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float send = b3 ? w[i] : w[i + 2];
        const float keep = b3 ? w[i + 2] : w[i];
        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    {
        const float send = b2 ? x[0] : x[1];
        const float keep = b2 ? x[1] : x[0];
        r8 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    r8 += __shfl_xor_sync(0xffffffffu, r8, 2);
    r8 += __shfl_xor_sync(0xffffffffu, r8, 1);
    r1 = v[8];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) r1 += __shfl_xor_sync(0xffffffffu, r1, d);
}

__global__ void __launch_bounds__(kBlendThreads)
    render_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                           int grid_x, const float* __restrict__ bg, const float2* __restrict__ means2D,
                           const float4* __restrict__ conic_opacity, const float* __restrict__ colors,
                           int color_stride, const float* __restrict__ final_T,
                           const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                           float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic,
                           float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolor) {
    __shared__ uint32_t s_id[kBlendBatch];
    __shared__ float2 s_xy[kBlendBatch];
    __shared__ float4 s_co[kBlendBatch];
    __shared__ float s_rgb[3][kBlendBatch];
    __shared__ float s_grad[kBlendBatch][9];
    __shared__ uint32_t s_max[kBlendThreads / 32];

    const unsigned tile_x = blockIdx.x, tile_y = blockIdx.y;
    const PixelMap pm = pixel_map(tile_x, tile_y);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool inside = pm.px < (unsigned)W && pm.py < (unsigned)H;
    const float pixfx = (float)pm.px, pixfy = (float)pm.py;
    const size_t pix_id = (size_t)W * pm.py + pm.px;
    const size_t plane = (size_t)W * H;

    const uint2 range = __ldg(ranges + tile_y * grid_x + tile_x);

    // In the forward, we stored the final value for T, the product of all (1 - alpha) factors,
    // and the (1-based) list position of the last contributing Gaussian of each pixel.
    const float T_final = inside ? __ldg(final_T + pix_id) : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? __ldg(n_contrib + pix_id) : 0u;

    // Largest contributor count of the tile: nothing behind it was blended by any pixel.
    {
        const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
        if (lane == 0) s_max[warp] = wmax;
    }
    __syncthreads();
    uint32_t max_contrib = 0;
#pragma unroll
    for (int w = 0; w < kBlendThreads / 32; w++) max_contrib = max(max_contrib, s_max[w]);
    const int todo = (int)min(max_contrib, range.y - range.x);
    if (todo == 0) return;

    float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f;
    if (inside) {
        dLp0 = __ldg(dL_dpix + pix_id);
        dLp1 = __ldg(dL_dpix + plane + pix_id);
        dLp2 = __ldg(dL_dpix + 2 * plane + pix_id);
    }
    const float bg_dot_dpixel = __ldg(bg + 0) * dLp0 + __ldg(bg + 1) * dLp1 + __ldg(bg + 2) * dLp2;

    float accum0 = 0.f, accum1 = 0.f, accum2 = 0.f;
    float last_alpha = 0.f, last_c0 = 0.f, last_c1 = 0.f, last_c2 = 0.f;

    // Gradient of pixel coordinate w.r.t. normalized screen-space viewport coordinates (-1 to 1).
    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;

    for (int base = 0; base < todo; base += kBlendBatch) {
        __syncthreads();
        // Stage back to front: slot j holds list position todo-1-(base+j).
        const int mine = base + (int)threadIdx.x;
        if (mine < todo) {
            const uint32_t id = __ldg(point_list + range.x + (uint32_t)(todo - 1 - mine));
            s_id[threadIdx.x] = id;
            s_xy[threadIdx.x] = __ldg(means2D + id);
            s_co[threadIdx.x] = __ldg(conic_opacity + id);
            if (color_stride == 4) {
                const float4 c = __ldg(reinterpret_cast<const float4*>(colors) + id);
                s_rgb[0][threadIdx.x] = c.x;
                s_rgb[1][threadIdx.x] = c.y;
                s_rgb[2][threadIdx.x] = c.z;
            } else {
                s_rgb[0][threadIdx.x] = __ldg(colors + (size_t)id * 3 + 0);
                s_rgb[1][threadIdx.x] = __ldg(colors + (size_t)id * 3 + 1);
                s_rgb[2][threadIdx.x] = __ldg(colors + (size_t)id * 3 + 2);
            }
        }
#pragma unroll
        for (int v = 0; v < 9; v++) s_grad[threadIdx.x][v] = 0.f;
        __syncthreads();

        const int nb = min(kBlendBatch, todo - base);
        for (int c = 0; c < nb; c += 32) {
            const int jj = c + (int)lane;
            bool keep = false;
            if (jj < nb) {
                const float2 g = s_xy[jj];
                const float4 co = s_co[jj];
                keep = gaussian_may_touch(g.x, g.y, co.x, co.y, co.z, co.w, pm.x0, pm.x1, pm.y0, pm.y1);
            }
            unsigned m = __ballot_sync(0xffffffffu, keep);
            while (m) {
                const int e = c + __ffs(m) - 1;
                m &= m - 1;
                const uint32_t pos = (uint32_t)(todo - 1 - (base + e));  // 0-based list position
                float g[9];
#pragma unroll
                for (int v = 0; v < 9; v++) g[v] = 0.f;
                bool hit = false;
                // Skip if this one is behind the last contributor of this pixel.
                if (pos < last_contributor) {
                    const float2 xy = s_xy[e];
                    const float4 co = s_co[e];
                    const float dx = __fadd_rn(xy.x, -pixfx);
                    const float dy = __fadd_rn(xy.y, -pixfy);
                    const float a = __fmaf_rn(dx, __fmul_rn(dx, co.x), __fmul_rn(dy, __fmul_rn(dy, co.z)));
                    const float power = __fmaf_rn(a, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, co.y)));
                    if (!(power > 0.0f)) {
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, __fmul_rn(co.w, G));
                        if (!(alpha < 1.0f / 255.0f)) {
                            hit = true;
                            T = T / (1.f - alpha);
                            const float dchannel_dcolor = alpha * T;
                            // Propagate gradients to per-Gaussian colours and keep gradients w.r.t.
                            // alpha (blending factor for a Gaussian/pixel pair).
                            const float c0 = s_rgb[0][e], c1 = s_rgb[1][e], c2 = s_rgb[2][e];
                            accum0 = last_alpha * last_c0 + (1.f - last_alpha) * accum0;
                            accum1 = last_alpha * last_c1 + (1.f - last_alpha) * accum1;
                            accum2 = last_alpha * last_c2 + (1.f - last_alpha) * accum2;
                            last_c0 = c0;
                            last_c1 = c1;
                            last_c2 = c2;
                            float dL_dalpha = (c0 - accum0) * dLp0 + (c1 - accum1) * dLp1 + (c2 - accum2) * dLp2;
                            g[0] = dchannel_dcolor * dLp0;
                            g[1] = dchannel_dcolor * dLp1;
                            g[2] = dchannel_dcolor * dLp2;
                            dL_dalpha *= T;
                            last_alpha = alpha;
                            // alpha also influences how much of the background colour is added.
                            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;

                            const float dL_dG = co.w * dL_dalpha;
                            const float gdx = G * dx;
                            const float gdy = G * dy;
                            const float dG_ddelx = -gdx * co.x - gdy * co.y;
                            const float dG_ddely = -gdy * co.z - gdx * co.y;
                            g[3] = dL_dG * dG_ddelx * ddelx_dx;
                            g[4] = dL_dG * dG_ddely * ddely_dy;
                            g[5] = -0.5f * gdx * dx * dL_dG;
                            g[6] = -0.5f * gdx * dy * dL_dG;
                            g[7] = -0.5f * gdy * dy * dL_dG;
                            g[8] = G * dL_dalpha;
                        }
                    }
                }
                if (__any_sync(0xffffffffu, hit)) {
                    float r8, r1;
                    warp_reduce9(g, r8, r1);
                    if ((lane & 3u) == 0u) atomicAdd(&s_grad[e][lane >> 2], r8);
                    if (lane == 1u) atomicAdd(&s_grad[e][8], r1);
                }
            }
        }
        __syncthreads();

        // One set of global reductions per (tile, Gaussian) pair.
        if ((int)threadIdx.x < nb) {
            float g[9];
            bool any = false;
#pragma unroll
            for (int v = 0; v < 9; v++) {
                g[v] = s_grad[threadIdx.x][v];
                any |= (g[v] != 0.f);
            }
            if (any) {
                const size_t id = s_id[threadIdx.x];
                atomicAdd(dL_dcolor + id * 3 + 0, g[0]);
                atomicAdd(dL_dcolor + id * 3 + 1, g[1]);
                atomicAdd(dL_dcolor + id * 3 + 2, g[2]);
                atomicAdd(dL_dmean2D + id * 3 + 0, g[3]);
                atomicAdd(dL_dmean2D + id * 3 + 1, g[4]);
                atomicAdd(dL_dconic + id * 4 + 0, g[5]);
                atomicAdd(dL_dconic + id * 4 + 1, g[6]);
                atomicAdd(dL_dconic + id * 4 + 3, g[7]);
                atomicAdd(dL_dopacity + id, g[8]);
            }
        }
    }
}
}  // namespace

int launch_render_backward(const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x, int grid_y,
                           const float* bg, const float2* means2D, const float4* conic_opacity, const float* colors,
                           int color_stride, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                           float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                           cudaStream_t stream) {
    const dim3 grid((unsigned)grid_x, (unsigned)grid_y, 1);
    render_backward_kernel<<<grid, kBlendThreads, 0, stream>>>(ranges, point_list, W, H, grid_x, bg, means2D,
                                                               conic_opacity, colors, color_stride, final_T, n_contrib,
                                                               dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor);
    HB_LAUNCH_OK("render_backward", stream, false);
    return 0;
}

}  // namespace hb
