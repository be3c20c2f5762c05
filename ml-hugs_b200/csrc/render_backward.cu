// Per-tile back-to-front blend backward for sm_100a.
// Replaces renderCUDA<3> backward (cuda_rasterizer/backward.cu:399-557).
//
// The reference issues 9 global float atomicAdds per (pixel, Gaussian) hit, all 256 pixels of a
// tile hammering the same 9 addresses in lock step.  This stage is bound by instruction issue, not
// by HBM, so the design is about warp instructions per blended (warp, Gaussian):
//
//   * a warp owns an 8x4 sub-tile and works on its own, with no CTA barrier anywhere: it walks ONLY
//     the Gaussians it blended in the forward pass (render_forward.cu leaves one 32-bit mask per
//     32 list positions and warp), back to front, starting at its own largest contributor count;
//   * the hits are taken 16 at a time and every batch runs in two phases with two thread layouts:
//       phase 1, lane = pixel.  The sequential part of backward.cu:486-535 (transmittance,
//         colour-behind recurrence, dL/dalpha).  Per (pixel, Gaussian) it leaves just two numbers
//         in shared memory (one 8-byte store): w = alpha T (the weight of the pixel's dL/dpixel in
//         dL/dcolour) and Z = G dL/dalpha (every geometric gradient is o Z times a polynomial in the
//         pixel offset).  A lane that is not hit runs the same instructions with alpha = 0 and G = 0,
//         which leaves its T and colour-behind unchanged and its two outputs zero without a select
//         each.  (Evaluating the exponent as a quadratic in the pixel's position inside the sub-tile,
//         five FMAs, was tried: its rounding error of ~1e-5 widens the band in which a pixel at
//         alpha = 1/255 is classified differently from the forward pass a hundredfold, and gradients
//         at full size left the 1e-3 bar.  The exponent is the forward pass's formula.)
//       phase 2, lane = Gaussian (half-warp h takes the pixels 16h..16h+15).  Each lane runs down
//         its Gaussian's row of (w, Z) and accumulates in registers -- no cross-lane reduction per
//         hit (round 1's kernel spent 43 of its 111 instructions per (warp, Gaussian) in a 9-value
//         shuffle butterfly).  The geometric sums are taken as RAW moments of Z over the integer
//         pixel position (sum Z, sum Z i, sum Z i^2 per row pair: one add and two FMAs with immediate
//         operands per pixel) and shifted to the Gaussian's centre once per lane and batch; one
//         shuffle per sum joins the two halves;
//   * a (warp, Gaussian) ends in two 16-byte vector reductions into a packed 32-byte gradient record
//     (colour 3, mean2D 2, conic 3) plus one scalar reduction for the opacity; the per-Gaussian
//     backward kernel reads the record (one sector) and writes the reference's arrays;
//   * gradients need 1e-3, not bit-exactness: exp and the reciprocal of (1 - alpha) are the bare
//     special-function instructions, and the conic is pre-scaled by -log2(e)/2 at staging so that
//     the exponent costs three multiplies and two FMAs.  Deviation from the reference: a pixel whose
//     alpha lies within rounding of 1/255 can be classified differently from the forward pass (which
//     uses the IEEE sequence); the pixel's earlier Gaussians then see T off by 1/255.  Measured:
//     below the 1e-3 bar at every size tested (tests/test_parity_gpu.py, element-wise metric).
// Per-hit arithmetic follows backward.cu:490-554.
#include "render_common.cuh"

namespace hb {

namespace {

constexpr int kBwdG = 16;                       // Gaussians per batch
constexpr int kBwdWarps = kBlendThreads / 32;  // 8

constexpr uint32_t kRecBytes = 48;   // (x, y, A', B') (C', opacity, r, g) (b, list position, Gaussian index, -):
                                     // 12-word stride, so the 16 lanes that store / read one record each do not collide
constexpr uint32_t kRowBytes = 272;  // 32 x (w, Z) + 16: rows 0..7 start 4 banks apart, so the 128-bit reads of
                                     // phase 2 (8 lanes = 8 rows, same column) touch every bank once

struct __align__(16) BwdWarpSmem {
    float4 rec[kBwdG][kRecBytes / 16];
    float2 wz[kBwdG][kRowBytes / 8];   // (w, Z) of (Gaussian k, pixel p) at [k][p]
    float4 pixc[32];                   // dL/dpixel of the warp's 32 pixels
    uint32_t slot[kBwdG];              // list positions of the batch being assembled, descending
};
static_assert(sizeof(BwdWarpSmem) * kBwdWarps <= 48 * 1024, "static shared memory");

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
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f2(uint32_t addr, float a, float b) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_f1(float* p, float a) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}

constexpr float kLog2e = 1.4426950408889634f;

// CLAMP: the incoming gradient is that of the CLAMPED image; channels the forward pass clamped (mask byte per
// pixel, render_forward.cu) get none -- torch.clamp's backward, without the elementwise pass.
template <bool CLAMP>
__global__ void __launch_bounds__(kBlendThreads, 4)
    render_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                           int grid_x, const float* __restrict__ bg, const float2* __restrict__ means2D,
                           const float4* __restrict__ conic_opacity, const float* __restrict__ colors,
                           int color_stride, const float* __restrict__ final_T,
                           const uint32_t* __restrict__ n_contrib, const uint32_t* __restrict__ hit_words,
                           const float* __restrict__ dL_dpix, float* __restrict__ grad_pack,
                           float* __restrict__ dL_dopacity, const uint8_t* __restrict__ clamp_mask) {
    __shared__ BwdWarpSmem s_all[kBwdWarps];

    const unsigned tile_x = blockIdx.x, tile_y = blockIdx.y;
    const unsigned tile_id = tile_y * grid_x + tile_x;
    const PixelMap pm = pixel_map(tile_x, tile_y);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool inside = pm.px < (unsigned)W && pm.py < (unsigned)H;
    const float pixfx = (float)pm.px, pixfy = (float)pm.py;
    const size_t pix_id = (size_t)W * pm.py + pm.px;
    const size_t plane = (size_t)W * H;

    // In the forward, we stored the final value for T, the product of all (1 - alpha) factors, and the
    // (1-based) list position of the pixel's last contributor.
    const uint32_t last_contributor = inside ? __ldg(n_contrib + pix_id) : 0u;
    // Nothing behind the warp's largest contributor count was blended by any of its pixels.
    const uint32_t warp_max = __reduce_max_sync(0xffffffffu, last_contributor);
    if (warp_max == 0u) return;  // the whole warp leaves; no CTA barrier follows

    BwdWarpSmem& s = s_all[warp];
    const uint32_t s_rec = (uint32_t)__cvta_generic_to_shared(&s.rec[0][0]);
    const uint32_t s_wz = (uint32_t)__cvta_generic_to_shared(&s.wz[0][0]);
    const uint32_t s_pixc = (uint32_t)__cvta_generic_to_shared(&s.pixc[0]);

    const uint2 range = __ldg(ranges + tile_id);
    const float T_final = inside ? __ldg(final_T + pix_id) : 0.f;
    float T = T_final;
    float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f;
    if (inside) {
        dLp0 = __ldg(dL_dpix + pix_id);
        dLp1 = __ldg(dL_dpix + plane + pix_id);
        dLp2 = __ldg(dL_dpix + 2 * plane + pix_id);
        if (CLAMP) {
            const unsigned m = __ldg(clamp_mask + pix_id);
            if (m & 1u) dLp0 = 0.f;
            if (m & 2u) dLp1 = 0.f;
            if (m & 4u) dLp2 = 0.f;
        }
    }
    sts_f4(s_pixc + lane * 16u, dLp0, dLp1, dLp2, 0.f);
    // alpha also decides how much of the background colour is added (backward.cu:531-534)
    const float bg_term = -T_final * (__ldg(bg + 0) * dLp0 + __ldg(bg + 1) * dLp1 + __ldg(bg + 2) * dLp2);
    // colour accumulated behind the Gaussian at hand, updated eagerly: after Gaussian j it is
    // alpha_j c_j + (1 - alpha_j) acc, which is what backward.cu:505-507 computes one step later
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f;

    // Gradient of pixel coordinate w.r.t. normalized screen-space viewport coordinates (-1 to 1).
    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;

    // phase-2 identity of this lane: Gaussian e of the batch, pixels 16 half .. 16 half + 15 (rows 2 half, 2 half + 1)
    const unsigned e = lane & 15u, half = lane >> 4;
    const uint32_t p2_row = s_wz + e * kRowBytes + half * 128u;   // this lane's 16 (w, Z) pairs
    const uint32_t p2_pix = s_pixc + half * 256u;                 // and the 16 pixels they belong to
    const float org_x = pm.x0, org_y = pm.y0 + 2.f * (float)half;

    // ---- the warp's hit masks, most distant list positions first ----------------------------------
    const uint32_t* const hit_base = hit_words + ((size_t)(range.x >> 5) + tile_id) * kBwdWarps + warp;
    const int nwords = (int)((warp_max + 31u) >> 5);
    int qb = nwords - 1;      // word index held by lane 0 of the current group of 32 words
    uint32_t mw = 0;          // lane i holds word qb - i
    unsigned nz = 0;          // lanes of the group whose word still has bits
    uint32_t m = 0;           // bits left in the current word
    int q = 0;                // index of the current word
    bool group_loaded = false;
    int fill = 0;

    for (;;) {
        // ---- assemble a batch: the next (up to) 16 set bits, descending list position -------------
        while (fill < kBwdG) {
            if (m == 0u) {
                if (nz == 0u) {
                    if (group_loaded) qb -= 32;
                    if (qb < 0) break;
                    const int myq = qb - (int)lane;
                    mw = (myq >= 0) ? __ldg(hit_base + (size_t)myq * kBwdWarps) : 0u;
                    if (myq == nwords - 1) {  // words are only valid below the warp's contributor count
                        const uint32_t rem = warp_max - 32u * (uint32_t)myq;
                        if (rem < 32u) mw &= (1u << rem) - 1u;
                    }
                    nz = __ballot_sync(0xffffffffu, mw != 0u);
                    group_loaded = true;
                    continue;
                }
                const int i = __ffs(nz) - 1;
                nz &= nz - 1u;
                m = __shfl_sync(0xffffffffu, mw, i);
                q = qb - i;
            }
            const bool mine = (m >> lane) & 1u;
            const int above = __popc((m >> lane) >> 1);  // set bits at higher list positions
            const bool take = mine && above < kBwdG - fill;
            if (take) s.slot[fill + above] = 32u * (uint32_t)q + lane;
            const unsigned taken = __ballot_sync(0xffffffffu, take);
            m &= ~taken;
            fill += __popc(taken);
        }
        if (fill == 0) break;
        const int n = fill;
        fill = 0;
        __syncwarp();

        // ---- stage the batch: lane k < n fetches Gaussian k ----------------------------------------
        if ((int)lane < n) {
            const uint32_t pos = s.slot[lane];
            const uint32_t id = __ldg(point_list + range.x + pos);
            const StagedGaussian g = load_gaussian(id, means2D, conic_opacity, colors, color_stride);
            const uint32_t a = s_rec + lane * kRecBytes;
            sts_f4(a, g.xy.x, g.xy.y, g.co.x * (-0.5f * kLog2e), g.co.y * (-kLog2e));
            sts_f4(a + 16, g.co.z * (-0.5f * kLog2e), g.co.w, g.r, g.g);
            sts_f4(a + 32, g.b, __uint_as_float(pos), __uint_as_float(id), 0.f);
        }
        __syncwarp();

        // ---- phase 1: lane = pixel --------------------------------------------------------------
#pragma unroll 2
        for (int k = 0; k < n; k++) {
            const uint32_t a = s_rec + (uint32_t)k * kRecBytes;
            const float4 g0 = lds_f4(a);       // x, y, A', B'
            const float4 g1 = lds_f4(a + 16);  // C', opacity, r, g
            const float2 g2 = lds_f2(a + 32);  // b, list position
            const float dx = g0.x - pixfx;
            const float dy = g0.y - pixfy;
            // log2(e) * power, power = -0.5 (A dx^2 + C dy^2) - B dx dy
            const float p2 = fmaf(g0.w * dx, dy, fmaf(g1.x * dy, dy, (g0.z * dx) * dx));
            // clamp: a lane outside the footprint must not make inf (1 - alpha >= 0.01; a G that
            // underflows belongs to a lane that is not hit)
            const float G = ex2_ftz(fminf(p2, 0.f));
            const float oG = g1.y * G;
            // Skip if this one is behind the last contributor of this pixel, or was skipped by it in
            // the forward pass (backward.cu:486-501: power > 0, alpha < 1/255).
            const bool hit = (__float_as_uint(g2.y) < last_contributor) && !(p2 > 0.0f) && !(oG < 1.0f / 255.0f);
            // a lane that is not hit runs the same instructions with alpha = 0 and G = 0: its T, its colour
            // behind and its two outputs come out unchanged / zero without a select each
            const float alpha = hit ? fminf(0.99f, oG) : 0.f;
            const float Gh = hit ? G : 0.f;
            const float rcp = rcp_ftz(1.f - alpha);
            if (hit) T *= rcp;
            const float d0 = g1.z - acc0, d1 = g1.w - acc1, d2 = g2.x - acc2;
            float dL_dalpha = fmaf(d2, dLp2, fmaf(d1, dLp1, d0 * dLp0));
            dL_dalpha = fmaf(dL_dalpha, T, bg_term * rcp);
            acc0 = fmaf(alpha, d0, acc0);
            acc1 = fmaf(alpha, d1, acc1);
            acc2 = fmaf(alpha, d2, acc2);
            sts_f2(s_wz + (uint32_t)k * kRowBytes + lane * 8u, alpha * T, Gh * dL_dalpha);
        }
        __syncwarp();

        // ---- phase 2: lane = (Gaussian e, half of the pixels) ----------------------------------
        {
            // raw moments of Z over the integer pixel position, rows 2 half (s0, t0) and 2 half + 1 (s1, t1):
            // s = sum Z, t = sum Z i, u = sum Z i^2 (both rows)
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, s0 = 0.f, s1 = 0.f, t0 = 0.f, t1 = 0.f, u = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 q = lds_f4(p2_row + 16u * (uint32_t)j);   // (w, Z) of pixels 2j, 2j + 1
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int sidx = 2 * j + h;  // pixel 16 half + sidx sits at (sidx & 7, 2 half + (sidx >> 3))
                    const float wv = h ? q.z : q.x, zv = h ? q.w : q.y;
                    const float4 pc = lds_f4(p2_pix + 16u * (uint32_t)sidx);
                    c0 = fmaf(wv, pc.x, c0);
                    c1 = fmaf(wv, pc.y, c1);
                    c2 = fmaf(wv, pc.z, c2);
                    const int i = sidx & 7;
                    if ((sidx >> 3) == 0) {
                        s0 += zv;
                        if (i) t0 = fmaf(zv, (float)i, t0);
                    } else {
                        s1 += zv;
                        if (i) t1 = fmaf(zv, (float)i, t1);
                    }
                    if (i) u = fmaf(zv, (float)(i * i), u);
                }
            }
            // shift to the Gaussian's centre: dx = gxl - i, dy = gyl - r (r = 0, 1 within this lane's two rows)
            const float4 r0 = lds_f4(s_rec + e * kRecBytes);       // x, y, A', B'
            const float gxl = r0.x - org_x, gyl = r0.y - org_y;
            const float m00 = s0 + s1, m10 = t0 + t1;   // m01 = m02 = s1, m11 = t1, m20 = u
            float M0 = m00;
            float Sx = fmaf(gxl, m00, -m10);
            float Sy = fmaf(gyl, m00, -s1);
            float Sxx = fmaf(gxl, fmaf(gxl, m00, -2.f * m10), u);
            float Syy = fmaf(gyl, fmaf(gyl, m00, -2.f * s1), s1);
            float Sxy = fmaf(gxl, Sy, fmaf(-gyl, m10, t1));
            c0 += __shfl_xor_sync(0xffffffffu, c0, 16);
            c1 += __shfl_xor_sync(0xffffffffu, c1, 16);
            c2 += __shfl_xor_sync(0xffffffffu, c2, 16);
            M0 += __shfl_xor_sync(0xffffffffu, M0, 16);
            Sx += __shfl_xor_sync(0xffffffffu, Sx, 16);
            Sy += __shfl_xor_sync(0xffffffffu, Sy, 16);
            Sxx += __shfl_xor_sync(0xffffffffu, Sxx, 16);
            Sxy += __shfl_xor_sync(0xffffffffu, Sxy, 16);
            Syy += __shfl_xor_sync(0xffffffffu, Syy, 16);
            if ((int)lane < n) {
                const float2 r1 = lds_f2(s_rec + e * kRecBytes + 16);        // C', opacity
                const float cA = r0.z * (-2.f / kLog2e), cB = r0.w * (-1.f / kLog2e), cC = r1.x * (-2.f / kLog2e);
                const float o = r1.y;
                // X = dL/dG G dx = o Z dx, Y likewise; dG/ddelx = -G dx A - G dy B,
                // dG/ddely = -G dy C - G dx B (backward.cu:520-521)
                const float X = o * Sx, Y = o * Sy;
                const size_t id = __float_as_uint(lds_f1(s_rec + e * kRecBytes + 40));
                float* rec = grad_pack + id * 8;
                red_add_v4(rec, c0, c1, c2, -(cA * X + cB * Y) * ddelx_dx);
                red_add_v4(rec + 4, -(cC * Y + cB * X) * ddely_dy, -0.5f * o * Sxx, -0.5f * o * Sxy, -0.5f * o * Syy);
                red_add_f1(dL_dopacity + id, M0);
            }
        }
        // the next batch's staging is separated from these reads by the barrier after its assembly
    }
}
}  // namespace

int launch_render_backward(const uint2* ranges, const uint32_t* point_list, int W, int H, int grid_x, int grid_y,
                           const float* bg, const float2* means2D, const float4* conic_opacity, const float* colors,
                           int color_stride, const float* final_T, const uint32_t* n_contrib,
                           const uint32_t* hit_words, const float* dL_dpix, float* grad_pack, float* dL_dopacity,
                           cudaStream_t stream, const uint8_t* clamp_mask) {
    const dim3 grid((unsigned)grid_x, (unsigned)grid_y, 1);
    if (clamp_mask != nullptr)
        launch_blend(render_backward_kernel<true>, grid, stream, ranges, point_list, W, H, grid_x, bg, means2D,
                     conic_opacity, colors, color_stride, final_T, n_contrib, hit_words, dL_dpix, grad_pack, dL_dopacity,
                     clamp_mask);
    else
        launch_blend(render_backward_kernel<false>, grid, stream, ranges, point_list, W, H, grid_x, bg, means2D,
                     conic_opacity, colors, color_stride, final_T, n_contrib, hit_words, dL_dpix, grad_pack, dL_dopacity,
                     (const uint8_t*)nullptr);
    HB_LAUNCH_OK("render_backward", stream, false);
    return 0;
}

}  // namespace hb
