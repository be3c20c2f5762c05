// Pieces shared by the forward and backward per-tile blend kernels.
#pragma once
#include "common.cuh"

namespace hb {

constexpr int kBlendThreads = 256;  // one thread per pixel of a 16x16 tile
constexpr int kBlendBatch = 256;    // Gaussians staged in shared memory per round

// Pixel <-> thread mapping.  The reference uses a (16,16) block, i.e. a warp covers a 16x2 strip
// (cuda_rasterizer/forward.cu:276-282).  Here a warp covers an 8x4 pixel sub-tile instead: the
// footprint is more compact, so the per-warp cull below rejects more Gaussians, and the 32 B row
// segments a warp writes are still sector aligned.  Tile geometry (16x16) is unchanged, so keys,
// ranges and per-pixel results are unaffected.
struct PixelMap {
    unsigned px, py;    // pixel coordinates
    float x0, x1, y0, y1;  // pixel-centre bounds of this warp's 8x4 sub-tile
};
__device__ __forceinline__ PixelMap pixel_map(unsigned tile_x, unsigned tile_y) {
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const unsigned sx = tile_x * HB_TILE_X + (warp & 1u) * 8u;
    const unsigned sy = tile_y * HB_TILE_Y + (warp >> 1) * 4u;
    PixelMap m;
    m.px = sx + (lane & 7u);
    m.py = sy + (lane >> 3);
    m.x0 = (float)sx;
    m.x1 = (float)(sx + 7u);
    m.y0 = (float)sy;
    m.y1 = (float)(sy + 3u);
    return m;
}

// Conservative reject test: returns false only when NO pixel centre inside the rectangle
// [x0,x1]x[y0,y1] can reach alpha >= 1/255 for this Gaussian, i.e. when the reference's
// per-pixel test `alpha < 1/255 -> continue` (forward.cu:343-345, backward.cu:499-501) would
// skip it for every pixel of the sub-tile.  Skipping such a Gaussian for the whole warp is
// therefore invisible in the results (colour, T, contributor counts).
//
// alpha = min(0.99, o*exp(-q/2)), q(u,v) = A u^2 + 2B uv + C v^2 over pixel offsets (u,v).  The
// minimum of the convex form q over the box lies on an edge facing the centre; each facing edge
// is a clamped 1-D quadratic.  Reject iff q_min > 2 ln(255 o) + margin, where the margin
// dominates every rounding error of this test and of the exact per-pixel evaluation.
// Anything non-finite or not positive definite is kept (evaluated exactly).
__device__ __forceinline__ bool gaussian_may_touch(float gx, float gy, float A, float B, float C, float opacity,
                                                   float x0, float x1, float y0, float y1) {
    const float u0 = x0 - gx, u1 = x1 - gx, v0 = y0 - gy, v1 = y1 - gy;
    const bool in_x = (u0 <= 0.f) && (u1 >= 0.f);
    const bool in_y = (v0 <= 0.f) && (v1 >= 0.f);
    if (in_x && in_y) return true;
    if (!(A > 0.f && C > 0.f && A * C > B * B)) return true;
    float q = 3.0e38f;
    if (!in_x) {
        const float ue = (u0 > 0.f) ? u0 : u1;
        const float v = fminf(fmaxf(__fdividef(-B * ue, C), v0), v1);
        q = A * ue * ue + 2.f * B * ue * v + C * v * v;
    }
    if (!in_y) {
        const float ve = (v0 > 0.f) ? v0 : v1;
        const float u = fminf(fmaxf(__fdividef(-B * ve, A), u0), u1);
        q = fminf(q, A * u * u + 2.f * B * u * ve + C * ve * ve);
    }
    const float U = fmaxf(fabsf(u0), fabsf(u1)), V = fmaxf(fabsf(v0), fabsf(v1));
    const float margin = 0.01f + 2e-5f * (A * U * U + C * V * V);
    const float thr = 2.f * __logf(255.f * opacity);
    return !(q > thr + margin);
}

}  // namespace hb
