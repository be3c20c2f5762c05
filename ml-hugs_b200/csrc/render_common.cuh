// Pieces shared by the forward and backward per-tile blend kernels.
#pragma once
#include "common.cuh"

namespace hb {

constexpr int kBlendThreads = 256;  // one thread per pixel of a 16x16 tile

// Launch of a blend kernel.  Both blend kernels fill every register of every SM for hundreds of microseconds; a
// caller that runs several views on several streams can ask for them to be launched at a LOWER priority than its
// streams' (hugs_b200_set_blend_priority), so that the short latency-bound kernels of the other views (binning,
// scans, per-Gaussian passes) get the CTA slots the blend kernel frees instead of queueing behind its grid.
bool blend_priority(int* priority);   // api.cu
template <typename... KArgs, typename... Args>
static inline void launch_blend(void (*kernel)(KArgs...), dim3 grid, cudaStream_t stream, Args... args) {
    int prio = 0;
    if (!blend_priority(&prio)) {
        kernel<<<grid, kBlendThreads, 0, stream>>>(args...);
        return;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kBlendThreads, 1, 1);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributePriority;
    attr[0].val.priority = prio;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);   // (the error, if any, is picked up by HB_LAUNCH_OK)
}
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

// ---- staged Gaussians --------------------------------------------------------------------------
// One 48-byte record per staged Gaussian: (x, y, A, B) (C, opacity, r, g) (b, -, -, -).  The blend
// loops read a record with three loads off ONE 32-bit shared-memory address (broadcast: every lane
// reads the same record); separate arrays cost an address computation each, and with generic
// addressing the compiler re-derives the shared window base inside the loop.
constexpr int kStageBytes = 48;

struct StagedGaussian {  // what a staging thread holds in registers between load and store
    float2 xy;
    float4 co;
    float r, g, b;
};

__device__ __forceinline__ StagedGaussian load_gaussian(uint32_t id, const float2* __restrict__ means2D,
                                                        const float4* __restrict__ conic_opacity,
                                                        const float* __restrict__ colors, int color_stride) {
    StagedGaussian s;
    s.xy = __ldg(means2D + id);
    s.co = __ldg(conic_opacity + id);
    if (color_stride == 4) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(colors) + id);
        s.r = c.x;
        s.g = c.y;
        s.b = c.z;
    } else {
        s.r = __ldg(colors + (size_t)id * 3 + 0);
        s.g = __ldg(colors + (size_t)id * 3 + 1);
        s.b = __ldg(colors + (size_t)id * 3 + 2);
    }
    return s;
}

__device__ __forceinline__ void sts_f4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float lds_f1(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void store_gaussian(uint32_t slot_addr, const StagedGaussian& s) {
    sts_f4(slot_addr, s.xy.x, s.xy.y, s.co.x, s.co.y);
    sts_f4(slot_addr + 16, s.co.z, s.co.w, s.r, s.g);
    sts_f4(slot_addr + 32, s.b, 0.f, 0.f, 0.f);
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
