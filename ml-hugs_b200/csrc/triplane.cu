// Tri-plane feature sampling, forward and backward, for sm_100a (the step right
// before the path).  Replaces the three F.grid_sample calls + cat of hugs/models/modules/triplane.py:25-39
// (bilinear, zero padding, align_corners=True) and their autograd.
//
// Layout: the planes stay logically [1,C,H,W] (the reference's parameters) but live CHANNELS-LAST in HBM,
// [H][W][C]: the C features of one texel are one contiguous run (128 B for C = 32), so a warp reads a
// corner with one coalesced request and the backward's scatter is one coalesced RED per corner.  In the
// reference's channel-major layout the same corner is C separate 32-byte sectors.
// One warp per point, lanes over channels; the 3*C features of a point leave as coalesced rows of the
// [N, 3C] output (xy | xz | yz, triplane.py:37).  The planes (3 x C x R^2 x 4 B = 25 MB at the defaults)
// stay L2-resident; HBM traffic is the compulsory plane read plus 12 B in and 12*C B out per point.
//
// Arithmetic follows ATen's grid_sampler_2d (unnormalise ((g+1)/2)*(size-1); weights from the
// floor corner; accumulation order nw, ne, sw, se) and the module's own normalisation
// ((x - center)/scale + 0.5)*2 - 1 with one rounding per operation.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kTriThreads = 256;

struct TriArgs {
    int N, C;
    int H[3], W[3];           // per plane: xy, xz, yz
    const float* plane[3];    // [H][W][C]
    float* gplane[3];         // backward: same layout, accumulated into
    const float* xyz;         // [N,3]
    float center, scale, eps;
    float* out;               // forward: [N,3C]
    const float* gout;        // backward: [N,3C]
    float* gxyz;              // backward: [N,3] or null
    int* flag;                // forward: set to 1 if a normalised coordinate leaves [-eps, 1+eps]
};

struct Corner {
    int x0, y0;
    float ix, iy;
    bool in_x0, in_x1, in_y0, in_y1;
};

__device__ __forceinline__ Corner corner_of(float u, float v, int H, int W) {
    Corner k;
    k.ix = __fmul_rn(__fdiv_rn(__fadd_rn(u, 1.f), 2.f), (float)(W - 1));
    k.iy = __fmul_rn(__fdiv_rn(__fadd_rn(v, 1.f), 2.f), (float)(H - 1));
    const float fx = floorf(k.ix), fy = floorf(k.iy);
    // clamp before the int conversion: far-outside coordinates only need "out of bounds"
    k.x0 = (int)fminf(fmaxf(fx, -2.f), (float)W);
    k.y0 = (int)fminf(fmaxf(fy, -2.f), (float)H);
    k.in_x0 = k.x0 >= 0 && k.x0 < W;
    k.in_x1 = k.x0 + 1 >= 0 && k.x0 + 1 < W;
    k.in_y0 = k.y0 >= 0 && k.y0 < H;
    k.in_y1 = k.y0 + 1 >= 0 && k.y0 + 1 < H;
    return k;
}

__device__ __forceinline__ float grid_coord(float x, float center, float scale) {
    const float t = __fadd_rn(__fdiv_rn(__fsub_rn(x, center), scale), 0.5f);
    return __fsub_rn(__fmul_rn(t, 2.f), 1.f);
}

__global__ void __launch_bounds__(kTriThreads) triplane_forward_kernel(const TriArgs a) {
    const int n = (blockIdx.x * kTriThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= a.N) return;
    float g[3];
    bool bad = false;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float x = __ldg(a.xyz + (size_t)n * 3 + d);
        const float t = __fadd_rn(__fdiv_rn(__fsub_rn(x, a.center), a.scale), 0.5f);
        bad |= !(t <= 1.f + a.eps && t >= -a.eps);
        g[d] = __fsub_rn(__fmul_rn(t, 2.f), 1.f);
    }
    if (bad && lane == 0 && a.flag) atomicOr(a.flag, 1);
    float* out = a.out + (size_t)n * 3 * a.C;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        const float u = p == 2 ? g[1] : g[0];
        const float v = p == 0 ? g[1] : g[2];
        const int H = a.H[p], W = a.W[p], C = a.C;
        const Corner k = corner_of(u, v, H, W);
        const float wx1 = k.ix - (float)k.x0, wx0 = (float)(k.x0 + 1) - k.ix;
        const float wy1 = k.iy - (float)k.y0, wy0 = (float)(k.y0 + 1) - k.iy;
        const float nw = wx0 * wy0, ne = wx1 * wy0, sw = wx0 * wy1, se = wx1 * wy1;
        const float* base = a.plane[p] + ((long long)k.y0 * W + k.x0) * C;   // only dereferenced in bounds
        const long long row = (long long)W * C;
        for (int c = lane; c < C; c += 32) {
            float acc = 0.f;
            if (k.in_y0 && k.in_x0) acc = __fmul_rn(__ldg(base + c), nw);
            if (k.in_y0 && k.in_x1) acc = __fmaf_rn(__ldg(base + C + c), ne, acc);
            if (k.in_y1 && k.in_x0) acc = __fmaf_rn(__ldg(base + row + c), sw, acc);
            if (k.in_y1 && k.in_x1) acc = __fmaf_rn(__ldg(base + row + C + c), se, acc);
            out[p * C + c] = acc;
        }
    }
}

__global__ void __launch_bounds__(kTriThreads) triplane_backward_kernel(const TriArgs a) {
    const int n = (blockIdx.x * kTriThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= a.N) return;
    float g[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) g[d] = grid_coord(__ldg(a.xyz + (size_t)n * 3 + d), a.center, a.scale);
    const float* gout = a.gout + (size_t)n * 3 * a.C;
    float gg[3] = {0.f, 0.f, 0.f};     // d loss / d grid coordinate (x, y, z), this lane's channels
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        const int du = p == 2 ? 1 : 0, dv = p == 0 ? 1 : 2;
        const int H = a.H[p], W = a.W[p], C = a.C;
        const Corner k = corner_of(g[du], g[dv], H, W);
        const float wx1 = k.ix - (float)k.x0, wx0 = (float)(k.x0 + 1) - k.ix;
        const float wy1 = k.iy - (float)k.y0, wy0 = (float)(k.y0 + 1) - k.iy;
        const float nw = wx0 * wy0, ne = wx1 * wy0, sw = wx0 * wy1, se = wx1 * wy1;
        const long long off = ((long long)k.y0 * W + k.x0) * C;
        const long long row = (long long)W * C;
        float* gp = a.gplane[p] ? a.gplane[p] + off : nullptr;
        const float* pl = a.plane[p] + off;
        float gix = 0.f, giy = 0.f;
        for (int c = lane; c < C; c += 32) {
            const float go = __ldg(gout + p * C + c);
            if (gp) {
                if (k.in_y0 && k.in_x0) atomicAdd(gp + c, go * nw);
                if (k.in_y0 && k.in_x1) atomicAdd(gp + C + c, go * ne);
                if (k.in_y1 && k.in_x0) atomicAdd(gp + row + c, go * sw);
                if (k.in_y1 && k.in_x1) atomicAdd(gp + row + C + c, go * se);
            }
            if (a.gxyz) {
                const float v00 = (k.in_y0 && k.in_x0) ? __ldg(pl + c) : 0.f;
                const float v01 = (k.in_y0 && k.in_x1) ? __ldg(pl + C + c) : 0.f;
                const float v10 = (k.in_y1 && k.in_x0) ? __ldg(pl + row + c) : 0.f;
                const float v11 = (k.in_y1 && k.in_x1) ? __ldg(pl + row + C + c) : 0.f;
                gix += go * ((v01 - v00) * wy0 + (v11 - v10) * wy1);
                giy += go * ((v10 - v00) * wx0 + (v11 - v01) * wx1);
            }
        }
        // d ix / d grid = (W-1)/2, d iy / d grid = (H-1)/2
        gg[du] += gix * (0.5f * (float)(W - 1));
        gg[dv] += giy * (0.5f * (float)(H - 1));
    }
    if (a.gxyz) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) gg[d] += __shfl_xor_sync(0xffffffffu, gg[d], o);
        }
        // d grid / d x = 2 / scale
        if (lane < 3) a.gxyz[(size_t)n * 3 + lane] = (lane == 0 ? gg[0] : lane == 1 ? gg[1] : gg[2]) * (2.f / a.scale);
    }
}

bool fill(TriArgs& a, int N, int C, int resX, int resY, int resZ) {
    a.N = N; a.C = C;
    a.H[0] = resX; a.W[0] = resY;     // plane_xy [1,C,resX,resY], grid (x, y): x indexes the LAST axis
    a.H[1] = resX; a.W[1] = resZ;     // plane_xz [1,C,resX,resZ], grid (x, z)
    a.H[2] = resY; a.W[2] = resZ;     // plane_yz [1,C,resY,resZ], grid (y, z)
    return N >= 0 && C > 0 && resX > 0 && resY > 0 && resZ > 0 && (long long)N * 32 < (1ll << 40);
}

}  // namespace

int launch_triplane_forward(int N, int C, int resX, int resY, int resZ, const float* plane_xy, const float* plane_xz,
                            const float* plane_yz, const float* xyz, float center, float scale, float eps,
                            float* out, int* flag, cudaStream_t stream) {
    TriArgs a{};
    if (!fill(a, N, C, resX, resY, resZ)) return -1;
    if (N == 0) return 0;
    a.plane[0] = plane_xy; a.plane[1] = plane_xz; a.plane[2] = plane_yz;
    a.xyz = xyz; a.center = center; a.scale = scale; a.eps = eps; a.out = out; a.flag = flag;
    const long long threads = (long long)N * 32;
    triplane_forward_kernel<<<(unsigned)((threads + kTriThreads - 1) / kTriThreads), kTriThreads, 0, stream>>>(a);
    HB_LAUNCH_OK("triplane_forward", stream, false);
    return 0;
}

int launch_triplane_backward(int N, int C, int resX, int resY, int resZ, const float* plane_xy, const float* plane_xz,
                             const float* plane_yz, const float* xyz, float center, float scale, const float* gout,
                             float* g_xy, float* g_xz, float* g_yz, float* gxyz, cudaStream_t stream) {
    TriArgs a{};
    if (!fill(a, N, C, resX, resY, resZ)) return -1;
    if (N == 0) return 0;
    a.plane[0] = plane_xy; a.plane[1] = plane_xz; a.plane[2] = plane_yz;
    a.gplane[0] = g_xy; a.gplane[1] = g_xz; a.gplane[2] = g_yz;
    a.xyz = xyz; a.center = center; a.scale = scale; a.gout = gout; a.gxyz = gxyz;
    const long long threads = (long long)N * 32;
    triplane_backward_kernel<<<(unsigned)((threads + kTriThreads - 1) / kTriThreads), kTriThreads, 0, stream>>>(a);
    HB_LAUNCH_OK("triplane_backward", stream, false);
    return 0;
}

}  // namespace hb
