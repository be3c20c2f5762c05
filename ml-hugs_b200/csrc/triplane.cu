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
    // ((g + 1) / 2) * (size - 1); the halving is exact, so a multiplication gives the same bits
    k.ix = __fmul_rn(__fmul_rn(__fadd_rn(u, 1.f), 0.5f), (float)(W - 1));
    k.iy = __fmul_rn(__fmul_rn(__fadd_rn(v, 1.f), 0.5f), (float)(H - 1));
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

// ---- vector path: C a multiple of 4 with C/4 a power of two <= 32 (C = 32: the reference's size) -----
// LPP = C/4 lanes own one point, four channels each (one 16-byte load per corner and lane), so a warp
// carries 32/LPP points: the per-point coordinate arithmetic is shared by LPP lanes instead of 32 and
// every memory instruction moves 16 bytes per lane.
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void red_add4(float* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

template <int LPP>
__global__ void __launch_bounds__(kTriThreads) triplane_forward_vec_kernel(const TriArgs a) {
    const long long t = (long long)blockIdx.x * kTriThreads + threadIdx.x;
    const int n = (int)(t / LPP);
    const int ch = (int)(threadIdx.x % LPP) * 4;
    if (n >= a.N) return;
    float g[3];
    bool bad = false;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float x = __ldg(a.xyz + (size_t)n * 3 + d);
        const float tt = __fadd_rn(__fdiv_rn(__fsub_rn(x, a.center), a.scale), 0.5f);
        bad |= !(tt <= 1.f + a.eps && tt >= -a.eps);
        g[d] = __fsub_rn(__fmul_rn(tt, 2.f), 1.f);
    }
    if (bad && ch == 0 && a.flag) atomicOr(a.flag, 1);
    constexpr int C = LPP * 4;
    float* out = a.out + (size_t)n * 3 * C + ch;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        const float u = p == 2 ? g[1] : g[0];
        const float v = p == 0 ? g[1] : g[2];
        const int H = a.H[p], W = a.W[p];
        const Corner k = corner_of(u, v, H, W);
        const float wx1 = k.ix - (float)k.x0, wx0 = (float)(k.x0 + 1) - k.ix;
        const float wy1 = k.iy - (float)k.y0, wy0 = (float)(k.y0 + 1) - k.iy;
        const float nw = wx0 * wy0, ne = wx1 * wy0, sw = wx0 * wy1, se = wx1 * wy1;
        const float* base = a.plane[p] + ((long long)k.y0 * W + k.x0) * C + ch;
        const long long row = (long long)W * C;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        // issue the four corner loads together, then combine in ATen's order nw, ne, sw, se
        const float4 v00 = (k.in_y0 && k.in_x0) ? ldg4(base) : z;
        const float4 v01 = (k.in_y0 && k.in_x1) ? ldg4(base + C) : z;
        const float4 v10 = (k.in_y1 && k.in_x0) ? ldg4(base + row) : z;
        const float4 v11 = (k.in_y1 && k.in_x1) ? ldg4(base + row + C) : z;
        float4 acc;
        // an out-of-range corner is skipped by ATen; adding 0 * w gives the same bits except for the sign
        // of an all-zero sum, which the explicit +0 start removes
        acc.x = __fmaf_rn(v11.x, se, __fmaf_rn(v10.x, sw, __fmaf_rn(v01.x, ne, __fmul_rn(v00.x, nw))));
        acc.y = __fmaf_rn(v11.y, se, __fmaf_rn(v10.y, sw, __fmaf_rn(v01.y, ne, __fmul_rn(v00.y, nw))));
        acc.z = __fmaf_rn(v11.z, se, __fmaf_rn(v10.z, sw, __fmaf_rn(v01.z, ne, __fmul_rn(v00.z, nw))));
        acc.w = __fmaf_rn(v11.w, se, __fmaf_rn(v10.w, sw, __fmaf_rn(v01.w, ne, __fmul_rn(v00.w, nw))));
        *reinterpret_cast<float4*>(out + p * C) = acc;
    }
}

template <int LPP>
__global__ void __launch_bounds__(kTriThreads) triplane_backward_vec_kernel(const TriArgs a) {
    const long long t = (long long)blockIdx.x * kTriThreads + threadIdx.x;
    const int n = (int)(t / LPP);
    const int ch = (int)(threadIdx.x % LPP) * 4;
    const unsigned mask = __ballot_sync(0xffffffffu, n < a.N);      // whole LPP-lane groups are in or out
    if (n >= a.N) return;
    float g[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) g[d] = grid_coord(__ldg(a.xyz + (size_t)n * 3 + d), a.center, a.scale);
    constexpr int C = LPP * 4;
    const float* gout = a.gout + (size_t)n * 3 * C + ch;
    float gg[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        const int du = p == 2 ? 1 : 0, dv = p == 0 ? 1 : 2;
        const int H = a.H[p], W = a.W[p];
        const Corner k = corner_of(g[du], g[dv], H, W);
        const float wx1 = k.ix - (float)k.x0, wx0 = (float)(k.x0 + 1) - k.ix;
        const float wy1 = k.iy - (float)k.y0, wy0 = (float)(k.y0 + 1) - k.iy;
        const long long off = ((long long)k.y0 * W + k.x0) * C + ch;
        const long long row = (long long)W * C;
        const float4 go = ldg4(gout + p * C);
        const bool b00 = k.in_y0 && k.in_x0, b01 = k.in_y0 && k.in_x1, b10 = k.in_y1 && k.in_x0,
                   b11 = k.in_y1 && k.in_x1;
        if (a.gplane[p]) {
            float* gp = a.gplane[p] + off;
            const float nw = wx0 * wy0, ne = wx1 * wy0, sw = wx0 * wy1, se = wx1 * wy1;
            if (b00) red_add4(gp, make_float4(go.x * nw, go.y * nw, go.z * nw, go.w * nw));
            if (b01) red_add4(gp + C, make_float4(go.x * ne, go.y * ne, go.z * ne, go.w * ne));
            if (b10) red_add4(gp + row, make_float4(go.x * sw, go.y * sw, go.z * sw, go.w * sw));
            if (b11) red_add4(gp + row + C, make_float4(go.x * se, go.y * se, go.z * se, go.w * se));
        }
        if (a.gxyz) {
            const float* pl = a.plane[p] + off;
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 v00 = b00 ? ldg4(pl) : z, v01 = b01 ? ldg4(pl + C) : z;
            const float4 v10 = b10 ? ldg4(pl + row) : z, v11 = b11 ? ldg4(pl + row + C) : z;
            float gix = go.x * ((v01.x - v00.x) * wy0 + (v11.x - v10.x) * wy1);
            gix += go.y * ((v01.y - v00.y) * wy0 + (v11.y - v10.y) * wy1);
            gix += go.z * ((v01.z - v00.z) * wy0 + (v11.z - v10.z) * wy1);
            gix += go.w * ((v01.w - v00.w) * wy0 + (v11.w - v10.w) * wy1);
            float giy = go.x * ((v10.x - v00.x) * wx0 + (v11.x - v01.x) * wx1);
            giy += go.y * ((v10.y - v00.y) * wx0 + (v11.y - v01.y) * wx1);
            giy += go.z * ((v10.z - v00.z) * wx0 + (v11.z - v01.z) * wx1);
            giy += go.w * ((v10.w - v00.w) * wx0 + (v11.w - v01.w) * wx1);
            gg[du] += gix * (0.5f * (float)(W - 1));
            gg[dv] += giy * (0.5f * (float)(H - 1));
        }
    }
    if (a.gxyz) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
#pragma unroll
            for (int o = LPP / 2; o > 0; o >>= 1) gg[d] += __shfl_xor_sync(mask, gg[d], o);
        }
        if (ch == 0) {
            const float s = 2.f / a.scale;
            a.gxyz[(size_t)n * 3 + 0] = gg[0] * s;
            a.gxyz[(size_t)n * 3 + 1] = gg[1] * s;
            a.gxyz[(size_t)n * 3 + 2] = gg[2] * s;
        }
    }
}

template <int LPP>
void launch_vec(bool backward, const TriArgs& a, cudaStream_t stream) {
    const long long threads = (long long)a.N * LPP;
    const unsigned grid = (unsigned)((threads + kTriThreads - 1) / kTriThreads);
    if (backward) triplane_backward_vec_kernel<LPP><<<grid, kTriThreads, 0, stream>>>(a);
    else triplane_forward_vec_kernel<LPP><<<grid, kTriThreads, 0, stream>>>(a);
}

// true if the vector path took the launch
bool try_vec(bool backward, const TriArgs& a, cudaStream_t stream) {
    switch (a.C) {
        case 4: launch_vec<1>(backward, a, stream); return true;
        case 8: launch_vec<2>(backward, a, stream); return true;
        case 16: launch_vec<4>(backward, a, stream); return true;
        case 32: launch_vec<8>(backward, a, stream); return true;
        case 64: launch_vec<16>(backward, a, stream); return true;
        case 128: launch_vec<32>(backward, a, stream); return true;
        default: return false;
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
    if (!try_vec(false, a, stream)) {
        const long long threads = (long long)N * 32;
        triplane_forward_kernel<<<(unsigned)((threads + kTriThreads - 1) / kTriThreads), kTriThreads, 0, stream>>>(a);
    }
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
    if (!try_vec(true, a, stream)) {
        const long long threads = (long long)N * 32;
        triplane_backward_kernel<<<(unsigned)((threads + kTriThreads - 1) / kTriThreads), kTriThreads, 0, stream>>>(a);
    }
    HB_LAUNCH_OK("triplane_backward", stream, false);
    return 0;
}

}  // namespace hb
