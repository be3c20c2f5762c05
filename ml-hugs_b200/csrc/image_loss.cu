// Fused image-space losses for sm_100a: L1 + SSIM (11x11 Gaussian window, sigma 1.5) forward and
// backward.  This is the step right after the rasterizer in the reference's training loop
// (SURVEY.md 8(f) N3): hugs/losses/utils.py:52-110 (l1_loss, gaussian, create_window, ssim, _ssim),
// called from hugs/losses/loss.py:87-107.  The reference evaluates SSIM with five depthwise
// conv2d calls on [3,H,W] images plus ~15 elementwise kernels, and autograd replays them; here
//   * one forward kernel reads both images once, runs the separable 11-tap window over
//     (x, y, x^2, y^2, xy) out of shared memory, evaluates the SSIM map, accumulates sum(SSIM) and
//     sum|x-y|, and leaves three per-pixel derivative planes (dS/dmu1, dS/dE[x^2], dS/dE[xy]);
//   * one backward kernel convolves those three planes with the same window and emits
//     dL/dx = g_ssim/N * (conv(Dm) + 2 x conv(D11) + y conv(D12)) + g_l1/N * sign(x - y)
//     directly in the CHW layout the blend backward reads.
// Zero padding (F.conv2d(padding=5)): taps outside the image contribute nothing, in both directions.
// Both kernels are HBM bound: forward 8 B read + 12 B written, backward 20 B read + 4 B written per
// pixel and channel.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kWin = 11;
constexpr int kHalf = kWin / 2;
constexpr int kTile = 16;
constexpr int kExt = kTile + 2 * kHalf;  // 26
constexpr int kLossThreads = kTile * kTile;

struct Window {
    float w[kWin];
};

// gaussian(11, 1.5) as hugs/losses/utils.py:63-65 builds it: exp() in double, stored as float32,
// normalised by the float32 sum.
static Window make_window() {
    Window win;
    float sum = 0.f;
    for (int i = 0; i < kWin; i++) {
        const double d = (double)(i - kHalf);
        win.w[i] = (float)exp(-(d * d) / (2.0 * 1.5 * 1.5));
    }
    for (int i = 0; i < kWin; i++) sum += win.w[i];
    for (int i = 0; i < kWin; i++) win.w[i] = win.w[i] / sum;
    return win;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// img1 / img2: [C,H,W].  dmap: [3][C][H][W] (may be null: no backward wanted).  sums: double[2] =
// { sum |img1 - img2|, sum SSIM }, zeroed by the launcher.
__global__ void __launch_bounds__(kLossThreads)
    l1_ssim_forward_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                           float* __restrict__ dmap, double* __restrict__ sums, const Window win) {
    __shared__ float s_x[kExt][kExt + 1];
    __shared__ float s_y[kExt][kExt + 1];
    __shared__ float s_h[5][kExt][kTile];
    __shared__ float s_red[kLossThreads / 32][2];
    const int tid = threadIdx.y * kTile + threadIdx.x;
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile, c = blockIdx.z;
    const size_t plane = (size_t)H * W;
    const float* p1 = img1 + (size_t)c * plane;
    const float* p2 = img2 + (size_t)c * plane;

    for (int i = tid; i < kExt * kExt; i += kLossThreads) {
        const int ly = i / kExt, lx = i - ly * kExt;
        const int gy = y0 - kHalf + ly, gx = x0 - kHalf + lx;
        float a = 0.f, b = 0.f;
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
            a = __ldg(p1 + (size_t)gy * W + gx);
            b = __ldg(p2 + (size_t)gy * W + gx);
        }
        s_x[ly][lx] = a;
        s_y[ly][lx] = b;
    }
    __syncthreads();
    // horizontal pass: 26 rows x 16 columns x 5 quantities
    for (int i = tid; i < kExt * kTile; i += kLossThreads) {
        const int r = i / kTile, col = i - r * kTile;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f, h4 = 0.f;
#pragma unroll
        for (int k = 0; k < kWin; k++) {
            const float a = s_x[r][col + k], b = s_y[r][col + k], w = win.w[k];
            const float wa = w * a, wb = w * b;
            h0 += wa;
            h1 += wb;
            h2 = fmaf(wa, a, h2);
            h3 = fmaf(wb, b, h3);
            h4 = fmaf(wa, b, h4);
        }
        s_h[0][r][col] = h0;
        s_h[1][r][col] = h1;
        s_h[2][r][col] = h2;
        s_h[3][r][col] = h3;
        s_h[4][r][col] = h4;
    }
    __syncthreads();
    // vertical pass for this thread's pixel
    const int lx = threadIdx.x, ly = threadIdx.y;
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < kWin; k++) {
        const float w = win.w[k];
        mu1 = fmaf(w, s_h[0][ly + k][lx], mu1);
        mu2 = fmaf(w, s_h[1][ly + k][lx], mu2);
        e11 = fmaf(w, s_h[2][ly + k][lx], e11);
        e22 = fmaf(w, s_h[3][ly + k][lx], e22);
        e12 = fmaf(w, s_h[4][ly + k][lx], e12);
    }
    const int gx = x0 + lx, gy = y0 + ly;
    const bool inside = gx < W && gy < H;
    float ssim_v = 0.f, l1_v = 0.f;
    if (inside) {
        // _ssim (hugs/losses/utils.py:81-103)
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
        const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1_mu2;
        const float A1 = 2.f * mu1_mu2 + C1, A2 = 2.f * sigma12 + C2;
        const float B1 = mu1_sq + mu2_sq + C1, B2 = sigma1_sq + sigma2_sq + C2;
        const float inv = 1.f / (B1 * B2);
        ssim_v = A1 * A2 * inv;
        l1_v = fabsf(s_x[ly + kHalf][lx + kHalf] - s_y[ly + kHalf][lx + kHalf]);
        if (dmap != nullptr) {
            // S as a function of (mu1, E11 = E[x^2], E12 = E[xy]) with the img2 statistics fixed
            const float dS_dE11 = -ssim_v / B2;
            const float dS_dE12 = 2.f * A1 * inv;
            const float dS_dmu1 = 2.f * mu2 * (A2 - A1) * inv + 2.f * mu1 * ssim_v * (1.f / B2 - 1.f / B1);
            const size_t o = (size_t)c * plane + (size_t)gy * W + gx;
            const size_t stride = (size_t)C * plane;
            dmap[o] = dS_dmu1;
            dmap[stride + o] = dS_dE11;
            dmap[2 * stride + o] = dS_dE12;
        }
    }
    const float a = warp_sum(l1_v), b = warp_sum(ssim_v);
    if ((tid & 31) == 0) {
        s_red[tid >> 5][0] = a;
        s_red[tid >> 5][1] = b;
    }
    __syncthreads();
    if (tid == 0) {
        double sa = 0.0, sb = 0.0;
#pragma unroll
        for (int w = 0; w < kLossThreads / 32; w++) {
            sa += (double)s_red[w][0];
            sb += (double)s_red[w][1];
        }
        atomicAdd(sums + 0, sa);
        atomicAdd(sums + 1, sb);
    }
}

// dL/dimg1 [C,H,W] from the three derivative planes of the forward pass.  g = { dL/d(l1 mean),
// dL/d(ssim mean) } lives on the device (it is the autograd gradient of two scalars: no host sync).
__global__ void __launch_bounds__(kLossThreads)
    l1_ssim_backward_kernel(int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                            const float* __restrict__ dmap, const float* __restrict__ g,
                            float* __restrict__ dL_dimg1, const Window win) {
    __shared__ float s_d[3][kExt][kExt + 1];
    __shared__ float s_h[3][kExt][kTile];
    const int tid = threadIdx.y * kTile + threadIdx.x;
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile, c = blockIdx.z;
    const size_t plane = (size_t)H * W;
    const size_t stride = (size_t)C * plane;
    const float* d0 = dmap + (size_t)c * plane;

    for (int i = tid; i < kExt * kExt; i += kLossThreads) {
        const int ly = i / kExt, lx = i - ly * kExt;
        const int gy = y0 - kHalf + ly, gx = x0 - kHalf + lx;
        float a = 0.f, b = 0.f, d = 0.f;
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
            const size_t o = (size_t)gy * W + gx;
            a = __ldg(d0 + o);
            b = __ldg(d0 + stride + o);
            d = __ldg(d0 + 2 * stride + o);
        }
        s_d[0][ly][lx] = a;
        s_d[1][ly][lx] = b;
        s_d[2][ly][lx] = d;
    }
    __syncthreads();
    for (int i = tid; i < kExt * kTile; i += kLossThreads) {
        const int r = i / kTile, col = i - r * kTile;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll
        for (int k = 0; k < kWin; k++) {
            const float w = win.w[k];
            h0 = fmaf(w, s_d[0][r][col + k], h0);
            h1 = fmaf(w, s_d[1][r][col + k], h1);
            h2 = fmaf(w, s_d[2][r][col + k], h2);
        }
        s_h[0][r][col] = h0;
        s_h[1][r][col] = h1;
        s_h[2][r][col] = h2;
    }
    __syncthreads();
    const int lx = threadIdx.x, ly = threadIdx.y;
    const int gx = x0 + lx, gy = y0 + ly;
    if (gx >= W || gy >= H) return;
    float cm = 0.f, c11 = 0.f, c12 = 0.f;
#pragma unroll
    for (int k = 0; k < kWin; k++) {
        const float w = win.w[k];
        cm = fmaf(w, s_h[0][ly + k][lx], cm);
        c11 = fmaf(w, s_h[1][ly + k][lx], c11);
        c12 = fmaf(w, s_h[2][ly + k][lx], c12);
    }
    const size_t o = (size_t)c * plane + (size_t)gy * W + gx;
    const float x = __ldg(img1 + o), y = __ldg(img2 + o);
    const float inv_n = 1.f / ((float)C * (float)H * (float)W);
    const float diff = x - y;
    const float sgn = (diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f);  // torch.abs has gradient 0 at 0
    const float g_l1 = __ldg(g + 0), g_ssim = __ldg(g + 1);
    dL_dimg1[o] = inv_n * (g_ssim * (cm + 2.f * x * c11 + y * c12) + g_l1 * sgn);
}
}  // namespace

int launch_l1_ssim_forward(int C, int H, int W, const float* img1, const float* img2, float* dmap, double* sums,
                           cudaStream_t stream) {
    static const Window win = make_window();
    HB_CUDA_OK(cudaMemsetAsync(sums, 0, 2 * sizeof(double), stream));
    if (C <= 0 || H <= 0 || W <= 0) return 0;
    const dim3 grid((unsigned)cdiv(W, kTile), (unsigned)cdiv(H, kTile), (unsigned)C);
    const dim3 block(kTile, kTile, 1);
    l1_ssim_forward_kernel<<<grid, block, 0, stream>>>(C, H, W, img1, img2, dmap, sums, win);
    HB_LAUNCH_OK("l1_ssim_forward", stream, false);
    return 0;
}

int launch_l1_ssim_backward(int C, int H, int W, const float* img1, const float* img2, const float* dmap,
                            const float* g, float* dL_dimg1, cudaStream_t stream) {
    static const Window win = make_window();
    if (C <= 0 || H <= 0 || W <= 0) return 0;
    const dim3 grid((unsigned)cdiv(W, kTile), (unsigned)cdiv(H, kTile), (unsigned)C);
    const dim3 block(kTile, kTile, 1);
    l1_ssim_backward_kernel<<<grid, block, 0, stream>>>(C, H, W, img1, img2, dmap, g, dL_dimg1, win);
    HB_LAUNCH_OK("l1_ssim_backward", stream, false);
    return 0;
}

}  // namespace hb
