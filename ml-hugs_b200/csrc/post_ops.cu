// Renderer post-ops: the densification statistics the trainer updates right after
// every backward pass, as ONE launch without the boolean-mask indexing (and its nonzero() host sync)
// of the reference's torch code:
//   hugs/trainer/gs_trainer.py:406-411, :429-435   max_radii2D[vis] = max(max_radii2D[vis], radii[vis])
//   hugs/models/scene.py:460-462                   xyz_gradient_accum[vis] += ||grad[:n][vis, :2]||
//                                                  denom[vis] += 1
// with vis = radii > 0 (hugs/renderer/gs_renderer.py:156).  HBM-bound: 16 B read (radius, 2 gradient
// components... the third is skipped but shares the sector) and up to 24 B read-modify-write per
// visible Gaussian; invisible ones cost their 4-byte radius only.
#include "common.cuh"

namespace hb {

__global__ void __launch_bounds__(256)
densification_stats_kernel(int n, const int* __restrict__ radii, const float* __restrict__ grad, int grad_stride,
                           float* __restrict__ grad_accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int r = __ldg(radii + i);
    if (r <= 0) return;
    if (max_radii) max_radii[i] = fmaxf(max_radii[i], (float)r);
    if (grad_accum) {
        const float gx = __ldg(grad + (size_t)i * grad_stride);
        const float gy = __ldg(grad + (size_t)i * grad_stride + 1);
        grad_accum[i] += sqrtf(fmaf(gy, gy, gx * gx));
    }
    if (denom) denom[i] += 1.f;
}

int launch_densification_stats(int n, const int* radii, const float* grad, int grad_stride, float* grad_accum,
                               float* denom, float* max_radii, cudaStream_t stream) {
    if (n == 0) return 0;
    densification_stats_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, radii, grad, grad_stride, grad_accum, denom,
                                                                    max_radii);
    HB_LAUNCH_OK("densification_stats", stream, false);
    return 0;
}

}  // namespace hb
