// Mean squared distance of every point to its 3 nearest neighbours, for sm_100a.
// Replaces simple-knn's distCUDA2 (submodules/simple-knn/simple_knn.cu:185-221 SimpleKNN::knn,
// kernels :65-183), which hugs/models/scene.py:181 uses to initialise the scene Gaussians' scales
// (SURVEY.md 8(f) N4).  The result is exact 3-NN (the reference prunes with bounding boxes, it does
// not approximate), so any traversal that visits every candidate gives the same three distances.
//
// Same outline as the reference -- Morton order, boxes of consecutive points with their AABBs, box
// pruning -- restructured around memory behaviour:
//   * the points are gathered ONCE into Morton order as float4 (the reference dereferences
//     points[indices[i]] for every candidate visit);
//   * the Morton argsort is this library's own radix sort (radix_sort.cu), not CUB;
//   * one CTA owns one box of 256 query points and walks the candidate boxes TOGETHER: a candidate
//     box that may improve any query of the CTA (box-to-box distance against the CTA's largest
//     current 3rd-best) is staged in shared memory with coalesced loads and every thread tests it
//     against its own bound; the reference has every thread walk all boxes and gather candidates on
//     its own.  Candidate boxes are visited outwards from the query box in Morton order, which
//     tightens the bounds early.
// dist = dx*dx + dy*dy + dz*dz is evaluated as fma(dz,dz, fma(dx,dx, rn(dy*dy))), the contraction nvcc
// applies to the reference's expression (:139).
#include "common.cuh"

#include <cfloat>

namespace hb {

namespace {

constexpr int kBox = 256;

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

// ---- 1. bounding box of the cloud: per-CTA partial min/max, finished by the Morton kernel -------
__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ pts, float* __restrict__ part) {
    __shared__ float s[6][8];
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float v = __ldg(pts + 3 * (size_t)i + c);
            mn[c] = fminf(mn[c], v);
            mx[c] = fmaxf(mx[c], v);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            s[c][warp] = mn[c];
            s[3 + c][warp] = mx[c];
        }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = s[threadIdx.x][0];
        for (int w = 1; w < 8; w++) v = (threadIdx.x < 3) ? fminf(v, s[threadIdx.x][w]) : fmaxf(v, s[threadIdx.x][w]);
        part[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

// ---- 2. Morton codes (10 bits per axis, simple_knn.cu:46-62) -------------------------------------
__global__ void __launch_bounds__(256)
    knn_morton_kernel(int P, const float* __restrict__ pts, const float* __restrict__ part, int nparts,
                      uint32_t* __restrict__ codes) {
    __shared__ float s_b[6];
    if (threadIdx.x < 6) {
        float v = part[threadIdx.x];
        for (int b = 1; b < nparts; b++) {
            const float o = part[b * 6 + threadIdx.x];
            v = (threadIdx.x < 3) ? fminf(v, o) : fmaxf(v, o);
        }
        s_b[threadIdx.x] = v;
    }
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t code = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const float ext = s_b[3 + c] - s_b[c];
        float t = (ext > 0.f) ? (__ldg(pts + 3 * (size_t)i + c) - s_b[c]) / ext : 0.f;
        t = fminf(fmaxf(t, 0.f), 1.f);
        code |= prep_morton((uint32_t)(t * 1023.f)) << c;
    }
    codes[i] = code;
}

// ---- 3. gather into Morton order + box AABBs -------------------------------------------------------
__global__ void __launch_bounds__(kBox)
    knn_gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ perm,
                      float4* __restrict__ sorted, float* __restrict__ boxes /*[nb][6]*/) {
    __shared__ float s[6][kBox / 32];
    const int i = blockIdx.x * kBox + threadIdx.x;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t src = __ldg(perm + i);
        const float x = __ldg(pts + 3 * (size_t)src), y = __ldg(pts + 3 * (size_t)src + 1), z = __ldg(pts + 3 * (size_t)src + 2);
        sorted[i] = make_float4(x, y, z, __uint_as_float(src));
        mn[0] = mx[0] = x;
        mn[1] = mx[1] = y;
        mn[2] = mx[2] = z;
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            s[c][warp] = mn[c];
            s[3 + c][warp] = mx[c];
        }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float v = s[threadIdx.x][0];
        for (int w = 1; w < kBox / 32; w++) v = (threadIdx.x < 3) ? fminf(v, s[threadIdx.x][w]) : fmaxf(v, s[threadIdx.x][w]);
        boxes[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

// ---- 4. the search ---------------------------------------------------------------------------------
__device__ __forceinline__ float axis_gap(float lo_a, float hi_a, float lo_b, float hi_b) {
    // distance between the intervals [lo_a, hi_a] and [lo_b, hi_b] (0 when they overlap)
    return fmaxf(0.f, fmaxf(lo_a - hi_b, lo_b - hi_a));
}

// distBoxPoint (simple_knn.cu:121-131)
__device__ __forceinline__ float dist_box_point(const float* __restrict__ box, float x, float y, float z) {
    float dx = 0.f, dy = 0.f, dz = 0.f;
    if (x < box[0] || x > box[3]) dx = fminf(fabsf(x - box[0]), fabsf(x - box[3]));
    if (y < box[1] || y > box[4]) dy = fminf(fabsf(y - box[1]), fabsf(y - box[4]));
    if (z < box[2] || z > box[5]) dz = fminf(fabsf(z - box[2]), fabsf(z - box[5]));
    return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(kBox)
    knn_search_kernel(int P, int nb, const float4* __restrict__ sorted, const float* __restrict__ boxes,
                      float* __restrict__ dists) {
    __shared__ float4 s_pts[kBox];
    __shared__ float s_wmax[kBox / 32];
    const int b = blockIdx.x, t = threadIdx.x;
    const int qi = b * kBox + t;
    const bool valid = qi < P;
    const float4 q = valid ? __ldg(sorted + qi) : make_float4(0.f, 0.f, 0.f, 0.f);
    float best0 = FLT_MAX, best1 = FLT_MAX, best2 = FLT_MAX;
    float qb[6];
#pragma unroll
    for (int k = 0; k < 6; k++) qb[k] = __ldg(boxes + 6 * (size_t)b + k);
    float bound = FLT_MAX;  // largest 3rd-best distance among this CTA's queries (uniform)

    for (int step = 0; step < nb; step++) {
#pragma unroll 1
        for (int side = 0; side < 2; side++) {
            if (step == 0 && side == 1) continue;
            const int c = side == 0 ? b + step : b - step;
            if (c < 0 || c >= nb) continue;
            // CTA-uniform prune: box against box
            float cb[6];
#pragma unroll
            for (int k = 0; k < 6; k++) cb[k] = __ldg(boxes + 6 * (size_t)c + k);
            const float gx = axis_gap(qb[0], qb[3], cb[0], cb[3]);
            const float gy = axis_gap(qb[1], qb[4], cb[1], cb[4]);
            const float gz = axis_gap(qb[2], qb[5], cb[2], cb[5]);
            // (slack: this bound only has to be conservative; the per-thread test below is the reference's)
            if (gx * gx + gy * gy + gz * gz > bound * 1.0001f) continue;

            __syncthreads();  // previous candidate fully consumed
            const int base = c * kBox;
            const int n = min(kBox, P - base);
            if (t < n) s_pts[t] = __ldg(sorted + base + t);
            __syncthreads();
            if (valid && !(dist_box_point(cb, q.x, q.y, q.z) > best2)) {
                const int self = (c == b) ? t : -1;
                for (int j = 0; j < n; j++) {
                    if (j == self) continue;
                    const float4 p = s_pts[j];
                    const float dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
                    float d = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
                    // updateKBest<3> (simple_knn.cu:133-147)
                    if (best0 > d) { const float tmp = best0; best0 = d; d = tmp; }
                    if (best1 > d) { const float tmp = best1; best1 = d; d = tmp; }
                    if (best2 > d) { best2 = d; }
                }
            }
            // refresh the CTA bound
            float v = valid ? best2 : 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
            if ((t & 31) == 0) s_wmax[t >> 5] = v;
            __syncthreads();
            bound = s_wmax[0];
#pragma unroll
            for (int w = 1; w < kBox / 32; w++) bound = fmaxf(bound, s_wmax[w]);
        }
    }
    if (valid) dists[__float_as_uint(q.w)] = (best0 + best1 + best2) / 3.0f;
}
}  // namespace

size_t knn_temp_bytes(int P) {
    const size_t n = (size_t)(P > 0 ? P : 0);
    const size_t nb = (n + kBox - 1) / kBox;
    size_t b = 0;
    b += align_up(sizeof(float) * 6 * 296);     // partial bounds
    b += 5 * align_up(sizeof(uint32_t) * n);    // codes, kB, vB, kC, vC (perm)
    b += align_up(sort_temp_bytes(P));
    b += align_up(sizeof(float4) * n);          // points in Morton order
    b += align_up(sizeof(float) * 6 * nb);      // box AABBs
    return b;
}

int launch_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* temp, size_t temp_bytes,
                          cudaStream_t stream) {
    if (P <= 0) return 0;
    if (temp == nullptr || temp_bytes < knn_temp_bytes(P)) {
        set_error("knn: temp storage too small (%zu < %zu)", temp_bytes, knn_temp_bytes(P));
        return -1;
    }
    const size_t n = (size_t)P;
    const int nb = cdiv(P, kBox);
    char* p = static_cast<char*>(temp);
    auto take = [&](size_t bytes) { char* r = p; p += align_up(bytes); return r; };
    float* part = reinterpret_cast<float*>(take(sizeof(float) * 6 * 296));
    uint32_t* codes = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * n));
    uint32_t* kB = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * n));
    uint32_t* vB = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * n));
    uint32_t* kC = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * n));
    uint32_t* vC = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * n));
    const size_t st_bytes = sort_temp_bytes(P);
    void* st = take(st_bytes);
    float4* sorted = reinterpret_cast<float4*>(take(sizeof(float4) * n));
    float* boxes = reinterpret_cast<float*>(take(sizeof(float) * 6 * (size_t)nb));

    const int nparts = min(296, cdiv(P, 256));
    knn_bounds_kernel<<<nparts, 256, 0, stream>>>(P, points, part);
    HB_LAUNCH_OK("knn_bounds", stream, false);
    knn_morton_kernel<<<cdiv(P, 256), 256, 0, stream>>>(P, points, part, nparts, codes);
    HB_LAUNCH_OK("knn_morton", stream, false);
    const int rc = launch_depth_argsort(codes, kB, vB, kC, vC, P, st, st_bytes, stream);
    if (rc) return rc;
    knn_gather_kernel<<<nb, kBox, 0, stream>>>(P, points, vC, sorted, boxes);
    HB_LAUNCH_OK("knn_gather", stream, false);
    knn_search_kernel<<<nb, kBox, 0, stream>>>(P, nb, sorted, boxes, mean_dist2);
    HB_LAUNCH_OK("knn_search", stream, false);
    return 0;
}

}  // namespace hb
