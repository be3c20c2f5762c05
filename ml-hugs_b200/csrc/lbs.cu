// Fused SMPL linear-blend-skinning deformation of the human Gaussians for sm_100a.
//
// The reference evaluates this as a chain of ~35 ATen kernels (cuBLAS SGEMM [N x J].[J x 16],
// cat, two bmm, and the boolean-mask gather in matrix_to_quaternion), materialising a 4x4
// transform per Gaussian:
//   lbs_extra                (hugs/models/modules/lbs.py:19-73; skinning at :59-72)
//   scale / translate / R'   (hugs/models/hugs_trimlp.py:510-518)
//   matrix_to_quaternion     (hugs/utils/rotations.py:105-156, _sqrt_positive_part :94-102)
// Here one thread handles one (pose frame, Gaussian): the frame's J joint transforms sit in shared
// memory, weights are read with 128-bit loads, and xyz' / quaternion / scales are written
// directly.  Algorithmic traffic per Gaussian and frame: 12 + 4J + 36 (+12) B read,
// 12 + 16 (+12) B written (J = 24: 172 B).
//
// The quaternion is NOT normalised and its sign is not standardised -- the blended 3x3 is not
// orthonormal and the rasterizer consumes the resulting scale.
#include "common.cuh"

namespace hb {

namespace {
constexpr int kLbsThreads = 128;
constexpr int kMaxJoints = 32;

// Blend the top three rows of the J joint matrices: T[r*4+c], r<3; T3 = sum_j w_j A_j[3][3].
__device__ __forceinline__ void blend_transform(const float* __restrict__ sA, const float* w, int J, float* T,
                                                float& T33) {
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = 0.f;
    T33 = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxJoints; j++) {
        if (j < J) {
            const float wj = w[j];
            const float* Aj = sA + 16 * j;
#pragma unroll
            for (int k = 0; k < 12; k++) T[k] = fmaf(wj, Aj[k], T[k]);
            T33 = fmaf(wj, Aj[15], T33);
        }
    }
}

__device__ __forceinline__ void load_weights(const float* __restrict__ weights, int i, int J, bool vec4, float* w) {
    if (vec4) {
        const float4* src = reinterpret_cast<const float4*>(weights + (size_t)i * J);
#pragma unroll
        for (int k = 0; k < kMaxJoints / 4; k++) {
            if (4 * k < J) {
                const float4 v = __ldg(src + k);
                w[4 * k + 0] = v.x;
                w[4 * k + 1] = v.y;
                w[4 * k + 2] = v.z;
                w[4 * k + 3] = v.w;
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < kMaxJoints; k++)
            if (k < J) w[k] = __ldg(weights + (size_t)i * J + k);
    }
}

// matrix_to_quaternion (rotations.py:105-156) of m (row-major 3x3).  Returns the chosen candidate
// index k, the four row entries before division and the denominator, for use by the backward.
__device__ __forceinline__ int mat_to_quat(const float* m, float* q, float* row_out, float& den_out,
                                           float& qabs_out) {
    const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6], m21 = m[7],
                m22 = m[8];
    float x[4];
    x[0] = 1.0f + m00 + m11 + m22;
    x[1] = 1.0f + m00 - m11 - m22;
    x[2] = 1.0f - m00 + m11 - m22;
    x[3] = 1.0f - m00 - m11 + m22;
    float qa[4];
#pragma unroll
    for (int k = 0; k < 4; k++) qa[k] = (x[k] > 0.f) ? sqrtf(x[k]) : 0.f;
    int best = 0;
#pragma unroll
    for (int k = 1; k < 4; k++)
        if (qa[k] > qa[best]) best = k;  // first maximal index, like torch.argmax
    float row[4];
    if (best == 0) {
        row[0] = qa[0] * qa[0]; row[1] = m21 - m12; row[2] = m02 - m20; row[3] = m10 - m01;
    } else if (best == 1) {
        row[0] = m21 - m12; row[1] = qa[1] * qa[1]; row[2] = m10 + m01; row[3] = m02 + m20;
    } else if (best == 2) {
        row[0] = m02 - m20; row[1] = m10 + m01; row[2] = qa[2] * qa[2]; row[3] = m12 + m21;
    } else {
        row[0] = m10 - m01; row[1] = m20 + m02; row[2] = m21 + m12; row[3] = qa[3] * qa[3];
    }
    const float den = 2.0f * fmaxf(qa[best], 0.1f);
#pragma unroll
    for (int c = 0; c < 4; c++) {
        q[c] = row[c] / den;
        if (row_out) row_out[c] = row[c];
    }
    den_out = den;
    qabs_out = qa[best];
    return best;
}

__global__ void __launch_bounds__(kLbsThreads)
    lbs_forward_kernel(int N, int J, const float* __restrict__ A, const float* __restrict__ xyz,
                       const float* __restrict__ weights, const float* __restrict__ rotmat,
                       const float* __restrict__ scales, const float* __restrict__ transl, float smpl_scale,
                       float* __restrict__ out_xyz, float* __restrict__ out_rotq, float* __restrict__ out_scales,
                       float* __restrict__ out_T, int w_vec4) {
    __shared__ float sA[kMaxJoints * 16];
    const int f = blockIdx.y;
    for (int k = threadIdx.x; k < J * 16; k += kLbsThreads) sA[k] = __ldg(A + (size_t)f * J * 16 + k);
    __syncthreads();
    const int i = blockIdx.x * kLbsThreads + threadIdx.x;
    if (i >= N) return;

    float w[kMaxJoints];
    load_weights(weights, i, J, w_vec4 != 0, w);
    float T[12], T33;
    blend_transform(sA, w, J, T, T33);

    const float x = __ldg(xyz + 3 * (size_t)i + 0), y = __ldg(xyz + 3 * (size_t)i + 1), z = __ldg(xyz + 3 * (size_t)i + 2);
    float tx = 0.f, ty = 0.f, tz = 0.f;
    if (transl != nullptr) {
        tx = __ldg(transl + 3 * f + 0);
        ty = __ldg(transl + 3 * f + 1);
        tz = __ldg(transl + 3 * f + 2);
    }
    const size_t o = (size_t)f * N + i;
    out_xyz[3 * o + 0] = (T[0] * x + T[1] * y + T[2] * z + T[3]) * smpl_scale + tx;
    out_xyz[3 * o + 1] = (T[4] * x + T[5] * y + T[6] * z + T[7]) * smpl_scale + ty;
    out_xyz[3 * o + 2] = (T[8] * x + T[9] * y + T[10] * z + T[11]) * smpl_scale + tz;

    float Rm[9];
#pragma unroll
    for (int k = 0; k < 9; k++) Rm[k] = __ldg(rotmat + 9 * (size_t)i + k);
    float Rp[9];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) Rp[3 * r + c] = T[4 * r + 0] * Rm[c] + T[4 * r + 1] * Rm[3 + c] + T[4 * r + 2] * Rm[6 + c];
    float q[4], den, qabs;
    mat_to_quat(Rp, q, nullptr, den, qabs);
    reinterpret_cast<float4*>(out_rotq)[o] = make_float4(q[0], q[1], q[2], q[3]);

    if (scales != nullptr && out_scales != nullptr) {
        out_scales[3 * o + 0] = __ldg(scales + 3 * (size_t)i + 0) * smpl_scale;
        out_scales[3 * o + 1] = __ldg(scales + 3 * (size_t)i + 1) * smpl_scale;
        out_scales[3 * o + 2] = __ldg(scales + 3 * (size_t)i + 2) * smpl_scale;
    }
    if (out_T != nullptr) {
        float4* dst = reinterpret_cast<float4*>(out_T) + 4 * o;
        dst[0] = make_float4(T[0], T[1], T[2], T[3]);
        dst[1] = make_float4(T[4], T[5], T[6], T[7]);
        dst[2] = make_float4(T[8], T[9], T[10], T[11]);
        // bottom row: blend of the joints' bottom rows, as torch.matmul(W, A) produces it
        float b0 = 0.f, b1 = 0.f, b2 = 0.f;
#pragma unroll
        for (int j = 0; j < kMaxJoints; j++) {
            if (j < J) {
                b0 = fmaf(w[j], sA[16 * j + 12], b0);
                b1 = fmaf(w[j], sA[16 * j + 13], b1);
                b2 = fmaf(w[j], sA[16 * j + 14], b2);
            }
        }
        dst[3] = make_float4(b0, b1, b2, T33);
    }
}

// Backward.  One thread per Gaussian, frames looped inside (F = 1 in training).  dA is reduced
// per warp through shared memory (lane j owns row j of the J x 12 matrix and sums w_j * dT over the
// warp's 32 Gaussians: one weight load + three broadcast 16-byte loads per 12 FMAs), accumulated in
// registers over the CTA's grid-stride chunks, and flushed with one atomicAdd per entry per warp.
__global__ void __launch_bounds__(kLbsThreads)
    lbs_backward_kernel(int F, int N, int J, const float* __restrict__ A, const float* __restrict__ xyz,
                        const float* __restrict__ weights, const float* __restrict__ rotmat,
                        const float* __restrict__ scales, const float* __restrict__ transl, float smpl_scale,
                        const float* __restrict__ g_xyz, const float* __restrict__ g_rotq,
                        const float* __restrict__ g_scales, float* __restrict__ dA, float* __restrict__ d_xyz,
                        float* __restrict__ d_weights, float* __restrict__ d_rotmat, float* __restrict__ d_scales,
                        float* __restrict__ d_transl, float* __restrict__ d_smpl_scale, int w_vec4) {
    __shared__ float sA[kMaxJoints * 16];
    __shared__ float s_w[kLbsThreads / 32][32][kMaxJoints + 1];
    __shared__ __align__(16) float s_dT[kLbsThreads / 32][32][12];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    (void)0;
    const int chunks = (N + kLbsThreads - 1) / kLbsThreads;
    (void)transl;

    for (int f = 0; f < F; f++) {
        __syncthreads();
        for (int k = threadIdx.x; k < J * 16; k += kLbsThreads) sA[k] = __ldg(A + (size_t)f * J * 16 + k);
        __syncthreads();
        float accA[12];  // lane j owns row j of the J x 12 gradient (J <= 32)
#pragma unroll
        for (int k = 0; k < 12; k++) accA[k] = 0.f;
        float acc_tr[3] = {0.f, 0.f, 0.f};
        float acc_scale = 0.f;

        for (int chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
            const int i = chunk * kLbsThreads + threadIdx.x;
            const bool valid = i < N;
            float w[kMaxJoints];
            float dT[12];
#pragma unroll
            for (int k = 0; k < 12; k++) dT[k] = 0.f;
#pragma unroll
            for (int k = 0; k < kMaxJoints; k++) w[k] = 0.f;
            if (valid) {
                load_weights(weights, i, J, w_vec4 != 0, w);
                float T[12], T33;
                blend_transform(sA, w, J, T, T33);
                const float x = __ldg(xyz + 3 * (size_t)i + 0), y = __ldg(xyz + 3 * (size_t)i + 1),
                            z = __ldg(xyz + 3 * (size_t)i + 2);
                const size_t o = (size_t)f * N + i;
                const float gx = __ldg(g_xyz + 3 * o + 0), gy = __ldg(g_xyz + 3 * o + 1), gz = __ldg(g_xyz + 3 * o + 2);
                // xyz' = (T3 x + t) * s + transl
                const float sgx = smpl_scale * gx, sgy = smpl_scale * gy, sgz = smpl_scale * gz;
                dT[0] = sgx * x; dT[1] = sgx * y; dT[2] = sgx * z; dT[3] = sgx;
                dT[4] = sgy * x; dT[5] = sgy * y; dT[6] = sgy * z; dT[7] = sgy;
                dT[8] = sgz * x; dT[9] = sgz * y; dT[10] = sgz * z; dT[11] = sgz;
                float dx = T[0] * sgx + T[4] * sgy + T[8] * sgz;
                float dy = T[1] * sgx + T[5] * sgy + T[9] * sgz;
                float dz = T[2] * sgx + T[6] * sgy + T[10] * sgz;
                acc_tr[0] += gx; acc_tr[1] += gy; acc_tr[2] += gz;
                acc_scale += gx * (T[0] * x + T[1] * y + T[2] * z + T[3]) + gy * (T[4] * x + T[5] * y + T[6] * z + T[7]) +
                             gz * (T[8] * x + T[9] * y + T[10] * z + T[11]);

                // R' = T3 R ; q = mat_to_quat(R')
                float Rm[9], Rp[9];
#pragma unroll
                for (int k = 0; k < 9; k++) Rm[k] = __ldg(rotmat + 9 * (size_t)i + k);
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++)
                        Rp[3 * r + c] = T[4 * r + 0] * Rm[c] + T[4 * r + 1] * Rm[3 + c] + T[4 * r + 2] * Rm[6 + c];
                float q[4], row[4], den, qabs;
                const int best = mat_to_quat(Rp, q, row, den, qabs);
                const float4 gq4 = __ldg(reinterpret_cast<const float4*>(g_rotq) + o);
                const float gq[4] = {gq4.x, gq4.y, gq4.z, gq4.w};
                float drow[4];
                float dden = 0.f;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    drow[c] = gq[c] / den;
                    dden -= gq[c] * row[c] / (den * den);
                }
                // x_k = 1 +- m00 +- m11 +- m22 feeds the diagonal entry (qabs^2 == x_k for x_k > 0) and,
                // when qabs > 0.1, the denominator 2*qabs.
                float dxk = 0.f;
                if (qabs > 0.f) {
                    dxk = drow[best];
                    if (qabs > 0.1f) dxk += dden * 2.0f * 0.5f / qabs;
                }
                float dRp[9];
#pragma unroll
                for (int k = 0; k < 9; k++) dRp[k] = 0.f;
                // signs of m00, m11, m22 in x_k
                const float s00 = (best == 0 || best == 1) ? 1.f : -1.f;
                const float s11 = (best == 0 || best == 2) ? 1.f : -1.f;
                const float s22 = (best == 0 || best == 3) ? 1.f : -1.f;
                dRp[0] += s00 * dxk;
                dRp[4] += s11 * dxk;
                dRp[8] += s22 * dxk;
                // off-diagonal entries; index map m01=1 m02=2 m10=3 m12=5 m20=6 m21=7
                if (best == 0) {
                    dRp[7] += drow[1]; dRp[5] -= drow[1];  // m21 - m12
                    dRp[2] += drow[2]; dRp[6] -= drow[2];  // m02 - m20
                    dRp[3] += drow[3]; dRp[1] -= drow[3];  // m10 - m01
                } else if (best == 1) {
                    dRp[7] += drow[0]; dRp[5] -= drow[0];  // m21 - m12
                    dRp[3] += drow[2]; dRp[1] += drow[2];  // m10 + m01
                    dRp[2] += drow[3]; dRp[6] += drow[3];  // m02 + m20
                } else if (best == 2) {
                    dRp[2] += drow[0]; dRp[6] -= drow[0];  // m02 - m20
                    dRp[3] += drow[1]; dRp[1] += drow[1];  // m10 + m01
                    dRp[5] += drow[3]; dRp[7] += drow[3];  // m12 + m21
                } else {
                    dRp[3] += drow[0]; dRp[1] -= drow[0];  // m10 - m01
                    dRp[6] += drow[1]; dRp[2] += drow[1];  // m20 + m02
                    dRp[7] += drow[2]; dRp[5] += drow[2];  // m21 + m12
                }
                // R' = T3 R  ->  dT3 += dR' R^T ; dR = T3^T dR'
                float dR[9];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        dT[4 * r + c] += dRp[3 * r + 0] * Rm[3 * c + 0] + dRp[3 * r + 1] * Rm[3 * c + 1] + dRp[3 * r + 2] * Rm[3 * c + 2];
                        dR[3 * r + c] = T[0 + r] * dRp[c] + T[4 + r] * dRp[3 + c] + T[8 + r] * dRp[6 + c];
                    }
                // per-Gaussian outputs (summed over frames when F > 1)
                if (F == 1) {
                    d_xyz[3 * (size_t)i + 0] = dx; d_xyz[3 * (size_t)i + 1] = dy; d_xyz[3 * (size_t)i + 2] = dz;
#pragma unroll
                    for (int k = 0; k < 9; k++) d_rotmat[9 * (size_t)i + k] = dR[k];
                } else {
                    atomicAdd(d_xyz + 3 * (size_t)i + 0, dx); atomicAdd(d_xyz + 3 * (size_t)i + 1, dy);
                    atomicAdd(d_xyz + 3 * (size_t)i + 2, dz);
#pragma unroll
                    for (int k = 0; k < 9; k++) atomicAdd(d_rotmat + 9 * (size_t)i + k, dR[k]);
                }
                if (scales != nullptr && g_scales != nullptr && d_scales != nullptr) {
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const float gs = __ldg(g_scales + 3 * o + c);
                        acc_scale += gs * __ldg(scales + 3 * (size_t)i + c);
                        if (F == 1) d_scales[3 * (size_t)i + c] = gs * smpl_scale;
                        else atomicAdd(d_scales + 3 * (size_t)i + c, gs * smpl_scale);
                    }
                }
                // dL/dw_j = <dT, A_j[:3,:4]>
                for (int j = 0; j < J; j++) {
                    const float* Aj = sA + 16 * j;
                    float s = 0.f;
#pragma unroll
                    for (int k = 0; k < 12; k++) s = fmaf(dT[k], Aj[k], s);
                    if (F == 1) d_weights[(size_t)i * J + j] = s;
                    else atomicAdd(d_weights + (size_t)i * J + j, s);
                }
            }
            // dA[j][k] += sum over the warp's Gaussians of w_j * dT[k]
#pragma unroll
            for (int k = 0; k < 12; k++) s_dT[warp][lane][k] = dT[k];
#pragma unroll
            for (int j = 0; j < kMaxJoints; j++)
                if (j < J) s_w[warp][lane][j] = w[j];
            __syncwarp();
            // lane j: accA[k] += sum_t w[t][j] * dT[t][k]; the dT row is a broadcast of three 16-byte loads,
            // the weight column a conflict-free stride-1 load (rows of 33 floats)
            if ((int)lane < J) {
#pragma unroll 4
                for (int t = 0; t < 32; t++) {
                    const float wj = s_w[warp][t][lane];
                    const float4* row = reinterpret_cast<const float4*>(&s_dT[warp][t][0]);
                    const float4 d0 = row[0], d1 = row[1], d2 = row[2];
                    accA[0] = fmaf(wj, d0.x, accA[0]); accA[1] = fmaf(wj, d0.y, accA[1]);
                    accA[2] = fmaf(wj, d0.z, accA[2]); accA[3] = fmaf(wj, d0.w, accA[3]);
                    accA[4] = fmaf(wj, d1.x, accA[4]); accA[5] = fmaf(wj, d1.y, accA[5]);
                    accA[6] = fmaf(wj, d1.z, accA[6]); accA[7] = fmaf(wj, d1.w, accA[7]);
                    accA[8] = fmaf(wj, d2.x, accA[8]); accA[9] = fmaf(wj, d2.y, accA[9]);
                    accA[10] = fmaf(wj, d2.z, accA[10]); accA[11] = fmaf(wj, d2.w, accA[11]);
                }
            }
            __syncwarp();
        }
        // flush dA for this frame (rows 0..2 of each joint's 4x4; the bottom row gets no gradient): the
        // CTA's warps are summed in shared memory first, then ONE atomic per entry and CTA.  (One atomic
        // per entry and WARP -- 1.4 M reductions onto 288 addresses when every Gaussian has a gradient --
        // made this kernel 4x slower on views where the human is unoccluded: same-address atomics
        // serialise at the L2.)
        __syncthreads();   // every warp is done with s_dT
        float* const s_red = &s_dT[0][0][0];   // [warps][kMaxJoints][12], exactly s_dT's size
        if ((int)lane < J) {
#pragma unroll
            for (int k = 0; k < 12; k++) s_red[(warp * kMaxJoints + lane) * 12 + k] = accA[k];
        }
        // translation / scale gradients: warp reduce
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            acc_tr[0] += __shfl_xor_sync(0xffffffffu, acc_tr[0], d);
            acc_tr[1] += __shfl_xor_sync(0xffffffffu, acc_tr[1], d);
            acc_tr[2] += __shfl_xor_sync(0xffffffffu, acc_tr[2], d);
            acc_scale += __shfl_xor_sync(0xffffffffu, acc_scale, d);
        }
        constexpr int kWarps = kLbsThreads / 32;
        float* const s_tail = &s_w[0][0][0];   // [warps][4]
        if (lane == 0) {
            s_tail[warp * 4 + 0] = acc_tr[0];
            s_tail[warp * 4 + 1] = acc_tr[1];
            s_tail[warp * 4 + 2] = acc_tr[2];
            s_tail[warp * 4 + 3] = acc_scale;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < J * 12; e += kLbsThreads) {
            float v = 0.f;
#pragma unroll
            for (int wp = 0; wp < kWarps; wp++) v += s_red[(wp * kMaxJoints + e / 12) * 12 + e % 12];
            if (v != 0.f) atomicAdd(dA + ((size_t)f * J + e / 12) * 16 + e % 12, v);
        }
        if (threadIdx.x < 4) {
            float v = 0.f;
#pragma unroll
            for (int wp = 0; wp < kWarps; wp++) v += s_tail[wp * 4 + threadIdx.x];
            if (threadIdx.x < 3) {
                if (d_transl != nullptr) atomicAdd(d_transl + 3 * f + threadIdx.x, v);
            } else if (d_smpl_scale != nullptr) {
                atomicAdd(d_smpl_scale, v);
            }
        }
    }
}
}  // namespace

static int check_lbs_args(int F, int N, int J) {
    if (F <= 0 || N < 0 || J <= 0 || J > kMaxJoints) {
        set_error("lbs: need F > 0, N >= 0, 0 < J <= %d (got F=%d N=%d J=%d)", kMaxJoints, F, N, J);
        return -1;
    }
    return 0;
}

int launch_lbs_forward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                       const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                       float* out_xyz, float* out_rotq, float* out_scales, float* out_T, cudaStream_t stream) {
    if (check_lbs_args(F, N, J)) return -1;
    if (N == 0) return 0;
    const int w_vec4 = (J % 4 == 0 && (reinterpret_cast<uintptr_t>(weights) & 15u) == 0) ? 1 : 0;
    const dim3 grid((unsigned)cdiv(N, kLbsThreads), (unsigned)F, 1);
    lbs_forward_kernel<<<grid, kLbsThreads, 0, stream>>>(N, J, A, xyz, weights, rotmat, scales, transl, smpl_scale,
                                                         out_xyz, out_rotq, out_scales, out_T, w_vec4);
    HB_LAUNCH_OK("lbs_forward", stream, false);
    return 0;
}

int launch_lbs_backward(int F, int N, int J, const float* A, const float* xyz, const float* weights,
                        const float* rotmat, const float* scales, const float* transl, float smpl_scale,
                        const float* g_xyz, const float* g_rotq, const float* g_scales, float* dA, float* d_xyz,
                        float* d_weights, float* d_rotmat, float* d_scales, float* d_transl, float* d_smpl_scale,
                        cudaStream_t stream) {
    if (check_lbs_args(F, N, J)) return -1;
    HB_CUDA_OK(cudaMemsetAsync(dA, 0, sizeof(float) * (size_t)F * J * 16, stream));
    if (d_transl) HB_CUDA_OK(cudaMemsetAsync(d_transl, 0, sizeof(float) * 3 * (size_t)F, stream));
    if (d_smpl_scale) HB_CUDA_OK(cudaMemsetAsync(d_smpl_scale, 0, sizeof(float), stream));
    if (N == 0) return 0;
    if (F > 1) {
        HB_CUDA_OK(cudaMemsetAsync(d_xyz, 0, sizeof(float) * 3 * (size_t)N, stream));
        HB_CUDA_OK(cudaMemsetAsync(d_weights, 0, sizeof(float) * (size_t)N * J, stream));
        HB_CUDA_OK(cudaMemsetAsync(d_rotmat, 0, sizeof(float) * 9 * (size_t)N, stream));
        if (d_scales) HB_CUDA_OK(cudaMemsetAsync(d_scales, 0, sizeof(float) * 3 * (size_t)N, stream));
    }
    const int w_vec4 = (J % 4 == 0 && (reinterpret_cast<uintptr_t>(weights) & 15u) == 0) ? 1 : 0;
    int blocks = cdiv(N, kLbsThreads);
    if (blocks > 148 * 4) blocks = 148 * 4;
    lbs_backward_kernel<<<blocks, kLbsThreads, 0, stream>>>(F, N, J, A, xyz, weights, rotmat, scales, transl,
                                                            smpl_scale, g_xyz, g_rotq, g_scales, dA, d_xyz, d_weights,
                                                            d_rotmat, d_scales, d_transl, d_smpl_scale, w_vec4);
    HB_LAUNCH_OK("lbs_backward", stream, false);
    return 0;
}

}  // namespace hb
