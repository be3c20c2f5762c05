// Fused per-Gaussian backward for sm_100a: one kernel replaces the reference's
//   computeCov2DCUDA      (cuda_rasterizer/backward.cu:144-274)  dL/dconic -> dL/dcov3D, dL/dmean
//   preprocessCUDA (bwd)  (backward.cu:346-396)                  projection path of dL/dmean
//   computeColorFromSH    (backward.cu:20-139)                   dL/dsh, view-direction path of dL/dmean
//   computeCov3D (bwd)    (backward.cu:278-341)                  dL/dscale, dL/drot
// and the nine torch::zeros memsets that precede them (rasterize_points.cu:151-159): every
// output element is written exactly once here, zeros included (culled Gaussians, SH bands above
// the active degree), so the caller may hand in uninitialised memory.
// The 3D covariance is recomputed from scale/rotation instead of being stored by the forward
// pass (saves 24 B written + 24 B read per Gaussian).
// Algorithmic traffic at SH degree 3: ~303 B read + 256 B written per visible Gaussian.
#include "common.cuh"

#include <cstring>

namespace hb {

namespace {

struct PreBwdArgs {
    int P, D, M;
    const float* means3D;
    const int* radii;
    const float* shs;
    const uint8_t* clamped;
    const float* scales;
    const float* rotations;
    float scale_modifier;
    const float* cov3D_precomp;
    const float* view;
    const float* proj;
    const float* campos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    const float* grad_pack;   // [P,8]  (input: accumulated by the blend backward) colour 3, mean2D 2, conic 3
    const float* grad_pack2;  // [n2,8] or null: a second image's share for the rows below n2 (layer.cu)
    int n2;
    int acc_mask;             // kAcc* bits: ADD the visible rows of that output into its buffer, leave culled rows alone
    const float* opac_scratch;  // [P] or null: the blend backward's dL/dopacity, moved / added to dL_dopacity here
    float* dL_dopacity;
    float* dL_dmean2D;        // [P,3]  (output: unpacked record, z = 0)
    float* dL_dconic;         // [P,4]  (output: unpacked record, (x, y, 0, w))
    float* dL_dmean3D;        // [P,3]
    float* dL_dcolor;         // [P,3]  (output: unpacked record)
    float* dL_dcov3D;         // [P,6]
    float* dL_dsh;            // [P,M,3]
    float* dL_dscale;         // [P,3]
    float* dL_drot;           // [P,4]
    int sh_vec4;
    int sh_accumulate;  // add visible rows of dL_dsh into the existing contents, leave culled rows alone
};

// Gradient sinks (view-batched training): an output with its kAcc bit set is a leaf parameter's accumulation
// buffer shared by all views of the step; fire-and-forget reductions at the L2 instead of a store, and no
// zero rows for culled Gaussians.
__device__ __forceinline__ void red_f1(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void put3(float* p, float x, float y, float z, bool acc) {
    if (acc) {
        red_f1(p, x);
        red_f1(p + 1, y);
        red_f1(p + 2, z);
    } else {
        p[0] = x;
        p[1] = y;
        p[2] = z;
    }
}

__device__ __forceinline__ void store_sh_grad(const PreBwdArgs& a, int idx, const float* g /*48*/) {
    if (a.sh_vec4) {
        float4* dst = reinterpret_cast<float4*>(a.dL_dsh) + 12 * (size_t)idx;
#pragma unroll
        for (int i = 0; i < 12; i++) dst[i] = make_float4(g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]);
    } else {
        float* dst = a.dL_dsh + (size_t)idx * a.M * 3;
        const int nf = a.M * 3;
#pragma unroll
        for (int i = 0; i < 48; i++)
            if (i < nf) dst[i] = g[i];
        for (int i = 48; i < nf; i++) dst[i] = 0.f;
    }
}

// SH backward for one Gaussian (backward.cu:20-139): dL/dsh row `gsh` (bands above the active degree are
// zero) and the view-direction path of dL/dmean (dnormvdv, auxiliary.h:107-117), added to dmean_*.
__device__ __forceinline__ void sh_backward(int deg, const float* sh, float ox, float oy, float oz, const float* dRGB,
                                            float* gsh, float& dmean_x, float& dmean_y, float& dmean_z) {
    const float inv_len = 1.f / sqrtf(ox * ox + oy * oy + oz * oz);
    const float x = ox * inv_len, y = oy * inv_len, z = oz * inv_len;
    float w[16];  // dRGB/dsh_k
#pragma unroll
    for (int k = 0; k < 16; k++) w[k] = 0.f;
    float ddir_x = 0.f, ddir_y = 0.f, ddir_z = 0.f;  // dL/ddir
    w[0] = HB_SH_C0;
    if (deg > 0) {
        w[1] = -HB_SH_C1 * y;
        w[2] = HB_SH_C1 * z;
        w[3] = -HB_SH_C1 * x;
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        if (deg > 1) {
            w[4] = HB_SH_C2_0 * xy;
            w[5] = HB_SH_C2_1 * yz;
            w[6] = HB_SH_C2_2 * (2.f * zz - xx - yy);
            w[7] = HB_SH_C2_3 * xz;
            w[8] = HB_SH_C2_4 * (xx - yy);
            if (deg > 2) {
                w[9] = HB_SH_C3_0 * y * (3.f * xx - yy);
                w[10] = HB_SH_C3_1 * xy * z;
                w[11] = HB_SH_C3_2 * y * (4.f * zz - xx - yy);
                w[12] = HB_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = HB_SH_C3_4 * x * (4.f * zz - xx - yy);
                w[14] = HB_SH_C3_5 * z * (xx - yy);
                w[15] = HB_SH_C3_6 * x * (xx - 3.f * yy);
            }
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float dx_ = -HB_SH_C1 * sh[3 * 3 + c];
            float dy_ = -HB_SH_C1 * sh[1 * 3 + c];
            float dz_ = HB_SH_C1 * sh[2 * 3 + c];
            if (deg > 1) {
                dx_ += HB_SH_C2_0 * y * sh[4 * 3 + c] + HB_SH_C2_2 * 2.f * -x * sh[6 * 3 + c] +
                       HB_SH_C2_3 * z * sh[7 * 3 + c] + HB_SH_C2_4 * 2.f * x * sh[8 * 3 + c];
                dy_ += HB_SH_C2_0 * x * sh[4 * 3 + c] + HB_SH_C2_1 * z * sh[5 * 3 + c] +
                       HB_SH_C2_2 * 2.f * -y * sh[6 * 3 + c] + HB_SH_C2_4 * 2.f * -y * sh[8 * 3 + c];
                dz_ += HB_SH_C2_1 * y * sh[5 * 3 + c] + HB_SH_C2_2 * 2.f * 2.f * z * sh[6 * 3 + c] +
                       HB_SH_C2_3 * x * sh[7 * 3 + c];
                if (deg > 2) {
                    dx_ += HB_SH_C3_0 * sh[9 * 3 + c] * 3.f * 2.f * xy + HB_SH_C3_1 * sh[10 * 3 + c] * yz +
                           HB_SH_C3_2 * sh[11 * 3 + c] * -2.f * xy + HB_SH_C3_3 * sh[12 * 3 + c] * -3.f * 2.f * xz +
                           HB_SH_C3_4 * sh[13 * 3 + c] * (-3.f * xx + 4.f * zz - yy) +
                           HB_SH_C3_5 * sh[14 * 3 + c] * 2.f * xz + HB_SH_C3_6 * sh[15 * 3 + c] * 3.f * (xx - yy);
                    dy_ += HB_SH_C3_0 * sh[9 * 3 + c] * 3.f * (xx - yy) + HB_SH_C3_1 * sh[10 * 3 + c] * xz +
                           HB_SH_C3_2 * sh[11 * 3 + c] * (-3.f * yy + 4.f * zz - xx) +
                           HB_SH_C3_3 * sh[12 * 3 + c] * -3.f * 2.f * yz + HB_SH_C3_4 * sh[13 * 3 + c] * -2.f * xy +
                           HB_SH_C3_5 * sh[14 * 3 + c] * -2.f * yz + HB_SH_C3_6 * sh[15 * 3 + c] * -3.f * 2.f * xy;
                    dz_ += HB_SH_C3_1 * sh[10 * 3 + c] * xy + HB_SH_C3_2 * sh[11 * 3 + c] * 4.f * 2.f * yz +
                           HB_SH_C3_3 * sh[12 * 3 + c] * 3.f * (2.f * zz - xx - yy) +
                           HB_SH_C3_4 * sh[13 * 3 + c] * 4.f * 2.f * xz + HB_SH_C3_5 * sh[14 * 3 + c] * (xx - yy);
                }
            }
            ddir_x += dx_ * dRGB[c];
            ddir_y += dy_ * dRGB[c];
            ddir_z += dz_ * dRGB[c];
        }
    }
    // dL/dsh[k][c] = w[k] * dRGB[c]; bands above the active degree have w == 0.
#pragma unroll
    for (int k = 0; k < 16; k++) {
        gsh[3 * k + 0] = w[k] * dRGB[0];
        gsh[3 * k + 1] = w[k] * dRGB[1];
        gsh[3 * k + 2] = w[k] * dRGB[2];
    }
    // dnormvdv (auxiliary.h:107-117): through the normalisation of the view direction.
    const float sum2 = ox * ox + oy * oy + oz * oz;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dmean_x += ((sum2 - ox * ox) * ddir_x - oy * ox * ddir_y - oz * ox * ddir_z) * invsum32;
    dmean_y += (-ox * oy * ddir_x + (sum2 - oy * oy) * ddir_y - oz * oy * ddir_z) * invsum32;
    dmean_z += (-ox * oz * ddir_x - oy * oz * ddir_y + (sum2 - oz * oz) * ddir_z) * invsum32;
}

constexpr int kPbThreads = 128;
constexpr int kPbRowBytes = 208;  // 192-B SH row padded to 52 words: conflict-free 128-bit row access
constexpr int kPbSmemBytes = kPbThreads * kPbRowBytes;

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// One thread per Gaussian.  Memory behaviour (the stage is HBM bound: ~303 B read + 256 B written
// per visible Gaussian, three quarters of it the SH row and its gradient):
//   * visibility is known from `radii` at once, so the SH rows of the visible Gaussians of a warp
//     are requested first thing with cp.async (global -> shared; the 32 rows of a warp are one
//     contiguous 6 KB block, copy j of lane l moves its 16-byte word 32 j + l) and arrive while the
//     covariance chain is being differentiated;
//   * every other input of the Gaussian is requested before the arithmetic starts;
//   * dL/dsh is assembled row by row in the same shared-memory slots and written by the whole warp
//     as twelve 512-byte contiguous stores (zero rows for culled Gaussians included), instead of
//     twelve 16-byte stores per thread at a 192-byte stride.
// Two segments in one launch: blocks [0, nb0) work on a0, the rest on a1 (see preprocess.cu).
__global__ void __launch_bounds__(kPbThreads, 5)
    preprocess_backward_kernel(const PreBwdArgs a0, const PreBwdArgs a1, const int nb0) {
    const bool second = (int)blockIdx.x >= nb0;
    const PreBwdArgs& a = second ? a1 : a0;
    const unsigned blk = second ? blockIdx.x - (unsigned)nb0 : blockIdx.x;
    extern __shared__ __align__(16) unsigned char s_rows[];  // [256][208 B] SH rows / dL_dsh rows (vec4 path)
    __shared__ float s_cam[36];
    if (threadIdx.x < 16) s_cam[threadIdx.x] = __ldg(a.view + threadIdx.x);
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = __ldg(a.proj + threadIdx.x - 16);
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = __ldg(a.campos + threadIdx.x - 32);

    const int idx = blk * kPbThreads + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool valid = idx < a.P;
    const bool vis = valid && (__ldg(a.radii + idx) > 0);
    const bool sh_out = (a.dL_dsh != nullptr && a.M > 0);
    const int deg = a.D;
    const int nf = 3 * (deg + 1) * (deg + 1);
    const int nv = (nf + 3) >> 2;
    const uint32_t s_warp = (uint32_t)__cvta_generic_to_shared(s_rows) + warp * 32u * kPbRowBytes;
    const size_t warp_row0 = (size_t)blk * kPbThreads + warp * 32u;
    const unsigned alive = __ballot_sync(0xffffffffu, vis);

    // ---- SH rows of the visible Gaussians of this warp: in flight from here on ------------------
    if (a.sh_vec4 && a.shs != nullptr && alive != 0u) {
        const float4* g_warp = reinterpret_cast<const float4*>(a.shs) + 12 * warp_row0;
        unsigned row = lane / 12u, col = lane % 12u;  // 16-byte word 32 j + lane = (row, col)
#pragma unroll
        for (int j = 0; j < 12; j++) {
            if (((alive >> row) & 1u) && (int)col < nv)
                cp_async16(s_warp + row * kPbRowBytes + col * 16u, g_warp + 32 * j + lane);
            col += 8u;
            row += 2u;
            if (col >= 12u) {
                col -= 12u;
                row += 1u;
            }
        }
    }
    __syncthreads();  // s_cam
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    float* out_mean = a.dL_dmean3D + 3 * (size_t)idx;
    float* out_cov = a.dL_dcov3D + 6 * (size_t)idx;
    const bool has_scale = (a.cov3D_precomp == nullptr);

    const bool acc_mean = (a.acc_mask & kAccMean3D) != 0, acc_scale = (a.acc_mask & kAccScale) != 0,
               acc_rot = (a.acc_mask & kAccRot) != 0, acc_opac = (a.acc_mask & kAccOpacity) != 0;
    if (valid && a.opac_scratch != nullptr) {
        if (vis) {
            const float v = __ldg(a.opac_scratch + idx);
            if (acc_opac) red_f1(a.dL_dopacity + idx, v);
            else a.dL_dopacity[idx] = v;
        } else if (!acc_opac) {
            a.dL_dopacity[idx] = 0.f;
        }
    }
    if (valid && !vis) {
        if (!acc_mean) out_mean[0] = out_mean[1] = out_mean[2] = 0.f;
        a.dL_dmean2D[3 * (size_t)idx + 0] = 0.f;
        a.dL_dmean2D[3 * (size_t)idx + 1] = 0.f;
        a.dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
        reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        a.dL_dcolor[3 * (size_t)idx + 0] = 0.f;
        a.dL_dcolor[3 * (size_t)idx + 1] = 0.f;
        a.dL_dcolor[3 * (size_t)idx + 2] = 0.f;
#pragma unroll
        for (int i = 0; i < 6; i++) out_cov[i] = 0.f;
        if (a.dL_dscale != nullptr && !acc_scale) {
            a.dL_dscale[3 * (size_t)idx + 0] = 0.f;
            a.dL_dscale[3 * (size_t)idx + 1] = 0.f;
            a.dL_dscale[3 * (size_t)idx + 2] = 0.f;
        }
        if (a.dL_drot != nullptr && !acc_rot) reinterpret_cast<float4*>(a.dL_drot)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }

    // state that crosses the cooperative SH phase
    float mx = 0.f, my = 0.f, mz = 0.f;
    float dmean_x = 0.f, dmean_y = 0.f, dmean_z = 0.f;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float sx = 0.f, sy = 0.f, sz = 0.f, qr = 0.f, qx = 0.f, qy = 0.f, qz = 0.f;
    float R[3][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};  // R[c][r], glm column-major
    float dRGB[3] = {0.f, 0.f, 0.f};

    if (vis) {
        // ---- every other per-Gaussian input, requested before the arithmetic ------------------
        mx = __ldg(a.means3D + 3 * (size_t)idx + 0);
        my = __ldg(a.means3D + 3 * (size_t)idx + 1);
        mz = __ldg(a.means3D + 3 * (size_t)idx + 2);
        // the blend backward's packed record: one 32-byte sector
        float4 pk0 = __ldg(reinterpret_cast<const float4*>(a.grad_pack) + 2 * (size_t)idx);
        float4 pk1 = __ldg(reinterpret_cast<const float4*>(a.grad_pack) + 2 * (size_t)idx + 1);
        // dL/dmean2D as REPORTED is the main image's alone: the reference's second render() has a
        // screen-space tensor of its own that nothing reads (gs_renderer.py:68-82)
        const float out2x = pk0.w, out2y = pk1.x;
        if (a.grad_pack2 != nullptr && idx < a.n2) {
            const float4 q0 = __ldg(reinterpret_cast<const float4*>(a.grad_pack2) + 2 * (size_t)idx);
            const float4 q1 = __ldg(reinterpret_cast<const float4*>(a.grad_pack2) + 2 * (size_t)idx + 1);
            pk0.x += q0.x; pk0.y += q0.y; pk0.z += q0.z; pk0.w += q0.w;
            pk1.x += q1.x; pk1.y += q1.y; pk1.z += q1.z; pk1.w += q1.w;
        }
        const float4 dconic4 = make_float4(pk1.y, pk1.z, 0.f, pk1.w);
        const float g2x = pk0.w, g2y = pk1.x;
        if (a.shs != nullptr) {
            const uint8_t cl = __ldg(a.clamped + idx);
            dRGB[0] = (cl & 1) ? 0.f : pk0.x;
            dRGB[1] = (cl & 2) ? 0.f : pk0.y;
            dRGB[2] = (cl & 4) ? 0.f : pk0.z;
        }
        // the reference's arrays (rasterize_points.cu:151-159 allocates them; dL_dmeans2D and dL_dcolors go
        // back to Python)
        a.dL_dmean2D[3 * (size_t)idx + 0] = out2x;
        a.dL_dmean2D[3 * (size_t)idx + 1] = out2y;
        a.dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
        reinterpret_cast<float4*>(a.dL_dconic)[idx] = dconic4;
        a.dL_dcolor[3 * (size_t)idx + 0] = pk0.x;
        a.dL_dcolor[3 * (size_t)idx + 1] = pk0.y;
        a.dL_dcolor[3 * (size_t)idx + 2] = pk0.z;
        float cov3[6];
        if (has_scale) {
            sx = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 0);
            sy = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 1);
            sz = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 2);
            qr = __ldg(a.rotations + 4 * (size_t)idx + 0);
            qx = __ldg(a.rotations + 4 * (size_t)idx + 1);
            qy = __ldg(a.rotations + 4 * (size_t)idx + 2);
            qz = __ldg(a.rotations + 4 * (size_t)idx + 3);
            R[0][0] = 1.f - 2.f * (qy * qy + qz * qz);
            R[0][1] = 2.f * (qx * qy - qr * qz);
            R[0][2] = 2.f * (qx * qz + qr * qy);
            R[1][0] = 2.f * (qx * qy + qr * qz);
            R[1][1] = 1.f - 2.f * (qx * qx + qz * qz);
            R[1][2] = 2.f * (qy * qz - qr * qx);
            R[2][0] = 2.f * (qx * qz - qr * qy);
            R[2][1] = 2.f * (qy * qz + qr * qx);
            R[2][2] = 1.f - 2.f * (qx * qx + qy * qy);
            const float s[3] = {sx, sy, sz};
            float Mm[3][3];
#pragma unroll
            for (int c = 0; c < 3; c++)
#pragma unroll
                for (int r = 0; r < 3; r++) Mm[c][r] = s[r] * R[c][r];
            cov3[0] = Mm[0][0] * Mm[0][0] + Mm[0][1] * Mm[0][1] + Mm[0][2] * Mm[0][2];
            cov3[1] = Mm[0][0] * Mm[1][0] + Mm[0][1] * Mm[1][1] + Mm[0][2] * Mm[1][2];
            cov3[2] = Mm[0][0] * Mm[2][0] + Mm[0][1] * Mm[2][1] + Mm[0][2] * Mm[2][2];
            cov3[3] = Mm[1][0] * Mm[1][0] + Mm[1][1] * Mm[1][1] + Mm[1][2] * Mm[1][2];
            cov3[4] = Mm[1][0] * Mm[2][0] + Mm[1][1] * Mm[2][1] + Mm[1][2] * Mm[2][2];
            cov3[5] = Mm[2][0] * Mm[2][0] + Mm[2][1] * Mm[2][1] + Mm[2][2] * Mm[2][2];
        } else {
#pragma unroll
            for (int i = 0; i < 6; i++) cov3[i] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + i);
        }

        // ---- cov2D backward (backward.cu:144-274) -------------------------------------------------
        const float dcon_x = dconic4.x, dcon_y = dconic4.y, dcon_z = dconic4.w;
        float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
        float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
        const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
        const float limx = 1.3f * a.tan_fovx;
        const float limy = 1.3f * a.tan_fovy;
        const float txtz = tx / tz;
        const float tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float hx = a.focal_x, hy = a.focal_y;
        const float J00 = hx / tz, J02 = -(hx * tx) / (tz * tz);
        const float J11 = hy / tz, J12 = -(hy * ty) / (tz * tz);
        // T = W J, T[c][r]; third column is zero.
        const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
        const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
        const float V00 = cov3[0], V01 = cov3[1], V02 = cov3[2], V11 = cov3[3], V12 = cov3[4], V22 = cov3[5];
        // rows of T times Vrk (used by both cov2D and dL/dT)
        const float TV00 = T00 * V00 + T01 * V01 + T02 * V02;
        const float TV01 = T00 * V01 + T01 * V11 + T02 * V12;
        const float TV02 = T00 * V02 + T01 * V12 + T02 * V22;
        const float TV10 = T10 * V00 + T11 * V01 + T12 * V02;
        const float TV11 = T10 * V01 + T11 * V11 + T12 * V12;
        const float TV12 = T10 * V02 + T11 * V12 + T12 * V22;
        const float ca = TV00 * T00 + TV01 * T01 + TV02 * T02 + 0.3f;
        const float cb = TV10 * T00 + TV11 * T01 + TV12 * T02;
        const float cc = TV10 * T10 + TV11 * T11 + TV12 * T12 + 0.3f;

        const float denom = ca * cc - cb * cb;
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-cc * cc * dcon_x + 2.f * cb * cc * dcon_y + (denom - ca * cc) * dcon_z);
            dL_dc = denom2inv * (-ca * ca * dcon_z + 2.f * ca * cb * dcon_y + (denom - ca * cc) * dcon_x);
            dL_db = denom2inv * 2.f * (cb * cc * dcon_x - (denom + 2.f * cb * cb) * dcon_y + ca * cb * dcon_z);
            dcov[0] = T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            dcov[3] = T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            dcov[5] = T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            dcov[1] = 2.f * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2.f * T10 * T11 * dL_dc;
            dcov[2] = 2.f * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2.f * T10 * T12 * dL_dc;
            dcov[4] = 2.f * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2.f * T11 * T12 * dL_dc;
        } else {
#pragma unroll
            for (int i = 0; i < 6; i++) dcov[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 6; i++) out_cov[i] = dcov[i];

        const float dL_dT00 = 2.f * TV00 * dL_da + TV10 * dL_db;
        const float dL_dT01 = 2.f * TV01 * dL_da + TV11 * dL_db;
        const float dL_dT02 = 2.f * TV02 * dL_da + TV12 * dL_db;
        const float dL_dT10 = 2.f * TV10 * dL_dc + TV00 * dL_db;
        const float dL_dT11 = 2.f * TV11 * dL_dc + TV01 * dL_db;
        const float dL_dT12 = 2.f * TV12 * dL_dc + TV02 * dL_db;
        const float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
        const float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
        const float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
        const float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
        const float itz = 1.f / tz;
        const float itz2 = itz * itz;
        const float itz3 = itz2 * itz;
        const float dL_dtx = x_grad_mul * -hx * itz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -hy * itz2 * dL_dJ12;
        const float dL_dtz = -hx * itz2 * dL_dJ00 - hy * itz2 * dL_dJ11 + (2.f * hx * tx) * itz3 * dL_dJ02 +
                             (2.f * hy * ty) * itz3 * dL_dJ12;
        // transformVec4x3Transpose
        dmean_x = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
        dmean_y = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
        dmean_z = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

        // ---- projection path (backward.cu:370-387) ---------------------------------------------
        {
            const float m_hw = Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15];
            const float m_w = 1.0f / (m_hw + 0.0000001f);
            const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
            const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
            dmean_x += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
            dmean_y += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
            dmean_z += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
        }
    }

    // ---- SH backward (backward.cu:20-139) ------------------------------------------------------
    float gsh[48];
    bool have_gsh = false;  // gsh holds this Gaussian's dL/dsh row (non-vec4 path stores it itself)
    if (a.sh_vec4 && sh_out) {
        // cooperative path: rows in shared memory
        if (a.shs != nullptr) cp_async_wait_all();
        __syncwarp();
        const uint32_t mine = s_warp + lane * kPbRowBytes;
        if (vis && a.shs != nullptr) {
            float sh[48];
#pragma unroll
            for (int i = 0; i < 12; i++) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < nv) v = lds128(mine + 16u * i);
                sh[4 * i + 0] = v.x;
                sh[4 * i + 1] = v.y;
                sh[4 * i + 2] = v.z;
                sh[4 * i + 3] = v.w;
            }
            sh_backward(deg, sh, mx - V[32], my - V[33], mz - V[34], dRGB, gsh, dmean_x, dmean_y, dmean_z);
#pragma unroll
            for (int i = 0; i < 12; i++) sts128(mine + 16u * i, gsh[4 * i], gsh[4 * i + 1], gsh[4 * i + 2], gsh[4 * i + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < 12; i++) sts128(mine + 16u * i, 0.f, 0.f, 0.f, 0.f);
        }
        __syncwarp();
        float4* g_out = reinterpret_cast<float4*>(a.dL_dsh) + 12 * warp_row0;
        const long long rows_left = (long long)a.P - (long long)warp_row0;  // rows of this warp inside [0, P)
        unsigned row = lane / 12u, col = lane % 12u;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            if ((long long)row < rows_left) {
                if (!a.sh_accumulate) {
                    g_out[32 * j + lane] = lds128(s_warp + row * kPbRowBytes + col * 16u);
                } else if ((alive >> row) & 1u) {
                    // view-batch accumulation: fire-and-forget 16-byte reductions at the L2, still
                    // 512 contiguous bytes per warp instruction; culled rows cost nothing
                    const float4 v = lds128(s_warp + row * kPbRowBytes + col * 16u);
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(g_out + 32 * j + lane),
                                 "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                                 : "memory");
                }
            }
            col += 8u;
            row += 2u;
            if (col >= 12u) {
                col -= 12u;
                row += 1u;
            }
        }
    } else if (valid) {
        if (vis && a.shs != nullptr) {
            float sh[48];
            const float* src = a.shs + (size_t)idx * a.M * 3;
#pragma unroll
            for (int i = 0; i < 48; i++) sh[i] = (i < nf) ? __ldg(src + i) : 0.f;
            sh_backward(deg, sh, mx - V[32], my - V[33], mz - V[34], dRGB, gsh, dmean_x, dmean_y, dmean_z);
            have_gsh = true;
        }
        if (sh_out && a.sh_accumulate) {
            if (have_gsh) {
                float* dst = a.dL_dsh + (size_t)idx * a.M * 3;
#pragma unroll
                for (int i = 0; i < 48; i++)   // other views may be adding to the same row on other streams
                    if (i < a.M * 3) atomicAdd(dst + i, gsh[i]);
            }
        } else if (sh_out) {
            if (!have_gsh) {
#pragma unroll
                for (int i = 0; i < 48; i++) gsh[i] = 0.f;
            }
            store_sh_grad(a, idx, gsh);
        }
    }

    if (!vis) return;
    put3(out_mean, dmean_x, dmean_y, dmean_z, acc_mean);

    // ---- cov3D backward (backward.cu:278-341) --------------------------------------------------
    if (has_scale && a.dL_dscale != nullptr) {
        const float s[3] = {sx, sy, sz};
        float Mm[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) Mm[c][r] = s[r] * R[c][r];
        // dL_dSigma (symmetric, glm column-major == row-major here)
        float dS[3][3];
        dS[0][0] = dcov[0];
        dS[0][1] = dS[1][0] = 0.5f * dcov[1];
        dS[0][2] = dS[2][0] = 0.5f * dcov[2];
        dS[1][1] = dcov[3];
        dS[1][2] = dS[2][1] = 0.5f * dcov[4];
        dS[2][2] = dcov[5];
        // dL_dM = 2 * M * dL_dSigma  (glm: result[c][r] = sum_k M[k][r] * dS[c][k])
        float dM[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++)
                dM[c][r] = 2.f * (Mm[0][r] * dS[c][0] + Mm[1][r] * dS[c][1] + Mm[2][r] * dS[c][2]);
        // Rt = transpose(R): Rt[c][r] = R[r][c];  dL_dMt[c][r] = dM[r][c]
        // dL_dscale.k = dot(Rt[k], dL_dMt[k]) = sum_r R[r][k] * dM[r][k]
        float dsc[3];
#pragma unroll
        for (int k = 0; k < 3; k++) dsc[k] = R[0][k] * dM[0][k] + R[1][k] * dM[1][k] + R[2][k] * dM[2][k];
        put3(a.dL_dscale + 3 * (size_t)idx, dsc[0], dsc[1], dsc[2], acc_scale);
        // dL_dMt[k] *= s[k]   ->  Mt[c][r] = dM[r][c] * s[c]
        float Mt[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) Mt[c][r] = dM[r][c] * s[c];
        float4 dq;
        dq.x = 2.f * qz * (Mt[0][1] - Mt[1][0]) + 2.f * qy * (Mt[2][0] - Mt[0][2]) + 2.f * qx * (Mt[1][2] - Mt[2][1]);
        dq.y = 2.f * qy * (Mt[1][0] + Mt[0][1]) + 2.f * qz * (Mt[2][0] + Mt[0][2]) + 2.f * qr * (Mt[1][2] - Mt[2][1]) -
               4.f * qx * (Mt[2][2] + Mt[1][1]);
        dq.z = 2.f * qx * (Mt[1][0] + Mt[0][1]) + 2.f * qr * (Mt[2][0] - Mt[0][2]) + 2.f * qz * (Mt[1][2] + Mt[2][1]) -
               4.f * qy * (Mt[2][2] + Mt[0][0]);
        dq.w = 2.f * qr * (Mt[0][1] - Mt[1][0]) + 2.f * qx * (Mt[2][0] + Mt[0][2]) + 2.f * qy * (Mt[1][2] + Mt[2][1]) -
               4.f * qz * (Mt[1][1] + Mt[0][0]);
        if (acc_rot) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(reinterpret_cast<float4*>(a.dL_drot) + idx),
                         "f"(dq.x), "f"(dq.y), "f"(dq.z), "f"(dq.w)
                         : "memory");
        } else {
            reinterpret_cast<float4*>(a.dL_drot)[idx] = dq;
        }
    } else {
        if (a.dL_dscale != nullptr && !acc_scale) {
            a.dL_dscale[3 * (size_t)idx + 0] = 0.f;
            a.dL_dscale[3 * (size_t)idx + 1] = 0.f;
            a.dL_dscale[3 * (size_t)idx + 2] = 0.f;
        }
        if (a.dL_drot != nullptr && !acc_rot) reinterpret_cast<float4*>(a.dL_drot)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// =========================================================================================================
// The scene segment's per-Gaussian backward for ALL views of a training step in one launch
// (hugs_b200_scene_backward_views).  A view-batched step runs the kernel above once per view: every launch
// reads the Gaussian's SH row (192 B) and its parameters again and adds its gradients to the parameters'
// accumulation buffers (read-modify-write of 236 B at the DRAM), 744 B per Gaussian and VIEW in all.  The scene
// Gaussians are leaves with gradient sinks and the same parameters in every view, so their share can wait
// until the step's last blend backward has run: this kernel then reads a Gaussian's parameters and SH row
// ONCE, walks the views (per view: the radius, the blend backward's 32-byte record, the clamp bits and the
// opacity scratch, 41 B), sums dL/dcov3D, dL/dmean and
// dL/dopacity in registers and dL/dsh in a shared-memory row, and adds each sum to its sink once: ~120 B per
// Gaussian and view at 8 views.  Same arithmetic as the per-view kernel, view by view; only the order of the
// additions differs.
// What a deferred view still owes its caller at once: the screen-space gradient it reports (means2D.grad,
// the densification statistic of hugs/trainer) -- the packed record's two mean2D words, zero for culled rows.
__global__ void __launch_bounds__(256) unpack_mean2d_kernel(int P, const int* __restrict__ radii,
                                                            const float* __restrict__ grad_pack,
                                                            float* __restrict__ dL_dmean2D) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    float x = 0.f, y = 0.f;
    if (__ldg(radii + idx) > 0) {
        x = __ldg(grad_pack + 8 * (size_t)idx + 3);
        y = __ldg(grad_pack + 8 * (size_t)idx + 4);
    }
    dL_dmean2D[3 * (size_t)idx + 0] = x;
    dL_dmean2D[3 * (size_t)idx + 1] = y;
    dL_dmean2D[3 * (size_t)idx + 2] = 0.f;
}

constexpr int kMaxViews = 16;
struct SceneViewRec {
    const float* view;
    const float* proj;
    const float* campos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    const int* radii;            // the scene rows of the view's radii
    const float* grad_pack;      // [P,8], scene rows of the view's geometry workspace
    const uint8_t* clamped;      // [P]
    const float* opac_scratch;   // [P]
};
struct SceneViewsArgs {
    int P, D, M, V;
    const float* means3D;
    const float* shs;
    const float* scales;
    const float* rotations;
    float scale_modifier;
    float* g_mean;    // sinks: [P,3]
    float* g_sh;      //        [P,16,3]
    float* g_scale;   //        [P,3]
    float* g_rot;     //        [P,4]
    float* g_opac;    //        [P]
    SceneViewRec v[kMaxViews];
};
constexpr int kSvSmemBytes = 2 * kPbThreads * kPbRowBytes;   // SH rows | dL/dsh accumulation rows

__global__ void __launch_bounds__(kPbThreads, 4) scene_backward_views_kernel(const SceneViewsArgs a) {
    extern __shared__ __align__(16) unsigned char s_rows[];
    __shared__ float s_cam[kMaxViews][36];
    for (int i = (int)threadIdx.x; i < a.V * 35; i += kPbThreads) {
        const int v = i / 35, j = i - 35 * v;
        s_cam[v][j] = j < 16 ? __ldg(a.v[v].view + j) : (j < 32 ? __ldg(a.v[v].proj + j - 16) : __ldg(a.v[v].campos + j - 32));
    }
    const int idx = blockIdx.x * kPbThreads + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const bool valid = idx < a.P;
    unsigned mask = 0;   // bit v: visible in view v
    if (valid)
        for (int v = 0; v < a.V; v++) mask |= (__ldg(a.v[v].radii + idx) > 0 ? 1u : 0u) << v;
    const int nf = 3 * (a.D + 1) * (a.D + 1);
    const int nv = (nf + 3) >> 2;
    const uint32_t s_warp = (uint32_t)__cvta_generic_to_shared(s_rows) + warp * 32u * kPbRowBytes;
    const uint32_t s_acc_warp = s_warp + kPbThreads * kPbRowBytes;
    const size_t warp_row0 = (size_t)blockIdx.x * kPbThreads + warp * 32u;
    const unsigned alive = __ballot_sync(0xffffffffu, mask != 0u);
    if (alive != 0u) {
        const float4* g_warp = reinterpret_cast<const float4*>(a.shs) + 12 * warp_row0;
        unsigned row = lane / 12u, col = lane % 12u;  // 16-byte word 32 j + lane = (row, col)
#pragma unroll
        for (int j = 0; j < 12; j++) {
            if (((alive >> row) & 1u) && (int)col < nv)
                cp_async16(s_warp + row * kPbRowBytes + col * 16u, g_warp + 32 * j + lane);
            col += 8u;
            row += 2u;
            if (col >= 12u) {
                col -= 12u;
                row += 1u;
            }
        }
    }
    const uint32_t mine = s_warp + lane * kPbRowBytes, mine_acc = s_acc_warp + lane * kPbRowBytes;
#pragma unroll
    for (int i = 0; i < 12; i++) sts128(mine_acc + 16u * i, 0.f, 0.f, 0.f, 0.f);
    __syncthreads();  // s_cam

    float mx = 0.f, my = 0.f, mz = 0.f, sx = 0.f, sy = 0.f, sz = 0.f, qr = 0.f, qx = 0.f, qy = 0.f, qz = 0.f;
    float R[3][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    float cov3[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (mask != 0u) {
        mx = __ldg(a.means3D + 3 * (size_t)idx + 0);
        my = __ldg(a.means3D + 3 * (size_t)idx + 1);
        mz = __ldg(a.means3D + 3 * (size_t)idx + 2);
        sx = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 0);
        sy = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 1);
        sz = a.scale_modifier * __ldg(a.scales + 3 * (size_t)idx + 2);
        qr = __ldg(a.rotations + 4 * (size_t)idx + 0);
        qx = __ldg(a.rotations + 4 * (size_t)idx + 1);
        qy = __ldg(a.rotations + 4 * (size_t)idx + 2);
        qz = __ldg(a.rotations + 4 * (size_t)idx + 3);
        R[0][0] = 1.f - 2.f * (qy * qy + qz * qz);
        R[0][1] = 2.f * (qx * qy - qr * qz);
        R[0][2] = 2.f * (qx * qz + qr * qy);
        R[1][0] = 2.f * (qx * qy + qr * qz);
        R[1][1] = 1.f - 2.f * (qx * qx + qz * qz);
        R[1][2] = 2.f * (qy * qz - qr * qx);
        R[2][0] = 2.f * (qx * qz - qr * qy);
        R[2][1] = 2.f * (qy * qz + qr * qx);
        R[2][2] = 1.f - 2.f * (qx * qx + qy * qy);
        const float s[3] = {sx, sy, sz};
        float Mm[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) Mm[c][r] = s[r] * R[c][r];
        cov3[0] = Mm[0][0] * Mm[0][0] + Mm[0][1] * Mm[0][1] + Mm[0][2] * Mm[0][2];
        cov3[1] = Mm[0][0] * Mm[1][0] + Mm[0][1] * Mm[1][1] + Mm[0][2] * Mm[1][2];
        cov3[2] = Mm[0][0] * Mm[2][0] + Mm[0][1] * Mm[2][1] + Mm[0][2] * Mm[2][2];
        cov3[3] = Mm[1][0] * Mm[1][0] + Mm[1][1] * Mm[1][1] + Mm[1][2] * Mm[1][2];
        cov3[4] = Mm[1][0] * Mm[2][0] + Mm[1][1] * Mm[2][1] + Mm[1][2] * Mm[2][2];
        cov3[5] = Mm[2][0] * Mm[2][0] + Mm[2][1] * Mm[2][1] + Mm[2][2] * Mm[2][2];
    }
    cp_async_wait_all();
    __syncwarp();

    float dmean_x = 0.f, dmean_y = 0.f, dmean_z = 0.f, dopac = 0.f;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int v = 0; v < a.V; v++) {
        const SceneViewRec& w = a.v[v];
        if (!((mask >> v) & 1u)) continue;   // culled in this view (or not a Gaussian at all)
        const float* V = s_cam[v];
        const float* Pm = s_cam[v] + 16;
        const float4 pk0 = __ldg(reinterpret_cast<const float4*>(w.grad_pack) + 2 * (size_t)idx);
        const float4 pk1 = __ldg(reinterpret_cast<const float4*>(w.grad_pack) + 2 * (size_t)idx + 1);
        const uint8_t cl = __ldg(w.clamped + idx);
        dopac += __ldg(w.opac_scratch + idx);
        const float g2x = pk0.w, g2y = pk1.x;
        float dRGB[3];
        dRGB[0] = (cl & 1) ? 0.f : pk0.x;
        dRGB[1] = (cl & 2) ? 0.f : pk0.y;
        dRGB[2] = (cl & 4) ? 0.f : pk0.z;

        // ---- cov2D backward (backward.cu:144-274), as in preprocess_backward_kernel ------------------
        const float dcon_x = pk1.y, dcon_y = pk1.z, dcon_z = pk1.w;
        float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
        float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
        const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
        const float limx = 1.3f * w.tan_fovx;
        const float limy = 1.3f * w.tan_fovy;
        const float txtz = tx / tz;
        const float tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float hx = w.focal_x, hy = w.focal_y;
        const float J00 = hx / tz, J02 = -(hx * tx) / (tz * tz);
        const float J11 = hy / tz, J12 = -(hy * ty) / (tz * tz);
        const float T00 = V[0] * J00 + V[2] * J02, T01 = V[4] * J00 + V[6] * J02, T02 = V[8] * J00 + V[10] * J02;
        const float T10 = V[1] * J11 + V[2] * J12, T11 = V[5] * J11 + V[6] * J12, T12 = V[9] * J11 + V[10] * J12;
        const float V00 = cov3[0], V01 = cov3[1], V02 = cov3[2], V11 = cov3[3], V12 = cov3[4], V22 = cov3[5];
        const float TV00 = T00 * V00 + T01 * V01 + T02 * V02;
        const float TV01 = T00 * V01 + T01 * V11 + T02 * V12;
        const float TV02 = T00 * V02 + T01 * V12 + T02 * V22;
        const float TV10 = T10 * V00 + T11 * V01 + T12 * V02;
        const float TV11 = T10 * V01 + T11 * V11 + T12 * V12;
        const float TV12 = T10 * V02 + T11 * V12 + T12 * V22;
        const float ca = TV00 * T00 + TV01 * T01 + TV02 * T02 + 0.3f;
        const float cb = TV10 * T00 + TV11 * T01 + TV12 * T02;
        const float cc = TV10 * T10 + TV11 * T11 + TV12 * T12 + 0.3f;
        const float denom = ca * cc - cb * cb;
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0.f) {
            dL_da = denom2inv * (-cc * cc * dcon_x + 2.f * cb * cc * dcon_y + (denom - ca * cc) * dcon_z);
            dL_dc = denom2inv * (-ca * ca * dcon_z + 2.f * ca * cb * dcon_y + (denom - ca * cc) * dcon_x);
            dL_db = denom2inv * 2.f * (cb * cc * dcon_x - (denom + 2.f * cb * cb) * dcon_y + ca * cb * dcon_z);
            dcov[0] += T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc;
            dcov[3] += T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc;
            dcov[5] += T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc;
            dcov[1] += 2.f * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2.f * T10 * T11 * dL_dc;
            dcov[2] += 2.f * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2.f * T10 * T12 * dL_dc;
            dcov[4] += 2.f * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2.f * T11 * T12 * dL_dc;
        }
        const float dL_dT00 = 2.f * TV00 * dL_da + TV10 * dL_db;
        const float dL_dT01 = 2.f * TV01 * dL_da + TV11 * dL_db;
        const float dL_dT02 = 2.f * TV02 * dL_da + TV12 * dL_db;
        const float dL_dT10 = 2.f * TV10 * dL_dc + TV00 * dL_db;
        const float dL_dT11 = 2.f * TV11 * dL_dc + TV01 * dL_db;
        const float dL_dT12 = 2.f * TV12 * dL_dc + TV02 * dL_db;
        const float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
        const float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
        const float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
        const float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
        const float itz = 1.f / tz;
        const float itz2 = itz * itz;
        const float itz3 = itz2 * itz;
        const float dL_dtx = x_grad_mul * -hx * itz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -hy * itz2 * dL_dJ12;
        const float dL_dtz = -hx * itz2 * dL_dJ00 - hy * itz2 * dL_dJ11 + (2.f * hx * tx) * itz3 * dL_dJ02 +
                             (2.f * hy * ty) * itz3 * dL_dJ12;
        float vm_x = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
        float vm_y = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
        float vm_z = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;
        {   // projection path (backward.cu:370-387)
            const float m_hw = Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15];
            const float m_w = 1.0f / (m_hw + 0.0000001f);
            const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
            const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
            vm_x += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
            vm_y += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
            vm_z += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
        }
        // ---- SH backward (backward.cu:20-139): the row is re-read from shared memory for every view ---
        {
            float sh[48], gsh[48];
#pragma unroll
            for (int i = 0; i < 12; i++) {
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < nv) t = lds128(mine + 16u * i);
                sh[4 * i + 0] = t.x;
                sh[4 * i + 1] = t.y;
                sh[4 * i + 2] = t.z;
                sh[4 * i + 3] = t.w;
            }
            sh_backward(a.D, sh, mx - V[32], my - V[33], mz - V[34], dRGB, gsh, vm_x, vm_y, vm_z);
#pragma unroll
            for (int i = 0; i < 12; i++) {
                const float4 t = lds128(mine_acc + 16u * i);
                sts128(mine_acc + 16u * i, t.x + gsh[4 * i], t.y + gsh[4 * i + 1], t.z + gsh[4 * i + 2], t.w + gsh[4 * i + 3]);
            }
        }
        dmean_x += vm_x;
        dmean_y += vm_y;
        dmean_z += vm_z;
    }
    __syncwarp();
    // dL/dsh: the accumulated rows of the warp's visible Gaussians, 512 contiguous bytes per warp instruction
    {
        float4* g_out = reinterpret_cast<float4*>(a.g_sh) + 12 * warp_row0;
        unsigned row = lane / 12u, col = lane % 12u;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            if ((alive >> row) & 1u) {
                const float4 t = lds128(s_acc_warp + row * kPbRowBytes + col * 16u);
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(g_out + 32 * j + lane), "f"(t.x),
                             "f"(t.y), "f"(t.z), "f"(t.w)
                             : "memory");
            }
            col += 8u;
            row += 2u;
            if (col >= 12u) {
                col -= 12u;
                row += 1u;
            }
        }
    }
    if (mask == 0u) return;
    red_f1(a.g_opac + idx, dopac);
    put3(a.g_mean + 3 * (size_t)idx, dmean_x, dmean_y, dmean_z, true);
    // ---- cov3D backward (backward.cu:278-341), once, on the views' summed dL/dcov3D --------------------
    {
        const float s[3] = {sx, sy, sz};
        float Mm[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) Mm[c][r] = s[r] * R[c][r];
        float dS[3][3];
        dS[0][0] = dcov[0];
        dS[0][1] = dS[1][0] = 0.5f * dcov[1];
        dS[0][2] = dS[2][0] = 0.5f * dcov[2];
        dS[1][1] = dcov[3];
        dS[1][2] = dS[2][1] = 0.5f * dcov[4];
        dS[2][2] = dcov[5];
        float dM[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++)
                dM[c][r] = 2.f * (Mm[0][r] * dS[c][0] + Mm[1][r] * dS[c][1] + Mm[2][r] * dS[c][2]);
        float dsc[3];
#pragma unroll
        for (int k = 0; k < 3; k++) dsc[k] = R[0][k] * dM[0][k] + R[1][k] * dM[1][k] + R[2][k] * dM[2][k];
        put3(a.g_scale + 3 * (size_t)idx, dsc[0], dsc[1], dsc[2], true);
        float Mt[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) Mt[c][r] = dM[r][c] * s[c];
        float4 dq;
        dq.x = 2.f * qz * (Mt[0][1] - Mt[1][0]) + 2.f * qy * (Mt[2][0] - Mt[0][2]) + 2.f * qx * (Mt[1][2] - Mt[2][1]);
        dq.y = 2.f * qy * (Mt[1][0] + Mt[0][1]) + 2.f * qz * (Mt[2][0] + Mt[0][2]) + 2.f * qr * (Mt[1][2] - Mt[2][1]) -
               4.f * qx * (Mt[2][2] + Mt[1][1]);
        dq.z = 2.f * qx * (Mt[1][0] + Mt[0][1]) + 2.f * qr * (Mt[2][0] - Mt[0][2]) + 2.f * qz * (Mt[1][2] + Mt[2][1]) -
               4.f * qy * (Mt[2][2] + Mt[0][0]);
        dq.w = 2.f * qr * (Mt[0][1] - Mt[1][0]) + 2.f * qx * (Mt[2][0] + Mt[0][2]) + 2.f * qy * (Mt[1][2] + Mt[2][1]) -
               4.f * qz * (Mt[1][1] + Mt[0][0]);
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(reinterpret_cast<float4*>(a.g_rot) + idx),
                     "f"(dq.x), "f"(dq.y), "f"(dq.z), "f"(dq.w)
                     : "memory");
    }
}
}  // namespace

int launch_preprocess_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs,
                               const uint8_t* clamped, const float* scales, const float* rotations,
                               float scale_modifier, const float* cov3D_precomp, const Camera& cam,
                               const float* grad_pack, float* dL_dmean2D, float* dL_dconic, float* dL_dmean3D,
                               float* dL_dcolor,
                               float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                               bool sh_accumulate, cudaStream_t stream, const PreBwdExtra* extra, SegBatch* batch) {
    static_assert(sizeof(PreBwdArgs) <= sizeof(SegBatch::args[0]), "SegBatch record too small");
    const float* grad_pack2 = extra ? extra->grad_pack2 : nullptr;
    const int n2 = extra ? extra->n2 : 0;
    PreBwdArgs a;
    a.P = P;
    a.D = D;
    a.M = M;
    a.means3D = means3D;
    a.radii = radii;
    a.shs = shs;
    a.clamped = clamped;
    a.scales = scales;
    a.rotations = rotations;
    a.scale_modifier = scale_modifier;
    a.cov3D_precomp = (scales != nullptr && rotations != nullptr) ? nullptr : cov3D_precomp;
    a.view = cam.view;
    a.proj = cam.proj;
    a.campos = cam.campos;
    a.tan_fovx = cam.tan_fovx;
    a.tan_fovy = cam.tan_fovy;
    a.focal_x = cam.focal_x;
    a.focal_y = cam.focal_y;
    a.grad_pack = grad_pack;
    a.grad_pack2 = n2 > 0 ? grad_pack2 : nullptr;
    a.n2 = n2;
    a.acc_mask = extra ? extra->acc_mask : 0;
    a.opac_scratch = extra ? extra->opac_scratch : nullptr;
    a.dL_dopacity = extra ? extra->dL_dopacity : nullptr;
    a.dL_dmean2D = dL_dmean2D;
    a.dL_dconic = dL_dconic;
    a.dL_dmean3D = dL_dmean3D;
    a.dL_dcolor = dL_dcolor;
    a.dL_dcov3D = dL_dcov3D;
    a.dL_dsh = dL_dsh;
    a.dL_dscale = dL_dscale;
    a.dL_drot = dL_drot;
    a.sh_accumulate = sh_accumulate ? 1 : 0;
    a.sh_vec4 = (M == 16 && (shs == nullptr || (reinterpret_cast<uintptr_t>(shs) & 15u) == 0) &&
                 (reinterpret_cast<uintptr_t>(dL_dsh) & 15u) == 0)
                    ? 1
                    : 0;
    if (a.cov3D_precomp == nullptr && (scales == nullptr || rotations == nullptr)) {
        set_error("preprocess_backward: need scales+rotations or cov3D_precomp");
        return -1;
    }
    {  // opt in to > 48 KB of dynamic shared memory, once per device
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(preprocess_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kPbSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    if (batch != nullptr) {   // a segment of a two-segment call: launched by flush_preprocess_backward
        if (batch->n >= 2) return -1;
        memcpy(batch->args[batch->n], &a, sizeof(a));
        batch->blocks[batch->n] = cdiv(P, kPbThreads);
        batch->smem[batch->n] = a.sh_vec4 ? kPbSmemBytes : 0;
        batch->n++;
        return 0;
    }
    preprocess_backward_kernel<<<cdiv(P, kPbThreads), kPbThreads, a.sh_vec4 ? kPbSmemBytes : 0, stream>>>(
        a, a, cdiv(P, kPbThreads));
    HB_LAUNCH_OK("preprocess_backward", stream, false);
    return 0;
}

int launch_unpack_mean2d(int P, const int* radii, const float* grad_pack, float* dL_dmean2D, cudaStream_t stream) {
    if (P <= 0) return 0;
    unpack_mean2d_kernel<<<cdiv(P, 256), 256, 0, stream>>>(P, radii, grad_pack, dL_dmean2D);
    HB_LAUNCH_OK("unpack_mean2d", stream, false);
    return 0;
}

int launch_scene_backward_views(int P, int D, int M, int n_views, const SceneViewHost* views, const float* means3D,
                                const float* shs, const float* scales, const float* rotations, float scale_modifier,
                                float* g_mean, float* g_sh, float* g_scale, float* g_rot, float* g_opac,
                                cudaStream_t stream) {
    if (P <= 0 || n_views <= 0) return 0;
    if (M != 16 || ((reinterpret_cast<uintptr_t>(shs) | reinterpret_cast<uintptr_t>(g_sh) |
                     reinterpret_cast<uintptr_t>(g_rot)) & 15u)) {
        set_error("scene_backward_views: needs M == 16 and 16-byte aligned shs / SH sink / rotation sink");
        return -3;
    }
    {  // opt in to > 48 KB of dynamic shared memory, once per device
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(scene_backward_views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kSvSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    for (int first = 0; first < n_views; first += kMaxViews) {
        SceneViewsArgs a;
        memset(&a, 0, sizeof(a));
        a.P = P;
        a.D = D;
        a.M = M;
        a.V = n_views - first < kMaxViews ? n_views - first : kMaxViews;
        a.means3D = means3D;
        a.shs = shs;
        a.scales = scales;
        a.rotations = rotations;
        a.scale_modifier = scale_modifier;
        a.g_mean = g_mean;
        a.g_sh = g_sh;
        a.g_scale = g_scale;
        a.g_rot = g_rot;
        a.g_opac = g_opac;
        for (int v = 0; v < a.V; v++) {
            const SceneViewHost& h = views[first + v];
            SceneViewRec& r = a.v[v];
            r.view = h.cam.view;
            r.proj = h.cam.proj;
            r.campos = h.cam.campos;
            r.tan_fovx = h.cam.tan_fovx;
            r.tan_fovy = h.cam.tan_fovy;
            r.focal_x = h.cam.focal_x;
            r.focal_y = h.cam.focal_y;
            r.radii = h.radii;
            r.grad_pack = h.grad_pack;
            r.clamped = h.clamped;
            r.opac_scratch = h.opac_scratch;
        }
        scene_backward_views_kernel<<<cdiv(P, kPbThreads), kPbThreads, kSvSmemBytes, stream>>>(a);
        HB_LAUNCH_OK("scene_backward_views", stream, false);
    }
    return 0;
}

int flush_preprocess_backward(SegBatch& b, cudaStream_t stream) {
    if (b.n == 0) return 0;
    PreBwdArgs a0, a1;
    memcpy(&a0, b.args[0], sizeof(a0));
    memcpy(&a1, b.args[b.n - 1], sizeof(a1));
    const int nb0 = b.blocks[0], nb = nb0 + (b.n == 2 ? b.blocks[1] : 0);
    const size_t smem = b.smem[0] > b.smem[b.n - 1] ? b.smem[0] : b.smem[b.n - 1];
    b.n = 0;
    preprocess_backward_kernel<<<nb, kPbThreads, smem, stream>>>(a0, a1, nb0);
    HB_LAUNCH_OK("preprocess_backward", stream, false);
    return 0;
}

}  // namespace hb
