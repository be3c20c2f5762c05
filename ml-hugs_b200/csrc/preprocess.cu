// Per-Gaussian forward preprocess for sm_100a.
//
// Behaviour follows the reference's preprocessCUDA (cuda_rasterizer/forward.cu:155-256) with
// in_frustum (auxiliary.h:139-164), computeCov3D (forward.cu:118-152), computeCov2D (:74-113),
// computeColorFromSH (:20-71), ndc2Pix / getRect (auxiliary.h:41-56).
//
// Bit-exactness: depth bits (-> sort keys), radius / tile rectangle (-> key duplication, ranges),
// pixel centre and conic (-> per-pixel contributor counts) must be bit-identical to what the
// reference build produces.  nvcc contracts the reference's source expressions into a particular
// sequence of FMUL / FFMA / FADD (verified in the sm_100 SASS of the reference build, see
// DESIGN.md "rounding sequence"); that sequence is spelled out here with explicitly rounded
// intrinsics, which the compiler never re-associates or re-contracts:
//     a*b + c*d + e*f      ->  fma(e,f, fma(a,b, rn(c*d)))
//     a*b + c*d + e*f + g  ->  rn(fma(e,f, fma(a,b, rn(c*d))) + g)
// Division, reciprocal and square root are the IEEE round-to-nearest versions (the reference
// is built without -use_fast_math).
#include "common.cuh"

#include <cstring>
#include <cstdio>

namespace hb {

#define RN_MUL(a, b) __fmul_rn((a), (b))
#define RN_ADD(a, b) __fadd_rn((a), (b))
#define RN_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define RN_DIV(a, b) __fdiv_rn((a), (b))

// a0*b0 + a1*b1 + a2*b2 as the reference build rounds it.
__device__ __forceinline__ float dot3_rn(float a0, float b0, float a1, float b1, float a2, float b2) {
    return RN_FMA(a2, b2, RN_FMA(a0, b0, RN_MUL(a1, b1)));
}

struct PreprocessArgs {
    int P, D, M;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* opacities;
    const float* shs;
    const float* cov3D_precomp;
    const float* colors_precomp;
    const float* view;
    const float* proj;
    const float* campos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y;
    int* radii;
    float* depths;
    float2* means2D;
    float4* conic_opacity;
    float4* rgb;
    uint8_t* clamped;
    uint32_t* tiles_touched;
    uint2* rects;            // tile rectangle (x0 | x1 << 16, y0 | y1 << 16), zero for culled Gaussians
    uint32_t* st_touched;    // super-tiles (8x4 tiles) the rectangle overlaps
    uint32_t* pair_counts;   // [0] += sum of tiles_touched, [1] += sum of st_touched (zeroed by the caller)
    int prefiltered;
    int sh_vec4;  // shs is [P,16,3] and 16-byte aligned: use 128-bit loads
    int rot_vec4;
};

// SH -> RGB (forward.cu:20-71).  `sh` holds the first 3*(D+1)^2 floats of this Gaussian.
__device__ __forceinline__ void eval_sh(int deg, const float* sh, float x, float y, float z, float* out) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float r = HB_SH_C0 * sh[0 * 3 + c];
        if (deg > 0) {
            r = r - HB_SH_C1 * y * sh[1 * 3 + c] + HB_SH_C1 * z * sh[2 * 3 + c] - HB_SH_C1 * x * sh[3 * 3 + c];
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z;
                float xy = x * y, yz = y * z, xz = x * z;
                r = r + HB_SH_C2_0 * xy * sh[4 * 3 + c] + HB_SH_C2_1 * yz * sh[5 * 3 + c] +
                    HB_SH_C2_2 * (2.0f * zz - xx - yy) * sh[6 * 3 + c] + HB_SH_C2_3 * xz * sh[7 * 3 + c] +
                    HB_SH_C2_4 * (xx - yy) * sh[8 * 3 + c];
                if (deg > 2) {
                    r = r + HB_SH_C3_0 * y * (3.0f * xx - yy) * sh[9 * 3 + c] + HB_SH_C3_1 * xy * z * sh[10 * 3 + c] +
                        HB_SH_C3_2 * y * (4.0f * zz - xx - yy) * sh[11 * 3 + c] +
                        HB_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12 * 3 + c] +
                        HB_SH_C3_4 * x * (4.0f * zz - xx - yy) * sh[13 * 3 + c] +
                        HB_SH_C3_5 * z * (xx - yy) * sh[14 * 3 + c] + HB_SH_C3_6 * x * (xx - 3.0f * yy) * sh[15 * 3 + c];
                }
            }
        }
        out[c] = r + 0.5f;
    }
}

// 3D covariance from scale + (unnormalised) quaternion, forward.cu:118-152, with the rounding
// sequence of the reference build.  q = (r, x, y, z).
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float r, float x,
                                                     float y, float z, float* cov) {
    const float xz = RN_MUL(x, z), rx = RN_MUL(r, x), rz = RN_MUL(r, z);
    const float yy = RN_MUL(y, y), zz = RN_MUL(z, z);
    const float xz_p_ry = RN_FMA(r, y, xz);
    const float xz_m_ry = RN_FMA(-r, y, xz);
    const float yz_m_rx = RN_FMA(y, z, -rx);
    const float yz_p_rx = RN_FMA(y, z, rx);
    const float xy_m_rz = RN_FMA(x, y, -rz);
    const float xy_p_rz = RN_FMA(x, y, rz);
    const float xx_yy = RN_FMA(x, x, yy);
    const float yy_zz = RN_ADD(yy, zz);
    const float xx_zz = RN_FMA(x, x, zz);
    // R[c][r], column-major as glm::mat3(...) fills it.
    const float R00 = RN_ADD(1.f, -RN_ADD(yy_zz, yy_zz));
    const float R01 = RN_ADD(xy_m_rz, xy_m_rz);
    const float R02 = RN_ADD(xz_p_ry, xz_p_ry);
    const float R10 = RN_ADD(xy_p_rz, xy_p_rz);
    const float R11 = RN_ADD(1.f, -RN_ADD(xx_zz, xx_zz));
    const float R12 = RN_ADD(yz_m_rx, yz_m_rx);
    const float R20 = RN_ADD(xz_m_ry, xz_m_ry);
    const float R21 = RN_ADD(yz_p_rx, yz_p_rx);
    const float R22 = RN_ADD(1.f, -RN_ADD(xx_yy, xx_yy));
    const float s0 = RN_MUL(sx, mod), s1 = RN_MUL(sy, mod), s2 = RN_MUL(sz, mod);
    // M = S * R  ->  M[c][r] = s_r * R[c][r]
    const float M00 = RN_MUL(s0, R00), M01 = RN_MUL(s1, R01), M02 = RN_MUL(s2, R02);
    const float M10 = RN_MUL(s0, R10), M11 = RN_MUL(s1, R11), M12 = RN_MUL(s2, R12);
    const float M20 = RN_MUL(s0, R20), M21 = RN_MUL(s1, R21), M22 = RN_MUL(s2, R22);
    // Sigma = M^T M; Sigma[c][r] = M[r][0]M[c][0] + M[r][1]M[c][1] + M[r][2]M[c][2]
    cov[0] = dot3_rn(M00, M00, M01, M01, M02, M02);
    cov[1] = dot3_rn(M00, M10, M01, M11, M02, M12);
    cov[2] = dot3_rn(M00, M20, M01, M21, M02, M22);
    cov[3] = dot3_rn(M10, M10, M11, M11, M12, M12);
    cov[4] = dot3_rn(M10, M20, M11, M21, M12, M22);
    cov[5] = dot3_rn(M20, M20, M21, M21, M22, M22);
}

// SH rows staged in shared memory: 192 B per Gaussian padded to 208 B (52 words), so that the
// 128-bit reads of a quarter warp (one row per lane) fall in disjoint banks.
constexpr int kPreThreads = 128;
constexpr int kShRowBytes = 208;
constexpr int kPreSmemBytes = kPreThreads * kShRowBytes;

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// One thread per Gaussian.  Memory behaviour (the stage is HBM bound, 311 B per visible Gaussian):
//   * position, scale, quaternion and opacity are requested together, before anything is computed
//     (one DRAM round trip, not three);
//   * the 192-B SH rows -- 62 % of the bytes -- are fetched only for the Gaussians that survive the
//     culls, by the whole warp: the 32 rows of a warp are one contiguous 6 KB block, lane l of copy j
//     moves 16-byte word 32 j + l of it with cp.async (global -> shared, no registers, all twelve
//     copies of a lane in flight at once), so every request is 512 contiguous bytes minus the rows of
//     culled Gaussians, and each thread then reads its own row out of shared memory.
//
// Two input segments in ONE launch (hugs_b200_rasterize_forward_seg2): blocks [0, nb0) work on a0, the rest
// on a1 -- the second segment's outputs are the same arrays moved to its first row, so a segment is just a
// second argument record; with one segment a1 == a0 and nb0 == gridDim.x.
__global__ void __launch_bounds__(kPreThreads, 6)
    preprocess_kernel(const PreprocessArgs a0, const PreprocessArgs a1, const int nb0) {
    const bool second = (int)blockIdx.x >= nb0;
    const PreprocessArgs& a = second ? a1 : a0;
    const unsigned blk = second ? blockIdx.x - (unsigned)nb0 : blockIdx.x;
    extern __shared__ __align__(16) unsigned char s_rows[];  // [256][208 B] SH rows (vec4 path)
    __shared__ float s_cam[36];  // view[16], proj[16], campos[3]
    if (threadIdx.x < 16) s_cam[threadIdx.x] = __ldg(a.view + threadIdx.x);
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = __ldg(a.proj + threadIdx.x - 16);
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = __ldg(a.campos + threadIdx.x - 32);
    __syncthreads();
    const float* V = s_cam;
    const float* Pm = s_cam + 16;

    const int idx = blk * kPreThreads + threadIdx.x;
    const bool valid = idx < a.P;
    bool vis = false;
    float px = 0.f, py = 0.f, pz = 0.f, depth = 0.f, pix_x = 0.f, pix_y = 0.f;
    float conic_x = 0.f, conic_y = 0.f, conic_z = 0.f, opacity = 0.f;
    int radius = 0;
    unsigned tiles = 0;
    unsigned rx0 = 0, rx1 = 0, ry0 = 0, ry1 = 0;

    if (valid) do {
        // ---- every per-Gaussian input except the SH row, requested up front -----------------
        px = __ldg(a.means3D + 3 * (size_t)idx + 0);
        py = __ldg(a.means3D + 3 * (size_t)idx + 1);
        pz = __ldg(a.means3D + 3 * (size_t)idx + 2);
        float cov3[6];
        float sx = 0.f, sy = 0.f, sz = 0.f;
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool precomp = a.cov3D_precomp != nullptr;
        if (precomp) {
#pragma unroll
            for (int i = 0; i < 6; i++) cov3[i] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + i);
        } else {
            sx = __ldg(a.scales + 3 * (size_t)idx + 0);
            sy = __ldg(a.scales + 3 * (size_t)idx + 1);
            sz = __ldg(a.scales + 3 * (size_t)idx + 2);
            if (a.rot_vec4) {
                q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
            } else {
                q.x = __ldg(a.rotations + 4 * (size_t)idx + 0);
                q.y = __ldg(a.rotations + 4 * (size_t)idx + 1);
                q.z = __ldg(a.rotations + 4 * (size_t)idx + 2);
                q.w = __ldg(a.rotations + 4 * (size_t)idx + 3);
            }
        }
        opacity = __ldg(a.opacities + idx);

        // in_frustum (auxiliary.h:139-164): near-plane cull on view-space z only.
        depth = RN_ADD(dot3_rn(V[2], px, V[6], py, V[10], pz), V[14]);
        if (depth <= 0.2f) {
            if (a.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
            break;
        }

        // Projection (forward.cu:197-200).
        const float hx = RN_ADD(dot3_rn(Pm[0], px, Pm[4], py, Pm[8], pz), Pm[12]);
        const float hy = RN_ADD(dot3_rn(Pm[1], px, Pm[5], py, Pm[9], pz), Pm[13]);
        const float hw = RN_ADD(dot3_rn(Pm[3], px, Pm[7], py, Pm[11], pz), Pm[15]);
        const float p_w = __frcp_rn(RN_ADD(hw, 0.0000001f));
        const float proj_x = RN_MUL(hx, p_w);
        const float proj_y = RN_MUL(hy, p_w);

        // 3D covariance (forward.cu:204-213).
        if (!precomp) cov3d_from_scale_rot(sx, sy, sz, a.scale_modifier, q.x, q.y, q.z, q.w, cov3);

        // 2D covariance, EWA (forward.cu:74-113).
        const float tz = depth;
        float tx = RN_ADD(dot3_rn(V[0], px, V[4], py, V[8], pz), V[12]);
        float ty = RN_ADD(dot3_rn(V[1], px, V[5], py, V[9], pz), V[13]);
        const float limx = RN_MUL(a.tan_fovx, 1.3f);
        const float limy = RN_MUL(a.tan_fovy, 1.3f);
        const float txtz = RN_DIV(tx, tz);
        const float tytz = RN_DIV(ty, tz);
        tx = RN_MUL(fminf(limx, fmaxf(-limx, txtz)), tz);
        ty = RN_MUL(fminf(limy, fmaxf(-limy, tytz)), tz);
        const float tz2 = RN_MUL(tz, tz);
        const float J00 = RN_DIV(a.focal_x, tz);
        const float J02 = RN_DIV(-RN_MUL(a.focal_x, tx), tz2);
        const float J11 = RN_DIV(a.focal_y, tz);
        const float J12 = RN_DIV(-RN_MUL(a.focal_y, ty), tz2);
        // T = W * J with W[c][r] = view[c + 4r]; only the two non-zero columns of J matter.
        const float T00 = RN_FMA(V[2], J02, RN_MUL(V[0], J00));
        const float T01 = RN_FMA(V[6], J02, RN_MUL(V[4], J00));
        const float T02 = RN_FMA(V[10], J02, RN_MUL(V[8], J00));
        const float T10 = RN_FMA(V[2], J12, RN_MUL(V[1], J11));
        const float T11 = RN_FMA(V[6], J12, RN_MUL(V[5], J11));
        const float T12 = RN_FMA(V[10], J12, RN_MUL(V[9], J11));
        // X = T^T Vrk^T : X[c][r] = T[r][0]Vrk[0][c] + T[r][1]Vrk[1][c] + T[r][2]Vrk[2][c]
        const float X00 = dot3_rn(T00, cov3[0], T01, cov3[1], T02, cov3[2]);
        const float X10 = dot3_rn(T00, cov3[1], T01, cov3[3], T02, cov3[4]);
        const float X20 = dot3_rn(T00, cov3[2], T01, cov3[4], T02, cov3[5]);
        const float X01 = dot3_rn(T10, cov3[0], T11, cov3[1], T12, cov3[2]);
        const float X11 = dot3_rn(T10, cov3[1], T11, cov3[3], T12, cov3[4]);
        const float X21 = dot3_rn(T10, cov3[2], T11, cov3[4], T12, cov3[5]);
        // cov = X T : cov[c][r] = X[0][r]T[c][0] + X[1][r]T[c][1] + X[2][r]T[c][2]; low-pass +0.3
        const float cx = RN_ADD(dot3_rn(X00, T00, X10, T01, X20, T02), 0.3f);
        const float cy = dot3_rn(X01, T00, X11, T01, X21, T02);
        const float cz = RN_ADD(dot3_rn(X01, T10, X11, T11, X21, T12), 0.3f);

        // Invert (forward.cu:219-223).
        const float det = RN_FMA(cx, cz, -RN_MUL(cy, cy));
        if (det == 0.0f) break;
        const float det_inv = __frcp_rn(det);
        conic_x = RN_MUL(cz, det_inv);
        conic_y = RN_MUL(cy, -det_inv);
        conic_z = RN_MUL(cx, det_inv);

        // Screen-space extent (forward.cu:229-237).
        const float mid = RN_MUL(RN_ADD(cx, cz), 0.5f);
        const float sq = __fsqrt_rn(fmaxf(0.1f, RN_FMA(mid, mid, -det)));
        const float lambda1 = RN_ADD(mid, sq);
        const float lambda2 = RN_ADD(mid, -sq);
        const float radius_f = ceilf(RN_MUL(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.f));
        radius = (int)radius_f;
        // ndc2Pix is evaluated in double in the reference (auxiliary.h:41-44).
        pix_x = (float)__dmul_rn(__fma_rn(__dadd_rn((double)proj_x, 1.0), (double)a.W, -1.0), 0.5);
        pix_y = (float)__dmul_rn(__fma_rn(__dadd_rn((double)proj_y, 1.0), (double)a.H, -1.0), 0.5);
        // getRect (auxiliary.h:46-56): float maths on an int radius, truncation, clamp to the grid.
        const float rf = (float)radius;
        const unsigned gx = (unsigned)a.grid_x, gy = (unsigned)a.grid_y;
        const unsigned min_x = min(gx, (unsigned)max(0, (int)RN_MUL(RN_ADD(pix_x, -rf), 0.0625f)));
        const unsigned min_y = min(gy, (unsigned)max(0, (int)RN_MUL(RN_ADD(pix_y, -rf), 0.0625f)));
        const unsigned max_x =
            min(gx, (unsigned)max(0, (int)RN_MUL(RN_ADD(RN_ADD(RN_ADD(pix_x, rf), 16.f), -1.f), 0.0625f)));
        const unsigned max_y =
            min(gy, (unsigned)max(0, (int)RN_MUL(RN_ADD(RN_ADD(RN_ADD(pix_y, rf), 16.f), -1.f), 0.0625f)));
        tiles = (max_x - min_x) * (max_y - min_y);
        if (tiles == 0) break;
        rx0 = min_x;
        rx1 = max_x;
        ry0 = min_y;
        ry1 = max_y;
        vis = true;
    } while (0);

    if (valid && !vis) {
        a.radii[idx] = 0;
        a.tiles_touched[idx] = 0;
        a.depths[idx] = __uint_as_float(0xFFFFFFFFu);  // depth-sort key of a culled Gaussian: last
    }

    // Colour (forward.cu:241-247).
    if (a.colors_precomp == nullptr) {
        float sh[48];
        const int nf = 3 * (a.D + 1) * (a.D + 1);  // floats actually used
        if (a.sh_vec4) {
            const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
            const unsigned alive = __ballot_sync(0xffffffffu, vis);
            if (alive != 0u) {
                const int nv = (nf + 3) >> 2;
                const uint32_t s_warp = (uint32_t)__cvta_generic_to_shared(s_rows) + warp * 32u * kShRowBytes;
                const float4* g_warp =
                    reinterpret_cast<const float4*>(a.shs) + 12 * ((size_t)blk * kPreThreads + warp * 32u);
                unsigned row = lane / 12u, col = lane % 12u;  // 16-byte word 32 j + lane = (row, col)
#pragma unroll
                for (int j = 0; j < 12; j++) {
                    if (((alive >> row) & 1u) && (int)col < nv)
                        cp_async16(s_warp + row * kShRowBytes + col * 16u, g_warp + 32 * j + lane);
                    col += 8u;
                    row += 2u;
                    if (col >= 12u) {
                        col -= 12u;
                        row += 1u;
                    }
                }
                cp_async_wait_all();
                __syncwarp();
                if (vis) {
                    const uint32_t mine = s_warp + lane * kShRowBytes;
#pragma unroll
                    for (int i = 0; i < 12; i++) {
                        if (i < nv) {
                            float4 v;
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                                         : "r"(mine + 16u * i));
                            sh[4 * i + 0] = v.x;
                            sh[4 * i + 1] = v.y;
                            sh[4 * i + 2] = v.z;
                            sh[4 * i + 3] = v.w;
                        }
                    }
                }
            }
        } else if (vis) {
            const float* src = a.shs + (size_t)idx * a.M * 3;
#pragma unroll
            for (int i = 0; i < 48; i++)
                if (i < nf) sh[i] = __ldg(src + i);
        }
        if (vis) {
            float dx = px - V[32], dy = py - V[33], dz = pz - V[34];
            const float len = sqrtf(dx * dx + dy * dy + dz * dz);
            dx = dx / len;
            dy = dy / len;
            dz = dz / len;
            float c[3];
            eval_sh(a.D, sh, dx, dy, dz, c);
            uint8_t mask = 0;
            if (c[0] < 0.f) mask |= 1;
            if (c[1] < 0.f) mask |= 2;
            if (c[2] < 0.f) mask |= 4;
            a.clamped[idx] = mask;
            a.rgb[idx] = make_float4(fmaxf(c[0], 0.f), fmaxf(c[1], 0.f), fmaxf(c[2], 0.f), 0.f);
        }
    }

    if (vis) {
        a.depths[idx] = depth;
        a.radii[idx] = radius;
        a.means2D[idx] = make_float2(pix_x, pix_y);
        a.conic_opacity[idx] = make_float4(conic_x, conic_y, conic_z, opacity);
        a.tiles_touched[idx] = tiles;
    }
    // Binning inputs (binning2.cu): the rectangle itself, the number of 8x4-tile super-tiles it overlaps,
    // and the two totals -- the pair count R the host needs, and the level-1 entry count.
    if (a.rects != nullptr) {
        unsigned st = 0;
        if (vis) st = (((rx1 + 7u) >> 3) - (rx0 >> 3)) * (((ry1 + 3u) >> 2) - (ry0 >> 2));
        if (valid) {
            a.rects[idx] = vis ? make_uint2(rx0 | (rx1 << 16), ry0 | (ry1 << 16)) : make_uint2(0u, 0u);
            a.st_touched[idx] = st;
        }
        const unsigned sum_t = __reduce_add_sync(0xffffffffu, vis ? tiles : 0u);
        const unsigned sum_s = __reduce_add_sync(0xffffffffu, st);
        if ((threadIdx.x & 31u) == 0u && sum_t != 0u) {
            atomicAdd(a.pair_counts + 0, sum_t);
            atomicAdd(a.pair_counts + 1, sum_s);
        }
    }
}

// =========================================================================================================
// The scene segment's preprocess for ALL views of a view-batched step in one launch
// (hugs_b200_preprocess_scene_views).  The per-view kernel above reads a Gaussian's parameters (44 B) and SH row
// (192 B) for every view and writes 75 B; the scene Gaussians are the same in every view, so this kernel reads
// them ONCE, builds the 3D covariance once, and walks the cameras: same arithmetic per (Gaussian, camera),
// operation for operation, hence bit-identical outputs.  The SH row is fetched (cp.async into shared memory)
// only if some view sees the Gaussian, and evaluated per view that does.
constexpr int kMaxPreViews = 16;
struct PreViewOut {
    const float* view;
    const float* proj;
    const float* campos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y;
    int* radii;
    float* depths;
    float2* means2D;
    float4* conic_opacity;
    float4* rgb;
    uint8_t* clamped;
    uint32_t* tiles_touched;
    uint2* rects;
    uint32_t* st_touched;
    uint32_t* pair_counts;
};
struct PreViewsArgs {
    int P, D, M, V;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* opacities;
    const float* shs;
    int prefiltered;
    int rot_vec4;
    PreViewOut v[kMaxPreViews];
};

__global__ void __launch_bounds__(kPreThreads, 5) preprocess_views_kernel(const PreViewsArgs a) {
    extern __shared__ __align__(16) unsigned char s_rows[];  // [128][208 B] SH rows
    __shared__ float s_cam[kMaxPreViews][36];
    for (int i = (int)threadIdx.x; i < a.V * 35; i += kPreThreads) {
        const int v = i / 35, j = i - 35 * v;
        s_cam[v][j] = j < 16 ? __ldg(a.v[v].view + j) : (j < 32 ? __ldg(a.v[v].proj + j - 16) : __ldg(a.v[v].campos + j - 32));
    }
    __syncthreads();
    const int idx = blockIdx.x * kPreThreads + threadIdx.x;
    const bool valid = idx < a.P;
    float px = 0.f, py = 0.f, pz = 0.f, opacity = 0.f;
    float cov3[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (valid) {
        px = __ldg(a.means3D + 3 * (size_t)idx + 0);
        py = __ldg(a.means3D + 3 * (size_t)idx + 1);
        pz = __ldg(a.means3D + 3 * (size_t)idx + 2);
        const float sx = __ldg(a.scales + 3 * (size_t)idx + 0);
        const float sy = __ldg(a.scales + 3 * (size_t)idx + 1);
        const float sz = __ldg(a.scales + 3 * (size_t)idx + 2);
        float4 q;
        if (a.rot_vec4) {
            q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        } else {
            q.x = __ldg(a.rotations + 4 * (size_t)idx + 0);
            q.y = __ldg(a.rotations + 4 * (size_t)idx + 1);
            q.z = __ldg(a.rotations + 4 * (size_t)idx + 2);
            q.w = __ldg(a.rotations + 4 * (size_t)idx + 3);
        }
        opacity = __ldg(a.opacities + idx);
        // 3D covariance (forward.cu:204-213): the same for every camera
        cov3d_from_scale_rot(sx, sy, sz, a.scale_modifier, q.x, q.y, q.z, q.w, cov3);
    }
    unsigned mask = 0;   // bit v: visible in view v
    for (int v = 0; v < a.V; v++) {
        const PreViewOut& w = a.v[v];
        const float* V = s_cam[v];
        const float* Pm = s_cam[v] + 16;
        bool vis = false;
        float depth = 0.f, pix_x = 0.f, pix_y = 0.f, conic_x = 0.f, conic_y = 0.f, conic_z = 0.f;
        int radius = 0;
        unsigned tiles = 0, rx0 = 0, rx1 = 0, ry0 = 0, ry1 = 0;
        if (valid) do {
            // in_frustum (auxiliary.h:139-164): near-plane cull on view-space z only.
            depth = RN_ADD(dot3_rn(V[2], px, V[6], py, V[10], pz), V[14]);
            if (depth <= 0.2f) {
                if (a.prefiltered) {
                    printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                    __trap();
                }
                break;
            }

            // Projection (forward.cu:197-200).
            const float hx = RN_ADD(dot3_rn(Pm[0], px, Pm[4], py, Pm[8], pz), Pm[12]);
            const float hy = RN_ADD(dot3_rn(Pm[1], px, Pm[5], py, Pm[9], pz), Pm[13]);
            const float hw = RN_ADD(dot3_rn(Pm[3], px, Pm[7], py, Pm[11], pz), Pm[15]);
            const float p_w = __frcp_rn(RN_ADD(hw, 0.0000001f));
            const float proj_x = RN_MUL(hx, p_w);
            const float proj_y = RN_MUL(hy, p_w);


            // 2D covariance, EWA (forward.cu:74-113).
            const float tz = depth;
            float tx = RN_ADD(dot3_rn(V[0], px, V[4], py, V[8], pz), V[12]);
            float ty = RN_ADD(dot3_rn(V[1], px, V[5], py, V[9], pz), V[13]);
            const float limx = RN_MUL(w.tan_fovx, 1.3f);
            const float limy = RN_MUL(w.tan_fovy, 1.3f);
            const float txtz = RN_DIV(tx, tz);
            const float tytz = RN_DIV(ty, tz);
            tx = RN_MUL(fminf(limx, fmaxf(-limx, txtz)), tz);
            ty = RN_MUL(fminf(limy, fmaxf(-limy, tytz)), tz);
            const float tz2 = RN_MUL(tz, tz);
            const float J00 = RN_DIV(w.focal_x, tz);
            const float J02 = RN_DIV(-RN_MUL(w.focal_x, tx), tz2);
            const float J11 = RN_DIV(w.focal_y, tz);
            const float J12 = RN_DIV(-RN_MUL(w.focal_y, ty), tz2);
            // T = W * J with W[c][r] = view[c + 4r]; only the two non-zero columns of J matter.
            const float T00 = RN_FMA(V[2], J02, RN_MUL(V[0], J00));
            const float T01 = RN_FMA(V[6], J02, RN_MUL(V[4], J00));
            const float T02 = RN_FMA(V[10], J02, RN_MUL(V[8], J00));
            const float T10 = RN_FMA(V[2], J12, RN_MUL(V[1], J11));
            const float T11 = RN_FMA(V[6], J12, RN_MUL(V[5], J11));
            const float T12 = RN_FMA(V[10], J12, RN_MUL(V[9], J11));
            // X = T^T Vrk^T : X[c][r] = T[r][0]Vrk[0][c] + T[r][1]Vrk[1][c] + T[r][2]Vrk[2][c]
            const float X00 = dot3_rn(T00, cov3[0], T01, cov3[1], T02, cov3[2]);
            const float X10 = dot3_rn(T00, cov3[1], T01, cov3[3], T02, cov3[4]);
            const float X20 = dot3_rn(T00, cov3[2], T01, cov3[4], T02, cov3[5]);
            const float X01 = dot3_rn(T10, cov3[0], T11, cov3[1], T12, cov3[2]);
            const float X11 = dot3_rn(T10, cov3[1], T11, cov3[3], T12, cov3[4]);
            const float X21 = dot3_rn(T10, cov3[2], T11, cov3[4], T12, cov3[5]);
            // cov = X T : cov[c][r] = X[0][r]T[c][0] + X[1][r]T[c][1] + X[2][r]T[c][2]; low-pass +0.3
            const float cx = RN_ADD(dot3_rn(X00, T00, X10, T01, X20, T02), 0.3f);
            const float cy = dot3_rn(X01, T00, X11, T01, X21, T02);
            const float cz = RN_ADD(dot3_rn(X01, T10, X11, T11, X21, T12), 0.3f);

            // Invert (forward.cu:219-223).
            const float det = RN_FMA(cx, cz, -RN_MUL(cy, cy));
            if (det == 0.0f) break;
            const float det_inv = __frcp_rn(det);
            conic_x = RN_MUL(cz, det_inv);
            conic_y = RN_MUL(cy, -det_inv);
            conic_z = RN_MUL(cx, det_inv);

            // Screen-space extent (forward.cu:229-237).
            const float mid = RN_MUL(RN_ADD(cx, cz), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(0.1f, RN_FMA(mid, mid, -det)));
            const float lambda1 = RN_ADD(mid, sq);
            const float lambda2 = RN_ADD(mid, -sq);
            const float radius_f = ceilf(RN_MUL(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.f));
            radius = (int)radius_f;
            // ndc2Pix is evaluated in double in the reference (auxiliary.h:41-44).
            pix_x = (float)__dmul_rn(__fma_rn(__dadd_rn((double)proj_x, 1.0), (double)w.W, -1.0), 0.5);
            pix_y = (float)__dmul_rn(__fma_rn(__dadd_rn((double)proj_y, 1.0), (double)w.H, -1.0), 0.5);
            // getRect (auxiliary.h:46-56): float maths on an int radius, truncation, clamp to the grid.
            const float rf = (float)radius;
            const unsigned gx = (unsigned)w.grid_x, gy = (unsigned)w.grid_y;
            const unsigned min_x = min(gx, (unsigned)max(0, (int)RN_MUL(RN_ADD(pix_x, -rf), 0.0625f)));
            const unsigned min_y = min(gy, (unsigned)max(0, (int)RN_MUL(RN_ADD(pix_y, -rf), 0.0625f)));
            const unsigned max_x =
                min(gx, (unsigned)max(0, (int)RN_MUL(RN_ADD(RN_ADD(RN_ADD(pix_x, rf), 16.f), -1.f), 0.0625f)));
            const unsigned max_y =
                min(gy, (unsigned)max(0, (int)RN_MUL(RN_ADD(RN_ADD(RN_ADD(pix_y, rf), 16.f), -1.f), 0.0625f)));
            tiles = (max_x - min_x) * (max_y - min_y);
            if (tiles == 0) break;
            rx0 = min_x;
            rx1 = max_x;
            ry0 = min_y;
            ry1 = max_y;
            vis = true;
        } while (0);
        if (valid && !vis) {
            w.radii[idx] = 0;
            w.tiles_touched[idx] = 0;
            w.depths[idx] = __uint_as_float(0xFFFFFFFFu);  // depth-sort key of a culled Gaussian: last
        }
        if (vis) {
            mask |= 1u << v;
            w.depths[idx] = depth;
            w.radii[idx] = radius;
            w.means2D[idx] = make_float2(pix_x, pix_y);
            w.conic_opacity[idx] = make_float4(conic_x, conic_y, conic_z, opacity);
            w.tiles_touched[idx] = tiles;
        }
        if (w.rects != nullptr) {
            unsigned st = 0;
            if (vis) st = (((rx1 + 7u) >> 3) - (rx0 >> 3)) * (((ry1 + 3u) >> 2) - (ry0 >> 2));
            if (valid) {
                w.rects[idx] = vis ? make_uint2(rx0 | (rx1 << 16), ry0 | (ry1 << 16)) : make_uint2(0u, 0u);
                w.st_touched[idx] = st;
            }
            const unsigned sum_t = __reduce_add_sync(0xffffffffu, vis ? tiles : 0u);
            const unsigned sum_s = __reduce_add_sync(0xffffffffu, st);
            if ((threadIdx.x & 31u) == 0u && sum_t != 0u) {
                atomicAdd(w.pair_counts + 0, sum_t);
                atomicAdd(w.pair_counts + 1, sum_s);
            }
        }
    }
    // Colour (forward.cu:241-247), per view that sees the Gaussian
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned alive = __ballot_sync(0xffffffffu, mask != 0u);
    if (alive == 0u) return;
    const int nf = 3 * (a.D + 1) * (a.D + 1);
    const int nv = (nf + 3) >> 2;
    const uint32_t s_warp = (uint32_t)__cvta_generic_to_shared(s_rows) + warp * 32u * kShRowBytes;
    {
        const float4* g_warp = reinterpret_cast<const float4*>(a.shs) + 12 * ((size_t)blockIdx.x * kPreThreads + warp * 32u);
        unsigned row = lane / 12u, col = lane % 12u;  // 16-byte word 32 j + lane = (row, col)
#pragma unroll
        for (int j = 0; j < 12; j++) {
            if (((alive >> row) & 1u) && (int)col < nv)
                cp_async16(s_warp + row * kShRowBytes + col * 16u, g_warp + 32 * j + lane);
            col += 8u;
            row += 2u;
            if (col >= 12u) {
                col -= 12u;
                row += 1u;
            }
        }
        cp_async_wait_all();
        __syncwarp();
    }
    if (mask == 0u) return;
    float sh[48];
    {
        const uint32_t mine = s_warp + lane * kShRowBytes;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < nv) {
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(mine + 16u * i));
            }
            sh[4 * i + 0] = t.x;
            sh[4 * i + 1] = t.y;
            sh[4 * i + 2] = t.z;
            sh[4 * i + 3] = t.w;
        }
    }
    for (int v = 0; v < a.V; v++) {
        if (!((mask >> v) & 1u)) continue;
        const PreViewOut& w = a.v[v];
        const float* V = s_cam[v];
        float dx = px - V[32], dy = py - V[33], dz = pz - V[34];
        const float len = sqrtf(dx * dx + dy * dy + dz * dz);
        dx = dx / len;
        dy = dy / len;
        dz = dz / len;
        float c[3];
        eval_sh(a.D, sh, dx, dy, dz, c);
        uint8_t m = 0;
        if (c[0] < 0.f) m |= 1;
        if (c[1] < 0.f) m |= 2;
        if (c[2] < 0.f) m |= 4;
        w.clamped[idx] = m;
        w.rgb[idx] = make_float4(fmaxf(c[0], 0.f), fmaxf(c[1], 0.f), fmaxf(c[2], 0.f), 0.f);
    }
}

int launch_preprocess_views(int P, int D, int M, int n_views, const PreViewHost* views, const float* means3D,
                            const float* scales, float scale_modifier, const float* rotations, const float* opacities,
                            const float* shs, bool prefiltered, cudaStream_t stream) {
    if (P <= 0 || n_views <= 0) return 0;
    if (M != 16 || (reinterpret_cast<uintptr_t>(shs) & 15u)) {
        set_error("preprocess_views: needs M == 16 and 16-byte aligned shs");
        return -3;
    }
    {  // opt in to > 48 KB of dynamic shared memory, once per device
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(preprocess_views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kPreSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    for (int first = 0; first < n_views; first += kMaxPreViews) {
        PreViewsArgs a;
        memset(&a, 0, sizeof(a));
        a.P = P;
        a.D = D;
        a.M = M;
        a.V = n_views - first < kMaxPreViews ? n_views - first : kMaxPreViews;
        a.means3D = means3D;
        a.scales = scales;
        a.scale_modifier = scale_modifier;
        a.rotations = rotations;
        a.opacities = opacities;
        a.shs = shs;
        a.prefiltered = prefiltered ? 1 : 0;
        a.rot_vec4 = (reinterpret_cast<uintptr_t>(rotations) & 15u) == 0 ? 1 : 0;
        for (int v = 0; v < a.V; v++) {
            const PreViewHost& h = views[first + v];
            PreViewOut& o = a.v[v];
            o.view = h.cam.view;
            o.proj = h.cam.proj;
            o.campos = h.cam.campos;
            o.W = h.cam.W;
            o.H = h.cam.H;
            o.tan_fovx = h.cam.tan_fovx;
            o.tan_fovy = h.cam.tan_fovy;
            o.focal_x = h.cam.focal_x;
            o.focal_y = h.cam.focal_y;
            o.grid_x = h.cam.grid_x;
            o.grid_y = h.cam.grid_y;
            o.radii = h.radii;
            o.depths = h.g.depths;
            o.means2D = h.g.means2D;
            o.conic_opacity = h.g.conic_opacity;
            o.rgb = h.g.rgb;
            o.clamped = h.g.clamped;
            o.tiles_touched = h.g.tiles_touched;
            o.rects = h.g.rects;
            o.st_touched = h.g.st_touched;
            o.pair_counts = h.g.pair_counts;
        }
        preprocess_views_kernel<<<cdiv(P, kPreThreads), kPreThreads, kPreSmemBytes, stream>>>(a);
        HB_LAUNCH_OK("preprocess_views", stream, false);
    }
    return 0;
}

int launch_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* opacities, const float* shs,
                      const float* cov3D_precomp, const float* colors_precomp, const Camera& cam, int* radii,
                      const GeomView& g, bool prefiltered, cudaStream_t stream, SegBatch* batch) {
    static_assert(sizeof(PreprocessArgs) <= sizeof(SegBatch::args[0]), "SegBatch record too small");
    PreprocessArgs a;
    a.P = P;
    a.D = D;
    a.M = M;
    a.means3D = means3D;
    a.scales = scales;
    a.scale_modifier = scale_modifier;
    a.rotations = rotations;
    a.opacities = opacities;
    a.shs = shs;
    a.cov3D_precomp = cov3D_precomp;
    a.colors_precomp = colors_precomp;
    a.view = cam.view;
    a.proj = cam.proj;
    a.campos = cam.campos;
    a.W = cam.W;
    a.H = cam.H;
    a.tan_fovx = cam.tan_fovx;
    a.tan_fovy = cam.tan_fovy;
    a.focal_x = cam.focal_x;
    a.focal_y = cam.focal_y;
    a.grid_x = cam.grid_x;
    a.grid_y = cam.grid_y;
    a.radii = radii;
    a.depths = g.depths;
    a.means2D = g.means2D;
    a.conic_opacity = g.conic_opacity;
    a.rgb = g.rgb;
    a.clamped = g.clamped;
    a.tiles_touched = g.tiles_touched;
    a.rects = g.rects;
    a.st_touched = g.st_touched;
    a.pair_counts = g.pair_counts;
    a.prefiltered = prefiltered ? 1 : 0;
    a.sh_vec4 = (shs != nullptr && M == 16 && (reinterpret_cast<uintptr_t>(shs) & 15u) == 0) ? 1 : 0;
    a.rot_vec4 = (rotations != nullptr && (reinterpret_cast<uintptr_t>(rotations) & 15u) == 0) ? 1 : 0;
    {  // opt in to > 48 KB of dynamic shared memory, once per device
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kPreSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    if (batch != nullptr) {   // a segment of a two-segment call: launched by flush_preprocess
        if (batch->n >= 2) return -1;
        memcpy(batch->args[batch->n], &a, sizeof(a));
        batch->blocks[batch->n] = cdiv(P, kPreThreads);
        batch->smem[batch->n] = a.sh_vec4 ? kPreSmemBytes : 0;
        batch->n++;
        return 0;
    }
    preprocess_kernel<<<cdiv(P, kPreThreads), kPreThreads, a.sh_vec4 ? kPreSmemBytes : 0, stream>>>(
        a, a, cdiv(P, kPreThreads));
    HB_LAUNCH_OK("preprocess", stream, false);
    return 0;
}

int flush_preprocess(SegBatch& b, cudaStream_t stream) {
    if (b.n == 0) return 0;
    PreprocessArgs a0, a1;
    memcpy(&a0, b.args[0], sizeof(a0));
    memcpy(&a1, b.args[b.n - 1], sizeof(a1));
    const int nb0 = b.blocks[0], nb = nb0 + (b.n == 2 ? b.blocks[1] : 0);
    const size_t smem = b.smem[0] > b.smem[b.n - 1] ? b.smem[0] : b.smem[b.n - 1];
    b.n = 0;
    preprocess_kernel<<<nb, kPreThreads, smem, stream>>>(a0, a1, nb0);
    HB_LAUNCH_OK("preprocess", stream, false);
    return 0;
}

// checkFrustum (rasterizer_impl.cu:54-66): present[i] = view-space z > 0.2.
__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ view,
                                                           uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float px = __ldg(means3D + 3 * (size_t)idx + 0);
    const float py = __ldg(means3D + 3 * (size_t)idx + 1);
    const float pz = __ldg(means3D + 3 * (size_t)idx + 2);
    const float depth = RN_ADD(dot3_rn(__ldg(view + 2), px, __ldg(view + 6), py, __ldg(view + 10), pz), __ldg(view + 14));
    present[idx] = (depth <= 0.2f) ? 0 : 1;
}

int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj, uint8_t* present,
                        cudaStream_t stream) {
    (void)proj;  // the reference computes p_proj but never tests it (auxiliary.h:154)
    mark_visible_kernel<<<cdiv(P, 256), 256, 0, stream>>>(P, means3D, view, present);
    HB_LAUNCH_OK("mark_visible", stream, false);
    return 0;
}

}  // namespace hb
