/*
 * splat_oracle.c -- CPU restatement of the reference's splat hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may load
 * this file's shared object.  The product (ml-hugs_b200/) never does.
 *
 * Every function restates one reference function (paths relative to the apple/ml-hugs tree,
 * CR = submodules/diff-gaussian-rasterization/cuda_rasterizer) in plain C, fp32, one Gaussian /
 * pixel at a time, OpenMP across Gaussians / tiles.
 *
 * Rounding.  The reference is CUDA compiled with nvcc's default -fmad=true, so its sort keys,
 * tile rectangles and contributor counts depend on which multiplies were fused into FMAs.  The
 * fused sequence was read off the sm_100 SASS of the unmodified reference build
 * (oracle/build_ref.py; see DESIGN.md "rounding sequence") and is restated here with explicit
 * fmaf(); this file must be compiled with -ffp-contract=off so the compiler adds none of its own.
 * Division, reciprocal and sqrt are IEEE in both.  The one thing that cannot be restated
 * bit-for-bit is expf(): CUDA's uses the MUFU.EX2 hardware approximation.  glibc's expf differs
 * from it by <= 2 ulp, so alpha thresholds can flip on a handful of pixels per million; the
 * parity tests bound that (n_contrib mismatches <= 2e-5 of pixels vs this oracle, zero vs the
 * reference build on the GPU).
 *
 * Pinned against: tests/golden/ fixtures produced by the reference build itself on a B200
 * (tests/golden/make_golden_raster.py) and, for LBS, by the reference's Python on CPU
 * (tests/golden/make_golden_lbs.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* a0*b0 + a1*b1 + a2*b2 as nvcc contracts it in the reference build: fma(a2,b2, fma(a0,b0, a1*b1)). */
static inline float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return fmaf(a2, b2, fmaf(a0, b0, a1 * b1));
}

/* CUDA float->int conversion saturates and maps NaN to 0; C's cast is undefined there. */
static inline int f2i_trunc(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return 2147483647;
    if (v <= -2147483648.0f) return (int)(-2147483647 - 1);
    return (int)v;
}
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline unsigned umin(unsigned a, unsigned b) { return a < b ? a : b; }

/* getHigherMsb, CR/rasterizer_impl.cu:35-50 */
uint32_t oracle_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step;
        else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* getRect, CR/auxiliary.h:46-56 */
static inline void get_rect(float px, float py, int radius, unsigned gx, unsigned gy, unsigned* min_x,
                            unsigned* min_y, unsigned* max_x, unsigned* max_y) {
    const float r = (float)radius;
    *min_x = umin(gx, (unsigned)imax(0, f2i_trunc((px - r) * 0.0625f)));
    *min_y = umin(gy, (unsigned)imax(0, f2i_trunc((py - r) * 0.0625f)));
    *max_x = umin(gx, (unsigned)imax(0, f2i_trunc((((px + r) + 16.0f) - 1.0f) * 0.0625f)));
    *max_y = umin(gy, (unsigned)imax(0, f2i_trunc((((py + r) + 16.0f) - 1.0f) * 0.0625f)));
}

/* computeCov3D, CR/forward.cu:118-152 (quaternion (r,x,y,z), NOT normalised) */
static void cov3d(const float* scale, float mod, const float* rot, float* cov) {
    const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    const float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    const float xz_p_ry = fmaf(r, y, xz), xz_m_ry = fmaf(-r, y, xz);
    const float yz_m_rx = fmaf(y, z, -rx), yz_p_rx = fmaf(y, z, rx);
    const float xy_m_rz = fmaf(x, y, -rz), xy_p_rz = fmaf(x, y, rz);
    const float xx_yy = fmaf(x, x, yy), yy_zz = yy + zz, xx_zz = fmaf(x, x, zz);
    /* R[c][r] as glm::mat3(...) lays it out (column-major) */
    const float R00 = 1.f - (yy_zz + yy_zz), R01 = xy_m_rz + xy_m_rz, R02 = xz_p_ry + xz_p_ry;
    const float R10 = xy_p_rz + xy_p_rz, R11 = 1.f - (xx_zz + xx_zz), R12 = yz_m_rx + yz_m_rx;
    const float R20 = xz_m_ry + xz_m_ry, R21 = yz_p_rx + yz_p_rx, R22 = 1.f - (xx_yy + xx_yy);
    const float s0 = scale[0] * mod, s1 = scale[1] * mod, s2 = scale[2] * mod;
    const float M00 = s0 * R00, M01 = s1 * R01, M02 = s2 * R02;
    const float M10 = s0 * R10, M11 = s1 * R11, M12 = s2 * R12;
    const float M20 = s0 * R20, M21 = s1 * R21, M22 = s2 * R22;
    cov[0] = dot3(M00, M00, M01, M01, M02, M02);
    cov[1] = dot3(M00, M10, M01, M11, M02, M12);
    cov[2] = dot3(M00, M20, M01, M21, M02, M22);
    cov[3] = dot3(M10, M10, M11, M11, M12, M12);
    cov[4] = dot3(M10, M20, M11, M21, M12, M22);
    cov[5] = dot3(M20, M20, M21, M21, M22, M22);
}

/* preprocessCUDA forward, CR/forward.cu:155-256 (+ in_frustum CR/auxiliary.h:139-164,
 * computeCov2D CR/forward.cu:74-113, computeColorFromSH CR/forward.cu:20-71, ndc2Pix CR/auxiliary.h:41-44).
 * Outputs for culled Gaussians: radii = tiles_touched = 0, everything else untouched. */
void oracle_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                       const float* rotations, const float* opacities, const float* shs,
                       const float* cov3D_precomp, const float* colors_precomp, const float* V, const float* Pm,
                       const float* campos, int W, int H, float tan_fovx, float tan_fovy, int* radii,
                       float* means2D, float* depths, float* cov3D_out, float* conic_opacity, float* rgb,
                       uint8_t* clamped, uint32_t* tiles_touched) {
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
    const unsigned gx = (unsigned)((W + TILE - 1) / TILE), gy = (unsigned)((H + TILE - 1) / TILE);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        const float px = means3D[3 * idx], py = means3D[3 * idx + 1], pz = means3D[3 * idx + 2];
        const float depth = dot3(V[2], px, V[6], py, V[10], pz) + V[14];
        if (depth <= 0.2f) continue;
        const float hx = dot3(Pm[0], px, Pm[4], py, Pm[8], pz) + Pm[12];
        const float hy = dot3(Pm[1], px, Pm[5], py, Pm[9], pz) + Pm[13];
        const float hw = dot3(Pm[3], px, Pm[7], py, Pm[11], pz) + Pm[15];
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float proj_x = hx * p_w, proj_y = hy * p_w;

        float c3[6];
        if (cov3D_precomp) {
            memcpy(c3, cov3D_precomp + 6 * (size_t)idx, sizeof(c3));
        } else {
            cov3d(scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, c3);
            if (cov3D_out) memcpy(cov3D_out + 6 * (size_t)idx, c3, sizeof(c3));
        }
        /* computeCov2D */
        float tx = dot3(V[0], px, V[4], py, V[8], pz) + V[12];
        float ty = dot3(V[1], px, V[5], py, V[9], pz) + V[13];
        const float tz = depth;
        const float limx = tan_fovx * 1.3f, limy = tan_fovy * 1.3f;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float tz2 = tz * tz;
        const float J00 = focal_x / tz, J02 = -(focal_x * tx) / tz2;
        const float J11 = focal_y / tz, J12 = -(focal_y * ty) / tz2;
        const float T00 = fmaf(V[2], J02, V[0] * J00), T01 = fmaf(V[6], J02, V[4] * J00), T02 = fmaf(V[10], J02, V[8] * J00);
        const float T10 = fmaf(V[2], J12, V[1] * J11), T11 = fmaf(V[6], J12, V[5] * J11), T12 = fmaf(V[10], J12, V[9] * J11);
        const float X00 = dot3(T00, c3[0], T01, c3[1], T02, c3[2]);
        const float X10 = dot3(T00, c3[1], T01, c3[3], T02, c3[4]);
        const float X20 = dot3(T00, c3[2], T01, c3[4], T02, c3[5]);
        const float X01 = dot3(T10, c3[0], T11, c3[1], T12, c3[2]);
        const float X11 = dot3(T10, c3[1], T11, c3[3], T12, c3[4]);
        const float X21 = dot3(T10, c3[2], T11, c3[4], T12, c3[5]);
        const float cx = dot3(X00, T00, X10, T01, X20, T02) + 0.3f;
        const float cy = dot3(X01, T00, X11, T01, X21, T02);
        const float cz = dot3(X01, T10, X11, T11, X21, T12) + 0.3f;
        const float det = fmaf(cx, cz, -(cy * cy));
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float conic_x = cz * det_inv, conic_y = cy * -det_inv, conic_z = cx * det_inv;
        const float mid = (cx + cz) * 0.5f;
        const float sq = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        const float lambda1 = mid + sq, lambda2 = mid - sq;
        const float my_radius = ceilf(sqrtf(fmaxf(lambda1, lambda2)) * 3.f);
        const int radius = f2i_trunc(my_radius);
        const float pix_x = (float)(fma((double)proj_x + 1.0, (double)W, -1.0) * 0.5);
        const float pix_y = (float)(fma((double)proj_y + 1.0, (double)H, -1.0) * 0.5);
        unsigned x0, y0, x1, y1;
        get_rect(pix_x, pix_y, radius, gx, gy, &x0, &y0, &x1, &y1);
        const unsigned tiles = (x1 - x0) * (y1 - y0);
        if (tiles == 0) continue;

        if (!colors_precomp) {
            float dx = px - campos[0], dy = py - campos[1], dz = pz - campos[2];
            const float len = sqrtf(fmaf(dz, dz, fmaf(dx, dx, dy * dy)));
            dx /= len; dy /= len; dz /= len;
            const float* sh = shs + (size_t)idx * M * 3;
            for (int c = 0; c < 3; c++) {
                const float x = dx, y = dy, z = dz;
                float res = SH_C0 * sh[c];
                if (D > 0) {
                    res = res - SH_C1 * y * sh[3 + c] + SH_C1 * z * sh[6 + c] - SH_C1 * x * sh[9 + c];
                    if (D > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        res = res + SH_C2[0] * xy * sh[12 + c] + SH_C2[1] * yz * sh[15 + c] +
                              SH_C2[2] * (2.0f * zz - xx - yy) * sh[18 + c] + SH_C2[3] * xz * sh[21 + c] +
                              SH_C2[4] * (xx - yy) * sh[24 + c];
                        if (D > 2) {
                            res = res + SH_C3[0] * y * (3.0f * xx - yy) * sh[27 + c] + SH_C3[1] * xy * z * sh[30 + c] +
                                  SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[33 + c] +
                                  SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + c] +
                                  SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[39 + c] + SH_C3[5] * z * (xx - yy) * sh[42 + c] +
                                  SH_C3[6] * x * (xx - 3.0f * yy) * sh[45 + c];
                        }
                    }
                }
                res += 0.5f;
                clamped[3 * (size_t)idx + c] = (res < 0);
                rgb[3 * (size_t)idx + c] = fmaxf(res, 0.0f);
            }
        }
        depths[idx] = depth;
        radii[idx] = radius;
        means2D[2 * (size_t)idx] = pix_x;
        means2D[2 * (size_t)idx + 1] = pix_y;
        conic_opacity[4 * (size_t)idx + 0] = conic_x;
        conic_opacity[4 * (size_t)idx + 1] = conic_y;
        conic_opacity[4 * (size_t)idx + 2] = conic_z;
        conic_opacity[4 * (size_t)idx + 3] = opacities[idx];
        tiles_touched[idx] = tiles;
    }
}

/* checkFrustum, CR/rasterizer_impl.cu:54-66 */
void oracle_mark_visible(int P, const float* means3D, const float* V, uint8_t* present) {
    for (int i = 0; i < P; i++) {
        const float d = dot3(V[2], means3D[3 * i], V[6], means3D[3 * i + 1], V[10], means3D[3 * i + 2]) + V[14];
        present[i] = (d <= 0.2f) ? 0 : 1;
    }
}

/* cub::DeviceScan::InclusiveSum call, CR/rasterizer_impl.cu:277 */
uint32_t oracle_inclusive_scan(int P, const uint32_t* in, uint32_t* out) {
    uint32_t s = 0;
    for (int i = 0; i < P; i++) {
        s += in[i];
        out[i] = s;
    }
    return s;
}

/* duplicateWithKeys, CR/rasterizer_impl.cu:70-111 */
void oracle_duplicate(int P, const float* means2D, const float* depths, const uint32_t* offsets, const int* radii,
                      int W, int H, uint64_t* keys, uint32_t* vals) {
    const unsigned gx = (unsigned)((W + TILE - 1) / TILE), gy = (unsigned)((H + TILE - 1) / TILE);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int idx = 0; idx < P; idx++) {
        if (radii[idx] > 0) {
            uint32_t off = (idx == 0) ? 0 : offsets[idx - 1];
            unsigned x0, y0, x1, y1;
            get_rect(means2D[2 * (size_t)idx], means2D[2 * (size_t)idx + 1], radii[idx], gx, gy, &x0, &y0, &x1, &y1);
            uint32_t dbits;
            memcpy(&dbits, depths + idx, 4);
            for (unsigned y = y0; y < y1; y++)
                for (unsigned x = x0; x < x1; x++) {
                    uint64_t key = (uint64_t)(y * gx + x);
                    key <<= 32;
                    key |= dbits;
                    keys[off] = key;
                    vals[off] = (uint32_t)idx;
                    off++;
                }
        }
    }
}

/* cub::DeviceRadixSort::SortPairs(..., 0, end_bit), CR/rasterizer_impl.cu:303-308: a STABLE
 * ascending sort on key bits [0, end_bit).  Restated as a plain LSD counting sort, 16-bit digits. */
void oracle_sort_pairs(int64_t n, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                       uint32_t* vals_out, int end_bit) {
    if (n <= 0) return;
    uint64_t* kA = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
    uint32_t* vA = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    uint64_t* kB = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
    uint32_t* vB = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    size_t* count = (size_t*)malloc(sizeof(size_t) * 65537);
    memcpy(kA, keys_in, sizeof(uint64_t) * (size_t)n);
    memcpy(vA, vals_in, sizeof(uint32_t) * (size_t)n);
    for (int shift = 0; shift < end_bit; shift += 16) {
        const int bits = (end_bit - shift) < 16 ? (end_bit - shift) : 16;
        const uint64_t mask = ((uint64_t)1 << bits) - 1;
        memset(count, 0, sizeof(size_t) * 65537);
        for (int64_t i = 0; i < n; i++) count[((kA[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 65536; d++) count[d + 1] += count[d];
        for (int64_t i = 0; i < n; i++) {
            const size_t dst = count[(kA[i] >> shift) & mask]++;
            kB[dst] = kA[i];
            vB[dst] = vA[i];
        }
        uint64_t* tk = kA; kA = kB; kB = tk;
        uint32_t* tv = vA; vA = vB; vB = tv;
    }
    memcpy(keys_out, kA, sizeof(uint64_t) * (size_t)n);
    memcpy(vals_out, vA, sizeof(uint32_t) * (size_t)n);
    free(kA); free(vA); free(kB); free(vB); free(count);
}

/* cudaMemset + identifyTileRanges, CR/rasterizer_impl.cu:310,116-138 */
void oracle_tile_ranges(int64_t L, const uint64_t* keys, int num_tiles, uint32_t* ranges /*[tiles][2]*/) {
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)num_tiles);
    for (int64_t idx = 0; idx < L; idx++) {
        const uint32_t currtile = (uint32_t)(keys[idx] >> 32);
        if (idx == 0) ranges[2 * currtile] = 0;
        else {
            const uint32_t prevtile = (uint32_t)(keys[idx - 1] >> 32);
            if (currtile != prevtile) {
                ranges[2 * prevtile + 1] = (uint32_t)idx;
                ranges[2 * currtile] = (uint32_t)idx;
            }
        }
        if (idx == L - 1) ranges[2 * currtile + 1] = (uint32_t)L;
    }
}

/* renderCUDA forward, CR/forward.cu:261-374.  features is [P,3]. */
void oracle_render_forward(const uint32_t* ranges, const uint32_t* point_list, int W, int H, const float* means2D,
                           const float* features, const float* conic_opacity, float* final_T, uint32_t* n_contrib,
                           const float* bg, float* out_color) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                if (pxi >= W || pyi >= H) continue;
                const float pixfx = (float)pxi, pixfy = (float)pyi;
                float T = 1.0f, C[3] = {0, 0, 0};
                uint32_t contributor = 0, last_contributor = 0;
                for (uint32_t k = r0; k < r1; k++) {
                    contributor++;
                    const uint32_t id = point_list[k];
                    const float dx = means2D[2 * (size_t)id] - pixfx, dy = means2D[2 * (size_t)id + 1] - pixfy;
                    const float* co = conic_opacity + 4 * (size_t)id;
                    const float a = fmaf(dx, dx * co[0], dy * (dy * co[2]));
                    const float power = fmaf(a, -0.5f, -(dy * (dx * co[1])));
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break; /* done = true */
                    for (int ch = 0; ch < 3; ch++) C[ch] = fmaf(T, alpha * features[3 * (size_t)id + ch], C[ch]);
                    T = test_T;
                    last_contributor = contributor;
                }
                const size_t pix = (size_t)W * pyi + pxi;
                final_T[pix] = T;
                n_contrib[pix] = last_contributor;
                for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pix] = fmaf(bg[ch], T, C[ch]);
            }
    }
}

/* renderCUDA backward, CR/backward.cu:399-557.  Per-hit arithmetic in fp32 as in the reference;
 * the per-Gaussian sums (the reference's float atomicAdd, order undefined) are kept in double so
 * the oracle does not depend on summation order.  Outputs are [P,3],[P,4],[P],[P,3] and must be
 * zero on entry. */
void oracle_render_backward(const uint32_t* ranges, const uint32_t* point_list, int W, int H, const float* bg,
                            const float* means2D, const float* conic_opacity, const float* colors,
                            const float* final_Ts, const uint32_t* n_contrib, const float* dL_dpixels, int P,
                            float* dL_dmean2D, float* dL_dconic2D, float* dL_dopacity, float* dL_dcolors) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    double* acc = (double*)calloc((size_t)P * 9, sizeof(double));
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel
    {
        double* loc = NULL;
        size_t loc_cap = 0;
#pragma omp for schedule(dynamic, 4)
        for (int tile = 0; tile < gx * gy; tile++) {
            const int tx = tile % gx, ty = tile / gx;
            const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
            const size_t len = r1 - r0;
            if (len == 0) continue;
            if (len * 9 > loc_cap) {
                free(loc);
                loc_cap = len * 9;
                loc = (double*)malloc(loc_cap * sizeof(double));
            }
            memset(loc, 0, len * 9 * sizeof(double));
            for (int ly = 0; ly < TILE; ly++)
                for (int lx = 0; lx < TILE; lx++) {
                    const int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                    if (pxi >= W || pyi >= H) continue;
                    const size_t pix = (size_t)W * pyi + pxi;
                    const float pixfx = (float)pxi, pixfy = (float)pyi;
                    const float T_final = final_Ts[pix];
                    float T = T_final;
                    const uint32_t last_contributor = n_contrib[pix];
                    float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0, dLp[3];
                    for (int c = 0; c < 3; c++) dLp[c] = dL_dpixels[(size_t)c * H * W + pix];
                    float bg_dot = 0;
                    for (int c = 0; c < 3; c++) bg_dot += bg[c] * dLp[c];
                    for (int64_t pos = (int64_t)last_contributor - 1; pos >= 0; pos--) {
                        const uint32_t id = point_list[r0 + pos];
                        const float dx = means2D[2 * (size_t)id] - pixfx, dy = means2D[2 * (size_t)id + 1] - pixfy;
                        const float* co = conic_opacity + 4 * (size_t)id;
                        const float a = fmaf(dx, dx * co[0], dy * (dy * co[2]));
                        const float power = fmaf(a, -0.5f, -(dy * (dx * co[1])));
                        if (power > 0.0f) continue;
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, co[3] * G);
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.f - alpha);
                        const float dchannel_dcolor = alpha * T;
                        float dL_dalpha = 0.0f;
                        double* g = loc + 9 * (size_t)pos;
                        for (int ch = 0; ch < 3; ch++) {
                            const float c = colors[3 * (size_t)id + ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = c;
                            dL_dalpha += (c - accum_rec[ch]) * dLp[ch];
                            g[ch] += dchannel_dcolor * dLp[ch];
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                        const float dL_dG = co[3] * dL_dalpha;
                        const float gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                        const float dG_ddely = -gdy * co[2] - gdx * co[1];
                        g[3] += dL_dG * dG_ddelx * ddelx_dx;
                        g[4] += dL_dG * dG_ddely * ddely_dy;
                        g[5] += -0.5f * gdx * dx * dL_dG;
                        g[6] += -0.5f * gdx * dy * dL_dG;
                        g[7] += -0.5f * gdy * dy * dL_dG;
                        g[8] += G * dL_dalpha;
                    }
                }
            for (size_t k = 0; k < len; k++) {
                const uint32_t id = point_list[r0 + k];
                for (int v = 0; v < 9; v++) {
                    const double d = loc[9 * k + v];
                    if (d != 0.0) {
#pragma omp atomic
                        acc[9 * (size_t)id + v] += d;
                    }
                }
            }
        }
        free(loc);
    }
    for (int i = 0; i < P; i++) {
        const double* g = acc + 9 * (size_t)i;
        dL_dcolors[3 * (size_t)i + 0] = (float)g[0];
        dL_dcolors[3 * (size_t)i + 1] = (float)g[1];
        dL_dcolors[3 * (size_t)i + 2] = (float)g[2];
        dL_dmean2D[3 * (size_t)i + 0] = (float)g[3];
        dL_dmean2D[3 * (size_t)i + 1] = (float)g[4];
        dL_dconic2D[4 * (size_t)i + 0] = (float)g[5];
        dL_dconic2D[4 * (size_t)i + 1] = (float)g[6];
        dL_dconic2D[4 * (size_t)i + 3] = (float)g[7];
        dL_dopacity[i] = (float)g[8];
    }
    free(acc);
}

/* BACKWARD::preprocess = computeCov2DCUDA (CR/backward.cu:144-274) then preprocessCUDA
 * (CR/backward.cu:346-396) with computeColorFromSH (:20-139) and computeCov3D (:278-341).
 * Outputs must be zero on entry (the reference allocates them with torch::zeros,
 * rasterize_points.cu:151-159).  clamped is [P,3] bytes. */
void oracle_preprocess_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs,
                                const uint8_t* clamped, const float* scales, const float* rotations,
                                float scale_modifier, const float* cov3Ds, const float* V, const float* proj,
                                const float* campos, int W, int H, float tan_fovx, float tan_fovy,
                                const float* dL_dmean2D, const float* dL_dconics, float* dL_dmeans,
                                const float* dL_dcolor, float* dL_dcov, float* dL_dsh, float* dL_dscale,
                                float* dL_drot) {
    const float h_y = H / (2.0f * tan_fovy), h_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float* cov3D = cov3Ds + 6 * (size_t)idx;
        const float mx = means3D[3 * (size_t)idx], my = means3D[3 * (size_t)idx + 1], mz = means3D[3 * (size_t)idx + 2];
        const float dcx = dL_dconics[4 * (size_t)idx], dcy = dL_dconics[4 * (size_t)idx + 1], dcz = dL_dconics[4 * (size_t)idx + 3];
        float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
        float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
        const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
        const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        /* glm column-major: X[c][r] */
        float J[3][3] = {{h_x / tz, 0.f, -(h_x * tx) / (tz * tz)}, {0.f, h_y / tz, -(h_y * ty) / (tz * tz)}, {0, 0, 0}};
        float Wm[3][3] = {{V[0], V[4], V[8]}, {V[1], V[5], V[9]}, {V[2], V[6], V[10]}};
        float Vrk[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
        float T[3][3];
        for (int c = 0; c < 3; c++)
            for (int r = 0; r < 3; r++) T[c][r] = Wm[0][r] * J[c][0] + Wm[1][r] * J[c][1] + Wm[2][r] * J[c][2];
        /* cov2D = T^T Vrk^T T */
        float X[3][3], cov2D[3][3];
        for (int c = 0; c < 3; c++)
            for (int r = 0; r < 3; r++) X[c][r] = T[r][0] * Vrk[0][c] + T[r][1] * Vrk[1][c] + T[r][2] * Vrk[2][c];
        for (int c = 0; c < 3; c++)
            for (int r = 0; r < 3; r++) cov2D[c][r] = X[0][r] * T[c][0] + X[1][r] * T[c][1] + X[2][r] * T[c][2];
        const float a = cov2D[0][0] + 0.3f, b = cov2D[0][1], c = cov2D[1][1] + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = dL_dcov + 6 * (size_t)idx;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
            dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
            dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
            dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
            dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        } else {
            for (int i = 0; i < 6; i++) dcov[i] = 0;
        }
        float dL_dT[2][3];
        for (int k = 0; k < 3; k++) {
            dL_dT[0][k] = 2 * (T[0][0] * Vrk[k][0] + T[0][1] * Vrk[k][1] + T[0][2] * Vrk[k][2]) * dL_da +
                          (T[1][0] * Vrk[k][0] + T[1][1] * Vrk[k][1] + T[1][2] * Vrk[k][2]) * dL_db;
            dL_dT[1][k] = 2 * (T[1][0] * Vrk[k][0] + T[1][1] * Vrk[k][1] + T[1][2] * Vrk[k][2]) * dL_dc +
                          (T[0][0] * Vrk[k][0] + T[0][1] * Vrk[k][1] + T[0][2] * Vrk[k][2]) * dL_db;
        }
        const float dL_dJ00 = Wm[0][0] * dL_dT[0][0] + Wm[0][1] * dL_dT[0][1] + Wm[0][2] * dL_dT[0][2];
        const float dL_dJ02 = Wm[2][0] * dL_dT[0][0] + Wm[2][1] * dL_dT[0][1] + Wm[2][2] * dL_dT[0][2];
        const float dL_dJ11 = Wm[1][0] * dL_dT[1][0] + Wm[1][1] * dL_dT[1][1] + Wm[1][2] * dL_dT[1][2];
        const float dL_dJ12 = Wm[2][0] * dL_dT[1][0] + Wm[2][1] * dL_dT[1][1] + Wm[2][2] * dL_dT[1][2];
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dL_dtx = x_grad_mul * -h_x * itz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * itz2 * dL_dJ12;
        const float dL_dtz = -h_x * itz2 * dL_dJ00 - h_y * itz2 * dL_dJ11 + (2 * h_x * tx) * itz3 * dL_dJ02 + (2 * h_y * ty) * itz3 * dL_dJ12;
        float dmean[3];
        dmean[0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
        dmean[1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
        dmean[2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

        /* projection path, CR/backward.cu:370-387 */
        {
            const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
            const float m_w = 1.0f / (m_hw + 0.0000001f);
            const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
            const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
            const float gx_ = dL_dmean2D[3 * (size_t)idx], gy_ = dL_dmean2D[3 * (size_t)idx + 1];
            dmean[0] += (proj[0] * m_w - proj[3] * mul1) * gx_ + (proj[1] * m_w - proj[3] * mul2) * gy_;
            dmean[1] += (proj[4] * m_w - proj[7] * mul1) * gx_ + (proj[5] * m_w - proj[7] * mul2) * gy_;
            dmean[2] += (proj[8] * m_w - proj[11] * mul1) * gx_ + (proj[9] * m_w - proj[11] * mul2) * gy_;
        }

        /* computeColorFromSH backward, CR/backward.cu:20-139 */
        if (shs) {
            const float ox = mx - campos[0], oy = my - campos[1], oz = mz - campos[2];
            const float len = sqrtf(ox * ox + oy * oy + oz * oz);
            const float x = ox / len, y = oy / len, z = oz / len;
            const float* sh = shs + (size_t)idx * M * 3;
            float dRGB[3];
            for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[3 * (size_t)idx + ch] * (clamped[3 * (size_t)idx + ch] ? 0.f : 1.f);
            float* dsh = dL_dsh + (size_t)idx * M * 3;
            float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
#define SH(k, ch) sh[3 * (k) + (ch)]
#define DSH(k, v)                                          \
    for (int ch = 0; ch < 3; ch++) dsh[3 * (k) + ch] = (v) * dRGB[ch];
            DSH(0, SH_C0);
            if (D > 0) {
                DSH(1, -SH_C1 * y);
                DSH(2, SH_C1 * z);
                DSH(3, -SH_C1 * x);
                for (int ch = 0; ch < 3; ch++) {
                    dRGBdx[ch] = -SH_C1 * SH(3, ch);
                    dRGBdy[ch] = -SH_C1 * SH(1, ch);
                    dRGBdz[ch] = SH_C1 * SH(2, ch);
                }
                if (D > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    DSH(4, SH_C2[0] * xy);
                    DSH(5, SH_C2[1] * yz);
                    DSH(6, SH_C2[2] * (2.f * zz - xx - yy));
                    DSH(7, SH_C2[3] * xz);
                    DSH(8, SH_C2[4] * (xx - yy));
                    for (int ch = 0; ch < 3; ch++) {
                        dRGBdx[ch] += SH_C2[0] * y * SH(4, ch) + SH_C2[2] * 2.f * -x * SH(6, ch) + SH_C2[3] * z * SH(7, ch) + SH_C2[4] * 2.f * x * SH(8, ch);
                        dRGBdy[ch] += SH_C2[0] * x * SH(4, ch) + SH_C2[1] * z * SH(5, ch) + SH_C2[2] * 2.f * -y * SH(6, ch) + SH_C2[4] * 2.f * -y * SH(8, ch);
                        dRGBdz[ch] += SH_C2[1] * y * SH(5, ch) + SH_C2[2] * 2.f * 2.f * z * SH(6, ch) + SH_C2[3] * x * SH(7, ch);
                    }
                    if (D > 2) {
                        DSH(9, SH_C3[0] * y * (3.f * xx - yy));
                        DSH(10, SH_C3[1] * xy * z);
                        DSH(11, SH_C3[2] * y * (4.f * zz - xx - yy));
                        DSH(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                        DSH(13, SH_C3[4] * x * (4.f * zz - xx - yy));
                        DSH(14, SH_C3[5] * z * (xx - yy));
                        DSH(15, SH_C3[6] * x * (xx - 3.f * yy));
                        for (int ch = 0; ch < 3; ch++) {
                            dRGBdx[ch] += (SH_C3[0] * SH(9, ch) * 3.f * 2.f * xy + SH_C3[1] * SH(10, ch) * yz + SH_C3[2] * SH(11, ch) * -2.f * xy +
                                           SH_C3[3] * SH(12, ch) * -3.f * 2.f * xz + SH_C3[4] * SH(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                                           SH_C3[5] * SH(14, ch) * 2.f * xz + SH_C3[6] * SH(15, ch) * 3.f * (xx - yy));
                            dRGBdy[ch] += (SH_C3[0] * SH(9, ch) * 3.f * (xx - yy) + SH_C3[1] * SH(10, ch) * xz +
                                           SH_C3[2] * SH(11, ch) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12, ch) * -3.f * 2.f * yz +
                                           SH_C3[4] * SH(13, ch) * -2.f * xy + SH_C3[5] * SH(14, ch) * -2.f * yz +
                                           SH_C3[6] * SH(15, ch) * -3.f * 2.f * xy);
                            dRGBdz[ch] += (SH_C3[1] * SH(10, ch) * xy + SH_C3[2] * SH(11, ch) * 4.f * 2.f * yz +
                                           SH_C3[3] * SH(12, ch) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13, ch) * 4.f * 2.f * xz +
                                           SH_C3[5] * SH(14, ch) * (xx - yy));
                        }
                    }
                }
            }
#undef SH
#undef DSH
            float ddir[3] = {0, 0, 0};
            for (int ch = 0; ch < 3; ch++) {
                ddir[0] += dRGBdx[ch] * dRGB[ch];
                ddir[1] += dRGBdy[ch] * dRGB[ch];
                ddir[2] += dRGBdz[ch] * dRGB[ch];
            }
            /* dnormvdv, CR/auxiliary.h:107-117 */
            const float sum2 = ox * ox + oy * oy + oz * oz;
            const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
            dmean[0] += ((+sum2 - ox * ox) * ddir[0] - oy * ox * ddir[1] - oz * ox * ddir[2]) * invsum32;
            dmean[1] += (-ox * oy * ddir[0] + (sum2 - oy * oy) * ddir[1] - oz * oy * ddir[2]) * invsum32;
            dmean[2] += (-ox * oz * ddir[0] - oy * oz * ddir[1] + (sum2 - oz * oz) * ddir[2]) * invsum32;
        }
        for (int k = 0; k < 3; k++) dL_dmeans[3 * (size_t)idx + k] = dmean[k];

        /* computeCov3D backward, CR/backward.cu:278-341 */
        if (scales) {
            const float* q = rotations + 4 * (size_t)idx;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {scale_modifier * scales[3 * (size_t)idx], scale_modifier * scales[3 * (size_t)idx + 1],
                                scale_modifier * scales[3 * (size_t)idx + 2]};
            float Mm[3][3];
            for (int c2 = 0; c2 < 3; c2++)
                for (int r2 = 0; r2 < 3; r2++) Mm[c2][r2] = s[r2] * R[c2][r2]; /* M = S * R */
            float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                              {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                              {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
            float dM[3][3]; /* 2 * M * dL_dSigma */
            for (int c2 = 0; c2 < 3; c2++)
                for (int r2 = 0; r2 < 3; r2++)
                    dM[c2][r2] = 2.0f * (Mm[0][r2] * dS[c2][0] + Mm[1][r2] * dS[c2][1] + Mm[2][r2] * dS[c2][2]);
            /* Rt[c][r] = R[r][c]; dMt[c][r] = dM[r][c] */
            for (int k = 0; k < 3; k++)
                dL_dscale[3 * (size_t)idx + k] = R[0][k] * dM[0][k] + R[1][k] * dM[1][k] + R[2][k] * dM[2][k];
            float Mt[3][3];
            for (int c2 = 0; c2 < 3; c2++)
                for (int r2 = 0; r2 < 3; r2++) Mt[c2][r2] = dM[r2][c2] * s[c2];
            float* dq = dL_drot + 4 * (size_t)idx;
            dq[0] = 2 * z * (Mt[0][1] - Mt[1][0]) + 2 * y * (Mt[2][0] - Mt[0][2]) + 2 * x * (Mt[1][2] - Mt[2][1]);
            dq[1] = 2 * y * (Mt[1][0] + Mt[0][1]) + 2 * z * (Mt[2][0] + Mt[0][2]) + 2 * r * (Mt[1][2] - Mt[2][1]) - 4 * x * (Mt[2][2] + Mt[1][1]);
            dq[2] = 2 * x * (Mt[1][0] + Mt[0][1]) + 2 * r * (Mt[2][0] - Mt[0][2]) + 2 * z * (Mt[1][2] + Mt[2][1]) - 4 * y * (Mt[2][2] + Mt[0][0]);
            dq[3] = 2 * r * (Mt[0][1] - Mt[1][0]) + 2 * x * (Mt[2][0] + Mt[0][2]) + 2 * y * (Mt[1][2] + Mt[2][1]) - 4 * z * (Mt[1][1] + Mt[0][0]);
        }
    }
}

/* LBS skinning + deformation tail + matrix_to_quaternion:
 *   hugs/models/modules/lbs.py:59-72, hugs/models/hugs_trimlp.py:510-518,
 *   hugs/utils/rotations.py:94-102,105-156.  A is [J,16] row-major 4x4, rotmat [N,9] row-major. */
void oracle_lbs_forward(int N, int J, const float* A, const float* xyz, const float* weights, const float* rotmat,
                        const float* scales, const float* transl, float smpl_scale, float* out_xyz, float* out_rotq,
                        float* out_scales, float* out_T) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; i++) {
        float T[16];
        for (int k = 0; k < 16; k++) {
            float s = 0.f;
            for (int j = 0; j < J; j++) s += weights[(size_t)i * J + j] * A[16 * j + k];
            T[k] = s;
        }
        if (out_T) memcpy(out_T + 16 * (size_t)i, T, sizeof(T));
        const float x = xyz[3 * (size_t)i], y = xyz[3 * (size_t)i + 1], z = xyz[3 * (size_t)i + 2];
        for (int r = 0; r < 3; r++) {
            float v = T[4 * r] * x + T[4 * r + 1] * y + T[4 * r + 2] * z + T[4 * r + 3];
            v = v * smpl_scale;
            if (transl) v += transl[r];
            out_xyz[3 * (size_t)i + r] = v;
        }
        float m[9];
        const float* Rm = rotmat + 9 * (size_t)i;
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) m[3 * r + c] = T[4 * r] * Rm[c] + T[4 * r + 1] * Rm[3 + c] + T[4 * r + 2] * Rm[6 + c];
        const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6], m21 = m[7], m22 = m[8];
        const float xs[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22, 1.0f - m00 - m11 + m22};
        float qa[4];
        for (int k = 0; k < 4; k++) qa[k] = xs[k] > 0 ? sqrtf(xs[k]) : 0.f;
        const float cand[4][4] = {{qa[0] * qa[0], m21 - m12, m02 - m20, m10 - m01},
                                  {m21 - m12, qa[1] * qa[1], m10 + m01, m02 + m20},
                                  {m02 - m20, m10 + m01, qa[2] * qa[2], m12 + m21},
                                  {m10 - m01, m20 + m02, m21 + m12, qa[3] * qa[3]}};
        int best = 0;
        for (int k = 1; k < 4; k++)
            if (qa[k] > qa[best]) best = k;
        const float den = 2.0f * fmaxf(qa[best], 0.1f);
        for (int c = 0; c < 4; c++) out_rotq[4 * (size_t)i + c] = cand[best][c] / den;
        if (scales && out_scales)
            for (int c = 0; c < 3; c++) out_scales[3 * (size_t)i + c] = scales[3 * (size_t)i + c] * smpl_scale;
    }
}
