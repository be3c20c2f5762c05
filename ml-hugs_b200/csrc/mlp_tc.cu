// Fused decoder MLPs on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
// Replaces the three GELU MLPs of hugs/models/modules/decoders.py:24-110 (AppearanceDecoder,
// DeformationDecoder, GeometryDecoder), evaluated on the 110 210 x 96 tri-plane feature rows in
// hugs/models/hugs_trimlp.py:205-238 -- in the reference ~12 cuBLAS GEMMs + ~10 elementwise kernels.
//
// One CTA owns a tile of 128 rows and runs a whole decoder on it: the activations never leave the SM.
//   * GEMM:  D[128 x N] (TMEM, fp32) = A[128 x K] (shared) . W[N x K]^T (shared), tcgen05.mma
//     kind::tf32, M = 128, N = the layer's width (16..128), K = 8 per instruction, issued by ONE thread;
//   * 3xTF32: the reference runs these layers in fp32 (torch's allow_tf32 is off by default) and the
//     skinning logits go through softmax(x / 0.1) afterwards, so plain TF32 (2^-11) is not enough.
//     Every operand is split x = hi + lo (hi = x with its low 13 mantissa bits cleared, lo = x - hi,
//     exact) and D += A_hi W_hi + A_lo W_hi + A_hi W_lo: three MMAs per K step, error ~2^-21 per product;
//   * operands live in the canonical K-major SWIZZLE_128B layout (rows of 32 tf32 = 128 bytes, 8-row
//     groups of 1024 bytes, 16-byte chunk c of row r stored at chunk c ^ (r & 7)); one K block of 32 is one
//     16 KB tile.  The weights are split and laid out ONCE per call by mlp_pack_kernel (they are 0.4 MB and
//     stay in L2), so a CTA stages a weight block with cp.async, three stages deep: blocks j + 1 and j + 2
//     (possibly the next layer's) are in flight while block j is multiplied (tcgen05.commit -> mbarrier says
//     when a stage may be overwritten);
//   * epilogue: a thread owns a row (TMEM lane) and a quarter of the columns: tcgen05.ld 16 columns at a
//     time, + bias, GELU with erf to 1.5e-7 (torch.nn.GELU() is the erf form), then either hi/lo split
//     straight back into the A tiles of the next layer
//     or, for the heads, row-contiguous stores to the outputs; optionally the pre-activations are saved
//     for the backward pass.
// Shared memory: A_hi 64 KB + A_lo 64 KB + three weight stages of 32 KB = 224 KB -> one CTA per SM;
// TMEM: 128 columns.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kMlpThreads = 512;   // 16 warps: warps w, w + 4, w + 8, w + 12 share TMEM lanes 32 (w & 3).. and
                                   // interleave the 16-column chunks of an epilogue
constexpr int kColGroups = kMlpThreads / 128;
constexpr int kTileRows = 128;
constexpr int kKB = 32;                       // tf32 per 128-byte swizzle row = one K block
constexpr uint32_t kBlockBytes = 128 * 128;   // one K block of a 128-row operand tile
constexpr uint32_t kOffAhi = 0, kOffAlo = 4 * kBlockBytes, kOffW = 8 * kBlockBytes;   // W: [2 stages][hi, lo]
constexpr int kFwdStages = 3;                  // weight stages of the forward kernel (the backward one uses two)
constexpr uint32_t kOffBias = kOffW + 2 * kFwdStages * kBlockBytes;
constexpr uint32_t kOffBar = kOffBias + 512;
constexpr uint32_t kMlpSmemBytes = kOffBar + 64 + 1024;   // + slack for the 1024-byte alignment of the base
constexpr int kTmemCols = 128;

__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t chunk) {   // byte offset inside a K-block tile
    return (row >> 3) * 1024u + (row & 7u) * 128u + ((chunk ^ (row & 7u)) << 4);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void st_shared_u4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4,
// leading byte offset 1 (unused for swizzled K-major), stride byte offset 1024 >> 4 between 8-row groups,
// version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A and B tf32, both K-major, M = 128.
__device__ __forceinline__ uint32_t instr_desc(int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kTileRows >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
        "[%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
// erf by Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, far inside the 2e-5 the outputs are held to):
// branch-free, one reciprocal and one exponential -- libdevice's erff is a two-branch routine of ~45
// instructions, and the epilogues (one GELU per activation) were bound by it.  e2 = exp(-x^2) is returned too:
// the GELU derivative needs it.
__device__ __forceinline__ float erf_as(float x, float& e2) {
    const float ax = fabsf(x);
    const float t = __fdividef(1.0f, fmaf(0.3275911f, ax, 1.0f));
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    e2 = __expf(-ax * ax);
    return copysignf(1.0f - p * t * e2, x);
}
__device__ __forceinline__ float gelu_exact(float x) {
    float e2;
    return 0.5f * x * (1.0f + erf_as(x * 0.70710678118654752f, e2));
}
__device__ __forceinline__ float sigmoid_exact(float x) { return 1.0f / (1.0f + expf(-x)); }

}  // namespace

// ---- weight packing ----------------------------------------------------------------------------------
// W [n_true x K] row-major (torch.nn.Linear.weight) -> for every K block of 32: a hi tile and a lo tile of
// n_pad rows in the swizzled layout above (rows >= n_true are zero); bias -> n_pad floats.
// packed layout per layer: [K / 32][2 (hi, lo)][n_pad * 128 bytes], then the padded bias.
size_t mlp_packed_floats(int K, int n_pad) { return (size_t)(K / kKB) * 2 * n_pad * kKB + (size_t)n_pad; }

namespace {
// round-to-nearest-even to tf32 precision (the backward pass splits with this one: its single-product
// GEMMs then carry an unbiased 2^-12 error instead of truncation's biased 2^-11)
__device__ __forceinline__ float tf32_rn(float x) {
    const uint32_t u = __float_as_uint(x);
    return __uint_as_float((u + 0x0fffu + ((u >> 13) & 1u)) & 0xffffe000u);
}

// element (n, k) of the packed matrix is W[n][k], or W[k][n] with `transpose` (the backward pass multiplies
// by W^T: rows = the layer's inputs, K = its outputs, zero beyond src_rows / src_cols)
__global__ void __launch_bounds__(256) mlp_pack_kernel(const float* __restrict__ W, const float* __restrict__ bias,
                                                       int src_rows, int src_cols, int transpose, int n_pad, int K,
                                                       float* __restrict__ packed) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;   // element (n, k) of the padded matrix
    const int total = n_pad * K;
    if (e < total) {
        const int n = e / K, k = e % K;
        const int r = transpose ? k : n, c = transpose ? n : k;
        const float w = (r < src_rows && c < src_cols) ? __ldg(W + (size_t)r * src_cols + c) : 0.f;
        const float hi = transpose ? tf32_rn(w) : tf32_hi(w), lo = w - hi;
        const int kb = k / kKB, kk = k % kKB;
        const uint32_t off = swz((uint32_t)n, (uint32_t)(kk >> 2)) + (uint32_t)(kk & 3) * 4u;   // bytes
        float* tile_hi = packed + (size_t)kb * 2 * n_pad * kKB;
        float* tile_lo = tile_hi + (size_t)n_pad * kKB;
        tile_hi[off >> 2] = hi;
        tile_lo[off >> 2] = lo;
    }
    if (e < n_pad && !transpose)
        packed[(size_t)(K / kKB) * 2 * n_pad * kKB + e] = (e < src_rows && bias != nullptr) ? __ldg(bias + e) : 0.f;
}
}  // namespace

int launch_mlp_pack(const float* W, const float* bias, int n_true, int n_pad, int K, float* packed,
                    cudaStream_t stream) {
    if (K <= 0 || K % kKB != 0 || K > 128 || n_pad % 16 != 0 || n_pad < 16 || n_pad > 128 || n_true > n_pad) {
        set_error("mlp_pack: need K in {32,64,96,128}, n_pad a multiple of 16 in [16,128] (K=%d n_pad=%d)", K, n_pad);
        return -1;
    }
    const int total = n_pad * K;
    mlp_pack_kernel<<<cdiv(total, 256), 256, 0, stream>>>(W, bias, n_true, K, 0, n_pad, K, packed);
    HB_LAUNCH_OK("mlp_pack", stream, false);
    return 0;
}

// W [n_true x k_true] -> packed W^T: k_true rows (the layer's inputs; a multiple of 32), K = n_pad columns
// (the layer's outputs rounded up to 32).
int launch_mlp_pack_t(const float* W, int n_true, int k_true, int n_pad, float* packed, cudaStream_t stream) {
    if (k_true <= 0 || k_true % 16 != 0 || k_true > 128 || n_pad % kKB != 0 || n_pad <= 0 || n_pad > 128 || n_true > n_pad) {
        set_error("mlp_pack_t: bad shape (n_true=%d k_true=%d n_pad=%d)", n_true, k_true, n_pad);
        return -1;
    }
    const int total = k_true * n_pad;
    mlp_pack_kernel<<<cdiv(total, 256), 256, 0, stream>>>(W, nullptr, n_true, k_true, 1, k_true, n_pad, packed);
    HB_LAUNCH_OK("mlp_pack_t", stream, false);
    return 0;
}

// ---- the fused chain ---------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ float apply_act(float x, int act) {
    if (act == HB_MLP_ACT_GELU) return gelu_exact(x);
    if (act == HB_MLP_ACT_SIGMOID) return sigmoid_exact(x);
    return x;
}

__global__ void __launch_bounds__(kMlpThreads, 1) mlp_chain_kernel(const MlpArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const MlpChainArgs& ch = a.chain[blockIdx.y];
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t rloc = (warp & 3u) * 32u + (tid & 31u), half = warp >> 2;   // epilogue: row of the tile, column group
    const int row0 = (int)blockIdx.x * kTileRows;
    const int row = row0 + (int)rloc;         // the row this thread owns in every epilogue
    const bool row_ok = row < a.rows;
    const uint32_t tmem_slot = base + kOffBar + 32;

    if (tid == 0) {
        for (uint32_t b = 0; b < (uint32_t)kFwdStages; b++) mbar_init(base + kOffBar + 8u * b, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {   // one warp allocates the accumulator columns and gives the permit back
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tmem_slot) : "memory");

    // ---- the input rows -> A tiles (hi / lo), coalesced 16-byte chunks ------------------------------
    {
        const int chunks = a.Kx >> 2;   // per row
        const int total = kTileRows * chunks;
        for (int i0 = (int)tid; i0 < total; i0 += 8 * kMlpThreads) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {   // eight independent loads in flight
                const int i = i0 + u * kMlpThreads;
                const int r = i / chunks, c = i - r * chunks;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < total && row0 + r < a.rows)
                    v[u] = __ldg(reinterpret_cast<const float4*>(a.x + (size_t)(row0 + r) * a.ldx) + c);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = i0 + u * kMlpThreads;
                if (i >= total) break;
                const int r = i / chunks, c = i - r * chunks;
                const uint32_t off = (uint32_t)(c >> 3) * kBlockBytes + swz((uint32_t)r, (uint32_t)(c & 7));
                const float hx = tf32_hi(v[u].x), hy = tf32_hi(v[u].y), hz = tf32_hi(v[u].z), hw = tf32_hi(v[u].w);
                st_shared_v4(base + kOffAhi + off, hx, hy, hz, hw);
                st_shared_v4(base + kOffAlo + off, v[u].x - hx, v[u].y - hy, v[u].z - hz, v[u].w - hw);
            }
        }
    }

    // ---- weight pipeline: block j of the chain (all layers' K blocks in order) lives in stage j & 1; block
    // j + 1 is requested (cp.async) before block j is consumed, the first block of the next layer while the
    // current layer's epilogue runs.  ncommit[b]: commits issued on stage b (uniform over the CTA).
    int first_blk[HB_MLP_MAX_LAYERS + 1];
    first_blk[0] = 0;
    for (int l = 0; l < ch.n_layers; l++) first_blk[l + 1] = first_blk[l] + ch.layer[l].K / kKB;
    const int n_blk = first_blk[ch.n_layers];
    // (scalars: an array indexed by the stage would live in local memory)
    uint32_t nc0 = 0u, nc1 = 0u, nc2 = 0u;
    auto ncommit = [&](uint32_t b) { return b == 0u ? nc0 : (b == 1u ? nc1 : nc2); };
    auto bar_of = [&](uint32_t b) { return base + kOffBar + 8u * b; };
    int issued = 0;
    auto request = [&](int j) {   // copies of chain block j -> stage j % 3 (after the MMAs that last read it)
        int l = 0;
        while (j >= first_blk[l + 1]) l++;
        const MlpLayerArgs& Lj = ch.layer[l];
        const uint32_t b = (uint32_t)j % (uint32_t)kFwdStages;
        const uint32_t nc = ncommit(b);
        if (nc > 0u) mbar_wait(bar_of(b), (nc - 1u) & 1u);
        const uint32_t t16 = (uint32_t)Lj.N * 8u;   // 16-byte words of one hi (or lo) weight tile
        const uint4* src = reinterpret_cast<const uint4*>(Lj.packed + (size_t)(j - first_blk[l]) * 2 * Lj.N * kKB);
        const uint32_t w_hi = base + kOffW + b * 2u * kBlockBytes, w_lo = w_hi + kBlockBytes;
        for (uint32_t i = tid; i < t16; i += kMlpThreads) {
            cp_async16(w_hi + i * 16u, src + i);
            cp_async16(w_lo + i * 16u, src + t16 + i);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    request(0);
    issued = 1;
    if (n_blk > 1) {
        request(1);
        issued = 2;
    }
    for (int l = 0; l < ch.n_layers; l++) {
        const MlpLayerArgs& L = ch.layer[l];
        const int nkb = L.K / kKB, N = L.N;
        const uint32_t idesc = instr_desc(N);
        uint32_t last_b = 0;
        for (int kb = 0; kb < nkb; kb++) {
            const int j = first_blk[l] + kb;
            const uint32_t b = (uint32_t)j % (uint32_t)kFwdStages;
            last_b = b;
            if (issued <= j + 2 && j + 2 < n_blk) {   // two blocks ahead (possibly the next layer's) on their way
                request(j + 2);
                issued = j + 3;
            }
            // groups still allowed in flight: the blocks after j that have been requested
            const int ahead = issued - (j + 1);
            if (ahead >= 2) asm volatile("cp.async.wait_group 2;" ::: "memory");
            else if (ahead == 1) asm volatile("cp.async.wait_group 1;" ::: "memory");
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
            const uint32_t w_hi = base + kOffW + b * 2u * kBlockBytes, w_lo = w_hi + kBlockBytes;
            // generic-proxy writes (these and the epilogue's A tiles) -> visible to the tensor core's async proxy
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_hi = base + kOffAhi + (uint32_t)kb * kBlockBytes;
                const uint32_t a_lo = base + kOffAlo + (uint32_t)kb * kBlockBytes;
#pragma unroll
                for (uint32_t ks = 0; ks < 4; ks++) {   // K = 8 per instruction: 32 bytes along the swizzled row
                    const uint64_t dAh = smem_desc(a_hi + ks * 32u), dAl = smem_desc(a_lo + ks * 32u);
                    const uint64_t dBh = smem_desc(w_hi + ks * 32u), dBl = smem_desc(w_lo + ks * 32u);
                    umma_tf32(tmem, dAh, dBh, idesc, (kb > 0 || ks > 0) ? 1u : 0u);
                    umma_tf32(tmem, dAl, dBh, idesc, 1u);
                    umma_tf32(tmem, dAh, dBl, idesc, 1u);
                }
                umma_commit(bar_of(b));   // arrives when every MMA issued so far has completed
            }
            if (b == 0u) nc0++;
            else if (b == 1u) nc1++;
            else nc2++;
        }
        // the layer's accumulator is complete when the last commit has arrived
        mbar_wait(bar_of(last_b), (ncommit(last_b) - 1u) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        // ---- epilogue: thread t <-> row t <-> TMEM lane t -----------------------------------------
        const float* bias = L.packed + (size_t)nkb * 2 * N * kKB;
        const bool last = (l == ch.n_layers - 1);
        for (int c0 = 16 * (int)half; c0 < N; c0 += 16 * kColGroups) {
            float v[16];
            tmem_ld16(tmem + (((warp & 3u) * 32u) << 16) + (uint32_t)c0, v);
            {
                const float4* bp = reinterpret_cast<const float4*>(bias + c0);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float4 t = __ldg(bp + q);
                    v[4 * q] += t.x; v[4 * q + 1] += t.y; v[4 * q + 2] += t.z; v[4 * q + 3] += t.w;
                }
            }
            if (L.z_out != nullptr && row_ok) {   // pre-activations for the backward pass
                float4* z = reinterpret_cast<float4*>(L.z_out + (size_t)row * N + c0);
#pragma unroll
                for (int q = 0; q < 4; q++) z[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            }
            if (!last) {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    float h[4], lo[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const float y = apply_act(v[4 * q + j], L.act);
                        h[j] = tf32_hi(y);
                        lo[j] = y - h[j];
                    }
                    const int k = c0 + 4 * q;   // column of this layer = K index of the next
                    const uint32_t off = (uint32_t)(k >> 5) * kBlockBytes + swz(rloc, (uint32_t)((k & 31) >> 2));
                    st_shared_v4(base + kOffAhi + off, h[0], h[1], h[2], h[3]);
                    st_shared_v4(base + kOffAlo + off, lo[0], lo[1], lo[2], lo[3]);
                }
            } else if (row_ok) {
                for (int s = 0; s < ch.n_outs; s++) {
                    const MlpOut& o = ch.out[s];
                    float* dst = o.ptr + (size_t)row * o.ld - o.col0;
                    const bool vec = ((o.col0 | o.ncols | o.ld) & 3) == 0 && (reinterpret_cast<uintptr_t>(o.ptr) & 15u) == 0;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const int col = c0 + 4 * q;
                        if (vec && col >= o.col0 && col + 4 <= o.col0 + o.ncols) {   // whole 16-byte words of the row
                            *reinterpret_cast<float4*>(dst + col) =
                                make_float4(apply_act(v[4 * q], o.act), apply_act(v[4 * q + 1], o.act),
                                            apply_act(v[4 * q + 2], o.act), apply_act(v[4 * q + 3], o.act));
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                if (col + j >= o.col0 && col + j < o.col0 + o.ncols) dst[col + j] = apply_act(v[4 * q + j], o.act);
                        }
                    }
                }
            }
        }
        // the next layer's MMAs overwrite the accumulator and read the A tiles written above
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
    }
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
}
}  // namespace

int launch_mlp_chains(const MlpArgs& a, cudaStream_t stream) {
    if (a.rows <= 0 || a.n_chains <= 0) return 0;
    if (a.n_chains > HB_MLP_MAX_CHAINS || a.Kx % kKB != 0 || a.Kx <= 0 || a.Kx > 128 || a.ldx % 4 != 0 ||
        (reinterpret_cast<uintptr_t>(a.x) & 15u)) {
        set_error("mlp_chains: need 1..%d chains, input width in {32,64,96,128}, 16-byte aligned rows", HB_MLP_MAX_CHAINS);
        return -1;
    }
    for (int c = 0; c < a.n_chains; c++) {
        const MlpChainArgs& ch = a.chain[c];
        if (ch.n_layers < 1 || ch.n_layers > HB_MLP_MAX_LAYERS || ch.n_outs < 0 || ch.n_outs > HB_MLP_MAX_OUTS) {
            set_error("mlp_chains: chain %d has a bad layer / output count", c);
            return -1;
        }
        int K = a.Kx;
        for (int l = 0; l < ch.n_layers; l++) {
            const MlpLayerArgs& L = ch.layer[l];
            const bool hidden = l + 1 < ch.n_layers;
            if (L.K != K || L.N % 16 != 0 || L.N < 16 || L.N > 128 || (hidden && L.N % kKB != 0) || !L.packed) {
                set_error("mlp_chains: chain %d layer %d: K=%d (expected %d) N=%d", c, l, L.K, K, L.N);
                return -1;
            }
            K = L.N;
        }
    }
    {
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kMlpSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    const dim3 grid((unsigned)cdiv(a.rows, kTileRows), (unsigned)a.n_chains, 1);
    mlp_chain_kernel<<<grid, kMlpThreads, kMlpSmemBytes, stream>>>(a);
    HB_LAUNCH_OK("mlp_chains", stream, false);
    return 0;
}


// ---- backward ------------------------------------------------------------------------------------------
// Same tile ownership as the forward pass: a CTA takes 128 rows through the whole chain, top layer first.
// Per layer l (input width K, output width N, dZ = dL/d(pre-activation) of the layer for the tile):
//   dW_l^T [K x N] = A_{l-1}^T dZ     the contraction runs over the tile's 128 ROWS, so both operands are
//                                     needed with the row index contiguous: A^T and dZ^T are kept as
//                                     TRANSPOSED K-major tiles ([feature][row], written with conflict-free
//                                     4-byte stores: a warp's 32 rows of one feature are one 128-byte line).
//                                     (tcgen05 reads 32-bit operands MN-major only in the SWIZZLE_128B_BASE32B
//                                     layout, which the K-major tiles of the other product cannot share.)
//                                     One TF32 product per term with operands rounded to nearest (unbiased
//                                     2^-12; the sum over 110 k rows averages it out); the tile's result is
//                                     added to the global dW^T with 16-byte vector reductions;
//   dA_{l-1} [rows x K] = dZ W_l      dZ_hi W_hi + dZ_hi W_lo against the packed W^T (the weights are exact;
//                                     dZ is rounded to nearest tf32 -- there is no room for its low half
//                                     next to the transposed copy); the epilogue multiplies by
//                                     GELU'(z_{l-1}) (saved by the forward pass) and leaves dZ_{l-1} in both
//                                     tile sets, in place;
//   db_l = column sums of dZ          warp shuffles in the epilogue, one reduction per warp and column.
// Shared memory: dZ (64 KB) + dZ^T (64 KB) + one 64 KB region that holds A^T during the dW product and the two
// W^T stages during the dA product.  TMEM: 128 columns for dA + 128 for dW^T.
namespace {

constexpr uint32_t kOffR2 = 8 * kBlockBytes;
constexpr int kBwdTmemCols = 256;

__device__ __forceinline__ float gelu_grad(float z) {
    float e2;   // exp(-z^2 / 2), shared by the two terms
    const float erf_v = erf_as(z * 0.70710678118654752f, e2);
    return fmaf(z * 0.3989422804014327f, e2, 0.5f * (1.0f + erf_v));
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void st_shared_f1(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// byte offset of element (feature f, row r) in a transposed tile set: K blocks of 32 ROWS, tile rows = features
__device__ __forceinline__ uint32_t swz_t(uint32_t f, uint32_t r) {
    return (r >> 5) * kBlockBytes + swz(f, (r & 31u) >> 2) + (r & 3u) * 4u;
}
// dz values of row `r`, columns k..k+3 (rounded to tf32) -> the K-major tile and the transposed one
__device__ __forceinline__ void put_dz4(uint32_t base, uint32_t r, int k, const float* dz) {
    float h[4];
#pragma unroll
    for (int j = 0; j < 4; j++) h[j] = tf32_rn(dz[j]);
    const uint32_t off = (uint32_t)(k >> 5) * kBlockBytes + swz(r, (uint32_t)((k & 31) >> 2));
    st_shared_v4(base + kOffAhi + off, h[0], h[1], h[2], h[3]);
#pragma unroll
    for (int j = 0; j < 4; j++) st_shared_f1(base + kOffAlo + swz_t((uint32_t)(k + j), r), h[j]);
}
// column sums over the warp's 32 rows, added to db[0..15]
__device__ __forceinline__ void warp_colsum16(const float* v, float* db, unsigned lane) {
#pragma unroll
    for (int j = 0; j < 16; j++) {
        float s = v[j];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (lane == (unsigned)j) atomicAdd(db + j, s);
    }
}

__global__ void __launch_bounds__(kMlpThreads, 1) mlp_chain_backward_kernel(const MlpBwdArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t base = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const MlpBwdChainArgs& ch = a.chain[blockIdx.y];
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t rloc = (warp & 3u) * 32u + lane, half = warp >> 2;   // row of the tile / TMEM lane, column half
    const int row0 = (int)blockIdx.x * kTileRows;
    const int row = row0 + (int)rloc;
    const bool row_ok = row < a.rows;
    const uint32_t bar[3] = {base + kOffBar, base + kOffBar + 8, base + kOffBar + 16};   // W stages 0 / 1, dW product
    const uint32_t tmem_slot = base + kOffBar + 32;

    if (tid == 0) {
        mbar_init(bar[0], 1);
        mbar_init(bar[1], 1);
        mbar_init(bar[2], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(kBwdTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tmem_slot) : "memory");
    const uint32_t tmem_dA = tmem, tmem_dW = tmem + 128u;

    // ---- dZ of the top layer from the output gradients: thread t <-> row t ---------------------------
    {
        const MlpBwdLayerArgs& L = ch.layer[ch.n_layers - 1];
        for (int c0 = 16 * (int)half; c0 < L.N; c0 += 16 * kColGroups) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] = 0.f;
            if (row_ok) {
                float z[16];
                const float4* zp = reinterpret_cast<const float4*>(ch.z_head + (size_t)row * L.N + c0);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float4 t = __ldg(zp + q);
                    z[4 * q] = t.x; z[4 * q + 1] = t.y; z[4 * q + 2] = t.z; z[4 * q + 3] = t.w;
                }
                for (int s = 0; s < ch.n_outs; s++) {
                    const MlpBwdOut& o = ch.out[s];
                    if (o.dptr == nullptr) continue;
                    const float* src = o.dptr + (size_t)row * o.ld - o.col0;
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const int col = c0 + j;
                        if (col >= o.col0 && col < o.col0 + o.ncols) {
                            float g = __ldg(src + col);
                            if (o.act == HB_MLP_ACT_GELU) g *= gelu_grad(z[j]);
                            else if (o.act == HB_MLP_ACT_SIGMOID) {
                                const float y = sigmoid_exact(z[j]);
                                g *= y * (1.0f - y);
                            }
                            v[j] = g;
                        }
                    }
                }
            }
            warp_colsum16(v, L.db + c0, lane);
#pragma unroll
            for (int q = 0; q < 4; q++) put_dz4(base, rloc, c0 + 4 * q, v + 4 * q);
        }
    }

    uint32_t ncommit[3] = {0u, 0u, 0u};
    uint32_t stage = 0;
    for (int l = ch.n_layers - 1; l >= 0; l--) {
        const MlpBwdLayerArgs& L = ch.layer[l];
        const int K = L.K, N = L.N;
        // ---- (1) A_{l-1}^T (the layer's input activations, rounded to tf32, [feature][row]) -> region 2 ----
        {
            const int chunks = K >> 2;   // 16-byte chunks of the row; the row's kColGroups threads interleave them
            const float* src = (l == 0) ? a.x + (size_t)row * a.ldx : L.z_prev + (size_t)row * K;
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {   // up to eight independent loads in flight (chunks <= 32)
                const int c = (int)half + u * kColGroups;
                v[u] = (row_ok && c < chunks) ? __ldg(reinterpret_cast<const float4*>(src) + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int c = (int)half + u * kColGroups;
                if (c >= chunks) break;
                if (l != 0) {
                    v[u].x = gelu_exact(v[u].x); v[u].y = gelu_exact(v[u].y);
                    v[u].z = gelu_exact(v[u].z); v[u].w = gelu_exact(v[u].w);
                }
                const uint32_t f = (uint32_t)(4 * c);
                st_shared_f1(base + kOffR2 + swz_t(f + 0u, rloc), tf32_rn(v[u].x));
                st_shared_f1(base + kOffR2 + swz_t(f + 1u, rloc), tf32_rn(v[u].y));
                st_shared_f1(base + kOffR2 + swz_t(f + 2u, rloc), tf32_rn(v[u].z));
                st_shared_f1(base + kOffR2 + swz_t(f + 3u, rloc), tf32_rn(v[u].w));
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        // ---- (2) dW^T = A^T dZ: M = K features (lanes), N = this layer's width, contraction over the rows ----
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t idesc_w = instr_desc(N);
#pragma unroll
            for (uint32_t rb = 0; rb < 4; rb++)        // K blocks of 32 rows
#pragma unroll
                for (uint32_t ks = 0; ks < 4; ks++)
                    umma_tf32(tmem_dW, smem_desc(base + kOffR2 + rb * kBlockBytes + ks * 32u),
                              smem_desc(base + kOffAlo + rb * kBlockBytes + ks * 32u), idesc_w, (rb | ks) ? 1u : 0u);
            umma_commit(bar[2]);
        }
        ncommit[2]++;
        mbar_wait(bar[2], (ncommit[2] - 1u) & 1u);   // region 2 is about to be overwritten by the W^T stages
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        // ---- (3) dA = dZ W: the forward loop on the packed W^T (K rows, K blocks = N / 32) --------------
        const int nkb = N / kKB;
        const uint32_t idesc = instr_desc(K);
        const uint32_t tile16 = (uint32_t)K * 8u;
        uint32_t last_b = 0;
        for (int kb = 0; kb < nkb; kb++, stage++) {
            const uint32_t b = stage & 1u;
            last_b = b;
            if (kb >= 2) mbar_wait(bar[b], (ncommit[b] - 1u) & 1u);   // (earlier layers' uses are long complete)
            const uint4* src = reinterpret_cast<const uint4*>(L.packed_t + (size_t)kb * 2 * K * kKB);
            const uint32_t w_hi = base + kOffR2 + b * 2u * kBlockBytes, w_lo = w_hi + kBlockBytes;
            // global -> shared without a register round trip: every copy of the block is in flight at once
            for (uint32_t i = tid; i < tile16; i += kMlpThreads) {
                cp_async16(w_hi + i * 16u, src + i);
                cp_async16(w_lo + i * 16u, src + tile16 + i);
            }
            cp_async_wait_all();
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t z_hi = base + kOffAhi + (uint32_t)kb * kBlockBytes;
#pragma unroll
                for (uint32_t ks = 0; ks < 4; ks++) {
                    const uint64_t dZh = smem_desc(z_hi + ks * 32u);
                    const uint64_t dWh = smem_desc(w_hi + ks * 32u), dWl = smem_desc(w_lo + ks * 32u);
                    umma_tf32(tmem_dA, dZh, dWh, idesc, (kb > 0 || ks > 0) ? 1u : 0u);
                    umma_tf32(tmem_dA, dZh, dWl, idesc, 1u);
                }
                umma_commit(bar[b]);
            }
            ncommit[b]++;
        }
        mbar_wait(bar[last_b], (ncommit[last_b] - 1u) & 1u);
        if (nkb >= 2) mbar_wait(bar[last_b ^ 1u], (ncommit[last_b ^ 1u] - 1u) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        // ---- (4) epilogue of dW^T: lane k holds row k of dW^T ------------------------------------------
        for (int c0 = 16 * (int)half; c0 < N; c0 += 16 * kColGroups) {
            float v[16];
            tmem_ld16(tmem_dW + (((warp & 3u) * 32u) << 16) + (uint32_t)c0, v);
            if ((int)rloc < K) {
                float* dst = L.dWt + (size_t)rloc * N + c0;
#pragma unroll
                for (int q = 0; q < 4; q++) red_add_v4(dst + 4 * q, v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            }
        }
        // ---- (5) epilogue of dA: times GELU'(z_{l-1}) -> dZ_{l-1}, in place; or dx for the first layer ----
        for (int c0 = 16 * (int)half; c0 < K; c0 += 16 * kColGroups) {
            float v[16];
            tmem_ld16(tmem_dA + (((warp & 3u) * 32u) << 16) + (uint32_t)c0, v);
            if (l == 0) {
                if (row_ok && ch.dx != nullptr) {
                    float4* dst = reinterpret_cast<float4*>(ch.dx + (size_t)row * K + c0);
#pragma unroll
                    for (int q = 0; q < 4; q++) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                }
            } else {
                if (row_ok) {
                    const float4* zp = reinterpret_cast<const float4*>(L.z_prev + (size_t)row * K + c0);
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float4 t = __ldg(zp + q);
                        v[4 * q] *= gelu_grad(t.x);
                        v[4 * q + 1] *= gelu_grad(t.y);
                        v[4 * q + 2] *= gelu_grad(t.z);
                        v[4 * q + 3] *= gelu_grad(t.w);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 16; j++) v[j] = 0.f;
                }
                warp_colsum16(v, ch.layer[l - 1].db + c0, lane);
#pragma unroll
                for (int q = 0; q < 4; q++) put_dz4(base, rloc, c0 + 4 * q, v + 4 * q);
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
    }
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kBwdTmemCols) : "memory");
}
}  // namespace

int launch_mlp_chains_backward(const MlpBwdArgs& a, cudaStream_t stream) {
    if (a.rows <= 0 || a.n_chains <= 0) return 0;
    if (a.n_chains > HB_MLP_MAX_CHAINS || a.Kx % 16 != 0 || a.Kx <= 0 || a.Kx > 128 || a.ldx % 4 != 0) {
        set_error("mlp_chains_backward: bad input shape");
        return -1;
    }
    for (int c = 0; c < a.n_chains; c++) {
        const MlpBwdChainArgs& ch = a.chain[c];
        if (ch.n_layers < 1 || ch.n_layers > HB_MLP_MAX_LAYERS || ch.n_outs < 0 || ch.n_outs > HB_MLP_MAX_OUTS || !ch.z_head) {
            set_error("mlp_chains_backward: chain %d is malformed", c);
            return -1;
        }
        int K = a.Kx;
        for (int l = 0; l < ch.n_layers; l++) {
            const MlpBwdLayerArgs& L = ch.layer[l];
            if (L.K != K || L.N % kKB != 0 || L.N < kKB || L.N > 128 || !L.packed_t || !L.dWt || !L.db || (l > 0 && !L.z_prev)) {
                set_error("mlp_chains_backward: chain %d layer %d: K=%d (expected %d) N=%d", c, l, L.K, K, L.N);
                return -1;
            }
            K = L.N;
        }
    }
    {
        static bool configured[64] = {false};
        int dev = 0;
        HB_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev]) {
            HB_CUDA_OK(cudaFuncSetAttribute(mlp_chain_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)kMlpSmemBytes));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    const dim3 grid((unsigned)cdiv(a.rows, kTileRows), (unsigned)a.n_chains, 1);
    mlp_chain_backward_kernel<<<grid, kMlpThreads, kMlpSmemBytes, stream>>>(a);
    HB_LAUNCH_OK("mlp_chains_backward", stream, false);
    return 0;
}

}  // namespace hb
