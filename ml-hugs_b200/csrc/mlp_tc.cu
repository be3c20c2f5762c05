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
//     stay in L2), so a CTA stages a weight block with plain 16-byte copies, double-buffered against the
//     MMAs of the previous block (tcgen05.commit -> mbarrier says when a buffer may be overwritten);
//   * epilogue: thread t owns row t (TMEM lane t): tcgen05.ld 16 columns at a time, + bias, exact GELU
//     (erf, as torch.nn.GELU()), then either hi/lo split straight back into the A tiles of the next layer
//     or, for the heads, row-contiguous stores to the outputs; optionally the pre-activations are saved
//     for the backward pass.
// Shared memory: A_hi 64 KB + A_lo 64 KB + two weight stages of 32 KB = 192 KB -> one CTA per SM;
// TMEM: 128 columns.
#include "common.cuh"

namespace hb {

namespace {

constexpr int kMlpThreads = 128;
constexpr int kTileRows = 128;
constexpr int kKB = 32;                       // tf32 per 128-byte swizzle row = one K block
constexpr uint32_t kBlockBytes = 128 * 128;   // one K block of a 128-row operand tile
constexpr uint32_t kOffAhi = 0, kOffAlo = 4 * kBlockBytes, kOffW = 8 * kBlockBytes;   // W: [2 stages][hi, lo]
constexpr uint32_t kOffBias = kOffW + 4 * kBlockBytes;
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
__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float sigmoid_exact(float x) { return 1.0f / (1.0f + expf(-x)); }

}  // namespace

// ---- weight packing ----------------------------------------------------------------------------------
// W [n_true x K] row-major (torch.nn.Linear.weight) -> for every K block of 32: a hi tile and a lo tile of
// n_pad rows in the swizzled layout above (rows >= n_true are zero); bias -> n_pad floats.
// packed layout per layer: [K / 32][2 (hi, lo)][n_pad * 128 bytes], then the padded bias.
size_t mlp_packed_floats(int K, int n_pad) { return (size_t)(K / kKB) * 2 * n_pad * kKB + (size_t)n_pad; }

namespace {
__global__ void __launch_bounds__(256) mlp_pack_kernel(const float* __restrict__ W, const float* __restrict__ bias,
                                                       int n_true, int n_pad, int K, float* __restrict__ packed) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;   // element (n, k) of the padded matrix
    const int total = n_pad * K;
    if (e < total) {
        const int n = e / K, k = e % K;
        const float w = (n < n_true) ? __ldg(W + (size_t)n * K + k) : 0.f;
        const float hi = tf32_hi(w), lo = w - hi;
        const int kb = k / kKB, kk = k % kKB;
        const uint32_t off = swz((uint32_t)n, (uint32_t)(kk >> 2)) + (uint32_t)(kk & 3) * 4u;   // bytes
        float* tile_hi = packed + (size_t)kb * 2 * n_pad * kKB;
        float* tile_lo = tile_hi + (size_t)n_pad * kKB;
        tile_hi[off >> 2] = hi;
        tile_lo[off >> 2] = lo;
    }
    if (e < n_pad) packed[(size_t)(K / kKB) * 2 * n_pad * kKB + e] = (e < n_true && bias != nullptr) ? __ldg(bias + e) : 0.f;
}
}  // namespace

int launch_mlp_pack(const float* W, const float* bias, int n_true, int n_pad, int K, float* packed,
                    cudaStream_t stream) {
    if (K <= 0 || K % kKB != 0 || K > 128 || n_pad % 16 != 0 || n_pad < 16 || n_pad > 128 || n_true > n_pad) {
        set_error("mlp_pack: need K in {32,64,96,128}, n_pad a multiple of 16 in [16,128] (K=%d n_pad=%d)", K, n_pad);
        return -1;
    }
    const int total = n_pad * K;
    mlp_pack_kernel<<<cdiv(total, 256), 256, 0, stream>>>(W, bias, n_true, n_pad, K, packed);
    HB_LAUNCH_OK("mlp_pack", stream, false);
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
    const int row0 = (int)blockIdx.x * kTileRows;
    const int row = row0 + (int)tid;          // the row this thread owns in every epilogue
    const bool row_ok = row < a.rows;
    const uint32_t bar[2] = {base + kOffBar, base + kOffBar + 8};
    const uint32_t tmem_slot = base + kOffBar + 16;

    if (tid == 0) {
        mbar_init(bar[0], 1);
        mbar_init(bar[1], 1);
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
        for (int i = (int)tid; i < kTileRows * chunks; i += kMlpThreads) {
            const int r = i / chunks, c = i - r * chunks;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row0 + r < a.rows) v = __ldg(reinterpret_cast<const float4*>(a.x + (size_t)(row0 + r) * a.ldx) + c);
            const uint32_t off = (uint32_t)(c >> 3) * kBlockBytes + swz((uint32_t)r, (uint32_t)(c & 7));
            const float hx = tf32_hi(v.x), hy = tf32_hi(v.y), hz = tf32_hi(v.z), hw = tf32_hi(v.w);
            st_shared_v4(base + kOffAhi + off, hx, hy, hz, hw);
            st_shared_v4(base + kOffAlo + off, v.x - hx, v.y - hy, v.z - hz, v.w - hw);
        }
    }

    uint32_t ncommit[2] = {0u, 0u};   // commits issued on each weight stage (uniform over the CTA)
    uint32_t stage = 0;
    for (int l = 0; l < ch.n_layers; l++) {
        const MlpLayerArgs& L = ch.layer[l];
        const int nkb = L.K / kKB, N = L.N;
        const uint32_t idesc = instr_desc(N);
        const uint32_t tile16 = (uint32_t)N * 8u;   // 16-byte words of one hi (or lo) weight tile
        uint32_t last_b = 0;
        for (int kb = 0; kb < nkb; kb++, stage++) {
            const uint32_t b = stage & 1u;
            last_b = b;
            // the MMAs that read this stage two blocks ago must have retired before it is overwritten
            if (ncommit[b] > 0u) mbar_wait(bar[b], (ncommit[b] - 1u) & 1u);
            const uint4* src = reinterpret_cast<const uint4*>(L.packed + (size_t)kb * 2 * N * kKB);
            const uint32_t w_hi = base + kOffW + b * 2u * kBlockBytes, w_lo = w_hi + kBlockBytes;
            for (uint32_t i = tid; i < tile16; i += kMlpThreads) {
                st_shared_u4(w_hi + i * 16u, __ldg(src + i));
                st_shared_u4(w_lo + i * 16u, __ldg(src + tile16 + i));
            }
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
                umma_commit(bar[b]);   // arrives when every MMA issued so far has completed
            }
            ncommit[b]++;
        }
        // the layer's accumulator is complete when the last commit has arrived
        mbar_wait(bar[last_b], (ncommit[last_b] - 1u) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        // ---- epilogue: thread t <-> row t <-> TMEM lane t -----------------------------------------
        const float* bias = L.packed + (size_t)nkb * 2 * N * kKB;
        const bool last = (l == ch.n_layers - 1);
        for (int c0 = 0; c0 < N; c0 += 16) {
            float v[16];
            tmem_ld16(tmem + ((warp * 32u) << 16) + (uint32_t)c0, v);
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] += __ldg(bias + c0 + j);
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
                    const uint32_t off = (uint32_t)(k >> 5) * kBlockBytes + swz(tid, (uint32_t)((k & 31) >> 2));
                    st_shared_v4(base + kOffAhi + off, h[0], h[1], h[2], h[3]);
                    st_shared_v4(base + kOffAlo + off, lo[0], lo[1], lo[2], lo[3]);
                }
            } else if (row_ok) {
                for (int s = 0; s < ch.n_outs; s++) {
                    const MlpOut& o = ch.out[s];
                    float* dst = o.ptr + (size_t)row * o.ld - o.col0;
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const int col = c0 + j;
                        if (col >= o.col0 && col < o.col0 + o.ncols) dst[col] = apply_act(v[j], o.act);
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

}  // namespace hb
