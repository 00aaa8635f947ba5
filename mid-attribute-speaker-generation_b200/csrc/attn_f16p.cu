// Masked multi-head self-attention of the decoder FFT blocks on planar fp16 operands (MASG_PREC_BF16, fp16 decoder mode) —
// transformer/SubLayers.py:29-57 (head split / concat) + transformer/Modules.py:14-25 (softmax(QK^T / sqrt(128) + key mask) V).
//
// The QKV projection leaves q | k | v as planar fp16 [B][768/8][S][8] — the UMMA no-swizzle operand layout — so a Q / K / V
// block of 128 rows x 128 head columns is 16 bulk async copies of 2 KB, no conversion pass:
//   * Q and K blocks are K-major operands ([16 column groups][128 rows][8]),
//   * the V block is the B operand of P V with N = head columns, K = keys: in the planar layout that is exactly the
//     MN-major no-swizzle form (16 bytes = 8 consecutive N, keys 16 bytes apart), so V needs no transpose either.
// One CTA = 128 queries of one (utterance, head), warp-specialised and asynchronous:
//   warp 8   producer: bulk copies into a 2-stage K ring and a 2-stage V ring (mbarrier complete_tx)
//   warp 9   one thread issues tcgen05.mma: S[i & 1] = Q K_i^T (TMEM columns 0-255), O += P_j V_j (columns 256-383)
//   warps 0-7 softmax: thread = (query row, 64-column half) of the 128 x 128 score block in TMEM
// Two sweeps over the key blocks instead of rescaling O in TMEM: sweep 1 takes the row maxima only (one fmax per score),
// sweep 2 recomputes the scores, writes P = exp(s - m) (un-normalised, fp16, <= 1) to a 2-stage shared-memory ring as the
// A operand and sums the denominators; O is divided by them in the epilogue.  QK^T twice costs 512 tensor cycles per key
// block; the K blocks of sweep 2 come from L2.
// Output: planar fp16 [B][256/8][S][8] (head h = groups 16h .. 16h+15) — the operand layout of the fc projection.
// Padded query rows are written as zeros (they are zeroed after the LayerNorm anyway, Layers.py:25).
#include "tc_ptx.cuh"
#include "tc.cuh"

namespace {
using namespace tcptx;

constexpr int AQ = 128;                       // queries per CTA = keys per block = head dim
constexpr int TILE = AQ * 128 * 2;            // 32 KB
constexpr int NSOFT = 256, NTHR = 320;
constexpr int HDR = 3072;                     // barriers, TMEM slot, row-max and row-sum exchange
constexpr int SMEM_BYTES = HDR + 7 * TILE;    // Q, K x2, V x2, P x2 = 227 KB: one CTA per SM

struct Bars {
    uint32_t b;
    __device__ uint32_t q_full() const { return b; }
    __device__ uint32_t k_full(int s) const { return b + 8 * (1 + s); }
    __device__ uint32_t k_empty(int s) const { return b + 8 * (3 + s); }
    __device__ uint32_t v_full(int s) const { return b + 8 * (5 + s); }
    __device__ uint32_t v_empty(int s) const { return b + 8 * (7 + s); }
    __device__ uint32_t s_full(int s) const { return b + 8 * (9 + s); }
    __device__ uint32_t s_empty(int s) const { return b + 8 * (11 + s); }
    __device__ uint32_t p_full(int s) const { return b + 8 * (13 + s); }
    __device__ uint32_t p_empty(int s) const { return b + 8 * (15 + s); }
    __device__ uint32_t o_full() const { return b + 8 * 17; }
};

// separate descriptor high words for A and B (the V operand is MN-major)
__device__ __forceinline__ void mma2(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
                 ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint4 pack8h(const float* v) {
    return make_uint4(pack_f16(v[0], v[1]), pack_f16(v[2], v[3]), pack_f16(v[4], v[5]), pack_f16(v[6], v[7]));
}

__global__ void __launch_bounds__(NTHR, 1)
k_attention_f16p(const __half* __restrict__ qkv, const int64_t* __restrict__ lens, const int S, __half* __restrict__ out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AQ;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int len = (int)min((int64_t)S, lens[b]);
    const __half* base = qkv + (size_t)b * 768 * S;                 // group g, row t: base + ((size_t)g * S + t) * 8
    __half* obase = out + ((size_t)b * 256 + h * 128) * S;
    if (q0 >= len) {                                                 // whole tile is padding
        for (int i = tid; i < 16 * AQ; i += NTHR) {
            const int g = i >> 7, r = i & (AQ - 1);
            if (q0 + r < S) *reinterpret_cast<uint4*>(obase + ((size_t)g * S + q0 + r) * 8) = make_uint4(0u, 0u, 0u, 0u);
        }
        return;
    }
    const Bars bars{smem_u32(smem)};
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 256);
    float* mx = reinterpret_cast<float*>(smem + 512);                // [2][128] row maxima of the two column halves
    float* lx = reinterpret_cast<float*>(smem + 1536);               // [2][128] row sums
    uint8_t* Qs = smem + HDR;
    uint8_t* Ks = Qs + TILE; uint8_t* Vs = Ks + 2 * TILE; uint8_t* Ps = Vs + 2 * TILE;
    if (tid == 0) {
        mbar_init(bars.q_full(), 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(bars.k_full(s), 1); mbar_init(bars.k_empty(s), 1); mbar_init(bars.v_full(s), 1); mbar_init(bars.v_empty(s), 1);
            mbar_init(bars.s_full(s), 1); mbar_init(bars.s_empty(s), NSOFT); mbar_init(bars.p_full(s), NSOFT); mbar_init(bars.p_empty(s), 1);
        }
        mbar_init(bars.o_full(), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32((const void*)tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tO = tmem + 256;
    const int nkb = (len + AQ - 1) / AQ;

    if (warp == 8) {
        // ------------------------------------------------ producer ------------------------------------------------
        const int qrows = min(AQ, S - q0);
        if (lane == 0) mbar_expect_tx(bars.q_full(), (uint32_t)qrows * 256u);
        __syncwarp();
        if (lane < 16) bulk_g2s(smem_u32(Qs) + lane * 2048, base + ((size_t)(h * 16 + lane) * S + q0) * 8, (uint32_t)qrows * 16u, bars.q_full());
        for (int i = 0; i < 2 * nkb; ++i) {
            const int kb = i < nkb ? i : i - nkb, st = i & 1;
            const int krows = min(AQ, S - kb * AQ);
            mbar_wait_relaxed(bars.k_empty(st), ((i >> 1) & 1) ^ 1);
            if (lane == 0) mbar_expect_tx(bars.k_full(st), (uint32_t)krows * 256u);
            __syncwarp();
            if (lane < 16) bulk_g2s(smem_u32(Ks) + st * TILE + lane * 2048, base + ((size_t)(32 + h * 16 + lane) * S + kb * AQ) * 8, (uint32_t)krows * 16u, bars.k_full(st));
            if (i >= nkb) {
                const int j = i - nkb, vs = j & 1;
                mbar_wait_relaxed(bars.v_empty(vs), ((j >> 1) & 1) ^ 1);
                if (krows < AQ) {                                    // rows past the buffer end: P is 0 there, but 0 x stale NaN is NaN
                    const int tail = AQ - krows;
                    for (int idx = lane; idx < 16 * tail; idx += 32) {
                        const int g = idx / tail, r = krows + idx % tail;
                        *reinterpret_cast<uint4*>(Vs + vs * TILE + g * 2048 + r * 16) = make_uint4(0u, 0u, 0u, 0u);
                    }
                    fence_proxy_async();
                }
                __syncwarp();
                if (lane == 0) mbar_expect_tx(bars.v_full(vs), (uint32_t)krows * 256u);
                __syncwarp();
                if (lane < 16) bulk_g2s(smem_u32(Vs) + vs * TILE + lane * 2048, base + ((size_t)(64 + h * 16 + lane) * S + kb * AQ) * 8, (uint32_t)krows * 16u, bars.v_full(vs));
            }
        }
    } else if (warp == 9) {
        // ------------------------------------------------ MMA issuer ----------------------------------------------
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // fp16 x fp16 -> fp32, M = N = 128
            const uint32_t hi_k = (128u >> 4) | (1u << 14);                    // K-major: SBO 128 B between 8-row groups
            const uint32_t lbo_k = (2048u >> 4) << 16;                         //          LBO 2048 B between 8-column groups
            // V, MN-major: 8-column (N) groups are 2048 B apart, 8-key (K) groups 128 B apart
            // (SBO = stride between 8-element MN groups, LBO = stride between 8-row K groups for this major-ness)
            const uint32_t hi_v = (2048u >> 4) | (1u << 14);
            const uint32_t lbo_v = (128u >> 4) << 16;
            const uint32_t qa = smem_u32(Qs) >> 4, ka = smem_u32(Ks) >> 4, va = smem_u32(Vs) >> 4, pa = smem_u32(Ps) >> 4;
            auto pv = [&](int j) {
                const int st = j & 1, ph = (j >> 1) & 1;
                mbar_wait(bars.p_full(st), ph);
                mbar_wait(bars.v_full(st), ph);
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < 8; ++ks)
                    mma2(tO, lbo_k | (pa + st * (TILE >> 4) + ks * 256), hi_k, lbo_v | (va + st * (TILE >> 4) + ks * 16), hi_v,
                         idesc | (1u << 16), (j > 0 || ks > 0) ? 1u : 0u);
                tc_commit(bars.p_empty(st));
                tc_commit(bars.v_empty(st));
            };
            mbar_wait(bars.q_full(), 0);
            for (int i = 0; i < 2 * nkb; ++i) {
                const int st = i & 1, ph = (i >> 1) & 1;
                mbar_wait(bars.k_full(st), ph);
                mbar_wait(bars.s_empty(st), ph ^ 1);
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < 8; ++ks)
                    mma2(tmem + st * 128, lbo_k | (qa + ks * 256), hi_k, lbo_k | (ka + st * (TILE >> 4) + ks * 256), hi_k, idesc, ks > 0 ? 1u : 0u);
                tc_commit(bars.s_full(st));
                tc_commit(bars.k_empty(st));
                if (i > nkb) pv(i - nkb - 1);
            }
            pv(nkb - 1);
            tc_commit(bars.o_full());
        }
    } else {
        // ------------------------------------------------ softmax -------------------------------------------------
        const int row = (warp & 3) * 32 + lane, half = warp >> 2, col0 = half * 64;
        const uint32_t trow = ((uint32_t)((warp & 3) * 32)) << 16;
        float m = -INFINITY;
        for (int i = 0; i < nkb; ++i) {
            const int st = i & 1, nvalid = min(AQ, len - i * AQ);
            mbar_wait(bars.s_full(st), (i >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 64; c += 16) {
                float v[16];
                tmem_ld16(tmem + st * 128 + trow + (uint32_t)(col0 + c), v);
#pragma unroll
                for (int k = 0; k < 16; ++k) if (col0 + c + k < nvalid) m = fmaxf(m, v[k]);
            }
            tc_fence_before();
            mbar_arrive_relaxed(bars.s_empty(st));
        }
        mx[half * AQ + row] = m;
        asm volatile("bar.sync 1, %0;" ::"n"(NSOFT) : "memory");
        m = fmaxf(mx[row], mx[AQ + row]);
        const float cs = 1.4426950408889634f / 11.313708498984761f;          // log2(e) / sqrt(d_k)
        const float mc = m * cs;
        float l = 0.f;
        for (int j = 0; j < nkb; ++j) {
            const int i = nkb + j, st = i & 1, ps = j & 1, nvalid = min(AQ, len - j * AQ);
            mbar_wait(bars.s_full(st), (i >> 1) & 1);
            tc_fence_after();
            mbar_wait(bars.p_empty(ps), ((j >> 1) & 1) ^ 1);
            uint8_t* P = Ps + ps * TILE;
#pragma unroll
            for (int c = 0; c < 64; c += 16) {
                float v[16];
                tmem_ld16(tmem + st * 128 + trow + (uint32_t)(col0 + c), v);
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    v[k] = (col0 + c + k < nvalid) ? ex2(fmaf(v[k], cs, -mc)) : 0.f;
                    l += v[k];
                }
                *reinterpret_cast<uint4*>(P + (((col0 + c) >> 3) * AQ + row) * 16) = pack8h(v);
                *reinterpret_cast<uint4*>(P + ((((col0 + c) >> 3) + 1) * AQ + row) * 16) = pack8h(v + 8);
            }
            tc_fence_before();
            mbar_arrive_relaxed(bars.s_empty(st));
            fence_proxy_async();
            mbar_arrive(bars.p_full(ps));
        }
        lx[half * AQ + row] = l;
        asm volatile("bar.sync 1, %0;" ::"n"(NSOFT) : "memory");
        const float inv_l = 1.f / (lx[row] + lx[AQ + row]);
        mbar_wait(bars.o_full(), 0);
        tc_fence_after();
        const int q = q0 + row;
        const bool live = q < len;
#pragma unroll
        for (int c = 0; c < 64; c += 16) {
            float v[16];
            tmem_ld16(tO + trow + (uint32_t)(col0 + c), v);
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = live ? v[k] * inv_l : 0.f;
            if (q < S) {
                __half* p = obase + ((size_t)((col0 + c) >> 3) * S + q) * 8;
                *reinterpret_cast<uint4*>(p) = pack8h(v);
                *reinterpret_cast<uint4*>(p + (size_t)S * 8) = pack8h(v + 8);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

}  // namespace

// qkv: planar fp16 [B][96][S][8] (q | k | v, heads of 128 columns); out: planar fp16 [B][32][S][8]
void launch_attention_f16p(masg_ctx* c, cudaStream_t s, const void* qkv, const int64_t* lens, int B, int S, void* out) {
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_attention_f16p, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    dim3 grid((S + AQ - 1) / AQ, 2, B);
    k_attention_f16p<<<grid, NTHR, SMEM_BYTES, s>>>((const __half*)qkv, lens, S, (__half*)out);
    CUDA_OK(cudaGetLastError());
    c->launches++;
}
