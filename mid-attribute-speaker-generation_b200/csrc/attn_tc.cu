// Masked multi-head self-attention on tcgen05 tensor cores (MASG_PREC_BF16 path) —
// transformer/SubLayers.py:29-57 (head split / concat) + transformer/Modules.py:14-25 (softmax(QK^T / sqrt(128) + key mask) V).
//
// One CTA = 128 queries of one (utterance, head); keys are walked in blocks of 128.  Scores and the output live in
// TMEM (S: columns 0-127, O: columns 128-255); nothing of size S x S ever reaches HBM.
//   * Operands are fp32 in HBM (the QKV projection output, [B, S, 768] = q | k | v, heads of 128) and are converted on
//     the way into shared memory to bf16 hi + lo halves in the UMMA no-swizzle K-major layout [K/8][rows][8]; every
//     product is three MMAs (hi hi + lo hi + hi lo, fp32 accumulate), so scores keep ~16 mantissa bits — the encoder's
//     attention feeds the pitch / energy bucketize and duration rounding decisions.
//   * Two passes over the key blocks instead of an online rescale of O in TMEM: pass 1 computes the row maxima and
//     the softmax denominators (online, in registers), pass 2 recomputes the scores, writes the normalised
//     probabilities P (bf16 hi + lo) to shared memory as the A operand and accumulates O += P V without ever
//     rescaling.  QK^T is cheap here (d_k = 128, S <= 3000): the extra MMAs cost less than a TMEM read-modify-write.
//   * Warps 0-3 own the 128 TMEM lanes (one query row per thread) for the softmax sweeps; in pass 2 warps 4-7 load
//     the transposed V block while warps 0-3 are busy with the softmax of the same key block.
// Padded query rows are written as zeros (they are zeroed after the LayerNorm anyway, Layers.py:25).
#include "common.cuh"
#include <cuda_fp16.h>

namespace {

constexpr int AQ = 128;                  // queries per CTA = keys per block = head dim
constexpr int AT_THREADS = 256;
constexpr int TILE_BYTES = AQ * 128 * 2; // one bf16 operand tile [16][128][8]
constexpr int AT_SMEM = 6 * TILE_BYTES + 1024;      // split-bf16: Q, K / V^T, P as hi + lo
constexpr int AT_SMEM_F16 = 3 * TILE_BYTES + 1024;  // fp16 single pass: two CTAs per SM

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void split8(const float* f, uint4& hi, uint4& lo) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h[i] = pack_bf16(f[2 * i], f[2 * i + 1]);
        const float a = __uint_as_float(h[i] << 16), b = __uint_as_float(h[i] & 0xffff0000u);
        l[i] = pack_bf16(f[2 * i] - a, f[2 * i + 1] - b);
    }
    hi = make_uint4(h[0], h[1], h[2], h[3]); lo = make_uint4(l[0], l[1], l[2], l[3]);
}
__device__ __forceinline__ uint4 half8(const float* f) {
    uint32_t h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { __half2 t = __floats2half2_rn(f[2 * i], f[2 * i + 1]); h[i] = *reinterpret_cast<uint32_t*>(&t); }
    return make_uint4(h[0], h[1], h[2], h[3]);
}
// operand conversion: split-bf16 (hi + lo tiles) or, F16, one fp16 tile (the decoder side: nothing downstream of it decides an integer)
template <bool F16>
__device__ __forceinline__ void put8(const float* f, uint8_t* hi, uint8_t* lo, size_t piece) {
    if (F16) { *reinterpret_cast<uint4*>(hi + piece * 16) = half8(f); return; }
    uint4 h, l;
    split8(f, h, l);
    *reinterpret_cast<uint4*>(hi + piece * 16) = h;
    *reinterpret_cast<uint4*>(lo + piece * 16) = l;
}
template <bool F16>
__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t accumulate) {
    // SWIZZLE_NONE K-major descriptors: LBO = 2048 B (between 8-column groups), SBO = 128 B (between 8-row groups)
    const uint32_t hi = (128u >> 4) | (1u << 14);
    const uint32_t a_lo = ((2048u >> 4) << 16) | (a_addr >> 4), b_lo = ((2048u >> 4) << 16) | (b_addr >> 4);
    const uint32_t idesc = (1u << 4) | (F16 ? 0u : ((1u << 7) | (1u << 10))) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
                 ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate) : "memory");
}
// D (+)= (Ah + Al) x (Bh + Bl) without the lo x lo term, K = 128 (8 steps of 16)
template <bool F16>
__device__ __forceinline__ void mma_split_k128(uint32_t tmem_d, uint32_t ah, uint32_t al, uint32_t bh, uint32_t bl, bool first) {
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
        const uint32_t o = (uint32_t)ks * 4096u;
        mma_bf16<F16>(tmem_d, ah + o, bh + o, (first && ks == 0) ? 0u : 1u);
        if (!F16) {
            mma_bf16<F16>(tmem_d, al + o, bh + o, 1u);
            mma_bf16<F16>(tmem_d, ah + o, bl + o, 1u);
        }
    }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void commit_and_wait(uint32_t bar, uint32_t& phase, bool issuer) {
    if (issuer) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(phase), "r"(100000u) : "memory");
    phase ^= 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// rows [row0, row0 + 128) x 128 columns of a row-major fp32 matrix (row stride 768) -> [16][128][8] bf16 hi / lo tiles;
// rows >= row_limit are zero.  Lane = row: conflict-free 16-byte shared stores, 32-byte global sectors fully used.
template <bool F16>
__device__ __forceinline__ void load_rows_tile(const float* __restrict__ src, int row0, int row_limit, float scale,
                                               uint8_t* hi, uint8_t* lo, int tid, int nthreads) {
    for (int i = tid; i < 16 * AQ; i += nthreads) {
        const int r = i & (AQ - 1), g = i >> 7;
        float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (row0 + r < row_limit) {
            const float4* p = reinterpret_cast<const float4*>(src + (size_t)(row0 + r) * 768 + g * 8);
            const float4 a = __ldg(p), b = __ldg(p + 1);
            f[0] = a.x * scale; f[1] = a.y * scale; f[2] = a.z * scale; f[3] = a.w * scale;
            f[4] = b.x * scale; f[5] = b.y * scale; f[6] = b.z * scale; f[7] = b.w * scale;
        }
        put8<F16>(f, hi, lo, (size_t)i);
    }
}
// transposed: B operand of P V is V^T, [16 key groups][128 value columns][8 keys]; thread = value column
template <bool F16>
__device__ __forceinline__ void load_vt_tile(const float* __restrict__ src, int key0, int key_limit, uint8_t* hi, uint8_t* lo, int n) {
#pragma unroll 2
    for (int kg = 0; kg < 16; ++kg) {
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int key = key0 + kg * 8 + i;
            f[i] = key < key_limit ? __ldg(src + (size_t)key * 768 + n) : 0.f;
        }
        put8<F16>(f, hi, lo, (size_t)(kg * AQ + n));
    }
}

template <bool F16>
__global__ void __launch_bounds__(AT_THREADS, F16 ? 2 : 1)
k_attention_tc(const float* __restrict__ qkv, const int64_t* __restrict__ lens, int S, float* __restrict__ out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int NP = F16 ? 1 : 2;                                        // tiles per operand (hi [+ lo])
    uint8_t* Qh = smem + 1024; uint8_t* Ql = Qh + TILE_BYTES;              // (the lo pointers are never dereferenced when F16)
    uint8_t* Kh = Qh + NP * TILE_BYTES; uint8_t* Kl = Kh + TILE_BYTES;     // K block in pass 1/2, then V^T block in pass 2
    uint8_t* Ph = Kh + NP * TILE_BYTES; uint8_t* Pl = Ph + TILE_BYTES;
    const uint32_t bar = smem_u32(smem);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 64);
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AQ;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int len = (int)min((int64_t)S, lens[b]);
    const float* base = qkv + (size_t)b * S * 768;
    float* obase = out + (size_t)b * S * 256 + h * 128;
    if (q0 >= len) {                         // whole tile is padding
        for (int i = tid; i < AQ * 32; i += AT_THREADS) {
            const int r = i >> 5, c4 = i & 31;
            if (q0 + r < S) reinterpret_cast<float4*>(obase + (size_t)(q0 + r) * 256)[c4] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32((const void*)tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    load_rows_tile<F16>(base + h * 128, q0, S, 1.0f / 11.313708498984761f, Qh, Ql, tid, AT_THREADS);     // q / np.power(128, 0.5)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    const uint32_t tS = tmem, tO = tmem + 128;
    const uint32_t trow = ((uint32_t)((warp & 3) * 32)) << 16;
    const uint32_t qh = smem_u32(Qh), ql = smem_u32(Ql), kh = smem_u32(Kh), kl = smem_u32(Kl), ph = smem_u32(Ph), pl = smem_u32(Pl);
    const int nkb = (len + AQ - 1) / AQ;
    uint32_t phase = 0;
    float m = -INFINITY, l = 0.f;

    // ---------------- pass 1: row maxima and softmax denominators ----------------
    for (int kb = 0; kb < nkb; ++kb) {
        load_rows_tile<F16>(base + 256 + h * 128, kb * AQ, len, 1.f, Kh, Kl, tid, AT_THREADS);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            mma_split_k128<F16>(tS, qh, ql, kh, kl, true);
        }
        commit_and_wait(bar, phase, tid == 0);
        if (warp < 4) {
            const int nvalid = min(AQ, len - kb * AQ);
            float mb = -INFINITY;
            for (int c = 0; c < AQ; c += 16) {
                float v[16];
                tmem_ld16(tS + trow + (uint32_t)c, v);
#pragma unroll
                for (int i = 0; i < 16; ++i) if (c + i < nvalid) mb = fmaxf(mb, v[i]);
            }
            const float mn = fmaxf(m, mb);
            float acc = 0.f;
            for (int c = 0; c < AQ; c += 16) {
                float v[16];
                tmem_ld16(tS + trow + (uint32_t)c, v);
#pragma unroll
                for (int i = 0; i < 16; ++i) if (c + i < nvalid) acc += __expf(v[i] - mn);
            }
            l = l * __expf(m - mn) + acc;
            m = mn;
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
    }
    const float inv_l = 1.f / l;

    // ---------------- pass 2: P = softmax rows (normalised), O += P V ----------------
    for (int kb = 0; kb < nkb; ++kb) {
        if (kb > 0) commit_and_wait(bar, phase, false);                     // P V of the previous block has read Kh/Kl and Ph/Pl
        load_rows_tile<F16>(base + 256 + h * 128, kb * AQ, len, 1.f, Kh, Kl, tid, AT_THREADS);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            mma_split_k128<F16>(tS, qh, ql, kh, kl, true);
        }
        commit_and_wait(bar, phase, tid == 0);
        if (warp < 4) {
            const int nvalid = min(AQ, len - kb * AQ);
            const int r = tid;                                              // query row of this thread = TMEM lane
            for (int c = 0; c < AQ; c += 16) {
                float v[16];
                tmem_ld16(tS + trow + (uint32_t)c, v);
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = (c + i < nvalid) ? __expf(v[i] - m) * inv_l : 0.f;
                put8<F16>(v, Ph, Pl, (size_t)((c >> 3) * AQ + r));
                put8<F16>(v + 8, Ph, Pl, (size_t)(((c >> 3) + 1) * AQ + r));
            }
        } else {
            load_vt_tile<F16>(base + 512 + h * 128, kb * AQ, len, Kh, Kl, tid - 128);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            mma_split_k128<F16>(tO, ph, pl, kh, kl, kb == 0);
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
        }
    }
    commit_and_wait(bar, phase, false);
    if (warp < 4) {
        const int q = q0 + tid;
        for (int c = 0; c < AQ; c += 16) {
            float v[16];
            tmem_ld16(tO + trow + (uint32_t)c, v);
            if (q < S) {
                float4* p = reinterpret_cast<float4*>(obase + (size_t)q * 256 + c);
                const bool live = q < len;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    p[i] = live ? make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

}  // namespace

void launch_attention_tc(masg_ctx* c, cudaStream_t s, const float* qkv, const int64_t* lens, int B, int S, float* out, int f16) {
    dim3 grid((S + AQ - 1) / AQ, 2, B);
    static PerDeviceOnce once;
    if (once.need()) {
        CUDA_OK(cudaFuncSetAttribute(k_attention_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
        CUDA_OK(cudaFuncSetAttribute(k_attention_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM_F16));
    }
    if (f16) k_attention_tc<true><<<grid, AT_THREADS, AT_SMEM_F16, s>>>(qkv, lens, S, out);
    else k_attention_tc<false><<<grid, AT_THREADS, AT_SMEM, s>>>(qkv, lens, S, out);
    CUDA_OK(cudaGetLastError());
    c->launches++;
}
