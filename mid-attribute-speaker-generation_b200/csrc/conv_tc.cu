// tcgen05 implicit-GEMM conv1d / linear / transposed-conv kernel for sm_100a (MASG_PREC_BF16 path).
//
// Same row-GEMM contract as conv_simt.cu (ConvDesc / ConvIO in common.cuh):
//   out[b, q*orow + ooff + n] = (act(sum_{j<taps, c<Cin} in_act(X[b, q + off0 + j*tstep, c]) * W[j*Cin + c, n] + bias[n % bias_mod])
//                                + R1 + R2) * out_scale
// It stands behind every dense contraction of the hot path: the FFT-block QKV / fc projections
// (transformer/SubLayers.py:39-41,54), the k=9 / k=1 position-wise FFN convs (SubLayers.py:88), the variance-
// predictor convs (model/modules.py:209-240), mel_linear (model/fastspeech2.py:242), PostNet (transformer/
// Layers.py:129-137) and every Conv1d / ConvTranspose1d of the HiFi-GAN generator (hifigan/models.py:149-165).
//
// Persistent, warp-specialised kernel: one CTA per SM walks a list of (128*msub rows) x (NT columns) output tiles
// (tile = one utterance's row range x one column tile).  Activation staging, weight streaming, MMA issue and the
// epilogue run as four concurrent roles connected by mbarrier rings, and the TMEM accumulators are double-buffered,
// so the epilogue of tile i overlaps the loads and MMAs of tile i+1:
//   * A operand (activations): the tile's rows PLUS the conv halo are staged ONCE per channel chunk in shared memory
//     in the UMMA no-swizzle K-major layout [KC/8][Rp rows][8 ch] (16-byte pieces).  In that layout a conv tap is a
//     16-byte-granular shift of the descriptor start address, so all taps (and dilations, and the two taps of a
//     phase-decomposed ConvTranspose1d) reuse the same resident tile — no im2col, no per-tap re-read from L2.
//     8 producer warps do the staging with 128-bit loads, converting fp32 -> bf16 and applying the input
//     leaky-relu on the way (hifigan/models.py:98,100,152), zero-filling rows outside [0, Tin) (= conv padding).
//   * B operand (weights): pre-tiled at commit (k_tile_w) into contiguous blobs already in UMMA layout, streamed by
//     one thread with 1-D bulk async copies (cp.async.bulk + mbarrier complete_tx) through a 4-stage ring.
//   * One thread issues tcgen05.mma (M=128, N=NT, K=16, bf16 x bf16 -> fp32) into msub TMEM accumulators of the
//     current accumulator set; tcgen05.commit releases weight stages / activation slots and signals the epilogue.
//   * split mode (FastSpeech2 layers): both operands are carried as bf16 hi + bf16 lo (x = hi + lo to 16 mantissa
//     bits) and each K step issues three MMAs, hi*hi + lo*hi + hi*lo, into the same fp32 accumulator.  That keeps
//     the bucketize / duration-rounding decisions and the mel within 1e-4 of the fp32 oracle while still running
//     on the bf16 tensor pipe (plain bf16 operands give a 0.45 max-abs mel error on random-init weights because
//     pitch/energy bucket indices flip).  HiFi-GAN layers use plain bf16.
//   * 8 epilogue warps (two per TMEM lane quadrant, half the columns each) prefetch the residual rows, read TMEM
//     (tcgen05.ld 32x32b), add the smem-staged bias, apply the activation, residuals and scale, store fp32 and/or bf16
//     rows and hand the accumulator set back to the MMA thread.
#include "common.cuh"
#include "tc.cuh"

namespace {

// warp roles of the persistent CTA (576 threads, one CTA per SM)
constexpr int EPI_WARPS = 8;            // warps 0-7: epilogue (warp % 4 = TMEM lane quadrant, warp / 4 = column half)
constexpr int MMA_WARP = 8;
constexpr int W_WARP = 9;
constexpr int A_WARP0 = 10;
constexpr int A_WARPS = 8;
constexpr int A_THREADS = A_WARPS * 32;
constexpr int NTHREADS = (A_WARP0 + A_WARPS) * 32;     // 576
constexpr int W_STAGES = 4;
constexpr int W_STAGE_BYTES = 16384;
constexpr int SMEM_HDR = 1024;
constexpr int BIAS_MAX = 2048;                          // floats staged per CTA (N <= 2048)
constexpr int A_SLOT_BUDGET = 72 * 1024;                // two activation slots
constexpr int MAX_SMEM = 227 * 1024;

struct TcArgs {
    const void* X; long x_bs; int x_ld; int Tin;
    const bf16* Wt; const float* bias; int bias_mod;
    int Cin, N, taps, off0, tstep;
    int KW, KC, G, a_bytes, stage_bytes, blob_bytes, n_blobs;
    int msub, rows, Rp, off_lo, tmem_cols, split, a_half;
    int B, row_tiles, n_ntiles, total_tiles;
    float* Y; bf16* Y16; long y_bs; long y_limit; long orow; long ooff; int Q;
    int in_act; float in_slope; int out_act; float out_slope;
    const void* R1; const void* R2; int r_bf16; float out_scale;
    int dbg;   // MASG_TC_DEBUG what-if switches (timing experiments only): 1 no weight stream, 2 no activation staging, 4 no epilogue I/O, 8 no MMA
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// one lane of the (converged) warp; the same lane every time, so MMAs and their commits come from one thread
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; both operands K-major
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns; the wait is part of the same statement so the registers are defined after it
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
// UMMA shared-memory descriptor, SWIZZLE_NONE, K-major: 8-row x 16-byte core matrices; SBO = stride between 8-row
// groups, LBO = stride between the two 8-element K chunks of one K=16 MMA.  version = 1 (Blackwell).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}
// streaming 128-bit global load (read-only path, no L1 allocation: every activation byte is used once per tile)
__device__ __forceinline__ uint4 ld_nc(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float act_apply(float v, int act, float slope) {
    switch (act) {
        case ACT_RELU: return fmaxf(v, 0.f);
        case ACT_TANH: return tanhf(v);
        case ACT_LRELU: return v > 0.f ? v : v * slope;
        default: return v;
    }
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t u) {
    __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&u);
    return __bfloat1622float2(h);
}

// ---------------------------------------------------------------------------------------------
// A staging: one 16-byte piece = 8 consecutive channels of one row.  Loads and transforms are separate so the
// producer loop can put U independent 128-bit loads in flight before the first use.
// ---------------------------------------------------------------------------------------------
template <bool XF32>
__device__ __forceinline__ uint4 transform_piece(const TcArgs& a, const uint4 (&raw)[XF32 ? 2 : 1], bool ok, uint4& lo) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    lo = o;
    if (!ok) return o;                                   // row outside [0, Tin): conv zero padding
    float f[8];
    if (XF32) {
        f[0] = __uint_as_float(raw[0].x); f[1] = __uint_as_float(raw[0].y); f[2] = __uint_as_float(raw[0].z); f[3] = __uint_as_float(raw[0].w);
        f[4] = __uint_as_float(raw[XF32 ? 1 : 0].x); f[5] = __uint_as_float(raw[XF32 ? 1 : 0].y);
        f[6] = __uint_as_float(raw[XF32 ? 1 : 0].z); f[7] = __uint_as_float(raw[XF32 ? 1 : 0].w);
    } else {
        if (a.in_act == ACT_NONE) return raw[0];
        float2 t0 = unpack_bf16(raw[0].x), t1 = unpack_bf16(raw[0].y), t2 = unpack_bf16(raw[0].z), t3 = unpack_bf16(raw[0].w);
        f[0] = t0.x; f[1] = t0.y; f[2] = t1.x; f[3] = t1.y; f[4] = t2.x; f[5] = t2.y; f[6] = t3.x; f[7] = t3.y;
    }
    if (a.in_act == ACT_LRELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = f[i] > 0.f ? f[i] : f[i] * a.in_slope;
    } else if (a.in_act != ACT_NONE) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = act_apply(f[i], a.in_act, a.in_slope);
    }
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    if (XF32 && a.split) {                                  // residual of the bf16 rounding, itself rounded to bf16
        float2 h0 = unpack_bf16(o.x), h1 = unpack_bf16(o.y), h2 = unpack_bf16(o.z), h3 = unpack_bf16(o.w);
        lo.x = pack_bf16(f[0] - h0.x, f[1] - h0.y); lo.y = pack_bf16(f[2] - h1.x, f[3] - h1.y);
        lo.z = pack_bf16(f[4] - h2.x, f[5] - h2.y); lo.w = pack_bf16(f[6] - h3.x, f[7] - h3.y);
    }
    return o;
}

template <int NT, bool XF32>
__global__ void __launch_bounds__(NTHREADS, 1) k_conv_tc(const TcArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // header: mbarriers + TMEM base; then the bias row, the two activation slots and the weight ring
    const uint32_t bar0 = smem_u32(smem);
    auto A_FULL = [&](int i) { return bar0 + 8u * i; };                       // 2
    auto A_EMPTY = [&](int i) { return bar0 + 8u * (2 + i); };                // 2
    auto W_FULL = [&](int i) { return bar0 + 8u * (4 + i); };                 // W_STAGES
    auto W_EMPTY = [&](int i) { return bar0 + 8u * (4 + W_STAGES + i); };     // W_STAGES
    auto ACC_FULL = [&](int i) { return bar0 + 8u * (4 + 2 * W_STAGES + i); };        // 2
    auto ACC_EMPTY = [&](int i) { return bar0 + 8u * (6 + 2 * W_STAGES + i); };       // 2
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 512);
    float* bias_s = reinterpret_cast<float*>(smem + SMEM_HDR);
    uint8_t* A_smem = smem + SMEM_HDR + BIAS_MAX * 4;
    uint8_t* W_smem = A_smem + 2 * (size_t)a.a_bytes;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_chunks = a.Cin / a.KC;
    const int blobs_per_chunk = (a.KC / a.KW) * a.taps;
    const int n_stages = (a.n_blobs + a.G - 1) / a.G;
    const int acc_cols = a.msub * NT;                      // TMEM columns of one accumulator set (two sets)

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(A_FULL(i), A_THREADS); mbar_init(A_EMPTY(i), 1);
            mbar_init(ACC_FULL(i), 1); mbar_init(ACC_EMPTY(i), EPI_WARPS);
        }
        for (int i = 0; i < W_STAGES; ++i) { mbar_init(W_FULL(i), 1); mbar_init(W_EMPTY(i), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < a.N; i += NTHREADS) bias_s[i] = a.bias ? a.bias[i % a.bias_mod] : 0.f;
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // tile id -> (batch b, row tile rt, column tile nt); a CTA walks ids blockIdx.x, blockIdx.x + gridDim.x, ...
    if (warp >= A_WARP0) {
        // ===== activation producers: stage rows [q0 + off_lo, q0 + off_lo + rows) x KC channels per chunk =====
        const int tid = threadIdx.x - A_WARP0 * 32;
        const int cpr = a.KC >> 3;                       // 16-byte pieces per row
        const int total = cpr * a.rows;
        const int dr = A_THREADS / cpr, dc = A_THREADS % cpr;
        int ia = 0;                                      // running chunk counter -> slot / parity
        for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
            const int rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
            const uint8_t* Xb = reinterpret_cast<const uint8_t*>(a.X) + (size_t)b * a.x_bs * (XF32 ? 4 : 2);
            const int q_lo = rt * a.msub * 128 + a.off_lo;
            for (int kc = 0; kc < n_chunks; ++kc, ++ia) {
                const int slot = ia & 1;
                if (a.dbg & 2) continue;
                mbar_wait(A_EMPTY(slot), ((ia >> 1) & 1) ^ 1);
                uint8_t* dst = A_smem + (size_t)slot * a.a_bytes;
                const int ch0 = kc * a.KC;
                int r = tid / cpr, cc = tid % cpr;
                constexpr int U = XF32 ? 4 : 8;            // independent 128-bit loads in flight per thread (x2 for fp32 pieces)
                constexpr int RW = XF32 ? 2 : 1;
                const int tmax = a.Tin - 1;
                for (int i = tid; i < total; i += U * A_THREADS) {
                    uint4 raw[U][RW]; int rr[U], cs[U]; bool ok[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        rr[u] = r; cs[u] = cc;
                        const int t = q_lo + r;
                        ok[u] = t >= 0 && t <= tmax;
                        const int tcl = min(max(t, 0), tmax);          // clamped: the load is always legal, the value is dropped when !ok
                        const uint4* p = reinterpret_cast<const uint4*>(Xb + ((size_t)tcl * a.x_ld + ch0 + cc * 8) * (XF32 ? 4 : 2));
                        raw[u][0] = ld_nc(p);
                        if (XF32) raw[u][RW - 1] = ld_nc(p + 1);
                        r += dr; cc += dc;
                        if (cc >= cpr) { cc -= cpr; ++r; }
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        if (i + u * A_THREADS < total) {
                            uint4 vl;
                            const uint4 v = transform_piece<XF32>(a, raw[u], ok[u], vl);
                            const size_t o = ((size_t)cs[u] * a.Rp + rr[u]) * 16;
                            *reinterpret_cast<uint4*>(dst + o) = v;
                            if (XF32 && a.split) *reinterpret_cast<uint4*>(dst + a.a_half + o) = vl;
                        }
                }
                fence_proxy_async();                      // generic-proxy stores -> visible to the tensor core (async proxy)
                mbar_arrive(A_FULL(slot));
            }
        }
    } else if (warp == W_WARP) {
        // ===== weight producer: 1-D bulk copies of pre-tiled blobs =====
        if (lane == 0 && !(a.dbg & 1)) {
            int iw = 0;
            for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
                const int nt = tile % a.n_ntiles;
                const uint8_t* src = reinterpret_cast<const uint8_t*>(a.Wt) + (size_t)nt * a.n_blobs * a.blob_bytes;
                for (int st = 0; st < n_stages; ++st, ++iw) {
                    const int slot = iw % W_STAGES;
                    mbar_wait(W_EMPTY(slot), ((iw / W_STAGES) & 1) ^ 1);
                    const int nb = min(a.G, a.n_blobs - st * a.G);
                    const uint32_t bytes = (uint32_t)nb * a.blob_bytes;
                    mbar_expect_tx(W_FULL(slot), bytes);
                    bulk_g2s(smem_u32(W_smem + (size_t)slot * a.stage_bytes), src + (size_t)st * a.G * a.blob_bytes, bytes, W_FULL(slot));
                }
            }
        }
    } else if (warp == MMA_WARP) {
        // ===== MMA issuer: the whole warp walks the (warp-uniform) loops, one elected lane issues tcgen05.mma / commit =====
        // instruction descriptor: D fp32, A/B bf16, both K-major, N = NT, M = 128
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t w_lbo = (uint32_t)NT * 16u;
        // descriptor halves: hi = SBO (128 B) | version 1; lo = start address | LBO — advancing an operand = adding to lo
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);
        const uint32_t a_lo_base = (((uint32_t)a.Rp * 16u) >> 4) << 16, w_lo_base = (w_lbo >> 4) << 16;
        int iw = 0, ia = 0, it = 0;
        for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
            const int accb = it & 1;
            mbar_wait(ACC_EMPTY(accb), ((it >> 1) & 1) ^ 1);          // epilogue has drained this accumulator set
            tc_fence_after();
            const uint32_t tacc = tmem_base + (uint32_t)(accb * acc_cols);
            int blob = 0;
            for (int st = 0; st < n_stages; ++st, ++iw) {
                const int slot = iw % W_STAGES;
                if (!(a.dbg & 1)) mbar_wait(W_FULL(slot), (iw / W_STAGES) & 1);
                tc_fence_after();
                const uint32_t wbase = smem_u32(W_smem + (size_t)slot * a.stage_bytes);
                for (int g = 0; g < a.G && blob < a.n_blobs; ++g, ++blob) {
                    const int rem = blob % blobs_per_chunk;
                    const int sub = rem / a.taps, j = rem % a.taps;
                    const int aslot = ia & 1;
                    if (rem == 0 && !(a.dbg & 2)) { mbar_wait(A_FULL(aslot), (ia >> 1) & 1); tc_fence_after(); }
                    const uint32_t abase = smem_u32(A_smem + (size_t)aslot * a.a_bytes);
                    const int rowoff = a.off0 + j * a.tstep - a.off_lo;
                    if (elect_one()) {
                        for (int ks = 0; ks < a.KW / 16 && !(a.dbg & 8); ++ks) {
                            const uint32_t aaddr = abase + (uint32_t)((((sub * a.KW + ks * 16) >> 3) * a.Rp + rowoff) * 16);
                            const uint32_t waddr = wbase + (uint32_t)g * a.blob_bytes + (uint32_t)ks * 2u * w_lbo;
                            const uint32_t a_lo = a_lo_base | (aaddr >> 4), w_lo = w_lo_base | (waddr >> 4);
                            const uint64_t bdesc = ((uint64_t)desc_hi << 32) | w_lo;
                            const uint32_t acc = (blob > 0 || ks > 0) ? 1u : 0u;
                            if (!a.split) {
#pragma unroll
                                for (int ms = 0; ms < 8; ++ms)
                                    if (ms < a.msub)
                                        tc_mma(tacc + (uint32_t)(ms * NT), ((uint64_t)desc_hi << 32) | (a_lo + (uint32_t)ms * 128u), bdesc, idesc, acc);
                            } else {
                                const uint64_t bdesc_lo = ((uint64_t)desc_hi << 32) | (w_lo + (uint32_t)(a.blob_bytes >> 5));
                                const uint32_t a_lo2 = a_lo + (uint32_t)(a.a_half >> 4);
#pragma unroll
                                for (int ms = 0; ms < 8; ++ms)
                                    if (ms < a.msub) {
                                        const uint64_t ad_hi = ((uint64_t)desc_hi << 32) | (a_lo + (uint32_t)ms * 128u);
                                        const uint64_t ad_lo = ((uint64_t)desc_hi << 32) | (a_lo2 + (uint32_t)ms * 128u);
                                        tc_mma(tacc + (uint32_t)(ms * NT), ad_hi, bdesc, idesc, acc);
                                        tc_mma(tacc + (uint32_t)(ms * NT), ad_lo, bdesc, idesc, 1u);
                                        tc_mma(tacc + (uint32_t)(ms * NT), ad_hi, bdesc_lo, idesc, 1u);
                                    }
                            }
                        }
                        if (rem == blobs_per_chunk - 1 && !(a.dbg & 2)) tc_commit(A_EMPTY(aslot));     // chunk consumed -> slot may be overwritten
                    }
                    __syncwarp();
                    if (rem == blobs_per_chunk - 1) ++ia;
                }
                if (!(a.dbg & 1) && elect_one()) tc_commit(W_EMPTY(slot));
                __syncwarp();
            }
            if (elect_one()) tc_commit(ACC_FULL(accb));
            __syncwarp();
        }
    } else {
        // ===== epilogue (8 warps): TMEM -> registers -> bias / activation / residual / scale -> global =====
        const int quad = warp & 3, half = warp >> 2;
        const int c_split = ((NT / 16 + 1) / 2) * 16;
        const int c_begin = half ? c_split : 0, c_end = half ? NT : c_split;
        const float* R1f = reinterpret_cast<const float*>(a.R1); const bf16* R1h = reinterpret_cast<const bf16*>(a.R1);
        const float* R2f = reinterpret_cast<const float*>(a.R2); const bf16* R2h = reinterpret_cast<const bf16*>(a.R2);
        const int gc = ((a.R1 || a.R2) && !a.r_bf16) ? 16 : 32;     // columns per residual prefetch group (4 x 16 bytes)
        int it = 0;
        for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
            const int nt = tile % a.n_ntiles, rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
            const int n0 = nt * NT, q0 = rt * a.msub * 128;
            const int accb = it & 1;
            mbar_wait(ACC_FULL(accb), (it >> 1) & 1);
            tc_fence_after();
            const long boff = (long)b * a.y_bs;
            for (int ms = 0; ms < a.msub; ++ms) {
                const int q = q0 + ms * 128 + quad * 32 + lane;
                const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols + ms * NT);
                for (int cg = c_begin; cg < c_end && !(a.dbg & 4); cg += gc) {
                    const int ncol = min(gc, c_end - cg);
                    const long eg = (long)q * a.orow + a.ooff + n0 + cg;
                    const bool rowok = q < a.Q;
                    // prefetch the residual values of this column group (up to 4 x 16 bytes each) before touching TMEM
                    uint4 r1[4], r2[4];
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        r1[h] = make_uint4(0u, 0u, 0u, 0u); r2[h] = r1[h];
                        const int col = a.r_bf16 ? h * 8 : h * 4;                       // first column covered by piece h
                        const long e = eg + (col & ~15);
                        const bool ok = rowok && col < ncol && e >= 0 && e + 16 <= a.y_limit;
                        if (ok && a.R1) r1[h] = a.r_bf16 ? *reinterpret_cast<const uint4*>(R1h + boff + eg + col)
                                                        : *reinterpret_cast<const uint4*>(R1f + boff + eg + col);
                        if (ok && a.R2) r2[h] = a.r_bf16 ? *reinterpret_cast<const uint4*>(R2h + boff + eg + col)
                                                        : *reinterpret_cast<const uint4*>(R2f + boff + eg + col);
                    }
#pragma unroll
                    for (int k16 = 0; k16 < 2; ++k16) {
                        const int cc = k16 * 16;
                        if (cc < ncol) {                                   // warp-uniform
                            float v[16];
                            tmem_ld16(trow + (uint32_t)(cg + cc), v);
                            const long e = eg + cc;
                            if (rowok && e >= 0 && e + 16 <= a.y_limit) {
                                const float* bp = bias_s + n0 + cg + cc;
#pragma unroll
                                for (int i = 0; i < 16; ++i) v[i] = act_apply(v[i] + bp[i], a.out_act, a.out_slope);
                                if (a.R1 || a.R2) {
                                    if (a.r_bf16) {
#pragma unroll
                                        for (int h = 0; h < 2; ++h) {
                                            const uint4 u = r1[(k16 * 2 + h) & 3], w = r2[(k16 * 2 + h) & 3];
                                            float2 t0 = unpack_bf16(u.x), t1 = unpack_bf16(u.y), t2 = unpack_bf16(u.z), t3 = unpack_bf16(u.w);
                                            float2 s0 = unpack_bf16(w.x), s1 = unpack_bf16(w.y), s2 = unpack_bf16(w.z), s3 = unpack_bf16(w.w);
                                            v[h * 8 + 0] += t0.x + s0.x; v[h * 8 + 1] += t0.y + s0.y; v[h * 8 + 2] += t1.x + s1.x; v[h * 8 + 3] += t1.y + s1.y;
                                            v[h * 8 + 4] += t2.x + s2.x; v[h * 8 + 5] += t2.y + s2.y; v[h * 8 + 6] += t3.x + s3.x; v[h * 8 + 7] += t3.y + s3.y;
                                        }
                                    } else {
#pragma unroll
                                        for (int h = 0; h < 4; ++h) {
                                            const uint4 u = r1[(k16 * 4 + h) & 3], w = r2[(k16 * 4 + h) & 3];
                                            v[h * 4 + 0] += __uint_as_float(u.x) + __uint_as_float(w.x); v[h * 4 + 1] += __uint_as_float(u.y) + __uint_as_float(w.y);
                                            v[h * 4 + 2] += __uint_as_float(u.z) + __uint_as_float(w.z); v[h * 4 + 3] += __uint_as_float(u.w) + __uint_as_float(w.w);
                                        }
                                    }
                                }
                                if (a.out_scale != 1.f) {
#pragma unroll
                                    for (int i = 0; i < 16; ++i) v[i] *= a.out_scale;
                                }
                                if (a.Y) {
                                    float4* p = reinterpret_cast<float4*>(a.Y + boff + e);
#pragma unroll
                                    for (int h = 0; h < 4; ++h) p[h] = make_float4(v[h * 4], v[h * 4 + 1], v[h * 4 + 2], v[h * 4 + 3]);
                                }
                                if (a.Y16) {
                                    uint4* p = reinterpret_cast<uint4*>(a.Y16 + boff + e);
#pragma unroll
                                    for (int h = 0; h < 2; ++h)
                                        p[h] = make_uint4(pack_bf16(v[h * 8], v[h * 8 + 1]), pack_bf16(v[h * 8 + 2], v[h * 8 + 3]),
                                                          pack_bf16(v[h * 8 + 4], v[h * 8 + 5]), pack_bf16(v[h * 8 + 6], v[h * 8 + 7]));
                                }
                            }
                        }
                    }
                }
            }
            tc_fence_before();                              // order this warp's tcgen05.ld before the release below
            __syncwarp();
            if (lane == 0) mbar_arrive(ACC_EMPTY(accb));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols) : "memory");
    }
}

// fp32 GEMM-B matrix [taps*Cin][N] -> bf16 blobs [N/NT][Cin/KW][taps][P][KW/8][NT][8]; P = 1, or 2 = (hi, lo) when split
__global__ void k_tile_w(const float* __restrict__ w32, bf16* __restrict__ out, int Cin, int N, int taps, int NT, int KW, int split) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long total = (long)taps * Cin * N;
    if (i >= total) return;
    int e = (int)(i % 8); long r = i / 8;
    int n = (int)(r % NT); r /= NT;
    int k8 = (int)(r % (KW / 8)); r /= (KW / 8);
    int j = (int)(r % taps); r /= taps;
    int ci = (int)(r % (Cin / KW)); int t = (int)(r / (Cin / KW));
    long row = (long)j * Cin + (long)ci * KW + k8 * 8 + e;
    float w = w32[row * N + (long)t * NT + n];
    bf16 hi = __float2bfloat16_rn(w);
    if (!split) { out[i] = hi; return; }
    long blob = i / ((long)KW * NT), within = i % ((long)KW * NT);
    out[blob * 2 * KW * NT + within] = hi;
    out[blob * 2 * KW * NT + (long)KW * NT + within] = __float2bfloat16_rn(w - __bfloat162float(hi));
}

template <int NT, bool XF32>
void launch_inst(const TcArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
    static bool attr = false;
    if (!attr) { CUDA_OK(cudaFuncSetAttribute(k_conv_tc<NT, XF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM)); attr = true; }
    k_conv_tc<NT, XF32><<<grid, NTHREADS, smem, s>>>(a);
}

}  // namespace

void tc_prepare_weights(masg_ctx* c, cudaStream_t s, ConvDesc& d, bool split) {
    MASG_CHECK(d.w32 != nullptr, "tc_prepare_weights: fp32 GEMM weights missing");
    int NT = 0;
    const int cands[5] = {256, 128, 80, 64, 32};
    for (int i = 0; i < 5 && !NT; ++i) if (d.N % cands[i] == 0) NT = cands[i];
    MASG_CHECK(NT != 0, "conv_tc: N=%d is not a multiple of 32/64/80/128/256 (unsupported configuration)", d.N);
    int KW = (d.Cin % 64 == 0) ? 64 : (d.Cin % 32 == 0 ? 32 : 16);
    if (NT == 256 && KW == 64) KW = 32;                     // keep one blob <= 16 KB
    while (split && NT * KW * 4 > W_STAGE_BYTES && KW > 16) KW /= 2;
    MASG_CHECK(d.Cin % KW == 0, "conv_tc: Cin=%d must be a multiple of 16", d.Cin);
    size_t n = (size_t)d.taps * d.Cin * d.N;
    bf16* w16 = nullptr;
    CUDA_OK(cudaMalloc(&w16, n * sizeof(bf16) * (split ? 2 : 1)));
    c->owned.push_back(w16);
    k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, w16, d.Cin, d.N, d.taps, NT, KW, split ? 1 : 0);
    c->launches++;
    d.w16 = w16; d.NT = NT; d.KW = KW; d.split = split ? 1 : 0;
}

void launch_conv_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io) {
    MASG_CHECK(d.w16 != nullptr && d.NT > 0, "conv_tc: bf16 weights not committed");
    MASG_CHECK(io.x_ld % 8 == 0 && io.x_bs % 8 == 0, "conv_tc: x_ld=%d / batch stride must be multiples of 8", io.x_ld);
    MASG_CHECK(io.Y != nullptr || io.Y16 != nullptr, "conv_tc: no output tensor");
    MASG_CHECK(d.N <= BIAS_MAX, "conv_tc: N=%d exceeds the staged bias row (%d)", d.N, BIAS_MAX);
    if (io.Q <= 0 || io.B <= 0) return;
    TcArgs a;
    a.X = io.X; a.x_bs = io.x_bs; a.x_ld = io.x_ld; a.Tin = io.Tin;
    a.Wt = d.w16; a.bias = d.bias; a.bias_mod = d.bias_mod ? d.bias_mod : d.N;
    a.Cin = d.Cin; a.N = d.N; a.taps = d.taps; a.off0 = d.off0; a.tstep = d.tstep;
    a.KW = d.KW; a.split = d.split;
    MASG_CHECK(!d.split || !io.x_bf16, "conv_tc: split-bf16 layers take fp32 activations");
    const int NT = d.NT;
    const size_t a_mul = d.split ? 2 : 1;
    int off_lo = d.off0, off_hi = d.off0;
    for (int j = 0; j < d.taps; ++j) { int o = d.off0 + j * d.tstep; off_lo = std::min(off_lo, o); off_hi = std::max(off_hi, o); }
    const int halo = off_hi - off_lo;
    a.off_lo = off_lo;
    // rows per tile: two accumulator sets share the 512 TMEM columns (2 * msub * NT <= 512); fewer sub-tiles when that
    // would leave SMs without a tile
    int msub = std::min(std::max(256 / NT, 1), (io.Q + 127) / 128);
    const int ntiles_n = d.N / NT;
    while (msub > 1 && (long)((io.Q + msub * 128 - 1) / (msub * 128)) * ntiles_n * io.B < 148) msub = (msub + 1) / 2;
    // channel chunk: the largest multiple of KW dividing Cin whose staged tile (rows + halo) fits one activation slot
    int KC = 0, Rp = 0;
    for (;;) {
        const int rows = msub * 128 + halo;
        auto pad_rows = [&](int kc) {
            const int cpr = kc / 8;
            const int want = cpr >= 8 ? 1 : (8 / std::max(cpr, 1)) % 8;    // bank-conflict-free staging stores
            int rp = rows;
            while ((rp % 8) != want && !(cpr >= 8 && (rp & 1))) ++rp;
            return rp;
        };
        for (int m = d.Cin / d.KW; m >= 1 && !KC; --m) {
            if ((d.Cin / d.KW) % m) continue;
            const int kc = m * d.KW;
            if ((size_t)pad_rows(kc) * kc * 2 * a_mul <= (size_t)A_SLOT_BUDGET) { KC = kc; Rp = pad_rows(kc); }
        }
        if (KC || msub == 1) break;
        msub = (msub + 1) / 2;
    }
    MASG_CHECK(KC > 0, "conv_tc: activation tile does not fit shared memory (Cin=%d taps=%d tstep=%d)", d.Cin, d.taps, d.tstep);
    a.msub = msub; a.rows = msub * 128 + halo; a.Rp = Rp; a.KC = KC;
    a.a_half = (((KC / 8) * Rp * 16) + 127) & ~127;
    a.a_bytes = a.a_half * (int)a_mul;
    a.blob_bytes = d.KW * NT * 2 * (int)a_mul;
    a.G = std::max(1, W_STAGE_BYTES / a.blob_bytes);
    a.stage_bytes = a.G * a.blob_bytes;
    a.n_blobs = (d.Cin / d.KW) * d.taps;
    int cols = 32; while (cols < 2 * msub * NT) cols *= 2;
    MASG_CHECK(cols <= 512, "conv_tc: %d TMEM columns needed", cols);
    a.tmem_cols = cols;
    a.B = io.B; a.row_tiles = (io.Q + msub * 128 - 1) / (msub * 128); a.n_ntiles = ntiles_n;
    a.total_tiles = io.B * a.row_tiles * ntiles_n;
    a.Y = io.Y; a.Y16 = io.Y16; a.y_bs = io.y_bs; a.y_limit = io.y_limit; a.orow = d.orow ? d.orow : d.N; a.ooff = d.ooff; a.Q = io.Q;
    a.in_act = io.in_act; a.in_slope = io.in_slope; a.out_act = io.out_act; a.out_slope = io.out_slope;
    a.R1 = io.R1; a.R2 = io.R2; a.r_bf16 = io.r_bf16; a.out_scale = io.out_scale;
    { static int dbg = -1; if (dbg < 0) { const char* e = getenv("MASG_TC_DEBUG"); dbg = e ? atoi(e) : 0; } a.dbg = dbg; }
    MASG_CHECK(a.orow % 16 == 0 && a.ooff % 16 == 0 && a.y_bs % 16 == 0 && a.bias_mod % 16 == 0,
               "conv_tc: output row stride/offset/bias period must be multiples of 16");
    const size_t smem = SMEM_HDR + BIAS_MAX * 4 + 2 * (size_t)a.a_bytes + (size_t)W_STAGES * a.stage_bytes;
    MASG_CHECK(smem <= (size_t)MAX_SMEM, "conv_tc: %zu bytes of shared memory needed", smem);
    static int n_sm = 0;
    if (!n_sm) { int dev = 0; CUDA_OK(cudaGetDevice(&dev)); CUDA_OK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)); }
    dim3 grid(std::min(a.total_tiles, n_sm), 1, 1);            // persistent: one CTA per SM walks the tile list
    const bool xf32 = !io.x_bf16;
#define MASG_TC_CASE(NTV)                                                                       \
    case NTV: if (xf32) launch_inst<NTV, true>(a, grid, smem, s); else launch_inst<NTV, false>(a, grid, smem, s); break;
    switch (NT) {
        MASG_TC_CASE(32) MASG_TC_CASE(64) MASG_TC_CASE(80) MASG_TC_CASE(128) MASG_TC_CASE(256)
        default: MASG_CHECK(false, "conv_tc: unsupported N tile %d", NT);
    }
#undef MASG_TC_CASE
    CUDA_OK(cudaGetLastError());
    c->launches++;
}

void launch_conv(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io) {
    masg_ctx::ProfRec rec;
    if (c->profiling) {
        CUDA_OK(cudaEventCreate(&rec.e0)); CUDA_OK(cudaEventCreate(&rec.e1));
        rec.flops = 2.0 * io.B * (double)io.Q * d.taps * d.Cin * d.N;      // algorithmic: one multiply-add per (row, k, n)
        CUDA_OK(cudaEventRecord(rec.e0, s));
    }
    if (c->precision == MASG_PREC_BF16) launch_conv_tc(c, s, d, io);
    else launch_conv_f32(c, s, d, io);
    if (c->profiling) { CUDA_OK(cudaEventRecord(rec.e1, s)); c->prof.push_back(rec); }
}
