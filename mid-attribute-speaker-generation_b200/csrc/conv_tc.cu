// tcgen05 implicit-GEMM conv1d / linear / transposed-conv kernel for sm_100a (MASG_PREC_BF16 path).
//
// Same row-GEMM contract as conv_simt.cu (ConvDesc / ConvIO in common.cuh):
//   out[b, q, n] = post( (act(sum_{j<taps, c<Cin} in_act(X[b, q + off0 + j*tstep, c]) * W[j*Cin + c, n] + bias[n % bias_mod])
//                         + R1 + R2) * out_scale )
// It stands behind every dense contraction of the hot path: the FFT-block QKV / fc projections
// (transformer/SubLayers.py:39-41,54), the k=9 / k=1 position-wise FFN convs (SubLayers.py:88), the variance-
// predictor convs (model/modules.py:209-240), mel_linear (model/fastspeech2.py:242), PostNet (transformer/
// Layers.py:129-137) and every Conv1d / ConvTranspose1d of the HiFi-GAN generator (hifigan/models.py:149-165).
//
// Persistent, warp-specialised kernel: one CTA per SM walks a list of (128*msub rows) x (NT columns) output tiles
// (tile = one utterance's row range x one column tile).  Five roles run concurrently, connected by mbarrier rings;
// the TMEM accumulators are double-buffered so the epilogue of tile i overlaps the loads and MMAs of tile i+1:
//   * A operand (activations), UMMA no-swizzle K-major layout (8-row x 16-byte core matrices), so that a conv tap is
//     a 16-byte-granular shift of the descriptor start address: all taps (and dilations, and the two taps of a
//     phase-decomposed ConvTranspose1d) reuse one resident tile — no im2col, no per-tap re-read.
//       - planar input (HiFi-GAN, the fp16 decoder side): activations live in HBM as bf16 / fp16 [B][C/8][T][8] — the operand
//         layout itself — so the rows [t, t + R) of one 8-channel group are ONE contiguous run and the copy warp moves a
//         (128 msub + halo)-row block with one 1-D bulk async copy per group (cp.async.bulk + mbarrier complete_tx, UBLKCP in
//         SASS; no tensor maps: a 4-D cp.async.bulk.tensor box with a 16-byte innermost dimension was measured at ~16 B/clk/SM
//         and removed in round 1).  Rows outside [0, T) (= conv padding, first / last tile only) are zero-filled by the lanes.
//         When the conv takes leaky_relu(x) (hifigan/models.py:98,152) 8 worker warps apply it in place in shared memory
//         before the tile is released.
//       - row-major fp32 input (FastSpeech2): 8 worker warps stage the tile with 128-bit loads, converting to bf16
//         (hi + lo halves in split mode).
//   * B operand (weights): pre-tiled at commit (k_tile_w) into contiguous blobs already in UMMA layout, streamed by
//     one thread with 1-D bulk async copies through a 3- or 4-stage ring, or resident for the life of the CTA when one N
//     tile's weights fit (<= 100 KB).
//   * One thread issues tcgen05.mma (M=128, N=NT, K=16, bf16 x bf16 -> fp32) into msub TMEM accumulators; the issue
//     loop is specialised on msub and keeps every descriptor a running add (a single thread issues one instruction
//     every few cycles, and an N=32 MMA only lasts 16); tcgen05.commit releases weight stages / activation slots.
//   * split mode (FastSpeech2 layers): both operands are carried as bf16 hi + bf16 lo (x = hi + lo to 16 mantissa
//     bits) and each K step issues three MMAs, hi*hi + lo*hi + hi*lo, into the same fp32 accumulator.  That keeps
//     the bucketize / duration-rounding decisions and the mel within 1e-4 of the fp32 oracle while still running
//     on the bf16 tensor pipe.  HiFi-GAN layers use plain bf16.
//   * 8 epilogue warps (two per TMEM lane quadrant, half the columns each) read TMEM (tcgen05.ld 32x32b), add the
//     smem-staged bias, apply activation, residuals, scale and the optional trailing leaky-relu, and store fp32
//     row-major rows or planar bf16 (lane = time row, so each store instruction writes 512 contiguous bytes).
#include "common.cuh"
#include "tc.cuh"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

// warp roles of the persistent CTA (704 threads, one CTA per SM)
constexpr int EPI_WARPS = 8;            // warps 0-7: epilogue (warp % 4 = TMEM lane quadrant, warp / 4 = column half)
constexpr int MMA_WARP = 8;             // warps 8-11: MMA issuers (one thread each, on the four different warp schedulers)
constexpr int MMA_WARPS = 4;
constexpr int W_WARP = 12;              // weight stream
constexpr int T_WARP = 13;              // bulk-copy issuer for planar activations
constexpr int A_WARP0 = 14;             // activation workers
constexpr int A_WARPS = 8;
constexpr int A_THREADS = A_WARPS * 32;
constexpr int NTHREADS = (A_WARP0 + A_WARPS) * 32;     // 704
// Per-role register budgets (setmaxnreg.dec in the issuer / copy groups, .inc in the epilogue groups, 768 threads = six aligned
// groups of four warps) were tried in round 2 and do not pay: ptxas then allocates each role's region at that role's cap, and
// every split of the 61,440 registers that was tried (epilogue / control / workers = 128/40/72, 112/56/72, 104/56/80, 96/64/80)
// spills MORE (300-3,000 bytes per thread against 64-144 at a uniform 80): the fully unrolled issue loops and the fp32 staging
// loop need ~80 registers themselves, so there is nothing to give to the epilogue.
constexpr int W_STAGES = 4;
constexpr int W_STAGE_BYTES = 16384;
constexpr int A_SLOTS = 3;
constexpr int A_SLOT_BYTES = 48 * 1024;
constexpr int SMEM_HDR = 1024;
constexpr int BIAS_MAX = 2048;                          // floats staged per CTA (N <= 2048)
constexpr int MAX_SMEM = 227 * 1024;

struct TcArgs {
    // input
    const void* X; long x_bs; int x_ld; int Tin;       // row-major fp32 input (worker-staged); unused with TMA
    const bf16* Wt; const float* bias; int bias_mod;
    int Cin, N, taps, off0, tstep;
    int KW, KC, G, a_bytes, stage_bytes, blob_bytes, n_blobs;
    int msub, rows, Rp, off_lo, tmem_cols, split, a_half;
    int y_f16;                                         // planar output is fp16
    int f16;                                           // operands are fp16 (single pass, 11-bit mantissa) instead of bf16
    int a_lbo16, a_ms16, a_tx, a_blk_bytes;            // operand strides in 16-byte units; TMA bytes per chunk; bytes per ms block
    int B, row_tiles, n_ntiles, total_tiles;
    int total_pairs;                                   // pair mode: tile pairs (two consecutive row blocks x one column tile)
    // output
    float* Y; bf16* Y16; long y_bs; long y_limit; long orow; long ooff; int Q;
    int y_planar, y_T, up, up_pad, cout_shift;         // planar bf16 output [B][Cout/8][y_T][8]; t = q*up + n/Cout - up_pad
    int in_act; float in_slope; int out_act; float out_slope;
    const void* R1; const void* R2; int r_bf16; float out_scale; float post_slope;
    int w_resident;                                    // the whole weight set is one stage, loaded once per CTA
    int res_pf;                                        // copy warp prefetches the tile's residual rows into L2 (MASG_TC_RESPF, default on)
    int w_stages;                                      // weight ring depth (3 with 32 KB stages, else 4 with 16 KB stages)
    // fused LayerNorm epilogue (NT == N == 256, row-major fp32 output): see epi_layernorm
    const float* ln_g; const float* ln_b; const int64_t* ln_lens; int ln_T; bf16* Y16p;
    int copiers;   // warps issuing the activation bulk copies (the copy warp + up to three idle MMA warps; MASG_TC_COPIERS forces a count)
    int dbg;   // MASG_TC_DEBUG what-if switches (timing experiments only): 1 no weight stream, 2 no activation staging, 4 no epilogue I/O, 8 no MMA
};

// ---------------------------------------------------------------------------------------------
// Worker staging of row-major fp32 activations: one 16-byte piece = 8 consecutive channels of one row.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 convert_piece(const TcArgs& a, const uint4 r0, const uint4 r1, bool ok, uint4& lo) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    lo = o;
    if (!ok) return o;                                   // row outside [0, Tin): conv zero padding
    float f[8] = {__uint_as_float(r0.x), __uint_as_float(r0.y), __uint_as_float(r0.z), __uint_as_float(r0.w),
                  __uint_as_float(r1.x), __uint_as_float(r1.y), __uint_as_float(r1.z), __uint_as_float(r1.w)};
    if (a.in_act != ACT_NONE) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = act_apply(f[i], a.in_act, a.in_slope);
    }
    if (a.f16) {
        o.x = pack_f16(f[0], f[1]); o.y = pack_f16(f[2], f[3]); o.z = pack_f16(f[4], f[5]); o.w = pack_f16(f[6], f[7]);
        return o;
    }
    o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    if (a.split) {                                       // residual of the bf16 rounding, itself rounded to bf16
        float2 h0 = unpack_bf16(o.x), h1 = unpack_bf16(o.y), h2 = unpack_bf16(o.z), h3 = unpack_bf16(o.w);
        lo.x = pack_bf16(f[0] - h0.x, f[1] - h0.y); lo.y = pack_bf16(f[2] - h1.x, f[3] - h1.y);
        lo.z = pack_bf16(f[4] - h2.x, f[5] - h2.y); lo.w = pack_bf16(f[6] - h3.x, f[7] - h3.y);
    }
    return o;
}

// MASG_TC_DEBUG & 16: cycles spent waiting, per role (issuer 0, epilogue warp 0, copy warp, weight warp), summed over CTAs
__device__ unsigned long long g_tc_stats[16];
#define STAT_WAIT(flag, idx, stmt) do { if (flag) { const long long _t = clock64(); stmt; atomicAdd(&g_tc_stats[idx], (unsigned long long)(clock64() - _t)); } else { stmt; } } while (0)

// MASG_TC_DEBUG & 32: timeline of CTA 0 — clock64 at the hand-over points of every role for its first TRACE_TILES tiles.
// roles: 0 copy warp, 1 worker 0, 2 issuer 0, 3 epilogue warp 0, 4 weight warp; events are role-specific (see tc_debug_trace users)
constexpr int TRACE_TILES = 96, TRACE_EVTS = 4, TRACE_ROLES = 5;
__device__ long long g_tc_trace[TRACE_ROLES * TRACE_TILES * TRACE_EVTS];
#define TRACE(on, role, it, evt) do { if ((on) && (it) < TRACE_TILES) g_tc_trace[((role) * TRACE_TILES + (it)) * TRACE_EVTS + (evt)] = clock64(); } while (0)

struct Bars {                                            // mbarrier addresses (shared window), 8 bytes each
    uint32_t base;
    __device__ uint32_t a_full(int i) const { return base + 8u * i; }
    __device__ uint32_t a_empty(int i) const { return base + 8u * (A_SLOTS + i); }
    __device__ uint32_t a_raw(int i) const { return base + 8u * (2 * A_SLOTS + i); }
    __device__ uint32_t w_full(int i) const { return base + 8u * (3 * A_SLOTS + i); }
    __device__ uint32_t w_empty(int i) const { return base + 8u * (3 * A_SLOTS + W_STAGES + i); }
    __device__ uint32_t acc_full(int i) const { return base + 8u * (3 * A_SLOTS + 2 * W_STAGES + i); }
    __device__ uint32_t acc_empty(int i) const { return base + 8u * (3 * A_SLOTS + 2 * W_STAGES + 2 + i); }
    __device__ uint32_t w_peer(int i) const { return base + 8u * (3 * A_SLOTS + 2 * W_STAGES + 4 + i); }   // pair mode, leader: the peer's weight stage has landed
};
static_assert(8 * (3 * A_SLOTS + 3 * W_STAGES + 4) <= 512, "barrier block overlaps the TMEM slot");

// ---------------------------------------------------------------------------------------------
// MMA issue loop (one thread).  Everything that changes per MMA is a running 32-bit add on the low descriptor word.
// ---------------------------------------------------------------------------------------------
// NI threads (one per MMA warp) share the work: issuer w owns accumulators w, w + NI, ... of the MS sub-tiles and walks
// the same barrier sequence as the others; every release barrier therefore expects NI commits.  (Measured: one
// thread sustains ~56 cycles per MMA, the tensor core needs 40 at N=32 and 48 at N=64 — shared-memory operand reads,
// (4096 + 32 N) bytes at 128 B/clk — so two to four issuers are needed to keep it busy.)
template <int NT, int MS, int NI, bool SPLIT, int W, bool PAIR>
__device__ __forceinline__ void mma_role(const TcArgs& a, const Bars bars, const uint32_t a_smem, const uint32_t w_smem, const uint32_t tmem_base) {
    constexpr int w = W;                                   // issuer index: compile-time, so every value below is warp-uniform
    constexpr int MSL = MS / NI;
    // instruction descriptor: D fp32, A/B bf16, both K-major, N = NT, M = 128
    const uint32_t ab_fmt = a.f16 ? 0u : ((1u << 7) | (1u << 10));                         // A / B format: 0 = fp16, 1 = bf16
    // pair mode (cta_group::2): M = 256 = 128 rows from each CTA of the pair; each CTA holds NT / 2 of the tile's weight columns
    constexpr int NTB = PAIR ? NT / 2 : NT;
    const uint32_t idesc = (1u << 4) | ab_fmt | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)((PAIR ? 256 : 128) >> 4) << 24);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);                     // SBO = 128 B, descriptor version 1
    const uint32_t a_fixed = (uint32_t)a.a_lbo16 << 16, w_fixed = (uint32_t)NTB << 16;    // LBO fields (16-byte units)
    const uint32_t a_ks = 2u * (uint32_t)a.a_lbo16, w_ks = 2u * (uint32_t)NTB;             // one K=16 step = two 8-channel groups
    const uint32_t a_sub = (uint32_t)(a.KW >> 3) * (uint32_t)a.a_lbo16;
    const uint32_t a_ms = (uint32_t)a.a_ms16;
    const uint32_t blob16 = (uint32_t)a.blob_bytes >> 4, a_half16 = (uint32_t)a.a_half >> 4, w_half16 = blob16 >> 1;
    const int n_chunks = a.Cin / a.KC, subs = a.KC / a.KW, ksteps = a.KW >> 4;
    const bool no_w = a.dbg & 1, no_a = a.dbg & 2, no_mma = a.dbg & 8;
    const bool st = (a.dbg & 16) && w == 0 && (threadIdx.x & 31) == 0;
    const long long t_begin = st ? clock64() : 0;
    const bool tr = (a.dbg & 32) && w == 0 && blockIdx.x == 0 && (threadIdx.x & 31) == 0;
    int wslot = 0, aslot = 0, it = 0;
    uint32_t wphase = 0, aphase = 0;
    const int w_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, w_stride = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int w_total = PAIR ? a.total_pairs : a.total_tiles;
    auto commit = [](uint32_t bar) { if (PAIR) tc_commit2(bar); else tc_commit(bar); };
    for (int tile = w_first; tile < w_total; tile += w_stride, ++it) {
        const int accb = it & 1;
        STAT_WAIT(st, 0, mbar_wait(bars.acc_empty(accb), ((it >> 1) & 1) ^ 1));   // epilogue has drained this accumulator set
        tc_fence_after();
        TRACE(tr, 2, it, 0);
        const uint32_t tacc_w = tmem_base + (uint32_t)(accb * MS * NT + w * NT);
        uint32_t accum = 0;
        // (Not issuing the MMAs of a sub-tile that starts beyond the utterance's last row — the tail tile of T = 600 covers rows
        // 512..767, its second sub-tile is empty: 1 of 6 — was tried: the test inside the issue loop slows EVERY layer of this
        // instantiation by ~10 % (stage-1 k = 11 0.335 -> 0.375 ms, w_1 0.192 -> 0.210): one thread issues an instruction every
        // few cycles and has no slack.  Removed.)
        int g = a.G;                                                       // forces a stage fetch before the first blob
        bool have_stage = false;
        uint32_t wbase16 = 0;
        for (int kc = 0; kc < n_chunks; ++kc) {
            if (!no_a) { STAT_WAIT(st, 1, mbar_wait(bars.a_full(aslot), aphase)); tc_fence_after(); }
            if (kc == 0) TRACE(tr, 2, it, 1);
            const uint32_t abase16 = (a_smem + (uint32_t)aslot * (uint32_t)a.a_bytes) >> 4;
            for (int sub = 0; sub < subs; ++sub) {
                uint32_t a_tap = abase16 + (uint32_t)sub * a_sub + (uint32_t)(a.off0 - a.off_lo);
                for (int j = 0; j < a.taps; ++j, a_tap += (uint32_t)a.tstep) {
                    if (g == a.G) {                                        // next weight stage
                        if (have_stage) {
                            if (!no_w && elect_one()) commit(bars.w_empty(wslot));
                            if (++wslot == a.w_stages) { wslot = 0; wphase ^= 1; }
                        }
                        if (!no_w) { STAT_WAIT(st, 2, mbar_wait(bars.w_full(wslot), wphase)); if (PAIR) mbar_wait(bars.w_peer(wslot), wphase); tc_fence_after(); }
                        wbase16 = (w_smem + (uint32_t)wslot * (uint32_t)a.stage_bytes) >> 4;
                        g = 0; have_stage = true;
                    }
                    uint32_t a_lo = (a_fixed | a_tap) + (uint32_t)w * a_ms, w_lo = w_fixed | (wbase16 + (uint32_t)g * blob16);
                    ++g;
                    if (no_mma) continue;
                    if (elect_one()) {
                        for (int ks = 0; ks < ksteps; ++ks, a_lo += a_ks, w_lo += w_ks) {
#pragma unroll
                            for (int i = 0; i < MSL; ++i) {
                                const uint32_t d = tacc_w + (uint32_t)(i * NI * NT), al = a_lo + (uint32_t)(i * NI) * a_ms;
                                if (PAIR) tc_mma2(d, al, w_lo, desc_hi, idesc, (ks == 0) ? accum : 1u);
                                else tc_mma(d, al, w_lo, desc_hi, idesc, (ks == 0) ? accum : 1u);
                                if (SPLIT) {
                                    tc_mma(d, al + a_half16, w_lo, desc_hi, idesc, 1u);
                                    tc_mma(d, al, w_lo + w_half16, desc_hi, idesc, 1u);
                                }
                            }
                        }
                    }
                    __syncwarp();
                    accum = 1u;
                }
            }
            if (!no_a && elect_one()) commit(bars.a_empty(aslot));                        // chunk consumed -> slot may be overwritten
            if (++aslot == A_SLOTS) { aslot = 0; aphase ^= 1; }
        }
        if (!no_w && elect_one()) commit(bars.w_empty(wslot));             // the weight stream restarts on a stage boundary per tile
        if (++wslot == a.w_stages) { wslot = 0; wphase ^= 1; }
        if (elect_one()) commit(bars.acc_full(accb));
        TRACE(tr, 2, it, 2);
    }
    if (st) atomicAdd(&g_tc_stats[3], (unsigned long long)(clock64() - t_begin));
}

// ---------------------------------------------------------------------------------------------
// Epilogue helpers
// ---------------------------------------------------------------------------------------------
// bias + activation on 16 accumulator columns
__device__ __forceinline__ void epi_bias_act(const TcArgs& a, float (&v)[16], const float* bp) {
    const float4* b4 = reinterpret_cast<const float4*>(bp);
#pragma unroll
    for (int h = 0; h < 4; ++h) {
        const float4 bb = b4[h];
        v[h * 4 + 0] += bb.x; v[h * 4 + 1] += bb.y; v[h * 4 + 2] += bb.z; v[h * 4 + 3] += bb.w;
    }
    if (a.out_act == ACT_LRELU) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * a.out_slope;
    } else if (a.out_act != ACT_NONE) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = act_apply(v[i], a.out_act, a.out_slope);
    }
}
__device__ __forceinline__ void add_bf16x8(float* v, const uint4 u) {
    const float2 t0 = unpack_bf16(u.x), t1 = unpack_bf16(u.y), t2 = unpack_bf16(u.z), t3 = unpack_bf16(u.w);
    add_f32x2(v[0], v[1], t0.x, t0.y); add_f32x2(v[2], v[3], t1.x, t1.y); add_f32x2(v[4], v[5], t2.x, t2.y); add_f32x2(v[6], v[7], t3.x, t3.y);
}
__device__ __forceinline__ void epi_scale_post(const TcArgs& a, float (&v)[16]) {
    if (a.out_scale != 1.f) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] *= a.out_scale;
    }
    if (a.post_slope != 1.f) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * a.post_slope;
    }
}
__device__ __forceinline__ uint4 pack8h(const float* v) {
    return make_uint4(pack_f16(v[0], v[1]), pack_f16(v[2], v[3]), pack_f16(v[4], v[5]), pack_f16(v[6], v[7]));
}
__device__ __forceinline__ uint4 pack8(const float* v) {
    return make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// planar bf16 output: element offset (within the batch) of the 8-channel group holding column n at row q, or -1
__device__ __forceinline__ long planar_off(const TcArgs& a, int q, int n) {
    const int co = n & ((1 << a.cout_shift) - 1), r = n >> a.cout_shift;
    const int t = q * a.up + r - a.up_pad;
    if (q >= a.Q || t < 0 || t >= a.y_T) return -1;
    return ((long)(co >> 3) * a.y_T + t) * 8;
}

// ---------------------------------------------------------------------------------------------
// Planar bf16 epilogue [Cout/8][y_T][8] (HiFi-GAN layers): lane = time row, one 16-byte store per 8 columns.  Columns
// outer (bias and column offsets in registers), sub-tiles inner.  Specialised on the residuals so that a plain conv
// (bias -> leaky-relu -> store) is ~70 instructions per 16 columns; with residuals the next sub-tile's rows are requested
// before this sub-tile's accumulator is read.  (The generic version of this loop cost ~200 instructions per 16 columns and
// set the pace of every HBM-bound layer: profiles/r1_conv_tc_trace.log.)
// ---------------------------------------------------------------------------------------------
template <int NT, bool HAS_R1, bool HAS_R2>
__device__ __forceinline__ void epi_planar(const TcArgs& a, const uint32_t tacc, const float* bias_s, const int n0, const int q0,
                                           const long boff, const int quad, const int half, const int lane) {
    constexpr int c_split = ((NT / 16 + 1) / 2) * 16;
    const int c_begin = half ? c_split : 0, c_end = half ? NT : c_split;
    const bf16* R1h = reinterpret_cast<const bf16*>(a.R1); const bf16* R2h = reinterpret_cast<const bf16*>(a.R2);
    const int q_lane = q0 + quad * 32 + lane;
    const long gstride = (long)a.y_T * 8;                                   // elements between 8-channel groups
    const int t_step = 128 * a.up;
    const bool lrelu_out = a.out_act == ACT_LRELU, other_act = a.out_act != ACT_NONE && a.out_act != ACT_LRELU;
    const bool scaled = a.out_scale != 1.f, post = a.post_slope != 1.f;
    const Slope2 sl_out = make_slope2(a.out_slope), sl_post = make_slope2(a.post_slope);
    for (int cc = c_begin; cc < c_end; cc += 16) {
        const int n = n0 + cc;
        const int co = n & ((1 << a.cout_shift) - 1), r = n >> a.cout_shift;
        const int t0 = q_lane * a.up + r - a.up_pad;
        const long cbase = boff + (long)(co >> 3) * gstride;
        const float4* b4 = reinterpret_cast<const float4*>(bias_s + n);
        const float4 ba = b4[0], bb = b4[1], bc = b4[2], bd = b4[3];
        uint4 r1a, r1b, r2a, r2b;
        bool ok = q_lane < a.Q && (unsigned)t0 < (unsigned)a.y_T;
        long off = cbase + (long)t0 * 8;
        if (HAS_R1) { r1a = r1b = make_uint4(0u, 0u, 0u, 0u); if (ok) { r1a = *reinterpret_cast<const uint4*>(R1h + off); r1b = *reinterpret_cast<const uint4*>(R1h + off + gstride); } }
        if (HAS_R2) { r2a = r2b = make_uint4(0u, 0u, 0u, 0u); if (ok) { r2a = *reinterpret_cast<const uint4*>(R2h + off); r2b = *reinterpret_cast<const uint4*>(R2h + off + gstride); } }
        for (int ms = 0; ms < a.msub; ++ms) {
            // rows of the next sub-tile
            const int tn = t0 + (ms + 1) * t_step;
            const bool okn = ms + 1 < a.msub && q_lane + (ms + 1) * 128 < a.Q && (unsigned)tn < (unsigned)a.y_T;
            const long offn = cbase + (long)tn * 8;
            uint4 x1a, x1b, x2a, x2b;
            if (HAS_R1) { x1a = x1b = make_uint4(0u, 0u, 0u, 0u); if (okn) { x1a = *reinterpret_cast<const uint4*>(R1h + offn); x1b = *reinterpret_cast<const uint4*>(R1h + offn + gstride); } }
            if (HAS_R2) { x2a = x2b = make_uint4(0u, 0u, 0u, 0u); if (okn) { x2a = *reinterpret_cast<const uint4*>(R2h + offn); x2b = *reinterpret_cast<const uint4*>(R2h + offn + gstride); } }
            float v[16];
            tmem_ld16(tacc + (uint32_t)(ms * NT + cc), v);
            add_f32x2(v[0], v[1], ba.x, ba.y); add_f32x2(v[2], v[3], ba.z, ba.w); add_f32x2(v[4], v[5], bb.x, bb.y); add_f32x2(v[6], v[7], bb.z, bb.w);
            add_f32x2(v[8], v[9], bc.x, bc.y); add_f32x2(v[10], v[11], bc.z, bc.w); add_f32x2(v[12], v[13], bd.x, bd.y); add_f32x2(v[14], v[15], bd.z, bd.w);
            if (other_act) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = act_apply(v[i], a.out_act, a.out_slope);
            }
            uint4 pa, pb;
            if (HAS_R1 || HAS_R2 || scaled) {
                if (lrelu_out) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], v[i] * a.out_slope);
                }
                if (HAS_R1) { add_bf16x8(v, r1a); add_bf16x8(v + 8, r1b); }
                if (HAS_R2) { add_bf16x8(v, r2a); add_bf16x8(v + 8, r2b); }
                if (scaled) {
#pragma unroll
                    for (int i = 0; i < 16; i += 2) mul_f32x2(v[i], v[i + 1], a.out_scale, a.out_scale);
                }
                if (a.y_f16) { pa = pack8h(v); pb = pack8h(v + 8); } else { pa = pack8(v); pb = pack8(v + 8); }
            } else {
                if (a.y_f16) { pa = pack8h(v); pb = pack8h(v + 8); } else { pa = pack8(v); pb = pack8(v + 8); }
                if (lrelu_out) {                                   // on the packed values: 3 instructions per pair
                    pa.x = lrelu_bf16x2(pa.x, sl_out); pa.y = lrelu_bf16x2(pa.y, sl_out); pa.z = lrelu_bf16x2(pa.z, sl_out); pa.w = lrelu_bf16x2(pa.w, sl_out);
                    pb.x = lrelu_bf16x2(pb.x, sl_out); pb.y = lrelu_bf16x2(pb.y, sl_out); pb.z = lrelu_bf16x2(pb.z, sl_out); pb.w = lrelu_bf16x2(pb.w, sl_out);
                }
            }
            if (post) {
                pa.x = lrelu_bf16x2(pa.x, sl_post); pa.y = lrelu_bf16x2(pa.y, sl_post); pa.z = lrelu_bf16x2(pa.z, sl_post); pa.w = lrelu_bf16x2(pa.w, sl_post);
                pb.x = lrelu_bf16x2(pb.x, sl_post); pb.y = lrelu_bf16x2(pb.y, sl_post); pb.z = lrelu_bf16x2(pb.z, sl_post); pb.w = lrelu_bf16x2(pb.w, sl_post);
            }
            if (ok) {
                *reinterpret_cast<uint4*>(a.Y16 + off) = pa;
                *reinterpret_cast<uint4*>(a.Y16 + off + gstride) = pb;
            }
            ok = okn; off = offn;
            if (HAS_R1) { r1a = x1a; r1b = x1b; }
            if (HAS_R2) { r2a = x2a; r2b = x2b; }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Planar epilogue with residuals, full tiles (msub * NT = 256): a thread owns 8 steps of (16 columns, sub-tile).  The R1
// rows of the first 4 steps are requested BEFORE the accumulator wait and the ring then stays 4 steps (8 x 16 bytes per
// thread) ahead, so the global-load latencies overlap the tile's MMAs and each other; epi_planar above runs one step
// ahead at best and eats most of one load latency per step — that made the second conv of every C = 128 / 256 ResBlock
// pair epilogue-bound (0.53 ms against 0.27 ms for the first conv of the same shape).  R2 (the ResBlock-sum accumulator,
// one launch in three) runs two steps ahead.
// ---------------------------------------------------------------------------------------------
constexpr int RES_STEPS = 8, RES_D2 = 2;
// per-tile addressing of one thread (plain convs only: up = 1): step st = (column group st / MS, sub-tile st % MS) sits at
// off0 + (st / MS) * 2 * gstride + (st % MS) * 1024 elements and is inside the utterance while (st % MS) * 128 < rows_left
struct ResTile { long off0, gstride; int rows_left, cc0; };
__device__ __forceinline__ ResTile epi_res_tile(const TcArgs& a, const int NT, const int n0, const int q0, const long boff, const int quad,
                                                const int half, const int lane) {
    ResTile t;
    const int q = q0 + quad * 32 + lane;
    t.cc0 = half ? NT / 2 : 0;
    t.gstride = (long)a.y_T * 8;
    t.off0 = boff + (long)((n0 + t.cc0) >> 3) * t.gstride + (long)q * 8;
    t.rows_left = min(a.Q, a.y_T) - q;
    return t;
}
template <int MS>
__device__ __forceinline__ void epi_res_load(const ResTile& t, const bf16* R, const int st, uint4 (&r)[2]) {
    const bf16* p = R + t.off0 + (long)(st / MS) * 2 * t.gstride + (st % MS) * 1024;
    r[0] = r[1] = make_uint4(0u, 0u, 0u, 0u);
    if ((st % MS) * 128 < t.rows_left) { r[0] = ld_nc(reinterpret_cast<const uint4*>(p)); r[1] = ld_nc(reinterpret_cast<const uint4*>(p + t.gstride)); }
}
template <int NT, int MS, bool HAS_R2>
__device__ __forceinline__ void epi_res_run(const TcArgs& a, const ResTile& t, const uint32_t acc_full_bar, const uint32_t parity,
                                            const uint32_t tacc, const float* bias_s, const int n0) {
    static_assert((NT / 32) * MS == RES_STEPS, "eight steps per thread");
    constexpr int RES_D1 = HAS_R2 ? 2 : 4;                  // 32 registers of residual rows in flight either way
    uint4 r1[RES_STEPS][2], r2[RES_STEPS][2];
#pragma unroll
    for (int st = 0; st < RES_D1; ++st) epi_res_load<MS>(t, reinterpret_cast<const bf16*>(a.R1), st, r1[st]);
    if (HAS_R2) {
#pragma unroll
        for (int st = 0; st < RES_D2; ++st) epi_res_load<MS>(t, reinterpret_cast<const bf16*>(a.R2), st, r2[st]);
    }
    mbar_wait_relaxed(acc_full_bar, parity);                // residual rows in flight while the tile's MMAs finish
    tc_fence_after();
    const bool lrelu_out = a.out_act == ACT_LRELU, other_act = a.out_act != ACT_NONE && a.out_act != ACT_LRELU;
    const bool scaled = a.out_scale != 1.f, post = a.post_slope != 1.f;
    const Slope2 sl_post = make_slope2(a.post_slope);
#pragma unroll
    for (int st = 0; st < RES_STEPS; ++st) {
        const int cc = t.cc0 + (st / MS) * 16, ms = st % MS;
        float v[16];
        tmem_ld16(tacc + (uint32_t)(ms * NT + cc), v);
        const float4* b4 = reinterpret_cast<const float4*>(bias_s + n0 + cc);
        const float4 ba = b4[0], bb = b4[1], bc = b4[2], bd = b4[3];
        add_f32x2(v[0], v[1], ba.x, ba.y); add_f32x2(v[2], v[3], ba.z, ba.w); add_f32x2(v[4], v[5], bb.x, bb.y); add_f32x2(v[6], v[7], bb.z, bb.w);
        add_f32x2(v[8], v[9], bc.x, bc.y); add_f32x2(v[10], v[11], bc.z, bc.w); add_f32x2(v[12], v[13], bd.x, bd.y); add_f32x2(v[14], v[15], bd.z, bd.w);
        if (other_act) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = act_apply(v[i], a.out_act, a.out_slope);
        }
        if (lrelu_out) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], v[i] * a.out_slope);
        }
        add_bf16x8(v, r1[st][0]); add_bf16x8(v + 8, r1[st][1]);
        if (HAS_R2) { add_bf16x8(v, r2[st][0]); add_bf16x8(v + 8, r2[st][1]); }
        if (st + RES_D1 < RES_STEPS) epi_res_load<MS>(t, reinterpret_cast<const bf16*>(a.R1), st + RES_D1, r1[st + RES_D1]);
        if (HAS_R2 && st + RES_D2 < RES_STEPS) epi_res_load<MS>(t, reinterpret_cast<const bf16*>(a.R2), st + RES_D2, r2[st + RES_D2]);
        if (scaled) {
#pragma unroll
            for (int i = 0; i < 16; i += 2) mul_f32x2(v[i], v[i + 1], a.out_scale, a.out_scale);
        }
        uint4 pa, pb;
        if (a.y_f16) { pa = pack8h(v); pb = pack8h(v + 8); } else { pa = pack8(v); pb = pack8(v + 8); }
        if (post) {
            pa.x = lrelu_bf16x2(pa.x, sl_post); pa.y = lrelu_bf16x2(pa.y, sl_post); pa.z = lrelu_bf16x2(pa.z, sl_post); pa.w = lrelu_bf16x2(pa.w, sl_post);
            pb.x = lrelu_bf16x2(pb.x, sl_post); pb.y = lrelu_bf16x2(pb.y, sl_post); pb.z = lrelu_bf16x2(pb.z, sl_post); pb.w = lrelu_bf16x2(pb.w, sl_post);
        }
        if (ms * 128 < t.rows_left) {
            bf16* p = a.Y16 + t.off0 + (long)(st / MS) * 2 * t.gstride + ms * 1024;
            *reinterpret_cast<uint4*>(p) = pa;
            *reinterpret_cast<uint4*>(p + t.gstride) = pb;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Fused LayerNorm epilogue: the CTA owns whole 256-column rows (NT = N = 256, one 128-row accumulator), so
// fc / w_2 (+ bias + residual) -> LayerNorm -> pad-row zeroing (SubLayers.py:54-55,88-91 + Layers.py:25,28) and the variance
// predictors' conv -> ReLU -> LayerNorm (modules.py:209-240) never leave the SM between the GEMM and the normalised output.
// Thread = (row = TMEM lane, column half).  Pass 1 adds bias, activation and the residual row, parks the result back in TMEM
// (tcgen05.st) and sums it; pass 2 re-reads it for the centred second moment (the two-pass form nn.LayerNorm uses); pass 3
// normalises and stores fp32 row-major and, optionally, the planar fp16 copy the next fp16 layer consumes.  The two column
// halves of a row meet through 2 KB of shared memory and a 64-thread named barrier per lane quadrant.
// ---------------------------------------------------------------------------------------------
// PARTS = 2: the eight epilogue warps (column halves).  PARTS = 4: the eight activation-worker warps join when they have nothing
// to stage (planar input without an input activation: the decoder side) — a row is then shared by four threads, which doubles
// the residual bytes in flight per SM; the epilogue is bound by exactly that (one 64-byte run per thread at ~1 us of latency:
// 16 KB per SM = 2.4 TB/s chip-wide; what-if timing: fc 26 us without its loads and stores, 69 us with them).
template <int PARTS>
__device__ __forceinline__ void epi_layernorm(const TcArgs& a, const uint32_t trow, float* smem_f, const int q, const long boff,
                                              const int b_tile, const int quad, const int part, const int lane) {
    constexpr int CW = 256 / PARTS;                        // columns per thread
    const float* bias_s = smem_f;                          // [256]
    const float* g_s = smem_f + 256; const float* b_s = smem_f + 512;
    float* psum = smem_f + 1024;                           // [4 parts][128 rows]
    float* psq = smem_f + 1536;                            // [4 parts][128 rows]
    const int c0 = part * CW, r = quad * 32 + lane;
    const bool rowok = q < a.Q;
    const float* R1f = reinterpret_cast<const float*>(a.R1);
    const long e0 = boff + (long)q * 256 + c0;
    float sum = 0.f;
#pragma unroll 1
    for (int cc = 0; cc < CW; cc += 16) {
        float v[16];
        float4 rr[4];
        if (a.R1) {
#pragma unroll
            for (int h = 0; h < 4; ++h) rr[h] = (rowok && !(a.dbg & 256)) ? *reinterpret_cast<const float4*>(R1f + e0 + cc + h * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        tmem_ld16(trow + (uint32_t)(c0 + cc), v);
        epi_bias_act(a, v, bias_s + c0 + cc);
        if (a.R1) {
#pragma unroll
            for (int h = 0; h < 4; ++h) { v[h * 4] += rr[h].x; v[h * 4 + 1] += rr[h].y; v[h * 4 + 2] += rr[h].z; v[h * 4 + 3] += rr[h].w; }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) sum += v[i];
        tmem_st16(trow + (uint32_t)(c0 + cc), v);
    }
    psum[part * 128 + r] = sum;
    asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(PARTS * 32) : "memory");
    float tot = 0.f;
#pragma unroll
    for (int p = 0; p < PARTS; ++p) tot += psum[p * 128 + r];                  // same order in every thread of the row
    const float mean = tot * (1.f / 256.f);
    float sq = 0.f;
#pragma unroll 1
    for (int cc = 0; cc < CW; cc += 16) {
        float v[16];
        tmem_ld16(trow + (uint32_t)(c0 + cc), v);
#pragma unroll
        for (int i = 0; i < 16; ++i) { const float d = v[i] - mean; sq += d * d; }
    }
    psq[part * 128 + r] = sq;
    asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(PARTS * 32) : "memory");
    float qt = 0.f;
#pragma unroll
    for (int p = 0; p < PARTS; ++p) qt += psq[p * 128 + r];
    const float rstd = rsqrtf(qt * (1.f / 256.f) + 1e-5f);
    // row -> (utterance, frame): k = 1 layers run flattened over the B * T rows
    int bq = b_tile, tq = q;
    if (a.ln_T > 0 && a.B == 1) { bq = q / a.ln_T; tq = q - bq * a.ln_T; }
    const bool keep = a.ln_lens == nullptr || (rowok && tq < (int)a.ln_lens[bq]);
    __half* p16 = a.Y16p ? reinterpret_cast<__half*>(a.Y16p) + (((long)bq * 32 + (c0 >> 3)) * a.ln_T + tq) * 8 : nullptr;
    const long gs16 = (long)a.ln_T * 8;
#pragma unroll 1
    for (int cc = 0; cc < CW; cc += 16) {
        float v[16];
        tmem_ld16(trow + (uint32_t)(c0 + cc), v);
        if (!rowok || (a.dbg & 512)) continue;
        const float4* g4 = reinterpret_cast<const float4*>(g_s + c0 + cc); const float4* b4 = reinterpret_cast<const float4*>(b_s + c0 + cc);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const float4 gg = g4[h], bb = b4[h];
            v[h * 4 + 0] = keep ? (v[h * 4 + 0] - mean) * rstd * gg.x + bb.x : 0.f;
            v[h * 4 + 1] = keep ? (v[h * 4 + 1] - mean) * rstd * gg.y + bb.y : 0.f;
            v[h * 4 + 2] = keep ? (v[h * 4 + 2] - mean) * rstd * gg.z + bb.z : 0.f;
            v[h * 4 + 3] = keep ? (v[h * 4 + 3] - mean) * rstd * gg.w + bb.w : 0.f;
        }
        float4* p = reinterpret_cast<float4*>(a.Y + e0 + cc);
#pragma unroll
        for (int h = 0; h < 4; ++h) p[h] = make_float4(v[h * 4], v[h * 4 + 1], v[h * 4 + 2], v[h * 4 + 3]);
        if (p16) {
            *reinterpret_cast<uint4*>(p16 + (long)(cc >> 3) * gs16) = pack8h(v);
            *reinterpret_cast<uint4*>(p16 + (long)((cc >> 3) + 1) * gs16) = pack8h(v + 8);
        }
    }
}

// PAIR: the kernel runs as clusters of two CTAs (launch attribute) whose MMAs are tcgen05 cta_group::2 instructions issued by the
// leader CTA (M = 256: 128 rows of A from each CTA's own tile; the NT weight columns split over the pair, NT / 2 per CTA) — the
// scheme of respair2_tc.cu.  A single-CTA N = 128 MMA reads 8 KB of operands per 64 cycles, all of the shared-memory bandwidth, so
// the weight ring's and the activation copies' writes slow the MMAs down; in pair mode each SM reads 6 KB per MMA and streams half
// of the weight bytes.  The two CTAs of a cluster take consecutive row blocks of the same column tile.  Barriers the issuers wait
// on live in the leader and count both CTAs' arrivals; the ones they signal are multicast commits.  Planar input only (TMA_IN).
template <int NT, bool TMA_IN, bool PAIR = false>
__global__ void __launch_bounds__(NTHREADS, 1) k_conv_tc(const TcArgs a) {
    static_assert(!PAIR || TMA_IN, "pair mode takes planar input");
    extern __shared__ __align__(1024) uint8_t smem[];
    // header: mbarriers + TMEM base; then the bias row, the activation slots and the weight ring
    const Bars bars{smem_u32(smem)};
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 512);
    float* bias_s = reinterpret_cast<float*>(smem + SMEM_HDR);
    uint8_t* A_smem = smem + SMEM_HDR + BIAS_MAX * 4;
    uint8_t* W_smem = A_smem + A_SLOTS * (size_t)a.a_bytes;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_chunks = a.Cin / a.KC;
    const int n_stages = (a.n_blobs + a.G - 1) / a.G;
    const int acc_cols = a.msub * NT;                      // TMEM columns of one accumulator set (two sets)
    const bool a_workers = PAIR || !TMA_IN || a.in_act != ACT_NONE;   // pair mode: the workers also forward "tile landed" to the leader
    const bool ln_helpers = NT == 256 && a.ln_g != nullptr && !a_workers && !(a.dbg & 1024);   // idle worker warps join the LayerNorm epilogue
    const int n_issuers = min(a.msub, MMA_WARPS);
    const int n_copiers = TMA_IN ? max(1, min(a.copiers, 1 + MMA_WARPS - max(n_issuers, 1))) : 1;   // copy warp + idle MMA warps
    // work items: tiles, or (pair mode) tile pairs: pair p = (row-block pair p / n_ntiles, column tile p % n_ntiles), rank r takes
    // row block 2 (p / n_ntiles) + r; an odd row-block count leaves the last peer tile empty (nothing loaded or stored)
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const int w_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, w_stride = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int w_total = PAIR ? a.total_pairs : a.total_tiles;
    auto tile_of = [&](int wi) { return PAIR ? (2 * (wi / a.n_ntiles) + (int)rank) * a.n_ntiles + wi % a.n_ntiles : wi; };

    if (threadIdx.x == 0) {
        for (int i = 0; i < A_SLOTS; ++i) {
            mbar_init(bars.a_full(i), PAIR ? 2 * A_WARPS : (a_workers ? A_THREADS : n_copiers)); mbar_init(bars.a_empty(i), n_issuers); mbar_init(bars.a_raw(i), n_copiers);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bars.acc_full(i), n_issuers);
            mbar_init(bars.acc_empty(i), PAIR ? 2 * EPI_WARPS : (ln_helpers ? EPI_WARPS + A_WARPS : EPI_WARPS));
        }
        for (int i = 0; i < W_STAGES; ++i) { mbar_init(bars.w_full(i), 1); mbar_init(bars.w_empty(i), n_issuers); mbar_init(bars.w_peer(i), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < a.N; i += NTHREADS) bias_s[i] = a.bias ? a.bias[i % a.bias_mod] : 0.f;
    if (NT == 256 && a.ln_g) for (int i = threadIdx.x; i < 256; i += NTHREADS) { bias_s[256 + i] = a.ln_g[i]; bias_s[512 + i] = a.ln_b[i]; }
    if (warp == MMA_WARP) {
        if (PAIR) {                                           // the same warp of both CTAs: the pair's TMEM is allocated together
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(a.tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(a.tmem_cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();                           // both CTAs' barriers are initialised before anyone arrives remotely
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // programmatic dependent launch: everything above touched constants only; the weight warp (constants) and the issuers (shared
    // memory / TMEM only) run ahead, every role that reads or writes activations waits for the previous grid of the stream
    pdl_trigger();
    if (warp != W_WARP && !(warp >= MMA_WARP && warp < MMA_WARP + MMA_WARPS)) pdl_wait();

    // ---- activation copies (planar input): copier ci of n_copiers moves the 8-channel groups ci, ci + n_copiers, ... of every chunk.
    // Copier 0 is the copy warp; the others are the MMA warps that issue nothing in this launch (msub = 1: three, msub = 2: two).
    // Why: a cp.async.bulk costs its issuing warp ~200-350 cycles whichever lane it comes from, so the 64 runs of a tile of the first
    // upsampler (Cin = 512) from one warp outlast its 128 MMAs.  Used only there (see the launch code).
    auto copy_role = [&](const int ci) {
        if (ci > 0) pdl_wait();
        int slot = 0, it = 0; uint32_t phase = 0;
        const bool tr = (a.dbg & 32) && blockIdx.x == 0 && lane == 0 && ci == 0;
        const int cgpc = a.KC >> 3;
        const int ng = ci < cgpc ? (cgpc - ci + n_copiers - 1) / n_copiers : 0;       // this copier's groups per chunk
        const uint8_t* X8 = reinterpret_cast<const uint8_t*>(a.X);
        const bool res_pf = ci == 0 && a.res_pf && a.y_planar && a.R1 != nullptr && a.r_bf16 && a.up == 1 && a.up_pad == 0;
        for (int wi = w_first; wi < w_total; wi += w_stride, ++it) {
            const int tile = tile_of(wi);
            const bool valid = !PAIR || tile < a.total_tiles;
            const int rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
            const int q_lo = rt * a.msub * 128 + a.off_lo;
            for (int kc = 0; kc < n_chunks; ++kc) {
                if (kc == 0) TRACE(tr, 0, it, 0);
                if (lane == 0) STAT_WAIT((a.dbg & 16) && ci == 0, 6, mbar_wait_relaxed(bars.a_empty(slot), phase ^ 1));
                __syncwarp();
                if (kc == 0) TRACE(tr, 0, it, 1);
                const uint32_t bar = (PAIR || a.in_act != ACT_NONE) ? bars.a_raw(slot) : bars.a_full(slot);
                uint8_t* dst_g = A_smem + (size_t)slot * a.a_bytes;
                const uint32_t dst = smem_u32(dst_g);
                // the tile is ONE [KC/8][Rp][8] block (its msub*128 rows + the halo): one contiguous run per 8-channel group
                const int t0 = q_lo, lo = max(t0, 0), hi = valid ? max(min(t0 + a.Rp, a.Tin), lo) : lo;
                const uint32_t run = (uint32_t)(hi - lo) * 16u;
                if (valid && (lo > t0 || hi < t0 + a.Rp)) {
                    for (int g = ci + n_copiers * lane; g < cgpc; g += 32 * n_copiers) {
                        uint4* blk = reinterpret_cast<uint4*>(dst_g + (size_t)g * a.Rp * 16);
                        for (int r = 0; r < min(lo - t0, a.Rp); ++r) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                        for (int r = max(hi - t0, 0); r < a.Rp; ++r) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                    }
                    fence_proxy_async();
                    __syncwarp();
                }
                if (lane == 0) {
                    if (run && ng) mbar_expect_tx(bar, run * (uint32_t)ng); else mbar_arrive(bar);
                }
                __syncwarp();
                if (run) {
                    for (int g = ci + n_copiers * lane; g < cgpc; g += 32 * n_copiers) {
                        const uint8_t* src = X8 + (((size_t)b * (a.Cin >> 3) + (size_t)(kc * cgpc + g)) * a.Tin + lo) * 16;
                        bulk_g2s(dst + (uint32_t)((g * a.Rp + (lo - t0)) * 16), src, run, bar);
                    }
                }
                if (kc == 0) TRACE(tr, 0, it, 2);
                if (++slot == A_SLOTS) { slot = 0; phase ^= 1; }
            }
            if (res_pf && valid) {
                // residual rows of this tile -> L2, one run per 8-channel group, a few tiles before the epilogue warps read them.
                // The activation copies above keep ~100 KB per SM (15 MB chip-wide) queued at the memory controllers, so a
                // residual load that has to go to DRAM waits 2.5-5 us behind them and the epilogue's two load rounds per tile
                // (16.8 kcycles at C = 64 k = 11) outlast the tile's MMAs (12.5 kcycles); from L2 they do not.
                const int nt = tile % a.n_ntiles, q0 = rt * a.msub * 128;
                const int rows = min(a.msub * 128, min(a.Q, a.y_T) - q0);
                if (rows > 0) {
                    const long gs = (long)a.y_T * 8;
                    const long off = (long)b * a.y_bs + (long)((nt * NT) >> 3) * gs + (long)q0 * 8;
                    for (int g = lane; g < NT / 8; g += 32) {
                        bulk_prefetch_l2(reinterpret_cast<const bf16*>(a.R1) + off + (long)g * gs, (uint32_t)rows * 16u);
                        if (a.R2) bulk_prefetch_l2(reinterpret_cast<const bf16*>(a.R2) + off + (long)g * gs, (uint32_t)rows * 16u);
                    }
                }
            }
        }
    };

    // tile id -> (batch b, row tile rt, column tile nt); a CTA walks ids blockIdx.x, blockIdx.x + gridDim.x, ...
    if (warp >= A_WARP0) {
        const int tid = threadIdx.x - A_WARP0 * 32;
        if (a.dbg & 2) {
            // what-if: no activation staging
        } else if (TMA_IN) {
            if (NT == 256 && ln_helpers) {
                // ===== nothing to stage: second half of the fused-LayerNorm epilogue (column quarters 2 and 3 of every row) =====
                const int hw = warp - A_WARP0, quad = warp & 3;
                // two worker warps share each TMEM lane quadrant: the one met first takes quarter 2, the other quarter 3
                int seen = 0;
                for (int k = 0; k < hw; ++k) if (((A_WARP0 + k) & 3) == quad) ++seen;
                const int my_part = 2 + seen;
                int it = 0;
                for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
                    const int rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
                    const int accb = it & 1;
                    mbar_wait_relaxed(bars.acc_full(accb), (it >> 1) & 1);
                    tc_fence_after();
                    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols);
                    epi_layernorm<4>(a, trow, bias_s, rt * a.msub * 128 + quad * 32 + lane, (long)b * a.y_bs, b, quad, my_part, lane);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bars.acc_empty(accb));
                }
            }
            // ===== workers, planar input: leaky-relu in place on the slot the TMA unit has filled (pair mode: also without an
            // activation, to tell the leader CTA that this CTA's tile has landed — one remote arrival per warp) =====
            if (PAIR || a.in_act != ACT_NONE) {
                const int pieces = a.in_act != ACT_NONE ? a.a_tx >> 4 : 0;
                const Slope2 sl = make_slope2(a.in_slope);
                int slot = 0, it = 0; uint32_t phase = 0;
                const bool tr = (a.dbg & 32) && blockIdx.x == 0 && tid == 0;
                for (int tile = w_first; tile < w_total; tile += w_stride, ++it) {
                    for (int kc = 0; kc < n_chunks; ++kc) {
                        STAT_WAIT((a.dbg & 16) && tid == 0, 8, mbar_wait_relaxed(bars.a_raw(slot), phase));
                        if (kc == 0) TRACE(tr, 1, it, 0);
                        uint4* p = reinterpret_cast<uint4*>(A_smem + (size_t)slot * a.a_bytes);
                        for (int i = tid; i < pieces; i += A_THREADS) {                // one piece per trip on purpose (respair_tc.cu)
                            uint4 v = p[i];
                            v.x = lrelu_bf16x2(v.x, sl); v.y = lrelu_bf16x2(v.y, sl);
                            v.z = lrelu_bf16x2(v.z, sl); v.w = lrelu_bf16x2(v.w, sl);
                            p[i] = v;
                        }
                        fence_proxy_async();                  // generic-proxy stores -> visible to the tensor core (async proxy)
                        if (PAIR) { __syncwarp(); if (lane == 0) mbar_arrive_cluster(map_to_rank(bars.a_full(slot), 0)); }
                        else mbar_arrive(bars.a_full(slot));
                        if (kc == 0) TRACE(tr, 1, it, 1);
                        if (++slot == A_SLOTS) { slot = 0; phase ^= 1; }
                    }
                }
            }
        } else {
            // ===== workers, row-major fp32 input: stage rows [q0 + off_lo, q0 + off_lo + rows) x KC channels per chunk =====
            const int cpr = a.KC >> 3;                       // 16-byte pieces per row
            const int total = cpr * a.rows;
            const int dr = A_THREADS / cpr, dc = A_THREADS % cpr;
            int slot = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x) {
                const int rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
                const uint8_t* Xb = reinterpret_cast<const uint8_t*>(a.X) + (size_t)b * a.x_bs * 4;
                const int q_lo = rt * a.msub * 128 + a.off_lo;
                for (int kc = 0; kc < n_chunks; ++kc) {
                    mbar_wait_relaxed(bars.a_empty(slot), phase ^ 1);
                    uint8_t* dst = A_smem + (size_t)slot * a.a_bytes;
                    const int ch0 = kc * a.KC;
                    int r = tid / cpr, cc = tid % cpr;
                    constexpr int U = 4;                       // rows in flight per thread (two 128-bit loads each)
                    const int tmax = a.Tin - 1;
                    for (int i = tid; i < total; i += U * A_THREADS) {
                        uint4 raw[U][2]; int rr[U], cs[U]; bool ok[U];
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            rr[u] = r; cs[u] = cc;
                            const int t = q_lo + r;
                            ok[u] = t >= 0 && t <= tmax;
                            const int tcl = min(max(t, 0), tmax);          // clamped: the load is always legal, the value is dropped when !ok
                            const uint4* p = reinterpret_cast<const uint4*>(Xb + ((size_t)tcl * a.x_ld + ch0 + cc * 8) * 4);
                            raw[u][0] = ld_nc(p);
                            raw[u][1] = ld_nc(p + 1);
                            r += dr; cc += dc;
                            if (cc >= cpr) { cc -= cpr; ++r; }
                        }
#pragma unroll
                        for (int u = 0; u < U; ++u)
                            if (i + u * A_THREADS < total) {
                                uint4 vl;
                                const uint4 v = convert_piece(a, raw[u][0], raw[u][1], ok[u], vl);
                                const size_t o = ((size_t)cs[u] * a.Rp + rr[u]) * 16;
                                *reinterpret_cast<uint4*>(dst + o) = v;
                                if (a.split) *reinterpret_cast<uint4*>(dst + a.a_half + o) = vl;
                            }
                    }
                    fence_proxy_async();                      // generic-proxy stores -> visible to the tensor core (async proxy)
                    mbar_arrive(bars.a_full(slot));
                    if (++slot == A_SLOTS) { slot = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == T_WARP) {
        // ===== activation copy warp (planar input).  In the planar layout the rows [t, t + Rb) of one 8-channel group are
        // one contiguous run of Rb*16 bytes, so every (sub-tile, channel group) block is a single 1-D bulk async copy
        // (cp.async.bulk -> UBLKCP), one per lane.  (A 4-D tensor-map copy of the same box has a 16-byte innermost
        // dimension and was measured at ~16 B/clk/SM, 3.9 TB/s chip-wide.)  Rows outside [0, Tin) — the conv zero padding,
        // only in the first / last tile of an utterance — are zero-filled by the lane before the copies are issued. =====
        if (TMA_IN && !(a.dbg & 2)) copy_role(0);
    } else if (warp == W_WARP) {
        // ===== weight producer: 1-D bulk copies of pre-tiled blobs =====
        if (lane == 0 && !(a.dbg & 1)) {
            int slot = 0; uint32_t phase = 0;
            for (int wi = w_first; wi < w_total; wi += w_stride) {
                const int nt = wi % a.n_ntiles;                      // (tile % n_ntiles; in pair mode both CTAs of a pair share the column tile)
                // pair mode: the blobs are tiled NT / 2 columns wide, column tile nt = blob sets 2 nt (leader) and 2 nt + 1 (peer)
                const uint8_t* src = reinterpret_cast<const uint8_t*>(a.Wt) + (size_t)(PAIR ? 2 * nt + (int)rank : nt) * a.n_blobs * a.blob_bytes;
                if (a.w_resident && wi != w_first) {                 // resident weights: only the hand-shake, no copy
                    mbar_wait_relaxed(bars.w_empty(0), phase ^ 1);
                    mbar_arrive(bars.w_full(0));
                    phase ^= 1;
                    continue;
                }
                for (int st = 0; st < n_stages; ++st) {
                    STAT_WAIT(a.dbg & 16, 7, mbar_wait_relaxed(bars.w_empty(slot), phase ^ 1));
                    const int nb = min(a.G, a.n_blobs - st * a.G);
                    const uint32_t bytes = (uint32_t)nb * a.blob_bytes;
                    mbar_expect_tx(bars.w_full(slot), bytes);
                    bulk_g2s(smem_u32(W_smem + (size_t)slot * a.stage_bytes), src + (size_t)st * a.G * a.blob_bytes, bytes, bars.w_full(slot));
                    if (++slot == a.w_stages) { slot = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= MMA_WARP && warp < MMA_WARP + MMA_WARPS) {
        // ===== MMA issuers: one thread per warp; specialised on the number of accumulator sub-tiles =====
        const int w = warp - MMA_WARP;
        if (PAIR && rank != 0) {
            // peer CTA: the leader issues every MMA of the pair; one thread here tells it when each weight stage of THIS CTA has landed
            if (w == 0 && lane == 0 && !(a.dbg & 1)) {
                int slot = 0; uint32_t phase = 0;
                for (int wi = w_first; wi < w_total; wi += w_stride) {
                    const int ns = (a.w_resident && wi != w_first) ? 1 : n_stages;     // the weight warp's sequence of w_full phases
                    for (int st = 0; st < ns; ++st) {
                        mbar_wait(bars.w_full(slot), phase);
                        mbar_arrive_cluster(map_to_rank(bars.w_peer(slot), 0));
                        if (++slot == a.w_stages) { slot = 0; phase ^= 1; }
                    }
                }
            }
        } else if (w < n_issuers) {
            const uint32_t as = smem_u32(A_smem), ws = smem_u32(W_smem);
#define MASG_ROLE(MSV, NIV, SPL, WV) mma_role<NT, MSV, NIV, SPL, WV, PAIR>(a, bars, as, ws, tmem_base)
            if (a.split) {
                if (a.msub == 1) MASG_ROLE(1, 1, true, 0);
                else if (a.msub == 2 && NT <= 128) { if (w == 0) MASG_ROLE(2, 2, true, 0); else MASG_ROLE(2, 2, true, 1); }
            } else if (a.msub == 1) {
                MASG_ROLE(1, 1, false, 0);
            } else if (a.msub == 2 && NT <= 128) {
                if (w == 0) MASG_ROLE(2, 2, false, 0); else MASG_ROLE(2, 2, false, 1);
            } else if (a.msub == 4 && NT <= 64) {
                switch (w) { case 0: MASG_ROLE(4, 4, false, 0); break; case 1: MASG_ROLE(4, 4, false, 1); break;
                             case 2: MASG_ROLE(4, 4, false, 2); break; default: MASG_ROLE(4, 4, false, 3); break; }
            } else if (a.msub == 8 && NT <= 32) {
                switch (w) { case 0: MASG_ROLE(8, 4, false, 0); break; case 1: MASG_ROLE(8, 4, false, 1); break;
                             case 2: MASG_ROLE(8, 4, false, 2); break; default: MASG_ROLE(8, 4, false, 3); break; }
            }
#undef MASG_ROLE
        }
        // MMA warps with nothing to issue help the copy warp (see copy_role)
        if (TMA_IN && !(a.dbg & 2) && w >= max(n_issuers, 1) && w - max(n_issuers, 1) + 1 < n_copiers) copy_role(w - max(n_issuers, 1) + 1);
        __syncwarp();
    } else {
        // ===== epilogue (8 warps): TMEM -> registers -> bias / activation / residual / scale -> global =====
        const int quad = warp & 3, half = warp >> 2;
        const int c_split = ((NT / 16 + 1) / 2) * 16;
        const int c_begin = half ? c_split : 0, c_end = half ? NT : c_split;
        const float* R1f = reinterpret_cast<const float*>(a.R1); const bf16* R1h = reinterpret_cast<const bf16*>(a.R1);
        const float* R2f = reinterpret_cast<const float*>(a.R2); const bf16* R2h = reinterpret_cast<const bf16*>(a.R2);
        constexpr bool RES_FAST = NT == 32 || NT == 64 || NT == 128;
        const bool res_fast = RES_FAST && a.y_planar && a.R1 != nullptr && a.up == 1 && a.up_pad == 0 && a.msub * NT == 256 && !(a.dbg & 64);
        int it = 0;
        const long long t_epi0 = clock64();
        auto release_acc = [&](int accb) {                  // this warp has read its part of the accumulator set
            tc_fence_before();                              // order this warp's tcgen05.ld before the release below
            __syncwarp();
            if (lane == 0) { if (PAIR) mbar_arrive_cluster(map_to_rank(bars.acc_empty(accb), 0)); else mbar_arrive(bars.acc_empty(accb)); }
        };
        for (int wi = w_first; wi < w_total; wi += w_stride, ++it) {
            const int tile = tile_of(wi);
            const int nt = tile % a.n_ntiles, rb = tile / a.n_ntiles, b = rb / a.row_tiles, rt = rb % a.row_tiles;
            const int n0 = nt * NT, q0 = rt * a.msub * 128;
            const int accb = it & 1;
            const bool tr = (a.dbg & 32) && blockIdx.x == 0 && threadIdx.x == 0;
            TRACE(tr, 3, it, 0);
            if (PAIR && tile >= a.total_tiles) {            // the empty peer tile of an odd row-block count
                mbar_wait_relaxed(bars.acc_full(accb), (it >> 1) & 1);
                tc_fence_after();
                release_acc(accb);
                continue;
            }
            const long boff = (long)b * a.y_bs;
            if constexpr (RES_FAST) {
                if (res_fast) {
                    const ResTile rtile = epi_res_tile(a, NT, n0, q0, boff, quad, half, lane);
                    const uint32_t tacc = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols);
                    if (a.R2 == nullptr) epi_res_run<NT, 256 / NT, false>(a, rtile, bars.acc_full(accb), (it >> 1) & 1, tacc, bias_s, n0);
                    else epi_res_run<NT, 256 / NT, true>(a, rtile, bars.acc_full(accb), (it >> 1) & 1, tacc, bias_s, n0);
                    TRACE(tr, 3, it, 1);
                    release_acc(accb);
                    TRACE(tr, 3, it, 2);
                    continue;
                }
            }
            STAT_WAIT((a.dbg & 16) && threadIdx.x == 0, 4, mbar_wait_relaxed(bars.acc_full(accb), (it >> 1) & 1));
            tc_fence_after();
            TRACE(tr, 3, it, 1);
            if (a.dbg & 4) {
                // what-if: no epilogue work
            } else if (NT == 256 && a.ln_g) {
                const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols);
                if (ln_helpers) epi_layernorm<4>(a, trow, bias_s, q0 + quad * 32 + lane, boff, b, quad, half, lane);
                else epi_layernorm<2>(a, trow, bias_s, q0 + quad * 32 + lane, boff, b, quad, half, lane);
            } else if (a.y_planar) {
                const uint32_t tacc = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols);
                if (a.R1 == nullptr) epi_planar<NT, false, false>(a, tacc, bias_s, n0, q0, boff, quad, half, lane);
                else if (a.R2 == nullptr) epi_planar<NT, true, false>(a, tacc, bias_s, n0, q0, boff, quad, half, lane);
                else epi_planar<NT, true, true>(a, tacc, bias_s, n0, q0, boff, quad, half, lane);
            } else {
                // ---- row-major output (fp32 and/or bf16): lane = row, 16 consecutive columns per store group ----
                const int gc = ((a.R1 || a.R2) && !a.r_bf16) ? 16 : 32;     // columns per residual prefetch group (4 x 16 bytes)
                for (int ms = 0; ms < a.msub; ++ms) {
                    const int q = q0 + ms * 128 + quad * 32 + lane;
                    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(accb * acc_cols + ms * NT);
                    for (int cg = c_begin; cg < c_end; cg += gc) {
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
                                    epi_bias_act(a, v, bias_s + n0 + cg + cc);
                                    if (a.R1 || a.R2) {
                                        if (a.r_bf16) {
#pragma unroll
                                            for (int h = 0; h < 2; ++h) { add_bf16x8(v + h * 8, r1[(k16 * 2 + h) & 3]); add_bf16x8(v + h * 8, r2[(k16 * 2 + h) & 3]); }
                                        } else {
#pragma unroll
                                            for (int h = 0; h < 4; ++h) {
                                                const uint4 u = r1[(k16 * 4 + h) & 3], w = r2[(k16 * 4 + h) & 3];
                                                v[h * 4 + 0] += __uint_as_float(u.x) + __uint_as_float(w.x); v[h * 4 + 1] += __uint_as_float(u.y) + __uint_as_float(w.y);
                                                v[h * 4 + 2] += __uint_as_float(u.z) + __uint_as_float(w.z); v[h * 4 + 3] += __uint_as_float(u.w) + __uint_as_float(w.w);
                                            }
                                        }
                                    }
                                    epi_scale_post(a, v);
                                    if (a.Y) {
                                        float4* p = reinterpret_cast<float4*>(a.Y + boff + e);
#pragma unroll
                                        for (int h = 0; h < 4; ++h) p[h] = make_float4(v[h * 4], v[h * 4 + 1], v[h * 4 + 2], v[h * 4 + 3]);
                                    }
                                    if (a.Y16) {
                                        uint4* p = reinterpret_cast<uint4*>(a.Y16 + boff + e);
                                        p[0] = pack8(v); p[1] = pack8(v + 8);
                                    }
                                }
                            }
                        }
                    }
                }
            }
            release_acc(accb);
            TRACE(tr, 3, it, 2);
        }
        if ((a.dbg & 16) && threadIdx.x == 0) atomicAdd(&g_tc_stats[5], (unsigned long long)(clock64() - t_epi0));
    }
    tc_fence_before();
    __syncthreads();
    if (PAIR) cluster_sync_all();                           // neither CTA's shared memory / TMEM goes away while the pair's MMAs can still touch it
    if (warp == MMA_WARP) {
        __syncwarp();
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols) : "memory");
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
    if (split == 2) { reinterpret_cast<__half*>(out)[i] = __float2half_rn(w); return; }      // fp16 single-pass layers
    bf16 hi = __float2bfloat16_rn(w);
    if (!split) { out[i] = hi; return; }
    long blob = i / ((long)KW * NT), within = i % ((long)KW * NT);
    out[blob * 2 * KW * NT + within] = hi;
    out[blob * 2 * KW * NT + (long)KW * NT + within] = __float2bfloat16_rn(w - __bfloat162float(hi));
}

// row-major fp32 [B][T][C] <-> planar bf16 [B][C/8][T][8] (the HiFi-GAN activation layout); one 16-byte piece per thread
__global__ void k_to_planar(const float* __restrict__ x, bf16* __restrict__ out, int T, int C, long total_pieces) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total_pieces) return;
    const int cg = C >> 3;
    long t = i % T; long r = i / T; int g = (int)(r % cg); long b = r / cg;
    const float4* p = reinterpret_cast<const float4*>(x + ((b * T + t) * C + g * 8));
    const float4 a = p[0], c = p[1];
    reinterpret_cast<uint4*>(out)[i] = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(c.x, c.y), pack_bf16(c.z, c.w));
}
__global__ void k_from_planar(const bf16* __restrict__ x, float* __restrict__ out, int T, int C, long total_pieces) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total_pieces) return;
    const int cg = C >> 3;
    long t = i % T; long r = i / T; int g = (int)(r % cg); long b = r / cg;
    const uint4 u = reinterpret_cast<const uint4*>(x)[i];
    const float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y), f2 = unpack_bf16(u.z), f3 = unpack_bf16(u.w);
    float4* p = reinterpret_cast<float4*>(out + ((b * T + t) * C + g * 8));
    p[0] = make_float4(f0.x, f0.y, f1.x, f1.y); p[1] = make_float4(f2.x, f2.y, f3.x, f3.y);
}

template <int NT, bool TMA_IN>
void launch_inst(const TcArgs& a, dim3 grid, size_t smem, cudaStream_t s, int n_sm) {
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_conv_tc<NT, TMA_IN>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    CUDA_OK(launch_pdl(k_conv_tc<NT, TMA_IN>, grid, NTHREADS, smem, s, a, a.total_tiles < n_sm));
}
// pair mode: clusters of two CTAs (one per SM of a TPC)
template <int NT>
void launch_inst_pair(const TcArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_conv_tc<NT, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(NTHREADS, 1, 1); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CUDA_OK(cudaLaunchKernelEx(&cfg, k_conv_tc<NT, true, true>, a));
}

}  // namespace

void tc_debug_stats(unsigned long long* out16, bool reset) {
    CUDA_OK(cudaMemcpyFromSymbol(out16, g_tc_stats, sizeof(unsigned long long) * 16));
    if (reset) { unsigned long long z[16] = {0}; CUDA_OK(cudaMemcpyToSymbol(g_tc_stats, z, sizeof(z))); }
}

int tc_debug_trace(long long* out, int max_elems, bool reset) {
    const int n = TRACE_ROLES * TRACE_TILES * TRACE_EVTS;
    if (max_elems < n) return -1;
    CUDA_OK(cudaMemcpyFromSymbol(out, g_tc_trace, sizeof(long long) * n));
    if (reset) { static long long z[TRACE_ROLES * TRACE_TILES * TRACE_EVTS]; CUDA_OK(cudaMemcpyToSymbol(g_tc_trace, z, sizeof(z))); }
    return TRACE_TILES;
}

void launch_to_planar(masg_ctx* c, cudaStream_t s, const float* x, bf16* out, int B, int T, int C) {
    MASG_CHECK(C % 8 == 0, "planar layout needs C %% 8 == 0 (C=%d)", C);
    const long n = (long)B * T * (C / 8);
    k_to_planar<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, out, T, C, n);
    c->launches++;
}
void launch_from_planar(masg_ctx* c, cudaStream_t s, const bf16* x, float* out, int B, int T, int C) {
    const long n = (long)B * T * (C / 8);
    k_from_planar<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, out, T, C, n);
    c->launches++;
}

void tc_prepare_weights(masg_ctx* c, cudaStream_t s, ConvDesc& d, int mode) {
    const bool wide = (mode & TC_MODE_WIDE) != 0;          // one 256-column tile per row block: the layers with a fused LayerNorm
    mode &= 0xff;
    const bool split = mode == 1;
    MASG_CHECK(d.w32 != nullptr, "tc_prepare_weights: fp32 GEMM weights missing");
    int NT = 0;
    const int cands[4] = {128, 80, 64, 32};
    for (int i = 0; i < 4 && !NT; ++i) if (d.N % cands[i] == 0) NT = cands[i];
    int KW = (d.Cin % 64 == 0) ? 64 : (d.Cin % 32 == 0 ? 32 : 16);
    if (wide) { MASG_CHECK(d.N == 256 && d.Cin % 32 == 0, "wide tiles need N == 256"); NT = 256; KW = 32; }   // 16 KB blobs
    while (split && NT * KW * 4 > W_STAGE_BYTES && KW > 16) KW /= 2;
    MASG_CHECK(d.Cin % KW == 0, "conv_tc: Cin=%d must be a multiple of 16", d.Cin);
    size_t n = (size_t)d.taps * d.Cin * d.N;
    bf16* w16 = nullptr;
    CUDA_OK(cudaMalloc(&w16, n * sizeof(bf16) * (split ? 2 : 1)));
    c->owned.push_back(w16);
    k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, w16, d.Cin, d.N, d.taps, NT, KW, mode);
    c->launches++;
    d.w16 = w16; d.NT = NT; d.KW = KW; d.split = split ? 1 : 0; d.f16 = mode == 2 ? 1 : 0;
    d.w16s = nullptr; d.NTs = 0; d.KWs = 0;
    static int small = -1;
    if (small < 0) { const char* e = getenv("MASG_TC_SMALL"); small = e ? atoi(e) : 1; }
    if (small && !wide && NT > 32 && d.N % 32 == 0) {
        const int KWs = (d.Cin % 64 == 0) ? 64 : (d.Cin % 32 == 0 ? 32 : 16);
        bf16* ws = nullptr;
        CUDA_OK(cudaMalloc(&ws, n * sizeof(bf16) * (split ? 2 : 1)));
        c->owned.push_back(ws);
        k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, ws, d.Cin, d.N, d.taps, 32, KWs, mode);
        c->launches++;
        d.w16s = ws; d.NTs = 32; d.KWs = KWs;
    }
    d.w16h = nullptr; d.w16h2 = nullptr;
    if (!split && !wide && NT == 128) {
        // pair mode of k_conv_tc: the same blobs 64 columns wide — column tile nt = blob sets 2 nt (leader CTA) and 2 nt + 1 (peer)
        bf16* wh = nullptr;
        CUDA_OK(cudaMalloc(&wh, n * sizeof(bf16)));
        c->owned.push_back(wh);
        // 128 input channels per blob where Cin allows it (16 KB blobs again): 8 K-steps per (tap, slab) block of the issue loop — at
        // 4, one thread issuing a whole 128-row tile (pair mode's msub = 1) spends ~79 cycles per 64-cycle MMA on the block overhead
        static int kw128 = -1;
        if (kw128 < 0) { const char* e = getenv("MASG_TC_PAIR_KW128"); kw128 = e ? atoi(e) : 1; }
        k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, wh, d.Cin, d.N, d.taps, NT / 2, KW, mode);
        c->launches++;
        d.w16h = wh;
        if (kw128 && d.Cin % 128 == 0 && KW < 128) {          // second copy, used by launches that take 128-row tiles (msub = 1)
            bf16* wh2 = nullptr;
            CUDA_OK(cudaMalloc(&wh2, n * sizeof(bf16)));
            c->owned.push_back(wh2);
            k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, wh2, d.Cin, d.N, d.taps, NT / 2, 128, mode);
            c->launches++;
            d.w16h2 = wh2;
        }
    }
    d.w16p = nullptr;
    if (!split && mode == 0 && !wide && d.Cin == d.N && ((d.N == 128 && d.taps == 3) || (d.N == 64 && (d.taps == 11 || d.taps == 7)) || (d.N == 32 && (d.taps == 3 || d.taps == 7 || d.taps == 11)))) {
        // square convs that respair2_tc.cu fuses on a CTA pair: [2 halves][1][taps][Cin/8][N/2][8]
        bf16* wp = nullptr;
        CUDA_OK(cudaMalloc(&wp, n * sizeof(bf16)));
        c->owned.push_back(wp);
        k_tile_w<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d.w32, wp, d.Cin, d.N, d.taps, d.N / 2, d.Cin, 0);
        c->launches++;
        d.w16p = wp;
    }
}

static void launch_conv_tc_tiled(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io);

// Tiling choice per launch: the commit-time tiling (128 / 80 / 64 / 32 / 256 columns), or — when that would put the layer on fewer
// than a third of the SMs (B = 1: a 256-channel HiFi-GAN conv is 10 tiles of 128 columns, each CTA streaming 0.7 MB of weights
// alone) — the 32-column tiling of the same weights: four times the CTAs, a quarter of the weight stream each.  N = 32 MMAs run at
// 40 % of the tensor peak, which does not matter where the launch is bound by weight latency (C2: 20-30 us per launch).
void launch_conv_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io) {
    if (d.w16s && io.Q > 0 && io.B > 0 && !io.ln_g) {
        static int thr = -1;
        if (thr < 0) { const char* e = getenv("MASG_TC_SMALL_TILES"); thr = e ? atoi(e) : c->n_sm / 3; }
        const long tiles = (long)io.B * ((io.Q + 127) / 128) * (d.N / d.NT);
        if (tiles < thr) {
            ConvDesc ds = d;
            ds.w16 = d.w16s; ds.NT = d.NTs; ds.KW = d.KWs;
            launch_conv_tc_tiled(c, s, ds, io);
            return;
        }
    }
    launch_conv_tc_tiled(c, s, d, io);
}

static void launch_conv_tc_tiled(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io) {
    MASG_CHECK(d.w16 != nullptr && d.NT > 0, "conv_tc: bf16 weights not committed");
    MASG_CHECK(io.Y != nullptr || io.Y16 != nullptr, "conv_tc: no output tensor");
    MASG_CHECK(d.N <= BIAS_MAX, "conv_tc: N=%d exceeds the staged bias row (%d)", d.N, BIAS_MAX);
    MASG_CHECK(!io.x_bf16 || io.x_planar, "conv_tc: bf16 activations must be planar [B][C/8][T][8]");
    MASG_CHECK(!d.split || !io.x_planar, "conv_tc: split-bf16 layers take row-major fp32 activations");
    if (io.Q <= 0 || io.B <= 0) return;
    const bool tma_in = io.x_planar != 0;
    TcArgs a;
    memset(&a, 0, sizeof(a));
    a.X = io.X; a.x_bs = io.x_bs; a.x_ld = io.x_ld; a.Tin = io.Tin;
    a.Wt = d.w16; a.bias = d.bias; a.bias_mod = d.bias_mod ? d.bias_mod : d.N;      // (pair mode: a.Wt is replaced below)
    a.Cin = d.Cin; a.N = d.N; a.taps = d.taps; a.off0 = d.off0; a.tstep = d.tstep;
    a.split = d.split; a.f16 = d.f16;                  // (a.KW: below, once pair mode is decided)
    MASG_CHECK(!io.x_planar || !io.in_act || !d.f16, "conv_tc: planar fp16 input has no input activation");
    a.y_f16 = io.y_f16;
    MASG_CHECK(!io.y_f16 || (io.y_planar && io.out_act != ACT_LRELU && io.post_slope == 1.f && !io.R1 && !io.R2),
               "conv_tc: fp16 planar output supports bias + relu / tanh only");
    const int NT = d.NT;
    const size_t a_mul = d.split ? 2 : 1;
    // pair mode (clusters of two CTAs, cta_group::2 MMAs): N = 128 tiles of planar, single-pass layers without a fused LayerNorm,
    // when the launch has at least two tiles per SM (MASG_TC_PAIR = 0: off)
    bool pair = false;
    {
        static int on = -1;
        if (on < 0) { const char* e = getenv("MASG_TC_PAIR"); on = e ? atoi(e) : 1; }
        // (not the transposed-conv upsamplers: two taps, many column tiles — measured 6-20 % slower in pair mode)
        pair = on && NT == 128 && tma_in && !d.split && d.w16h != nullptr && !io.ln_g && io.y_planar && (d.up == 0 || io.pair_mode > 0) &&
               io.pair_mode >= 0 && (io.pair_mode > 0 || (long)io.B * ((io.Q + 255) / 256) * (d.N / NT) >= 2L * c->n_sm);
    }
    const int NTB = pair ? NT / 2 : NT;                    // weight columns per CTA
    int off_lo = d.off0, off_hi = d.off0;
    for (int j = 0; j < d.taps; ++j) { int o = d.off0 + j * d.tstep; off_lo = std::min(off_lo, o); off_hi = std::max(off_hi, o); }
    const int halo = off_hi - off_lo;
    a.off_lo = off_lo;
    // rows per tile: two accumulator sets share the 512 TMEM columns (2 * msub * NT <= 512), msub a power of two;
    // fewer sub-tiles when that would leave SMs without a tile
    int msub = 1;
    while (msub * 2 * NT <= 256 && msub * 128 < io.Q) msub *= 2;
    if (d.split) msub = std::min(msub, 2);
    {
        // k = 1 layers (QKV / fc / w_2, flattened over the batch rows): 128-row tiles.  Their A block is the whole K extent
        // (256 rows x 256..1024 channels = 4-16 chunks for 3 slots), so a 256-row tile cannot be prefetched across the tile
        // boundary and every tile starts with an exposed load latency; at 128 rows the chunks are half as long and the tile
        // count doubles (4.05 -> 8.1 waves for fc / w_2): QKV / fc / w_2 8-10 % faster in isolation, -0.1 ms per step (launch list)
        // (MASG_TC_MSUB_K1: cap on the sub-tiles per tile, 0 = none)
        static int cap = -1;
        if (cap < 0) { const char* e = getenv("MASG_TC_MSUB_K1"); cap = e ? atoi(e) : 1; }
        if (cap > 0 && d.taps == 1) msub = std::min(msub, cap);
    }
    const int ntiles_n = d.N / NT;
    while (msub > 1 && (long)((io.Q + msub * 128 - 1) / (msub * 128)) * ntiles_n * io.B < c->n_sm) msub /= 2;
    if (pair && msub == 2) {
        // pair mode streams half of a tile's weights per CTA, so a 128-row tile per CTA costs the L2 what a 256-row tile costs the
        // single-CTA form — and pads less: T = 600 is 5 tiles of 128 rows (640) against 3 of 256 (768), 17 % fewer MMAs in the
        // decoder's w_1 and the PostNet.  Taken when it saves at least 8 % of the rows.
        const long r2 = (long)((io.Q + 255) / 256) * 256, r1 = (long)((io.Q + 127) / 128) * 128;
        static int m1 = -1;
        if (m1 < 0) { const char* e = getenv("MASG_TC_PAIR_M1"); m1 = e ? atoi(e) : 1; }
        if (m1 && r1 * 100 <= r2 * 92) msub = 1;
    }
    // input channels per weight blob: the commit-time 64 (or less), or 128 for pair-mode launches with 128-row tiles
    const bool kw128 = pair && msub == 1 && d.w16h2 != nullptr;
    const int KW = kw128 ? 128 : d.KW;
    a.KW = KW;
    int KC = 0, grid_cap = c->n_sm;
    if (tma_in) {
        // one [KC/8][Rb][8] block per tile (its msub*128 rows + the halo); sub-tile ms starts 128 rows further down; every
        // 8-channel group is one contiguous run in HBM, moved by a single 1-D bulk copy
        MASG_CHECK(io.x_ld == d.Cin, "conv_tc: planar input must have exactly Cin=%d channels", d.Cin);
        int Rb = 0;
        for (;;) {
            Rb = (msub * 128 + halo + 7) & ~7;
            for (int m = d.Cin / KW; m >= 1 && !KC; --m) {
                if ((d.Cin / KW) % m) continue;
                const int kc = m * KW;
                if ((size_t)Rb * kc * 2 <= (size_t)A_SLOT_BYTES) KC = kc;
            }
            if (KC || msub == 1) break;
            msub /= 2;
        }
        MASG_CHECK(KC > 0, "conv_tc: activation tile does not fit shared memory (Cin=%d taps=%d tstep=%d)", d.Cin, d.taps, d.tstep);
        a.Rp = Rb; a.rows = Rb; a.a_blk_bytes = Rb * KC * 2; a.a_tx = a.a_blk_bytes;
        a.a_bytes = (a.a_tx + 1023) & ~1023; a.a_half = 0;
        a.a_lbo16 = Rb; a.a_ms16 = 128;
    } else {
        // one shared [KC/8][Rp][8] tile (rows + halo) staged by the worker warps; sub-tile ms starts 128 rows further down
        MASG_CHECK(io.x_ld % 8 == 0 && io.x_bs % 8 == 0, "conv_tc: x_ld=%d / batch stride must be multiples of 8", io.x_ld);
        int Rp = 0;
        for (;;) {
            const int rows = msub * 128 + halo;
            auto pad_rows = [&](int kc) {
                const int cpr = kc / 8;
                const int want = cpr >= 8 ? 1 : (8 / std::max(cpr, 1)) % 8;    // bank-conflict-free staging stores
                int rp = rows;
                while ((rp % 8) != want && !(cpr >= 8 && (rp & 1))) ++rp;
                return rp;
            };
            for (int m = d.Cin / KW; m >= 1 && !KC; --m) {
                if ((d.Cin / KW) % m) continue;
                const int kc = m * KW;
                if ((size_t)pad_rows(kc) * kc * 2 * a_mul + 256 <= (size_t)A_SLOT_BYTES) { KC = kc; Rp = pad_rows(kc); }
            }
            if (KC || msub == 1) break;
            msub /= 2;
        }
        MASG_CHECK(KC > 0, "conv_tc: activation tile does not fit shared memory (Cin=%d taps=%d tstep=%d)", d.Cin, d.taps, d.tstep);
        a.rows = msub * 128 + halo; a.Rp = Rp;
        a.a_half = (((KC / 8) * Rp * 16) + 127) & ~127;
        a.a_bytes = a.a_half * (int)a_mul;
        a.a_lbo16 = Rp; a.a_ms16 = 128;
    }
    a.msub = msub; a.KC = KC;
    a.blob_bytes = KW * NTB * 2 * (int)a_mul;
    if (pair) a.Wt = kw128 ? d.w16h2 : d.w16h;
    a.n_blobs = (d.Cin / KW) * d.taps;
    // weight ring: every stage costs the issuers one barrier round trip, so stages hold two 16 KB blobs (three stages)
    // when that fits beside the activation slots and the stream is long, else 16 KB (four stages)
    a.w_stages = W_STAGES;
    a.G = std::max(1, W_STAGE_BYTES / a.blob_bytes);
    {
        static int big = -1;
        if (big < 0) { const char* e = getenv("MASG_TC_WSTAGE32"); big = e ? atoi(e) : 6; }     // minimum blobs per tile (huge = never)
        const int g2 = std::max(1, 2 * W_STAGE_BYTES / a.blob_bytes);
        const size_t need = SMEM_HDR + BIAS_MAX * 4 + A_SLOTS * (size_t)a.a_bytes + 3 * (size_t)g2 * a.blob_bytes;
        // measured at the C3 shapes (with the lean planar epilogue): +6 % at k = 7, +8 % / +13 % at k = 11 (stage 2 / stage 1),
        // neutral for k = 3 and the upsamplers
        const int per16 = W_STAGE_BYTES / a.blob_bytes;                       // blobs per 16 KB (1, or 2 for pair mode's 64-channel blobs)
        if (big && !d.split && per16 >= 1 && per16 * a.blob_bytes == W_STAGE_BYTES && a.n_blobs >= big * per16 && need <= (size_t)MAX_SMEM) { a.G = g2; a.w_stages = 3; }
    }
    {
        // small weight sets (C = 64 k = 11: 90 KB; C = 128 k = 3: 96 KB) stay resident: one stage holding every blob, loaded by
        // the first tile of the CTA; later tiles only exchange the barrier
        static int res = -1;
        if (res < 0) { const char* e = getenv("MASG_TC_WRESIDENT"); res = e ? atoi(e) : 1; }
        const size_t wbytes = (size_t)a.n_blobs * a.blob_bytes;
        const size_t need = SMEM_HDR + BIAS_MAX * 4 + A_SLOTS * (size_t)a.a_bytes + wbytes;
        a.w_resident = 0;
        // several N tiles (decoder QKV: 6, fc: 2): the grid is rounded down to a multiple of the N-tile count, so that
        // tile % n_ntiles is the same for every tile a CTA walks and its one N tile's weights can stay resident too
        const int grid_r = (c->n_sm / ntiles_n) * ntiles_n;
        const long tiles = (long)io.B * ((io.Q + msub * 128 - 1) / (msub * 128)) * ntiles_n;
        const bool multi_ok = ntiles_n > 1 && ntiles_n <= 8 && grid_r * 10 >= c->n_sm * 9 && tiles >= 2L * grid_r;
        if (res && tma_in && !d.split && (ntiles_n == 1 || multi_ok) && wbytes <= 100 * 1024 && need <= (size_t)MAX_SMEM) {
            a.w_resident = 1; a.G = a.n_blobs; a.w_stages = 1;
            if (ntiles_n > 1) grid_cap = grid_r;
        }
    }
    a.stage_bytes = a.G * a.blob_bytes;
    a.n_blobs = (d.Cin / KW) * d.taps;
    int cols = 32; while (cols < 2 * msub * NT) cols *= 2;
    MASG_CHECK(cols <= 512, "conv_tc: %d TMEM columns needed", cols);
    a.tmem_cols = cols;
    a.B = io.B; a.row_tiles = (io.Q + msub * 128 - 1) / (msub * 128); a.n_ntiles = ntiles_n;
    a.total_tiles = io.B * a.row_tiles * ntiles_n;
    a.total_pairs = ((io.B * a.row_tiles + 1) / 2) * ntiles_n;
    a.Y = io.Y; a.Y16 = io.Y16; a.y_bs = io.y_bs; a.y_limit = io.y_limit; a.orow = d.orow ? d.orow : d.N; a.ooff = d.ooff; a.Q = io.Q;
    a.in_act = io.in_act; a.in_slope = io.in_slope; a.out_act = io.out_act; a.out_slope = io.out_slope;
    a.R1 = io.R1; a.R2 = io.R2; a.r_bf16 = io.r_bf16; a.out_scale = io.out_scale; a.post_slope = io.post_slope;
    { static int dbg = -1; if (dbg < 0) { const char* e = getenv("MASG_TC_DEBUG"); dbg = e ? atoi(e) : 0; } a.dbg = dbg; }
    { static int pf = -1; if (pf < 0) { const char* e = getenv("MASG_TC_RESPF"); pf = e ? atoi(e) : 1; } a.res_pf = pf; }
    {   // helper copiers only where a tile has many runs and few MMAs: the first upsampler (64 groups, 2 taps: 0.290 -> 0.238 ms).
        // Everywhere else four copiers measured 1-3 % SLOWER than the copy warp alone (more warps polling the slot barriers), and the
        // k = 1 layers (w_2: 128 runs per 128-row tile), the suspects, did not move: they are not bound by the copy issue.
        static int nc = -1;
        if (nc < 0) { const char* e = getenv("MASG_TC_COPIERS"); nc = e ? atoi(e) : 0; }
        a.copiers = nc > 0 ? nc : ((d.Cin >= 512 && d.taps <= 2) ? 4 : 1);
    }
    a.ln_g = io.ln_g; a.ln_b = io.ln_b; a.ln_lens = io.ln_lens; a.ln_T = io.ln_T; a.Y16p = io.Y16p;
    if (io.ln_g) {
        MASG_CHECK(NT == 256 && d.N == 256 && io.ln_b, "conv_tc: the fused LayerNorm needs a 256-column tile (weights prepared wide)");
        MASG_CHECK(io.Y && !io.Y16 && !io.y_planar && !io.R2 && !io.r_bf16 && io.out_scale == 1.f && io.post_slope == 1.f && a.orow == 256 && a.ooff == 0,
                   "conv_tc: the fused LayerNorm writes fp32 row-major rows of 256 and takes one fp32 residual");
        MASG_CHECK((!io.ln_lens && !io.Y16p) || io.ln_T > 0, "conv_tc: ln_T is required with ln_lens / Y16p");
    }
    a.y_planar = io.y_planar;
    if (io.y_planar) {
        MASG_CHECK(io.Y16 != nullptr && io.Y == nullptr && io.y_T > 0, "conv_tc: planar output is bf16 only and needs y_T");
        // transposed convs split a column into (phase, channel) with a shift; plain convs only need whole 8-channel groups
        MASG_CHECK(a.bias_mod >= 8 && (d.up > 1 ? (a.bias_mod & (a.bias_mod - 1)) == 0 : a.bias_mod % 16 == 0),
                   "conv_tc: planar output needs a power-of-two (transposed conv) / multiple-of-16 channel count (%d)", a.bias_mod);
        MASG_CHECK((io.R1 == nullptr && io.R2 == nullptr) || io.r_bf16, "conv_tc: planar residuals must be bf16");
        int sh = 0; while ((1 << sh) < a.bias_mod) ++sh;
        a.cout_shift = sh; a.y_T = io.y_T; a.up = d.up > 0 ? d.up : 1; a.up_pad = d.up_pad;
        MASG_CHECK(d.N == a.bias_mod * a.up || d.up == 0, "conv_tc: transposed-conv column count mismatch");
    } else {
        MASG_CHECK(a.orow % 16 == 0 && a.ooff % 16 == 0 && a.y_bs % 16 == 0 && a.bias_mod % 16 == 0,
                   "conv_tc: output row stride/offset/bias period must be multiples of 16");
    }
    MASG_CHECK(a.in_act == ACT_NONE || a.in_act == ACT_LRELU || !tma_in, "conv_tc: planar input supports leaky-relu only");
    const size_t smem = SMEM_HDR + BIAS_MAX * 4 + A_SLOTS * (size_t)a.a_bytes + (size_t)a.w_stages * a.stage_bytes;
    MASG_CHECK(smem <= (size_t)MAX_SMEM, "conv_tc: %zu bytes of shared memory needed", smem);
    MASG_CHECK(!tma_in || ((uintptr_t)io.X & 15) == 0, "conv_tc: planar input must be 16-byte aligned");
    if (pair) {
        // resident weights with several column tiles: the cluster count is a multiple of the column-tile count (see grid_r above)
        int clusters = c->n_sm / 2;
        if (a.w_resident && ntiles_n > 1) clusters = (clusters / ntiles_n) * ntiles_n;
        MASG_CHECK(clusters > 0, "conv_tc: pair mode needs at least %d SMs", 2 * ntiles_n);
        const dim3 grid2(2 * std::min(a.total_pairs, clusters), 1, 1);
        launch_inst_pair<128>(a, grid2, smem, s);
        CUDA_OK(cudaGetLastError());
        c->launches++;
        return;
    }
    dim3 grid(std::min(a.total_tiles, grid_cap), 1, 1);            // persistent: one CTA per SM walks the tile list
#define MASG_TC_CASE(NTV)                                                                       \
    case NTV: if (tma_in) launch_inst<NTV, true>(a, grid, smem, s, c->n_sm); else launch_inst<NTV, false>(a, grid, smem, s, c->n_sm); break;
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
