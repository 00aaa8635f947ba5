// Fused HiFi-GAN ResBlock1 pair for sm_100a (MASG_PREC_BF16 path): one kernel computes
//     y = ( c2( lrelu( c1( lrelu(x) ) + b1 ) ) + b2 + x [+ acc] ) * scale      [then an optional trailing leaky-relu]
// i.e. one (convs1[m], convs2[m]) step of ResBlock1.forward (hifigan/models.py:96-103), with the MRF sum, the /3 and
// the leaky-relu that follows the block folded in as in hifigan.cu.  The intermediate h = lrelu(c1(..)) never leaves
// the SM: conv 1 accumulates in TMEM, the epilogue warps write h as bf16 straight into shared memory in the UMMA
// no-swizzle K-major layout, and conv 2 consumes it from there as its A operand.  Per pair the HBM traffic drops from
// five tensor passes (x, h out, h in, residual x, y) to two (x in, y out).
//
// Used for the narrow stages (C = 32: k = 3, 7, 11; C = 64: k = 3, 7 — the pairs whose weights fit), which are HBM-bound as separate convs.  Both convs' weights
// (k C C bf16 each, pre-tiled by tc_prepare_weights) stay RESIDENT in shared memory for the life of the CTA, so the MMA
// issue loop is a fully unrolled run of taps x K-steps with descriptors advanced by uniform adds.
//
// Tile = 256 rows of h (two M = 128 sub-tiles), from which R_out = 256 - (k - 1) output rows are valid:
//     x rows [q0 - p2 - p1, +256 + 2 p1)  ->  h rows [q0 - p2, +256)  ->  y rows [q0, q0 + R_out),   p1 = (k-1) d / 2, p2 = (k-1) / 2.
// h rows outside [0, T) are written as zeros (conv 2's zero padding applies to h, not to c1 of a padded x).
//
// Persistent, warp-specialised, one 768-thread CTA per SM; mbarrier rings connect the roles:
//   2 copy warps: one 1-D bulk async copy per 8-channel group of the planar bf16 input [B][C/8][T][8] (3 slots)
//   4 workers   : leaky-relu in place on the landed tile (the raw x is also the residual, so the producer cannot apply it)
//   2 issuers   : issuer w owns sub-tile w of both convs; software-pipelined  P1(0) | P1(i+1) P2(i) ...
//   8 + 8 epilogue warps: E1 (TMEM -> bias, lrelu, bf16 -> shared h) and E2 (TMEM -> bias, residuals, scale -> planar bf16 y)
// TMEM: acc1[2 buffers][2 sub-tiles][C] | acc2[2][2][C] = 8 C columns (256 / 512).
#include "common.cuh"
#include "tc.cuh"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int RP_EPI_WARPS = 8;             // warps 0-7: E1, warps 8-15: E2 (warp % 4 = TMEM lane quadrant, (warp / 4) % 2 = channel half)
constexpr int RP_MMA_WARP = 16;             // warps 16, 17: MMA issuers
constexpr int RP_ISSUERS = 2;
constexpr int RP_COPY_WARP = 18;            // warps 18, 19: copy warps (even / odd tiles of the CTA)
constexpr int RP_WORK_WARP0 = 20;           // warps 20-23: leaky-relu workers
constexpr int RP_WORK_WARPS = 4;
constexpr int RP_WORK_THREADS = RP_WORK_WARPS * 32;
constexpr int RP_THREADS = (RP_WORK_WARP0 + RP_WORK_WARPS) * 32;     // 768
constexpr int RP_HROWS = 272;               // 256 rows of h + the (k - 1) rows the second sub-tile's taps reach past them
constexpr int RP_HDR = 1024;
constexpr int RP_MAX_SMEM = 227 * 1024;

struct RpArgs {
    const bf16* X; bf16* Y; const bf16* R2;
    const bf16* W1; const bf16* W2; const float* b1; const float* b2;
    int B, T, dil, tiles_per_b, total_tiles, Rx, r_out, p1, p2;
    float slope, out_scale, post_slope;
    int dbg, res_pf;
};

// MASG_TC_DEBUG & 32: timeline of CTA 0 (clock64 at the hand-over points of its first RP_TRACE_TILES tiles).
// roles: 0 copy warp (wait-begin, slot-free, issued) | 1 worker 0 (landed, done) | 2 issuer 0 (P1 ready, P1 committed, P2 ready,
// P2 committed) | 3 E1 warp 0 (acc1 full, h written) | 4 E2 warp 0 (residuals requested, acc2 full, stored)
constexpr int RP_TRACE_TILES = 96, RP_TRACE_EVTS = 4, RP_TRACE_ROLES = 5;
__device__ long long g_rp_trace[RP_TRACE_ROLES * RP_TRACE_TILES * RP_TRACE_EVTS];
#define RP_TRACE(on, role, it, evt) do { if ((on) && (it) < RP_TRACE_TILES) g_rp_trace[((role) * RP_TRACE_TILES + (it)) * RP_TRACE_EVTS + (evt)] = clock64(); } while (0)

constexpr int RP_MAX_XS = 6;                // activation slots the barrier block has room for
struct RpBars {
    uint32_t base;
    __device__ uint32_t x_raw(int i) const { return base + 8u * i; }
    __device__ uint32_t x_full(int i) const { return base + 8u * (RP_MAX_XS + i); }
    __device__ uint32_t x_empty(int i) const { return base + 8u * (2 * RP_MAX_XS + i); }
    __device__ uint32_t acc1_full(int i) const { return base + 8u * (3 * RP_MAX_XS + i); }
    __device__ uint32_t acc1_empty(int i) const { return base + 8u * (3 * RP_MAX_XS + 2 + i); }
    __device__ uint32_t h_full(int i) const { return base + 8u * (3 * RP_MAX_XS + 4 + i); }
    __device__ uint32_t h_empty(int i) const { return base + 8u * (3 * RP_MAX_XS + 6 + i); }
    __device__ uint32_t acc2_full(int i) const { return base + 8u * (3 * RP_MAX_XS + 8 + i); }
    __device__ uint32_t acc2_empty(int i) const { return base + 8u * (3 * RP_MAX_XS + 10 + i); }
    __device__ uint32_t w_ready() const { return base + 8u * (3 * RP_MAX_XS + 12); }
};
static_assert(8 * (3 * RP_MAX_XS + 13) <= 256, "barrier block overlaps the TMEM slot");

// XS activation slots, NH buffers of h (3 / 2 when shared memory allows, else 2 / 1)
template <int C, int K, int XS, int NH>
__global__ void __launch_bounds__(RP_THREADS, 1) k_respair_tc(const RpArgs a) {
    constexpr int RP_XSLOTS = XS;
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int CG = C / 8;                               // 8-channel groups
    constexpr int KS = C / 16;                              // K = 16 steps per tap
    constexpr int WBYTES = K * C * C * 2;
    constexpr int HBYTES = CG * RP_HROWS * 16;
    const RpBars bars{smem_u32(smem)};
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 256);
    float* bias_s = reinterpret_cast<float*>(smem + 512);   // b1[C] | b2[C]
    const int x_bytes = CG * a.Rx * 16;
    uint8_t* Xs = smem + RP_HDR;
    uint8_t* Hs = Xs + RP_XSLOTS * (size_t)x_bytes;
    uint8_t* W1s = Hs + NH * HBYTES;
    uint8_t* W2s = W1s + WBYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool trace = (a.dbg & 32) && blockIdx.x == 0;
    const int n_my = a.total_tiles > (int)blockIdx.x ? (a.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

    if (threadIdx.x == 0) {
        for (int i = 0; i < RP_XSLOTS; ++i) { mbar_init(bars.x_raw(i), 1); mbar_init(bars.x_full(i), RP_WORK_THREADS); mbar_init(bars.x_empty(i), RP_ISSUERS); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bars.acc1_full(i), RP_ISSUERS); mbar_init(bars.acc1_empty(i), RP_EPI_WARPS);
            mbar_init(bars.h_full(i), RP_EPI_WARPS); mbar_init(bars.h_empty(i), RP_ISSUERS);
            mbar_init(bars.acc2_full(i), RP_ISSUERS); mbar_init(bars.acc2_empty(i), RP_EPI_WARPS);
        }
        mbar_init(bars.w_ready(), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 2 * C; i += RP_THREADS) bias_s[i] = i < C ? a.b1[i] : a.b2[i - C];
    if (warp == RP_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(8 * C) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // programmatic dependent launch (see conv_tc.cu): the resident weights are requested before the wait
    pdl_trigger();
    if (warp == RP_COPY_WARP && lane == 0) {
        mbar_expect_tx(bars.w_ready(), 2u * WBYTES);
        bulk_g2s(smem_u32(W1s), a.W1, WBYTES, bars.w_ready());
        bulk_g2s(smem_u32(W2s), a.W2, WBYTES, bars.w_ready());
    }
    if (!(warp >= RP_MMA_WARP && warp < RP_MMA_WARP + RP_ISSUERS)) pdl_wait();

    if (warp >= RP_WORK_WARP0) {
        // ===== workers: leaky-relu in place on the landed x tile =====
        const int tid = threadIdx.x - RP_WORK_WARP0 * 32;
        const int pieces = x_bytes >> 4;
        const Slope2 sl = make_slope2(a.slope);
        for (int it = 0; it < n_my; ++it) {
            const int slot = it % RP_XSLOTS, u = it / RP_XSLOTS;
            mbar_wait_relaxed(bars.x_raw(slot), u & 1);
            RP_TRACE(trace && tid == 0, 1, it, 0);
            uint4* p = reinterpret_cast<uint4*>(Xs + (size_t)slot * x_bytes);
            // one piece per thread per trip, on purpose: four pieces in flight shortened this pass from ~2000 to ~1250 cycles but the
            // bursts compete with the MMAs' operand reads for the shared-memory pipe (k = 7 0.401 -> 0.425 ms, and the MMA-bound
            // stage-2 convs of conv_tc, which share the loop shape, +4 %)
            for (int i = tid; i < pieces && !(a.dbg & 128); i += RP_WORK_THREADS) {   // dbg 128: what-if, no leaky-relu pass
                uint4 v = p[i];
                v.x = lrelu_bf16x2(v.x, sl); v.y = lrelu_bf16x2(v.y, sl); v.z = lrelu_bf16x2(v.z, sl); v.w = lrelu_bf16x2(v.w, sl);
                p[i] = v;
            }
            fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core (async proxy)
            mbar_arrive(bars.x_full(slot));
            RP_TRACE(trace && tid == 0, 1, it, 1);
        }
    } else if (warp == RP_COPY_WARP || warp == RP_COPY_WARP + 1) {
        // ===== copy warp: resident weights once, then one contiguous run per 8-channel group per tile =====
        const uint8_t* X8 = reinterpret_cast<const uint8_t*>(a.X);
        for (int it = warp - RP_COPY_WARP; it < n_my; it += 2) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int slot = it % RP_XSLOTS, u = it / RP_XSLOTS;
            const int b = tile / a.tiles_per_b, i = tile - b * a.tiles_per_b;
            RP_TRACE(trace && lane == 0, 0, it, 0);
            if (lane == 0) mbar_wait_relaxed(bars.x_empty(slot), (u & 1) ^ 1);
            __syncwarp();
            RP_TRACE(trace && lane == 0, 0, it, 1);
            uint8_t* dst_g = Xs + (size_t)slot * x_bytes;
            const uint32_t dst = smem_u32(dst_g), bar = bars.x_raw(slot);
            const int t0 = i * a.r_out - a.p2 - a.p1;
            const int lo = max(t0, 0), hi = max(min(t0 + a.Rx, a.T), lo);
            const uint32_t run = (uint32_t)(hi - lo) * 16u;
            if (lo > t0 || hi < t0 + a.Rx) {                 // conv zero padding at the utterance edges
                if (lane < CG) {
                    uint4* blk = reinterpret_cast<uint4*>(dst_g + (size_t)lane * a.Rx * 16);
                    for (int r = 0; r < min(lo - t0, a.Rx); ++r) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                    for (int r = max(hi - t0, 0); r < a.Rx; ++r) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                }
                fence_proxy_async();
                __syncwarp();
            }
            if (lane == 0) {
                if (run) mbar_expect_tx(bar, run * (uint32_t)CG); else mbar_arrive(bar);
            }
            __syncwarp();
            if (run && lane < CG) {
                const uint8_t* src = X8 + (((size_t)b * CG + lane) * a.T + lo) * 16;
                bulk_g2s(dst + (uint32_t)((lane * a.Rx + (lo - t0)) * 16), src, run, bar);
            }
            if (a.R2 != nullptr && a.res_pf && lane < CG) {
                // MRF accumulator rows of this tile -> L2 ahead of the E2 warps (see conv_tc.cu: DRAM loads queue behind the copies)
                const int q0 = i * a.r_out, rows = min(a.r_out, a.T - q0);
                if (rows > 0) bulk_prefetch_l2(a.R2 + (((size_t)b * CG + lane) * a.T + q0) * 8, (uint32_t)rows * 16u);
            }
            RP_TRACE(trace && lane == 0, 0, it, 2);
        }
    } else if (warp >= RP_MMA_WARP && warp < RP_MMA_WARP + RP_ISSUERS) {
        // ===== MMA issuers: issuer w owns sub-tile w (rows [128 w, 128 w + 128) of the tile) of both convs =====
        const uint32_t w = (uint32_t)(warp - RP_MMA_WARP);
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(C >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);                  // SBO = 128 B, descriptor version 1
        const uint32_t lbo_x = (uint32_t)a.Rx, dil = (uint32_t)a.dil;
        const uint32_t xs16 = smem_u32(Xs) >> 4, hs16 = smem_u32(Hs) >> 4, x16 = (uint32_t)x_bytes >> 4;
        const uint32_t w1_lo = ((uint32_t)C << 16) | (smem_u32(W1s) >> 4), w2_lo = ((uint32_t)C << 16) | (smem_u32(W2s) >> 4);
        mbar_wait(bars.w_ready(), 0);
        auto P1 = [&](int it) {
            const int slot = it % RP_XSLOTS, u = it / RP_XSLOTS, b = it & 1, ub = it >> 1;
            mbar_wait(bars.x_full(slot), u & 1);
            mbar_wait(bars.acc1_empty(b), (ub & 1) ^ 1);
            tc_fence_after();
            RP_TRACE(trace && w == 0 && lane == 0, 2, it, 0);
            if (elect_one()) {
                const uint32_t a0 = (lbo_x << 16) | (xs16 + (uint32_t)slot * x16 + w * 128u);
                const uint32_t d = tmem_base + (uint32_t)((b * 2) * C) + w * (uint32_t)C;
#pragma unroll
                for (int j = 0; j < K; ++j) {
#pragma unroll
                    for (int ks = 0; ks < KS; ++ks)
                        tc_mma(d, a0 + (uint32_t)j * dil + (uint32_t)ks * 2u * lbo_x, w1_lo + (uint32_t)((j * KS + ks) * 2 * C), desc_hi, idesc, (j | ks) ? 1u : 0u);
                }
                tc_commit(bars.x_empty(slot));
                tc_commit(bars.acc1_full(b));
            }
            __syncwarp();
            RP_TRACE(trace && w == 0 && lane == 0, 2, it, 1);
        };
        auto P2 = [&](int it) {
            const int b = it & 1, ub = it >> 1;
            const int hbuf = NH == 2 ? b : 0, hu = NH == 2 ? ub : it;
            mbar_wait(bars.h_full(hbuf), hu & 1);
            mbar_wait(bars.acc2_empty(b), (ub & 1) ^ 1);
            tc_fence_after();
            RP_TRACE(trace && w == 0 && lane == 0, 2, it, 2);
            if (elect_one()) {
                const uint32_t a0 = ((uint32_t)RP_HROWS << 16) | (hs16 + (uint32_t)hbuf * (uint32_t)(HBYTES >> 4) + w * 128u);
                const uint32_t d = tmem_base + (uint32_t)(4 * C + (b * 2) * C) + w * (uint32_t)C;
#pragma unroll
                for (int j = 0; j < K; ++j) {
#pragma unroll
                    for (int ks = 0; ks < KS; ++ks)
                        tc_mma(d, a0 + (uint32_t)j + (uint32_t)(ks * 2 * RP_HROWS), w2_lo + (uint32_t)((j * KS + ks) * 2 * C), desc_hi, idesc, (j | ks) ? 1u : 0u);
                }
                tc_commit(bars.h_empty(hbuf));
                tc_commit(bars.acc2_full(b));
            }
            __syncwarp();
            RP_TRACE(trace && w == 0 && lane == 0, 2, it, 3);
        };
        if (n_my > 0) P1(0);
        for (int it = 0; it < n_my; ++it) {
            if (it + 1 < n_my) P1(it + 1);
            P2(it);
        }
    } else if (warp < RP_EPI_WARPS) {
        // ===== E1 warps: acc1 (TMEM) -> + b1 -> bf16 -> leaky-relu -> h in shared memory (UMMA K-major layout) =====
        const int quad = warp & 3, hf = warp >> 2;
        constexpr int CH = C / 2;                            // channels per warp (this half)
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const Slope2 sl = make_slope2(a.slope);
        for (int it = 0; it < n_my; ++it) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int b = it & 1, ub = it >> 1;
            const int i = tile % a.tiles_per_b;
            const int tq = i * a.r_out - a.p2;               // time of h row 0
            const bool edge = tq < 0 || tq + 256 > a.T;      // some h rows fall outside [0, T): they must be zeros (conv 2 padding)
            const int hbuf = NH == 2 ? b : 0, hu = NH == 2 ? ub : it;
            mbar_wait_relaxed(bars.acc1_full(b), ub & 1);
            mbar_wait_relaxed(bars.h_empty(hbuf), (hu & 1) ^ 1);
            tc_fence_after();
            RP_TRACE(trace && threadIdx.x == 0, 3, it, 0);
            uint8_t* hb = Hs + (size_t)hbuf * HBYTES;
#pragma unroll
            for (int cg = 0; cg < CH / 16; ++cg) {
                const int col0 = hf * CH + cg * 16;
                float v[2][16];
                tmem_ld16x2(tlane + (uint32_t)((b * 2) * C + col0), tlane + (uint32_t)((b * 2 + 1) * C + col0), v[0], v[1]);
                const float4* b4 = reinterpret_cast<const float4*>(bias_s + col0);
                const float4 ba = b4[0], bb = b4[1], bc = b4[2], bd = b4[3];
#pragma unroll
                for (int ms = 0; ms < 2; ++ms) {
                    float* x = v[ms];
                    add_f32x2(x[0], x[1], ba.x, ba.y); add_f32x2(x[2], x[3], ba.z, ba.w); add_f32x2(x[4], x[5], bb.x, bb.y); add_f32x2(x[6], x[7], bb.z, bb.w);
                    add_f32x2(x[8], x[9], bc.x, bc.y); add_f32x2(x[10], x[11], bc.z, bc.w); add_f32x2(x[12], x[13], bd.x, bd.y); add_f32x2(x[14], x[15], bd.z, bd.w);
                    uint4 pa = make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
                    uint4 pb = make_uint4(pack_bf16(x[8], x[9]), pack_bf16(x[10], x[11]), pack_bf16(x[12], x[13]), pack_bf16(x[14], x[15]));
                    pa.x = lrelu_bf16x2(pa.x, sl); pa.y = lrelu_bf16x2(pa.y, sl); pa.z = lrelu_bf16x2(pa.z, sl); pa.w = lrelu_bf16x2(pa.w, sl);
                    pb.x = lrelu_bf16x2(pb.x, sl); pb.y = lrelu_bf16x2(pb.y, sl); pb.z = lrelu_bf16x2(pb.z, sl); pb.w = lrelu_bf16x2(pb.w, sl);
                    const int r = ms * 128 + quad * 32 + lane;
                    if (edge) {
                        const int t = tq + r;
                        if (t < 0 || t >= a.T) { pa = make_uint4(0u, 0u, 0u, 0u); pb = pa; }
                    }
                    uint4* dst = reinterpret_cast<uint4*>(hb + ((size_t)(col0 >> 3) * RP_HROWS + r) * 16);
                    dst[0] = pa;
                    dst[RP_HROWS] = pb;
                }
            }
            fence_proxy_async();                            // h stores -> visible to the tensor core
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { mbar_arrive(bars.h_full(hbuf)); mbar_arrive(bars.acc1_empty(b)); }
            RP_TRACE(trace && threadIdx.x == 0, 3, it, 1);
        }
    } else if (warp < 2 * RP_EPI_WARPS) {
        // ===== E2 warps: acc2 (TMEM) -> + b2 + x [+ acc] -> scale -> trailing leaky-relu -> planar bf16 y =====
        const int quad = warp & 3, hf = (warp - RP_EPI_WARPS) >> 2;
        constexpr int CH = C / 2;
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const long gstride = (long)a.T * 8;                  // elements between 8-channel groups
        const bool has_r2 = a.R2 != nullptr;
        const Slope2 sl_post = make_slope2(a.post_slope);
        for (int it = 0; it < n_my; ++it) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int b = it & 1, ub = it >> 1;
            const int bb = tile / a.tiles_per_b, i = tile - bb * a.tiles_per_b;
            const int q0 = i * a.r_out;
            const long boff = ((long)bb * CG + hf * (CH / 8)) * gstride;
            // residuals of both sub-tiles are requested before the accumulator is waited for (L2 hits: the copy warp read
            // the same rows a few tiles ago).  MASG_TC_DEBUG what-ifs: 2 = no residual loads / stores, 4 = no stores
            constexpr bool PRE_RR = CH / 8 <= 2;            // wide halves: the MRF accumulator is read at use instead (registers)
            uint4 rx[2][CH / 8], rr[2][PRE_RR ? CH / 8 : 1];
            bool ok[2];
#pragma unroll
            for (int ms = 0; ms < 2; ++ms) {
                const int r = ms * 128 + quad * 32 + lane, t = q0 + r;
                ok[ms] = r < a.r_out && t < a.T && !(a.dbg & 2);
#pragma unroll
                for (int g = 0; g < CH / 8; ++g) {
                    const long off = boff + (long)g * gstride + (long)t * 8;
                    rx[ms][g] = ok[ms] ? *reinterpret_cast<const uint4*>(a.X + off) : make_uint4(0u, 0u, 0u, 0u);
                    if (PRE_RR) rr[ms][g] = (ok[ms] && has_r2) ? *reinterpret_cast<const uint4*>(a.R2 + off) : make_uint4(0u, 0u, 0u, 0u);
                }
            }
            RP_TRACE(trace && threadIdx.x == RP_EPI_WARPS * 32, 4, it, 0);
            mbar_wait_relaxed(bars.acc2_full(b), ub & 1);
            tc_fence_after();
            RP_TRACE(trace && threadIdx.x == RP_EPI_WARPS * 32, 4, it, 1);
#pragma unroll
            for (int ms = 0; ms < 2; ++ms) {
                const int t = q0 + ms * 128 + quad * 32 + lane;
#pragma unroll
                for (int cg = 0; cg < CH / 16; ++cg) {
                    const int col0 = hf * CH + cg * 16;
                    float v[16];
                    tmem_ld16(tlane + (uint32_t)(4 * C + (b * 2 + ms) * C + col0), v);
                    const float4* b4 = reinterpret_cast<const float4*>(bias_s + C + col0);
                    const float4 ba = b4[0], bc = b4[1], bd = b4[2], be = b4[3];
                    add_f32x2(v[0], v[1], ba.x, ba.y); add_f32x2(v[2], v[3], ba.z, ba.w); add_f32x2(v[4], v[5], bc.x, bc.y); add_f32x2(v[6], v[7], bc.z, bc.w);
                    add_f32x2(v[8], v[9], bd.x, bd.y); add_f32x2(v[10], v[11], bd.z, bd.w); add_f32x2(v[12], v[13], be.x, be.y); add_f32x2(v[14], v[15], be.z, be.w);
#pragma unroll
                    for (int h8 = 0; h8 < 2; ++h8) {
                        const uint4 ux = rx[ms][cg * 2 + h8];
                        const float2 x0 = unpack_bf16(ux.x), x1 = unpack_bf16(ux.y), x2 = unpack_bf16(ux.z), x3 = unpack_bf16(ux.w);
                        float* vv = v + h8 * 8;
                        add_f32x2(vv[0], vv[1], x0.x, x0.y); add_f32x2(vv[2], vv[3], x1.x, x1.y); add_f32x2(vv[4], vv[5], x2.x, x2.y); add_f32x2(vv[6], vv[7], x3.x, x3.y);
                        if (has_r2) {
                            uint4 ur = make_uint4(0u, 0u, 0u, 0u);
                            if (PRE_RR) ur = rr[ms][cg * 2 + h8];
                            else if (ok[ms]) ur = *reinterpret_cast<const uint4*>(a.R2 + boff + (long)(cg * 2 + h8) * gstride + (long)t * 8);
                            const float2 r0 = unpack_bf16(ur.x), r1 = unpack_bf16(ur.y), r2 = unpack_bf16(ur.z), r3 = unpack_bf16(ur.w);
                            add_f32x2(vv[0], vv[1], r0.x, r0.y); add_f32x2(vv[2], vv[3], r1.x, r1.y); add_f32x2(vv[4], vv[5], r2.x, r2.y); add_f32x2(vv[6], vv[7], r3.x, r3.y);
                        }
                    }
                    if (a.out_scale != 1.f) {
#pragma unroll
                        for (int k = 0; k < 16; k += 2) mul_f32x2(v[k], v[k + 1], a.out_scale, a.out_scale);
                    }
                    uint4 pa = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                    uint4 pb = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                    if (a.post_slope != 1.f) {
                        pa.x = lrelu_bf16x2(pa.x, sl_post); pa.y = lrelu_bf16x2(pa.y, sl_post); pa.z = lrelu_bf16x2(pa.z, sl_post); pa.w = lrelu_bf16x2(pa.w, sl_post);
                        pb.x = lrelu_bf16x2(pb.x, sl_post); pb.y = lrelu_bf16x2(pb.y, sl_post); pb.z = lrelu_bf16x2(pb.z, sl_post); pb.w = lrelu_bf16x2(pb.w, sl_post);
                    }
                    if (ok[ms] && !(a.dbg & 4)) {
                        const long off = boff + (long)(cg * 2) * gstride + (long)t * 8;
                        *reinterpret_cast<uint4*>(a.Y + off) = pa;
                        *reinterpret_cast<uint4*>(a.Y + off + gstride) = pb;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_relaxed(bars.acc2_empty(b));   // orders only the TMEM reads above, not the global stores
            RP_TRACE(trace && threadIdx.x == RP_EPI_WARPS * 32, 4, it, 2);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == RP_MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(8 * C) : "memory");
    }
}

template <int C, int K, int XS, int NH>
void launch_rp(const RpArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_respair_tc<C, K, XS, NH>, cudaFuncAttributeMaxDynamicSharedMemorySize, RP_MAX_SMEM));
    CUDA_OK(launch_pdl(k_respair_tc<C, K, XS, NH>, grid, RP_THREADS, smem, s, a, a.total_tiles <= (int)grid.x));
}

size_t rp_smem(int C, int K, int Rx, int xs, int nh) {
    return (size_t)RP_HDR + (size_t)xs * (C / 8) * Rx * 16 + (size_t)nh * (C / 8) * RP_HROWS * 16 + 2 * (size_t)K * C * C * 2;
}
// C = 32: 6 activation slots + 2 h buffers (the slot cycle — copy issue 1.5 k + copy latency 2.8 k + leaky-relu pass 2.4 k + conv 1 —
// is ~7.6 kcycles, so 3 slots capped the tile period at 2.6 kcycles whatever k; MASG_RP_XS overrides); C = 64: 3 + 2 when they
// fit beside the resident weights, else 2 + 1 (k = 7); false: no fit
bool rp_config(int C, int K, int Rx, int& xs, int& nh) {
    static int want = -1;
    if (want < 0) { const char* e = getenv("MASG_RP_XS"); want = e ? atoi(e) : RP_MAX_XS; }
    {   // C = 64, k = 3: experiment switch MASG_RP_C64K3=41 -> 4 activation slots + 1 h buffer instead of 3 + 2
        static int alt = -1;
        if (alt < 0) { const char* e = getenv("MASG_RP_C64K3"); alt = e ? atoi(e) : 0; }
        if (C == 64 && K == 3 && alt == 41 && rp_smem(C, K, Rx, 4, 1) <= (size_t)RP_MAX_SMEM) { xs = 4; nh = 1; return true; }
    }
    xs = (C == 32 && want == RP_MAX_XS) ? RP_MAX_XS : 3; nh = 2;
    if (rp_smem(C, K, Rx, xs, nh) <= (size_t)RP_MAX_SMEM) return true;
    xs = 3;
    if (rp_smem(C, K, Rx, xs, nh) <= (size_t)RP_MAX_SMEM) return true;
    xs = 2; nh = 1;
    return rp_smem(C, K, Rx, xs, nh) <= (size_t)RP_MAX_SMEM;
}
int rp_rx(const ConvDesc& d1) { return (256 + (d1.taps - 1) * d1.tstep + 7) & ~7; }

}  // namespace

int respair_debug_trace(long long* out, int max_elems) {
    const int n = RP_TRACE_ROLES * RP_TRACE_TILES * RP_TRACE_EVTS;
    if (max_elems < n) return -1;
    CUDA_OK(cudaMemcpyFromSymbol(out, g_rp_trace, sizeof(long long) * n));
    return RP_TRACE_TILES;
}

bool respair_tc_supported(const masg_ctx* c, const ConvDesc& d1, const ConvDesc& d2) {
    static int off = -1;
    if (off < 0) { const char* e = getenv("MASG_NO_RESPAIR"); off = e ? atoi(e) : 0; }
    if (off || c->precision != MASG_PREC_BF16) return false;
    const int C = d1.Cin, K = d1.taps;
    if (!(C == 32 || C == 64) || d1.N != C || d2.Cin != C || d2.N != C) return false;
    if (!(K == 3 || K == 7 || K == 11) || d2.taps != K || d2.tstep != 1 || d1.tstep < 1) return false;
    if (d1.off0 != -((K - 1) * d1.tstep) / 2 || d2.off0 != -(K - 1) / 2) return false;
    if (d1.split || d2.split || d1.NT != C || d2.NT != C || d1.KW != C || d2.KW != C || !d1.w16 || !d2.w16) return false;
    if (d1.up || d2.up || (d1.bias_mod && d1.bias_mod != C) || (d2.bias_mod && d2.bias_mod != C)) return false;
    int xs, nh;
    return rp_config(C, K, rp_rx(d1), xs, nh);
}

void launch_respair_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d1, const ConvDesc& d2, const bf16* X, int B, int T,
                       bf16* Y, const bf16* R2, float out_scale, float post_slope) {
    MASG_CHECK(respair_tc_supported(c, d1, d2), "respair_tc: unsupported pair (C=%d k=%d)", d1.Cin, d1.taps);
    MASG_CHECK(X != Y && (((uintptr_t)X | (uintptr_t)Y | (uintptr_t)R2) & 15) == 0, "respair_tc: tensors must be distinct and 16-byte aligned");
    if (B <= 0 || T <= 0) return;
    const int C = d1.Cin, K = d1.taps;
    RpArgs a;
    memset(&a, 0, sizeof(a));
    a.X = X; a.Y = Y; a.R2 = R2; a.W1 = d1.w16; a.W2 = d2.w16; a.b1 = d1.bias; a.b2 = d2.bias;
    a.B = B; a.T = T; a.dil = d1.tstep; a.p1 = -d1.off0; a.p2 = -d2.off0;
    a.r_out = 256 - (K - 1); a.Rx = rp_rx(d1);
    a.tiles_per_b = (T + a.r_out - 1) / a.r_out; a.total_tiles = B * a.tiles_per_b;
    a.slope = 0.1f; a.out_scale = out_scale; a.post_slope = post_slope;
    { static int dbg = -1; if (dbg < 0) { const char* e = getenv("MASG_TC_DEBUG"); dbg = e ? atoi(e) : 0; } a.dbg = dbg; }
    { static int pf = -1; if (pf < 0) { const char* e = getenv("MASG_TC_RESPF"); pf = e ? atoi(e) : 1; } a.res_pf = pf; }
    const dim3 grid(std::min(a.total_tiles, c->n_sm), 1, 1);
    int xs = 0, nh = 0;
    rp_config(C, K, a.Rx, xs, nh);
    const size_t smem = rp_smem(C, K, a.Rx, xs, nh);
    masg_ctx::ProfRec rec;
    if (c->profiling) {
        CUDA_OK(cudaEventCreate(&rec.e0)); CUDA_OK(cudaEventCreate(&rec.e1));
        rec.flops = 2.0 * 2.0 * B * (double)T * K * C * C;               // both convs, algorithmic
        CUDA_OK(cudaEventRecord(rec.e0, s));
    }
#define RP_CASE(CV, KV, XV, NV) if (C == CV && K == KV && xs == XV && nh == NV) launch_rp<CV, KV, XV, NV>(a, grid, smem, s);
    RP_CASE(32, 3, 3, 2) RP_CASE(32, 7, 3, 2) RP_CASE(32, 11, 3, 2) RP_CASE(64, 3, 3, 2) RP_CASE(64, 7, 2, 1)
    RP_CASE(32, 3, 6, 2) RP_CASE(32, 7, 6, 2) RP_CASE(32, 11, 6, 2) RP_CASE(64, 3, 4, 1)
#undef RP_CASE
    CUDA_OK(cudaGetLastError());
    c->launches++;
    if (c->profiling) { CUDA_OK(cudaEventRecord(rec.e1, s)); c->prof.push_back(rec); }
}
