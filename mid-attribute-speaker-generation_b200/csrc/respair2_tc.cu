// Fused HiFi-GAN ResBlock1 pair on a CTA PAIR (thread-block cluster of 2, tcgen05 cta_group::2) for sm_100a, MASG_PREC_BF16:
//     y = ( c2( lrelu( c1( lrelu(x) ) + b1 ) ) + b2 + x [+ acc] ) * scale      [then an optional trailing leaky-relu]
// — the same contract as respair_tc.cu (hifigan/models.py:96-103), for the pairs whose two weight sets do NOT fit one SM's
// shared memory beside the activation tiles: C = 128 k = 3 (192 KB of weights) and C = 64 k = 11 (176 KB).  As separate convs
// these move five tensor passes through HBM per pair (x, h out, h in, residual x, y) and the C = 128 k = 3 ones are HBM-bound
// (5.9 TB/s, 36 % tensor pipe); fused, h never leaves the SM and the pair moves two passes.
//
// How the weights fit: one tcgen05.mma.cta_group::2 instruction (M = 256, issued by one thread of the LEADER CTA, rank 0 of the
// cluster) multiplies 128 rows of A from EACH CTA's shared memory with a B operand whose N columns are split over the pair — each
// CTA holds the weights of C/2 output channels only (tiled at commit as [half][tap][C/8][C/2][8], d.w16p) — and leaves a
// 128 x C accumulator in each CTA's own TMEM.  So each CTA works on its own time tile (own x copy, own h, own epilogues, own
// stores) exactly as in respair_tc.cu, with half the weight bytes resident; only the issue loop is shared.
//
// Barriers: the ones the issuer waits on ("x tile ready", "h written", "accumulator drained") live in the leader CTA and count
// arrivals from both CTAs (the peer arrives through mapa + mbarrier.arrive.release.cluster); the ones the issuer signals
// ("slot free", "accumulator full", "h consumed") are tcgen05.commit.cta_group::2 ... multicast::cluster arrivals on the same
// barrier offset in both CTAs, so every other role waits on its local copy.
//
// Tile = MS x 128 rows of h per CTA (MS = 2 at C = 64, MS = 1 at C = 128: the 512 TMEM columns hold acc1[2][MS][C] | acc2[2][MS][C]),
// r_out = 128 MS - (k - 1) output rows.  The cluster walks PAIRS of consecutive tiles: rank r of cluster c takes tile
// 2 (c + it * n_clusters) + r; an odd tile count leaves the last peer tile empty (its MMAs run on whatever the slot holds, nothing
// is loaded or stored).
// Roles per CTA (768 threads): 8 E1 warps (which also issue the bulk copies of the x tiles), 8 E2 warps, 2 issuer warps (leader
// only), 6 leaky-relu worker warps.
#include "common.cuh"
#include "tc.cuh"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int R2_EPI_WARPS = 8;             // warps 0-7: E1, warps 8-15: E2 (warp % 4 = TMEM lane quadrant, (warp / 4) % 2 = channel half)
constexpr int R2_MMA_WARP = 16;             // warps 16, 17: MMA issuers (leader CTA); warp 16 also allocates TMEM in both CTAs
constexpr int R2_WORK_WARP0 = 18;           // warps 18-23: leaky-relu workers
constexpr int R2_WORK_WARPS = 6;
constexpr int R2_WORK_THREADS = R2_WORK_WARPS * 32;
constexpr int R2_THREADS = (R2_WORK_WARP0 + R2_WORK_WARPS) * 32;     // 768
constexpr int R2_HDR = 2048;                // barriers [0, 256) | TMEM slot 256 | biases b1[C] | b2[C] from 512
constexpr int R2_MAX_SMEM = 227 * 1024;
constexpr int R2_MAX_XS = 4;

struct Rp2Args {
    const bf16* X; bf16* Y; const bf16* R2;
    const bf16* W1; const bf16* W2; const float* b1; const float* b2;     // W1 / W2: [2 halves][K][C/8][C/2][8]
    int B, T, dil, tiles_per_b, total_tiles, n_pairs, Rx, r_out, p1, p2;
    float slope, out_scale, post_slope;
    int res_pf, dbg;
};

// MASG_TC_DEBUG & 32: timeline of the leader CTA of cluster 0 (clock64 at the hand-over points of its first R2_TRACE_TILES tiles).
// roles: 0 unused | 1 worker 0 (loads converted, stored) | 2 issuer 0 (P1 ready, P1 issued, P2 ready,
// P2 issued) | 3 E1 warp 0 (acc1 full, drained, h free, h written) | 4 E2 warp 0 (residuals requested, acc2 full, stored)
constexpr int R2_TRACE_TILES = 96, R2_TRACE_EVTS = 4, R2_TRACE_ROLES = 5;
__device__ long long g_rp2_trace[R2_TRACE_ROLES * R2_TRACE_TILES * R2_TRACE_EVTS];
#define R2_TRACE(on, role, it, evt) do { if ((on) && (it) < R2_TRACE_TILES) g_rp2_trace[((role) * R2_TRACE_TILES + (it)) * R2_TRACE_EVTS + (evt)] = clock64(); } while (0)

struct Rp2Bars {
    uint32_t base;
    __device__ uint32_t x_raw(int i) const { return base + 8u * i; }                          // local: bulk copy landed
    __device__ uint32_t x_full(int i) const { return base + 8u * (R2_MAX_XS + i); }            // leader: leaky-relu done in both CTAs
    __device__ uint32_t x_empty(int i) const { return base + 8u * (2 * R2_MAX_XS + i); }       // multicast: conv 1 has read the slot
    __device__ uint32_t acc1_full(int i) const { return base + 8u * (3 * R2_MAX_XS + i); }     // multicast
    __device__ uint32_t acc1_empty(int i) const { return base + 8u * (3 * R2_MAX_XS + 2 + i); }   // leader
    __device__ uint32_t h_full(int i) const { return base + 8u * (3 * R2_MAX_XS + 4 + i); }    // leader
    __device__ uint32_t h_empty(int i) const { return base + 8u * (3 * R2_MAX_XS + 6 + i); }   // multicast
    __device__ uint32_t acc2_full(int i) const { return base + 8u * (3 * R2_MAX_XS + 8 + i); } // multicast
    __device__ uint32_t acc2_empty(int i) const { return base + 8u * (3 * R2_MAX_XS + 10 + i); }  // leader
    __device__ uint32_t w_ready() const { return base + 8u * (3 * R2_MAX_XS + 12); }           // local: this CTA's weight halves landed
    __device__ uint32_t w_peer() const { return base + 8u * (3 * R2_MAX_XS + 13); }            // leader: the peer's have
};
static_assert(8 * (3 * R2_MAX_XS + 14) <= 256, "barrier block overlaps the TMEM slot");

// XS activation slots, NH buffers of h, MS sub-tiles of 128 rows per CTA tile
template <int C, int K, int MS, int XS, int NH>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(R2_THREADS, 1) k_respair2_tc(const Rp2Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int CG = C / 8;                               // 8-channel groups
    constexpr int KS = C / 16;                              // K = 16 steps per tap
    constexpr int NHALF = C / 2;                            // weight columns held by this CTA
    constexpr int WBYTES = K * C * NHALF * 2;               // one conv's half
    constexpr int HROWS = MS * 128 + 16;                    // rows of h + the (k - 1) rows the last sub-tile's taps reach past them
    constexpr int HBYTES = CG * HROWS * 16;
    constexpr int NI = MS >= 2 ? 2 : 1;                     // commits per multicast barrier: MS >= 2 both issuers (sub-tiles w, w + 2, ... each), MS = 1 the one issuer of that conv
    constexpr int MSL = MS >= 2 ? MS / 2 : 1;               // sub-tiles per issuer
    static_assert(4 * MS * C == 512, "TMEM: acc1[2][MS][C] | acc2[2][MS][C] must be the 512 columns");
    static_assert(XS <= R2_MAX_XS && NH <= 2 && 512 + 2 * C * 4 <= R2_HDR, "header layout");
    const Rp2Bars bars{smem_u32(smem)};
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + 256);
    float* bias_s = reinterpret_cast<float*>(smem + 512);   // b1[C] | b2[C]
    const int x_bytes = CG * a.Rx * 16;
    uint8_t* Xs = smem + R2_HDR;
    uint8_t* Hs = Xs + XS * (size_t)x_bytes;
    uint8_t* W1s = Hs + NH * HBYTES;
    uint8_t* W2s = W1s + WBYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int cid = (int)blockIdx.x >> 1, n_clusters = (int)gridDim.x >> 1;
    const bool trace = (a.dbg & 32) && (int)blockIdx.x == ((a.dbg >> 8) & 0xff);   // MASG_TC_DEBUG = 32 + 256 n: trace CTA n
    const int n_my = a.n_pairs > cid ? (a.n_pairs - cid + n_clusters - 1) / n_clusters : 0;    // tile pairs of this cluster: the same in both CTAs

    if (threadIdx.x == 0) {
        for (int i = 0; i < XS; ++i) { mbar_init(bars.x_raw(i), CG >= R2_EPI_WARPS ? R2_EPI_WARPS : CG); mbar_init(bars.x_full(i), 2 * R2_WORK_WARPS); mbar_init(bars.x_empty(i), NI); }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bars.acc1_full(i), NI); mbar_init(bars.acc1_empty(i), 2 * R2_EPI_WARPS);
            mbar_init(bars.h_full(i), 2 * R2_EPI_WARPS); mbar_init(bars.h_empty(i), NI);
            mbar_init(bars.acc2_full(i), NI); mbar_init(bars.acc2_empty(i), 2 * R2_EPI_WARPS);
        }
        mbar_init(bars.w_ready(), 1); mbar_init(bars.w_peer(), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 2 * C; i += R2_THREADS) bias_s[i] = i < C ? a.b1[i] : a.b2[i - C];
    if (warp == R2_MMA_WARP) {                              // the same warp of both CTAs: the pair's TMEM is allocated together
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                     // both CTAs' barriers are initialised before anyone arrives remotely
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp == R2_WORK_WARP0 && lane == 0) {               // this CTA's half of both weight sets, resident for the life of the CTA
        mbar_expect_tx(bars.w_ready(), 2u * WBYTES);
        bulk_g2s(smem_u32(W1s), reinterpret_cast<const uint8_t*>(a.W1) + (size_t)rank * WBYTES, WBYTES, bars.w_ready());
        bulk_g2s(smem_u32(W2s), reinterpret_cast<const uint8_t*>(a.W2) + (size_t)rank * WBYTES, WBYTES, bars.w_ready());
    }

    if (warp >= R2_WORK_WARP0) {
        // ===== workers: leaky-relu in place on the landed x tile (the raw x is also the residual, so the producer cannot apply it);
        // one remote arrival per warp on the leader's barrier.
        // (Loading the tile global -> registers -> leaky-relu -> shared memory instead — one shared-memory pass where copy + in-place
        // pass + operand read cost three — was tried: the loads of a tile can only be issued once the previous tile's registers are
        // free, their latency under load (~5,000 cycles from DRAM, still > 2,000 with the rows prefetched into L2) then sits in every
        // tile period: C = 128 k = 3 0.41 -> 0.49 ms.  The bulk copies run XS tiles ahead for free.) =====
        const int tid = threadIdx.x - R2_WORK_WARP0 * 32;
        const int pieces = x_bytes >> 4;
        const Slope2 sl = make_slope2(a.slope);
        for (int it = 0; it < n_my; ++it) {
            const int slot = it % XS, u = it / XS;
            mbar_wait_relaxed(bars.x_raw(slot), u & 1);
            R2_TRACE(trace && tid == 0, 1, it, 0);
            uint4* p = reinterpret_cast<uint4*>(Xs + (size_t)slot * x_bytes);
            // six pieces per thread in flight: the slot cycle (copy -> leaky-relu -> conv 1) is part of what sets the tile period
            for (int i = tid; i < pieces; i += 6 * R2_WORK_THREADS) {
                uint4 v[6];
#pragma unroll
                for (int e = 0; e < 6; ++e) if (i + e * R2_WORK_THREADS < pieces) v[e] = p[i + e * R2_WORK_THREADS];
#pragma unroll
                for (int e = 0; e < 6; ++e) {
                    v[e].x = lrelu_bf16x2(v[e].x, sl); v[e].y = lrelu_bf16x2(v[e].y, sl); v[e].z = lrelu_bf16x2(v[e].z, sl); v[e].w = lrelu_bf16x2(v[e].w, sl);
                }
#pragma unroll
                for (int e = 0; e < 6; ++e) if (i + e * R2_WORK_THREADS < pieces) p[i + e * R2_WORK_THREADS] = v[e];
            }
            fence_proxy_async_all();                        // generic-proxy stores -> visible to the tensor cores (async proxy)
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_rank(bars.x_full(slot), 0));
            R2_TRACE(trace && tid == 0, 1, it, 1);
        }
    } else if (warp >= R2_MMA_WARP && warp < R2_MMA_WARP + 2) {
        const uint32_t w = (uint32_t)(warp - R2_MMA_WARP);
        if (!leader) {
            // peer: tell the leader when this CTA's weight halves have landed; the leader's threads issue every MMA of the pair
            if (w == 0 && lane == 0) { mbar_wait(bars.w_ready(), 0); mbar_arrive_cluster(map_to_rank(bars.w_peer(), 0)); }
        } else {
            // ===== MMA issuers (leader CTA).  MS = 2: issuer w owns sub-tile w (rows [128 w, 128 w + 128) of each CTA's tile) of both
            // convs.  MS = 1: issuer 0 issues every conv 1, issuer 1 every conv 2 — one thread sustains ~54 cycles per MMA, so a single
            // issuer would spend 2,600 of the ~3,100 cycles the pair's MMAs take just issuing, and, issuing in order, would hold conv 2
            // (it) behind its wait for the x tile of conv 1 (it + 1) and vice versa =====
            // instruction descriptor: D fp32, A / B bf16, K-major, N = C, M = 256 (128 rows from each CTA of the pair)
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(C >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            const uint32_t desc_hi = (128u >> 4) | (1u << 14);                  // SBO = 128 B, descriptor version 1
            const uint32_t lbo_x = (uint32_t)a.Rx, dil = (uint32_t)a.dil;
            const uint32_t xs16 = smem_u32(Xs) >> 4, hs16 = smem_u32(Hs) >> 4, x16 = (uint32_t)x_bytes >> 4;
            // B: this CTA's C/2 columns, 8-channel groups C/2 rows apart; the peer's half is read at the same offsets of its memory
            const uint32_t w1_lo = ((uint32_t)NHALF << 16) | (smem_u32(W1s) >> 4), w2_lo = ((uint32_t)NHALF << 16) | (smem_u32(W2s) >> 4);
            const uint32_t sub = MS >= 2 ? w : 0u;
            mbar_wait(bars.w_ready(), 0);
            mbar_wait_cluster(bars.w_peer(), 0);
            auto P1 = [&](int it) {
                const int slot = it % XS, u = it / XS, b = it & 1, ub = it >> 1;
                mbar_wait_cluster(bars.x_full(slot), u & 1);
                mbar_wait_cluster(bars.acc1_empty(b), (ub & 1) ^ 1);
                tc_fence_after();
                R2_TRACE(trace && sub == 0 && lane == 0, 2, it, 0);
                if (elect_one()) {
                    const uint32_t a0 = (lbo_x << 16) | (xs16 + (uint32_t)slot * x16 + sub * 128u);
                    const uint32_t d = tmem_base + (uint32_t)((b * MS) * C) + sub * (uint32_t)C;
#pragma unroll
                    for (int j = 0; j < K; ++j) {
#pragma unroll
                        for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
                            for (int i = 0; i < MSL; ++i)       // this issuer's sub-tiles sub, sub + 2, ...: the same weight block back to back
                                tc_mma2(d + (uint32_t)(2 * i * C), a0 + (uint32_t)(2 * i) * 128u + (uint32_t)j * dil + (uint32_t)ks * 2u * lbo_x,
                                        w1_lo + (uint32_t)((j * KS + ks) * 2 * NHALF), desc_hi, idesc, (j | ks) ? 1u : 0u);
                        }
                    }
                    tc_commit2(bars.x_empty(slot));
                    tc_commit2(bars.acc1_full(b));
                }
                __syncwarp();
                R2_TRACE(trace && sub == 0 && lane == 0, 2, it, 1);
            };
            auto P2 = [&](int it) {
                const int b = it & 1, ub = it >> 1;
                const int hbuf = NH == 2 ? b : 0, hu = NH == 2 ? ub : it;
                mbar_wait_cluster(bars.h_full(hbuf), hu & 1);
                mbar_wait_cluster(bars.acc2_empty(b), (ub & 1) ^ 1);
                tc_fence_after();
                R2_TRACE(trace && sub == 0 && lane == 0, 2, it, 2);
                if (elect_one()) {
                    const uint32_t a0 = ((uint32_t)HROWS << 16) | (hs16 + (uint32_t)hbuf * (uint32_t)(HBYTES >> 4) + sub * 128u);
                    const uint32_t d = tmem_base + (uint32_t)(2 * MS * C + (b * MS) * C) + sub * (uint32_t)C;
#pragma unroll
                    for (int j = 0; j < K; ++j) {
#pragma unroll
                        for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
                            for (int i = 0; i < MSL; ++i)
                                tc_mma2(d + (uint32_t)(2 * i * C), a0 + (uint32_t)(2 * i) * 128u + (uint32_t)j + (uint32_t)(ks * 2 * HROWS),
                                        w2_lo + (uint32_t)((j * KS + ks) * 2 * NHALF), desc_hi, idesc, (j | ks) ? 1u : 0u);
                        }
                    }
                    tc_commit2(bars.h_empty(hbuf));
                    tc_commit2(bars.acc2_full(b));
                }
                __syncwarp();
                R2_TRACE(trace && sub == 0 && lane == 0, 2, it, 3);
            };
            if (MS >= 2) {
                if (n_my > 0) P1(0);
                for (int it = 0; it < n_my; ++it) {
                    if (it + 1 < n_my) P1(it + 1);
                    P2(it);
                }
            } else if (w == 0) {
                for (int it = 0; it < n_my; ++it) P1(it);
            } else {
                for (int it = 0; it < n_my; ++it) P2(it);
            }
        }
    } else if (warp < R2_EPI_WARPS) {
        // ===== E1 warps: acc1 (TMEM) -> + b1 -> bf16 -> leaky-relu -> h in shared memory (UMMA K-major layout) =====
        const int quad = warp & 3, hf = warp >> 2;
        constexpr int CH = C / 2;                            // channels per warp (this half)
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const Slope2 sl = make_slope2(a.slope);
        // ---- x tile copies: E1 warp w moves the GW 8-channel groups [w GW, (w + 1) GW) of every tile, one contiguous run per group
        // (planar bf16 [B][C/8][T][8]).  A cp.async.bulk costs its issuing warp ~350 cycles, whichever lane it comes from (2,600-3,400
        // cycles for a tile's 8 / 16 runs from one or two dedicated copy warps, all of it in the slot cycle), so the runs are spread
        // over these 8 warps; the copy of tile it + XS goes out when acc1 (it) is full: conv 1 (it) has then read the slot.
        constexpr int GW = CG >= R2_EPI_WARPS ? CG / R2_EPI_WARPS : 1;     // C = 32: four groups, E1 warps 0-3 issue one each
        constexpr int CW = CG >= R2_EPI_WARPS ? R2_EPI_WARPS : CG;          // warps that issue copies
        const uint8_t* X8 = reinterpret_cast<const uint8_t*>(a.X);
        auto issue_copy = [&](int itc) {
            if (warp >= CW) return;
            const int tile = 2 * (cid + itc * n_clusters) + (int)rank;
            const bool valid = tile < a.total_tiles;
            const int slot = itc % XS;
            const int bb = tile / a.tiles_per_b, i = tile - bb * a.tiles_per_b;
            uint8_t* dst_g = Xs + (size_t)slot * x_bytes;
            const uint32_t dst = smem_u32(dst_g), bar = bars.x_raw(slot);
            const int t0 = i * a.r_out - a.p2 - a.p1;
            const int lo = max(t0, 0), hi = valid ? max(min(t0 + a.Rx, a.T), lo) : lo;
            const uint32_t run = (uint32_t)(hi - lo) * 16u;
            const int g0 = warp * GW;
            if (valid && (lo > t0 || hi < t0 + a.Rx)) {      // conv zero padding at the utterance edges
                const int nlo = min(lo - t0, a.Rx), nhi0 = max(hi - t0, 0);
                for (int j = 0; j < GW; ++j) {
                    uint4* blk = reinterpret_cast<uint4*>(dst_g + (size_t)(g0 + j) * a.Rx * 16);
                    for (int r = lane; r < nlo; r += 32) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                    for (int r = nhi0 + lane; r < a.Rx; r += 32) blk[r] = make_uint4(0u, 0u, 0u, 0u);
                }
                fence_proxy_async_all();
                __syncwarp();
            }
            if (lane == 0) {
                if (run) {
                    mbar_expect_tx(bar, run * (uint32_t)GW);
                    const uint8_t* src = X8 + (((size_t)bb * CG + g0) * a.T + lo) * 16;
                    const uint32_t d0 = dst + (uint32_t)((g0 * a.Rx + (lo - t0)) * 16);
#pragma unroll
                    for (int j = 0; j < GW; ++j) bulk_g2s(d0 + (uint32_t)(j * a.Rx * 16), src + (size_t)j * a.T * 16, run, bar);
                } else mbar_arrive(bar);
                if (valid && a.R2 != nullptr && a.res_pf) {
                    // MRF accumulator rows of this tile -> L2 ahead of the E2 warps (DRAM loads queue behind the copies)
                    const int q0 = i * a.r_out, rows = min(a.r_out, a.T - q0);
                    if (rows > 0)
                        for (int j = 0; j < GW; ++j) bulk_prefetch_l2(a.R2 + (((size_t)bb * CG + g0 + j) * a.T + q0) * 8, (uint32_t)rows * 16u);
                }
            }
            __syncwarp();
        };
        for (int itc = 0; itc < XS && itc < n_my; ++itc) issue_copy(itc);
        for (int it = 0; it < n_my; ++it) {
            const int tile = 2 * (cid + it * n_clusters) + (int)rank;
            const int b = it & 1, ub = it >> 1;
            const int i = tile % a.tiles_per_b;
            const int tq = i * a.r_out - a.p2;               // time of h row 0
            const bool edge = tq < 0 || tq + MS * 128 > a.T; // some h rows fall outside [0, T): they must be zeros (conv 2 padding)
            const int hbuf = NH == 2 ? b : 0, hu = NH == 2 ? ub : it;
            // The accumulator is drained into packed bf16 registers BEFORE the h buffer is waited for: with one h buffer the chain
            // conv 2 (it - 1) done -> E1 (it) -> conv 2 (it) sets the tile period, and this keeps the TMEM loads, the bias and the
            // leaky-relu out of it (acc1 (it) has been full since before conv 2 (it - 1) was issued)
            mbar_wait_relaxed(bars.acc1_full(b), ub & 1);
            tc_fence_after();
            R2_TRACE(trace && threadIdx.x == 0, 3, it, 0);
            if (it + XS < n_my) {                            // the slot conv 1 (it) read is free: start the copy of tile it + XS
                mbar_wait_relaxed(bars.x_empty(it % XS), (it / XS) & 1);
                issue_copy(it + XS);
                R2_TRACE(trace && threadIdx.x == 0, 0, it + XS, 2);
            }
            uint4 pk[MS][CH / 8];
            constexpr int NLD = MS * (CH / 16) / 2;             // two 16-column loads in flight per wait: items (cg, ms) = (t / MS, t % MS), t = 2 q, 2 q + 1
#pragma unroll
            for (int q = 0; q < NLD; ++q) {
                float v[2][16];
                const int cga = (2 * q) / MS, cgb = (2 * q + 1) / MS;
                const int msa = (2 * q) % MS, msb = (2 * q + 1) % MS;
                tmem_ld16x2(tlane + (uint32_t)((b * MS + msa) * C + hf * CH + cga * 16), tlane + (uint32_t)((b * MS + msb) * C + hf * CH + cgb * 16), v[0], v[1]);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int cg = e ? cgb : cga, ms = e ? msb : msa;
                    const float4* b4 = reinterpret_cast<const float4*>(bias_s + hf * CH + cg * 16);
                    const float4 ba = b4[0], bb = b4[1], bc = b4[2], bd = b4[3];
                    float* x = v[e];
                    add_f32x2(x[0], x[1], ba.x, ba.y); add_f32x2(x[2], x[3], ba.z, ba.w); add_f32x2(x[4], x[5], bb.x, bb.y); add_f32x2(x[6], x[7], bb.z, bb.w);
                    add_f32x2(x[8], x[9], bc.x, bc.y); add_f32x2(x[10], x[11], bc.z, bc.w); add_f32x2(x[12], x[13], bd.x, bd.y); add_f32x2(x[14], x[15], bd.z, bd.w);
                    uint4 pa = make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
                    uint4 pb = make_uint4(pack_bf16(x[8], x[9]), pack_bf16(x[10], x[11]), pack_bf16(x[12], x[13]), pack_bf16(x[14], x[15]));
                    pa.x = lrelu_bf16x2(pa.x, sl); pa.y = lrelu_bf16x2(pa.y, sl); pa.z = lrelu_bf16x2(pa.z, sl); pa.w = lrelu_bf16x2(pa.w, sl);
                    pb.x = lrelu_bf16x2(pb.x, sl); pb.y = lrelu_bf16x2(pb.y, sl); pb.z = lrelu_bf16x2(pb.z, sl); pb.w = lrelu_bf16x2(pb.w, sl);
                    if (edge) {
                        const int t = tq + ms * 128 + quad * 32 + lane;
                        if (t < 0 || t >= a.T) { pa = make_uint4(0u, 0u, 0u, 0u); pb = pa; }
                    }
                    pk[ms][cg * 2] = pa; pk[ms][cg * 2 + 1] = pb;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_relaxed(map_to_rank(bars.acc1_empty(b), 0));   // orders only the TMEM reads above
            R2_TRACE(trace && threadIdx.x == 0, 3, it, 1);
            mbar_wait_relaxed(bars.h_empty(hbuf), (hu & 1) ^ 1);
            R2_TRACE(trace && threadIdx.x == 0, 3, it, 2);
            uint8_t* hb = Hs + (size_t)hbuf * HBYTES;
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
                const int r = ms * 128 + quad * 32 + lane;
#pragma unroll
                for (int g = 0; g < CH / 8; ++g)
                    *reinterpret_cast<uint4*>(hb + ((size_t)(hf * (CH / 8) + g) * HROWS + r) * 16) = pk[ms][g];
            }
            fence_proxy_async_all();                        // h stores -> visible to the tensor cores
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_rank(bars.h_full(hbuf), 0));
            R2_TRACE(trace && threadIdx.x == 0, 3, it, 3);
        }
    } else if (warp < 2 * R2_EPI_WARPS) {
        // ===== E2 warps: acc2 (TMEM) -> + b2 + x [+ acc] -> scale -> trailing leaky-relu -> planar bf16 y =====
        const int quad = warp & 3, hf = (warp - R2_EPI_WARPS) >> 2;
        constexpr int CH = C / 2;
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const long gstride = (long)a.T * 8;                  // elements between 8-channel groups
        const bool has_r2 = a.R2 != nullptr;
        const Slope2 sl_post = make_slope2(a.post_slope);
        // (Two groups of four warps on alternate tiles, each warp covering all channels of its quadrant in two passes, was tried when
        // this pass looked like the limiter — 4,400 cycles per tile, never waiting: no change; the kernel is bound by shared-memory
        // bandwidth, see the header.)
        for (int it = 0; it < n_my; ++it) {
            const int tile = 2 * (cid + it * n_clusters) + (int)rank;
            const bool valid = tile < a.total_tiles;
            const int b = it & 1, ub = it >> 1;
            const int bb = tile / a.tiles_per_b, i = tile - bb * a.tiles_per_b;
            const int q0 = i * a.r_out;
            const long boff = ((long)bb * CG + hf * (CH / 8)) * gstride;
            // the residual rows (x again: L2 hits, the copy read them a few tiles ago) are requested before the accumulator is waited for
            uint4 rx[MS][CH / 8];
            bool ok[MS];
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
                const int r = ms * 128 + quad * 32 + lane, t = q0 + r;
                ok[ms] = valid && r < a.r_out && t < a.T;
#pragma unroll
                for (int g = 0; g < CH / 8; ++g)
                    rx[ms][g] = ok[ms] ? *reinterpret_cast<const uint4*>(a.X + boff + (long)g * gstride + (long)t * 8) : make_uint4(0u, 0u, 0u, 0u);
            }
            R2_TRACE(trace && threadIdx.x == R2_EPI_WARPS * 32, 4, it, 0);
            mbar_wait_relaxed(bars.acc2_full(b), ub & 1);
            tc_fence_after();
            R2_TRACE(trace && threadIdx.x == R2_EPI_WARPS * 32, 4, it, 1);
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
                const int t = q0 + ms * 128 + quad * 32 + lane;
#pragma unroll
                for (int cg = 0; cg < CH / 16; ++cg) {
                    const int col0 = hf * CH + cg * 16;
                    float v[16];
                    tmem_ld16(tlane + (uint32_t)(2 * MS * C + (b * MS + ms) * C + col0), v);
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
                            if (ok[ms]) ur = *reinterpret_cast<const uint4*>(a.R2 + boff + (long)(cg * 2 + h8) * gstride + (long)t * 8);
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
                    if (ok[ms]) {
                        const long off = boff + (long)(cg * 2) * gstride + (long)t * 8;
                        *reinterpret_cast<uint4*>(a.Y + off) = pa;
                        *reinterpret_cast<uint4*>(a.Y + off + gstride) = pb;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster_relaxed(map_to_rank(bars.acc2_empty(b), 0));   // orders only the TMEM reads above, not the global stores
            R2_TRACE(trace && threadIdx.x == R2_EPI_WARPS * 32, 4, it, 2);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                     // neither CTA's shared memory / TMEM goes away while the pair's MMAs can still touch it
    if (warp == R2_MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

struct Rp2Config { int ms, xs, nh; };
bool rp2_config(int C, int K, Rp2Config& cf) {
    if (C == 128 && K == 3) { cf = {1, 2, 1}; return true; }
    if (C == 64 && K == 11) { cf = {2, 2, 1}; return true; }
    if (C == 32 && (K == 3 || K == 7 || K == 11)) {          // fits one SM (respair_tc.cu, 6 slots + 2 h buffers, 256-row tiles); here 512-row tiles, 3 + 2.  MASG_RP2_C32=0: off
        static int on = -1;
        if (on < 0) { const char* e = getenv("MASG_RP2_C32"); on = e ? atoi(e) : 0; }   // measured: k = 3 / 11 2 % faster, k = 7 6 % slower than respair_tc: off
        if (on) { cf = {4, 3, 2}; return true; }
    }
    if (C == 64 && K == 7) {                                // fits one SM too (respair_tc.cu, 2 slots + 1 h buffer); here 2 + 2.  MASG_RP2_K7=0: off
        static int on = -1;
        if (on < 0) { const char* e = getenv("MASG_RP2_K7"); on = e ? atoi(e) : 1; }
        if (on) { cf = {2, 2, 2}; return true; }
    }
    return false;
}
int rp2_rx(int ms, const ConvDesc& d1) { return (ms * 128 + (d1.taps - 1) * d1.tstep + 7) & ~7; }
size_t rp2_smem(int C, int K, int Rx, const Rp2Config& cf) {
    return (size_t)R2_HDR + (size_t)cf.xs * (C / 8) * Rx * 16 + (size_t)cf.nh * (C / 8) * (cf.ms * 128 + 16) * 16 + 2 * (size_t)K * C * (C / 2) * 2;
}

template <int C, int K, int MS, int XS, int NH>
void launch_rp2(const Rp2Args& a, dim3 grid, size_t smem, cudaStream_t s) {
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_respair2_tc<C, K, MS, XS, NH>, cudaFuncAttributeMaxDynamicSharedMemorySize, R2_MAX_SMEM));
    k_respair2_tc<C, K, MS, XS, NH><<<grid, R2_THREADS, smem, s>>>(a);
}

}  // namespace

int respair2_debug_trace(long long* out, int max_elems) {
    const int n = R2_TRACE_ROLES * R2_TRACE_TILES * R2_TRACE_EVTS;
    if (max_elems < n) return -1;
    CUDA_OK(cudaMemcpyFromSymbol(out, g_rp2_trace, sizeof(long long) * n));
    return R2_TRACE_TILES;
}

bool respair2_tc_supported(const masg_ctx* c, const ConvDesc& d1, const ConvDesc& d2) {
    static int off = -1;
    if (off < 0) { const char* e = getenv("MASG_NO_RESPAIR2"); off = e ? atoi(e) : 0; }
    if (off || c->precision != MASG_PREC_BF16) return false;
    const int C = d1.Cin, K = d1.taps;
    Rp2Config cf;
    if (!rp2_config(C, K, cf)) return false;
    if (d1.N != C || d2.Cin != C || d2.N != C || d2.taps != K || d2.tstep != 1 || d1.tstep < 1) return false;
    if (d1.off0 != -((K - 1) * d1.tstep) / 2 || d2.off0 != -(K - 1) / 2) return false;
    if (d1.split || d2.split || d1.f16 || d2.f16 || !d1.w16p || !d2.w16p) return false;
    if (d1.up || d2.up || (d1.bias_mod && d1.bias_mod != C) || (d2.bias_mod && d2.bias_mod != C)) return false;
    if ((K - 1) > 16) return false;
    return rp2_smem(C, K, rp2_rx(cf.ms, d1), cf) <= (size_t)R2_MAX_SMEM && c->n_sm >= 2;
}

void launch_respair2_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d1, const ConvDesc& d2, const bf16* X, int B, int T,
                        bf16* Y, const bf16* R2, float out_scale, float post_slope) {
    MASG_CHECK(respair2_tc_supported(c, d1, d2), "respair2_tc: unsupported pair (C=%d k=%d)", d1.Cin, d1.taps);
    MASG_CHECK(X != Y && (((uintptr_t)X | (uintptr_t)Y | (uintptr_t)R2) & 15) == 0, "respair2_tc: tensors must be distinct and 16-byte aligned");
    if (B <= 0 || T <= 0) return;
    const int C = d1.Cin, K = d1.taps;
    Rp2Config cf;
    rp2_config(C, K, cf);
    Rp2Args a;
    memset(&a, 0, sizeof(a));
    a.X = X; a.Y = Y; a.R2 = R2; a.W1 = d1.w16p; a.W2 = d2.w16p; a.b1 = d1.bias; a.b2 = d2.bias;
    a.B = B; a.T = T; a.dil = d1.tstep; a.p1 = -d1.off0; a.p2 = -d2.off0;
    a.r_out = cf.ms * 128 - (K - 1); a.Rx = rp2_rx(cf.ms, d1);
    a.tiles_per_b = (T + a.r_out - 1) / a.r_out; a.total_tiles = B * a.tiles_per_b; a.n_pairs = (a.total_tiles + 1) / 2;
    a.slope = 0.1f; a.out_scale = out_scale; a.post_slope = post_slope;
    { static int pf = -1; if (pf < 0) { const char* e = getenv("MASG_TC_RESPF"); pf = e ? atoi(e) : 1; } a.res_pf = pf; }
    { static int dbg = -1; if (dbg < 0) { const char* e = getenv("MASG_TC_DEBUG"); dbg = e ? atoi(e) : 0; } a.dbg = dbg; }
    const dim3 grid(2 * std::min(a.n_pairs, c->n_sm / 2), 1, 1);           // one CTA pair per TPC, persistent
    const size_t smem = rp2_smem(C, K, a.Rx, cf);
    masg_ctx::ProfRec rec;
    if (c->profiling) {
        CUDA_OK(cudaEventCreate(&rec.e0)); CUDA_OK(cudaEventCreate(&rec.e1));
        rec.flops = 2.0 * 2.0 * B * (double)T * K * C * C;               // both convs, algorithmic
        CUDA_OK(cudaEventRecord(rec.e0, s));
    }
    if (C == 128 && K == 3) launch_rp2<128, 3, 1, 2, 1>(a, grid, smem, s);
    else if (C == 64 && K == 11) launch_rp2<64, 11, 2, 2, 1>(a, grid, smem, s);
    else if (C == 64 && K == 7) launch_rp2<64, 7, 2, 2, 2>(a, grid, smem, s);
    else if (C == 32 && K == 3) launch_rp2<32, 3, 4, 3, 2>(a, grid, smem, s);
    else if (C == 32 && K == 7) launch_rp2<32, 7, 4, 3, 2>(a, grid, smem, s);
    else if (C == 32 && K == 11) launch_rp2<32, 11, 4, 3, 2>(a, grid, smem, s);
    CUDA_OK(cudaGetLastError());
    c->launches++;
    if (c->profiling) { CUDA_OK(cudaEventRecord(rec.e1, s)); c->prof.push_back(rec); }
}
