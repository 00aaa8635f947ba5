// Fused HiFi-GAN ResBlock1 pair for sm_100a (MASG_PREC_BF16 path): one kernel computes
//     y = ( c2( lrelu( c1( lrelu(x) ) + b1 ) ) + b2 + x [+ acc] ) * scale      [then an optional trailing leaky-relu]
// i.e. one (convs1[m], convs2[m]) step of ResBlock1.forward (hifigan/models.py:96-103), with the MRF sum, the /3 and
// the leaky-relu that follows the block folded in as in hifigan.cu.  The intermediate h = lrelu(c1(..)) never leaves
// the SM: conv 1 accumulates in TMEM, the epilogue warps write h as bf16 straight into shared memory in the UMMA
// no-swizzle K-major layout, and conv 2 consumes it from there as its A operand.  Per pair the HBM traffic drops from
// five tensor passes (x, h out, h in, residual x, y) to two (x in, y out).
//
// Used for the narrow stages (C = 32, 64; k = 3, 7, 11), which are HBM-bound as separate convs.  Both convs' weights
// (k C C bf16 each, pre-tiled by tc_prepare_weights) stay RESIDENT in shared memory for the life of the CTA, so the MMA
// issue loop is a fully unrolled run of taps x K-steps with descriptors advanced by uniform adds.
//
// Tile = 256 rows of h (two M = 128 sub-tiles), from which R_out = 256 - (k - 1) output rows are valid:
//     x rows [q0 - p2 - p1, +256 + 2 p1)  ->  h rows [q0 - p2, +256)  ->  y rows [q0, q0 + R_out),   p1 = (k-1) d / 2, p2 = (k-1) / 2.
// h rows outside [0, T) are written as zeros (conv 2's zero padding applies to h, not to c1 of a padded x).
//
// Persistent, warp-specialised, one 640-thread CTA per SM; mbarrier rings connect the roles:
//   copy warp   : one 1-D bulk async copy per 8-channel group of the planar bf16 input [B][C/8][T][8] (3 slots)
//   8 workers   : leaky-relu in place on the landed tile (the raw x is also the residual, so the producer cannot apply it)
//   2 issuers   : issuer w owns sub-tile w of both convs; software-pipelined  P1(0) | P1(i+1) P2(i) ...
//   8 epilogue  : E1 (TMEM -> bias, lrelu, bf16 -> shared h) and E2 (TMEM -> bias, residuals, scale -> planar bf16 y)
// TMEM: acc1[2 buffers][2 sub-tiles][C] | acc2[2][2][C] = 8 C columns (256 / 512).
#include "common.cuh"
#include "tc.cuh"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int RP_EPI_WARPS = 8;             // warps 0-7 (warp % 4 = TMEM lane quadrant, warp / 4 = channel half)
constexpr int RP_MMA_WARP = 8;              // warps 8, 9: MMA issuers
constexpr int RP_ISSUERS = 2;
constexpr int RP_COPY_WARP = 10;
constexpr int RP_WORK_WARP0 = 12;           // warps 12-19: leaky-relu workers
constexpr int RP_WORK_WARPS = 8;
constexpr int RP_WORK_THREADS = RP_WORK_WARPS * 32;
constexpr int RP_THREADS = (RP_WORK_WARP0 + RP_WORK_WARPS) * 32;     // 640
constexpr int RP_XSLOTS = 3;
constexpr int RP_HROWS = 272;               // 256 rows of h + the (k - 1) rows the second sub-tile's taps reach past them
constexpr int RP_HDR = 1024;
constexpr int RP_MAX_SMEM = 227 * 1024;

struct RpArgs {
    const bf16* X; bf16* Y; const bf16* R2;
    const bf16* W1; const bf16* W2; const float* b1; const float* b2;
    int B, T, dil, tiles_per_b, total_tiles, Rx, r_out, p1, p2;
    float slope, out_scale, post_slope;
};

struct RpBars {
    uint32_t base;
    __device__ uint32_t x_raw(int i) const { return base + 8u * i; }
    __device__ uint32_t x_full(int i) const { return base + 8u * (3 + i); }
    __device__ uint32_t x_empty(int i) const { return base + 8u * (6 + i); }
    __device__ uint32_t acc1_full(int i) const { return base + 8u * (9 + i); }
    __device__ uint32_t acc1_empty(int i) const { return base + 8u * (11 + i); }
    __device__ uint32_t h_full(int i) const { return base + 8u * (13 + i); }
    __device__ uint32_t h_empty(int i) const { return base + 8u * (15 + i); }
    __device__ uint32_t acc2_full(int i) const { return base + 8u * (17 + i); }
    __device__ uint32_t acc2_empty(int i) const { return base + 8u * (19 + i); }
    __device__ uint32_t w_ready() const { return base + 8u * 21; }
};

template <int C, int K>
__global__ void __launch_bounds__(RP_THREADS, 1) k_respair_tc(const RpArgs a) {
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
    uint8_t* W1s = Hs + 2 * HBYTES;
    uint8_t* W2s = W1s + WBYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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

    if (warp >= RP_WORK_WARP0) {
        // ===== workers: leaky-relu in place on the landed x tile =====
        const int tid = threadIdx.x - RP_WORK_WARP0 * 32;
        const int pieces = x_bytes >> 4;
        const Slope2 sl = make_slope2(a.slope);
        for (int it = 0; it < n_my; ++it) {
            const int slot = it % RP_XSLOTS, u = it / RP_XSLOTS;
            mbar_wait_relaxed(bars.x_raw(slot), u & 1);
            uint4* p = reinterpret_cast<uint4*>(Xs + (size_t)slot * x_bytes);
            for (int i = tid; i < pieces; i += RP_WORK_THREADS) {
                uint4 v = p[i];
                v.x = lrelu_bf16x2(v.x, sl); v.y = lrelu_bf16x2(v.y, sl); v.z = lrelu_bf16x2(v.z, sl); v.w = lrelu_bf16x2(v.w, sl);
                p[i] = v;
            }
            fence_proxy_async();                            // generic-proxy stores -> visible to the tensor core (async proxy)
            mbar_arrive(bars.x_full(slot));
        }
    } else if (warp == RP_COPY_WARP) {
        // ===== copy warp: resident weights once, then one contiguous run per 8-channel group per tile =====
        if (lane == 0) {
            mbar_expect_tx(bars.w_ready(), 2u * WBYTES);
            bulk_g2s(smem_u32(W1s), a.W1, WBYTES, bars.w_ready());
            bulk_g2s(smem_u32(W2s), a.W2, WBYTES, bars.w_ready());
        }
        const uint8_t* X8 = reinterpret_cast<const uint8_t*>(a.X);
        for (int it = 0; it < n_my; ++it) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int slot = it % RP_XSLOTS, u = it / RP_XSLOTS;
            const int b = tile / a.tiles_per_b, i = tile - b * a.tiles_per_b;
            if (lane == 0) mbar_wait_relaxed(bars.x_empty(slot), (u & 1) ^ 1);
            __syncwarp();
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
        };
        auto P2 = [&](int it) {
            const int b = it & 1, ub = it >> 1;
            mbar_wait(bars.h_full(b), ub & 1);
            mbar_wait(bars.acc2_empty(b), (ub & 1) ^ 1);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t a0 = ((uint32_t)RP_HROWS << 16) | (hs16 + (uint32_t)b * (uint32_t)(HBYTES >> 4) + w * 128u);
                const uint32_t d = tmem_base + (uint32_t)(4 * C + (b * 2) * C) + w * (uint32_t)C;
#pragma unroll
                for (int j = 0; j < K; ++j) {
#pragma unroll
                    for (int ks = 0; ks < KS; ++ks)
                        tc_mma(d, a0 + (uint32_t)j + (uint32_t)(ks * 2 * RP_HROWS), w2_lo + (uint32_t)((j * KS + ks) * 2 * C), desc_hi, idesc, (j | ks) ? 1u : 0u);
                }
                tc_commit(bars.h_empty(b));
                tc_commit(bars.acc2_full(b));
            }
            __syncwarp();
        };
        if (n_my > 0) P1(0);
        for (int it = 0; it < n_my; ++it) {
            if (it + 1 < n_my) P1(it + 1);
            P2(it);
        }
    } else if (warp < RP_EPI_WARPS) {
        // ===== epilogue warps =====
        const int quad = warp & 3, hf = warp >> 2;
        constexpr int CH = C / 2;                            // channels per warp (this half)
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const long gstride = (long)a.T * 8;                  // elements between 8-channel groups
        const bool has_r2 = a.R2 != nullptr;
        auto E1 = [&](int it) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int b = it & 1, ub = it >> 1;
            const int i = tile % a.tiles_per_b;
            mbar_wait_relaxed(bars.acc1_full(b), ub & 1);
            mbar_wait_relaxed(bars.h_empty(b), (ub & 1) ^ 1);
            tc_fence_after();
            const int tq = i * a.r_out - a.p2;               // time of h row 0
            uint8_t* hb = Hs + (size_t)b * HBYTES;
#pragma unroll
            for (int ms = 0; ms < 2; ++ms) {
                const int r = ms * 128 + quad * 32 + lane;
                const int t = tq + r;
                const bool inb = t >= 0 && t < a.T;
#pragma unroll
                for (int cg = 0; cg < CH / 16; ++cg) {
                    const int col0 = hf * CH + cg * 16;
                    float v[16];
                    tmem_ld16(tlane + (uint32_t)((b * 2 + ms) * C + col0), v);
#pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        float x = v[k] + bias_s[col0 + k];
                        x = fmaxf(x, x * a.slope);
                        v[k] = inb ? x : 0.f;
                    }
                    uint4* dst = reinterpret_cast<uint4*>(hb + ((size_t)(col0 >> 3) * RP_HROWS + r) * 16);
                    dst[0] = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                    dst[RP_HROWS] = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                }
            }
            fence_proxy_async();                            // h stores -> visible to the tensor core
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { mbar_arrive(bars.h_full(b)); mbar_arrive(bars.acc1_empty(b)); }
        };
        auto E2 = [&](int it) {
            const int tile = blockIdx.x + it * gridDim.x;
            const int b = it & 1, ub = it >> 1;
            const int bb = tile / a.tiles_per_b, i = tile - bb * a.tiles_per_b;
            const int q0 = i * a.r_out;
            const long boff = (long)bb * CG * gstride;
            // residuals of both sub-tiles are requested before the accumulator is waited for
            uint4 rx[2][CH / 8], rr[2][CH / 8];
            bool ok[2];
#pragma unroll
            for (int ms = 0; ms < 2; ++ms) {
                const int r = ms * 128 + quad * 32 + lane, t = q0 + r;
                ok[ms] = r < a.r_out && t < a.T;
#pragma unroll
                for (int g = 0; g < CH / 8; ++g) {
                    const long off = boff + (long)(hf * (CH / 8) + g) * gstride + (long)t * 8;
                    rx[ms][g] = make_uint4(0u, 0u, 0u, 0u); rr[ms][g] = rx[ms][g];
                    if (ok[ms]) {
                        rx[ms][g] = *reinterpret_cast<const uint4*>(a.X + off);
                        if (has_r2) rr[ms][g] = *reinterpret_cast<const uint4*>(a.R2 + off);
                    }
                }
            }
            mbar_wait_relaxed(bars.acc2_full(b), ub & 1);
            tc_fence_after();
#pragma unroll
            for (int ms = 0; ms < 2; ++ms) {
                const int r = ms * 128 + quad * 32 + lane, t = q0 + r;
#pragma unroll
                for (int cg = 0; cg < CH / 16; ++cg) {
                    const int col0 = hf * CH + cg * 16;
                    float v[16];
                    tmem_ld16(tlane + (uint32_t)(4 * C + (b * 2 + ms) * C + col0), v);
#pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] += bias_s[C + col0 + k];
#pragma unroll
                    for (int h8 = 0; h8 < 2; ++h8) {
                        const uint4 ux = rx[ms][cg * 2 + h8], ur = rr[ms][cg * 2 + h8];
                        const float2 x0 = unpack_bf16(ux.x), x1 = unpack_bf16(ux.y), x2 = unpack_bf16(ux.z), x3 = unpack_bf16(ux.w);
                        float* vv = v + h8 * 8;
                        vv[0] += x0.x; vv[1] += x0.y; vv[2] += x1.x; vv[3] += x1.y; vv[4] += x2.x; vv[5] += x2.y; vv[6] += x3.x; vv[7] += x3.y;
                        if (has_r2) {
                            const float2 r0 = unpack_bf16(ur.x), r1 = unpack_bf16(ur.y), r2 = unpack_bf16(ur.z), r3 = unpack_bf16(ur.w);
                            vv[0] += r0.x; vv[1] += r0.y; vv[2] += r1.x; vv[3] += r1.y; vv[4] += r2.x; vv[5] += r2.y; vv[6] += r3.x; vv[7] += r3.y;
                        }
                    }
                    if (a.out_scale != 1.f) {
#pragma unroll
                        for (int k = 0; k < 16; ++k) v[k] *= a.out_scale;
                    }
                    if (a.post_slope != 1.f) {
#pragma unroll
                        for (int k = 0; k < 16; ++k) v[k] = fmaxf(v[k], v[k] * a.post_slope);
                    }
                    if (ok[ms]) {
                        const long off = boff + (long)(col0 >> 3) * gstride + (long)t * 8;
                        *reinterpret_cast<uint4*>(a.Y + off) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                        *reinterpret_cast<uint4*>(a.Y + off + gstride) = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bars.acc2_empty(b));
        };
        if (n_my > 0) E1(0);
        for (int it = 0; it < n_my; ++it) {
            if (it + 1 < n_my) E1(it + 1);
            E2(it);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == RP_MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(8 * C) : "memory");
    }
}

template <int C, int K>
void launch_rp(const RpArgs& a, dim3 grid, size_t smem, cudaStream_t s) {
    static bool attr = false;
    if (!attr) { CUDA_OK(cudaFuncSetAttribute(k_respair_tc<C, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, RP_MAX_SMEM)); attr = true; }
    k_respair_tc<C, K><<<grid, RP_THREADS, smem, s>>>(a);
}

size_t rp_smem(int C, int K, int Rx) {
    return (size_t)RP_HDR + (size_t)RP_XSLOTS * (C / 8) * Rx * 16 + 2 * (size_t)(C / 8) * RP_HROWS * 16 + 2 * (size_t)K * C * C * 2;
}
int rp_rx(const ConvDesc& d1) { return (256 + (d1.taps - 1) * d1.tstep + 7) & ~7; }

}  // namespace

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
    return rp_smem(C, K, rp_rx(d1)) <= (size_t)RP_MAX_SMEM;
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
    static int n_sm = 0;
    if (!n_sm) { int dev = 0; CUDA_OK(cudaGetDevice(&dev)); CUDA_OK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)); }
    const dim3 grid(std::min(a.total_tiles, n_sm), 1, 1);
    const size_t smem = rp_smem(C, K, a.Rx);
    masg_ctx::ProfRec rec;
    if (c->profiling) {
        CUDA_OK(cudaEventCreate(&rec.e0)); CUDA_OK(cudaEventCreate(&rec.e1));
        rec.flops = 2.0 * 2.0 * B * (double)T * K * C * C;               // both convs, algorithmic
        CUDA_OK(cudaEventRecord(rec.e0, s));
    }
#define RP_CASE(CV, KV) if (C == CV && K == KV) launch_rp<CV, KV>(a, grid, smem, s);
    RP_CASE(32, 3) RP_CASE(32, 7) RP_CASE(32, 11) RP_CASE(64, 3) RP_CASE(64, 7) RP_CASE(64, 11)
#undef RP_CASE
    CUDA_OK(cudaGetLastError());
    c->launches++;
    if (c->profiling) { CUDA_OK(cudaEventRecord(rec.e1, s)); c->prof.push_back(rec); }
}
