// tcgen05 (bf16 operands, fp32 TMEM accumulators) implicit-GEMM path — conv_tc.cu.
#pragma once
#include "common.cuh"

// bf16 pre-tiled weight copy of a ConvDesc (called at commit in MASG_PREC_BF16 mode): fills d.w16/NT/KW.
// split: carry the weights as bf16 hi + lo and issue 3 MMAs per K step (FastSpeech2 layers; ~fp32 accuracy)
// mode 2: fp16 operands, one MMA per K step (11-bit mantissa; the FastSpeech2 decoder side, after the duration / bucket decisions)
// mode | TC_MODE_WIDE: N == 256 handled as ONE 256-column tile (the layers whose epilogue is a fused LayerNorm over the row)
constexpr int TC_MODE_WIDE = 0x100;
void tc_prepare_weights(masg_ctx* c, cudaStream_t s, ConvDesc& d, int mode);
// same row-GEMM contract as launch_conv_f32 (see ConvDesc / ConvIO in common.cuh)
void launch_conv_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io);
// MASG_TC_DEBUG & 16 wait-cycle counters of k_conv_tc (summed over CTAs): [0..3] issuer 0: acc_empty, a_full, w_full waits, total;
// [4,5] epilogue warp 0: acc_full wait, total; [6] copy warp a_empty wait; [7] weight warp w_empty wait; [8] worker a_raw wait
void tc_debug_stats(unsigned long long* out16, bool reset);
int tc_debug_trace(long long* out, int max_elems, bool reset);   // MASG_TC_DEBUG & 32: [5 roles][96 tiles][4 events] clock64 of CTA 0
// fused ResBlock1 pair y = (c2(lrelu(c1(lrelu(x)))) + x [+ R2]) * out_scale [-> lrelu(post_slope)] on planar bf16 [B][C/8][T][8]
// (respair_tc.cu); supported for C = 32 (k = 3, 7, 11) and C = 64 (k = 3, 7) — the pairs whose weights fit shared memory
bool respair_tc_supported(const masg_ctx* c, const ConvDesc& d1, const ConvDesc& d2);
void launch_respair_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d1, const ConvDesc& d2, const bf16* X, int B, int T,
                       bf16* Y, const bf16* R2, float out_scale, float post_slope);
int respair_debug_trace(long long* out, int max_elems);      // MASG_TC_DEBUG & 32: [5 roles][96 tiles][4 events] clock64 of CTA 0
// the same fused pair on a CTA pair (cluster of 2, tcgen05 cta_group::2, the weights' columns split over the two CTAs) for the
// pairs whose weights do not fit one SM: C = 128 k = 3 and C = 64 k = 11 (respair2_tc.cu)
bool respair2_tc_supported(const masg_ctx* c, const ConvDesc& d1, const ConvDesc& d2);
void launch_respair2_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d1, const ConvDesc& d2, const bf16* X, int B, int T,
                        bf16* Y, const bf16* R2, float out_scale, float post_slope);
int respair2_debug_trace(long long* out, int max_elems);     // MASG_TC_DEBUG & 32: the same timeline for the leader CTA of cluster 0
