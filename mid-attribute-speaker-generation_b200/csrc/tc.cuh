// tcgen05 (bf16 operands, fp32 TMEM accumulators) implicit-GEMM path — conv_tc.cu.
#pragma once
#include "common.cuh"

// bf16 pre-tiled weight copy of a ConvDesc (called at commit in MASG_PREC_BF16 mode): fills d.w16/NT/KW.
// split: carry the weights as bf16 hi + lo and issue 3 MMAs per K step (FastSpeech2 layers; ~fp32 accuracy)
void tc_prepare_weights(masg_ctx* c, cudaStream_t s, ConvDesc& d, bool split);
// same row-GEMM contract as launch_conv_f32 (see ConvDesc / ConvIO in common.cuh)
void launch_conv_tc(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io);
// MASG_TC_DEBUG & 16 wait-cycle counters of k_conv_tc (summed over CTAs): [0..3] issuer 0: acc_empty, a_full, w_full waits, total;
// [4,5] epilogue warp 0: acc_full wait, total; [6] copy warp a_empty wait; [7] weight warp w_empty wait; [8] worker a_raw wait
void tc_debug_stats(unsigned long long* out16, bool reset);
