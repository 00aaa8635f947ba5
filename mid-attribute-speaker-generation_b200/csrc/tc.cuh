// tcgen05 (bf16) path entry points — implemented in conv_tc.cu / fs2_tc.cu / hifigan_tc.cu.
#pragma once
#include "common.cuh"

// bf16 K-major weight copy + TMA descriptor for a ConvDesc (called at commit in MASG_PREC_BF16 mode)
void tc_prepare_weights(masg_ctx* c, cudaStream_t s, ConvDesc& d);
void tc_fs2_encoder(masg_ctx* c, cudaStream_t s, Arena& ar, const int64_t* texts, const int64_t* accents, int acc1d,
                    const int64_t* src_lens, int B, int L, const float* pos, float* enc_out_f32);
void tc_predictor(masg_ctx* c, cudaStream_t s, Arena& ar, const Predictor& P, const float* x, const int64_t* lens,
                  int B, int L, float scale, float* out);
void tc_fs2_decoder_mel(masg_ctx* c, cudaStream_t s, Arena& ar, const float* lr_out, const float* pos, const int64_t* lens,
                        int B, int T, float* dec_out_f32, float* mel, float* post);
void tc_hifigan_forward(masg_ctx* c, cudaStream_t s, Arena& ar, const float* mel_btc, int B, int T, float* wav, int16_t* w16);
