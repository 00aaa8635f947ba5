// TEMPORARY: bf16 tcgen05 path not linked yet.
#include "tc.cuh"
void tc_prepare_weights(masg_ctx*, cudaStream_t, ConvDesc&) {}
void tc_fs2_encoder(masg_ctx*, cudaStream_t, Arena&, const int64_t*, const int64_t*, int, const int64_t*, int, int, const float*, float*) { MASG_CHECK(false, "bf16 path not built"); }
void tc_predictor(masg_ctx*, cudaStream_t, Arena&, const Predictor&, const float*, const int64_t*, int, int, float, float*) { MASG_CHECK(false, "bf16 path not built"); }
void tc_fs2_decoder_mel(masg_ctx*, cudaStream_t, Arena&, const float*, const float*, const int64_t*, int, int, float*, float*, float*) { MASG_CHECK(false, "bf16 path not built"); }
void tc_hifigan_forward(masg_ctx*, cudaStream_t, Arena&, const float*, int, int, float*, int16_t*) { MASG_CHECK(false, "bf16 path not built"); }
