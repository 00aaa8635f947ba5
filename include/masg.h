/*
 * masg.h — C ABI of libmasg.so: the B200 (sm_100a) batched synthesis path for
 * sarulab-speech/Mid-Attribute-Speaker-Generation.
 *
 * The reference has no FFI/plugin interface of its own (it is pure Python/PyTorch);
 * the drop-in boundary is the Python call surface of two nn.Modules and three distribution helpers
 * (SURVEY.md §8(b)).  Each entry point below names the reference callable it stands behind
 * (file:line under the reference tree).  The Python shims in
 * mid-attribute-speaker-generation_b200/{fastspeech2,hifigan,distributions}.py keep the reference
 * signatures and forward here through ctypes; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain C symbols, no exceptions cross the boundary; every call returns 0 on success, non-zero
 *     on error, and masg_last_error(ctx) returns a message (the shim raises RuntimeError).
 *   - all pointers are DEVICE pointers on the context's device unless the name starts with h_.
 *   - all work is enqueued on the cudaStream_t passed as `void* stream` (0 = legacy default
 *     stream); no hidden device-wide synchronisation.  The only host syncs are the documented ones
 *     (phase A publishes T_max; the *_host entry points copy results back).
 *   - inputs/outputs are caller-owned; the library owns its repacked weights and a workspace arena.
 *   - a context is not thread-safe; contexts are independent.  One context per (device, stream).
 *   - no CPU fallback and no alternate backends: an unsupported configuration is an error.
 */
#ifndef MASG_H_
#define MASG_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct masg_ctx masg_ctx;

enum { MASG_PREC_BF16 = 0,   /* bf16 / fp16 operands on tcgen05 tensor cores, fp32 accumulate (mel max-abs <= 1e-2): the benchmarked mode */
       MASG_PREC_FP32 = 1,   /* fp32 CUDA-core path (mel max-abs <= 1e-4); the strict-parity mode                                          */
       MASG_PREC_BF16X3 = 2 }; /* the 1e-4 mode ON the tensor cores: every FastSpeech2 contraction in split-bf16 (x = hi + lo on both
                                  operands, three MMAs per K step, ~16 mantissa bits, fp32 accumulate); HiFi-GAN as in MASG_PREC_BF16 */

/* library / build identification (no reference counterpart) */
const char* masg_version(void);

/* utils/model.py:11-34 get_model → one context per device.  precision: MASG_PREC_*. */
int  masg_ctx_create(int device, int precision, masg_ctx** out);
void masg_ctx_destroy(masg_ctx* ctx);
const char* masg_last_error(masg_ctx* ctx);   /* ctx may be NULL: returns the last ctx-less error */

/* nn.Module.load_state_dict (utils/model.py:21 for FastSpeech2, :66 for hifigan.Generator):
 * register one caller-owned fp32 (or int64 for num_batches_tracked: ignored) device tensor under its
 * reference state_dict key.  `model` is "fs2" or "hifigan".  HiFi-GAN accepts
 * either weight-normed keys (weight_g/weight_v) or folded ones (weight).  Tensors are read at commit. */
int masg_load_tensor(masg_ctx* ctx, const char* model, const char* key, const float* dev_ptr,
                     int ndim, const int64_t* shape);
/* Repack the registered tensors into kernel layouts (QKV concat, conv weights to [k][Cin][Cout],
 * BatchNorm folded (transformer/Layers.py:129-137), weight-norm folded = Generator.remove_weight_norm,
 * hifigan/models.py:167-174).  Re-callable after another load_state_dict. */
int masg_fs2_commit(masg_ctx* ctx, void* stream);
int masg_hifigan_commit(masg_ctx* ctx, void* stream);

/* FastSpeech2.forward / synthesize_from_speaker_emb, first half (model/fastspeech2.py:187-241):
 * get_mask_from_lengths → Encoder → + speaker embedding → duration/pitch/energy predictors →
 * bucketize+embedding → duration rounding → mel lengths.  Performs ONE stream synchronisation to
 * publish max(mel_lens) to *h_tmax (host).
 *   texts, accents : int64 [B, L]   (accents_is_1d != 0: accents is [L], broadcast over the batch,
 *                                    synthesize.py:234)
 *   src_lens       : int64 [B];  spk_emb: fp32 [B,256]
 *   p_ctl, e_ctl, d_ctl: controls; e_ctl is accepted and IGNORED, energy is scaled by p_ctl exactly
 *                        as model/modules.py:123-125 does.
 *   d_target       : nullable fp32 [B, L]: teacher-forced durations (modules.py:128-130)
 *   p_target, e_target: nullable fp32 [B, L]: teacher-forced pitch / energy — the embedding is then looked up from
 *                        bucketize(target) and the prediction is returned unscaled (modules.py:80-100)
 *   out: pitch, energy, log_d, d_rounded fp32 [B, L]; mel_lens int64 [B]; src_mask uint8 [B, L]
 *        (1 = pad). */
int masg_fs2_phase_a(masg_ctx* ctx, void* stream, const int64_t* texts, const int64_t* accents,
                     int accents_is_1d, const int64_t* src_lens, int B, int L, const float* spk_emb,
                     float p_ctl, float e_ctl, float d_ctl, const float* d_target,
                     const float* p_target, const float* e_target,
                     float* pitch, float* energy, float* log_d, float* d_rounded, int64_t* mel_lens,
                     uint8_t* src_mask, int64_t* h_tmax);

/* Second half (model/fastspeech2.py:232-244): LengthRegulator (model/modules.py:167-194) → Decoder →
 * mel_linear → PostNet + residual.  Uses the state left by the preceding phase A on the same ctx.
 *   out: mel, postnet_mel fp32 [B, T, 80]; mel_mask uint8 [B, T] (1 = pad); T = *h_tmax of phase A,
 *        or the caller's max_mel_len. */
int masg_fs2_phase_b(masg_ctx* ctx, void* stream, int T, float* mel, float* postnet_mel, uint8_t* mel_mask);

/* LengthRegulator.forward alone (model/modules.py:167-194 + utils/tools.py:363-381) and the duration
 * rule of VarianceAdaptor.forward (modules.py:132-135) — the bit-exact integer kernels, exposed for
 * parity tests.  masg_duration_round: d = clamp(rint(exp(log_d) - 1) * d_ctl, 0).  masg_length_regulate:
 * x fp32 [B, L, C], d fp32 [B, L] → out fp32 [B, T, C] zero-padded, mel_lens int64 [B], mask uint8 [B,T]. */
int masg_duration_round(masg_ctx* ctx, void* stream, const float* log_d, float d_ctl, int64_t n, float* d_rounded);
int masg_mel_lengths(masg_ctx* ctx, void* stream, const float* d, int B, int L, int64_t* mel_lens);
int masg_length_regulate(masg_ctx* ctx, void* stream, const float* x, const float* d, int B, int L, int C,
                         int T, float* out, int64_t* mel_lens, uint8_t* mel_mask);
/* VarianceAdaptor.get_pitch_embedding tail (modules.py:85-88): x[row] += table[bucketize(v[row], bins)] */
int masg_bucketize_embed_add(masg_ctx* ctx, void* stream, const float* v, const float* bins, int n_bins_m1,
                             const float* table, int64_t rows, int C, float* x, int64_t* idx_out /*nullable*/);

/* MultiHeadAttention core (transformer/SubLayers.py:43-52 + transformer/Modules.py:14-25) alone, for parity tests:
 * qkv fp32 [B, S, 768] = q | k | v (2 heads x 128 each, as the concatenated projection writes them), lens int64 [B]
 * (keys >= lens[b] are masked with -inf before the softmax; query rows >= lens[b] are written as zeros — the
 * reference computes and then zeroes them, Layers.py:25) -> out fp32 [B, S, 256] (heads concatenated).
 * MASG_PREC_BF16: tcgen05 kernel (bf16 hi+lo operands, scores and output in TMEM); MASG_PREC_FP32: fp32 SIMT kernel. */
int masg_attention(masg_ctx* ctx, void* stream, const float* qkv, const int64_t* lens, int B, int S, float* out);

/* The same attention core as the decoder FFT blocks run it in MASG_PREC_BF16 (fp16 operands, planar hand-overs):
 * qkv16 fp16 planar [B, 768/8, S, 8] (channel c of row t at ((c/8)*S + t)*8 + c%8; channels = q | k | v, 2 heads x 128),
 * out16 fp16 planar [B, 256/8, S, 8].  Masking as masg_attention.  MASG_PREC_BF16 contexts only. */
int masg_attention_planar(masg_ctx* ctx, void* stream, const void* qkv16, const int64_t* lens, int B, int S, void* out16);

/* hifigan.Generator.forward (hifigan/models.py:149-165): mel fp32 [B, 80, T] → wav fp32 [B, 1, 256 T].
 * wav_i16 (nullable): int16 [B, 256 T] = C-cast of wav*32768 as utils/model.py:82-85 (wrap, not saturate).
 * mel_is_channels_last != 0: mel is [B, T, 80] (postnet output layout; skips the transpose that
 * utils/tools.py:266 performs). */
int masg_hifigan_forward(masg_ctx* ctx, void* stream, const float* mel, int mel_is_channels_last, int B, int T,
                         float* wav, int16_t* wav_i16);

/* hifigan.Generator.forward + vocoder_infer's int16 cast AND its per-utterance trim (utils/model.py:82-90), fused into the
 * generator's last kernel: row b keeps its first mel_lens[b] * hop samples and the rows are written back to back.
 * mel_lens: device int64 [B]; offsets: device int64 [B + 1] out (offsets[b] = first sample of row b, offsets[B] = total);
 * wav_packed: device int16 with room for hop * sum(mel_lens) samples.  One D2H copy of offsets[B] samples then moves
 * exactly the valid audio. */
int masg_hifigan_forward_packed(masg_ctx* ctx, void* stream, const float* mel, int mel_is_channels_last, int B, int T,
                                const int64_t* mel_lens, int hop, int16_t* wav_packed, int64_t* offsets);

/* SpeakerMetaEncoder.forward (model/fastspeech2.py:332-341): meta fp32 [N,4] → pi [N,3], mu/sigma [N,3,256]. */
int masg_gmm_params(masg_ctx* ctx, void* stream, const float* meta, int N, float* pi, float* mu, float* sigma);
/* MixtureSameFamily.sample() (model/fastspeech2.py:176-180, distributions.py .sample()): per row n draw
 * k ~ Categorical(pi[n or 0]) and x = mu[.,k] + scale[.,k]·eps.  Philox counter RNG (seed, offset).
 * pi [Np,K], mu/scale [Np,K,D] with Np = N or 1 (broadcast).  k_inject (int32 [N]) / eps_inject (fp32
 * [N,D]) nullable: inject the draws for exact tests.  k_out nullable. */
int masg_gmm_sample(masg_ctx* ctx, void* stream, const float* pi, const float* mu, const float* scale,
                    int N, int Np, int K, int D, uint64_t seed, uint64_t offset,
                    const int32_t* k_inject, const float* eps_inject, float* out, int32_t* k_out);
/* InterpolateGMM.__init__/interpolate_rate (model/distributions.py:12-62): device closed forms for the
 * K×K cost (as written, Appendix B-9), component means and McCann scales; the K×K exact EMD
 * (ot.emd, :22) runs on the host inside this call (one small D2H of the cost matrix).
 * a/b: pi [K], mu/sigma [K,D].  out: weights [K*K] (gamma.flatten()), comp_mu/comp_scale [K*K, D]
 * in the reference's (j,i) order; h_cost (nullable host double [K*K]). */
int masg_gmm_interpolate(masg_ctx* ctx, void* stream, const float* pi_a, const float* mu_a, const float* sig_a,
                         const float* pi_b, const float* mu_b, const float* sig_b, int K, int D, float t,
                         int recompute_plan, float* weights, float* comp_mu, float* comp_scale, double* h_cost);
/* BarycenterGMM maths (model/distributions.py:136-192): A attribute GMMs (pi [A,K], mu/sigma [A,K,D]),
 * rate [A] → all K^A barycentres (mean [K^A,D], std [K^A,D]) and per-barycentre weights [K^A]
 * (zero for unused ones); the shim keeps the reference's first-use component order. */
int masg_gmm_barycenter(masg_ctx* ctx, void* stream, const float* pi, const float* mu, const float* sigma,
                        const float* h_rate, int A, int K, int D, float* bary_mean, float* bary_std,
                        float* bary_weight, int32_t* assign /* [A*K] nearest barycentre index */);

/* masg_gmm_barycenter for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, same results
 * (bit for bit) as masg_gmm_barycenter, but the call allocates nothing — no cudaMalloc / cudaFree, no host heap — and never
 * synchronises the stream: the rate vector and the per-barycentre weight accumulators live in one fixed file-scope device array
 * (16.25 KB, held by every GPU that loads the library until exit) and the weights are summed on the device.
 * Limit: K^A <= MASG_NOALLOC_MAX_BARYCENTRES (and A <= 8); beyond it the call fails with a message naming the limit.
 * All contexts and streams of one device share that array: calls that overlap on the GPU race on it (serialise them on one
 * stream, or with events).  h_rate is read on the host before the call returns. */
#define MASG_NOALLOC_MAX_BARYCENTRES 4096
int masg_gmm_barycenter_noalloc(masg_ctx* ctx, void* stream, const float* pi, const float* mu, const float* sigma,
                                const float* h_rate, int A, int K, int D, float* bary_mean, float* bary_std,
                                float* bary_weight, int32_t* assign /* [A*K] nearest barycentre index */);

/* End-to-end with HOST buffers (examples_gen.py:114-140 synthesize() + utils/model.py:74-92
 * vocoder_infer): H2D of ids/lens/speaker embeddings, FastSpeech2, HiFi-GAN, int16 conversion and D2H
 * of the waveforms, all inside the call.  h_wav_i16 must hold B * 256 * t_cap samples; rows are
 * written with stride 256 * t_cap (t_cap >= resulting T_max, else error).  h_mel_lens int64 [B]. */
int masg_synthesize_host(masg_ctx* ctx, void* stream, const int64_t* h_texts, const int64_t* h_accents,
                         const int64_t* h_src_lens, int B, int L, const float* h_spk_emb,
                         float p_ctl, float e_ctl, float d_ctl, int t_cap,
                         int16_t* h_wav_i16, int64_t* h_mel_lens, int64_t* h_tmax);

/* The same with the trim of utils/model.py:88-90 done on the device: h_wav_packed receives the utterances back to back (row b =
 * h_mel_lens[b] * 256 samples starting at h_offsets[b]; h_offsets has B + 1 entries, the last is the total) in ONE D2H copy of
 * exactly the valid samples; cap_samples = capacity of h_wav_packed in samples (error if the batch produces more). */
int masg_synthesize_host_packed(masg_ctx* ctx, void* stream, const int64_t* h_texts, const int64_t* h_accents,
                                const int64_t* h_src_lens, int B, int L, const float* h_spk_emb,
                                float p_ctl, float e_ctl, float d_ctl, int64_t cap_samples,
                                int16_t* h_wav_packed, int64_t* h_offsets, int64_t* h_mel_lens, int64_t* h_tmax);

/* Deferred-copy form for a stream of batches (a generation job, a service): the call returns once the batch is enqueued — after
 * phase A's one synchronisation — with the D2H copy of the audio queued on a copy stream of the context's own; h_offsets and
 * h_mel_lens are valid on return, h_wav_packed after masg_synthesize_host_wait(ctx, *ticket).  Two output buffers alternate, so
 * the copy of batch n overlaps phase A of batch n + 1; at most two tickets are outstanding (the third call waits, on the device,
 * for the first copy).  The caller alternates its own pinned destination buffers. */
int masg_synthesize_host_packed_async(masg_ctx* ctx, void* stream, const int64_t* h_texts, const int64_t* h_accents,
                                      const int64_t* h_src_lens, int B, int L, const float* h_spk_emb,
                                      float p_ctl, float e_ctl, float d_ctl, int64_t cap_samples,
                                      int16_t* h_wav_packed, int64_t* h_offsets, int64_t* h_mel_lens, int64_t* h_tmax,
                                      int* ticket);
int masg_synthesize_host_wait(masg_ctx* ctx, int ticket);

/* One Conv1d (torch.nn.Conv1d as used throughout transformer/SubLayers.py:72-83, model/modules.py:253-296,
 * hifigan/models.py:24-91) or, with stride_transposed = u > 0, one ConvTranspose1d(k = 2u, stride u,
 * padding u/2) (hifigan/models.py:127-133) through the context's dense-contraction kernel — the tcgen05
 * implicit GEMM in MASG_PREC_BF16, the fp32 CUDA-core kernel in MASG_PREC_FP32.  Diagnostic / micro-benchmark
 * hook: x fp32 [B, T, Cin] channels-last, w fp32 in the torch layout ([Cout, Cin, k], or [Cin, Cout, k] when
 * transposed), padding = (k*dilation - dilation)/2, optional leaky-relu on the input and activation
 * (0 none, 1 relu, 2 tanh) on the output; y fp32 [B, T (or u*T), Cout].  flags (bf16 mode): 1 = the hi/lo
 * split-bf16 variant used for the FastSpeech2 layers (three MMAs per K step, ~fp32 accuracy); 2 = keep the
 * activations in bf16 in HBM as the HiFi-GAN path does (x is converted before, y after the timed launches);
 * 4 = add x as residual (Cin == Cout), as ResBlock.forward does (hifigan/models.py:102); 8 = single-pass fp16 operands (the
 * FastSpeech2 decoder side); 16 = with 2 | 4: the residual is read from its own copy of x (DRAM reads, as in the generator,
 * where the residual is not the conv's input); 32 / 64 = with 2: run the CTA-pair form of the kernel (clusters of two CTAs, tcgen05
 * cta_group::2; N = 128 column tiles only) whatever the size / never (default: for launches of at least two tiles per SM).  The launch is repeated `reps`
 * times between CUDA events on `stream`; *h_ms (nullable) receives the mean milliseconds per launch. */
int masg_conv1d(masg_ctx* ctx, void* stream, const float* x, int B, int T, int Cin, const float* w, const float* bias,
                int Cout, int k, int dilation, int stride_transposed, float in_lrelu_slope, int out_act, int flags,
                int reps, float* y, float* h_ms);

/* One (convs1[m], convs2[m]) step of ResBlock1.forward (hifigan/models.py:96-103) through the fused bf16 kernel:
 *     y = (c2(leaky_relu(c1(leaky_relu(x, 0.1)), 0.1)) + x [+ acc]) * out_scale,  then leaky_relu(post_slope) unless 1.
 * Diagnostic / micro-benchmark hook for the vocoder's fused stages (MASG_PREC_BF16 only; C = 32 with k = 3, 7, 11, C = 64 with
 * k = 3, 7, 11 or C = 128 with k = 3; anything else is an error).  The call takes the kernel the generator takes: the one-SM kernel
 * (respair_tc.cu) or, for C = 64 k = 7 / 11 and C = 128 k = 3, the CTA-pair kernel (respair2_tc.cu: clusters of two CTAs, tcgen05
 * cta_group::2, the weight columns split over the pair).  x, acc (nullable), y fp32 [B, T, C] channels-last (converted to the
 * planar bf16 layout before and after the timed launches); w1, w2 fp32 [C, C, k] (torch layout, weight-norm already
 * folded); c1 has `dilation`, c2 dilation 1, both "same" padding.  *h_ms as in masg_conv1d. */
int masg_respair(masg_ctx* ctx, void* stream, const float* x, const float* acc, int B, int T, int C, const float* w1,
                 const float* b1, const float* w2, const float* b2, int k, int dilation, float out_scale, float post_slope,
                 int reps, float* y, float* h_ms);

/* Roofline accounting for bench.py (no reference counterpart): while enabled, every launch of the dense-
 * contraction kernel is bracketed by CUDA events on its stream and its algorithmic FLOPs (2 x rows x K x N) are
 * recorded.  masg_profile_read synchronises the stream, returns the summed device milliseconds, FLOPs and the
 * launch count since the last read, and clears the records. */
int masg_profile(masg_ctx* ctx, int enable);
int masg_profile_read(masg_ctx* ctx, void* stream, double* h_ms_total, double* h_flops_total, int64_t* h_launches);

/* Counters for the benchmark harness: number of kernels this context launched since the last reset. */
int64_t masg_launch_count(masg_ctx* ctx, int reset);
/* Intermediate taps for parity tests (off by default: they cost two extra activation copies per forward).  After
 * masg_debug_enable_taps(ctx, 1), masg_debug_tap copies a named internal fp32 buffer of the LAST FastSpeech2 forward
 * ("encoder_out" [B,L,256], "va_out" [B,L,256], "lr_out" [B,T,256], "decoder_out" [B,T,256]) into dst (device, fp32); the
 * phase-B taps are valid until the next call that uses the phase-B workspace (vocoder, GMM interpolation / barycentre). */
int masg_debug_enable_taps(masg_ctx* ctx, int enable);
int masg_debug_tap(masg_ctx* ctx, void* stream, const char* name, float* dst, int64_t n_elems);

/* ---- text front-end (host only, no device work): the Japanese branch of the reference's tokenisation, batched ----
 * prepare_tg_accent_jsut.py:18-70 pp_symbols (HTS full-context label -> phoneme + accent mark), convert_label.py:52-59
 * openjtalk2julius, and the id lookups of synthesize.py:217-218 ({s: i for i, s in enumerate(symbols)} — the LAST occurrence
 * of a duplicated symbol wins — and {'0':0, '[':1, ']':2, '#':3}).
 * masg_frontend_create: `symbols` = the model's symbol inventory (text/symbols.py:25-36), index = id.
 * masg_frontend_tokenize: `blob` holds one label per '\n'-separated line for all utterances back to back,
 * labels_per_utt[u] lines each; writes texts / accents int64 [n_utts, L_cap] (zero padded, caller-owned HOST memory,
 * typically pinned) and src_lens int64 [n_utts].  n_threads host threads split the utterances.
 * Errors (non-zero, message via masg_frontend_last_error): 11 = 'sil' inside an utterance (the reference asserts),
 * 12 = symbol not in the inventory (KeyError in the reference), 13 = malformed label / no trailing 'sil' (IndexError),
 * 14 = utterance longer than L_cap. */
typedef struct masg_frontend masg_frontend;
int  masg_frontend_create(const char* const* symbols, int n_symbols, masg_frontend** out);
void masg_frontend_destroy(masg_frontend* fe);
const char* masg_frontend_last_error(masg_frontend* fe);      /* fe may be NULL */
int  masg_frontend_tokenize(masg_frontend* fe, const char* blob, int64_t blob_bytes, const int64_t* labels_per_utt, int n_utts,
                            int L_cap, int drop_unvoiced_vowels, int n_threads, int64_t* texts, int64_t* accents,
                            int64_t* src_lens);

#ifdef __cplusplus
}
#endif
#endif /* MASG_H_ */
