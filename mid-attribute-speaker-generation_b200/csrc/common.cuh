// Shared declarations for libmasg.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include <unordered_map>

#include "../../include/masg.h"

#define MASG_D 256          // encoder/decoder hidden
#define MASG_NMEL 80
#define MASG_DFF 1024

typedef __nv_bfloat16 bf16;

struct MasgError { std::string msg; };

#define MASG_CHECK(cond, ...)                                                              \
    do { if (!(cond)) { char _b[512]; snprintf(_b, sizeof(_b), __VA_ARGS__);                \
         throw MasgError{std::string(_b) + " [" #cond "] at " __FILE__ ":" + std::to_string(__LINE__)}; } } while (0)
#define CUDA_OK(call)                                                                      \
    do { cudaError_t _e = (call); if (_e != cudaSuccess) {                                  \
         throw MasgError{std::string("CUDA error: ") + cudaGetErrorString(_e) + " in " #call " at " __FILE__ ":" + std::to_string(__LINE__)}; } } while (0)

// The dynamic shared-memory opt-in (cudaFuncSetAttribute) is a per-DEVICE attribute of a kernel: one process driving several GPUs
// has to set it on each.  `static PerDeviceOnce once; if (once.need()) cudaFuncSetAttribute(...)` in every launcher.
struct PerDeviceOnce {
    bool done[64] = {};
    bool need() {
        int d = 0;
        if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
        if (done[d]) return false;
        done[d] = true;                                   // (a race between host threads only repeats an idempotent call)
        return true;
    }
};

// ---- device arena: grown on demand, never freed mid-stream (stream-ordered reuse by one ctx) ----
struct Arena {
    char* base = nullptr; size_t cap = 0, off = 0;
    std::vector<char*> retired;
    void reset() { off = 0; }
    size_t mark() const { return off; }
    void restore(size_t m) { off = m; }
    void* alloc(size_t bytes);
    template <typename T> T* get(size_t n) { return (T*)alloc(n * sizeof(T)); }
    void release();
};

struct TensorRef { const float* ptr; std::vector<int64_t> shape; int64_t numel() const { int64_t n = 1; for (auto s : shape) n *= s; return n; } };

// ---- generic row-GEMM / conv description shared by the fp32 SIMT kernel and the tcgen05 kernel ----
// out[b, q*orow + ooff + n] = act( sum_{j<taps, c<Cin} X[b, q + off0 + j*tstep, c] * W[(j*Cin + c), n] + bias[n % bias_mod] )
//                             (+ R1[...] + R2[...]) * out_scale
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2, ACT_LRELU = 3 };

struct ConvDesc {
    int Cin = 0, N = 0, taps = 1, off0 = 0, tstep = 1;
    int bias_mod = 0;                 // 0 → N
    long orow = 0, ooff = 0;          // output row stride / offset in elements (0 → N / 0)
    int up = 0, up_pad = 0;           // phase-decomposed ConvTranspose1d: column n = r*Cout + co lands at time q*up + r - up_pad (planar output)
    const float* w32 = nullptr;       // [taps*Cin][N] fp32 (SIMT path)
    const float* bias = nullptr;      // [bias_mod]
    // tcgen05 path (conv_tc.cu): bf16 weights pre-tiled into UMMA no-swizzle K-major blobs
    // [n_tile = N/NT][ci = Cin/KW][tap][KW/8][NT][8]; one blob = one (channel slab, tap) B operand
    const bf16* w16 = nullptr; int NT = 0, KW = 0, split = 0, f16 = 0;   // f16: the blobs hold fp16, not bf16
    // second tiling of the same weights at 32 columns per tile, for launches that would otherwise run on a handful of CTAs (small
    // batches: the latency path): 4x the CTAs, each streaming a quarter of the weights
    const bf16* w16s = nullptr; int NTs = 0, KWs = 0;
    // third tiling, for the fused ResBlock pair on a CTA pair (respair2_tc.cu): the columns split in two halves of N/2, each
    // [tap][Cin/8][N/2][8] — one half per CTA of the cluster (tcgen05 cta_group::2 B operand)
    const bf16* w16p = nullptr;
    // the commit-time tiling at NT / 2 columns per blob set, for the pair mode of k_conv_tc (NT = 128 layers)
    const bf16* w16h = nullptr;
    const bf16* w16h2 = nullptr;      // the same at 128 input channels per blob (launches with 128-row tiles: one issuing thread)
};

struct ConvWeights {                  // owned storage for a ConvDesc
    float* w32 = nullptr; float* bias = nullptr; bf16* w16 = nullptr;
};

struct FFTLayer {
    ConvDesc qkv, fc, w1, w2;
    ConvDesc fc_n, w2_n;              // the same layers tiled 128 columns wide (small batches: more CTAs, separate LayerNorm)
    const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
};
struct Predictor {
    ConvDesc c1, c2, c1_n, c2_n;
    const float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *lin_w; float lin_b;
};

// One captured CUDA graph of a phase of the host-buffer entry points (api.cu): the launch-bound small-batch path replays
// ~50-120 kernel launches as one graph launch.  Valid while the library-owned buffers it was captured on have not moved.
struct GraphEntry {
    cudaGraphExec_t exec = nullptr;
    void* snap[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int64_t launches = 0; int seen = 0;
    // phase-A -> phase-B context state as the captured body left it
    float* va_out = nullptr; float* d_rounded_dev = nullptr; int64_t* mel_lens_dev = nullptr; const int64_t* src_lens_dev = nullptr;
};

struct masg_ctx {
    int device = 0, precision = 0, n_sm = 148;
    bool x3 = false;                  // MASG_PREC_BF16X3: split-bf16 (3 MMAs per K step) through the decoder / mel_linear / PostNet too
    std::map<std::vector<long>, GraphEntry> graphs;   // key: (phase, B, L, T, packed)
    cudaStream_t cap_stream = nullptr;                // graphs are recorded here (the caller's stream may be the legacy stream)
    int64_t weights_gen = 0;                          // bumped by every commit: captured graphs bake weight pointers
    std::string err;
    int64_t launches = 0;
    // per-launch device timing of the dense-contraction kernel (masg_profile): event pairs + algorithmic FLOPs
    bool profiling = false;
    struct ProfRec { cudaEvent_t e0, e1; double flops; };
    std::vector<ProfRec> prof;
    Arena arena;                      // per-forward workspace
    std::vector<void*> owned;         // repacked weights being built / temporaries of the diagnostic hooks
    std::vector<void*> owned_fs2, owned_hg;   // ... of the last commit of each model (freed by the next commit of that model)
    bool taps = false;                // keep copies of the intermediate tensors for masg_debug_tap (masg_debug_enable_taps)
    std::map<std::string, TensorRef> fs2_t, hg_t;
    bool fs2_ready = false, hg_ready = false;
    // ---- FastSpeech2 ----
    int n_enc = 0, n_dec = 0, max_seq_len = 1000;
    const float *word_emb = nullptr, *accent_emb = nullptr, *enc_pos = nullptr, *dec_pos = nullptr;
    int n_vocab = 0, n_accent = 0;
    float* pos_long = nullptr; int pos_long_n = 0;           // rebuilt table for len > max_seq_len
    std::vector<FFTLayer> enc, dec;
    Predictor pred[3];                                       // duration, pitch, energy
    const float *pitch_bins = nullptr, *energy_bins = nullptr, *pitch_emb = nullptr, *energy_emb = nullptr;
    ConvDesc mel_linear, postnet[5];
    const float *gmm_pi_w, *gmm_pi_b, *gmm_sig_w, *gmm_sig_b, *gmm_mu_w, *gmm_mu_b;
    // phase A → phase B state
    int B = 0, L = 0; float* va_out = nullptr; float* d_rounded_dev = nullptr; int64_t* mel_lens_dev = nullptr;
    const int64_t* src_lens_dev = nullptr;
    float* tap_enc = nullptr; float* tap_lr = nullptr; float* tap_dec = nullptr; int tap_T = 0;
    Arena arena_b;                    // phase-B / vocoder workspace (arena holds phase-A state)
    int64_t* h_pinned = nullptr; int h_pinned_n = 0;   // pinned host scratch: [mel_lens (B) | T_max] of the last phase A
    void* io_buf = nullptr; size_t io_cap = 0;   // device staging of the host-buffer entry points: inputs + phase-A outputs
    void* out_buf = nullptr; size_t out_cap = 0; // ... and mel / postnet mel / audio (sized after phase A)
    // deferred-copy mode of the host-buffer entry points: a second output buffer, a copy stream and per-buffer events, so that the
    // D2H of call n runs while call n + 1 computes
    void* out_alt = nullptr; size_t out_alt_cap = 0; int out_idx = 0;
    cudaStream_t copy_stream = nullptr; cudaEvent_t ev_done[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
    // ---- HiFi-GAN ----
    ConvDesc hg_pre, hg_ups[4], hg_c1[12][3], hg_c2[12][3];
    const float* hg_post_w = nullptr; float hg_post_b = 0.f;   // [7][32] fp32
    int hg_up_rate[4] = {8, 8, 2, 2}, hg_up_k[4] = {16, 16, 4, 4}, hg_ch[5] = {512, 256, 128, 64, 32};
    int hg_rk[3] = {3, 7, 11}, hg_rd[3] = {1, 3, 5};
};

// ------------------------------------------------------------------------------------------------
// kernel launchers (each increments ctx->launches)
// ------------------------------------------------------------------------------------------------
struct ConvIO {
    const void* X = nullptr; int x_bf16 = 0; long x_bs = 0; int x_ld = 0; int Tin = 0;   // input rows [B][Tin][x_ld] fp32, or
    int x_planar = 0;                                                        // bf16 planar [B][x_ld/8][Tin][8] (tcgen05 path only)
    int y_planar = 0, y_T = 0;                                               // Y16 is planar [B][Cout/8][y_T][8] (residuals too)
    int y_f16 = 0;                                                           // planar output holds fp16 (feeds an fp16 layer), not bf16
    float post_slope = 1.f;                                                  // trailing leaky-relu after residuals and scale (1 = none)
    float* Y = nullptr; bf16* Y16 = nullptr; long y_bs = 0; long y_limit = 0;  // fp32 and/or bf16 output, valid elements per batch
    int B = 1, Q = 0;                                                        // rows computed per batch
    int in_act = ACT_NONE; float in_slope = 0.f;
    int out_act = ACT_NONE; float out_slope = 0.f;
    const void* R1 = nullptr; const void* R2 = nullptr; int r_bf16 = 0; float out_scale = 1.f;   // residuals (fp32 or bf16)
    // fused LayerNorm epilogue (tcgen05 path, N == 256 == one tile row): Y = mask(LN(act(acc + bias) + R1) * ln_g + ln_b), rows with
    // t >= ln_lens[b] written as zeros (Layers.py:25,28); Y16p (nullable) = the same as planar fp16 [B][32][ln_T][8]
    const float* ln_g = nullptr; const float* ln_b = nullptr; const int64_t* ln_lens = nullptr; int ln_T = 0; bf16* Y16p = nullptr;
    int pair_mode = 0;                                                       // tcgen05 path: 1 = CTA-pair kernel whatever the size (tests), -1 = never, 0 = by size
};
void launch_conv_f32(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io);
// precision dispatch: MASG_PREC_FP32 → conv_simt.cu, MASG_PREC_BF16 → conv_tc.cu (tcgen05)
void launch_conv(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io);

void launch_embed(masg_ctx* c, cudaStream_t s, const int64_t* texts, const int64_t* accents, int acc1d,
                  const float* wemb, const float* aemb, const float* pos, int B, int L, int n_vocab, int n_accent,
                  int64_t* err /*nullable device word: bit 0 = phoneme id, bit 1 = accent id out of range*/, float* out);
void launch_add_rowvec(masg_ctx* c, cudaStream_t s, float* x, const float* v, int B, int L, int C);   // x[b,l,:] += v[b,:]
void launch_add_pos(masg_ctx* c, cudaStream_t s, float* x, const float* pos, int B, int T, int C);
// y = LN(x (+res)) * g + b ; rows with t >= lens[b] are zeroed when lens != null
void launch_layernorm(masg_ctx* c, cudaStream_t s, const float* x, const float* res, const float* g, const float* b,
                      const int64_t* lens, int B, int T, float* y, bf16* y16_planar = nullptr);   // + planar fp16 copy [B][32][T][8]
// out[b,t] = (t < lens[b]) ? dot(x[b,t,:], w) + bias : 0
void launch_rowdot_mask(masg_ctx* c, cudaStream_t s, const float* x, const float* w, float bias, const int64_t* lens,
                        int B, int T, float scale, float* out);
void launch_mask(masg_ctx* c, cudaStream_t s, const int64_t* lens, int B, int T, uint8_t* mask);
void launch_bucket_add(masg_ctx* c, cudaStream_t s, const float* v, const float* bins, int nb, const float* table,
                       int64_t rows, int C, float* x, int64_t* idx_out);
void launch_duration(masg_ctx* c, cudaStream_t s, const float* log_d, float d_ctl, int64_t n, float* d);
void launch_mel_lens(masg_ctx* c, cudaStream_t s, const float* d, int B, int L, int64_t* mel_lens, int64_t* tmax_out);
void launch_length_regulate(masg_ctx* c, cudaStream_t s, const float* x, const float* d, int B, int L, int C, int T,
                            const float* pos /*nullable, added to valid+pad rows*/, float* out, int64_t* mel_lens,
                            uint8_t* mask);
void launch_attention_f32(masg_ctx* c, cudaStream_t s, const float* qkv, const int64_t* lens, int B, int S, float* out);
// tcgen05 masked attention (attn_tc.cu), MASG_PREC_BF16
// f16: fp16 single-pass operands, two CTAs per SM (decoder side); else split-bf16 hi + lo, 3 MMAs per product
void launch_attention_tc(masg_ctx* c, cudaStream_t s, const float* qkv, const int64_t* lens, int B, int S, float* out, int f16 = 0);
// planar fp16 in ([B][96][S][8]) and out ([B][32][S][8]): attn_f16p.cu
void launch_attention_f16p(masg_ctx* c, cudaStream_t s, const void* qkv, const int64_t* lens, int B, int S, void* out);
void launch_transpose_mel(masg_ctx* c, cudaStream_t s, const float* mel_bct, int B, int C, int T, float* out_btc);
// row_len / row_off (nullable, device int64 [B]): store only the first row_len[b] samples of row b, at wav + row_off[b]
void launch_conv_post(masg_ctx* c, cudaStream_t s, const void* x, int x_bf16, const float* w, float bias, int B, long S,
                      float in_slope, float* wav, int16_t* wav16, const int64_t* row_len = nullptr, const int64_t* row_off = nullptr);
void launch_wav_offsets(masg_ctx* c, cudaStream_t s, const int64_t* mel_lens, int B, int hop, int64_t* off /*[B+1]*/, int64_t* len /*[B], nullable*/);
void launch_pos_table(masg_ctx* c, cudaStream_t s, int n, float* out);
void launch_add2(masg_ctx* c, cudaStream_t s, const float* a, const float* b, float* out, int64_t n);
void launch_gmm_params(masg_ctx* c, cudaStream_t s, const float* meta, int N, float* pi, float* mu, float* sigma);
void launch_gmm_sample(masg_ctx* c, cudaStream_t s, const float* pi, const float* mu, const float* scale, int N, int Np,
                       int K, int D, uint64_t seed, uint64_t offset, const int32_t* kin, const float* ein, float* out,
                       int32_t* kout);
void launch_gmm_interp_cost(masg_ctx* c, cudaStream_t s, const float* mu_a, const float* sig_a, const float* mu_b,
                            const float* sig_b, int K, int D, double* cost);
void launch_gmm_interp_comps(masg_ctx* c, cudaStream_t s, const float* mu_a, const float* sig_a, const float* mu_b,
                             const float* sig_b, int K, int D, float t, float* comp_mu, float* comp_scale);
void launch_gmm_bary(masg_ctx* c, cudaStream_t s, const float* mu, const float* sigma, const float* rate, int A, int K,
                     int D, float* bmean, float* bstd);
void launch_gmm_bary_assign(masg_ctx* c, cudaStream_t s, const float* mu, const float* sigma, const float* bmean,
                            const float* bstd, int A, int K, int D, int nb, int32_t* assign);

// row-major fp32 [B][T][C] <-> planar bf16 [B][C/8][T][8] (conv_tc.cu)
void launch_to_planar(masg_ctx* c, cudaStream_t s, const float* x, bf16* out, int B, int T, int C);
void launch_from_planar(masg_ctx* c, cudaStream_t s, const bf16* x, float* out, int B, int T, int C);
void launch_f32_to_bf16(masg_ctx* c, cudaStream_t s, const float* in, bf16* out, long n);
void launch_bf16_to_f32(masg_ctx* c, cudaStream_t s, const bf16* in, float* out, long n);
// repack helpers (device side)
void launch_repack_conv_w(masg_ctx* c, cudaStream_t s, const float* w /*[Cout][Cin][k]*/, int Cout, int Cin, int k,
                          const float* scale /*nullable per-Cout*/, float* out /*[k*Cin][ldn] at col0*/, int ldn, int col0);
void launch_repack_convT_w(masg_ctx* c, cudaStream_t s, const float* w /*[Cin][Cout][k]*/, int Cin, int Cout, int k, int u,
                           float* out /*[2*Cin][u*Cout]*/);
void launch_weight_norm_fold(masg_ctx* c, cudaStream_t s, const float* v, const float* g, int n0, int inner, float* w);
void launch_bn_fold(masg_ctx* c, cudaStream_t s, const float* gamma, const float* beta, const float* mean, const float* var,
                    const float* conv_bias, int C, float* scale, float* bias);

// fs2.cu / hifigan.cu / gmm.cu
void fs2_commit(masg_ctx* c, cudaStream_t s);
void hifigan_commit(masg_ctx* c, cudaStream_t s);
void fs2_phase_a(masg_ctx* c, cudaStream_t s, const int64_t* texts, const int64_t* accents, int acc1d,
                 const int64_t* src_lens, int B, int L, const float* spk, float p_ctl, float e_ctl, float d_ctl,
                 const float* d_target, const float* p_target, const float* e_target, float* pitch, float* energy, float* log_d,
                 float* d_rounded, int64_t* mel_lens, uint8_t* src_mask, int64_t* h_tmax);
void fs2_phase_a_publish(masg_ctx* c, cudaStream_t s, int64_t* h_tmax);   // D2H of [mel_lens | T_max | error word] + the one sync
void fs2_phase_b(masg_ctx* c, cudaStream_t s, int T, float* mel, float* post, uint8_t* mel_mask);
// row_len / row_off (nullable, device int64 [B]): per-utterance trim + packed placement of the output rows (launch_conv_post)
void hifigan_forward(masg_ctx* c, cudaStream_t s, const float* mel, int cl, int B, int T, float* wav, int16_t* w16,
                     const int64_t* row_len = nullptr, const int64_t* row_off = nullptr);
