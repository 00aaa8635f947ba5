// FastSpeech2 inference orchestration (model/fastspeech2.py:187-303): weight repacking at commit and
// the two-phase forward.  Host code only; every tensor op is a kernel from kernels_mem.cu and, for the
// dense contractions, conv_simt.cu (MASG_PREC_FP32) or the tcgen05 kernel of conv_tc.cu (MASG_PREC_BF16:
// bf16 operands, fp32 accumulation; activations stay fp32 in HBM and are converted while staged).
#include "common.cuh"
#include "tc.cuh"
#include <cmath>

void* Arena::alloc(size_t bytes) {
    bytes = (bytes + 255) & ~size_t(255);
    if (off + bytes > cap) {
        // grow: new block sized for everything allocated so far plus this request; old block is retired
        // (kept alive until release(): kernels already enqueued on the stream may still use it).
        size_t ncap = cap ? cap : (size_t)64 << 20;
        while (ncap < off + bytes) ncap *= 2;
        char* nb = nullptr;
        CUDA_OK(cudaMalloc(&nb, ncap));
        // Pointers handed out earlier keep pointing into the retired block, which stays allocated (and keeps
        // its contents) until release(); only new requests come from the new block.
        if (base) retired.push_back(base);
        base = nb; cap = ncap;
    }
    void* p = base + off;
    off += bytes;
    return p;
}
void Arena::release() {
    for (char* r : retired) cudaFree(r);
    retired.clear();
    if (base) cudaFree(base);
    base = nullptr; cap = off = 0;
}

static const TensorRef& T_(std::map<std::string, TensorRef>& m, const std::string& key, std::initializer_list<int64_t> shape) {
    auto it = m.find(key);
    MASG_CHECK(it != m.end(), "missing state_dict tensor '%s'", key.c_str());
    const TensorRef& t = it->second;
    std::vector<int64_t> want(shape);
    bool ok = want.size() == t.shape.size();
    for (size_t i = 0; ok && i < want.size(); ++i) ok = want[i] < 0 || want[i] == t.shape[i];
    MASG_CHECK(ok, "tensor '%s' has an unsupported shape (unsupported configuration; no fallback)", key.c_str());
    return t;
}

template <typename T> static T* dev_alloc(masg_ctx* c, size_t n) {
    T* p = nullptr;
    CUDA_OK(cudaMalloc(&p, n * sizeof(T)));
    c->owned.push_back(p);
    return p;
}

// Build a ConvDesc from a Conv1d/Linear weight [Cout][Cin][k] (+ optional per-Cout scale and replacement bias)
ConvDesc make_conv(masg_ctx* c, cudaStream_t s, const float* w, const float* bias, int Cout, int Cin, int k, int pad, int dil,
                   const float* scale, const float* bias_override, int mode) {
    ConvDesc d;
    d.Cin = Cin; d.N = Cout; d.taps = k; d.off0 = -pad; d.tstep = dil; d.bias_mod = Cout; d.orow = Cout; d.ooff = 0;
    float* w32 = dev_alloc<float>(c, (size_t)k * Cin * Cout);
    launch_repack_conv_w(c, s, w, Cout, Cin, k, scale, w32, Cout, 0);
    d.w32 = w32;
    d.bias = bias_override ? bias_override : bias;
    if (c->precision == MASG_PREC_BF16) tc_prepare_weights(c, s, d, mode);
    return d;
}

// MASG_LN_FUSE (default 1): fc / w_2 / the predictor convs are tiled as one 256-column tile per row block and their epilogue is
// the LayerNorm that follows them (conv_tc.cu epi_layernorm); 0 = separate k_layernorm256 launches (the round-1 graph)
static bool ln_fuse_on() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("MASG_LN_FUSE"); on = e ? atoi(e) : 1; }
    return on != 0;
}

static int ln_fuse_min_rows() {                             // MASG_LN_FUSE_ROWS: batches with fewer rows use the 128-column tiling
    static int v = -1;
    if (v < 0) { const char* e = getenv("MASG_LN_FUSE_ROWS"); v = e ? atoi(e) : 4096; }
    return v;
}

void fs2_commit(masg_ctx* c, cudaStream_t s) {
    auto& m = c->fs2_t;
    const int wide = ln_fuse_on() ? TC_MODE_WIDE : 0;
    // Operand mode of the dense layers (MASG_PREC_BF16): the encoder and the variance predictors decide integers (durations,
    // pitch / energy buckets) and run split-bf16 (hi + lo, 3 MMAs per K step, ~fp32).  Everything after the length regulator
    // only has to keep the mel within 1e-2: single-pass fp16 operands (11-bit mantissa, fp32 accumulation; measured 4e-3 where
    // plain bf16 gives 3e-2).  MASG_DEC_MODE = 1 (split) / 0 (plain bf16) / 2 (fp16, default) for experiments.
    const char* ds_env = getenv("MASG_DEC_MODE");
    const int dec_mode = ds_env ? atoi(ds_env) : (c->x3 ? 1 : 2);
    c->fs2_ready = false;
    // the repacked buffers of the previous commit are dead once the stream has drained (re-commit after load_state_dict / .to())
    CUDA_OK(cudaStreamSynchronize(s));
    for (void* p : c->owned_fs2) cudaFree(p);
    c->owned_fs2.clear();
    const size_t owned_mark = c->owned.size();
    c->enc.clear(); c->dec.clear();
    const TensorRef& we = T_(m, "encoder.src_word_emb.weight", {-1, 256});
    const TensorRef& ae = T_(m, "encoder.src_accent_emb.weight", {-1, 256});
    c->word_emb = we.ptr; c->n_vocab = (int)we.shape[0];
    c->accent_emb = ae.ptr; c->n_accent = (int)ae.shape[0];
    const TensorRef& pe = T_(m, "encoder.position_enc", {1, -1, 256});
    c->enc_pos = pe.ptr; c->max_seq_len = (int)pe.shape[1] - 1;
    c->dec_pos = T_(m, "decoder.position_enc", {1, pe.shape[1], 256}).ptr;
    for (int part = 0; part < 2; ++part) {
        const char* pn = part ? "decoder" : "encoder";
        auto& vec = part ? c->dec : c->enc;
        for (int n = 0;; ++n) {
            std::string p = std::string(pn) + ".layer_stack." + std::to_string(n) + ".";
            if (!m.count(p + "slf_attn.fc.weight")) break;
            FFTLayer L;
            // QKV: three Linear(256,256) concatenated along N (transformer/SubLayers.py:39-41)
            {
                ConvDesc d; d.Cin = 256; d.N = 768; d.taps = 1; d.off0 = 0; d.tstep = 1; d.bias_mod = 768; d.orow = 768;
                float* w32 = dev_alloc<float>(c, 256 * 768); float* b = dev_alloc<float>(c, 768);
                const char* nm[3] = {"w_qs", "w_ks", "w_vs"};
                for (int i = 0; i < 3; ++i) {
                    launch_repack_conv_w(c, s, T_(m, p + "slf_attn." + nm[i] + ".weight", {256, 256}).ptr, 256, 256, 1, nullptr, w32, 768, 256 * i);
                    CUDA_OK(cudaMemcpyAsync(b + 256 * i, T_(m, p + "slf_attn." + nm[i] + ".bias", {256}).ptr, 256 * 4, cudaMemcpyDeviceToDevice, s));
                }
                d.w32 = w32; d.bias = b;
                if (c->precision == MASG_PREC_BF16) tc_prepare_weights(c, s, d, part == 0 ? 1 : dec_mode);
                L.qkv = d;
            }
            L.fc = make_conv(c, s, T_(m, p + "slf_attn.fc.weight", {256, 256}).ptr, T_(m, p + "slf_attn.fc.bias", {256}).ptr, 256, 256, 1, 0, 1, nullptr, nullptr, (part == 0 ? 1 : dec_mode) | wide);
            L.fc_n = wide ? make_conv(c, s, T_(m, p + "slf_attn.fc.weight", {256, 256}).ptr, T_(m, p + "slf_attn.fc.bias", {256}).ptr, 256, 256, 1, 0, 1, nullptr, nullptr, part == 0 ? 1 : dec_mode) : L.fc;
            L.ln1_g = T_(m, p + "slf_attn.layer_norm.weight", {256}).ptr; L.ln1_b = T_(m, p + "slf_attn.layer_norm.bias", {256}).ptr;
            const TensorRef& w1 = T_(m, p + "pos_ffn.w_1.weight", {1024, 256, -1});
            int k1 = (int)w1.shape[2];
            L.w1 = make_conv(c, s, w1.ptr, T_(m, p + "pos_ffn.w_1.bias", {1024}).ptr, 1024, 256, k1, (k1 - 1) / 2, 1, nullptr, nullptr, part == 0 ? 1 : dec_mode);
            const TensorRef& w2 = T_(m, p + "pos_ffn.w_2.weight", {256, 1024, -1});
            int k2 = (int)w2.shape[2];
            L.w2 = make_conv(c, s, w2.ptr, T_(m, p + "pos_ffn.w_2.bias", {256}).ptr, 256, 1024, k2, (k2 - 1) / 2, 1, nullptr, nullptr, (part == 0 ? 1 : dec_mode) | (k2 == 1 ? wide : 0));
            L.w2_n = (wide && k2 == 1) ? make_conv(c, s, w2.ptr, T_(m, p + "pos_ffn.w_2.bias", {256}).ptr, 256, 1024, k2, (k2 - 1) / 2, 1, nullptr, nullptr, part == 0 ? 1 : dec_mode) : L.w2;
            L.ln2_g = T_(m, p + "pos_ffn.layer_norm.weight", {256}).ptr; L.ln2_b = T_(m, p + "pos_ffn.layer_norm.bias", {256}).ptr;
            vec.push_back(L);
        }
    }
    MASG_CHECK(!c->enc.empty() && !c->dec.empty(), "no encoder/decoder layers registered");
    const char* pnames[3] = {"duration", "pitch", "energy"};
    std::vector<float> lin_b(3);
    for (int i = 0; i < 3; ++i) {
        std::string p = std::string("variance_adaptor.") + pnames[i] + "_predictor.";
        Predictor& P = c->pred[i];
        const TensorRef& w1 = T_(m, p + "conv_layer.conv1d_1.conv.weight", {256, 256, 3});
        P.c1 = make_conv(c, s, w1.ptr, T_(m, p + "conv_layer.conv1d_1.conv.bias", {256}).ptr, 256, 256, 3, 1, 1, nullptr, nullptr, 1 | wide);
        P.c1_n = wide ? make_conv(c, s, w1.ptr, T_(m, p + "conv_layer.conv1d_1.conv.bias", {256}).ptr, 256, 256, 3, 1, 1, nullptr, nullptr, 1) : P.c1;
        P.c2 = make_conv(c, s, T_(m, p + "conv_layer.conv1d_2.conv.weight", {256, 256, 3}).ptr,
                         T_(m, p + "conv_layer.conv1d_2.conv.bias", {256}).ptr, 256, 256, 3, 1, 1, nullptr, nullptr, 1 | wide);   // padding=1 (modules.py:229-231)
        P.c2_n = wide ? make_conv(c, s, T_(m, p + "conv_layer.conv1d_2.conv.weight", {256, 256, 3}).ptr,
                                  T_(m, p + "conv_layer.conv1d_2.conv.bias", {256}).ptr, 256, 256, 3, 1, 1, nullptr, nullptr, 1) : P.c2;
        P.ln1_g = T_(m, p + "conv_layer.layer_norm_1.weight", {256}).ptr; P.ln1_b = T_(m, p + "conv_layer.layer_norm_1.bias", {256}).ptr;
        P.ln2_g = T_(m, p + "conv_layer.layer_norm_2.weight", {256}).ptr; P.ln2_b = T_(m, p + "conv_layer.layer_norm_2.bias", {256}).ptr;
        P.lin_w = T_(m, p + "linear_layer.weight", {1, 256}).ptr;
        CUDA_OK(cudaMemcpyAsync(&lin_b[i], T_(m, p + "linear_layer.bias", {1}).ptr, 4, cudaMemcpyDeviceToHost, s));
    }
    const TensorRef& pb = T_(m, "variance_adaptor.pitch_bins", {255});
    c->pitch_bins = pb.ptr;
    c->energy_bins = T_(m, "variance_adaptor.energy_bins", {255}).ptr;
    c->pitch_emb = T_(m, "variance_adaptor.pitch_embedding.weight", {256, 256}).ptr;
    c->energy_emb = T_(m, "variance_adaptor.energy_embedding.weight", {256, 256}).ptr;
    c->mel_linear = make_conv(c, s, T_(m, "mel_linear.weight", {80, 256}).ptr, T_(m, "mel_linear.bias", {80}).ptr, 80, 256, 1, 0, 1, nullptr, nullptr, dec_mode);
    int chans[6] = {80, 512, 512, 512, 512, 80};
    for (int i = 0; i < 5; ++i) {
        std::string p = "postnet.convolutions." + std::to_string(i) + ".";
        int ci = chans[i], co = chans[i + 1];
        float* scale = dev_alloc<float>(c, co); float* fb = dev_alloc<float>(c, co);
        launch_bn_fold(c, s, T_(m, p + "1.weight", {co}).ptr, T_(m, p + "1.bias", {co}).ptr, T_(m, p + "1.running_mean", {co}).ptr,
                       T_(m, p + "1.running_var", {co}).ptr, T_(m, p + "0.conv.bias", {co}).ptr, co, scale, fb);
        c->postnet[i] = make_conv(c, s, T_(m, p + "0.conv.weight", {co, ci, 5}).ptr, nullptr, co, ci, 5, 2, 1, scale, fb, dec_mode);
    }
    c->gmm_pi_w = T_(m, "speaker_enc.pi_linear.0.weight", {3, 4}).ptr; c->gmm_pi_b = T_(m, "speaker_enc.pi_linear.0.bias", {3}).ptr;
    c->gmm_sig_w = T_(m, "speaker_enc.sigma_linear.0.weight", {768, 4}).ptr; c->gmm_sig_b = T_(m, "speaker_enc.sigma_linear.0.bias", {768}).ptr;
    c->gmm_mu_w = T_(m, "speaker_enc.mu_linear.weight", {768, 4}).ptr; c->gmm_mu_b = T_(m, "speaker_enc.mu_linear.bias", {768}).ptr;
    CUDA_OK(cudaStreamSynchronize(s));
    for (int i = 0; i < 3; ++i) c->pred[i].lin_b = lin_b[i];
    c->weights_gen++;
    c->owned_fs2.assign(c->owned.begin() + owned_mark, c->owned.end());
    c->owned.resize(owned_mark);
    c->fs2_ready = true;
}

static const float* pos_table(masg_ctx* c, cudaStream_t s, bool decoder, int n) {
    if (n <= c->max_seq_len) return decoder ? c->dec_pos : c->enc_pos;
    if (n > c->pos_long_n) {                      // eval-mode rebuild for long sequences (Models.py:86-98,157-164)
        int cap = n + 1024;
        float* p = nullptr;
        CUDA_OK(cudaMalloc(&p, (size_t)cap * 256 * sizeof(float)));
        if (c->pos_long) { CUDA_OK(cudaStreamSynchronize(s)); cudaFree(c->pos_long); }     // the shorter table is dead after the sync
        launch_pos_table(c, s, cap, p);
        c->pos_long = p; c->pos_long_n = cap;
    }
    return c->pos_long;
}

// LayerNorm fused into the producing conv's epilogue (conv_tc.cu): gamma / beta, optional pad-row zeroing, optional planar copy
struct LnSpec { const float* g = nullptr; const float* b = nullptr; const int64_t* lens = nullptr; bf16* y16p = nullptr; };
static bool ln_fusable(const masg_ctx* c, const ConvDesc& d) { return c->precision == MASG_PREC_BF16 && d.NT == 256; }
static void ln_apply(ConvIO& io, const LnSpec* ln, int T) {
    if (!ln) return;
    io.ln_g = ln->g; io.ln_b = ln->b; io.ln_lens = ln->lens; io.Y16p = ln->y16p; io.ln_T = T;
}

// ---- fp32 building blocks ---------------------------------------------------------------------
static void conv32(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const float* X, int x_ld, int B, int T, float* Y,
                   int out_act = ACT_NONE, const float* R1 = nullptr, const LnSpec* ln = nullptr) {
    ConvIO io;
    ln_apply(io, ln, T);
    io.X = X; io.x_ld = x_ld; io.x_bs = (long)T * x_ld; io.Tin = T;
    io.Y = Y; io.y_bs = (long)T * d.N; io.y_limit = (long)T * d.N; io.B = B; io.Q = T; io.out_act = out_act; io.R1 = R1;
    if (d.taps == 1 && c->precision == MASG_PREC_BF16) {
        // a k = 1 layer is a plain GEMM over the B*T rows: no utterance boundaries, so no partly filled last tile per
        // utterance (T = 600 fills 2.3 tiles of 256 rows)
        io.B = 1; io.Q = B * T; io.Tin = B * T; io.x_bs = 0; io.y_bs = 0; io.y_limit = (long)B * T * d.N;
    }
    launch_conv(c, s, d, io);
}

// fp16 layer -> fp16 layer hand-over (decoder side, MASG_PREC_BF16): the producer's epilogue writes the activation as planar
// fp16 [B][C/8][T][8] — the tcgen05 operand layout — and the consumer moves it with bulk async copies, no conversion pass.
static void conv_to_planar16(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const void* X, bool x_planar, int x_ld, int B, int T,
                             bf16* Y16, int out_act) {
    ConvIO io;
    io.X = X; io.x_ld = x_ld; io.x_bs = (long)T * x_ld; io.Tin = T; io.x_planar = x_planar; io.x_bf16 = x_planar;
    io.Y16 = Y16; io.y_planar = 1; io.y_f16 = 1; io.y_T = T; io.y_bs = (long)T * d.N; io.y_limit = io.y_bs;
    io.B = B; io.Q = T; io.out_act = out_act;
    launch_conv(c, s, d, io);
}
static void conv_from_planar16(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const bf16* X16, int B, int T, float* Y,
                               int out_act = ACT_NONE, const float* R1 = nullptr, const LnSpec* ln = nullptr) {
    ConvIO io;
    ln_apply(io, ln, T);
    io.X = X16; io.x_ld = d.Cin; io.x_bs = (long)T * d.Cin; io.Tin = T; io.x_planar = 1; io.x_bf16 = 1;
    io.Y = Y; io.y_bs = (long)T * d.N; io.y_limit = (long)T * d.N; io.B = B; io.Q = T; io.out_act = out_act; io.R1 = R1;
    launch_conv(c, s, d, io);
}
static bool ln_planar(int bit) {                 // MASG_LN_PLANAR: 1 = w_1 input, 2 = QKV input as planar fp16 from the LayerNorm
    static int on = -1;
    if (on < 0) { const char* e = getenv("MASG_LN_PLANAR"); on = e ? atoi(e) : 3; }
    return (on & bit) != 0;
}
static bool f16_chain(const masg_ctx* c, const ConvDesc& a, const ConvDesc& b) {
    static int on = -1;
    if (on < 0) { const char* e = getenv("MASG_F16_PLANAR"); on = e ? atoi(e) : 1; }
    return on && c->precision == MASG_PREC_BF16 && a.f16 && b.f16 && (a.N & (a.N - 1)) == 0 && a.N >= 64;
}

static bool attn_planar(const masg_ctx* c, const FFTLayer& L) {      // MASG_ATTN_PLANAR=0: fp32 row-major q|k|v and k_attention_tc
    static int on = -1;
    if (on < 0) { const char* e = getenv("MASG_ATTN_PLANAR"); on = e ? atoi(e) : 1; }
    return on && c->precision == MASG_PREC_BF16 && L.qkv.f16 && L.fc.f16;
}

// xp_in: planar fp16 copy of x from the previous block's LayerNorm (nullable); xp_out: where this block's last LayerNorm
// leaves the same for the next block (nullable).  Decoder side in MASG_PREC_BF16 only (fp16 layers).
static void fft_block_f32(masg_ctx* c, cudaStream_t s, Arena& ar, const FFTLayer& L, float* x, const int64_t* lens, int B, int S,
                          const bf16* xp_in = nullptr, bf16* xp_out = nullptr) {
    // in place on x: every consumer of the old x has been enqueued before the LayerNorm overwrites it
    size_t rows = (size_t)B * S, mk = ar.mark();
    float* qkv = ar.get<float>(rows * 768);
    float* att = ar.get<float>(rows * 256);
    const bool chain = f16_chain(c, L.w1, L.w2);
    bf16* xp1 = (chain && ln_planar(1)) ? ar.get<bf16>(rows * 256) : nullptr;   // LN output as w_1's operand
    // fc (+ bias + residual x) -> LayerNorm -> masked_fill (SubLayers.py:54-55, Layers.py:25): one kernel when fc is tiled wide —
    // the epilogue reads the residual row of x and writes the normalised row back in place (same thread, same elements)
    // small batches (the latency path) keep the 128-column tiling: twice the CTAs per layer, LayerNorm as its own launch
    const bool big = rows >= (size_t)ln_fuse_min_rows();
    const ConvDesc& fc = big ? L.fc : L.fc_n; const ConvDesc& w2 = big ? L.w2 : L.w2_n;
    const bool fuse1 = ln_fusable(c, fc), fuse2 = ln_fusable(c, w2);
    const LnSpec ln1{L.ln1_g, L.ln1_b, lens, xp1}, ln2{L.ln2_g, L.ln2_b, lens, xp_out};
    float* y = fuse1 ? nullptr : ar.get<float>(rows * 256);
    if (attn_planar(c, L)) {
        // fp16 layers: q | k | v and the attention output stay planar fp16 (operand layout end to end, attn_f16p.cu)
        if (xp_in) conv_to_planar16(c, s, L.qkv, xp_in, true, 256, B, S, (bf16*)qkv, ACT_NONE);
        else conv_to_planar16(c, s, L.qkv, x, false, 256, B, S, (bf16*)qkv, ACT_NONE);
        launch_attention_f16p(c, s, qkv, lens, B, S, att);
        if (fuse1) conv_from_planar16(c, s, fc, (const bf16*)att, B, S, x, ACT_NONE, x, &ln1);
        else conv_from_planar16(c, s, fc, (const bf16*)att, B, S, y, ACT_NONE, x); // fc(out) + residual
    } else {
        if (xp_in) conv_from_planar16(c, s, L.qkv, xp_in, B, S, qkv);
        else conv32(c, s, L.qkv, x, 256, B, S, qkv);
        if (c->precision == MASG_PREC_BF16) launch_attention_tc(c, s, qkv, lens, B, S, att, L.qkv.f16);
        else launch_attention_f32(c, s, qkv, lens, B, S, att);
        if (fuse1) conv32(c, s, fc, att, 256, B, S, x, ACT_NONE, x, &ln1);
        else conv32(c, s, fc, att, 256, B, S, y, ACT_NONE, x);                       // fc(out) + residual
    }
    if (!fuse1) launch_layernorm(c, s, y, nullptr, L.ln1_g, L.ln1_b, lens, B, S, x, xp1);   // LN + masked_fill (Layers.py:25)
    float* h = ar.get<float>(rows * 1024);
    float* y2 = fuse2 ? nullptr : ar.get<float>(rows * 256);
    if (chain) {                                                                // h only feeds w_2: planar fp16, never fp32
        if (xp1) conv_to_planar16(c, s, L.w1, xp1, true, 256, B, S, (bf16*)h, ACT_RELU);
        else conv_to_planar16(c, s, L.w1, x, false, 256, B, S, (bf16*)h, ACT_RELU);
        if (fuse2) conv_from_planar16(c, s, w2, (const bf16*)h, B, S, x, ACT_NONE, x, &ln2);
        else conv_from_planar16(c, s, w2, (const bf16*)h, B, S, y2, ACT_NONE, x);
    } else {
        conv32(c, s, L.w1, x, 256, B, S, h, ACT_RELU);
        if (fuse2) conv32(c, s, w2, h, 1024, B, S, x, ACT_NONE, x, &ln2);
        else conv32(c, s, w2, h, 1024, B, S, y2, ACT_NONE, x);
    }
    if (!fuse2) launch_layernorm(c, s, y2, nullptr, L.ln2_g, L.ln2_b, lens, B, S, x, xp_out);   // LN + masked_fill (Layers.py:28)
    ar.restore(mk);
}

static void predictor_f32(masg_ctx* c, cudaStream_t s, Arena& ar, const Predictor& P, const float* x, const int64_t* lens,
                          int B, int L, float scale, float* out) {
    size_t rows = (size_t)B * L, mk = ar.mark();
    float* a = ar.get<float>(rows * 256); float* b = ar.get<float>(rows * 256);
    // Conv k3 -> ReLU -> LayerNorm, twice (modules.py:209-240); the LayerNorm rides in the conv's epilogue when tiled wide
    const LnSpec l1{P.ln1_g, P.ln1_b, nullptr, nullptr}, l2{P.ln2_g, P.ln2_b, nullptr, nullptr};
    const bool big = rows >= (size_t)ln_fuse_min_rows();
    const ConvDesc& c1 = big ? P.c1 : P.c1_n; const ConvDesc& c2 = big ? P.c2 : P.c2_n;
    if (ln_fusable(c, c1)) conv32(c, s, c1, x, 256, B, L, b, ACT_RELU, nullptr, &l1);
    else { conv32(c, s, c1, x, 256, B, L, a, ACT_RELU); launch_layernorm(c, s, a, nullptr, P.ln1_g, P.ln1_b, nullptr, B, L, b); }
    if (ln_fusable(c, c2)) conv32(c, s, c2, b, 256, B, L, a, ACT_RELU, nullptr, &l2);
    else { conv32(c, s, c2, b, 256, B, L, a, ACT_RELU); launch_layernorm(c, s, a, nullptr, P.ln2_g, P.ln2_b, nullptr, B, L, b); std::swap(a, b); }
    launch_rowdot_mask(c, s, a, P.lin_w, P.lin_b, lens, B, L, scale, out);
    ar.restore(mk);
}

void fs2_phase_a(masg_ctx* c, cudaStream_t s, const int64_t* texts, const int64_t* accents, int acc1d,
                 const int64_t* src_lens, int B, int L, const float* spk, float p_ctl, float e_ctl, float d_ctl,
                 const float* d_target, const float* p_target, const float* e_target, float* pitch, float* energy, float* log_d,
                 float* d_rounded, int64_t* mel_lens, uint8_t* src_mask, int64_t* h_tmax) {
    MASG_CHECK(c->fs2_ready, "FastSpeech2 weights not committed (call masg_fs2_commit)");
    MASG_CHECK(B > 0 && L > 0, "empty batch B=%d L=%d", B, L);
    MASG_CHECK(accents != nullptr, "accents is required (transformer/Models.py:101 would raise on None)");
    (void)e_ctl;   // accepted and ignored: energy is scaled by p_control (model/modules.py:123-125)
    Arena& ar = c->arena;
    ar.reset();
    size_t rows = (size_t)B * L;
    // persistent phase-A → phase-B state first
    c->va_out = ar.get<float>(rows * 256);
    c->d_rounded_dev = ar.get<float>(rows);
    c->mel_lens_dev = ar.get<int64_t>(B + 2);                                   // [mel_lens (B) | T_max | error word]
    c->tap_enc = c->taps ? ar.get<float>(rows * 256) : nullptr;
    c->B = B; c->L = L; c->src_lens_dev = src_lens;
    if (src_mask) launch_mask(c, s, src_lens, B, L, src_mask);
    const float* pos = pos_table(c, s, false, L);
    float* x = c->va_out;
    CUDA_OK(cudaMemsetAsync(c->mel_lens_dev + B + 1, 0, sizeof(int64_t), s));
    launch_embed(c, s, texts, accents, acc1d, c->word_emb, c->accent_emb, pos, B, L, c->n_vocab, c->n_accent, c->mel_lens_dev + B + 1, x);
    for (auto& Lr : c->enc) fft_block_f32(c, s, ar, Lr, x, src_lens, B, L);
    if (c->taps) CUDA_OK(cudaMemcpyAsync(c->tap_enc, x, rows * 256 * sizeof(float), cudaMemcpyDeviceToDevice, s));   // opt-in debug copy
    launch_add_rowvec(c, s, x, spk, B, L, 256);                                 // pads too (fastspeech2.py:215-218)
    auto predict = [&](int which, float scale, float* out) {
        predictor_f32(c, s, ar, c->pred[which], x, src_lens, B, L, scale, out);
    };
    predict(0, 1.0f, log_d);
    // with a target the embedding comes from bucketize(target) and the prediction is returned unscaled (modules.py:80-100)
    predict(1, p_target ? 1.0f : p_ctl, pitch);
    launch_bucket_add(c, s, p_target ? p_target : pitch, c->pitch_bins, 255, c->pitch_emb, (int64_t)rows, 256, x, nullptr);
    predict(2, e_target ? 1.0f : p_ctl, energy);                                // p_control, sic (modules.py:123-125)
    launch_bucket_add(c, s, e_target ? e_target : energy, c->energy_bins, 255, c->energy_emb, (int64_t)rows, 256, x, nullptr);
    if (d_target) CUDA_OK(cudaMemcpyAsync(c->d_rounded_dev, d_target, rows * sizeof(float), cudaMemcpyDeviceToDevice, s));
    else launch_duration(c, s, log_d, d_ctl, (int64_t)rows, c->d_rounded_dev);
    if (d_rounded) CUDA_OK(cudaMemcpyAsync(d_rounded, c->d_rounded_dev, rows * sizeof(float), cudaMemcpyDeviceToDevice, s));
    launch_mel_lens(c, s, c->d_rounded_dev, B, L, c->mel_lens_dev, c->mel_lens_dev + B);
    if (mel_lens) CUDA_OK(cudaMemcpyAsync(mel_lens, c->mel_lens_dev, B * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
    if (h_tmax) fs2_phase_a_publish(c, s, h_tmax);
}

// the single documented synchronisation of phase A: [mel_lens (B) | T_max | error word] -> pinned host scratch
void fs2_phase_a_publish(masg_ctx* c, cudaStream_t s, int64_t* h_tmax) {
    const int B = c->B;
    {
        if (c->h_pinned_n < B + 2) {
            if (c->h_pinned) cudaFreeHost(c->h_pinned);
            c->h_pinned = nullptr; c->h_pinned_n = 0;
            CUDA_OK(cudaMallocHost(&c->h_pinned, (size_t)(B + 2 + 64) * sizeof(int64_t)));
            c->h_pinned_n = B + 2 + 64;
        }
        // [mel_lens (B) | T_max | error word] in one copy: the host-buffer entry points size their trimmed D2H from the lengths
        CUDA_OK(cudaMemcpyAsync(c->h_pinned, c->mel_lens_dev, (size_t)(B + 2) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));                                      // the single documented sync
        *h_tmax = c->h_pinned[B];
        MASG_CHECK(c->h_pinned[B + 1] == 0, "%s id outside its embedding table (nn.Embedding raises IndexError: transformer/Models.py:101-103)",
                   (c->h_pinned[B + 1] & 1) ? "phoneme" : "accent");
    }
}

void fs2_phase_b(masg_ctx* c, cudaStream_t s, int T, float* mel, float* post, uint8_t* mel_mask) {
    MASG_CHECK(c->fs2_ready && c->va_out, "phase B called without a preceding phase A");
    MASG_CHECK(T > 0, "T=%d: no frames to synthesise (all durations zero)", T);
    int B = c->B, L = c->L;
    Arena& ar = c->arena_b;
    ar.reset();
    size_t rows = (size_t)B * T;
    float* x = ar.get<float>(rows * 256);
    c->tap_lr = c->taps ? ar.get<float>(rows * 256) : nullptr;
    c->tap_dec = c->taps ? x : nullptr;
    c->tap_T = T;
    const float* pos = pos_table(c, s, true, T);
    const int64_t* lens = c->mel_lens_dev;
    // LengthRegulator with the decoder's position encoding added on the way out (valid and padded rows: Models.py:176-180 adds
    // it to the zero-padded batch); the un-encoded copy exists only for the opt-in debug tap
    if (c->taps) launch_length_regulate(c, s, c->va_out, c->d_rounded_dev, B, L, 256, T, nullptr, c->tap_lr, nullptr, nullptr);
    launch_length_regulate(c, s, c->va_out, c->d_rounded_dev, B, L, 256, T, pos, x, nullptr, mel_mask);
    {
        // between decoder blocks the LayerNorm output also travels as planar fp16 (the next block's QKV operand)
        const bool pl = c->precision == MASG_PREC_BF16 && !c->dec.empty() && c->dec[0].qkv.f16 && ln_planar(2);
        bf16* xp[2] = {pl ? ar.get<bf16>((size_t)B * T * 256) : nullptr, pl ? ar.get<bf16>((size_t)B * T * 256) : nullptr};
        for (size_t i = 0; i < c->dec.size(); ++i)
            fft_block_f32(c, s, ar, c->dec[i], x, lens, B, T, (pl && i > 0) ? xp[(i - 1) & 1] : nullptr, (pl && i + 1 < c->dec.size()) ? xp[i & 1] : nullptr);
    }
    conv32(c, s, c->mel_linear, x, 256, B, T, mel);                             // pads become bias (fastspeech2.py:242)
    float* h1 = ar.get<float>(rows * 512); float* h2 = ar.get<float>(rows * 512);
    if (f16_chain(c, c->postnet[0], c->postnet[1]) && f16_chain(c, c->postnet[3], c->postnet[4])) {
        conv_to_planar16(c, s, c->postnet[0], mel, false, 80, B, T, (bf16*)h1, ACT_TANH);
        conv_to_planar16(c, s, c->postnet[1], h1, true, 512, B, T, (bf16*)h2, ACT_TANH);
        conv_to_planar16(c, s, c->postnet[2], h2, true, 512, B, T, (bf16*)h1, ACT_TANH);
        conv_to_planar16(c, s, c->postnet[3], h1, true, 512, B, T, (bf16*)h2, ACT_TANH);
        conv_from_planar16(c, s, c->postnet[4], (const bf16*)h2, B, T, post, ACT_NONE, mel);
    } else {
        conv32(c, s, c->postnet[0], mel, 80, B, T, h1, ACT_TANH);
        conv32(c, s, c->postnet[1], h1, 512, B, T, h2, ACT_TANH);
        conv32(c, s, c->postnet[2], h2, 512, B, T, h1, ACT_TANH);
        conv32(c, s, c->postnet[3], h1, 512, B, T, h2, ACT_TANH);
        conv32(c, s, c->postnet[4], h2, 512, B, T, post, ACT_NONE, mel);        // + residual (fastspeech2.py:244)
    }
}
