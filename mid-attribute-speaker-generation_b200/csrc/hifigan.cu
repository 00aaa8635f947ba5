// HiFi-GAN V1 generator (hifigan/models.py:112-174): weight-norm folding + repacking at commit, and
// the forward graph.  Activations are channels-last [B, S, C]; ConvTranspose1d runs as a 2-tap
// phase-decomposed row GEMM (see conv_simt.cu header); leaky-relu is applied on operand load, the
// residual add, the MRF sum and the 1/3 scale are folded into conv epilogues.
#include "common.cuh"
#include "tc.cuh"

ConvDesc make_conv(masg_ctx* c, cudaStream_t s, const float* w, const float* bias, int Cout, int Cin, int k, int pad, int dil,
                   const float* scale, const float* bias_override, int mode);

template <typename T> static T* dev_alloc(masg_ctx* c, size_t n) {
    T* p = nullptr;
    CUDA_OK(cudaMalloc(&p, n * sizeof(T)));
    c->owned.push_back(p);
    return p;
}

// folded fp32 weight for `name` ([n0][...]): accepts weight_g/weight_v (checkpoint layout) or weight
static const float* folded_weight(masg_ctx* c, cudaStream_t s, const std::string& name, std::vector<int64_t>& shape) {
    auto& m = c->hg_t;
    auto iw = m.find(name + ".weight");
    if (iw != m.end()) { shape = iw->second.shape; return iw->second.ptr; }
    auto iv = m.find(name + ".weight_v"), ig = m.find(name + ".weight_g");
    MASG_CHECK(iv != m.end() && ig != m.end(), "missing HiFi-GAN tensor '%s.weight' (or weight_g/weight_v)", name.c_str());
    shape = iv->second.shape;
    MASG_CHECK(shape.size() == 3 && ig->second.numel() == shape[0], "bad weight-norm shapes for '%s'", name.c_str());
    int n0 = (int)shape[0], inner = (int)(shape[1] * shape[2]);
    float* w = dev_alloc<float>(c, (size_t)n0 * inner);
    launch_weight_norm_fold(c, s, iv->second.ptr, ig->second.ptr, n0, inner, w);
    return w;
}
static const float* bias_of(masg_ctx* c, const std::string& name, int n) {
    auto it = c->hg_t.find(name + ".bias");
    MASG_CHECK(it != c->hg_t.end() && it->second.numel() == n, "missing/bad HiFi-GAN bias '%s.bias'", name.c_str());
    return it->second.ptr;
}

void hifigan_commit(masg_ctx* c, cudaStream_t s) {
    c->hg_ready = false;
    CUDA_OK(cudaStreamSynchronize(s));                       // previous generation of repacked buffers: dead once the stream drained
    for (void* p : c->owned_hg) cudaFree(p);
    c->owned_hg.clear();
    const size_t owned_mark = c->owned.size();
    std::vector<int64_t> sh;
    const float* w = folded_weight(c, s, "conv_pre", sh);
    MASG_CHECK(sh[0] == 512 && sh[1] == 80 && sh[2] == 7, "conv_pre must be (512,80,7): unsupported generator config");
    c->hg_pre = make_conv(c, s, w, bias_of(c, "conv_pre", 512), 512, 80, 7, 3, 1, nullptr, nullptr, false);
    for (int i = 0; i < 4; ++i) {
        std::string name = "ups." + std::to_string(i);
        int cin = c->hg_ch[i], cout = c->hg_ch[i + 1], k = c->hg_up_k[i], u = c->hg_up_rate[i];
        w = folded_weight(c, s, name, sh);
        MASG_CHECK(sh[0] == cin && sh[1] == cout && sh[2] == k && k == 2 * u, "%s must be (%d,%d,%d) with k = 2*stride", name.c_str(), cin, cout, k);
        ConvDesc d;
        d.Cin = cin; d.N = u * cout; d.taps = 2; d.off0 = 0; d.tstep = -1; d.bias_mod = cout;
        d.orow = (long)u * cout; d.ooff = -(long)((k - u) / 2) * cout; d.up = u; d.up_pad = (k - u) / 2;
        float* w32 = dev_alloc<float>(c, (size_t)2 * cin * u * cout);
        launch_repack_convT_w(c, s, w, cin, cout, k, u, w32);
        d.w32 = w32; d.bias = bias_of(c, name, cout);
        if (c->precision == MASG_PREC_BF16) tc_prepare_weights(c, s, d, false);
        c->hg_ups[i] = d;
        for (int j = 0; j < 3; ++j) {
            int rb = i * 3 + j, kk = c->hg_rk[j];
            for (int mth = 0; mth < 3; ++mth) {
                int dil = c->hg_rd[mth];
                std::string n1 = "resblocks." + std::to_string(rb) + ".convs1." + std::to_string(mth);
                std::string n2 = "resblocks." + std::to_string(rb) + ".convs2." + std::to_string(mth);
                w = folded_weight(c, s, n1, sh);
                MASG_CHECK(sh[0] == cout && sh[1] == cout && sh[2] == kk, "%s has an unsupported shape", n1.c_str());
                c->hg_c1[rb][mth] = make_conv(c, s, w, bias_of(c, n1, cout), cout, cout, kk, (kk * dil - dil) / 2, dil, nullptr, nullptr, false);
                w = folded_weight(c, s, n2, sh);
                MASG_CHECK(sh[0] == cout && sh[1] == cout && sh[2] == kk, "%s has an unsupported shape", n2.c_str());
                c->hg_c2[rb][mth] = make_conv(c, s, w, bias_of(c, n2, cout), cout, cout, kk, (kk - 1) / 2, 1, nullptr, nullptr, false);
            }
        }
    }
    w = folded_weight(c, s, "conv_post", sh);
    MASG_CHECK(sh[0] == 1 && sh[1] == 32 && sh[2] == 7, "conv_post must be (1,32,7)");
    float* pw = dev_alloc<float>(c, 7 * 32);
    launch_repack_conv_w(c, s, w, 1, 32, 7, nullptr, pw, 1, 0);               // → [7][32]
    c->hg_post_w = pw;
    float pb = 0.f;
    CUDA_OK(cudaMemcpyAsync(&pb, bias_of(c, "conv_post", 1), 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    c->hg_post_b = pb;
    c->weights_gen++;
    c->owned_hg.assign(c->owned.begin() + owned_mark, c->owned.end());
    c->owned.resize(owned_mark);
    c->hg_ready = true;
}

// activations are fp32 (MASG_PREC_FP32) or bf16 (MASG_PREC_BF16) channels-last [B, S, C]
static void convS(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const void* X, bool x_bf16, int B, long Sin, void* Y, long Sout,
                  int Cout, int in_act, float in_slope, const void* R1, const void* R2, float out_scale, int Q) {
    const bool bf = c->precision == MASG_PREC_BF16;
    ConvIO io;
    io.X = X; io.x_bf16 = x_bf16; io.x_ld = d.Cin; io.x_bs = Sin * d.Cin; io.Tin = (int)Sin;
    if (bf) io.Y16 = (bf16*)Y; else io.Y = (float*)Y;
    io.y_bs = Sout * Cout; io.y_limit = Sout * Cout; io.B = B; io.Q = Q;
    io.in_act = in_act; io.in_slope = in_slope; io.R1 = R1; io.R2 = R2; io.r_bf16 = bf; io.out_scale = out_scale;
    launch_conv(c, s, d, io);
}

// MASG_PREC_BF16 graph: activations are planar bf16 [B][C/8][S][8] (the tcgen05 operand layout, moved by TMA).  Where a
// tensor is only ever consumed through a leaky-relu (conv_pre output, convs1 outputs, the MRF mean) the producer's
// epilogue applies it; where the raw tensor is also a residual (ups / convs2 outputs) the consumer applies it in
// shared memory.
static void convP(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const void* X, bool x_planar, int B, long Sin, bf16* Y, long Sout,
                  int in_act, int out_act, const bf16* R1, const bf16* R2, float out_scale, float post_slope, int Q) {
    ConvIO io;
    io.X = X; io.x_bf16 = x_planar; io.x_planar = x_planar; io.x_ld = d.Cin; io.x_bs = Sin * d.Cin; io.Tin = (int)Sin;
    io.Y16 = Y; io.y_planar = 1; io.y_T = (int)Sout; io.y_bs = Sout * (d.bias_mod ? d.bias_mod : d.N); io.y_limit = io.y_bs;
    io.B = B; io.Q = Q;
    io.in_act = in_act; io.in_slope = 0.1f; io.out_act = out_act; io.out_slope = 0.1f;
    io.R1 = R1; io.R2 = R2; io.r_bf16 = 1; io.out_scale = out_scale; io.post_slope = post_slope;
    launch_conv(c, s, d, io);
}

static void hifigan_forward_bf16(masg_ctx* c, cudaStream_t s, const float* melT, int B, int T, float* wav, int16_t* w16,
                                 const int64_t* row_len, const int64_t* row_off) {
    Arena& ar = c->arena_b;
    long S = T;
    bf16* x = ar.get<bf16>((size_t)B * S * 512);
    // conv_pre (models.py:151); its output only feeds leaky_relu -> ups[0] (models.py:152-153)
    convP(c, s, c->hg_pre, melT, false, B, S, x, S, ACT_NONE, ACT_LRELU, nullptr, nullptr, 1.f, 1.f, (int)S);
    for (int i = 0; i < 4; ++i) {
        const int u = c->hg_up_rate[i], C = c->hg_ch[i + 1];
        const long So = S * u;
        const size_t n = (size_t)B * So * C;
        bf16* xu = ar.get<bf16>(n); bf16* xs = ar.get<bf16>(n); bf16* h = ar.get<bf16>(n);
        bf16* ya = ar.get<bf16>(n); bf16* yb = ar.get<bf16>(n);
        convP(c, s, c->hg_ups[i], x, true, B, S, xu, So, ACT_NONE, ACT_NONE, nullptr, nullptr, 1.f, 1.f, (int)S + 1);
        for (int j = 0; j < 3; ++j) {
            const int rb = i * 3 + j;
            const bf16* y = xu;
            for (int m = 0; m < 3; ++m) {
                // h = lrelu(c1(lrelu(y))); y' = c2(h) + y   (models.py:96-103).  Last pair of the block: fold the MRF sum,
                // the /3 and the leaky-relu that follows it (models.py:155-161: slope 0.1 before the next upsampler,
                // F.leaky_relu's default 0.01 before conv_post)
                const bool last = m == 2;
                bf16* yn = last ? xs : ((y == ya) ? yb : ya);
                const bf16* acc = (last && j > 0) ? xs : nullptr;
                const float scale = (last && j == 2) ? (1.0f / 3.0f) : 1.f;
                const float post = (last && j == 2) ? (i < 3 ? 0.1f : 0.01f) : 1.f;
                if (respair2_tc_supported(c, c->hg_c1[rb][m], c->hg_c2[rb][m])) {
                    // both convs in one kernel on a CTA pair, the weights' columns split over the two SMs (respair2_tc.cu): the pairs
                    // whose weights do not fit one SM (C = 128 k = 3, C = 64 k = 11) or leave it too few buffers (C = 64 k = 7)
                    launch_respair2_tc(c, s, c->hg_c1[rb][m], c->hg_c2[rb][m], y, B, (int)So, yn, acc, scale, post);
                } else if (respair_tc_supported(c, c->hg_c1[rb][m], c->hg_c2[rb][m])) {
                    // narrow stages: both convs in one kernel, h stays in shared memory (respair_tc.cu)
                    launch_respair_tc(c, s, c->hg_c1[rb][m], c->hg_c2[rb][m], y, B, (int)So, yn, acc, scale, post);
                } else {
                    convP(c, s, c->hg_c1[rb][m], y, true, B, So, h, So, ACT_LRELU, ACT_LRELU, nullptr, nullptr, 1.f, 1.f, (int)So);
                    convP(c, s, c->hg_c2[rb][m], h, true, B, So, yn, So, ACT_NONE, ACT_NONE, y, acc, scale, post, (int)So);
                }
                y = yn;
            }
        }
        x = xs; S = So;
    }
    launch_conv_post(c, s, x, 1, c->hg_post_w, c->hg_post_b, B, S, 1.0f, wav, w16, row_len, row_off);   // the leaky-relu is already applied
}

void hifigan_forward(masg_ctx* c, cudaStream_t s, const float* mel, int cl, int B, int T, float* wav, int16_t* w16,
                     const int64_t* row_len, const int64_t* row_off) {
    MASG_CHECK(c->hg_ready, "HiFi-GAN weights not committed (call masg_hifigan_commit)");
    MASG_CHECK(B > 0 && T > 0, "empty vocoder input B=%d T=%d", B, T);
    const bool bf = c->precision == MASG_PREC_BF16;
    const size_t esz = bf ? 2 : 4;
    Arena& ar = c->arena_b;
    ar.reset();
    c->tap_lr = c->tap_dec = nullptr;                        // the phase-B debug taps lived in this arena
    const float* melT = mel;
    if (!cl) {
        float* t = ar.get<float>((size_t)B * T * 80);
        launch_transpose_mel(c, s, mel, B, 80, T, t);
        melT = t;
    }
    if (bf) { hifigan_forward_bf16(c, s, melT, B, T, wav, w16, row_len, row_off); return; }
    long S = T;
    char* x = (char*)ar.alloc((size_t)B * S * 512 * esz);
    convS(c, s, c->hg_pre, melT, false, B, S, x, S, 512, ACT_NONE, 0.f, nullptr, nullptr, 1.f, (int)S);
    for (int i = 0; i < 4; ++i) {
        int u = c->hg_up_rate[i], C = c->hg_ch[i + 1];
        long So = S * u;
        size_t n = (size_t)B * So * C * esz;
        char* xu = (char*)ar.alloc(n); char* xs = (char*)ar.alloc(n); char* h = (char*)ar.alloc(n);
        char* ya = (char*)ar.alloc(n); char* yb = (char*)ar.alloc(n);
        // x = ups[i](leaky_relu(x, 0.1))   (models.py:152-153)
        convS(c, s, c->hg_ups[i], x, bf, B, S, xu, So, C, ACT_LRELU, 0.1f, nullptr, nullptr, 1.f, (int)S + 1);
        for (int j = 0; j < 3; ++j) {
            int rb = i * 3 + j;
            const char* y = xu;
            for (int m = 0; m < 3; ++m) {
                // xt = c2(lrelu(c1(lrelu(y)))); y = xt + y   (models.py:96-103)
                convS(c, s, c->hg_c1[rb][m], y, bf, B, So, h, So, C, ACT_LRELU, 0.1f, nullptr, nullptr, 1.f, (int)So);
                if (m < 2) {
                    char* yn = (y == ya) ? yb : ya;
                    convS(c, s, c->hg_c2[rb][m], h, bf, B, So, yn, So, C, ACT_LRELU, 0.1f, y, nullptr, 1.f, (int)So);
                    y = yn;
                } else {
                    // last pair of the block: fold the MRF sum (xs += block) and, on the last block, the /3 (models.py:155-160)
                    convS(c, s, c->hg_c2[rb][m], h, bf, B, So, xs, So, C, ACT_LRELU, 0.1f, y, j > 0 ? xs : nullptr,
                          j == 2 ? (1.0f / 3.0f) : 1.f, (int)So);
                }
            }
        }
        x = xs; S = So;
    }
    launch_conv_post(c, s, x, bf, c->hg_post_w, c->hg_post_b, B, S, 0.01f, wav, w16, row_len, row_off);   // F.leaky_relu default slope (models.py:161)
}
