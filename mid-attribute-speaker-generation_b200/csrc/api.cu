// extern "C" boundary of libmasg.so (include/masg.h).  No exceptions cross it: every entry point
// catches MasgError and returns non-zero with the message retrievable through masg_last_error().
#include "common.cuh"
#include "tc.cuh"
#include <cstring>
#include <cmath>
#include <algorithm>

ConvDesc make_conv(masg_ctx* c, cudaStream_t s, const float* w, const float* bias, int Cout, int Cin, int k, int pad, int dil,
                   const float* scale, const float* bias_override, int mode);

static thread_local std::string g_err;

#define API_BEGIN(ctx)                                                                   \
    if (!(ctx)) { g_err = "null masg_ctx"; return 1; }                                     \
    try {                                                                                  \
        int _dev = -1; cudaGetDevice(&_dev);                                               \
        if (_dev != (ctx)->device) CUDA_OK(cudaSetDevice((ctx)->device));
#define API_END(ctx)                                                                     \
        return 0;                                                                          \
    } catch (const MasgError& e) { (ctx)->err = e.msg; g_err = e.msg; return 1; }          \
      catch (const std::exception& e) { (ctx)->err = e.what(); g_err = e.what(); return 2; }

extern "C" {

const char* masg_version(void) { return "masg-b200 0.1 (sm_100a)"; }

int masg_ctx_create(int device, int precision, masg_ctx** out) {
    try {
        MASG_CHECK(out != nullptr, "null out pointer");
        MASG_CHECK(precision == MASG_PREC_BF16 || precision == MASG_PREC_FP32 || precision == MASG_PREC_BF16X3, "unknown precision %d", precision);
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        MASG_CHECK(e == cudaSuccess && n > 0, "no CUDA device available (%s): this library has no CPU fallback", cudaGetErrorString(e));
        MASG_CHECK(device >= 0 && device < n, "device %d out of range (%d devices)", device, n);
        cudaDeviceProp prop;
        CUDA_OK(cudaGetDeviceProperties(&prop, device));
        MASG_CHECK(prop.major == 10, "device %d is sm_%d%d; libmasg is built for sm_100a (B200) only", device, prop.major, prop.minor);
        CUDA_OK(cudaSetDevice(device));
        masg_ctx* c = new masg_ctx();
        // MASG_PREC_BF16X3 is the tensor-core path with every FastSpeech2 contraction in split-bf16: internally the bf16 graph + a flag
        c->device = device; c->precision = precision == MASG_PREC_BF16X3 ? MASG_PREC_BF16 : precision; c->x3 = precision == MASG_PREC_BF16X3;
        c->n_sm = prop.multiProcessorCount;
        *out = c;
        return 0;
    } catch (const MasgError& e) { g_err = e.msg; return 1; }
      catch (const std::exception& e) { g_err = e.what(); return 2; }
      catch (...) { g_err = "unknown exception in masg_ctx_create"; return 3; }
}

static void free_owned(masg_ctx* c) {
    for (auto* v : {&c->owned, &c->owned_fs2, &c->owned_hg}) { for (void* p : *v) cudaFree(p); v->clear(); }
    if (c->pos_long) cudaFree(c->pos_long);
    c->pos_long = nullptr; c->pos_long_n = 0;
}

void masg_ctx_destroy(masg_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_owned(c);
    c->arena.release(); c->arena_b.release();
    for (auto& kv : c->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (c->cap_stream) cudaStreamDestroy(c->cap_stream);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    if (c->io_buf) cudaFree(c->io_buf);
    if (c->out_buf) cudaFree(c->out_buf);
    if (c->out_alt) cudaFree(c->out_alt);
    for (int i = 0; i < 2; ++i) { if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]); if (c->ev_copied[i]) cudaEventDestroy(c->ev_copied[i]); }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    delete c;
}

const char* masg_last_error(masg_ctx* c) { return c ? c->err.c_str() : g_err.c_str(); }

int masg_load_tensor(masg_ctx* c, const char* model, const char* key, const float* dev_ptr, int ndim, const int64_t* shape) {
    API_BEGIN(c)
    MASG_CHECK(model && key && dev_ptr, "null argument");
    cudaPointerAttributes at;
    CUDA_OK(cudaPointerGetAttributes(&at, dev_ptr));
    MASG_CHECK(at.type == cudaMemoryTypeDevice && at.device == c->device, "tensor '%s' is not on cuda:%d (CPU tensors are an error: no fallback)", key, c->device);
    TensorRef t; t.ptr = dev_ptr; t.shape.assign(shape, shape + ndim);
    if (!strcmp(model, "fs2")) { c->fs2_t[key] = t; c->fs2_ready = false; }
    else if (!strcmp(model, "hifigan")) { c->hg_t[key] = t; c->hg_ready = false; }
    else MASG_CHECK(false, "unknown model '%s' (expected 'fs2' or 'hifigan')", model);
    API_END(c)
}

int masg_fs2_commit(masg_ctx* c, void* stream) {
    API_BEGIN(c)
    fs2_commit(c, (cudaStream_t)stream);
    API_END(c)
}
int masg_hifigan_commit(masg_ctx* c, void* stream) {
    API_BEGIN(c)
    hifigan_commit(c, (cudaStream_t)stream);
    API_END(c)
}

int masg_fs2_phase_a(masg_ctx* c, void* stream, const int64_t* texts, const int64_t* accents, int accents_is_1d,
                     const int64_t* src_lens, int B, int L, const float* spk_emb, float p_ctl, float e_ctl, float d_ctl,
                     const float* d_target, const float* p_target, const float* e_target, float* pitch, float* energy, float* log_d,
                     float* d_rounded, int64_t* mel_lens, uint8_t* src_mask, int64_t* h_tmax) {
    API_BEGIN(c)
    MASG_CHECK(texts && src_lens && spk_emb && pitch && energy && log_d, "null argument");
    fs2_phase_a(c, (cudaStream_t)stream, texts, accents, accents_is_1d, src_lens, B, L, spk_emb, p_ctl, e_ctl, d_ctl, d_target,
                p_target, e_target, pitch, energy, log_d, d_rounded, mel_lens, src_mask, h_tmax);
    API_END(c)
}
int masg_fs2_phase_b(masg_ctx* c, void* stream, int T, float* mel, float* postnet_mel, uint8_t* mel_mask) {
    API_BEGIN(c)
    MASG_CHECK(mel && postnet_mel, "null argument");
    fs2_phase_b(c, (cudaStream_t)stream, T, mel, postnet_mel, mel_mask);
    API_END(c)
}

int masg_duration_round(masg_ctx* c, void* stream, const float* log_d, float d_ctl, int64_t n, float* d_rounded) {
    API_BEGIN(c)
    if (n > 0) launch_duration(c, (cudaStream_t)stream, log_d, d_ctl, n, d_rounded);
    API_END(c)
}
int masg_mel_lengths(masg_ctx* c, void* stream, const float* d, int B, int L, int64_t* mel_lens) {
    API_BEGIN(c)
    if (B > 0) launch_mel_lens(c, (cudaStream_t)stream, d, B, L, mel_lens, nullptr);
    API_END(c)
}
int masg_length_regulate(masg_ctx* c, void* stream, const float* x, const float* d, int B, int L, int C, int T, float* out,
                         int64_t* mel_lens, uint8_t* mel_mask) {
    API_BEGIN(c)
    MASG_CHECK(C % 4 == 0, "C=%d must be a multiple of 4", C);
    launch_length_regulate(c, (cudaStream_t)stream, x, d, B, L, C, T, nullptr, out, mel_lens, mel_mask);
    API_END(c)
}
int masg_bucketize_embed_add(masg_ctx* c, void* stream, const float* v, const float* bins, int n_bins_m1, const float* table,
                             int64_t rows, int C, float* x, int64_t* idx_out) {
    API_BEGIN(c)
    MASG_CHECK(C % 4 == 0, "C=%d must be a multiple of 4", C);
    if (rows > 0) launch_bucket_add(c, (cudaStream_t)stream, v, bins, n_bins_m1, table, rows, C, x, idx_out);
    API_END(c)
}

int masg_hifigan_forward(masg_ctx* c, void* stream, const float* mel, int mel_is_channels_last, int B, int T, float* wav,
                         int16_t* wav_i16) {
    API_BEGIN(c)
    MASG_CHECK(mel && (wav || wav_i16), "null argument");
    hifigan_forward(c, (cudaStream_t)stream, mel, mel_is_channels_last, B, T, wav, wav_i16);
    API_END(c)
}

int masg_hifigan_forward_packed(masg_ctx* c, void* stream, const float* mel, int mel_is_channels_last, int B, int T,
                                const int64_t* mel_lens, int hop, int16_t* wav_packed, int64_t* offsets) {
    API_BEGIN(c)
    MASG_CHECK(mel && mel_lens && wav_packed && offsets && hop > 0, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    // offsets [B + 1] are the caller's; the per-row lengths live in the I/O staging buffer (the arenas are reset by the forward)
    if (c->io_cap < (size_t)B * 8 + 256) {
        if (c->io_buf) { CUDA_OK(cudaStreamSynchronize(s)); cudaFree(c->io_buf); c->io_buf = nullptr; c->io_cap = 0; }
        CUDA_OK(cudaMalloc(&c->io_buf, (size_t)B * 8 + 4096)); c->io_cap = (size_t)B * 8 + 4096;
    }
    int64_t* len = (int64_t*)c->io_buf;
    launch_wav_offsets(c, s, mel_lens, B, hop, offsets, len);
    hifigan_forward(c, s, mel, mel_is_channels_last, B, T, nullptr, wav_packed, len, offsets);
    API_END(c)
}

int masg_gmm_params(masg_ctx* c, void* stream, const float* meta, int N, float* pi, float* mu, float* sigma) {
    API_BEGIN(c)
    MASG_CHECK(c->fs2_ready, "FastSpeech2 weights not committed");
    if (N > 0) launch_gmm_params(c, (cudaStream_t)stream, meta, N, pi, mu, sigma);
    API_END(c)
}
int masg_gmm_sample(masg_ctx* c, void* stream, const float* pi, const float* mu, const float* scale, int N, int Np, int K, int D,
                    uint64_t seed, uint64_t offset, const int32_t* k_inject, const float* eps_inject, float* out, int32_t* k_out) {
    API_BEGIN(c)
    MASG_CHECK(Np == 1 || Np == N, "Np=%d must be 1 or N=%d", Np, N);
    if (N > 0) launch_gmm_sample(c, (cudaStream_t)stream, pi, mu, scale, N, Np, K, D, seed, offset, k_inject, eps_inject, out, k_out);
    API_END(c)
}

// Exact earth-mover plan for an n x m problem (n, m <= 4): enumerate the basic solutions of the
// transportation polytope (supports of n+m-1 cells), keep the cheapest feasible one; ties keep the
// lexicographically first support.  Stands in for POT ot.emd (model/distributions.py:22).
static void emd_small(const double* a, const double* b_in, const double* M, int n, int m, double* G) {
    double b[8], sa = 0, sb = 0;
    for (int i = 0; i < n; ++i) sa += a[i];
    for (int j = 0; j < m; ++j) sb += b_in[j];
    for (int j = 0; j < m; ++j) b[j] = b_in[j] * (sa / sb);
    int nc = n * m, ns = n + m - 1;
    std::vector<int> idx(ns);
    for (int i = 0; i < ns; ++i) idx[i] = i;
    double best = INFINITY; std::vector<int> bsup; std::vector<double> bx;
    while (true) {
        // solve rows: n row-sum constraints + first m-1 column constraints
        double A[8][9];
        for (int r = 0; r < ns; ++r) {
            for (int t = 0; t < ns; ++t) {
                int ci = idx[t] / m, cj = idx[t] % m;
                A[r][t] = r < n ? (ci == r ? 1.0 : 0.0) : (cj == r - n ? 1.0 : 0.0);
            }
            A[r][ns] = r < n ? a[r] : b[r - n];
        }
        bool sing = false;
        for (int col = 0; col < ns && !sing; ++col) {
            int piv = col; double mx = fabs(A[col][col]);
            for (int r = col + 1; r < ns; ++r) if (fabs(A[r][col]) > mx) { mx = fabs(A[r][col]); piv = r; }
            if (mx < 1e-12) { sing = true; break; }
            if (piv != col) for (int t = 0; t <= ns; ++t) std::swap(A[piv][t], A[col][t]);
            for (int r = 0; r < ns; ++r) if (r != col) {
                double f = A[r][col] / A[col][col];
                if (f != 0.0) for (int t = col; t <= ns; ++t) A[r][t] -= f * A[col][t];
            }
        }
        if (!sing) {
            double x[8]; bool ok = true; double lastcol = 0;
            for (int t = 0; t < ns; ++t) { x[t] = A[t][ns] / A[t][t]; if (x[t] < -1e-12) ok = false; }
            for (int t = 0; t < ns && ok; ++t) if (idx[t] % m == m - 1) lastcol += x[t];
            if (ok && fabs(lastcol - b[m - 1]) > 1e-9) ok = false;
            if (ok) {
                double cost = 0;
                for (int t = 0; t < ns; ++t) cost += x[t] * M[idx[t]];
                if (cost < best - 1e-15) { best = cost; bsup = idx; bx.assign(x, x + ns); }
            }
        }
        int t = ns - 1;
        while (t >= 0 && idx[t] == nc - ns + t) --t;
        if (t < 0) break;
        ++idx[t];
        for (int r = t + 1; r < ns; ++r) idx[r] = idx[r - 1] + 1;
    }
    for (int i = 0; i < nc; ++i) G[i] = 0.0;
    for (size_t t = 0; t < bsup.size(); ++t) G[bsup[t]] = bx[t] > 0 ? bx[t] : 0.0;
}

int masg_gmm_interpolate(masg_ctx* c, void* stream, const float* pi_a, const float* mu_a, const float* sig_a, const float* pi_b,
                         const float* mu_b, const float* sig_b, int K, int D, float t, int recompute_plan, float* weights,
                         float* comp_mu, float* comp_scale, double* h_cost) {
    API_BEGIN(c)
    cudaStream_t s = (cudaStream_t)stream;
    MASG_CHECK(K >= 1 && K <= 4, "K=%d unsupported (exact EMD enumerates supports for K <= 4)", K);
    launch_gmm_interp_comps(c, s, mu_a, sig_a, mu_b, sig_b, K, D, t, comp_mu, comp_scale);
    if (recompute_plan) {
        c->arena_b.reset(); c->tap_lr = c->tap_dec = nullptr;             // the phase-B taps lived in this arena
        double* cost = c->arena_b.get<double>(K * K);
        launch_gmm_interp_cost(c, s, mu_a, sig_a, mu_b, sig_b, K, D, cost);
        double hc[16], G[16], ha[4], hb[4]; float fa[4], fb[4], fw[16];
        CUDA_OK(cudaMemcpyAsync(hc, cost, K * K * sizeof(double), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaMemcpyAsync(fa, pi_a, K * sizeof(float), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaMemcpyAsync(fb, pi_b, K * sizeof(float), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        for (int i = 0; i < K; ++i) { ha[i] = fa[i]; hb[i] = fb[i]; }
        emd_small(ha, hb, hc, K, K, G);
        for (int i = 0; i < K * K; ++i) fw[i] = (float)G[i];
        CUDA_OK(cudaMemcpyAsync(weights, fw, K * K * sizeof(float), cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaStreamSynchronize(s));          // fw is a stack buffer
        if (h_cost) memcpy(h_cost, hc, K * K * sizeof(double));
    }
    API_END(c)
}

int masg_gmm_barycenter(masg_ctx* c, void* stream, const float* pi, const float* mu, const float* sigma, const float* h_rate,
                        int A, int K, int D, float* bary_mean, float* bary_std, float* bary_weight, int32_t* assign) {
    API_BEGIN(c)
    cudaStream_t s = (cudaStream_t)stream;
    MASG_CHECK(A >= 1 && A <= 8 && K >= 1, "A=%d K=%d unsupported", A, K);
    int nb = 1; for (int a = 0; a < A; ++a) nb *= K;
    c->arena_b.reset(); c->tap_lr = c->tap_dec = nullptr;
    float* rate = c->arena_b.get<float>(A);
    CUDA_OK(cudaMemcpyAsync(rate, h_rate, A * sizeof(float), cudaMemcpyHostToDevice, s));
    launch_gmm_bary(c, s, mu, sigma, rate, A, K, D, bary_mean, bary_std);
    launch_gmm_bary_assign(c, s, mu, sigma, bary_mean, bary_std, A, K, D, nb, assign);
    std::vector<int32_t> ha(A * K); std::vector<float> hp(A * K), hw(nb, 0.f);
    CUDA_OK(cudaMemcpyAsync(ha.data(), assign, A * K * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(hp.data(), pi, A * K * sizeof(float), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    for (int i = 0; i < A; ++i) for (int j = 0; j < K; ++j) hw[ha[i * K + j]] += h_rate[i] * hp[i * K + j];
    CUDA_OK(cudaMemcpyAsync(bary_weight, hw.data(), nb * sizeof(float), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaStreamSynchronize(s));
    API_END(c)
}

}  // extern "C"

// Replay `body` (a sequence of kernel launches on library-owned buffers) as a CUDA graph when the small-batch latency path asks
// for it (MASG_GRAPH=1 / 0, or automatically for batches of <= MASG_GRAPH_ROWS = 4096 phoneme rows; measured: nothing at C2 —
// its 161 kernels take 20-30 us each, the launches are not the limit — and 4 % at 2 x 1000 phonemes, the C5-on-8-GPUs shard).  First sight of a key: plain run
// (sizes the arenas, sets kernel attributes); second: capture + instantiate + launch; afterwards: one cudaGraphLaunch, as long
// as no buffer the captured kernels point into has moved and the weights have not been re-committed.
static int graph_rows_limit() {
    static int lim = -2;
    if (lim == -2) {
        const char* g = getenv("MASG_GRAPH"); const char* r = getenv("MASG_GRAPH_ROWS");
        lim = g ? (atoi(g) ? (1 << 30) : -1) : (r ? atoi(r) : 4096);
    }
    return lim;
}
template <typename F>
static void run_graphed(masg_ctx* c, cudaStream_t s, const std::vector<long>& key, long rows, F body) {
    if (rows > graph_rows_limit() || c->profiling || c->taps) { body(s); return; }
    GraphEntry& e = c->graphs[key];
    void* snap[6] = {c->arena.base, c->arena_b.base, c->io_buf, c->out_buf, (void*)(intptr_t)c->weights_gen, c->pos_long};
    const bool same = !memcmp(snap, e.snap, sizeof(snap));
    if (e.exec && same) {
        CUDA_OK(cudaGraphLaunch(e.exec, s));
        c->launches += e.launches;
        if (key[0] == 0) { c->va_out = e.va_out; c->d_rounded_dev = e.d_rounded_dev; c->mel_lens_dev = e.mel_lens_dev; c->src_lens_dev = e.src_lens_dev; c->B = (int)key[1]; c->L = (int)key[2]; }
        return;
    }
    if (e.exec) { cudaGraphExecDestroy(e.exec); e.exec = nullptr; }
    if (e.seen++ == 0 || !same) {                             // plain run; remember where the buffers are now
        body(s);
        void* after[6] = {c->arena.base, c->arena_b.base, c->io_buf, c->out_buf, (void*)(intptr_t)c->weights_gen, c->pos_long};
        memcpy(e.snap, after, sizeof(after));
        return;
    }
    // the caller's stream may be the legacy default stream, which cannot be captured: the body is recorded on a stream of the
    // context's own and the instantiated graph is then launched into the caller's stream
    if (!c->cap_stream) CUDA_OK(cudaStreamCreateWithFlags(&c->cap_stream, cudaStreamNonBlocking));
    const int64_t l0 = c->launches;
    CUDA_OK(cudaStreamBeginCapture(c->cap_stream, cudaStreamCaptureModeRelaxed));
    try { body(c->cap_stream); }
    catch (...) { cudaGraph_t g = nullptr; cudaStreamEndCapture(c->cap_stream, &g); if (g) cudaGraphDestroy(g); throw; }
    cudaGraph_t g = nullptr;
    CUDA_OK(cudaStreamEndCapture(c->cap_stream, &g));
    void* after[6] = {c->arena.base, c->arena_b.base, c->io_buf, c->out_buf, (void*)(intptr_t)c->weights_gen, c->pos_long};
    if (memcmp(after, snap, sizeof(after))) {                 // a buffer moved while capturing: the graph is stale already
        cudaGraphDestroy(g);
        memcpy(e.snap, after, sizeof(after));
        body(s);
        return;
    }
    cudaError_t ie = cudaGraphInstantiate(&e.exec, g, 0);
    cudaGraphDestroy(g);
    CUDA_OK(ie);
    e.launches = c->launches - l0;
    e.va_out = c->va_out; e.d_rounded_dev = c->d_rounded_dev; e.mel_lens_dev = c->mel_lens_dev; e.src_lens_dev = c->src_lens_dev;
    CUDA_OK(cudaGraphLaunch(e.exec, s));
}

// shared body of the two host-buffer entry points.  packed == 0: rows of stride 256 * t_cap in h_wav (every row is copied up to
// T_max frames).  packed != 0: the vocoder's last kernel writes the trimmed rows back to back (row b = its mel_len[b] * 256
// samples at h_offsets[b]) and ONE D2H copy moves exactly the valid samples.
static void synthesize_host(masg_ctx* c, cudaStream_t s, const int64_t* h_texts, const int64_t* h_accents, const int64_t* h_src_lens,
                            int B, int L, const float* h_spk_emb, float p_ctl, float e_ctl, float d_ctl, int packed, int64_t cap,
                            int16_t* h_wav, int64_t* h_offsets, int64_t* h_mel_lens, int64_t* h_tmax, int* ticket = nullptr) {
    MASG_CHECK(h_texts && h_accents && h_src_lens && h_spk_emb && h_wav, "null argument");
    MASG_CHECK(B > 0 && L > 0, "empty batch B=%d L=%d", B, L);
    const bool defer = ticket != nullptr;
    if (defer) {                                             // the other output buffer: the previous call's audio may still be in flight
        std::swap(c->out_buf, c->out_alt); std::swap(c->out_cap, c->out_alt_cap); c->out_idx ^= 1;
        if (!c->copy_stream) {
            CUDA_OK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) { CUDA_OK(cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming)); CUDA_OK(cudaEventCreateWithFlags(&c->ev_copied[i], cudaEventDisableTiming)); }
        }
    }
    const int oi = c->out_idx;
    // whatever was last copied out of THIS buffer (two deferred calls ago) must have left before anything overwrites it
    if (c->ev_copied[oi]) CUDA_OK(cudaStreamWaitEvent(s, c->ev_copied[oi], 0));
    // two growth-only staging buffers of their own (arena / arena_b are reset by the phases): io_buf for the inputs and phase A's
    // small outputs (sized by B and L, before anything is enqueued), out_buf for mel / postnet mel / audio (sized once T_max and
    // the lengths are known, right after phase A's sync — growing it there costs no extra synchronisation)
    const size_t rows = (size_t)B * L;
    const size_t need_in = rows * 8 * 2 + (size_t)B * 8 * 4 + 8 + (size_t)B * 256 * 4 + rows * 4 * 4 + 8192;
    if (c->io_cap < need_in) {
        if (c->io_buf) { CUDA_OK(cudaStreamSynchronize(s)); cudaFree(c->io_buf); c->io_buf = nullptr; c->io_cap = 0; }
        CUDA_OK(cudaMalloc(&c->io_buf, need_in * 2)); c->io_cap = need_in * 2;
    }
    char* p = (char*)c->io_buf;
    auto take = [&](size_t bytes) { void* r = p; p += (bytes + 255) & ~size_t(255); return r; };
    int64_t* texts = (int64_t*)take(rows * 8); int64_t* acc = (int64_t*)take(rows * 8);
    int64_t* lens = (int64_t*)take(B * 8); int64_t* mel_lens = (int64_t*)take(B * 8);
    int64_t* woff = (int64_t*)take((size_t)(B + 1) * 8); int64_t* wlen = (int64_t*)take(B * 8);
    float* spk = (float*)take((size_t)B * 256 * 4);
    float* pitch = (float*)take(rows * 4); float* energy = (float*)take(rows * 4); float* logd = (float*)take(rows * 4);
    float* drnd = (float*)take(rows * 4);
    CUDA_OK(cudaMemcpyAsync(texts, h_texts, rows * 8, cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(acc, h_accents, rows * 8, cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(lens, h_src_lens, B * 8, cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(spk, h_spk_emb, (size_t)B * 256 * 4, cudaMemcpyHostToDevice, s));
    int64_t tmax = 0;
    // the controls are baked into a captured graph: they are part of the key (as float bit patterns)
    auto fbits = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (long)u; };
    run_graphed(c, s, {0, B, L, 0, oi, fbits(p_ctl), fbits(d_ctl)}, (long)rows, [&](cudaStream_t bs) {
        fs2_phase_a(c, bs, texts, acc, 0, lens, B, L, spk, p_ctl, e_ctl, d_ctl, nullptr, nullptr, nullptr, pitch, energy, logd, drnd, mel_lens,
                    nullptr, nullptr);
    });
    fs2_phase_a_publish(c, s, &tmax);
    MASG_CHECK(tmax > 0, "no frames to synthesise (all durations are zero)");
    const int T = (int)tmax;
    // phase A's single sync left [mel_lens | T_max] in the pinned scratch: the host knows every row's length from here on
    int64_t total = 0;
    for (int b = 0; b < B; ++b) { if (h_mel_lens) h_mel_lens[b] = c->h_pinned[b]; if (h_offsets) h_offsets[b] = total * 256; total += std::max<int64_t>(c->h_pinned[b], 0); }
    if (h_offsets) h_offsets[B] = total * 256;
    if (packed) MASG_CHECK(total * 256 <= cap, "the batch produces %lld samples, capacity is %lld", (long long)(total * 256), (long long)cap);
    else MASG_CHECK(tmax <= cap, "T_max=%lld exceeds t_cap=%lld", (long long)tmax, (long long)cap);
    const size_t out_bytes = ((size_t)B * T * 80 * 4 + 256) * 2 + (packed ? (size_t)total * 512 : (size_t)B * T * 512) + 1024;
    if (c->out_cap < out_bytes) {                            // the stream is idle here (phase A just synchronised)
        if (c->ev_copied[oi]) CUDA_OK(cudaEventSynchronize(c->ev_copied[oi]));
        if (c->out_buf) { cudaFree(c->out_buf); c->out_buf = nullptr; c->out_cap = 0; }
        CUDA_OK(cudaMalloc(&c->out_buf, out_bytes + out_bytes / 4)); c->out_cap = out_bytes + out_bytes / 4;
    }
    p = (char*)c->out_buf;
    float* mel = (float*)take((size_t)B * T * 80 * 4); float* post = (float*)take((size_t)B * T * 80 * 4);
    int16_t* w16 = (int16_t*)take(packed ? (size_t)total * 512 : (size_t)B * T * 512);
    run_graphed(c, s, {1, B, L, T, packed, oi}, (long)rows, [&](cudaStream_t bs) {
        fs2_phase_b(c, bs, T, mel, post, nullptr);
        if (packed) {
            launch_wav_offsets(c, bs, mel_lens, B, 256, woff, wlen);
            hifigan_forward(c, bs, post, 1, B, T, nullptr, w16, wlen, woff);
        } else {
            hifigan_forward(c, bs, post, 1, B, T, nullptr, w16);
        }
    });
    // deferred: the copy goes to the context's copy stream behind an event and the call returns; masg_synthesize_host_wait(ticket)
    // makes the audio visible.  The next call's phase A (and its one synchronisation) then overlaps this copy.
    cudaStream_t cs = s;
    if (defer) {
        CUDA_OK(cudaEventRecord(c->ev_done[oi], s));
        CUDA_OK(cudaStreamWaitEvent(c->copy_stream, c->ev_done[oi], 0));
        cs = c->copy_stream;
    }
    if (packed) CUDA_OK(cudaMemcpyAsync(h_wav, w16, (size_t)total * 512, cudaMemcpyDeviceToHost, cs));
    else CUDA_OK(cudaMemcpy2DAsync(h_wav, (size_t)cap * 256 * 2, w16, (size_t)T * 256 * 2, (size_t)T * 256 * 2, B, cudaMemcpyDeviceToHost, cs));
    if (defer) { CUDA_OK(cudaEventRecord(c->ev_copied[oi], cs)); *ticket = oi; }
    else CUDA_OK(cudaStreamSynchronize(s));
    if (h_tmax) *h_tmax = tmax;
}

extern "C" {

int masg_synthesize_host(masg_ctx* c, void* stream, const int64_t* h_texts, const int64_t* h_accents, const int64_t* h_src_lens,
                         int B, int L, const float* h_spk_emb, float p_ctl, float e_ctl, float d_ctl, int t_cap,
                         int16_t* h_wav_i16, int64_t* h_mel_lens, int64_t* h_tmax) {
    API_BEGIN(c)
    synthesize_host(c, (cudaStream_t)stream, h_texts, h_accents, h_src_lens, B, L, h_spk_emb, p_ctl, e_ctl, d_ctl, 0, t_cap, h_wav_i16,
                    nullptr, h_mel_lens, h_tmax);
    API_END(c)
}

int masg_synthesize_host_packed(masg_ctx* c, void* stream, const int64_t* h_texts, const int64_t* h_accents, const int64_t* h_src_lens,
                                int B, int L, const float* h_spk_emb, float p_ctl, float e_ctl, float d_ctl, int64_t cap_samples,
                                int16_t* h_wav_packed, int64_t* h_offsets, int64_t* h_mel_lens, int64_t* h_tmax) {
    API_BEGIN(c)
    MASG_CHECK(h_offsets != nullptr, "null h_offsets");
    synthesize_host(c, (cudaStream_t)stream, h_texts, h_accents, h_src_lens, B, L, h_spk_emb, p_ctl, e_ctl, d_ctl, 1, cap_samples,
                    h_wav_packed, h_offsets, h_mel_lens, h_tmax);
    API_END(c)
}

int masg_synthesize_host_packed_async(masg_ctx* c, void* stream, const int64_t* h_texts, const int64_t* h_accents, const int64_t* h_src_lens,
                                      int B, int L, const float* h_spk_emb, float p_ctl, float e_ctl, float d_ctl, int64_t cap_samples,
                                      int16_t* h_wav_packed, int64_t* h_offsets, int64_t* h_mel_lens, int64_t* h_tmax, int* ticket) {
    API_BEGIN(c)
    MASG_CHECK(h_offsets != nullptr && ticket != nullptr, "null h_offsets / ticket");
    synthesize_host(c, (cudaStream_t)stream, h_texts, h_accents, h_src_lens, B, L, h_spk_emb, p_ctl, e_ctl, d_ctl, 1, cap_samples,
                    h_wav_packed, h_offsets, h_mel_lens, h_tmax, ticket);
    API_END(c)
}

int masg_synthesize_host_wait(masg_ctx* c, int ticket) {
    API_BEGIN(c)
    MASG_CHECK((ticket == 0 || ticket == 1) && c->ev_copied[ticket], "unknown ticket %d", ticket);
    CUDA_OK(cudaEventSynchronize(c->ev_copied[ticket]));
    API_END(c)
}

// Single Conv1d / ConvTranspose1d through the context's dense-contraction kernel (diagnostic + micro-benchmark hook).
int masg_conv1d(masg_ctx* c, void* stream, const float* x, int B, int T, int Cin, const float* w, const float* bias, int Cout,
                int k, int dilation, int stride_transposed, float in_lrelu_slope, int out_act, int flags, int reps, float* y,
                float* ms_out) {
    API_BEGIN(c)
    cudaStream_t s = (cudaStream_t)stream;
    MASG_CHECK(x && w && y && B > 0 && T > 0 && reps >= 1, "bad argument");
    const int split = (flags & 8) ? 2 : (flags & 1), act_bf16 = (flags & 2) && c->precision == MASG_PREC_BF16, residual = flags & 4;
    MASG_CHECK(!residual || (Cin == Cout && stride_transposed == 0), "residual needs Cin == Cout and a plain conv");
    const size_t keep = c->owned.size();
    ConvDesc d; int Q = T; long So = T;
    if (stride_transposed > 0) {
        const int u = stride_transposed;
        MASG_CHECK(k == 2 * u, "transposed conv needs k = 2*stride");
        d.Cin = Cin; d.N = u * Cout; d.taps = 2; d.off0 = 0; d.tstep = -1; d.bias_mod = Cout;
        d.orow = (long)u * Cout; d.ooff = -(long)((k - u) / 2) * Cout; d.up = u; d.up_pad = (k - u) / 2;
        float* w32 = nullptr; CUDA_OK(cudaMalloc(&w32, (size_t)2 * Cin * u * Cout * 4)); c->owned.push_back(w32);
        launch_repack_convT_w(c, s, w, Cin, Cout, k, u, w32);
        d.w32 = w32; d.bias = bias;
        if (c->precision == MASG_PREC_BF16) tc_prepare_weights(c, s, d, split);
        Q = T + 1; So = (long)T * u;
    } else {
        d = make_conv(c, s, w, bias, Cout, Cin, k, (k * dilation - dilation) / 2, dilation, nullptr, nullptr, split);
    }
    ConvIO io;
    io.X = x; io.x_ld = Cin; io.x_bs = (long)T * Cin; io.Tin = T; io.Y = y; io.y_bs = So * Cout; io.y_limit = So * Cout;
    io.B = B; io.Q = Q; io.out_act = out_act;
    bf16 *x16 = nullptr, *y16 = nullptr;
    if (act_bf16) {                                           // HiFi-GAN style: bf16 activations in HBM
        CUDA_OK(cudaMalloc(&x16, (size_t)B * T * Cin * 2)); c->owned.push_back(x16);
        CUDA_OK(cudaMalloc(&y16, (size_t)B * So * Cout * 2)); c->owned.push_back(y16);
        launch_to_planar(c, s, x, x16, B, T, Cin);                  // planar [B][C/8][T][8], moved by TMA
        io.X = x16; io.x_bf16 = 1; io.x_planar = 1; io.Y = nullptr; io.Y16 = y16; io.y_planar = 1; io.y_T = (int)So;
    }
    if (residual) { io.R1 = io.X; io.r_bf16 = act_bf16; }
    if (residual && (flags & 16) && act_bf16) {               // residual in its own buffer (as in the generator: x, not the conv's input h)
        bf16* r16 = nullptr; CUDA_OK(cudaMalloc(&r16, (size_t)B * T * Cin * 2)); c->owned.push_back(r16);
        CUDA_OK(cudaMemcpyAsync(r16, x16, (size_t)B * T * Cin * 2, cudaMemcpyDeviceToDevice, s));
        io.R1 = r16;
    }
    if (in_lrelu_slope != 0.f) { io.in_act = ACT_LRELU; io.in_slope = in_lrelu_slope; }
    io.pair_mode = (flags & 32) ? 1 : ((flags & 64) ? -1 : 0);
    launch_conv(c, s, d, io);                                  // warm-up / result
    cudaEvent_t e0, e1; CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1));
    CUDA_OK(cudaEventRecord(e0, s));
    for (int r = 0; r < reps; ++r) launch_conv(c, s, d, io);
    CUDA_OK(cudaEventRecord(e1, s));
    CUDA_OK(cudaStreamSynchronize(s));
    float ms = 0.f; CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms_out) *ms_out = ms / reps;
    if (getenv("MASG_TC_DEBUG") && (atoi(getenv("MASG_TC_DEBUG")) & 16)) {
        unsigned long long st[16]; tc_debug_stats(st, true);
        const double n = (double)c->n_sm * (reps + 1);
        fprintf(stderr, "  [tc stats, kcycles per CTA per launch] issuer: wait acc_empty %.0f a_full %.0f w_full %.0f of total %.0f | epilogue: wait acc_full %.0f of %.0f | copy warp wait a_empty %.0f | weight warp wait w_empty %.0f | worker wait a_raw %.0f\n",
                st[0] / n / 1e3, st[1] / n / 1e3, st[2] / n / 1e3, st[3] / n / 1e3, st[4] / n / 1e3, st[5] / n / 1e3, st[6] / n / 1e3, st[7] / n / 1e3, st[8] / n / 1e3);
    }
    if (getenv("MASG_TC_DEBUG") && (atoi(getenv("MASG_TC_DEBUG")) & 32)) {
        // timeline of CTA 0 (last launch): one line per tile, cycles relative to the first printed event
        static long long tr[5 * 96 * 4];
        const int nt = tc_debug_trace(tr, 5 * 96 * 4, true);
        auto at = [&](int role, int it, int e) { return tr[(role * nt + it) * 4 + e]; };
        const int i0 = 40, i1 = 56;
        const long long base = at(0, i0, 0);
        fprintf(stderr, "  [tc trace, CTA 0, cycles] tile | copy: wait-begin slot-free issued | worker: landed done | issuer: acc-free a-ready committed | epilogue: idle acc-full done\n");
        for (int it = i0; it < i1; ++it) {
            fprintf(stderr, "   %3d |", it);
            const int ne[4] = {3, 2, 3, 3};
            for (int role = 0; role < 4; ++role) {
                for (int e = 0; e < ne[role]; ++e) fprintf(stderr, " %7lld", at(role, it, e) ? at(role, it, e) - base : -1LL);
                fprintf(stderr, " |");
            }
            fprintf(stderr, "\n");
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (act_bf16) { launch_from_planar(c, s, y16, y, B, (int)So, Cout); CUDA_OK(cudaStreamSynchronize(s)); }
    while (c->owned.size() > keep) { cudaFree(c->owned.back()); c->owned.pop_back(); }
    API_END(c)
}

int masg_respair(masg_ctx* c, void* stream, const float* x, const float* acc, int B, int T, int C, const float* w1,
                 const float* b1, const float* w2, const float* b2, int k, int dilation, float out_scale, float post_slope,
                 int reps, float* y, float* ms_out) {
    API_BEGIN(c)
    cudaStream_t s = (cudaStream_t)stream;
    MASG_CHECK(x && w1 && w2 && b1 && b2 && y && B > 0 && T > 0 && reps >= 1, "bad argument");
    MASG_CHECK(c->precision == MASG_PREC_BF16, "masg_respair: the fused pair exists in MASG_PREC_BF16 only");
    const size_t keep = c->owned.size();
    ConvDesc d1 = make_conv(c, s, w1, b1, C, C, k, (k * dilation - dilation) / 2, dilation, nullptr, nullptr, false);
    ConvDesc d2 = make_conv(c, s, w2, b2, C, C, k, (k - 1) / 2, 1, nullptr, nullptr, false);
    const bool two_sm = respair2_tc_supported(c, d1, d2), one_sm = !two_sm && respair_tc_supported(c, d1, d2);   // the order hifigan.cu uses
    MASG_CHECK(one_sm || two_sm, "masg_respair: unsupported pair C=%d k=%d dilation=%d", C, k, dilation);
    auto run = [&](const bf16* xi, bf16* yo, const bf16* ai) {
        if (one_sm) launch_respair_tc(c, s, d1, d2, xi, B, T, yo, ai, out_scale, post_slope);
        else launch_respair2_tc(c, s, d1, d2, xi, B, T, yo, ai, out_scale, post_slope);
    };
    const size_t n = (size_t)B * T * C;
    bf16 *x16 = nullptr, *y16 = nullptr, *a16 = nullptr;
    CUDA_OK(cudaMalloc(&x16, n * 2)); c->owned.push_back(x16);
    CUDA_OK(cudaMalloc(&y16, n * 2)); c->owned.push_back(y16);
    launch_to_planar(c, s, x, x16, B, T, C);
    if (acc) { CUDA_OK(cudaMalloc(&a16, n * 2)); c->owned.push_back(a16); launch_to_planar(c, s, acc, a16, B, T, C); }
    run(x16, y16, a16);
    cudaEvent_t e0, e1; CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1));
    CUDA_OK(cudaEventRecord(e0, s));
    for (int r = 0; r < reps; ++r) run(x16, y16, a16);
    CUDA_OK(cudaEventRecord(e1, s));
    CUDA_OK(cudaStreamSynchronize(s));
    float ms = 0.f; CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms_out) *ms_out = ms / reps;
    if (getenv("MASG_TC_DEBUG") && (atoi(getenv("MASG_TC_DEBUG")) & 32)) {
        static long long tr[5 * 96 * 4];
        const int nt = one_sm ? respair_debug_trace(tr, 5 * 96 * 4) : respair2_debug_trace(tr, 5 * 96 * 4);
        auto at = [&](int role, int it, int e) { return tr[(role * nt + it) * 4 + e]; };
        const int i0 = 40, i1 = 52;
        const long long base = at(0, i0, 0);
        fprintf(stderr, "  [respair trace, CTA 0, cycles] tile | copy: wait-begin slot-free issued | worker: landed done | issuer: P1-ready P1-committed P2-ready P2-committed | E1: acc1-full h-written | E2: requested acc2-full stored\n");
        for (int it = i0; it < i1; ++it) {
            fprintf(stderr, "   %3d |", it);
            const int ne[5] = {3, 2, 4, one_sm ? 2 : 4, 3};
            for (int role = 0; role < 5; ++role) {
                for (int e = 0; e < ne[role]; ++e) fprintf(stderr, " %7lld", at(role, it, e) ? at(role, it, e) - base : -1LL);
                fprintf(stderr, " |");
            }
            fprintf(stderr, "\n");
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    launch_from_planar(c, s, y16, y, B, T, C);
    CUDA_OK(cudaStreamSynchronize(s));
    while (c->owned.size() > keep) { cudaFree(c->owned.back()); c->owned.pop_back(); }
    API_END(c)
}

int masg_attention(masg_ctx* c, void* stream, const float* qkv, const int64_t* lens, int B, int S, float* out) {
    API_BEGIN(c)
    MASG_CHECK(qkv && lens && out && B > 0 && S > 0, "bad argument");
    if (c->precision == MASG_PREC_BF16) launch_attention_tc(c, (cudaStream_t)stream, qkv, lens, B, S, out);
    else launch_attention_f32(c, (cudaStream_t)stream, qkv, lens, B, S, out);
    API_END(c)
}

int masg_attention_planar(masg_ctx* c, void* stream, const void* qkv16, const int64_t* lens, int B, int S, void* out16) {
    API_BEGIN(c)
    MASG_CHECK(qkv16 && lens && out16 && B > 0 && S > 0, "bad argument");
    MASG_CHECK(c->precision == MASG_PREC_BF16, "masg_attention_planar is the tcgen05 fp16 kernel: MASG_PREC_BF16 contexts only");
    launch_attention_f16p(c, (cudaStream_t)stream, qkv16, lens, B, S, out16);
    API_END(c)
}

int masg_profile(masg_ctx* c, int enable) {
    API_BEGIN(c)
    c->profiling = enable != 0;
    API_END(c)
}
int masg_profile_read(masg_ctx* c, void* stream, double* h_ms_total, double* h_flops_total, int64_t* h_launches) {
    API_BEGIN(c)
    CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    double ms = 0, fl = 0;
    for (auto& r : c->prof) {
        float t = 0.f; CUDA_OK(cudaEventElapsedTime(&t, r.e0, r.e1));
        ms += t; fl += r.flops;
        cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
    }
    if (h_ms_total) *h_ms_total = ms;
    if (h_flops_total) *h_flops_total = fl;
    if (h_launches) *h_launches = (int64_t)c->prof.size();
    c->prof.clear();
    API_END(c)
}

int64_t masg_launch_count(masg_ctx* c, int reset) {
    if (!c) return -1;
    int64_t n = c->launches;
    if (reset) c->launches = 0;
    return n;
}

int masg_debug_enable_taps(masg_ctx* c, int enable) {
    API_BEGIN(c)
    c->taps = enable != 0;
    if (!c->taps) { c->tap_enc = c->tap_lr = c->tap_dec = nullptr; }
    API_END(c)
}

int masg_debug_tap(masg_ctx* c, void* stream, const char* name, float* dst, int64_t n_elems) {
    API_BEGIN(c)
    MASG_CHECK(c->taps, "intermediate taps are off: call masg_debug_enable_taps(ctx, 1) before the forward");
    const float* src = nullptr; int64_t n = 0;
    if (!strcmp(name, "encoder_out")) { src = c->tap_enc; n = (int64_t)c->B * c->L * 256; }
    else if (!strcmp(name, "va_out")) { src = c->va_out; n = (int64_t)c->B * c->L * 256; }
    else if (!strcmp(name, "lr_out")) { src = c->tap_lr; n = (int64_t)c->B * c->tap_T * 256; }
    else if (!strcmp(name, "decoder_out")) { src = c->tap_dec; n = (int64_t)c->B * c->tap_T * 256; }
    MASG_CHECK(src != nullptr, "unknown or unavailable tap '%s'", name);
    MASG_CHECK(n_elems == n, "tap '%s' has %lld elements, caller asked for %lld", name, (long long)n, (long long)n_elems);
    CUDA_OK(cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    API_END(c)
}

}  // extern "C"
