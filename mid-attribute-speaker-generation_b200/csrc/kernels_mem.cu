// Memory-bound kernels of the hot path: embedding+pos, LayerNorm, bucketize+embedding, duration
// rounding, LengthRegulator (prefix-sum + gather), masks, conv_post+tanh(+int16), GMM prior kernels,
// weight repacking.  All are coalesced 128-bit accesses with warp-shuffle reductions; one warp owns
// one 256-channel row (2 x float4 per lane).
#include "common.cuh"
#include <cuda_fp16.h>
#include <curand_kernel.h>

#define FULL 0xffffffffu

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
static inline int cdiv(long a, long b) { return (int)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------------
// K1: word_emb[ids] + accent_emb[acc] + pos[l]   (transformer/Models.py:101-103)
// ------------------------------------------------------------------------------------------------
__global__ void k_embed(const int64_t* __restrict__ texts, const int64_t* __restrict__ accents, int acc1d,
                        const float4* __restrict__ wemb, const float4* __restrict__ aemb,
                        const float4* __restrict__ pos, int B, int L, int n_vocab, int n_accent,
                        unsigned long long* __restrict__ err, float4* __restrict__ out) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= B * L) return;
    int l = row % L;
    long id = texts[row];
    long ac = acc1d ? accents[l] : accents[row];
    // nn.Embedding raises IndexError on an id outside its table; here the row is read as id 0 and the error word is raised
    // (checked by the host at phase A's synchronisation)
    if (id < 0 || id >= n_vocab) { if (lane == 0 && err) atomicOr(err, 1ull); id = 0; }
    if (ac < 0 || ac >= n_accent) { if (lane == 0 && err) atomicOr(err, 2ull); ac = 0; }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        int c = lane + 32 * h;
        float4 w = wemb[id * 64 + c], a = aemb[ac * 64 + c], p = pos[(long)l * 64 + c];
        float4 o;
        o.x = (w.x + a.x) + p.x; o.y = (w.y + a.y) + p.y; o.z = (w.z + a.z) + p.z; o.w = (w.w + a.w) + p.w;
        out[(long)row * 64 + c] = o;
    }
}
void launch_embed(masg_ctx* c, cudaStream_t s, const int64_t* texts, const int64_t* accents, int acc1d,
                  const float* wemb, const float* aemb, const float* pos, int B, int L, int n_vocab, int n_accent, int64_t* err,
                  float* out) {
    k_embed<<<cdiv((long)B * L, 8), 256, 0, s>>>(texts, accents, acc1d, (const float4*)wemb, (const float4*)aemb,
                                                 (const float4*)pos, B, L, n_vocab, n_accent, (unsigned long long*)err, (float4*)out);
    c->launches++;
}

// x[b,l,:] += v[b,:]   (speaker embedding on ALL positions, model/fastspeech2.py:215-218)
__global__ void k_add_rowvec(float4* __restrict__ x, const float4* __restrict__ v, long rows, int L, int C4) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * C4) return;
    long row = i / C4; int c = (int)(i % C4);
    float4 a = x[i], b = v[(row / L) * C4 + c];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    x[i] = a;
}
void launch_add_rowvec(masg_ctx* c, cudaStream_t s, float* x, const float* v, int B, int L, int C) {
    long n = (long)B * L * (C / 4);
    k_add_rowvec<<<cdiv(n, 256), 256, 0, s>>>((float4*)x, (const float4*)v, (long)B * L, L, C / 4);
    c->launches++;
}
// x[b,t,:] += pos[t,:]   (transformer/Models.py:157-174)
__global__ void k_add_pos(float4* __restrict__ x, const float4* __restrict__ pos, long rows, int T, int C4) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * C4) return;
    long row = i / C4; int c = (int)(i % C4);
    float4 a = x[i], b = pos[(row % T) * C4 + c];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    x[i] = a;
}
void launch_add_pos(masg_ctx* c, cudaStream_t s, float* x, const float* pos, int B, int T, int C) {
    long n = (long)B * T * (C / 4);
    k_add_pos<<<cdiv(n, 256), 256, 0, s>>>((float4*)x, (const float4*)pos, (long)B * T, T, C / 4);
    c->launches++;
}
__global__ void k_add2(const float4* __restrict__ a, const float4* __restrict__ b, float4* __restrict__ o, long n4) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 x = a[i], y = b[i];
    x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
    o[i] = x;
}
void launch_add2(masg_ctx* c, cudaStream_t s, const float* a, const float* b, float* out, int64_t n) {
    k_add2<<<cdiv(n / 4, 256), 256, 0, s>>>((const float4*)a, (const float4*)b, (float4*)out, n / 4);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm over 256 channels (+ residual, + pad-row zeroing): SubLayers.py:55,91 and Layers.py:25,28
// ------------------------------------------------------------------------------------------------
__global__ void k_layernorm256(const float4* __restrict__ x, const float4* __restrict__ res,
                               const float4* __restrict__ g, const float4* __restrict__ b,
                               const int64_t* __restrict__ lens, long rows, int T, float4* __restrict__ y,
                               uint2* __restrict__ y16) {
    long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= rows) return;
    // optional second output: planar fp16 [B][32][T][8] (the tcgen05 operand layout of the next fp16 layer).  Lane l holds
    // channels 4l..4l+3 and 128+4l..: half an 8-channel group each, written as 8-byte pieces
    const long bq = row / T, tq = row % T;
    uint2* p16a = y16 ? y16 + (((bq * 32 + (lane >> 1)) * T + tq) * 2 + (lane & 1)) : nullptr;
    uint2* p16b = y16 ? y16 + (((bq * 32 + 16 + (lane >> 1)) * T + tq) * 2 + (lane & 1)) : nullptr;
    if (lens && (row % T) >= lens[row / T]) {
        float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        y[row * 64 + lane] = z; y[row * 64 + lane + 32] = z;
        if (y16) { *p16a = make_uint2(0u, 0u); *p16b = make_uint2(0u, 0u); }
        return;
    }
    float4 v0 = x[row * 64 + lane], v1 = x[row * 64 + lane + 32];
    if (res) {
        float4 r0 = res[row * 64 + lane], r1 = res[row * 64 + lane + 32];
        v0.x += r0.x; v0.y += r0.y; v0.z += r0.z; v0.w += r0.w;
        v1.x += r1.x; v1.y += r1.y; v1.z += r1.z; v1.w += r1.w;
    }
    float mean = warp_sum(v0.x + v0.y + v0.z + v0.w + v1.x + v1.y + v1.z + v1.w) * (1.f / 256.f);
    float d, q = 0.f;
    d = v0.x - mean; q += d * d; d = v0.y - mean; q += d * d; d = v0.z - mean; q += d * d; d = v0.w - mean; q += d * d;
    d = v1.x - mean; q += d * d; d = v1.y - mean; q += d * d; d = v1.z - mean; q += d * d; d = v1.w - mean; q += d * d;
    float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + 1e-5f);
    float4 g0 = g[lane], g1 = g[lane + 32], b0 = b[lane], b1 = b[lane + 32], o0, o1;
    o0.x = (v0.x - mean) * rstd * g0.x + b0.x; o0.y = (v0.y - mean) * rstd * g0.y + b0.y;
    o0.z = (v0.z - mean) * rstd * g0.z + b0.z; o0.w = (v0.w - mean) * rstd * g0.w + b0.w;
    o1.x = (v1.x - mean) * rstd * g1.x + b1.x; o1.y = (v1.y - mean) * rstd * g1.y + b1.y;
    o1.z = (v1.z - mean) * rstd * g1.z + b1.z; o1.w = (v1.w - mean) * rstd * g1.w + b1.w;
    y[row * 64 + lane] = o0; y[row * 64 + lane + 32] = o1;
    if (y16) {
        __half2 h0 = __floats2half2_rn(o0.x, o0.y), h1 = __floats2half2_rn(o0.z, o0.w);
        __half2 h2 = __floats2half2_rn(o1.x, o1.y), h3 = __floats2half2_rn(o1.z, o1.w);
        *p16a = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
        *p16b = make_uint2(*reinterpret_cast<uint32_t*>(&h2), *reinterpret_cast<uint32_t*>(&h3));
    }
}
void launch_layernorm(masg_ctx* c, cudaStream_t s, const float* x, const float* res, const float* g, const float* b,
                      const int64_t* lens, int B, int T, float* y, bf16* y16_planar) {
    long rows = (long)B * T;
    k_layernorm256<<<cdiv(rows, 8), 256, 0, s>>>((const float4*)x, (const float4*)res, (const float4*)g,
                                                 (const float4*)b, lens, rows, T, (float4*)y, (uint2*)y16_planar);
    c->launches++;
}

// Linear(256 -> 1) + masked_fill(pad, 0)  (model/modules.py:244-248), times `scale` (p_control)
__global__ void k_rowdot_mask(const float4* __restrict__ x, const float4* __restrict__ w, float bias,
                              const int64_t* __restrict__ lens, long rows, int T, float scale, float* __restrict__ out) {
    long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float4 a = x[row * 64 + lane], b = x[row * 64 + lane + 32], wa = w[lane], wb = w[lane + 32];
    float acc = a.x * wa.x + a.y * wa.y + a.z * wa.z + a.w * wa.w + b.x * wb.x + b.y * wb.y + b.z * wb.z + b.w * wb.w;
    acc = warp_sum(acc);
    if (lane == 0) {
        bool pad = lens && (row % T) >= lens[row / T];
        out[row] = pad ? 0.f * scale : (acc + bias) * scale;
    }
}
void launch_rowdot_mask(masg_ctx* c, cudaStream_t s, const float* x, const float* w, float bias, const int64_t* lens,
                        int B, int T, float scale, float* out) {
    long rows = (long)B * T;
    k_rowdot_mask<<<cdiv(rows, 8), 256, 0, s>>>((const float4*)x, (const float4*)w, bias, lens, rows, T, scale, out);
    c->launches++;
}

// get_mask_from_lengths (utils/tools.py:155-163): 1 = pad
__global__ void k_mask(const int64_t* __restrict__ lens, int B, int T, uint8_t* __restrict__ mask) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)B * T) return;
    mask[i] = (i % T) >= lens[i / T] ? 1 : 0;
}
void launch_mask(masg_ctx* c, cudaStream_t s, const int64_t* lens, int B, int T, uint8_t* mask) {
    k_mask<<<cdiv((long)B * T, 256), 256, 0, s>>>(lens, B, T, mask);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// K8: torch.bucketize(v, bins) (right=False → lower bound) + Embedding + add  (model/modules.py:85-88,121)
// ------------------------------------------------------------------------------------------------
__global__ void k_bucket_add(const float* __restrict__ v, const float* __restrict__ bins, int nb,
                             const float4* __restrict__ table, long rows, int C4, float4* __restrict__ x,
                             int64_t* __restrict__ idx_out) {
    long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float val = v[row];
    int lo = 0, hi = nb;                 // first i with bins[i] >= val; NaN compares false everywhere: torch.bucketize sends it to nb
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (!(bins[mid] >= val)) lo = mid + 1; else hi = mid;
    }
    if (idx_out && lane == 0) idx_out[row] = lo;
    for (int c = lane; c < C4; c += 32) {
        float4 a = x[row * C4 + c], e = table[(long)lo * C4 + c];
        a.x += e.x; a.y += e.y; a.z += e.z; a.w += e.w;
        x[row * C4 + c] = a;
    }
}
void launch_bucket_add(masg_ctx* c, cudaStream_t s, const float* v, const float* bins, int nb, const float* table,
                       int64_t rows, int C, float* x, int64_t* idx_out) {
    k_bucket_add<<<cdiv(rows, 8), 256, 0, s>>>(v, bins, nb, (const float4*)table, rows, C / 4, (float4*)x, idx_out);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// K9a: d = clamp(round(exp(log_d) - 1) * d_control, min=0)   (model/modules.py:132-135)
// exp evaluated in double and rounded to fp32 (correctly rounded in all but ~2^-29 of cases) so the
// half-to-even rounding decision does not depend on a 1-2 ulp fp32 expf.
// ------------------------------------------------------------------------------------------------
__global__ void k_duration(const float* __restrict__ log_d, float d_ctl, long n, float* __restrict__ d) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float e = (float)exp((double)log_d[i]);
    float r = rintf(e - 1.0f) * d_ctl;
    d[i] = fmaxf(r, 0.0f);
}
void launch_duration(masg_ctx* c, cudaStream_t s, const float* log_d, float d_ctl, int64_t n, float* d) {
    k_duration<<<cdiv(n, 256), 256, 0, s>>>(log_d, d_ctl, n, d);
    c->launches++;
}

// mel_len[b] = sum_i max(int(d[b,i]), 0); optional running max into *tmax (must be zeroed before)
__global__ void k_mel_lens(const float* __restrict__ d, int B, int L, int64_t* __restrict__ mel_lens,
                           unsigned long long* __restrict__ tmax) {
    int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (b >= B) return;
    long acc = 0;
    for (int i = lane; i < L; i += 32) {
        long r = (long)d[(long)b * L + i];           // truncation toward zero == Python int()
        acc += r > 0 ? r : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
    if (lane == 0) {
        mel_lens[b] = acc;
        if (tmax) atomicMax(tmax, (unsigned long long)acc);
    }
}
void launch_mel_lens(masg_ctx* c, cudaStream_t s, const float* d, int B, int L, int64_t* mel_lens, int64_t* tmax_out) {
    if (tmax_out) CUDA_OK(cudaMemsetAsync(tmax_out, 0, sizeof(int64_t), s));
    k_mel_lens<<<cdiv(B, 8), 256, 0, s>>>(d, B, L, mel_lens, (unsigned long long*)tmax_out);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// K9b: LengthRegulator (model/modules.py:167-194 + utils/tools.py:363-381).
// One block per (32 output frames, utterance): block-wide inclusive prefix sum of the integer repeats
// in shared memory (warp-shuffle scans), then each warp binary-searches the source phoneme of its
// frames and copies the 256-channel row with 2 x float4 per lane.  No host loop, no .item().
// ------------------------------------------------------------------------------------------------
#define LR_ROWS 32
__global__ void __launch_bounds__(256)
k_length_regulate(const float* __restrict__ x, const float* __restrict__ d, int L, int C4, int T,
                  const float4* __restrict__ pos, float4* __restrict__ out, int64_t* __restrict__ mel_lens,
                  uint8_t* __restrict__ mask) {
    extern __shared__ int s_pref[];          // inclusive prefix [L]
    __shared__ int s_part[256];
    int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float* db = d + (long)b * L;
    int per = (L + 255) / 256;
    int beg = tid * per, end = min(beg + per, L);
    int local = 0;
    for (int i = beg; i < end; ++i) { long r = (long)db[i]; local += r > 0 ? (int)r : 0; }
    // block exclusive scan of `local`
    int v = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int n = __shfl_up_sync(FULL, v, o); if (lane >= o) v += n; }
    if (lane == 31) s_part[wid] = v;
    __syncthreads();
    if (wid == 0) {
        int w = lane < 8 ? s_part[lane] : 0;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) { int n = __shfl_up_sync(FULL, w, o); if (lane >= o) w += n; }
        if (lane < 8) s_part[lane] = w;          // inclusive over warps
    }
    __syncthreads();
    int run = v - local + (wid > 0 ? s_part[wid - 1] : 0);
    for (int i = beg; i < end; ++i) { long r = (long)db[i]; run += r > 0 ? (int)r : 0; s_pref[i] = run; }
    __syncthreads();
    int total = L > 0 ? s_pref[L - 1] : 0;
    if (blockIdx.x == 0 && tid == 0 && mel_lens) mel_lens[b] = total;
    int t0 = blockIdx.x * LR_ROWS;
    for (int r = wid; r < LR_ROWS; r += 8) {
        int t = t0 + r;
        if (t >= T) break;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (t < total) {
            int lo = 0, hi = L - 1;              // smallest i with pref[i] > t
            while (lo < hi) { int mid = (lo + hi) >> 1; if (s_pref[mid] > t) hi = mid; else lo = mid + 1; }
            const float4* src = (const float4*)x + ((long)b * L + lo) * C4;
            for (int cc = lane; cc < C4; cc += 32) {
                float4 a = src[cc];
                if (pos) { float4 p = pos[(long)t * C4 + cc]; a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w; }
                out[((long)b * T + t) * C4 + cc] = a;
            }
        } else {
            for (int cc = lane; cc < C4; cc += 32) {
                float4 a = a0;
                if (pos) a = pos[(long)t * C4 + cc];
                out[((long)b * T + t) * C4 + cc] = a;
            }
        }
        if (mask && lane == 0) mask[(long)b * T + t] = t >= total ? 1 : 0;
    }
}
void launch_length_regulate(masg_ctx* c, cudaStream_t s, const float* x, const float* d, int B, int L, int C, int T,
                            const float* pos, float* out, int64_t* mel_lens, uint8_t* mask) {
    if (T <= 0 || B <= 0) {
        if (mel_lens && B > 0) launch_mel_lens(c, s, d, B, L, mel_lens, nullptr);
        return;
    }
    dim3 grid(cdiv(T, LR_ROWS), B);
    size_t smem = (size_t)(L > 0 ? L : 1) * sizeof(int);
    MASG_CHECK(smem <= 160 * 1024, "LengthRegulator: L=%d too long for the shared-memory prefix", L);
    if (smem > 48 * 1024) CUDA_OK(cudaFuncSetAttribute(k_length_regulate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_length_regulate<<<grid, 256, smem, s>>>(x, d, L, C / 4, T, (const float4*)pos, (float4*)out, mel_lens, mask);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// fp32 masked multi-head attention (transformer/SubLayers.py:29-57, Modules.py:14-25), flash-style:
// CTA = 4 warps x 8 queries; K/V streamed through shared memory in 32-key tiles; online softmax with
// warp-shuffle max/sum; scores never materialised.  qkv: [B,S,768] = q|k|v, heads of 128.
// Padded query rows are written as zeros (they are zeroed after the LayerNorm anyway, Layers.py:25).
// ------------------------------------------------------------------------------------------------
#define AT_QW 8
#define AT_Q (4 * AT_QW)
__global__ void __launch_bounds__(128)
k_attention_f32(const float* __restrict__ qkv, const int64_t* __restrict__ lens, int S, float* __restrict__ out) {
    extern __shared__ __align__(16) float at_smem[];
    float (*Qs)[128] = (float (*)[128])at_smem;                       // [AT_Q][128]
    float (*Vs)[128] = (float (*)[128])(at_smem + AT_Q * 128);        // [32][128]
    float (*Ks)[129] = (float (*)[129])(at_smem + AT_Q * 128 + 32 * 128);   // [32][129]
    int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AT_Q;
    int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int len = (int)min((int64_t)S, lens[b]);
    const float* base = qkv + (long)b * S * 768;
    if (q0 >= len) {                         // whole tile is padding
        for (int i = tid; i < AT_Q * 32; i += 128) {
            int r = i >> 5, c4 = i & 31;
            if (q0 + r < S) ((float4*)(out + ((long)b * S + q0 + r) * 256 + h * 128))[c4] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    for (int i = tid; i < AT_Q * 32; i += 128) {
        int r = i >> 5, c4 = i & 31;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q0 + r < S) v = ((const float4*)(base + (long)(q0 + r) * 768 + h * 128))[c4];
        ((float4*)&Qs[r][0])[c4] = v;
    }
    float m[AT_QW], l[AT_QW], o[AT_QW][4];
#pragma unroll
    for (int i = 0; i < AT_QW; ++i) { m[i] = -INFINITY; l[i] = 0.f; o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    const float inv_temp = 1.0f / 11.313708498984761f;     // np.power(128, 0.5)
    for (int j0 = 0; j0 < len; j0 += 32) {
        __syncthreads();
        for (int i = tid; i < 32 * 32; i += 128) {
            int r = i >> 5, c4 = i & 31;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (j0 + r < len) {
                kv = ((const float4*)(base + (long)(j0 + r) * 768 + 256 + h * 128))[c4];
                vv = ((const float4*)(base + (long)(j0 + r) * 768 + 512 + h * 128))[c4];
            }
            Ks[r][c4 * 4 + 0] = kv.x; Ks[r][c4 * 4 + 1] = kv.y; Ks[r][c4 * 4 + 2] = kv.z; Ks[r][c4 * 4 + 3] = kv.w;
            ((float4*)&Vs[r][0])[c4] = vv;
        }
        __syncthreads();
        float sc[AT_QW];
#pragma unroll
        for (int i = 0; i < AT_QW; ++i) sc[i] = 0.f;
        for (int dd = 0; dd < 128; dd += 4) {
            float k0 = Ks[lane][dd], k1 = Ks[lane][dd + 1], k2 = Ks[lane][dd + 2], k3 = Ks[lane][dd + 3];
#pragma unroll
            for (int i = 0; i < AT_QW; ++i) {
                float4 qv = *(const float4*)&Qs[wid * AT_QW + i][dd];
                sc[i] += qv.x * k0 + qv.y * k1 + qv.z * k2 + qv.w * k3;
            }
        }
        bool valid = (j0 + lane) < len;
        float p[AT_QW];
#pragma unroll
        for (int i = 0; i < AT_QW; ++i) {
            float sv = valid ? sc[i] * inv_temp : -INFINITY;
            float mx = sv;
#pragma unroll
            for (int of = 16; of > 0; of >>= 1) mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, of));
            float mnew = fmaxf(m[i], mx);
            float corr = expf(m[i] - mnew);
            p[i] = valid ? expf(sv - mnew) : 0.f;
            l[i] = l[i] * corr + warp_sum(p[i]);
            o[i][0] *= corr; o[i][1] *= corr; o[i][2] *= corr; o[i][3] *= corr;
            m[i] = mnew;
        }
        for (int j = 0; j < 32; ++j) {
            float4 vv = *(const float4*)&Vs[j][lane * 4];
#pragma unroll
            for (int i = 0; i < AT_QW; ++i) {
                float pj = __shfl_sync(FULL, p[i], j);
                o[i][0] += pj * vv.x; o[i][1] += pj * vv.y; o[i][2] += pj * vv.z; o[i][3] += pj * vv.w;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < AT_QW; ++i) {
        int q = q0 + wid * AT_QW + i;
        if (q >= S) continue;
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < len) { float inv = 1.f / l[i]; r = make_float4(o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv); }
        ((float4*)(out + ((long)b * S + q) * 256 + h * 128))[lane] = r;
    }
}
void launch_attention_f32(masg_ctx* c, cudaStream_t s, const float* qkv, const int64_t* lens, int B, int S, float* out) {
    dim3 grid(cdiv(S, AT_Q), 2, B);
    const int smem = (AT_Q * 128 + 32 * 128 + 32 * 129) * sizeof(float);
    static PerDeviceOnce once;
    if (once.need()) CUDA_OK(cudaFuncSetAttribute(k_attention_f32, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    k_attention_f32<<<grid, 128, smem, s>>>(qkv, lens, S, out);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// mel [B,C,T] → [B,T,C]  (the transpose utils/tools.py:266 undoes; smem-tiled, coalesced both ways)
// ------------------------------------------------------------------------------------------------
__global__ void k_transpose_bct(const float* __restrict__ in, int C, int T, float* __restrict__ out) {
    __shared__ float tile[32][33];
    int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    int tx = threadIdx.x, ty = threadIdx.y;
    for (int i = ty; i < 32; i += 8) {
        int cc = c0 + i, t = t0 + tx;
        tile[i][tx] = (cc < C && t < T) ? in[((long)b * C + cc) * T + t] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        int t = t0 + i, cc = c0 + tx;
        if (t < T && cc < C) out[((long)b * T + t) * C + cc] = tile[tx][i];
    }
}
void launch_transpose_mel(masg_ctx* c, cudaStream_t s, const float* mel_bct, int B, int C, int T, float* out_btc) {
    dim3 grid(cdiv(T, 32), cdiv(C, 32), B), blk(32, 8);
    k_transpose_bct<<<grid, blk, 0, s>>>(mel_bct, C, T, out_btc);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// K16: leaky_relu(0.01) → conv_post (32→1, k7, p3) → tanh (+ int16)   (hifigan/models.py:161-163,
// utils/model.py:82-85).  256 samples per CTA; the (256+6)x32 input window is staged once in shared
// memory (stride 33: conflict-free) with the activation applied on the way in.
// ------------------------------------------------------------------------------------------------
template <bool XBF>
__global__ void __launch_bounds__(256)
k_conv_post(const void* __restrict__ xv, const float* __restrict__ w, float bias, long S, float slope,
            float* __restrict__ wav, int16_t* __restrict__ wav16, const int64_t* __restrict__ row_len, const int64_t* __restrict__ row_off) {
    constexpr int XS = 36;                                   // row stride in floats: 16-byte aligned rows, conflict-free 128-bit accesses
    __shared__ __align__(16) float xs[262 * XS];
    __shared__ __align__(16) float ws[7 * 32];
    int b = blockIdx.y; long s0 = (long)blockIdx.x * 256;
    const long n_valid = row_len ? min((long)row_len[b], S) : S;          // vocoder_infer's trim (utils/model.py:88-90), fused
    if (s0 >= n_valid) return;
    const long obase = row_off ? (long)row_off[b] : (long)b * S;
    int tid = threadIdx.x;
    if (tid < 224) ws[tid] = w[tid];
    if (XBF) {
        // planar bf16 input [B][4][S][8]: one 16-byte piece = 8 channels of one sample.  All of a thread's loads are issued
        // before the first is used (5 x 16 bytes in flight per thread instead of 1)
        constexpr int NP = (262 * 4 + 255) / 256;
        uint4 u[NP];
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int i = tid + k * 256;
            const int g = i / 262, r = i % 262;
            const long srow = s0 + r - 3;
            u[k] = make_uint4(0u, 0u, 0u, 0u);
            if (i < 262 * 4 && srow >= 0 && srow < S)
                u[k] = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(xv) + (((long)b * 4 + g) * S + srow) * 8);
        }
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int i = tid + k * 256;
            if (i >= 262 * 4) break;
            const int g = i / 262, r = i % 262;
            const uint32_t w4[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
            float f[8];
#pragma unroll
            for (int q = 0; q < 4; ++q) { f[2 * q] = __uint_as_float(w4[q] << 16); f[2 * q + 1] = __uint_as_float(w4[q] & 0xffff0000u); }
#pragma unroll
            for (int q = 0; q < 8; ++q) f[q] = f[q] > 0.f ? f[q] : f[q] * slope;
            float4* dst = reinterpret_cast<float4*>(&xs[r * XS + g * 8]);
            dst[0] = make_float4(f[0], f[1], f[2], f[3]); dst[1] = make_float4(f[4], f[5], f[6], f[7]);
        }
    } else {
        for (int i = tid; i < 262 * 8; i += 256) {
            int r = i >> 3, c4 = i & 7;
            long srow = s0 + r - 3;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (srow >= 0 && srow < S) v = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(xv) + ((long)b * S + srow) * 32)[c4];
            v.x = v.x > 0.f ? v.x : v.x * slope; v.y = v.y > 0.f ? v.y : v.y * slope;
            v.z = v.z > 0.f ? v.z : v.z * slope; v.w = v.w > 0.f ? v.w : v.w * slope;
            *reinterpret_cast<float4*>(&xs[r * XS + c4 * 4]) = v;
        }
    }
    __syncthreads();
    long sidx = s0 + tid;
    if (sidx >= n_valid) return;
    float acc = bias;
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        const float4* row = reinterpret_cast<const float4*>(&xs[(tid + j) * XS]);
        const float4* wr = reinterpret_cast<const float4*>(&ws[j * 32]);
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {                       // same summation order as before: channels ascending within a tap
            const float4 xv4 = row[c4], wv = wr[c4];
            acc += xv4.x * wv.x; acc += xv4.y * wv.y; acc += xv4.z * wv.z; acc += xv4.w * wv.w;
        }
    }
    float y = tanhf(acc);
    if (wav) wav[obase + sidx] = y;
    if (wav16) wav16[obase + sidx] = (int16_t)__float2int_rz(y * 32768.0f);
}
// Planar bf16 input (the bf16 graph): thread = (8-channel group g, two runs of 4 consecutive samples, 256 samples apart).  The CTA
// stages its 512 + 6 rows as raw bf16 pieces (16 bytes = 8 channels of one sample, plane pitch 521 pieces so a quarter-warp's
// 128-bit reads cover all banks), reads 2 x 10 pieces per thread (1/2.8 of the shared-memory wavefronts of the one-sample-per-thread
// form, which ran at the LDS limit) and the 4 group partials meet in two shuffles.  9 x 16-byte global loads per thread are in
// flight before the first is used (33 KB per CTA); the 7 x 32 weights sit in shared memory (same-address broadcast within a
// group) so that three CTAs (100 KB of loads in flight) fit an SM.
// Output addressing (vocoder_infer's trim, utils/model.py:88-90, fused): row b is written at wav16 + row_off[b] (or b * S) and only
// its first row_len[b] samples (or S) are stored — with offsets = prefix sums of the lengths this is the packed layout.
template <bool ACT>
__global__ void __launch_bounds__(256, 3)
k_conv_post_planar(const bf16* __restrict__ x, const float* __restrict__ w, float bias, long S, float slope,
                   float* __restrict__ wav, int16_t* __restrict__ wav16, const int64_t* __restrict__ row_len,
                   const int64_t* __restrict__ row_off) {
    constexpr int TILE = 512, ROWS = TILE + 6, PP = 521;     // plane pitch in 16-byte pieces (== 1 mod 8)
    __shared__ uint4 xs[4 * PP];
    __shared__ __align__(16) float ws[7 * 32];
    const int b = blockIdx.y; const long s0 = (long)blockIdx.x * TILE;
    const long n_valid = row_len ? min((long)row_len[b], S) : S;
    if (s0 >= n_valid) return;                               // the whole tile lies in the trimmed tail
    const int tid = threadIdx.x;
    if (tid < 224) ws[tid] = w[tid];
    {
        constexpr int NP = (ROWS * 4 + 255) / 256;
        uint4 u[NP];
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int i = tid + k * 256;
            const int g = i / ROWS, r = i % ROWS;
            const long srow = s0 + r - 3;
            u[k] = make_uint4(0u, 0u, 0u, 0u);
            if (i < ROWS * 4 && srow >= 0 && srow < S) u[k] = *reinterpret_cast<const uint4*>(x + (((long)b * 4 + g) * S + srow) * 8);
        }
#pragma unroll
        for (int k = 0; k < NP; ++k) {
            const int i = tid + k * 256;
            if (i < ROWS * 4) xs[(i / ROWS) * PP + i % ROWS] = u[k];
        }
    }
    const int g = tid & 3, og = tid >> 2;
    __syncthreads();
    const long obase = row_off ? (long)row_off[b] : (long)b * S;
#pragma unroll
    for (int run = 0; run < 2; ++run) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int r0 = run * 256 + og * 4;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const uint4 u = xs[g * PP + r0 + r];
            const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
            float f[8];
#pragma unroll
            for (int q = 0; q < 4; ++q) { f[2 * q] = __uint_as_float(w4[q] << 16); f[2 * q + 1] = __uint_as_float(w4[q] & 0xffff0000u); }
#pragma unroll
            for (int q = 0; q < 8; ++q) if (ACT) f[q] = fmaxf(f[q], f[q] * slope);    // leaky-relu, 0 < slope < 1
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int j = r - o;
                if (j >= 0 && j < 7) {
                    const float4 wa = *reinterpret_cast<const float4*>(&ws[j * 32 + g * 8]), wb = *reinterpret_cast<const float4*>(&ws[j * 32 + g * 8 + 4]);
                    acc[o] += f[0] * wa.x; acc[o] += f[1] * wa.y; acc[o] += f[2] * wa.z; acc[o] += f[3] * wa.w;
                    acc[o] += f[4] * wb.x; acc[o] += f[5] * wb.y; acc[o] += f[6] * wb.z; acc[o] += f[7] * wb.w;
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], 1);
            acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], 2);
        }
        // lane g writes sample r0 + g
        const float mine = g == 0 ? acc[0] : g == 1 ? acc[1] : g == 2 ? acc[2] : acc[3];
        const long sidx = s0 + r0 + g;
        if (sidx < n_valid) {
            const float y = tanhf(mine + bias);
            if (wav) wav[obase + sidx] = y;
            if (wav16) wav16[obase + sidx] = (int16_t)__float2int_rz(y * 32768.0f);
        }
    }
}
// packed-output offsets: off[b] = hop * sum_{i<b} max(mel_lens[i], 0), off[B] = total; len[b] = hop * mel_lens[b]   (one block)
__global__ void k_wav_offsets(const int64_t* __restrict__ mel_lens, int B, int hop, int64_t* __restrict__ off, int64_t* __restrict__ len) {
    __shared__ long part[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    long carry = 0;
    for (int base = 0; base < B; base += blockDim.x) {
        const int i = base + tid;
        const long v = i < B ? max((long)mel_lens[i], 0L) * hop : 0;
        long sc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const long n = __shfl_up_sync(FULL, sc, o); if (lane >= o) sc += n; }
        if (lane == 31) part[wid] = sc;
        __syncthreads();
        if (wid == 0) {
            long p = lane < (int)(blockDim.x >> 5) ? part[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const long n = __shfl_up_sync(FULL, p, o); if (lane >= o) p += n; }
            part[lane] = p;
        }
        __syncthreads();
        const long incl = carry + sc + (wid > 0 ? part[wid - 1] : 0);
        if (i < B) { off[i] = incl - v; if (len) len[i] = v; }
        carry += part[(blockDim.x >> 5) - 1];
        __syncthreads();
    }
    if (tid == 0) off[B] = carry;
}
void launch_wav_offsets(masg_ctx* c, cudaStream_t s, const int64_t* mel_lens, int B, int hop, int64_t* off, int64_t* len) {
    k_wav_offsets<<<1, 256, 0, s>>>(mel_lens, B, hop, off, len);
    c->launches++;
}
void launch_conv_post(masg_ctx* c, cudaStream_t s, const void* x, int x_bf16, const float* w, float bias, int B, long S,
                      float in_slope, float* wav, int16_t* wav16, const int64_t* row_len, const int64_t* row_off) {
    if (x_bf16 && in_slope > 0.f && in_slope <= 1.f) {
        dim3 grid(cdiv(S, 512), B);
        if (in_slope == 1.f) k_conv_post_planar<false><<<grid, 256, 0, s>>>((const bf16*)x, w, bias, S, in_slope, wav, wav16, row_len, row_off);
        else k_conv_post_planar<true><<<grid, 256, 0, s>>>((const bf16*)x, w, bias, S, in_slope, wav, wav16, row_len, row_off);
    } else {
        dim3 grid(cdiv(S, 256), B);
        if (x_bf16) k_conv_post<true><<<grid, 256, 0, s>>>(x, w, bias, S, in_slope, wav, wav16, row_len, row_off);
        else k_conv_post<false><<<grid, 256, 0, s>>>(x, w, bias, S, in_slope, wav, wav16, row_len, row_off);
    }
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// K17: SpeakerMetaEncoder (model/fastspeech2.py:332-341) and MixtureSameFamily.sample
// ------------------------------------------------------------------------------------------------
__global__ void k_gmm_params(const float* __restrict__ meta, int N, const float* __restrict__ pw, const float* __restrict__ pb,
                             const float* __restrict__ sw, const float* __restrict__ sb, const float* __restrict__ mw,
                             const float* __restrict__ mb, float* __restrict__ pi, float* __restrict__ mu,
                             float* __restrict__ sigma) {
    int n = blockIdx.x, tid = threadIdx.x;
    float m0 = meta[n * 4], m1 = meta[n * 4 + 1], m2 = meta[n * 4 + 2], m3 = meta[n * 4 + 3];
    for (int j = tid; j < 768; j += blockDim.x) {
        float a = ((sw[j * 4] * m0 + sw[j * 4 + 1] * m1) + sw[j * 4 + 2] * m2) + sw[j * 4 + 3] * m3 + sb[j];
        sigma[(long)n * 768 + j] = a > 20.f ? a : log1pf(expf(a));           // nn.Softplus(beta=1, threshold=20)
        float u = ((mw[j * 4] * m0 + mw[j * 4 + 1] * m1) + mw[j * 4 + 2] * m2) + mw[j * 4 + 3] * m3 + mb[j];
        mu[(long)n * 768 + j] = u;
    }
    if (tid == 0) {
        float z[3], mx = -INFINITY, sum = 0.f;
        for (int k = 0; k < 3; ++k) {
            z[k] = ((pw[k * 4] * m0 + pw[k * 4 + 1] * m1) + pw[k * 4 + 2] * m2) + pw[k * 4 + 3] * m3 + pb[k];
            mx = fmaxf(mx, z[k]);
        }
        for (int k = 0; k < 3; ++k) { z[k] = expf(z[k] - mx); sum += z[k]; }
        for (int k = 0; k < 3; ++k) pi[n * 3 + k] = z[k] / sum;
    }
}
void launch_gmm_params(masg_ctx* c, cudaStream_t s, const float* meta, int N, float* pi, float* mu, float* sigma) {
    k_gmm_params<<<N, 256, 0, s>>>(meta, N, c->gmm_pi_w, c->gmm_pi_b, c->gmm_sig_w, c->gmm_sig_b, c->gmm_mu_w, c->gmm_mu_b,
                                   pi, mu, sigma);
    c->launches++;
}

// one CTA (D/4 threads) per sampled speaker; Philox4x32-10 counter RNG: subsequence 2n for the categorical
// draw, subsequence (2n+1)*64 + thread for the 4 normals of each thread; `offset` advances both.
__global__ void k_gmm_sample(const float* __restrict__ pi, const float* __restrict__ mu, const float* __restrict__ scale,
                             int Np, int K, int D, unsigned long long seed, unsigned long long offset,
                             const int32_t* __restrict__ kin, const float* __restrict__ ein, float* __restrict__ out,
                             int32_t* __restrict__ kout) {
    __shared__ int s_k;
    int n = blockIdx.x, tid = threadIdx.x;
    int np = Np == 1 ? 0 : n;
    if (tid == 0) {
        int k;
        if (kin) k = kin[n];
        else {
            curandStatePhilox4_32_10_t st;
            curand_init(seed, 2ull * n, offset, &st);
            float u = curand_uniform(&st);                      // (0, 1]
            float tot = 0.f;
            for (int i = 0; i < K; ++i) tot += pi[np * K + i];
            float acc = 0.f; k = K - 1;
            for (int i = 0; i < K; ++i) { acc += pi[np * K + i]; if (u * tot <= acc) { k = i; break; } }
        }
        s_k = k;
        if (kout) kout[n] = k;
    }
    __syncthreads();
    int k = s_k;
    int d0 = tid * 4;
    if (d0 >= D) return;
    float4 e;
    if (ein) e = *(const float4*)(ein + (long)n * D + d0);
    else {
        curandStatePhilox4_32_10_t st;
        curand_init(seed, (2ull * n + 1ull) * 64ull + tid, offset, &st);
        e = curand_normal4(&st);
    }
    const float* m = mu + ((long)np * K + k) * D + d0;
    const float* sc = scale + ((long)np * K + k) * D + d0;
    float4 r;
    // separate multiply and add (no FMA contraction): bit-identical to loc + scale * eps evaluated by torch
    r.x = __fadd_rn(m[0], __fmul_rn(sc[0], e.x)); r.y = __fadd_rn(m[1], __fmul_rn(sc[1], e.y));
    r.z = __fadd_rn(m[2], __fmul_rn(sc[2], e.z)); r.w = __fadd_rn(m[3], __fmul_rn(sc[3], e.w));
    *(float4*)(out + (long)n * D + d0) = r;
}
void launch_gmm_sample(masg_ctx* c, cudaStream_t s, const float* pi, const float* mu, const float* scale, int N, int Np,
                       int K, int D, uint64_t seed, uint64_t offset, const int32_t* kin, const float* ein, float* out,
                       int32_t* kout) {
    MASG_CHECK(D % 4 == 0 && D / 4 <= 64, "gmm_sample: D=%d unsupported", D);
    k_gmm_sample<<<N, 64, 0, s>>>(pi, mu, scale, Np, K, D, seed, offset, kin, ein, out, kout);
    c->launches++;
}

// InterpolateGMM._w2sq as written (model/distributions.py:64-77), on diagonals, in double:
// ||mu_a - mu_b||^2 + sum(va + vb - 2 * sa * sqrt(sa * vb * sa) * sa),  sa = sqrt(va)
__global__ void k_gmm_interp_cost(const float* __restrict__ mu_a, const float* __restrict__ sig_a, const float* __restrict__ mu_b,
                                  const float* __restrict__ sig_b, int K, int D, double* __restrict__ cost) {
    int i = blockIdx.x / K, j = blockIdx.x % K, tid = threadIdx.x;
    double acc_mu = 0.0, acc_tr = 0.0;
    for (int d = tid; d < D; d += blockDim.x) {
        float va_f = sig_a[i * D + d] * sig_a[i * D + d], vb_f = sig_b[j * D + d] * sig_b[j * D + d];   // Normal.variance = scale^2 (fp32)
        double va = va_f, vb = vb_f, sa = sqrt(va);
        double inner = sqrt(sa * vb * sa);
        acc_tr += va + vb + (-2.0 * sa * inner * sa);
        float dm = mu_a[i * D + d] - mu_b[j * D + d];
        acc_mu += (double)dm * (double)dm;
    }
    __shared__ double sm[2][8];
    acc_mu = warp_sum_d(acc_mu); acc_tr = warp_sum_d(acc_tr);
    if ((tid & 31) == 0) { sm[0][tid >> 5] = acc_mu; sm[1][tid >> 5] = acc_tr; }
    __syncthreads();
    if (tid == 0) {
        double a = 0, b = 0;
        for (int w = 0; w < (blockDim.x >> 5); ++w) { a += sm[0][w]; b += sm[1][w]; }
        cost[i * K + j] = a + b;
    }
}
void launch_gmm_interp_cost(masg_ctx* c, cudaStream_t s, const float* mu_a, const float* sig_a, const float* mu_b,
                            const float* sig_b, int K, int D, double* cost) {
    k_gmm_interp_cost<<<K * K, 256, 0, s>>>(mu_a, sig_a, mu_b, sig_b, K, D, cost);
    c->launches++;
}
// components in the reference's (j outer, i inner) order: mu = (1-t) mu_a[i] + t mu_b[j];
// "scale" = S0^-1/2 ((1-t) S0 + t (S0^1/2 S1 S0^1/2)^1/2)^2 S0^-1/2 on the diagonal (distributions.py:45-62)
__global__ void k_gmm_interp_comps(const float* __restrict__ mu_a, const float* __restrict__ sig_a, const float* __restrict__ mu_b,
                                   const float* __restrict__ sig_b, int K, int D, float t, float* __restrict__ cmu,
                                   float* __restrict__ csc) {
    int comp = blockIdx.x, j = comp / K, i = comp % K;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        cmu[comp * D + d] = (1.f - t) * mu_a[i * D + d] + t * mu_b[j * D + d];
        float va_f = sig_a[i * D + d] * sig_a[i * D + d], vb_f = sig_b[j * D + d] * sig_b[j * D + d];
        double va = va_f, vb = vb_f, s0 = sqrt(va);
        double inner = sqrt(s0 * vb * s0);
        double mid = (1.0 - (double)t) * va + (double)t * inner;
        csc[comp * D + d] = (float)((1.0 / s0) * (mid * mid) * (1.0 / s0));
    }
}
void launch_gmm_interp_comps(masg_ctx* c, cudaStream_t s, const float* mu_a, const float* sig_a, const float* mu_b,
                             const float* sig_b, int K, int D, float t, float* comp_mu, float* comp_scale) {
    k_gmm_interp_comps<<<K * K, 256, 0, s>>>(mu_a, sig_a, mu_b, sig_b, K, D, t, comp_mu, comp_scale);
    c->launches++;
}
// BarycenterGMM._one_barycenter_gaussian (distributions.py:144-163): all K^A index tuples (itertools.product order)
__global__ void k_gmm_bary(const float* __restrict__ mu, const float* __restrict__ sigma, const float* __restrict__ rate,
                           int A, int K, int D, float* __restrict__ bmean, float* __restrict__ bstd) {
    int pos = blockIdx.x;
    int idx[8]; int p = pos;
    for (int a = A - 1; a >= 0; --a) { idx[a] = p % K; p /= K; }
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float mean = 0.f;
        float orig[8];
        for (int a = 0; a < A; ++a) {
            mean += rate[a] * mu[((long)a * K + idx[a]) * D + d];
            orig[a] = sigma[((long)a * K + idx[a]) * D + d];
        }
        float sd = orig[0];
        for (int it = 0; it < 60; ++it) {
            float acc = 0.f;
            for (int a = 0; a < A; ++a) acc += (rate[a] * sd) * orig[a];
            sd = (1.f / sd) * acc;
        }
        bmean[(long)pos * D + d] = mean;
        bstd[(long)pos * D + d] = sd;
    }
}
void launch_gmm_bary(masg_ctx* c, cudaStream_t s, const float* mu, const float* sigma, const float* rate, int A, int K,
                     int D, float* bmean, float* bstd) {
    int nb = 1; for (int a = 0; a < A; ++a) nb *= K;
    k_gmm_bary<<<nb, 256, 0, s>>>(mu, sigma, rate, A, K, D, bmean, bstd);
    c->launches++;
}
// BarycenterGMM._determine_pi nearest-barycentre search (distributions.py:165-192): standard diagonal W2^2
__global__ void k_gmm_bary_assign(const float* __restrict__ mu, const float* __restrict__ sigma, const float* __restrict__ bmean,
                                  const float* __restrict__ bstd, int D, int nb, int32_t* __restrict__ assign) {
    int comp = blockIdx.x, tid = threadIdx.x;
    __shared__ float sm[8];
    __shared__ float s_best; __shared__ int s_arg;
    if (tid == 0) { s_best = INFINITY; s_arg = 0; }
    for (int p = 0; p < nb; ++p) {
        float acc = 0.f;
        for (int d = tid; d < D; d += blockDim.x) {
            float dm = mu[(long)comp * D + d] - bmean[(long)p * D + d];
            float ds = sigma[(long)comp * D + d] - bstd[(long)p * D + d];
            acc += dm * dm + ds * ds;
        }
        acc = warp_sum(acc);
        __syncthreads();
        if ((tid & 31) == 0) sm[tid >> 5] = acc;
        __syncthreads();
        if (tid == 0) {
            float tot = 0.f;
            for (int w = 0; w < (blockDim.x >> 5); ++w) tot += sm[w];
            if (tot < s_best) { s_best = tot; s_arg = p; }
        }
    }
    __syncthreads();
    if (tid == 0) assign[comp] = s_arg;
}
void launch_gmm_bary_assign(masg_ctx* c, cudaStream_t s, const float* mu, const float* sigma, const float* bmean,
                            const float* bstd, int A, int K, int D, int nb, int32_t* assign) {
    k_gmm_bary_assign<<<A * K, 256, 0, s>>>(mu, sigma, bmean, bstd, D, nb, assign);
    c->launches++;
}

// fp32 <-> bf16 tensor conversion (diagnostic hook masg_conv1d with bf16 activations)
__global__ void k_f32_to_bf16(const float* __restrict__ in, bf16* __restrict__ out, long n) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2bfloat16_rn(in[i]);
}
__global__ void k_bf16_to_f32(const bf16* __restrict__ in, float* __restrict__ out, long n) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __bfloat162float(in[i]);
}
void launch_f32_to_bf16(masg_ctx* c, cudaStream_t s, const float* in, bf16* out, long n) {
    k_f32_to_bf16<<<cdiv(n, 256), 256, 0, s>>>(in, out, n); c->launches++;
}
void launch_bf16_to_f32(masg_ctx* c, cudaStream_t s, const bf16* in, float* out, long n) {
    k_bf16_to_f32<<<cdiv(n, 256), 256, 0, s>>>(in, out, n); c->launches++;
}

// ------------------------------------------------------------------------------------------------
// weight repacking (load time only)
// ------------------------------------------------------------------------------------------------
// Conv1d weight [Cout][Cin][k] (* per-Cout scale) → GEMM B matrix rows (j*Cin + c), column col0 + co
__global__ void k_repack_conv_w(const float* __restrict__ w, int Cout, int Cin, int k, const float* __restrict__ scale,
                                float* __restrict__ out, int ldn, int col0) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long n = (long)Cout * Cin * k;
    if (i >= n) return;
    int co = (int)(i % Cout); long r = i / Cout; int ci = (int)(r % Cin); int j = (int)(r / Cin);
    float v = w[((long)co * Cin + ci) * k + j];
    if (scale) v *= scale[co];
    out[((long)j * Cin + ci) * ldn + col0 + co] = v;
}
void launch_repack_conv_w(masg_ctx* c, cudaStream_t s, const float* w, int Cout, int Cin, int k, const float* scale,
                          float* out, int ldn, int col0) {
    long n = (long)Cout * Cin * k;
    k_repack_conv_w<<<cdiv(n, 256), 256, 0, s>>>(w, Cout, Cin, k, scale, out, ldn, col0);
    c->launches++;
}
// ConvTranspose1d weight [Cin][Cout][k], k = 2u, stride u, padding u/2  →  B matrix [2*Cin][u*Cout]:
// row (tap*Cin + ci), column (r*Cout + co) = w[ci][co][r + tap*u]   (tap 0 reads x[q], tap 1 reads x[q-1])
__global__ void k_repack_convT_w(const float* __restrict__ w, int Cin, int Cout, int k, int u, float* __restrict__ out) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long n = (long)2 * Cin * u * Cout;
    if (i >= n) return;
    int col = (int)(i % ((long)u * Cout)); long row = i / ((long)u * Cout);
    int co = col % Cout, r = col / Cout; int ci = (int)(row % Cin), tap = (int)(row / Cin);
    out[i] = w[((long)ci * Cout + co) * k + r + tap * u];
}
void launch_repack_convT_w(masg_ctx* c, cudaStream_t s, const float* w, int Cin, int Cout, int k, int u, float* out) {
    long n = (long)2 * Cin * u * Cout;
    k_repack_convT_w<<<cdiv(n, 256), 256, 0, s>>>(w, Cin, Cout, k, u, out);
    c->launches++;
}
// weight_norm fold: w[n0][inner] = v * g[n0] / ||v[n0,:]||   (hifigan/models.py:167-174)
__global__ void k_weight_norm_fold(const float* __restrict__ v, const float* __restrict__ g, int inner, float* __restrict__ w) {
    int r = blockIdx.x, tid = threadIdx.x;
    __shared__ float sm[8]; __shared__ float s_scale;
    float acc = 0.f;
    for (int i = tid; i < inner; i += blockDim.x) { float a = v[(long)r * inner + i]; acc += a * a; }
    acc = warp_sum(acc);
    if ((tid & 31) == 0) sm[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) { float t = 0.f; for (int q = 0; q < (blockDim.x >> 5); ++q) t += sm[q]; s_scale = g[r] / sqrtf(t); }
    __syncthreads();
    float sc = s_scale;
    for (int i = tid; i < inner; i += blockDim.x) w[(long)r * inner + i] = v[(long)r * inner + i] * sc;
}
void launch_weight_norm_fold(masg_ctx* c, cudaStream_t s, const float* v, const float* g, int n0, int inner, float* w) {
    k_weight_norm_fold<<<n0, 256, 0, s>>>(v, g, inner, w);
    c->launches++;
}
// BatchNorm1d(eval) fold into the preceding conv: y = (conv + cb - mean) * gamma / sqrt(var + 1e-5) + beta
__global__ void k_bn_fold(const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                          const float* __restrict__ var, const float* __restrict__ cb, int C, float* __restrict__ scale,
                          float* __restrict__ bias) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= C) return;
    float sc = gamma[i] / sqrtf(var[i] + 1e-5f);
    scale[i] = sc;
    bias[i] = (cb[i] - mean[i]) * sc + beta[i];
}
void launch_bn_fold(masg_ctx* c, cudaStream_t s, const float* gamma, const float* beta, const float* mean, const float* var,
                    const float* conv_bias, int C, float* scale, float* bias) {
    k_bn_fold<<<cdiv(C, 256), 256, 0, s>>>(gamma, beta, mean, var, conv_bias, C, scale, bias);
    c->launches++;
}

// ------------------------------------------------------------------------------------------------
// sinusoid position table for lengths beyond max_seq_len (transformer/Models.py:10-30, rebuilt per call
// by the reference at :86-98 / :157-164): angle = pos / 10000^(2*(j/2)/256) in double, sin on even and
// cos on odd columns, cast to fp32.
// ------------------------------------------------------------------------------------------------
__global__ void k_pos_table(int n, float* __restrict__ out) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)n * 256) return;
    int j = (int)(i & 255); long pos = i >> 8;
    double denom = pow(10000.0, (double)(2 * (j / 2)) / 256.0);
    double ang = (double)pos / denom;
    out[i] = (float)((j & 1) ? cos(ang) : sin(ang));
}
void launch_pos_table(masg_ctx* c, cudaStream_t s, int n, float* out) {
    k_pos_table<<<cdiv((long)n * 256, 256), 256, 0, s>>>(n, out);
    c->launches++;
}
