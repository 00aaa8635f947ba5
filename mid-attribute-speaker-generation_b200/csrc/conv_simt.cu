// fp32 CUDA-core implicit-GEMM conv1d / linear / transposed-conv kernel (MASG_PREC_FP32 path: the
// strict-parity mode, mel max-abs <= 1e-4).  Channels-last activations; the same "row GEMM" form as
// the tcgen05 kernel (see ConvDesc in common.cuh):
//   out[b, q*orow + ooff + n] = (act(sum_{j,c} in_act(X[b, q + off0 + j*tstep, c]) W[j*Cin + c, n] + bias[n % bias_mod])
//                                + R1 + R2) * out_scale
// Conv1d(k, dil, pad): taps = k, off0 = -pad, tstep = dil.  ConvTranspose1d(k = 2u, stride u, pad u/2)
// is the 2-tap case off0 = 0, tstep = -1 with N = u*Cout, orow = u*Cout, ooff = -pad*Cout
// (phase decomposition) — no atomics, no col2im.
#include "common.cuh"

#define TM 64
#define TN 64
#define TK 16

__device__ __forceinline__ float act_apply(float v, int act, float slope) {
    switch (act) {
        case ACT_RELU: return fmaxf(v, 0.f);
        case ACT_TANH: return tanhf(v);
        case ACT_LRELU: return v > 0.f ? v : v * slope;
        default: return v;
    }
}

struct ConvArgs {
    const float* X; long x_bs; int x_ld; int Tin;
    const float* W; const float* bias; int bias_mod;
    int Cin, N, taps, off0, tstep;
    float* Y; long y_bs; long y_limit; long orow; long ooff; int Q;
    int in_act; float in_slope; int out_act; float out_slope;
    const float* R1; const float* R2; float out_scale;
};

__global__ void __launch_bounds__(256)
k_conv_f32(ConvArgs a) {
    __shared__ __align__(16) float As[TK][TM + 4];
    __shared__ __align__(16) float Bs[TK][TN];
    int b = blockIdx.z, q0 = blockIdx.x * TM, n0 = blockIdx.y * TN;
    int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const float* Xb = a.X + (long)b * a.x_bs;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    int a_row = tid >> 2, a_c4 = tid & 3;            // A tile: 64 rows x 4 float4
    int b_k = tid >> 4, b_n4 = tid & 15;             // B tile: 16 rows x 16 float4
    for (int j = 0; j < a.taps; ++j) {
        int trow = q0 + a_row + a.off0 + j * a.tstep;
        bool rowok = trow >= 0 && trow < a.Tin;
        const float* xr = Xb + (long)trow * a.x_ld;
        for (int c0 = 0; c0 < a.Cin; c0 += TK) {
            float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rowok) av = *(const float4*)(xr + c0 + a_c4 * 4);
            if (a.in_act) {
                av.x = act_apply(av.x, a.in_act, a.in_slope); av.y = act_apply(av.y, a.in_act, a.in_slope);
                av.z = act_apply(av.z, a.in_act, a.in_slope); av.w = act_apply(av.w, a.in_act, a.in_slope);
            }
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            int ncol = n0 + b_n4 * 4;
            if (ncol < a.N) bv = *(const float4*)(a.W + ((long)j * a.Cin + c0 + b_k) * a.N + ncol);
            __syncthreads();
            As[a_c4 * 4 + 0][a_row] = av.x; As[a_c4 * 4 + 1][a_row] = av.y;
            As[a_c4 * 4 + 2][a_row] = av.z; As[a_c4 * 4 + 3][a_row] = av.w;
            *(float4*)&Bs[b_k][b_n4 * 4] = bv;
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
                float4 x4 = *(const float4*)&As[kk][ty * 4];
                float4 w4 = *(const float4*)&Bs[kk][tx * 4];
                float xa[4] = {x4.x, x4.y, x4.z, x4.w}, wb[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) acc[i][jj] += xa[i] * wb[jj];
            }
        }
    }
    int bias_mod = a.bias_mod;
    float* Yb = a.Y + (long)b * a.y_bs;
    const float* R1b = a.R1 ? a.R1 + (long)b * a.y_bs : nullptr;
    const float* R2b = a.R2 ? a.R2 + (long)b * a.y_bs : nullptr;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int q = q0 + ty * 4 + i;
        if (q >= a.Q) continue;
        int n = n0 + tx * 4;
        if (n >= a.N) continue;
        long e = (long)q * a.orow + a.ooff + n;
        float v[4];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            float bsv = a.bias ? a.bias[(n + jj) % bias_mod] : 0.f;
            v[jj] = act_apply(acc[i][jj] + bsv, a.out_act, a.out_slope);
        }
        bool full = (n + 3 < a.N) && e >= 0 && (e + 3 < a.y_limit) && ((e & 3) == 0);
        if (full) {
            if (R1b) { float4 r = *(const float4*)(R1b + e); v[0] += r.x; v[1] += r.y; v[2] += r.z; v[3] += r.w; }
            if (R2b) { float4 r = *(const float4*)(R2b + e); v[0] += r.x; v[1] += r.y; v[2] += r.z; v[3] += r.w; }
            *(float4*)(Yb + e) = make_float4(v[0] * a.out_scale, v[1] * a.out_scale, v[2] * a.out_scale, v[3] * a.out_scale);
        } else {
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                long ee = e + jj;
                if (n + jj >= a.N || ee < 0 || ee >= a.y_limit) continue;
                float o = v[jj];
                if (R1b) o += R1b[ee];
                if (R2b) o += R2b[ee];
                Yb[ee] = o * a.out_scale;
            }
        }
    }
}

void launch_conv_f32(masg_ctx* c, cudaStream_t s, const ConvDesc& d, const ConvIO& io) {
    MASG_CHECK(d.w32 != nullptr, "conv_f32: fp32 weights not committed");
    MASG_CHECK(!io.x_bf16 && !io.r_bf16 && io.Y16 == nullptr && io.Y != nullptr, "conv_f32: the fp32 path takes fp32 tensors only");
    MASG_CHECK(d.Cin % TK == 0 && d.N % 4 == 0, "conv_f32: Cin=%d must be a multiple of 16 and N=%d of 4", d.Cin, d.N);
    MASG_CHECK(io.x_ld % 4 == 0, "conv_f32: x_ld=%d must be a multiple of 4", io.x_ld);
    if (io.Q <= 0 || io.B <= 0) return;
    ConvArgs a;
    a.X = (const float*)io.X; a.x_bs = io.x_bs; a.x_ld = io.x_ld; a.Tin = io.Tin;
    a.W = d.w32; a.bias = d.bias; a.bias_mod = d.bias_mod ? d.bias_mod : d.N;
    a.Cin = d.Cin; a.N = d.N; a.taps = d.taps; a.off0 = d.off0; a.tstep = d.tstep;
    a.Y = io.Y; a.y_bs = io.y_bs; a.y_limit = io.y_limit; a.orow = d.orow ? d.orow : d.N; a.ooff = d.ooff; a.Q = io.Q;
    a.in_act = io.in_act; a.in_slope = io.in_slope; a.out_act = io.out_act; a.out_slope = io.out_slope;
    a.R1 = (const float*)io.R1; a.R2 = (const float*)io.R2; a.out_scale = io.out_scale;
    dim3 grid((io.Q + TM - 1) / TM, (d.N + TN - 1) / TN, io.B);
    k_conv_f32<<<grid, 256, 0, s>>>(a);
    c->launches++;
}
