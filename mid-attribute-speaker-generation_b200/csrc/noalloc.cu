// masg_gmm_barycenter_noalloc: the BarycenterGMM maths of masg_gmm_barycenter for callers that may not allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The allocating entry point keeps its `rate` vector in the context's growable arena and sums the per-barycentre weights on
// the host in heap vectors.  Here both live in ONE file-scope device array of fixed size, the weights are summed by a
// single-thread kernel in the same (attribute, component) order and with the same unfused fp32 multiply and add as the host
// loop of api.cu, so the two entry points agree bit for bit, and the call never synchronises: no cudaMalloc, no cudaFree, no
// host heap, no stream synchronisation.
//
// Every call on a device shares g_bary_scratch (all contexts, all streams of that device): calls that overlap on the GPU
// race on it.  It keeps the last call's contents; the weight part is cleared at the start of each call.
#include "common.cuh"

#define NOALLOC_RATE_FLOATS 64                                    // rate[A] lives in the first 256 bytes (A <= 8)
#define NOALLOC_SCRATCH_FLOATS (NOALLOC_RATE_FLOATS + MASG_NOALLOC_MAX_BARYCENTRES)

static __device__ __align__(256) float g_bary_scratch[NOALLOC_SCRATCH_FLOATS];   // 16.25 KB per GPU that loads the library, until exit

// In host code `g_bary_scratch` is the host-side stand-in of the symbol, not a device address: the address the kernels get is
// looked up once per device, with that device current, and cached here.
#define NOALLOC_MAX_DEVICES 64
static float* g_scratch_dev[NOALLOC_MAX_DEVICES];                 // zero-initialised: nullptr = not looked up yet on that device

static float* scratch_for(int device) {
    MASG_CHECK(device >= 0 && device < NOALLOC_MAX_DEVICES, "device %d beyond the %d the no-alloc scratch table holds", device,
               NOALLOC_MAX_DEVICES);
    float* p = __atomic_load_n(&g_scratch_dev[device], __ATOMIC_ACQUIRE);
    if (p) return p;
    int cur = -1;
    CUDA_OK(cudaGetDevice(&cur));
    MASG_CHECK(cur == device, "no-alloc scratch lookup for cuda:%d with cuda:%d current", device, cur);
    void* sym = nullptr;
    CUDA_OK(cudaGetSymbolAddress(&sym, g_bary_scratch));
    MASG_CHECK(sym != nullptr, "cudaGetSymbolAddress returned a null address for the no-alloc scratch on cuda:%d", device);
    __atomic_store_n(&g_scratch_dev[device], (float*)sym, __ATOMIC_RELEASE);      // idempotent: a racing thread stores the same value
    return (float*)sym;
}

struct RateArg { float r[8]; };

// rate (by value, so the host vector needs no H2D copy) -> scratch[0..A); zero the weight accumulators scratch[64 .. 64 + nb)
__global__ void k_noalloc_prepare(RateArg rate, int A, int nb, float* __restrict__ scratch) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < A) scratch[i] = rate.r[i];
    if (i < nb) scratch[NOALLOC_RATE_FLOATS + i] = 0.f;
}

// BarycenterGMM._determine_pi accumulation (model/distributions.py:165-192): weight[assign[a, k]] += rate[a] * pi[a, k], in
// (a, k) order by one thread with separately rounded multiply and add — the order and rounding of the host loop in
// masg_gmm_barycenter — then the block copies the nb weights to the caller's buffer.
__global__ void k_noalloc_weights(const float* __restrict__ pi, const int32_t* __restrict__ assign, int AK, int K, int nb,
                                  float* __restrict__ scratch, float* __restrict__ bary_weight) {
    float* w = scratch + NOALLOC_RATE_FLOATS;
    if (threadIdx.x == 0) {
        for (int i = 0; i < AK; ++i) {
            int t = assign[i];
            w[t] = __fadd_rn(w[t], __fmul_rn(scratch[i / K], pi[i]));
        }
    }
    __syncthreads();
    for (int p = threadIdx.x; p < nb; p += blockDim.x) bary_weight[p] = w[p];
}

extern "C" int masg_gmm_barycenter_noalloc(masg_ctx* c, void* stream, const float* pi, const float* mu, const float* sigma,
                                           const float* h_rate, int A, int K, int D, float* bary_mean, float* bary_std,
                                           float* bary_weight, int32_t* assign) {
    if (!c) return 1;
    try {
        int dev = -1;
        CUDA_OK(cudaGetDevice(&dev));
        if (dev != c->device) CUDA_OK(cudaSetDevice(c->device));
        cudaStream_t s = (cudaStream_t)stream;
        MASG_CHECK(pi && mu && sigma && h_rate && bary_mean && bary_std && bary_weight && assign, "null argument");
        MASG_CHECK(A >= 1 && A <= 8 && K >= 1 && D >= 1, "A=%d K=%d D=%d unsupported", A, K, D);
        long nb = 1;
        for (int a = 0; a < A && nb <= MASG_NOALLOC_MAX_BARYCENTRES; ++a) nb *= K;
        MASG_CHECK(nb <= MASG_NOALLOC_MAX_BARYCENTRES,
                   "K^A = %d^%d barycentres exceed MASG_NOALLOC_MAX_BARYCENTRES = %d, the fixed scratch of the no-alloc entry point; "
                   "use masg_gmm_barycenter", K, A, MASG_NOALLOC_MAX_BARYCENTRES);
        float* scratch = scratch_for(c->device);
        RateArg r;
        for (int a = 0; a < 8; ++a) r.r[a] = a < A ? h_rate[a] : 0.f;
        int n0 = (int)(nb > A ? nb : A);
        k_noalloc_prepare<<<(n0 + 255) / 256, 256, 0, s>>>(r, A, (int)nb, scratch);
        c->launches++;
        launch_gmm_bary(c, s, mu, sigma, scratch, A, K, D, bary_mean, bary_std);
        launch_gmm_bary_assign(c, s, mu, sigma, bary_mean, bary_std, A, K, D, (int)nb, assign);
        k_noalloc_weights<<<1, 256, 0, s>>>(pi, assign, A * K, K, (int)nb, scratch, bary_weight);
        c->launches++;
        CUDA_OK(cudaGetLastError());
        return 0;
    } catch (const MasgError& e) { c->err = e.msg; return 1; }
      catch (const std::exception& e) { c->err = e.what(); return 2; }
}
