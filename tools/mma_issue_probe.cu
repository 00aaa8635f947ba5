// Micro-benchmark: how fast can tcgen05.mma (M=128, N=32, K=16, bf16) be ISSUED, by issue pattern?
//   A  one thread (threadIdx.x == 0 branch), descriptors advanced by running adds, 8 MMAs unrolled per step
//   B  converged warp, elect.sync around every MMA (the CUTLASS atom pattern)
//   C  converged warp, one elect.sync branch around a group of 8 MMAs
//   D  like A but W warps issue concurrently into disjoint accumulators (aggregate rate)
// Operand contents are irrelevant (timing only).  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mma(uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
                 ::"r"(d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}

template <int N>
__global__ void __launch_bounds__(256, 1) k_issue(int mode, int iters, int nwarps, long long* out, int a_shift16, int b_shift16) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar[8];
    __shared__ uint32_t tslot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 8; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tslot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = tslot;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t hi = (128u >> 4) | (1u << 14);
    const uint32_t a_fixed = 1040u << 16, b_fixed = (uint32_t)N << 16;
    const uint32_t a0 = a_fixed | ((smem_u32(smem) >> 4) + a_shift16), b0 = b_fixed | (((smem_u32(smem) + 140 * 1024) >> 4) + b_shift16);
    constexpr int MS = 512 / N > 8 ? 8 : 512 / N;
    long long t0 = 0, t1 = 0;
    if (mode == 0 || mode == 3) {
        // A / D: one thread per issuing warp
        if (warp < nwarps && lane == 0) {
            const int ms_per = MS / nwarps > 0 ? MS / nwarps : 1;
            t0 = clock64();
            uint32_t a_lo = a0, b_lo = b0, acc = 0;
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int ms = 0; ms < 8; ++ms) {
                    const int m = (warp * ms_per + (ms % ms_per)) % MS;
                    mma(tbase + m * N, a_lo + m * 128, b_lo, hi, idesc, acc);
                }
                acc = 1; a_lo += 2; b_lo += 2;
                if ((it & 3) == 3) { a_lo -= 8; b_lo -= 8; }
            }
            commit(smem_u32(&bar[warp]));
            wait(smem_u32(&bar[warp]), 0);
            t1 = clock64();
            out[warp * 2] = t1 - t0;
        }
    } else if (mode == 1) {
        // B: converged warp, elect per MMA
        if (warp == 0) {
            t0 = clock64();
            uint32_t a_lo = a0, b_lo = b0, acc = 0;
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int ms = 0; ms < 8; ++ms) {
                    const int m = ms % MS;
                    if (elect_one()) mma(tbase + m * N, a_lo + m * 128, b_lo, hi, idesc, acc);
                }
                acc = 1; a_lo += 2; b_lo += 2;
                if ((it & 3) == 3) { a_lo -= 8; b_lo -= 8; }
            }
            if (elect_one()) commit(smem_u32(&bar[0]));
            __syncwarp();
            wait(smem_u32(&bar[0]), 0);
            t1 = clock64();
            if (lane == 0) out[0] = t1 - t0;
        }
    } else if (mode == 2) {
        // C: converged warp, one elect per group of 8
        if (warp == 0) {
            t0 = clock64();
            uint32_t a_lo = a0, b_lo = b0, acc = 0;
            for (int it = 0; it < iters; ++it) {
                if (elect_one()) {
#pragma unroll
                    for (int ms = 0; ms < 8; ++ms) {
                        const int m = ms % MS;
                        mma(tbase + m * N, a_lo + m * 128, b_lo, hi, idesc, acc);
                    }
                }
                __syncwarp();
                acc = 1; a_lo += 2; b_lo += 2;
                if ((it & 3) == 3) { a_lo -= 8; b_lo -= 8; }
            }
            if (elect_one()) commit(smem_u32(&bar[0]));
            __syncwarp();
            wait(smem_u32(&bar[0]), 0);
            t1 = clock64();
            if (lane == 0) out[0] = t1 - t0;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

template <int N>
static void run(const char* name, int mode, int nwarps, int a_shift16 = 0, int b_shift16 = 0) {
    long long* d; cudaMalloc(&d, 128); cudaMemset(d, 0, 128);
    cudaFuncSetAttribute(k_issue<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    const int iters = 2048;
    for (int r = 0; r < 2; ++r) k_issue<N><<<1, 256, 160 * 1024>>>(mode, iters, nwarps, d, a_shift16, b_shift16);
    long long h[16]; cudaMemcpy(h, d, 128, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaDeviceSynchronize();
    long long mx = 0; for (int w = 0; w < (mode == 3 ? nwarps : 1); ++w) mx = h[w * 2] > mx ? h[w * 2] : mx;
    const double total = (double)iters * 8 * (mode == 3 ? nwarps : 1);
    printf("N=%3d %-52s %7.1f cyc per MMA (aggregate), ideal %d %s\n", N, name, (double)mx / total, N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(d);
}

template <int N> static void all() {
    run<N>("A one thread, running adds, 8 unrolled", 0, 1);
    run<N>("B converged warp, elect per MMA", 1, 1);
    run<N>("C converged warp, elect per 8 MMAs", 2, 1);
    run<N>("D 2 warps x one thread, disjoint accumulators", 3, 2);
    run<N>("D 4 warps x one thread, disjoint accumulators", 3, 4);
    run<N>("D 4 warps, A start shifted 1 row (16 B)", 3, 4, 1, 0);
    run<N>("D 4 warps, A start shifted 3 rows", 3, 4, 3, 0);
    run<N>("D 4 warps, A start shifted 4 rows", 3, 4, 4, 0);
    run<N>("D 4 warps, B start shifted 3 rows", 3, 4, 0, 3);
    run<N>("D 4 warps, A and B shifted 3 rows", 3, 4, 3, 3);
}
int main() { all<32>(); all<64>(); all<128>(); all<256>(); return 0; }
