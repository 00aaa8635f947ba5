// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, cta_group::1, M=128) from shared-memory operands as a
// function of N, the shared-memory layout (no swizzle vs 128-byte swizzle), operand start alignment and the
// number of accumulators in rotation.  Operand contents are whatever is in shared memory (timing only).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe tools/mma_probe.cu ; run: ./mma_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct P {
    int N, n_mma, n_acc, swz;          // swz: 0 none, 1 = 128B swizzle
    uint32_t a_lbo, a_sbo, b_lbo, b_sbo, a_shift, a_kstep, b_kstep, a_acc_step;
    int ksteps;                         // K steps cycled over before wrapping (distinct operand addresses)
};

__global__ void __launch_bounds__(128, 1) k_probe(P p, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tslot;
    const uint32_t barA = smem_u32(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barA) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // fill operands with small finite bf16 values
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tslot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = tslot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint64_t hi_a = ((uint64_t)((p.a_sbo >> 4) & 0x3FFF) | (1ull << 14) | ((uint64_t)(p.swz ? 2 : 0) << 29)) << 32;
        const uint64_t hi_b = ((uint64_t)((p.b_sbo >> 4) & 0x3FFF) | (1ull << 14) | ((uint64_t)(p.swz ? 2 : 0) << 29)) << 32;
        const uint32_t a0 = smem_u32(smem) + p.a_shift, b0 = smem_u32(smem) + 128 * 1024;
        const uint32_t lo_a = ((p.a_lbo >> 4) & 0x3FFF) << 16, lo_b = ((p.b_lbo >> 4) & 0x3FFF) << 16;
        long long t0 = clock64();
        int ks = 0, acc = 0;
        for (int i = 0; i < p.n_mma; ++i) {
            const uint64_t ad = hi_a | lo_a | (((a0 + ks * p.a_kstep + acc * p.a_acc_step) >> 4) & 0x3FFF);
            const uint64_t bd = hi_b | lo_b | (((b0 + ks * p.b_kstep) >> 4) & 0x3FFF);
            const uint32_t d = tbase + (uint32_t)(acc * p.N);
            asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q;\n\t}"
                         ::"r"(d), "l"(ad), "l"(bd), "r"(idesc), "r"(i >= p.n_acc ? 1u : 0u) : "memory");
            if (++acc == p.n_acc) { acc = 0; if (++ks == p.ksteps) ks = 0; }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(barA) : "memory");
        long long t1 = clock64();
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                         : "=r"(ok) : "r"(barA) : "memory");
        long long t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

static void run(const char* name, P p, int grid) {
    long long* d; cudaMalloc(&d, 16);
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int r = 0; r < 2; ++r) k_probe<<<grid, 128, 200 * 1024>>>(p, d);
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%-44s N=%3d acc=%d grid=%3d: issue %7.1f cyc/mma, complete %7.1f cyc/mma (ideal %d)%s\n", name, p.N, p.n_acc, grid,
           (double)h[0] / p.n_mma, (double)h[1] / p.n_mma, p.N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    const int Ns[4] = {32, 64, 128, 256};
    for (int gi = 0; gi < 2; ++gi) {
        const int grid = gi ? 148 : 1;
        for (int ni = 0; ni < 4; ++ni) {
            const int N = Ns[ni];
            const int n_acc = 512 / N > 8 ? 8 : 512 / N;
            P p{}; p.N = N; p.n_mma = 4096; p.ksteps = 3;
            // no swizzle, conv_tc layout: A [KC/8][Rp][8] with Rp = 1024+16 rows, B [KW/8][N][8]
            p.swz = 0; p.a_lbo = 1040 * 16; p.a_sbo = 128; p.b_lbo = N * 16; p.b_sbo = 128;
            p.a_kstep = 2 * p.a_lbo; p.b_kstep = 2 * p.b_lbo; p.a_acc_step = 128 * 16;
            p.n_acc = 1; p.a_shift = 0; run("noswz aligned, 1 acc", p, grid);
            p.n_acc = n_acc; run("noswz aligned, rotating acc", p, grid);
            p.a_shift = 48; run("noswz A shifted 3 rows, rotating acc", p, grid);
            p.n_acc = 1; run("noswz A shifted 3 rows, 1 acc", p, grid);
            // 128B swizzle, K-major: rows of 128 B (64 bf16), 8-row groups 1024 B; K step = 32 B inside the atom
            p.swz = 1; p.a_lbo = 16; p.a_sbo = 1024; p.b_lbo = 16; p.b_sbo = 1024; p.a_kstep = 32; p.b_kstep = 32;
            p.a_acc_step = 128 * 128; p.a_shift = 0;
            p.n_acc = 1; run("swz128 1 acc", p, grid);
            p.n_acc = n_acc > 4 ? 4 : n_acc; run("swz128 rotating acc", p, grid);
        }
    }
    return 0;
}
