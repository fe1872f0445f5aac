// Micro-benchmark: cycles per tcgen05.mma (cta_group::1, kind::f16, M=128, K=16) as a function of N and of the
// shared-memory swizzle of the operands, one CTA per SM, all SMs busy.  Answers: is N=96 (three merged dz taps at
// Co=32) slower per column than N=128/192/256?   nvcc -arch=sm_100a -o mma_shapes mma_shapes.cu && ./mma_shapes
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1) bench(int N, int swz64, int iters, int a_views, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = 0x3c003c00u + i;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (warp == 1) {   // warp-uniform loop, one elected lane issues (a divergent branch makes ptxas wrap each MMA in an election loop)
        const uint32_t row_bytes = swz64 ? 64 : 128;
        const uint32_t sbo16 = (8 * row_bytes) >> 4;
        const uint32_t desc_hi = sbo16 | (1u << 14) | ((swz64 ? 4u : 2u) << 29);
        const uint32_t a_lo0 = ((base & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t b_lo0 = (((base + 96 * 1024) & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t line16 = (8 * row_bytes) >> 4;
        uint32_t pred;
        asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
        const long long t0 = clock64();
        for (int i = 0; i < iters; i += 8) {
            if (pred) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t a_lo = a_lo0 + (uint32_t)(j >> 1) * line16 * (uint32_t)a_views + (j & 1) * 2;
                    const uint32_t b_lo = b_lo0 + (j & 1) * 2;
                    const uint32_t d = tmem + (uint32_t)((j >> 2) * 256);
                    asm volatile(
                        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                        ::"r"(d), "l"(((uint64_t)desc_hi << 32) | a_lo), "l"(((uint64_t)desc_hi << 32) | b_lo), "r"(idesc), "r"(i ? 1u : 0u)
                        : "memory");
                }
            }
            __syncwarp();
        }
        if (pred) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        __syncwarp();
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        const long long t1 = clock64();
        if (blockIdx.x == 0 && pred) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

int main() {
    long long* d;
    cudaMalloc(&d, 8);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 8192;
    printf("swizzle N cycles/mma ideal(N/2) flop/clk/SM\n");
    for (int swz64 = 0; swz64 < 2; ++swz64)
        for (int N : {16, 32, 48, 64, 96, 128, 160, 192, 224, 256}) {
            for (int rep = 0; rep < 2; ++rep) {
                bench<<<148, 128, 200 * 1024>>>(N, swz64, iters, 2, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            long long c;
            cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
            printf("%s %3d %7.1f %6.1f %8.0f\n", swz64 ? "64B " : "128B", N, (double)c / iters, N / 2.0, 2.0 * 128 * N * 16 * iters / (double)c);
        }
    return 0;
}
