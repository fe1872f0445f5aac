// Experiment: can one haloed TMA box [lines][10 voxels][32 ch] (SWIZZLE_64B, line pitch 640 B) serve the three dx taps
// through UMMA descriptors whose start address is shifted by dx*64 B and whose 8-row group stride (SBO) is 640 B?
// Prints the max abs error of D = A_view(dx, dy) x W^T against a host reference for each (dx, dy, base_offset mode).
//   nvcc -gencode arch=compute_100a,code=sm_100a -o shifted_view shifted_view.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int LINES = 18, XW = 10, KC = 32, NCO = 32;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct Params { CUtensorMap tmA, tmB; float* out; int dx, dy, mode; };

__global__ void __launch_bounds__(128, 1) kern(const __grid_constant__ Params P) {
    extern __shared__ uint8_t raw[];
    __shared__ __align__(8) uint64_t bar_ld, bar_mma;
    __shared__ uint32_t tmem_slot;
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t a_smem = base, b_smem = base + 16384;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_ld)) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_mma)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        const uint32_t bytes = LINES * XW * KC * 2 + NCO * KC * 2;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar_ld)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(a_smem), "l"(&P.tmA), "r"(smem_u32(&bar_ld)), "r"(0), "r"(0), "r"(0) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(b_smem), "l"(&P.tmB), "r"(smem_u32(&bar_ld)), "r"(0), "r"(0) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_u32(&bar_ld)) : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (warp == 1) {
        uint32_t pred;
        asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
        if (pred) {
            const uint32_t row_bytes = KC * 2;                       // 64
            const uint32_t a_start = a_smem + (uint32_t)P.dy * XW * row_bytes + (uint32_t)P.dx * row_bytes;
            const uint32_t sbo_a = (XW * row_bytes) >> 4, sbo_b = (8 * row_bytes) >> 4;
            uint64_t a_hi = sbo_a | (1u << 14) | (4u << 29), b_hi = sbo_b | (1u << 14) | (4u << 29);
            if (P.mode == 1) a_hi |= (uint64_t)((a_start >> 7) & 7) << (49 - 32);
            const uint32_t a_lo = ((a_start & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_lo = ((b_smem & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NCO >> 3) << 17) | ((128u >> 4) << 24);
            for (int k = 0; k < 2; ++k)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             ::"r"(tmem), "l"((a_hi << 32) | (a_lo + k * 2)), "l"((b_hi << 32) | (b_lo + k * 2)), "r"(idesc), "r"(k ? 1u : 0u)
                             : "memory");
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_mma)) : "memory");
        }
        __syncwarp();
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(smem_u32(&bar_mma)) : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
        const int row = threadIdx.x;                                  // warp w reads TMEM lanes 32w..32w+31
        for (int c0 = 0; c0 < NCO; c0 += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                           "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 16; ++i) P.out[row * NCO + c0 + i] = __uint_as_float(r[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    std::vector<__nv_bfloat16> hA(LINES * XW * KC), hW(NCO * KC);
    std::vector<float> fA(hA.size()), fW(hW.size());
    srand(1);
    for (size_t i = 0; i < hA.size(); ++i) { hA[i] = __float2bfloat16((rand() % 17 - 8) / 8.f); fA[i] = __bfloat162float(hA[i]); }
    for (size_t i = 0; i < hW.size(); ++i) { hW[i] = __float2bfloat16((rand() % 13 - 6) / 8.f); fW[i] = __bfloat162float(hW[i]); }
    __nv_bfloat16 *dA, *dW;
    float* dO;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dW, hW.size() * 2); cudaMalloc(&dO, 128 * NCO * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice);
    Params P;
    {
        cuuint64_t dims[3] = {KC, XW, LINES};
        cuuint64_t strides[2] = {KC * 2, (cuuint64_t)KC * 2 * XW};
        cuuint32_t box[3] = {KC, XW, LINES}, es[3] = {1, 1, 1};
        CUresult r = enc(&P.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, dA, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode A failed %d\n", r); return 1; }
        cuuint64_t dimsB[2] = {KC, NCO};
        cuuint64_t stridesB[1] = {KC * 2};
        cuuint32_t boxB[2] = {KC, NCO};
        r = enc(&P.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dW, dimsB, stridesB, boxB, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r) { printf("encode B failed %d\n", r); return 1; }
    }
    P.out = dO;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    std::vector<float> out(128 * NCO);
    for (int mode = 0; mode < 2; ++mode)
        for (int dy = 0; dy < 3; ++dy)
            for (int dx = 0; dx < 3; ++dx) {
                P.dx = dx; P.dy = dy; P.mode = mode;
                cudaMemset(dO, 0, out.size() * 4);
                kern<<<1, 128, 64 * 1024>>>(P);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("mode %d dy %d dx %d: CUDA error %s\n", mode, dy, dx, cudaGetErrorString(e)); return 1; }
                cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost);
                double worst = 0;
                for (int r = 0; r < 128; ++r)
                    for (int n = 0; n < NCO; ++n) {
                        const int l = r / 8 + dy, x = r % 8 + dx;
                        double ref = 0;
                        for (int c = 0; c < KC; ++c) ref += (double)fA[(l * XW + x) * KC + c] * fW[n * KC + c];
                        const double d = fabs(ref - out[r * NCO + n]);
                        if (d > worst) worst = d;
                    }
                printf("base_offset_mode %d  dy %d dx %d  max|err| %.4g\n", mode, dy, dx, worst);
            }
    return 0;
}
