// Microbenchmark: how fast does one SM push an epilogue tile to global memory?
//   (a) cp.async.bulk.tensor stores (shared -> global) of a [C ch][8 x][16 y] box (128 voxels) for inner extents
//       C*2 B = 32 / 64 / 128 bytes, voxel pitch ld*2 bytes in global memory
//   (b) the per-lane 32-byte st.global.v8 pattern of the round-1 epilogue (one voxel per lane)
// Every CTA (one per SM, 128 store threads) writes `iters` tiles of a [B][N][64][64][ld] bf16 tensor, tile after tile.
// Output: bytes per clock per SM.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_store_rows tma_store_rows.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int kSlots>
__global__ void __launch_bounds__(128) tma_store_kernel(const __grid_constant__ CUtensorMap tm, int C, int tiles_x, int tiles_y, int planes,
                                                        int iters, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t slot_bytes = 128u * C * 2u;
    // fill the slots once (contents do not matter)
    for (uint32_t i = threadIdx.x; i < kSlots * slot_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(smem + (base - smem_u32(smem)))[i] = i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const long long t0 = clock64();
    int tile = blockIdx.x;
    const int total = tiles_x * tiles_y * planes;
    for (int it = 0; it < iters; ++it, tile += gridDim.x) {
        if (tile >= total) tile -= total;
        const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, pl = tile / (tiles_x * tiles_y);
        // every thread "writes" its row of the slot (32..128 bytes), as the epilogue would
        const uint32_t slot = base + (it % kSlots) * slot_bytes;
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kSlots - 1) : "memory");
        __syncthreads();
        for (int c = 0; c < C * 2; c += 16)
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(slot + threadIdx.x * C * 2 + (c ^ (((threadIdx.x >> 1) & 3) << 4)) % (C * 2)), "r"(it));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(&tm), "r"(slot), "r"(0),
                         "r"(tx * 8), "r"(ty * 16), "r"(pl), "r"(0)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

// round-1 pattern: lane = voxel, 32 bytes (16 channels) per st.global.v8, C/16 instructions per voxel
__global__ void __launch_bounds__(128) lane_store_kernel(__nv_bfloat16* out, int C, int ld, int tiles_x, int tiles_y, int planes, int iters,
                                                         long long* cycles) {
    const long long t0 = clock64();
    int tile = blockIdx.x;
    const int total = tiles_x * tiles_y * planes;
    const int x_in = threadIdx.x % 8, y_in = threadIdx.x / 8;
    for (int it = 0; it < iters; ++it, tile += gridDim.x) {
        if (tile >= total) tile -= total;
        const int tx = tile % tiles_x, ty = (tile / tiles_x) % tiles_y, pl = tile / (tiles_x * tiles_y);
        const size_t vox = ((size_t)pl * 64 + ty * 16 + y_in) * 64 + tx * 8 + x_in;
        for (int c = 0; c < C; c += 16)
            asm volatile("st.global.v8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"l"(out + vox * ld + c), "r"(it) : "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    EncodeTiledFn enc = (EncodeTiledFn)fn;
    const int planes = 256, H = 64, W = 64, iters = 400;
    long long* d_cyc;
    CK(cudaMalloc(&d_cyc, 148 * sizeof(long long)));
    printf("%-34s %8s %10s %12s\n", "variant", "C", "ld", "B/clk/SM");
    for (int ld : {32, 96}) {
        __nv_bfloat16* out;
        CK(cudaMalloc(&out, (size_t)planes * H * W * ld * 2));
        for (int C : {16, 32, 64}) {
            if (C > ld) continue;
            CUtensorMap tm;
            cuuint64_t dims[5] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes, 1};
            cuuint64_t strides[4] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * W, (cuuint64_t)ld * 2 * W * H, (cuuint64_t)ld * 2 * W * H * planes};
            cuuint32_t box[5] = {(cuuint32_t)C, 8, 16, 1, 1}, estr[5] = {1, 1, 1, 1, 1};
            const CUtensorMapSwizzle sw = C == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : (C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
            if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                    CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
            for (int slots : {2, 4}) {
                const size_t smem = (size_t)slots * 128 * C * 2 + 1024;
                for (int rep = 0; rep < 2; ++rep) {
                    if (slots == 2) tma_store_kernel<2><<<148, 128, smem>>>(tm, C, W / 8, H / 16, planes, iters, d_cyc);
                    else tma_store_kernel<4><<<148, 128, smem>>>(tm, C, W / 8, H / 16, planes, iters, d_cyc);
                    CK(cudaDeviceSynchronize());
                }
                std::vector<long long> c(148);
                CK(cudaMemcpy(c.data(), d_cyc, 148 * sizeof(long long), cudaMemcpyDeviceToHost));
                long long mx = 0;
                for (auto v : c) mx = v > mx ? v : mx;
                char name[64];
                snprintf(name, sizeof(name), "tma tensor store, %d slots", slots);
                printf("%-34s %8d %10d %12.1f\n", name, C, ld, (double)iters * 128 * C * 2 / mx);
            }
            for (int rep = 0; rep < 2; ++rep) {
                lane_store_kernel<<<148, 128>>>(out, C, ld, W / 8, H / 16, planes, iters, d_cyc);
                CK(cudaDeviceSynchronize());
            }
            std::vector<long long> c(148);
            CK(cudaMemcpy(c.data(), d_cyc, 148 * sizeof(long long), cudaMemcpyDeviceToHost));
            long long mx = 0;
            for (auto v : c) mx = v > mx ? v : mx;
            printf("%-34s %8d %10d %12.1f\n", "st.global.v8 per lane", C, ld, (double)iters * 128 * C * 2 / mx);
        }
        CK(cudaFree(out));
    }
    return 0;
}
