// Microbenchmark: what bounds the conv epilogue's global stores?  One CTA per SM, 384 store threads (the 12 epilogue warps),
// every CTA writes `iters` tiles ([16 y][8 x] voxels x C bf16 channels, voxel pitch ld) of a [planes][64][64][ld] tensor.
//   lane32   : lane = voxel, 32 bytes per st.global.v8, C/16 instructions per voxel (the kernel's pattern)
//   coal16   : 32 lanes write 512 contiguous bytes per instruction (16 B per lane), rows of 8 voxels (needs ld == C)
// swept over the number of CTAs (per-SM limit or a shared one?) and the tensor size (L2-resident or not).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o store_patterns store_patterns.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(384) lane32_kernel(__nv_bfloat16* out, int C, int ld, int planes, int iters, long long* cycles) {
    const long long t0 = clock64();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = warp & 3, sub = warp >> 2;      // 3 warps per 32-voxel quarter
    const int r = q * 32 + lane, x_in = r % 8, y_in = r / 8;
    const int total = 8 * 4 * planes;
    int tile = blockIdx.x;
    for (int it = 0; it < iters; ++it, tile += gridDim.x) {
        if (tile >= total) tile -= total;
        const int tx = tile % 8, ty = (tile / 8) % 4, pl = tile / 32;
        const size_t vox = ((size_t)pl * 64 + ty * 16 + y_in) * 64 + tx * 8 + x_in;
        for (int c = sub * 16; c < C; c += 48)
            asm volatile("st.global.v8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"l"(out + vox * ld + c), "r"(it) : "memory");
        if (C == 32 && sub == 2) {   // 2 slabs over 3 warps: the third warp takes every other tile's slabs instead (same bytes per tile on average)
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

__global__ void __launch_bounds__(384) coal16_kernel(__nv_bfloat16* out, int C, int planes, int iters, long long* cycles) {
    const long long t0 = clock64();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = 8 * 4 * planes;
    const int row_bytes = 8 * C * 2, per_instr = 512, instr_per_row = row_bytes / per_instr;      // C = 32: one instruction per row of 8 voxels
    int tile = blockIdx.x;
    for (int it = 0; it < iters; ++it, tile += gridDim.x) {
        if (tile >= total) tile -= total;
        const int tx = tile % 8, ty = (tile / 8) % 4, pl = tile / 32;
        for (int j = warp; j < 16 * instr_per_row; j += 12) {
            const int y_in = j / instr_per_row, part = j % instr_per_row;
            char* p = reinterpret_cast<char*>(out) + (((size_t)pl * 64 + ty * 16 + y_in) * 64 + tx * 8) * C * 2 + part * 512 + lane * 16;
            asm volatile("st.global.v4.b32 [%0], {%1, %1, %1, %1};" ::"l"(p), "r"(it) : "memory");
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
    long long* d_cyc;
    CK(cudaMalloc(&d_cyc, 148 * sizeof(long long)));
    const int iters = 600;
    printf("%-10s %4s %4s %7s %5s %12s %10s\n", "pattern", "C", "ld", "MB", "CTAs", "B/clk/SM", "TB/s@1.85");
    for (int planes : {64, 1024}) {              // 16 MB (L2-resident) / 268 MB at C = 32
        for (int C : {32, 64}) {
            for (int ld : {C, 3 * C}) {
                __nv_bfloat16* out;
                const size_t bytes = (size_t)planes * 4096 * ld * 2;
                CK(cudaMalloc(&out, bytes));
                for (int ctas : {148, 74, 37}) {
                    for (int pat = 0; pat < 2; ++pat) {
                        if (pat == 1 && ld != C) continue;
                        for (int rep = 0; rep < 2; ++rep) {
                            if (pat == 0) lane32_kernel<<<ctas, 384>>>(out, C, ld, planes, iters, d_cyc);
                            else coal16_kernel<<<ctas, 384>>>(out, C, planes, iters, d_cyc);
                            CK(cudaDeviceSynchronize());
                        }
                        std::vector<long long> c(ctas);
                        CK(cudaMemcpy(c.data(), d_cyc, ctas * sizeof(long long), cudaMemcpyDeviceToHost));
                        long long mx = 0;
                        for (auto v : c) mx = v > mx ? v : mx;
                        const double bpc = (double)iters * 128 * C * 2 / mx;
                        printf("%-10s %4d %4d %7.0f %5d %12.1f %10.2f\n", pat ? "coal16" : "lane32", C, ld, bytes / 1e6, ctas, bpc, bpc * ctas * 1.85e9 / 1e12);
                    }
                }
                CK(cudaFree(out));
            }
        }
    }
    return 0;
}
