// fp32 CUDA-core implicit-GEMM 3-D convolution (the parity / fp32 mode of the conv path).
//
// Replaces nn.Conv3d + F.conv3d(groups=B) call sites of the reference denoiser
// (GeneralModel/Hyper_unet_general.py:74, :161-164, :224-231, :344-361).  Arithmetic is plain
// fp32 FMA so the result stays within ~1e-6 of the reference's fp32 path.
//
// Tiling: one CTA = one (sample, band-plane), a TH x TW window of that plane and BN output
// channels.  256 threads; a thread owns a 2x2 spatial quad x 4 channels, which makes the Haar
// pooling / upsampling epilogues thread-local.  K loop: band tap dz (outer) x 16-channel chunk;
// per chunk the haloed input window and the 9 in-plane taps' weights are staged in shared memory.
#include "conv_op.h"
#include "kernels.h"

namespace tmd {

constexpr int KC = 16;        // input channels per staged chunk
constexpr int AS = KC + 4;    // smem row stride of the input window (bank spread, 16B aligned)

template <typename T, int BN, bool kFast>
__global__ void __launch_bounds__(256) conv_simt_kernel(ConvOp op, int TH, int TW, int tiles_x) {
    constexpr int NTX = BN / 4;
    extern __shared__ __align__(16) float smem[];
    const int halo = op.taps == 27 ? 1 : 0;
    const int ntp = op.taps == 27 ? 9 : 1;
    const int PH = TH + 2 * halo, PW = TW + 2 * halo;
    float* sA = smem;
    float* sB = smem + ((PH * PW * AS + 3) & ~3);

    const int tid = threadIdx.x;
    const int tx = tid % NTX, ty = tid / NTX;
    const int qcols = TW / 2;
    const int qr = ty / qcols, qc = ty % qcols;
    const int tile = blockIdx.x;
    const int y0 = (tile / tiles_x) * TH, x0 = (tile % tiles_x) * TW;
    const int n = blockIdx.y % op.N, b = blockIdx.y / op.N;
    const int co_base = blockIdx.z * BN;
    const T* __restrict__ src = reinterpret_cast<const T*>(op.src);

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int dz = -halo; dz <= halo; ++dz) {
        const int nz = n + dz;
        if (nz < 0 || nz >= op.N) continue;  // zero padding along the band axis (block-uniform)
        const size_t plane = ((size_t)b * op.N + nz) * op.H;
        for (int c0 = 0; c0 < op.Ci; c0 += KC) {
            __syncthreads();
            for (int idx = tid; idx < PH * PW * (KC / 4); idx += 256) {
                const int p = idx / (KC / 4), cq = idx % (KC / 4);
                const int gy = y0 + p / PW - halo, gx = x0 + p % PW - halo;
                float v[4] = {0.f, 0.f, 0.f, 0.f};
                if (gy >= 0 && gy < op.H && gx >= 0 && gx < op.W && c0 + cq * 4 < op.Ci)
                    load4<T>(src + ((plane + gy) * op.W + gx) * op.ld_in + c0 + cq * 4, v);
                *reinterpret_cast<float4*>(sA + p * AS + cq * 4) = make_float4(v[0], v[1], v[2], v[3]);
            }
            const int tap0 = op.taps == 27 ? (dz + 1) * 9 : 0;
            for (int idx = tid; idx < ntp * KC * (BN / 4); idx += 256) {
                const int cq = idx % (BN / 4), k = (idx / (BN / 4)) % KC, t = idx / (BN / 4) / KC;
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c0 + k < op.Ci && co_base + cq * 4 < op.Co)
                    w = *reinterpret_cast<const float4*>(op.w_f32 + (size_t)b * op.w_sample_stride + ((size_t)(tap0 + t) * op.Ci + c0 + k) * op.Co + co_base + cq * 4);
                *reinterpret_cast<float4*>(sB + (t * KC + k) * BN + cq * 4) = w;
            }
            __syncthreads();
            for (int t = 0; t < ntp; ++t) {
                const int dy = ntp == 9 ? t / 3 : 0, dx = ntp == 9 ? t % 3 : 0;  // already shifted by +halo
                const float* a0 = sA + ((2 * qr + dy) * PW + 2 * qc + dx) * AS;
                const float* a2 = a0 + PW * AS;
                const float* bp = sB + t * KC * BN + tx * 4;
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    const float4 w = *reinterpret_cast<const float4*>(bp + k * BN);
                    const float a[4] = {a0[k], a0[AS + k], a2[k], a2[AS + k]};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        acc[i][0] = fmaf(a[i], w.x, acc[i][0]);
                        acc[i][1] = fmaf(a[i], w.y, acc[i][1]);
                        acc[i][2] = fmaf(a[i], w.z, acc[i][2]);
                        acc[i][3] = fmaf(a[i], w.w, acc[i][3]);
                    }
                }
            }
        }
    }

    // ---------------- epilogue ----------------
    const int co = co_base + tx * 4;
    const int oy = y0 + 2 * qr, ox = x0 + 2 * qc;
    if (co >= op.Co || oy >= op.H || ox >= op.W) return;
    const ConvEpilogue& e = op.epi;
    float bias[4] = {0.f, 0.f, 0.f, 0.f};
    if (e.bias) {
#pragma unroll
        for (int i = 0; i < 4; ++i) bias[i] = e.bias[(size_t)b * e.ld_bias + co + i];
    }
    const size_t bn = (size_t)b * op.N + n;
    if (e.mode == CONV_NORMAL) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int y = oy + (q >> 1), x = ox + (q & 1);
            if (y >= op.H || x >= op.W) continue;
            float v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = acc[q][i] + bias[i];
            const size_t ovox = (bn * op.H + y) * op.W + x;
            const size_t hvox = (bn * (op.H / 2) + y / 2) * (op.W / 2) + x / 2;
            epilogue_store4<T, kFast>(e, b, ovox, hvox, co, v);
        }
    } else if (e.mode == CONV_POOL) {
        const size_t ovox = (bn * (op.H / 2) + oy / 2) * (op.W / 2) + ox / 2;
        float v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float a = acc[0][i], bb = acc[1][i], c = acc[2][i], d = acc[3][i];
            v[i] = ((a + bb) + (c + d)) * 0.25f + bias[i];
            if (e.hf_out[0]) {
                // un-halved Haar high bands: LH, HL, HH
                T* lh = reinterpret_cast<T*>(e.hf_out[0]) + ovox * e.ld_hf + co + i;
                T* hl = reinterpret_cast<T*>(e.hf_out[1]) + ovox * e.ld_hf + co + i;
                T* hh = reinterpret_cast<T*>(e.hf_out[2]) + ovox * e.ld_hf + co + i;
                *lh = from_f32<T>(((a - bb) + (c - d)) * 0.5f);
                *hl = from_f32<T>(((a + bb) - (c + d)) * 0.5f);
                *hh = from_f32<T>(((a - bb) - (c - d)) * 0.5f);
            }
        }
        epilogue_store4<T, kFast>(e, b, ovox, 0, co, v);
    } else {  // CONV_UPSAMPLE
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int y = oy + (q >> 1), x = ox + (q & 1);
            if (y >= op.H || x >= op.W) continue;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                float v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = acc[q][i] + bias[i];
                const size_t ovox = (bn * (2 * op.H) + 2 * y + (r >> 1)) * (2 * op.W) + 2 * x + (r & 1);
                epilogue_store4<T, kFast>(e, b, ovox, 0, co, v);
            }
        }
    }
}

template <typename T, bool kFast>
static cudaError_t launch_simt(const ConvOp& op, cudaStream_t stream) {
    const int BN = op.Co <= 32 ? 32 : 64;
    const int NTY = 256 / (BN / 4);            // quads per CTA
    int TW = op.W >= 16 ? 16 : 8;
    if (op.W < 8) TW = (op.W + 1) & ~1;
    const int TH = NTY * 4 / TW;                // thread map needs (TH/2)*(TW/2) == NTY; rows past H are predicated
    const int tiles_x = (op.W + TW - 1) / TW, tiles_y = (op.H + TH - 1) / TH;
    const int halo = op.taps == 27 ? 1 : 0, ntp = op.taps == 27 ? 9 : 1;
    const int PH = TH + 2 * halo, PW = TW + 2 * halo;
    const size_t smem = (size_t)(((PH * PW * AS + 3) & ~3) + ntp * KC * BN) * sizeof(float);
    dim3 grid(tiles_x * tiles_y, op.N * op.B, (op.Co + BN - 1) / BN);
    if (BN == 32) {
        auto k = conv_simt_kernel<T, 32, kFast>;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, 256, smem, stream>>>(op, TH, TW, tiles_x);
    } else {
        auto k = conv_simt_kernel<T, 64, kFast>;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, 256, smem, stream>>>(op, TH, TW, tiles_x);
    }
    return cudaGetLastError();
}

cudaError_t conv_simt_f32(const ConvOp& op, cudaStream_t stream) { return launch_simt<float, false>(op, stream); }
// bf16-stored activations through the same CUDA-core kernel: the checker for the tcgen05 path and
// the route for layer shapes the tensor-core kernel does not take (channel counts not multiples of 16).
cudaError_t conv_simt_bf16(const ConvOp& op, cudaStream_t stream) { return launch_simt<__nv_bfloat16, true>(op, stream); }

}  // namespace tmd
