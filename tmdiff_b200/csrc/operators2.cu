// Stand-alone operators outside the sampling path, second file (8(f) rank 3):
//   LayerNorm over token rows and the GEGLU gate of BasicTransformerBlock (core/Attention.py:69-96, 266-296),
//   the learned-PSF blur + subsample of core/kernelAdaption.py:117-180,
//   band-pair moments behind the torch metrics of core/metrics.py:421-503 and utils/util.py:352-386.
// All memory-bound: one pass over the data, reductions in registers / shared memory.
#include "kernels.h"

#include <math.h>

namespace tmd {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- LayerNorm(C, eps) over rows of x [rows, C]: one warp per row, the row held in registers when C <= 1024 ----
constexpr int LN_WARPS = 8;
template <int kVec>   // float4 pieces per lane held in registers (C <= 128 * kVec), 0: re-read the row
__global__ void __launch_bounds__(LN_WARPS * 32) layernorm_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                  const float* __restrict__ gamma, const float* __restrict__ beta, int rows,
                                                                  int C, float eps) {
    const int row = blockIdx.x * LN_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* xr = x + (size_t)row * C;
    float* yr = y + (size_t)row * C;
    if (kVec) {
        float4 v[kVec ? kVec : 1];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kVec; ++i) {
            const int c = (i * 32 + lane) * 4;
            v[i] = c < C ? *reinterpret_cast<const float4*>(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            s += v[i].x + v[i].y + v[i].z + v[i].w;
        }
        const float mean = warp_sum(s) / C;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < kVec; ++i) {
            const int c = (i * 32 + lane) * 4;
            if (c < C) {
                const float a = v[i].x - mean, b = v[i].y - mean, d = v[i].z - mean, e = v[i].w - mean;
                q += a * a + b * b + d * d + e * e;
            }
        }
        const float rstd = rsqrtf(warp_sum(q) / C + eps);       // biased variance, like nn.LayerNorm
#pragma unroll
        for (int i = 0; i < kVec; ++i) {
            const int c = (i * 32 + lane) * 4;
            if (c >= C) continue;
            const float4 g = gamma ? *reinterpret_cast<const float4*>(gamma + c) : make_float4(1.f, 1.f, 1.f, 1.f);
            const float4 b = beta ? *reinterpret_cast<const float4*>(beta + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            float4 o;
            o.x = (v[i].x - mean) * rstd * g.x + b.x; o.y = (v[i].y - mean) * rstd * g.y + b.y;
            o.z = (v[i].z - mean) * rstd * g.z + b.z; o.w = (v[i].w - mean) * rstd * g.w + b.w;
            *reinterpret_cast<float4*>(yr + c) = o;
        }
    } else {
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += xr[c];
        const float mean = warp_sum(s) / C;
        float q = 0.f;
        for (int c = lane; c < C; c += 32) { const float d = xr[c] - mean; q += d * d; }
        const float rstd = rsqrtf(warp_sum(q) / C + eps);
        for (int c = lane; c < C; c += 32) yr[c] = (xr[c] - mean) * rstd * (gamma ? gamma[c] : 1.f) + (beta ? beta[c] : 0.f);
    }
}

// ---- GEGLU: out[r, j] = p[r, j] * gelu(p[r, D + j]) for p [rows, 2D] (exact erf GELU, F.gelu default) ----
__global__ void geglu_kernel(const float* __restrict__ p, float* __restrict__ out, size_t rows, int D) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= rows * (size_t)D) return;
    const size_t r = i / D;
    const int j = (int)(i - r * D);
    const float4 a = *reinterpret_cast<const float4*>(p + r * 2 * D + j);
    const float4 g = *reinterpret_cast<const float4*>(p + r * 2 * D + D + j);
    auto gelu = [](float v) { return 0.5f * v * (1.f + erff(v * 0.70710678118654752440f)); };
    float4 o;
    o.x = a.x * gelu(g.x); o.y = a.y * gelu(g.y); o.z = a.z * gelu(g.z); o.w = a.w * gelu(g.w);
    *reinterpret_cast<float4*>(out + i) = o;
}

// ---- band-pair moments ----
// For image b: per band of a and of b the sums (v, v^2), and for every band pair the sum a_i * b_j, over the H x W pixels, or
// over the (H-2) x (W-2) outputs of the 3x3 high-pass [[-1,-1,-1],[-1,8,-1],[-1,-1,-1]] (sobel_torch, core/metrics.py:411-422)
// applied to both first.  out[b] = [sa(Ca) | saa(Ca) | sb(Cb) | sbb(Cb) | sab(Ca*Cb) | count].
constexpr int BM_MAXC = 8, BM_THREADS = 256;
template <bool kHighpass>
__global__ void __launch_bounds__(BM_THREADS) band_moments_kernel(const float* __restrict__ A, const float* __restrict__ Bm, double* __restrict__ ws,
                                                                  int Ca, int Cb, int H, int W) {
    __shared__ double acc[4 * BM_MAXC + BM_MAXC * BM_MAXC];
    const int b = blockIdx.y, nacc = 2 * Ca + 2 * Cb + Ca * Cb;
    const size_t HW = (size_t)H * W;
    const float* a0 = A + (size_t)b * Ca * HW;
    const float* b0 = Bm + (size_t)b * Cb * HW;
    for (int i = threadIdx.x; i < nacc; i += BM_THREADS) acc[i] = 0.0;
    __syncthreads();
    float sa[BM_MAXC], saa[BM_MAXC], sb[BM_MAXC], sbb[BM_MAXC], sab[BM_MAXC][BM_MAXC];
#pragma unroll
    for (int i = 0; i < BM_MAXC; ++i) {
        sa[i] = saa[i] = sb[i] = sbb[i] = 0.f;
#pragma unroll
        for (int j = 0; j < BM_MAXC; ++j) sab[i][j] = 0.f;
    }
    const int oh = kHighpass ? H - 2 : H, ow = kHighpass ? W - 2 : W;
    const int count = oh * ow;
    auto fetch = [&](const float* img, int y, int x) {
        if (!kHighpass) return img[(size_t)y * W + x];
        const float* p = img + (size_t)y * W + x;             // (y, x): top-left of the 3x3 window
        float s = 0.f;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) s += p[(size_t)dy * W + dx];
        return 9.f * p[(size_t)W + 1] - s;                     // 8*centre - the eight neighbours
    };
    // fp32 partial sums per thread over a short stride (a few dozen terms each), doubles across threads and CTAs
    for (int p = blockIdx.x * BM_THREADS + threadIdx.x; p < count; p += gridDim.x * BM_THREADS) {
        const int y = p / ow, x = p - y * ow;
        float va[BM_MAXC], vb[BM_MAXC];
#pragma unroll
        for (int i = 0; i < BM_MAXC; ++i) va[i] = i < Ca ? fetch(a0 + (size_t)i * HW, y, x) : 0.f;
#pragma unroll
        for (int j = 0; j < BM_MAXC; ++j) vb[j] = j < Cb ? fetch(b0 + (size_t)j * HW, y, x) : 0.f;
#pragma unroll
        for (int i = 0; i < BM_MAXC; ++i) {
            sa[i] += va[i]; saa[i] = fmaf(va[i], va[i], saa[i]);
            sb[i] += vb[i]; sbb[i] = fmaf(vb[i], vb[i], sbb[i]);
#pragma unroll
            for (int j = 0; j < BM_MAXC; ++j) sab[i][j] = fmaf(va[i], vb[j], sab[i][j]);
        }
    }
    // warp reduction in double, then one shared-memory atomic per warp and value (shared double atomics are CAS loops)
    const bool lead = (threadIdx.x & 31) == 0;
#pragma unroll
    for (int i = 0; i < BM_MAXC; ++i) {
        if (i < Ca) {                                            // warp-uniform conditions
            const double u = warp_sum_d(sa[i]), v = warp_sum_d(saa[i]);
            if (lead) { atomicAdd(&acc[i], u); atomicAdd(&acc[Ca + i], v); }
        }
        if (i < Cb) {
            const double u = warp_sum_d(sb[i]), v = warp_sum_d(sbb[i]);
            if (lead) { atomicAdd(&acc[2 * Ca + i], u); atomicAdd(&acc[2 * Ca + Cb + i], v); }
        }
#pragma unroll
        for (int j = 0; j < BM_MAXC; ++j)
            if (i < Ca && j < Cb) {
                const double u = warp_sum_d(sab[i][j]);
                if (lead) atomicAdd(&acc[2 * Ca + 2 * Cb + i * Cb + j], u);
            }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nacc; i += BM_THREADS) atomicAdd(&ws[(size_t)b * nacc + i], acc[i]);
}

__global__ void band_moments_finish_kernel(const double* __restrict__ ws, float* __restrict__ out, int B, int nacc, float count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * (nacc + 1)) return;
    const int b = i / (nacc + 1), k = i - b * (nacc + 1);
    out[i] = k < nacc ? (float)ws[(size_t)b * nacc + k] : count;
}

// ---- learned PSF blur + subsample (EntireKernel / GaussianKernel forward) ----
// out[b,c,y,x] = sum_{i,j} psf[i][j] * in[b,c, y*s + i, x*s + j]  (valid cross-correlation, every s-th output kept).
// mode 0: psf = softmax(param[k*k]) (EntireKernel: k == scale);  mode 1: Gaussian from the 2x2 parameter (kernel_generator,
// core/kernelAdaption.py:12-47, shift 'center').  The PSF is built in shared memory by the block and also written to psf_out.
constexpr int PSF_MAXK = 16;
__global__ void __launch_bounds__(256) psf_downsample_kernel(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ param,
                                                             float* __restrict__ psf_out, int mode, int H, int W, int k, int s, int OH, int OW) {
    __shared__ float psf[PSF_MAXK * PSF_MAXK];
    __shared__ float red[2];
    const int kk = k * k;
    if (threadIdx.x < 32) {
        // raw kernel values by the first warp, then one normalisation
        float mx = -INFINITY;
        if (mode == 0)
            for (int i = threadIdx.x; i < kk; i += 32) mx = fmaxf(mx, param[i]);
#pragma unroll
        for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        // INV_SIGMA = M^T M with M = Q * [[1,0],[1,1]]
        const float q00 = mode ? param[0] : 0.f, q10 = mode ? param[2] : 0.f, q11 = mode ? param[3] : 0.f;
        const float s00 = q00 * q00 + q10 * q10, s01 = q10 * q11, s11 = q11 * q11;
        const float mu = (float)(k / 2);
        for (int i = threadIdx.x; i < kk; i += 32) {
            float v;
            if (mode == 0) v = expf(param[i] - mx);
            else {
                const float zx = (float)(i / k) - mu, zy = (float)(i % k) - mu;     // meshgrid 'ij': X = row index
                v = expf(-0.5f * (zx * (s00 * zx + s01 * zy) + zy * (s01 * zx + s11 * zy)));
            }
            psf[i] = v;
            sum += v;
        }
        sum = warp_sum(sum);
        if (threadIdx.x == 0) red[0] = sum;
    }
    __syncthreads();
    const float inv = 1.f / red[0];
    for (int i = threadIdx.x; i < kk; i += blockDim.x) {
        if (blockIdx.x == 0 && blockIdx.y == 0 && psf_out) psf_out[i] = psf[i] * inv;
    }
    const float* img = in + (size_t)blockIdx.y * H * W;
    float* o = out + (size_t)blockIdx.y * OH * OW;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < OH * OW; p += gridDim.x * blockDim.x) {
        const int y = p / OW, x = p - y * OW;
        const float* w0 = img + (size_t)(y * s) * W + x * s;
        float a = 0.f;
        for (int i = 0; i < k; ++i)
            for (int j = 0; j < k; ++j) a = fmaf(psf[i * k + j], w0[(size_t)i * W + j], a);
        o[p] = a * inv;
    }
}

}  // namespace

cudaError_t layernorm_rows(const float* x, float* y, const float* gamma, const float* beta, int rows, int C, float eps, cudaStream_t s) {
    const unsigned grid = (unsigned)((rows + LN_WARPS - 1) / LN_WARPS);
    const bool vec = (C % 4 == 0) && !((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(gamma) |
                                        reinterpret_cast<uintptr_t>(beta)) & 15);
    if (vec && C <= 128) layernorm_kernel<1><<<grid, LN_WARPS * 32, 0, s>>>(x, y, gamma, beta, rows, C, eps);
    else if (vec && C <= 512) layernorm_kernel<4><<<grid, LN_WARPS * 32, 0, s>>>(x, y, gamma, beta, rows, C, eps);
    else if (vec && C <= 1024) layernorm_kernel<8><<<grid, LN_WARPS * 32, 0, s>>>(x, y, gamma, beta, rows, C, eps);
    else layernorm_kernel<0><<<grid, LN_WARPS * 32, 0, s>>>(x, y, gamma, beta, rows, C, eps);
    return cudaGetLastError();
}

cudaError_t geglu_rows(const float* p, float* out, size_t rows, int D, cudaStream_t s) {
    if (D % 4 || ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(out)) & 15)) return cudaErrorInvalidValue;
    const size_t n4 = rows * (size_t)D / 4;
    geglu_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(p, out, rows, D);
    return cudaGetLastError();
}

cudaError_t band_moments(const float* a, const float* b, float* out, int B, int Ca, int Cb, int H, int W, bool highpass, cudaStream_t s) {
    if (Ca > BM_MAXC || Cb > BM_MAXC || Ca < 1 || Cb < 1) return cudaErrorInvalidValue;
    if (highpass && (H < 3 || W < 3)) return cudaErrorInvalidValue;
    const int nacc = 2 * Ca + 2 * Cb + Ca * Cb;
    const int count = highpass ? (H - 2) * (W - 2) : H * W;
    // enough CTAs for two per SM, each with at least a few pixels per thread
    int split = (2 * 148 + B - 1) / B, most = (count + 4 * BM_THREADS - 1) / (4 * BM_THREADS);
    if (split > most) split = most;
    if (split < 1) split = 1;
    double* ws = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&ws, (size_t)B * nacc * sizeof(double), s);
    if (e != cudaSuccess) return e;
    cudaMemsetAsync(ws, 0, (size_t)B * nacc * sizeof(double), s);
    const dim3 grid((unsigned)split, (unsigned)B);
    if (highpass) band_moments_kernel<true><<<grid, BM_THREADS, 0, s>>>(a, b, ws, Ca, Cb, H, W);
    else band_moments_kernel<false><<<grid, BM_THREADS, 0, s>>>(a, b, ws, Ca, Cb, H, W);
    band_moments_finish_kernel<<<(B * (nacc + 1) + 255) / 256, 256, 0, s>>>(ws, out, B, nacc, (float)count);
    e = cudaGetLastError();
    cudaFreeAsync(ws, s);
    return e;
}

cudaError_t psf_downsample(const float* in, float* out, const float* param, float* psf_out, int mode, int planes, int H, int W, int k, int scale,
                           cudaStream_t s) {
    if (k < 1 || k > PSF_MAXK || scale < 1 || H < k || W < k) return cudaErrorInvalidValue;
    const int OH = (H - k) / scale + 1, OW = (W - k) / scale + 1;
    unsigned gx = (unsigned)((OH * OW + 255) / 256);
    if (gx > 64) gx = 64;
    psf_downsample_kernel<<<dim3(gx, (unsigned)planes), 256, 0, s>>>(in, out, param, psf_out, mode, H, W, k, scale, OH, OW);
    return cudaGetLastError();
}

}  // namespace tmd
