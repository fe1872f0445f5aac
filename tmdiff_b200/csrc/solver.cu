// DPM-Solver++ device arithmetic: x_start -> eps -> x0 conversion with the exact per-sample
// 99.5% dynamic threshold, and the update family.  (core/dpm_solver_pytorch.py:296-306,
// :430-456, :589-592, :648-658, :744-760.)  Every operation is rounded exactly where the
// reference's fp32 tensor expression rounds (no FMA contraction), because x0 = (x - s*eps)/a
// divides by alpha_t ~ 5e-5 on the first steps and amplifies any re-association.
#include <float.h>

#include "kernels.h"

namespace tmd {

constexpr int TH_THREADS = 1024;

__device__ __forceinline__ float x0_of(float x, float out, float alpha, float sigma) {
    const float eps = __fdiv_rn(__fsub_rn(x, __fmul_rn(alpha, out)), sigma);   // model_wrapper x_start -> noise
    return __fdiv_rn(__fsub_rn(x, __fmul_rn(sigma, eps)), alpha);              // data_prediction_fn
}

// In-block bitonic sort, descending, P = power of two.
__device__ void bitonic_desc(float* v, int P) {
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const float a = v[i], b = v[ixj];
                    const bool desc = (i & k) == 0;
                    if (desc ? (a < b) : (a > b)) { v[i] = b; v[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
}

// One CTA per sample.  kFromModel: compute x0 from (x, out) first; otherwise threshold dst in place.
//
// s = max(quantile(|x0|, q), max_val) with torch.quantile's linear interpolation between sorted
// positions q_lo and q_lo+1.  The exact order statistics are only needed when at least
// n-q_lo-1 values exceed max_val; otherwise both neighbours are <= max_val and s == max_val.
// In the needed case only the values > max_val take part: they are gathered into shared memory
// and sorted there (worst case all n of them).
template <bool kFromModel>
__global__ void __launch_bounds__(TH_THREADS) x0_threshold_kernel(const float* x, const float* out,
                                                                   float* dst, float alpha, float sigma, int n,
                                                                   int q_lo, float q_w, float max_val) {
    extern __shared__ float big[];
    __shared__ int s_count;
    __shared__ float s_red[32];
    __shared__ float s_scale;
    const int b = blockIdx.x;
    const size_t base = (size_t)b * n;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();

    int cnt = 0;
    float m1 = 0.f;  // largest |x0| that is <= max_val
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float v;
        if (kFromModel) {
            v = x0_of(x[base + i], out[base + i], alpha, sigma);
            dst[base + i] = v;
        } else {
            v = dst[base + i];
        }
        const float a = fabsf(v);
        if (a > max_val) ++cnt; else m1 = fmaxf(m1, a);
    }
    // block reductions
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&s_count, cnt); s_red[threadIdx.x >> 5] = m1; }
    __syncthreads();
    const int total = s_count;
    if (threadIdx.x < 32) {
        float m = threadIdx.x < (blockDim.x >> 5) ? s_red[threadIdx.x] : 0.f;
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) s_red[0] = m;
    }
    __syncthreads();
    m1 = s_red[0];

    const int hi_pos = min(q_lo + 1, n - 1);
    const int need_hi = n - hi_pos;   // v[hi_pos] is the need_hi-th largest
    const int need_lo = n - q_lo;     // v[q_lo]  is the need_lo-th largest
    if (total < need_hi) {
        if (threadIdx.x == 0) s_scale = max_val;
    } else {
        __syncthreads();
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
        int P = 1;
        while (P < total) P <<= 1;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const float a = fabsf(dst[base + i]);
            if (a > max_val) big[atomicAdd(&s_count, 1)] = a;
        }
        for (int i = total + threadIdx.x; i < P; i += blockDim.x) big[i] = -FLT_MAX;
        __syncthreads();
        bitonic_desc(big, P);
        if (threadIdx.x == 0) {
            const float v_hi = big[need_hi - 1];
            const float v_lo = (total >= need_lo) ? big[need_lo - 1] : m1;
            // at::lerp: weight < 0.5 ? a + w*(b-a) : b - (b-a)*(1-w)
            const float diff = __fsub_rn(v_hi, v_lo);
            const float qv = (fabsf(q_w) < 0.5f) ? __fadd_rn(v_lo, __fmul_rn(q_w, diff))
                                                 : __fsub_rn(v_hi, __fmul_rn(diff, __fsub_rn(1.f, q_w)));
            s_scale = fmaxf(qv, max_val);
        }
    }
    __syncthreads();
    const float s = s_scale;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = dst[base + i];
        dst[base + i] = __fdiv_rn(fminf(fmaxf(v, -s), s), s);   // clamp(x0, -s, s) / s
    }
}

static cudaError_t launch_threshold(bool from_model, const float* x, const float* out, float* dst, float alpha, float sigma, int B,
                                    int n, int q_lo, float q_w, float max_val, cudaStream_t s) {
    int P = 1;
    while (P < n) P <<= 1;
    const size_t smem = (size_t)P * sizeof(float);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;  // > 51200 values per sample: not supported by the in-smem select
    if (from_model) {
        cudaFuncSetAttribute(x0_threshold_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        x0_threshold_kernel<true><<<B, TH_THREADS, smem, s>>>(x, out, dst, alpha, sigma, n, q_lo, q_w, max_val);
    } else {
        cudaFuncSetAttribute(x0_threshold_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        x0_threshold_kernel<false><<<B, TH_THREADS, smem, s>>>(nullptr, nullptr, dst, 1.f, 1.f, n, q_lo, q_w, max_val);
    }
    return cudaGetLastError();
}

cudaError_t x0_threshold(const float* x, const float* out, float* dst, float alpha, float sigma, int B, int n, int q_lo, float q_w,
                         float max_val, cudaStream_t s) {
    return launch_threshold(true, x, out, dst, alpha, sigma, B, n, q_lo, q_w, max_val, s);
}
cudaError_t dynamic_threshold_inplace(float* x0, int B, int n, int q_lo, float q_w, float max_val, cudaStream_t s) {
    return launch_threshold(false, nullptr, nullptr, x0, 1.f, 1.f, B, n, q_lo, q_w, max_val, s);
}

// dst = (c1*x - c2*m) [+ c3*(m1 - m)], 128-bit vectorised; dst may alias x.
__global__ void solver_update_kernel(float* dst, const float* x, const float* m, const float* m1, float c1, float c2, float c3, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 X = reinterpret_cast<const float4*>(x)[i], M = reinterpret_cast<const float4*>(m)[i];
    float r[4] = {__fsub_rn(__fmul_rn(c1, X.x), __fmul_rn(c2, M.x)), __fsub_rn(__fmul_rn(c1, X.y), __fmul_rn(c2, M.y)),
                  __fsub_rn(__fmul_rn(c1, X.z), __fmul_rn(c2, M.z)), __fsub_rn(__fmul_rn(c1, X.w), __fmul_rn(c2, M.w))};
    if (m1) {
        const float4 N = reinterpret_cast<const float4*>(m1)[i];
        r[0] = __fadd_rn(r[0], __fmul_rn(c3, __fsub_rn(N.x, M.x)));
        r[1] = __fadd_rn(r[1], __fmul_rn(c3, __fsub_rn(N.y, M.y)));
        r[2] = __fadd_rn(r[2], __fmul_rn(c3, __fsub_rn(N.z, M.z)));
        r[3] = __fadd_rn(r[3], __fmul_rn(c3, __fsub_rn(N.w, M.w)));
    }
    reinterpret_cast<float4*>(dst)[i] = make_float4(r[0], r[1], r[2], r[3]);
}

cudaError_t solver_update(float* dst, const float* x, const float* m, const float* m1, float c1, float c2, float c3, size_t n,
                          cudaStream_t s) {
    const size_t n4 = n / 4;
    solver_update_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(dst, x, m, m1, c1, c2, c3, n4);
    return cudaGetLastError();
}

}  // namespace tmd
