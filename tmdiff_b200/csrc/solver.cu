// DPM-Solver++ device arithmetic: x_start -> eps -> x0 conversion with the exact per-sample
// 99.5% dynamic threshold, and the update family.  (core/dpm_solver_pytorch.py:296-306,
// :430-456, :589-592, :648-658, :744-760.)  Every operation is rounded exactly where the
// reference's fp32 tensor expression rounds (no FMA contraction), because x0 = (x - s*eps)/a
// divides by alpha_t ~ 5e-5 on the first steps and amplifies any re-association.
#include <float.h>

#include "kernels.h"

namespace tmd {

constexpr int TH_THREADS = 1024;

// noise_pred_fn (:296-306): the network output as eps.  noise_type: the output already is eps.
__device__ __forceinline__ float eps_of(float x, float out, float alpha, float sigma, bool noise_type) {
    return noise_type ? out : __fdiv_rn(__fsub_rn(x, __fmul_rn(alpha, out)), sigma);
}
__device__ __forceinline__ float x0_of_eps(float x, float eps, float alpha, float sigma) {
    return __fdiv_rn(__fsub_rn(x, __fmul_rn(sigma, eps)), alpha);              // data_prediction_fn (:447-456)
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
//
// out_c != nullptr: classifier-free guidance (model_wrapper :336-344) -- `out` is the unconditional
// half, `out_c` the conditional one, eps = eps_u + g*(eps_c - eps_u).
template <bool kFromModel>
__global__ void __launch_bounds__(TH_THREADS) x0_threshold_kernel(const float* x, const float* out, const float* out_c,
                                                                   float* dst, float alpha, float sigma, bool noise_type,
                                                                   float g, int n, int q_lo, float q_w, float max_val) {
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
            const float xv = x[base + i];
            float eps = eps_of(xv, out[base + i], alpha, sigma, noise_type);
            if (out_c) eps = __fadd_rn(eps, __fmul_rn(g, __fsub_rn(eps_of(xv, out_c[base + i], alpha, sigma, noise_type), eps)));
            v = x0_of_eps(xv, eps, alpha, sigma);
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

static cudaError_t launch_threshold(bool from_model, const float* x, const float* out, const float* out_c, float* dst, float alpha,
                                    float sigma, bool noise_type, float g, int B, int n, int q_lo, float q_w, float max_val,
                                    cudaStream_t s) {
    int P = 1;
    while (P < n) P <<= 1;
    const size_t smem = (size_t)P * sizeof(float);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;  // > 51200 values per sample: not supported by the in-smem select
    if (from_model) {
        cudaFuncSetAttribute(x0_threshold_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        x0_threshold_kernel<true><<<B, TH_THREADS, smem, s>>>(x, out, out_c, dst, alpha, sigma, noise_type, g, n, q_lo, q_w, max_val);
    } else {
        cudaFuncSetAttribute(x0_threshold_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        x0_threshold_kernel<false><<<B, TH_THREADS, smem, s>>>(nullptr, nullptr, nullptr, dst, 1.f, 1.f, false, 0.f, n, q_lo, q_w, max_val);
    }
    return cudaGetLastError();
}

cudaError_t x0_threshold(const float* x, const float* out, const float* out_c, float* dst, float alpha, float sigma, bool noise_type,
                         float g, int B, int n, int q_lo, float q_w, float max_val, cudaStream_t s) {
    return launch_threshold(true, x, out, out_c, dst, alpha, sigma, noise_type, g, B, n, q_lo, q_w, max_val, s);
}
cudaError_t dynamic_threshold_inplace(float* x0, int B, int n, int q_lo, float q_w, float max_val, cudaStream_t s) {
    return launch_threshold(false, nullptr, nullptr, nullptr, x0, 1.f, 1.f, false, 0.f, B, n, q_lo, q_w, max_val, s);
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

// Multistep updates (:817-927), dpmsolver++ / solver_type 'dpmsolver'; `m` newest model value.
//   form 1: D1_0 = k0*(m - m1);                      dst = c1*x - c2*m - c3*D1_0
//   form 2: D1_0 = k0*(m - m1); D1_1 = k1*(m1 - m2); D1 = D1_0 + k2*(D1_0 - D1_1); D2 = k3*(D1_0 - D1_1)
//           dst = c1*x - c2*m + c3*D1 - k4*D2
struct MultiCoef { float c1, c2, c3, k[5]; };
template <int kForm>
__global__ void solver_update_multistep_kernel(float* dst, const float* x, const float* m, const float* m1, const float* m2,
                                               MultiCoef q, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 X4 = reinterpret_cast<const float4*>(x)[i], M4 = reinterpret_cast<const float4*>(m)[i];
    const float4 N4 = reinterpret_cast<const float4*>(m1)[i];
    float4 O4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kForm == 2) O4 = reinterpret_cast<const float4*>(m2)[i];
    const float X[4] = {X4.x, X4.y, X4.z, X4.w}, M[4] = {M4.x, M4.y, M4.z, M4.w}, N[4] = {N4.x, N4.y, N4.z, N4.w},
                O[4] = {O4.x, O4.y, O4.z, O4.w};
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float d10 = __fmul_rn(q.k[0], __fsub_rn(M[j], N[j]));
        const float base = __fsub_rn(__fmul_rn(q.c1, X[j]), __fmul_rn(q.c2, M[j]));
        if (kForm == 1) {
            r[j] = __fsub_rn(base, __fmul_rn(q.c3, d10));
        } else {
            const float d11 = __fmul_rn(q.k[1], __fsub_rn(N[j], O[j]));
            const float dd = __fsub_rn(d10, d11);
            const float D1 = __fadd_rn(d10, __fmul_rn(q.k[2], dd));
            const float D2 = __fmul_rn(q.k[3], dd);
            r[j] = __fsub_rn(__fadd_rn(base, __fmul_rn(q.c3, D1)), __fmul_rn(q.k[4], D2));
        }
    }
    reinterpret_cast<float4*>(dst)[i] = make_float4(r[0], r[1], r[2], r[3]);
}

cudaError_t solver_update_multistep(int form, float* dst, const float* x, const float* m, const float* m1, const float* m2, float c1,
                                    float c2, float c3, const float* k, size_t n, cudaStream_t s) {
    const size_t n4 = n / 4;
    MultiCoef q{c1, c2, c3, {k[0], k[1], k[2], k[3], k[4]}};
    const unsigned grid = (unsigned)((n4 + 255) / 256);
    if (form == 1) solver_update_multistep_kernel<1><<<grid, 256, 0, s>>>(dst, x, m, m1, nullptr, q, n4);
    else if (form == 2) solver_update_multistep_kernel<2><<<grid, 256, 0, s>>>(dst, x, m, m1, m2, q, n4);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}


// Ancestral DDPM step (GeneralDiffusion.p_sample, diffusion_general.py:134-138, 155-174, 196-201) fused into one pass:
//   x0 = clamp(c_recip*x - c_recipm1*eps, -1, 1);  mean = coef1*x0 + coef2*x;  out = mean + sigma*z
// z: the caller's noise tensor (torch.randn_like in the reference) or, when noise == nullptr, Philox4x32-10 normals drawn in
// the kernel (counter = element index / 4, subsequence per call) -- no noise tensor in HBM.  Same operation order as the
// reference's torch expressions.
}  // namespace tmd
#include <curand_kernel.h>
namespace tmd {
__global__ void ancestral_update_kernel(float* out, const float* x, const float* eps, const float* noise, float c_recip, float c_recipm1,
                                        float coef1, float coef2, float sigma, int clip, unsigned long long seed,
                                        unsigned long long step, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 X4 = reinterpret_cast<const float4*>(x)[i], E4 = reinterpret_cast<const float4*>(eps)[i];
    float4 Z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sigma != 0.f) {
        if (noise) Z4 = reinterpret_cast<const float4*>(noise)[i];
        else {
            curandStatePhilox4_32_10_t st;
            curand_init(seed, /*subsequence*/ i, /*offset*/ step * 4ull, &st);
            Z4 = curand_normal4(&st);
        }
    }
    const float X[4] = {X4.x, X4.y, X4.z, X4.w}, E[4] = {E4.x, E4.y, E4.z, E4.w}, Z[4] = {Z4.x, Z4.y, Z4.z, Z4.w};
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float x0 = __fsub_rn(__fmul_rn(c_recip, X[j]), __fmul_rn(c_recipm1, E[j]));
        if (clip) x0 = fminf(fmaxf(x0, -1.f), 1.f);
        const float mean = __fadd_rn(__fmul_rn(coef1, x0), __fmul_rn(coef2, X[j]));
        r[j] = __fadd_rn(mean, __fmul_rn(Z[j], sigma));
    }
    reinterpret_cast<float4*>(out)[i] = make_float4(r[0], r[1], r[2], r[3]);
}

cudaError_t ancestral_update(float* out, const float* x, const float* eps, const float* noise, float c_recip, float c_recipm1, float coef1,
                             float coef2, float sigma, bool clip, unsigned long long seed, unsigned long long step, size_t n,
                             cudaStream_t s) {
    if (n % 4) return cudaErrorInvalidValue;
    const size_t n4 = n / 4;
    ancestral_update_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(out, x, eps, noise, c_recip, c_recipm1, coef1, coef2, sigma, clip ? 1 : 0,
                                                                        seed, step, n4);
    return cudaGetLastError();
}

}  // namespace tmd
