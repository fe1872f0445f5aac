// DPM-Solver++ device arithmetic: x_start -> eps -> x0 conversion with the exact per-sample
// 99.5% dynamic threshold, and the update family.  (core/dpm_solver_pytorch.py:296-306,
// :430-456, :589-592, :648-658, :744-760.)  Every operation is rounded exactly where the
// reference's fp32 tensor expression rounds (no FMA contraction), because x0 = (x - s*eps)/a
// divides by alpha_t ~ 5e-5 on the first steps and amplifies any re-association.
#include <float.h>
#include <stdint.h>

#include <mutex>

#include "../../include/tmdiff_b200.h"
#include "kernels.h"

namespace tmd {

// noise_pred_fn (:296-306): the network output as eps.  noise_type: the output already is eps.
__device__ __forceinline__ float eps_of(float x, float out, float alpha, float sigma, bool noise_type) {
    return noise_type ? out : __fdiv_rn(__fsub_rn(x, __fmul_rn(alpha, out)), sigma);
}
__device__ __forceinline__ float x0_of_eps(float x, float eps, float alpha, float sigma) {
    return __fdiv_rn(__fsub_rn(x, __fmul_rn(sigma, eps)), alpha);              // data_prediction_fn (:447-456)
}

// ------------------------------------------------------------------------------------------------
// dynamic_thresholding_fn (:430-439): s = max(quantile(|x0|, q), max_val), x0 = clamp(x0, -s, s) / s
//
// torch.quantile interpolates linearly between the sorted positions q_lo and q_lo + 1, so two exact
// order statistics per sample are needed.  They are found by a three-level radix select over the
// bit pattern of |x0| (a non-negative float orders like its bits): 10 + 11 + 10 bits, one histogram
// pass per level over the (L2-resident) x0, several CTAs per sample so that a batch of 64 patches
// fills the machine and a patch of any size fits (no shared-memory copy of the sample).
//
//   th_first   x0 from (x, out) -> dst, histogram of bits [30:21]                    (level 0)
//   th_level1  picks the level-0 bins of both ranks, histograms bits [20:10] of the values in them
//   th_level2  picks the level-1 bins, histograms bits [9:0]
//   th_apply   picks the level-2 bins = the exact values, interpolates as at::lerp does, clamps and
//              divides; optionally also evaluates the solver update that consumes this x0
//              (dpmsolver++ singlestep family, :589-592, :648-658, :744-760) in the same pass.
//
// Workspace per sample: TH_WS_INTS ints = hist0 | hist1[2] | hist2[2] | sel[8].  hist0 is zeroed by a
// memset node in front of th_first, which zeroes hist1 / hist2 itself.
// ------------------------------------------------------------------------------------------------
constexpr int TH_THREADS = 256;
constexpr int TH_BINS = 2048;
constexpr int TH_WS_INTS = 5 * TH_BINS + 8;

__device__ __forceinline__ uint32_t th_key(float v) { return __float_as_uint(fabsf(v)); }

// Block-wide: the bin holding the k-th largest entry (k >= 1) of a TH_BINS histogram in shared memory, and k's rank inside
// that bin.  All threads get the result.  scratch: 16 ints of shared memory.
__device__ void th_select(const int* h, int k, int& bin, int& res, int* scratch) {
    constexpr int PER = TH_BINS / TH_THREADS;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    int c[PER], sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { c[i] = h[TH_BINS - 1 - (t * PER + i)]; sum += c[i]; }   // thread t: bins in descending order
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    __syncthreads();                       // scratch may still be read from a previous call
    if (lane == 31) scratch[w] = incl;
    __syncthreads();
    int before_w = 0;
    for (int i = 0; i < w; ++i) before_w += scratch[i];
    incl += before_w;
    const int before = incl - sum;
    if (before < k && k <= incl) {
        int acc = before;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            if (acc < k && k <= acc + c[i]) { scratch[8] = TH_BINS - 1 - (t * PER + i); scratch[9] = k - acc; }
            acc += c[i];
        }
    }
    __syncthreads();
    bin = scratch[8];
    res = scratch[9];
}

__device__ __forceinline__ void th_load_hist(int* dst, const int* src) {
    for (int i = threadIdx.x; i < TH_BINS; i += TH_THREADS) dst[i] = src[i];
}
__device__ __forceinline__ void th_flush_hist(const int* src, int* dst) {
    for (int i = threadIdx.x; i < TH_BINS; i += TH_THREADS) {
        const int v = src[i];
        if (v) atomicAdd(dst + i, v);
    }
}
// this CTA's share [lo, hi) of a sample of n values; multiples of 4 when n is one
__device__ __forceinline__ void th_range(int n, int& lo, int& hi) {
    const int S = gridDim.x;
    int per = (n + S - 1) / S;
    per = (per + 3) & ~3;
    lo = min(n, (int)blockIdx.x * per);
    hi = min(n, lo + per);
}

template <bool kFromModel>
__global__ void __launch_bounds__(TH_THREADS) th_first_kernel(const float* __restrict__ x, const float* __restrict__ out,
                                                               const float* __restrict__ out_c, float* __restrict__ dst, float alpha,
                                                               float sigma, bool noise_type, float g, int n, int* __restrict__ ws) {
    __shared__ int h[TH_BINS];
    const int b = blockIdx.y;
    const size_t base = (size_t)b * n;
    int* wsb = ws + (size_t)b * TH_WS_INTS;
    for (int i = threadIdx.x; i < TH_BINS; i += TH_THREADS) h[i] = 0;
    {   // zero the deeper histograms of this sample (4 * TH_BINS ints, split over the sample's CTAs)
        const int tot = 4 * TH_BINS, per = (tot + gridDim.x - 1) / gridDim.x;
        const int z0 = blockIdx.x * per, z1 = min(tot, z0 + per);
        for (int i = z0 + threadIdx.x; i < z1; i += TH_THREADS) wsb[TH_BINS + i] = 0;
    }
    __syncthreads();
    int lo, hi;
    th_range(n, lo, hi);
    auto one = [&](float xv, float ov, float ocv) {
        float eps = eps_of(xv, ov, alpha, sigma, noise_type);
        if (out_c) eps = __fadd_rn(eps, __fmul_rn(g, __fsub_rn(eps_of(xv, ocv, alpha, sigma, noise_type), eps)));
        return x0_of_eps(xv, eps, alpha, sigma);
    };
    if ((n & 3) == 0) {
        for (int i = lo + 4 * threadIdx.x; i < hi; i += 4 * TH_THREADS) {
            float4 v;
            if (kFromModel) {
                const float4 X = *reinterpret_cast<const float4*>(x + base + i), O = *reinterpret_cast<const float4*>(out + base + i);
                float4 Oc = make_float4(0.f, 0.f, 0.f, 0.f);
                if (out_c) Oc = *reinterpret_cast<const float4*>(out_c + base + i);
                v = make_float4(one(X.x, O.x, Oc.x), one(X.y, O.y, Oc.y), one(X.z, O.z, Oc.z), one(X.w, O.w, Oc.w));
                *reinterpret_cast<float4*>(dst + base + i) = v;
            } else {
                v = *reinterpret_cast<const float4*>(dst + base + i);
            }
            atomicAdd(&h[th_key(v.x) >> 21], 1);
            atomicAdd(&h[th_key(v.y) >> 21], 1);
            atomicAdd(&h[th_key(v.z) >> 21], 1);
            atomicAdd(&h[th_key(v.w) >> 21], 1);
        }
    } else {
        for (int i = lo + threadIdx.x; i < hi; i += TH_THREADS) {
            float v;
            if (kFromModel) {
                v = one(x[base + i], out[base + i], out_c ? out_c[base + i] : 0.f);
                dst[base + i] = v;
            } else {
                v = dst[base + i];
            }
            atomicAdd(&h[th_key(v) >> 21], 1);
        }
    }
    __syncthreads();
    th_flush_hist(h, wsb);
}

// kLevel 1: select in hist0, histogram bits [20:10] of the values whose bits [30:21] match; kLevel 2: select in hist1,
// histogram bits [9:0] of the values whose bits [30:10] match.  The two ranks (sorted positions q_lo + 1 and q_lo) share
// a histogram while their prefixes agree.
template <int kLevel>
__global__ void __launch_bounds__(TH_THREADS) th_level_kernel(const float* __restrict__ dst, int n, int k_hi, int k_lo, int* __restrict__ ws) {
    __shared__ int hA[TH_BINS], hB[TH_BINS];
    __shared__ int scratch[16];
    const int b = blockIdx.y;
    const size_t base = (size_t)b * n;
    int* wsb = ws + (size_t)b * TH_WS_INTS;
    int* sel = wsb + 5 * TH_BINS;
    uint32_t pA, pB;       // prefixes the values must match (compared after the shift below)
    int rA, rB;            // remaining ranks inside those prefixes
    if (kLevel == 1) {
        th_load_hist(hA, wsb);
        __syncthreads();
        int bA, bB;
        th_select(hA, k_hi, bA, rA, scratch);
        th_select(hA, k_lo, bB, rB, scratch);
        pA = (uint32_t)bA; pB = (uint32_t)bB;
        if (blockIdx.x == 0 && threadIdx.x == 0) { sel[0] = bA; sel[1] = rA; sel[2] = bB; sel[3] = rB; }
    } else {
        const int bA0 = sel[0], rA0 = sel[1], bB0 = sel[2], rB0 = sel[3];
        th_load_hist(hA, wsb + TH_BINS);
        if (bB0 != bA0) th_load_hist(hB, wsb + 2 * TH_BINS);
        __syncthreads();
        int bA, bB;
        th_select(hA, rA0, bA, rA, scratch);
        th_select(bB0 != bA0 ? hB : hA, rB0, bB, rB, scratch);
        pA = ((uint32_t)bA0 << 11) | (uint32_t)bA;
        pB = ((uint32_t)bB0 << 11) | (uint32_t)bB;
        if (blockIdx.x == 0 && threadIdx.x == 0) { sel[4] = (int)pA; sel[5] = rA; sel[6] = (int)pB; sel[7] = rB; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < TH_BINS; i += TH_THREADS) { hA[i] = 0; hB[i] = 0; }
    __syncthreads();
    constexpr int kShift = kLevel == 1 ? 21 : 10;
    constexpr uint32_t kMask = kLevel == 1 ? 2047u : 1023u;
    constexpr int kSub = kLevel == 1 ? 10 : 0;
    const bool two = pA != pB;
    auto one = [&](float v) {
        const uint32_t key = th_key(v), top = key >> kShift;
        if (top == pA) atomicAdd(&hA[(key >> kSub) & kMask], 1);
        else if (two && top == pB) atomicAdd(&hB[(key >> kSub) & kMask], 1);
    };
    int lo, hi;
    th_range(n, lo, hi);
    if ((n & 3) == 0) {
        for (int i = lo + 4 * threadIdx.x; i < hi; i += 4 * TH_THREADS) {
            const float4 v = *reinterpret_cast<const float4*>(dst + base + i);
            one(v.x); one(v.y); one(v.z); one(v.w);
        }
    } else {
        for (int i = lo + threadIdx.x; i < hi; i += TH_THREADS) one(dst[base + i]);
    }
    __syncthreads();
    int* g = wsb + (kLevel == 1 ? 1 : 3) * TH_BINS;
    th_flush_hist(hA, g);
    if (two) th_flush_hist(hB, g + TH_BINS);
    if (kLevel == 2) {   // hist0 has been consumed by level 1: leave it zeroed for the next call on this workspace (no memset node)
        const int per = (TH_BINS + gridDim.x - 1) / gridDim.x;
        const int z0 = blockIdx.x * per, z1 = min(TH_BINS, z0 + per);
        for (int i = z0 + threadIdx.x; i < z1; i += TH_THREADS) wsb[i] = 0;
    }
}

// The dpmsolver++ singlestep update that follows an evaluation, evaluated by th_apply on the x0 it has in registers:
//   upd.dst = (c1*x - c2*m) [+ c3*(m1 - m)]      (same rounding as solver_update_kernel)
// m / m1 equal to the threshold's dst mean "the fresh x0".  dst == nullptr: no fused update.
struct ThUpdate {
    float* dst;
    const float* x;
    const float* m;
    const float* m1;
    float c1, c2, c3;
};

__global__ void __launch_bounds__(TH_THREADS) th_apply_kernel(float* dst, int n, int q_pair, float q_w, float max_val,
                                                               const int* __restrict__ ws, ThUpdate upd) {
    __shared__ int hA[TH_BINS], hB[TH_BINS];
    __shared__ int scratch[16];
    const int b = blockIdx.y;
    const size_t base = (size_t)b * n;
    const int* wsb = ws + (size_t)b * TH_WS_INTS;
    const int* sel = wsb + 5 * TH_BINS;
    const uint32_t pA = (uint32_t)sel[4], pB = (uint32_t)sel[6];
    const int rA = sel[5], rB = sel[7];
    th_load_hist(hA, wsb + 3 * TH_BINS);
    if (pB != pA) th_load_hist(hB, wsb + 4 * TH_BINS);
    __syncthreads();
    int bA, bB, dummy;
    th_select(hA, rA, bA, dummy, scratch);
    th_select(pB != pA ? hB : hA, rB, bB, dummy, scratch);
    const float v_hi = __uint_as_float((pA << 10) | (uint32_t)bA);     // sorted position q_lo + 1
    const float v_lo = q_pair ? __uint_as_float((pB << 10) | (uint32_t)bB) : v_hi;
    // at::lerp: weight < 0.5 ? a + w*(b-a) : b - (b-a)*(1-w)
    const float diff = __fsub_rn(v_hi, v_lo);
    const float qv = (fabsf(q_w) < 0.5f) ? __fadd_rn(v_lo, __fmul_rn(q_w, diff)) : __fsub_rn(v_hi, __fmul_rn(diff, __fsub_rn(1.f, q_w)));
    const float s = fmaxf(qv, max_val);
    const bool fuse = upd.dst != nullptr;
    const bool m_fresh = upd.m == dst, m1_fresh = upd.m1 == dst;
    auto one = [&](float v, size_t gi) -> float {
        const float r = __fdiv_rn(fminf(fmaxf(v, -s), s), s);   // clamp(x0, -s, s) / s
        if (fuse) {
            const float mv = m_fresh ? r : upd.m[gi];
            float u = __fsub_rn(__fmul_rn(upd.c1, upd.x[gi]), __fmul_rn(upd.c2, mv));
            if (upd.m1) u = __fadd_rn(u, __fmul_rn(upd.c3, __fsub_rn(m1_fresh ? r : upd.m1[gi], mv)));
            upd.dst[gi] = u;
        }
        return r;
    };
    int lo, hi;
    th_range(n, lo, hi);
    if ((n & 3) == 0) {
        for (int i = lo + 4 * threadIdx.x; i < hi; i += 4 * TH_THREADS) {
            float4 v = *reinterpret_cast<const float4*>(dst + base + i);
            const size_t gi = base + i;
            float r[4] = {0.f, 0.f, 0.f, 0.f}, u[4];
            const float vv[4] = {v.x, v.y, v.z, v.w};
            if (fuse) {
                const float4 X = *reinterpret_cast<const float4*>(upd.x + gi);
                float4 M = make_float4(0.f, 0.f, 0.f, 0.f), M1 = M;
                if (!m_fresh) M = *reinterpret_cast<const float4*>(upd.m + gi);
                if (upd.m1 && !m1_fresh) M1 = *reinterpret_cast<const float4*>(upd.m1 + gi);
                const float xs[4] = {X.x, X.y, X.z, X.w}, ms[4] = {M.x, M.y, M.z, M.w}, m1s[4] = {M1.x, M1.y, M1.z, M1.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    r[j] = __fdiv_rn(fminf(fmaxf(vv[j], -s), s), s);
                    const float mv = m_fresh ? r[j] : ms[j];
                    u[j] = __fsub_rn(__fmul_rn(upd.c1, xs[j]), __fmul_rn(upd.c2, mv));
                    if (upd.m1) u[j] = __fadd_rn(u[j], __fmul_rn(upd.c3, __fsub_rn(m1_fresh ? r[j] : m1s[j], mv)));
                }
                *reinterpret_cast<float4*>(upd.dst + gi) = make_float4(u[0], u[1], u[2], u[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) r[j] = __fdiv_rn(fminf(fmaxf(vv[j], -s), s), s);
            }
            *reinterpret_cast<float4*>(dst + gi) = make_float4(r[0], r[1], r[2], r[3]);
        }
    } else {
        for (int i = lo + threadIdx.x; i < hi; i += TH_THREADS) dst[base + i] = one(dst[base + i], base + i);
    }
}

size_t threshold_workspace_bytes(int B) { return (size_t)B * TH_WS_INTS * sizeof(int); }

static cudaError_t launch_threshold(bool from_model, const float* x, const float* out, const float* out_c, float* dst, float alpha,
                                    float sigma, bool noise_type, float g, int B, int n, int q_lo, float q_w, float max_val, void* ws,
                                    bool ws_clean, const ThUpdate& upd, cudaStream_t s) {
    if (!ws || B <= 0 || n <= 0) return cudaErrorInvalidValue;
    int* w = reinterpret_cast<int*>(ws);
    // CTAs per sample: two waves of the machine over the batch, at least 2048 values each
    int S = (2 * 148 + B - 1) / B;
    const int max_s = (n + 2047) / 2048;
    if (S > max_s) S = max_s;
    if (S < 1) S = 1;
    const dim3 grid((unsigned)S, (unsigned)B);
    if (!ws_clean) {     // hist0 of every sample (a workspace that went through a call before is left clean by its level-2 pass)
        cudaError_t e = cudaMemset2DAsync(w, TH_WS_INTS * sizeof(int), 0, TH_BINS * sizeof(int), (size_t)B, s);
        if (e != cudaSuccess) return e;
    }
    if (from_model) th_first_kernel<true><<<grid, TH_THREADS, 0, s>>>(x, out, out_c, dst, alpha, sigma, noise_type, g, n, w);
    else th_first_kernel<false><<<grid, TH_THREADS, 0, s>>>(nullptr, nullptr, nullptr, dst, 1.f, 1.f, false, 0.f, n, w);
    const int hi_pos = q_lo + 1 < n - 1 ? q_lo + 1 : n - 1;
    const int k_hi = n - hi_pos, k_lo = n - q_lo;      // sorted position p is the (n - p)-th largest
    th_level_kernel<1><<<grid, TH_THREADS, 0, s>>>(dst, n, k_hi, k_lo, w);
    th_level_kernel<2><<<grid, TH_THREADS, 0, s>>>(dst, n, k_hi, k_lo, w);
    th_apply_kernel<<<grid, TH_THREADS, 0, s>>>(dst, n, k_hi != k_lo ? 1 : 0, q_w, max_val, w, upd);
    return cudaGetLastError();
}

cudaError_t x0_threshold(const float* x, const float* out, const float* out_c, float* dst, float alpha, float sigma, bool noise_type,
                         float g, int B, int n, int q_lo, float q_w, float max_val, void* ws, bool ws_clean, const FusedUpdate* fu, cudaStream_t s) {
    ThUpdate u{};
    if (fu) { u.dst = fu->dst; u.x = fu->x; u.m = fu->m; u.m1 = fu->m1; u.c1 = fu->c1; u.c2 = fu->c2; u.c3 = fu->c3; }
    return launch_threshold(true, x, out, out_c, dst, alpha, sigma, noise_type, g, B, n, q_lo, q_w, max_val, ws, ws_clean, u, s);
}
cudaError_t dynamic_threshold_inplace(float* x0, int B, int n, int q_lo, float q_w, float max_val, void* ws, cudaStream_t s) {
    return launch_threshold(false, nullptr, nullptr, nullptr, x0, 1.f, 1.f, false, 0.f, B, n, q_lo, q_w, max_val, ws, false, ThUpdate{}, s);
}

// ------------------------------------------------------------------------------------------------
// tmd_dynamic_threshold_static: the same four kernels with the radix-select workspace in a file-scope device array instead
// of a stream-ordered allocation, for callers that may not allocate (graph capture, tight serving loops).  No counterpart in
// the reference.
//
// The array holds the workspace of TMD_THRESHOLD_STATIC_MAX_BATCH samples (64 * 10248 ints = 2 623 488 bytes) on every GPU
// that loads the library, until exit.  In host code `g_th_static_ws` is only the host-side stand-in of the symbol; its device
// address is looked up once per device with cudaGetSymbolAddress (that device current) and that pointer goes to the kernels.
// Every call shares the array and finds the previous call's contents in it: hist0 of the B samples is cleared by a kernel in
// front (th_first clears hist1 / hist2 itself, sel is written before it is read).  Two calls that overlap on the GPU -- two
// streams, or two host threads -- race on it; callers serialise them.
// ------------------------------------------------------------------------------------------------
constexpr int kThStaticMaxBatch = TMD_THRESHOLD_STATIC_MAX_BATCH;
static __device__ __align__(256) int g_th_static_ws[(size_t)kThStaticMaxBatch * TH_WS_INTS];

// device code names the array directly
__global__ void __launch_bounds__(TH_THREADS) th_static_clear_kernel() {
    int* h0 = g_th_static_ws + (size_t)blockIdx.x * TH_WS_INTS;
    for (int i = threadIdx.x; i < TH_BINS; i += TH_THREADS) h0[i] = 0;
}

// Device address of g_th_static_ws on the current device (looked up once per device).
static cudaError_t th_static_ws_address(int** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static int* addr[kMaxDevices] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_th_static_ws);
        if (e != cudaSuccess) return e;
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) return cudaErrorInvalidSymbol;
        addr[dev] = static_cast<int*>(p);
    }
    *out = addr[dev];
    return cudaSuccess;
}

}  // namespace tmd

extern "C" int tmd_dynamic_threshold_static(float* x0, int32_t B, int32_t n, int32_t q_lo, float q_weight, float max_val, void* stream) {
    using namespace tmd;
    if (!x0) return set_last_error("tmd_dynamic_threshold_static: null argument");
    if (B <= 0 || n <= 0 || q_lo < 0 || q_lo >= n) return set_last_error("tmd_dynamic_threshold_static: bad sizes");
    if (B > kThStaticMaxBatch)
        return set_last_error("tmd_dynamic_threshold_static: B = %d exceeds the static workspace (TMD_THRESHOLD_STATIC_MAX_BATCH = %d samples); "
                              "split the batch or use tmd_dynamic_threshold",
                              (int)B, kThStaticMaxBatch);
    cudaStream_t s = (cudaStream_t)stream;
    int* ws = nullptr;
    cudaError_t e = th_static_ws_address(&ws);
    if (e != cudaSuccess) return set_last_error("tmd_dynamic_threshold_static: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
    th_static_clear_kernel<<<(unsigned)B, TH_THREADS, 0, s>>>();
    e = cudaGetLastError();
    if (e == cudaSuccess)
        e = launch_threshold(false, nullptr, nullptr, nullptr, x0, 1.f, 1.f, false, 0.f, B, n, q_lo, q_weight, max_val, ws, /*ws_clean=*/true,
                             ThUpdate{}, s);
    if (e != cudaSuccess) return set_last_error("tmd_dynamic_threshold_static: %s", cudaGetErrorString(e));
    return 0;
}

namespace tmd {

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
