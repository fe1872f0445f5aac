// Stand-alone operators named by the north star but not executed on the reference's sampling path (SURVEY.md 8(a)
// n1-n4): group normalisation with (adaptive) affine + Swish, and fused softmax attention.  fp32 in the reference
// layout; these are the per-operator sweep's kernels, not part of patches/s.
//
//   GroupNorm / AdaptiveGroupNorm:  torch.nn.GroupNorm(G, C, eps) over [B, C, S] followed by
//       y = gamma[b?, c] * x_hat + beta[b?, c]   (+ Swish)
//     core/Attention.py:108-109 (Normalize), GeneralModel/Hyper_unet_general.py:311-331 (AdaptiveGroupNorm),
//     :487-488 (AttnBlockpp.GroupNorm_0).  Memory-bound: one statistics pass + one apply pass, each group split over
//     several CTAs so small batches still fill the machine; warp-shuffle reductions, Chan's merge of the partials.
//
//   attention:  out[bh, i, :] = softmax_j(scale * q[bh, i, :] . k[bh, j, :]) v[bh, j, :]
//     core/Attention.py:140-159 (SpatialSelfAttention: one head of width C), :185-214 (CrossAttention: 8 heads of 64
//     against 77 text tokens), Hyper_unet_general.py:500-506 (AttnBlockpp).  One pass over the keys with a running
//     row maximum / sum (no [Lq, Lk] matrix in HBM), any head width up to 2048.
#include <cooperative_groups.h>
#include <float.h>
#include <stdlib.h>

#include "kernels.h"

namespace cg = cooperative_groups;

namespace tmd {

namespace {

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------
// group norm
// ------------------------------------------------------------------------------------------------
constexpr int GN_THREADS = 256;

// partial[(bg * split + s)] = {count, mean, M2} of slice s of group bg  (group = n contiguous floats)
__global__ void __launch_bounds__(GN_THREADS) gn_stats_kernel(const float* __restrict__ x, float* __restrict__ partial, size_t n, int split) {
    const size_t bg = blockIdx.x;
    const int s = blockIdx.y;
    const size_t per = ((n + split - 1) / split + 3) & ~(size_t)3;
    const size_t lo = (size_t)s * per, hi = lo + per < n ? lo + per : n;
    const float* p = x + bg * n;
    __shared__ float red[2][GN_THREADS / 32];
    float cnt = 0.f, mean = 0.f, m2 = 0.f;
    if (lo < hi) {
        const float shift = p[lo];                      // shifted sums: no cancellation for data with a large mean
        float a = 0.f, q = 0.f;
        if ((n & 3) == 0) {
            for (size_t i = lo + 4 * threadIdx.x; i < hi; i += 4 * GN_THREADS) {
                const float4 v = *reinterpret_cast<const float4*>(p + i);
                const float d0 = v.x - shift, d1 = v.y - shift, d2 = v.z - shift, d3 = v.w - shift;
                a += (d0 + d1) + (d2 + d3);
                q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
            }
        } else {
            for (size_t i = lo + threadIdx.x; i < hi; i += GN_THREADS) {
                const float d = p[i] - shift;
                a += d;
                q += d * d;
            }
        }
        a = warp_sum_f(a);
        q = warp_sum_f(q);
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = q; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float A = 0.f, Q = 0.f;
            for (int w = 0; w < GN_THREADS / 32; ++w) { A += red[0][w]; Q += red[1][w]; }
            cnt = (float)(hi - lo);
            mean = shift + A / cnt;
            m2 = fmaxf(Q - A * A / cnt, 0.f);
        }
    }
    if (threadIdx.x == 0) {
        float* o = partial + (bg * split + s) * 3;
        o[0] = cnt; o[1] = mean; o[2] = m2;
    }
}

// y = gamma * (x - mean) * rstd + beta (+ swish);  gamma/beta indexed [b * gb_stride + c] (gb_stride = 0: shared affine)
__global__ void __launch_bounds__(GN_THREADS) gn_apply_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                              const float* __restrict__ partial, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, int gb_stride, int G, int cpg, size_t S,
                                                              int split, float eps, int act) {
    const size_t bg = blockIdx.x;
    const int s = blockIdx.y;
    const size_t n = (size_t)cpg * S;
    // Chan's pairwise merge of the slices' (count, mean, M2)
    float cnt = 0.f, mean = 0.f, m2 = 0.f;
    for (int i = 0; i < split; ++i) {
        const float* q = partial + (bg * split + i) * 3;
        const float nb = q[0];
        if (nb == 0.f) continue;
        const float d = q[1] - mean, tot = cnt + nb;
        mean += d * nb / tot;
        m2 += q[2] + d * d * cnt * nb / tot;
        cnt = tot;
    }
    const float rstd = rsqrtf(m2 / cnt + eps);
    const int b = (int)(bg / G), g = (int)(bg % G);
    const size_t per = ((n + split - 1) / split + 3) & ~(size_t)3;
    const size_t lo = (size_t)s * per, hi = lo + per < n ? lo + per : n;
    const float* p = x + bg * n;
    float* o = y + bg * n;
    auto one = [&](size_t i, float v) {
        const int c = g * cpg + (int)(i / S);
        float r = (v - mean) * rstd;
        const float ga = gamma ? gamma[(size_t)b * gb_stride + c] : 1.f, be = beta ? beta[(size_t)b * gb_stride + c] : 0.f;
        r = ga * r + be;
        if (act) r = r * (1.f / (1.f + expf(-r)));
        return r;
    };
    if ((S & 3) == 0) {
        for (size_t i = lo + 4 * threadIdx.x; i < hi; i += 4 * GN_THREADS) {
            const float4 v = *reinterpret_cast<const float4*>(p + i);
            *reinterpret_cast<float4*>(o + i) = make_float4(one(i, v.x), one(i + 1, v.y), one(i + 2, v.z), one(i + 3, v.w));
        }
    } else {
        for (size_t i = lo + threadIdx.x; i < hi; i += GN_THREADS) o[i] = one(i, p[i]);
    }
}

// Single-read variant: a thread-block cluster owns one group, every CTA keeps its slice of the group in shared memory
// between the statistics and the apply phase, and the slices' (count, mean, M2) are exchanged through distributed
// shared memory.  HBM traffic = one read + one write (the algorithmic bytes).  Needs n % 4 == 0 and S % 4 == 0.
__global__ void __launch_bounds__(GN_THREADS) gn_cluster_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                int gb_stride, int G, int cpg, size_t S, size_t slice, float eps, int act) {
    extern __shared__ __align__(16) float buf[];
    __shared__ float part_all[8][4];      // every rank's (count, mean, M2), pushed here by the owners
    __shared__ float red[2][GN_THREADS / 32];
    __shared__ float stat[2];
    cg::cluster_group cl = cg::this_cluster();
    const unsigned r = cl.block_rank(), CS = cl.num_blocks();
    const size_t bg = blockIdx.x / CS;
    const size_t n = (size_t)cpg * S;
    const size_t lo = (size_t)r * slice, hi = lo + slice < n ? lo + slice : n;
    const float* p = x + bg * n;
    float a = 0.f, q = 0.f;
    const float shift = lo < hi ? p[lo] : 0.f;
    {
        // four 16-byte loads in flight per thread before the first use
        const unsigned cnt = lo < hi ? (unsigned)(hi - lo) : 0u;
        const float* ps = p + lo;
        unsigned i = 4 * threadIdx.x;
        for (; i + 12 * GN_THREADS < cnt; i += 16 * GN_THREADS) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldcs(reinterpret_cast<const float4*>(ps + i + u * 4 * GN_THREADS));
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                *reinterpret_cast<float4*>(buf + i + u * 4 * GN_THREADS) = v[u];
                const float d0 = v[u].x - shift, d1 = v[u].y - shift, d2 = v[u].z - shift, d3 = v[u].w - shift;
                a += (d0 + d1) + (d2 + d3);
                q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
            }
        }
        for (; i < cnt; i += 4 * GN_THREADS) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(ps + i));
            *reinterpret_cast<float4*>(buf + i) = v;
            const float d0 = v.x - shift, d1 = v.y - shift, d2 = v.z - shift, d3 = v.w - shift;
            a += (d0 + d1) + (d2 + d3);
            q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
        }
    }
    a = warp_sum_f(a);
    q = warp_sum_f(q);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = a; red[1][threadIdx.x >> 5] = q; }
    __syncthreads();
    if (threadIdx.x < CS) {               // thread t pushes this CTA's partial into rank t's shared memory (DSMEM store)
        float A = 0.f, Q = 0.f;
        for (int w = 0; w < GN_THREADS / 32; ++w) { A += red[0][w]; Q += red[1][w]; }
        const float cnt = lo < hi ? (float)(hi - lo) : 0.f;
        float* dst = cl.map_shared_rank(&part_all[r][0], threadIdx.x);
        dst[0] = cnt;
        dst[1] = cnt > 0.f ? shift + A / cnt : 0.f;
        dst[2] = cnt > 0.f ? fmaxf(Q - A * A / cnt, 0.f) : 0.f;
    }
    cl.sync();                            // release/acquire at cluster scope: all partials are now local
    if (threadIdx.x == 0) {               // every CTA merges the same partials in the same order: identical statistics
        float cnt = 0.f, mean = 0.f, m2 = 0.f;
        for (unsigned i = 0; i < CS; ++i) {
            const float nb = part_all[i][0];
            if (nb == 0.f) continue;
            const float d = part_all[i][1] - mean, tot = cnt + nb;
            mean += d * nb / tot;
            m2 += part_all[i][2] + d * d * cnt * nb / tot;
            cnt = tot;
        }
        stat[0] = mean;
        stat[1] = rsqrtf(m2 / cnt + eps);
    }
    __syncthreads();
    const float mean = stat[0], rstd = stat[1];
    const int b = (int)(bg / G), g = (int)(bg % G);
    float* o = y + bg * n;
    // 32-bit indices inside the group (n < 2^31): the channel of a float4 costs one 32-bit division
    const unsigned lo32 = (unsigned)lo, cnt32 = (unsigned)(hi - lo), S32 = (unsigned)S;
    const float* gp = gamma ? gamma + (size_t)b * gb_stride + g * cpg : nullptr;
    const float* bp = beta ? beta + (size_t)b * gb_stride + g * cpg : nullptr;
    for (unsigned i = 4 * threadIdx.x; i < cnt32; i += 4 * GN_THREADS) {
        const float4 v = *reinterpret_cast<const float4*>(buf + i);
        const unsigned c = (lo32 + i) / S32;            // S % 4 == 0: the four values share a channel
        const float ga = (gp ? gp[c] : 1.f) * rstd, be = (bp ? bp[c] : 0.f);
        float rr[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t = fmaf(ga, rr[j] - mean, be);
            if (act) t = __fdividef(t, 1.f + __expf(-t));
            rr[j] = t;
        }
        __stcs(reinterpret_cast<float4*>(o + lo + i), make_float4(rr[0], rr[1], rr[2], rr[3]));
    }
    // no trailing cluster barrier: after the one above nobody touches a peer's shared memory again
}

// ------------------------------------------------------------------------------------------------
// fused softmax attention (fp32): 8 queries per CTA (one per warp for the row statistics), keys in tiles of 32,
// the head dimension in chunks of 64 for q.k and striped over the 256 threads for p.v
// ------------------------------------------------------------------------------------------------
constexpr int AT_BQ = 8, AT_BK = 32, AT_DC = 64, AT_THREADS = 256, AT_MAXDPT = 8;   // head width <= 2048

template <int DPT>
__global__ void __launch_bounds__(AT_THREADS) attention_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                               const float* __restrict__ v, float* __restrict__ out, int Lq, int Lk,
                                                               int D, float scale) {
    __shared__ float Qs[AT_BQ][AT_DC];
    __shared__ float Ks[AT_BK][AT_DC + 1];
    __shared__ float Ps[AT_BQ][AT_BK];
    __shared__ float alpha_s[AT_BQ], l_s[AT_BQ];
    const int bh = blockIdx.y, q0 = blockIdx.x * AT_BQ;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const float* qb = q + ((size_t)bh * Lq + q0) * D;
    const float* kb = k + (size_t)bh * Lk * D;
    const float* vb = v + (size_t)bh * Lk * D;
    float acc[AT_BQ][DPT];
#pragma unroll
    for (int i = 0; i < AT_BQ; ++i)
#pragma unroll
        for (int c = 0; c < DPT; ++c) acc[i][c] = 0.f;
    float m_run = -FLT_MAX, l_run = 0.f;            // row w's statistics (replicated across the warp)
    for (int j0 = 0; j0 < Lk; j0 += AT_BK) {
        // ---- s[w][lane] = scale * q[q0+w] . k[j0+lane]
        float sacc = 0.f;
        for (int d0 = 0; d0 < D; d0 += AT_DC) {
            __syncthreads();
            for (int e = tid; e < AT_BQ * AT_DC; e += AT_THREADS) {
                const int i = e / AT_DC, d = e % AT_DC;
                Qs[i][d] = (q0 + i < Lq && d0 + d < D) ? qb[(size_t)i * D + d0 + d] : 0.f;
            }
            for (int e = tid; e < AT_BK * AT_DC; e += AT_THREADS) {
                const int j = e / AT_DC, d = e % AT_DC;
                Ks[j][d] = (j0 + j < Lk && d0 + d < D) ? kb[(size_t)(j0 + j) * D + d0 + d] : 0.f;
            }
            __syncthreads();
#pragma unroll 16
            for (int d = 0; d < AT_DC; ++d) sacc = fmaf(Qs[w][d], Ks[lane][d], sacc);
        }
        const bool live = j0 + lane < Lk;
        const float sv = live ? sacc * scale : -FLT_MAX;
        float mx = sv;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        const float m_new = fmaxf(m_run, mx);
        const float pv = live ? expf(sv - m_new) : 0.f;
        const float a = expf(m_run - m_new);
        l_run = l_run * a + warp_sum_f(pv);
        m_run = m_new;
        Ps[w][lane] = pv;
        if (lane == 0) alpha_s[w] = a;
        __syncthreads();
        // ---- acc[i][:] = acc[i][:] * alpha[i] + sum_j P[i][j] * v[j0+j][:]
#pragma unroll
        for (int i = 0; i < AT_BQ; ++i) {
            const float al = alpha_s[i];
#pragma unroll
            for (int c = 0; c < DPT; ++c) acc[i][c] *= al;
        }
        const int jn = Lk - j0 < AT_BK ? Lk - j0 : AT_BK;
        for (int j = 0; j < jn; ++j) {
            float vv[DPT];
#pragma unroll
            for (int c = 0; c < DPT; ++c) {
                const int d = tid + c * AT_THREADS;
                vv[c] = d < D ? vb[(size_t)(j0 + j) * D + d] : 0.f;
            }
#pragma unroll
            for (int i = 0; i < AT_BQ; ++i) {
                const float pj = Ps[i][j];
#pragma unroll
                for (int c = 0; c < DPT; ++c) acc[i][c] = fmaf(pj, vv[c], acc[i][c]);
            }
        }
    }
    if (lane == 0) l_s[w] = l_run;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < AT_BQ; ++i) {
        if (q0 + i >= Lq) break;
        const float inv = 1.f / l_s[i];
#pragma unroll
        for (int c = 0; c < DPT; ++c) {
            const int d = tid + c * AT_THREADS;
            if (d < D) out[((size_t)bh * Lq + q0 + i) * D + d] = acc[i][c] * inv;
        }
    }
}

}  // namespace

cudaError_t attention_f32(const float* q, const float* k, const float* v, float* out, int BH, int Lq, int Lk, int D, float scale,
                          cudaStream_t s) {
    if (D > AT_MAXDPT * AT_THREADS) return cudaErrorInvalidValue;
    dim3 grid((unsigned)((Lq + AT_BQ - 1) / AT_BQ), (unsigned)BH);
    const int dpt = (D + AT_THREADS - 1) / AT_THREADS;
    if (dpt <= 1) attention_kernel<1><<<grid, AT_THREADS, 0, s>>>(q, k, v, out, Lq, Lk, D, scale);
    else if (dpt <= 2) attention_kernel<2><<<grid, AT_THREADS, 0, s>>>(q, k, v, out, Lq, Lk, D, scale);
    else if (dpt <= 4) attention_kernel<4><<<grid, AT_THREADS, 0, s>>>(q, k, v, out, Lq, Lk, D, scale);
    else attention_kernel<8><<<grid, AT_THREADS, 0, s>>>(q, k, v, out, Lq, Lk, D, scale);
    return cudaGetLastError();
}

namespace {
__global__ void __launch_bounds__(256) spatial_mean_kernel(const float* __restrict__ x, float* __restrict__ out, size_t S) {
    const float* p = x + (size_t)blockIdx.x * S;
    __shared__ float red[8];
    float a = 0.f;
    for (size_t i = threadIdx.x; i < S; i += 256) a += p[i];
    a = warp_sum_f(a);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += red[w];
        out[blockIdx.x] = t / (float)S;
    }
}
__global__ void softmax_temperature_kernel(float* x, int rows, int K, float temperature) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    float* p = x + (size_t)r * K;
    float mx = -FLT_MAX;
    for (int k = 0; k < K; ++k) { p[k] = p[k] / temperature; mx = fmaxf(mx, p[k]); }
    float sum = 0.f;
    for (int k = 0; k < K; ++k) { p[k] = expf(p[k] - mx); sum += p[k]; }
    for (int k = 0; k < K; ++k) p[k] /= sum;
}
__global__ void mix_experts_kernel(const float* __restrict__ alpha, const float* __restrict__ w, float* __restrict__ out, int K, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = blockIdx.y;
    float a = 0.f;
    for (int k = 0; k < K; ++k) a = fmaf(alpha[b * K + k], w[(size_t)k * n + i], a);    // torch.mm(softmax_attention, weight)
    out[(size_t)b * n + i] = a;
}
}  // namespace

// ------------------------------------------------------------------------------------------------
// full-reference quality metrics as one fused reduction per image:
// SAM_numpy (core/metrics.py:88-111), ERGAS_numpy (:221-241), UIQC_numpy / QIndex_numpy (:244-284)
// ------------------------------------------------------------------------------------------------
namespace {
constexpr int QM_MAXC = 8, QM_THREADS = 256;
// stage 1: every CTA reduces a pixel range of one image into ws[b] = {per band: sum a, sum b, sum a^2, sum b^2, sum ab} + {sum of angles}
__global__ void __launch_bounds__(QM_THREADS) quality_moments_kernel(const float* __restrict__ ref, const float* __restrict__ img,
                                                                     double* __restrict__ ws, int C, int HW) {
    __shared__ double acc[QM_MAXC][5];     // per band: sum a, sum b, sum a^2, sum b^2, sum ab   (a = ref, b = img)
    __shared__ double sam_acc;
    const int b = blockIdx.y;
    const float* A = ref + (size_t)b * C * HW;
    const float* Bm = img + (size_t)b * C * HW;
    for (int i = threadIdx.x; i < QM_MAXC * 5; i += QM_THREADS) (&acc[0][0])[i] = 0.0;
    if (threadIdx.x == 0) sam_acc = 0.0;
    __syncthreads();
    double loc[QM_MAXC][5];
#pragma unroll
    for (int c = 0; c < QM_MAXC; ++c)
#pragma unroll
        for (int k = 0; k < 5; ++k) loc[c][k] = 0.0;
    double sam = 0.0;
    for (int p = blockIdx.x * QM_THREADS + threadIdx.x; p < HW; p += gridDim.x * QM_THREADS) {
        float dot = 0.f, na = 0.f, nb = 0.f;
#pragma unroll
        for (int c = 0; c < QM_MAXC; ++c) {
            if (c >= C) break;
            const float a = A[(size_t)c * HW + p], v = Bm[(size_t)c * HW + p];
            dot = fmaf(a, v, dot); na = fmaf(a, a, na); nb = fmaf(v, v, nb);
            loc[c][0] += a; loc[c][1] += v; loc[c][2] += (double)a * a; loc[c][3] += (double)v * v; loc[c][4] += (double)a * v;
        }
        const float cs = dot / sqrtf(nb) / sqrtf(na);
        const float ang = acosf(cs);                       // |cs| > 1 or 0/0 -> NaN, counted as 0 like the reference
        if (ang == ang) sam += ang;
    }
    auto wsum = [](double v) {
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    };
    const bool lead = (threadIdx.x & 31) == 0;      // one shared-memory atomic per warp and value (shared double atomics are CAS loops)
#pragma unroll
    for (int c = 0; c < QM_MAXC; ++c) {
        if (c >= C) break;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double u = wsum(loc[c][k]);
            if (lead) atomicAdd(&acc[c][k], u);
        }
    }
    sam = wsum(sam);
    if (lead) atomicAdd(&sam_acc, sam);
    __syncthreads();
    double* w = ws + (size_t)b * (QM_MAXC * 5 + 1);
    for (int i = threadIdx.x; i < C * 5; i += QM_THREADS) atomicAdd(&w[i], (&acc[0][0])[i]);
    if (threadIdx.x == 0) atomicAdd(&w[QM_MAXC * 5], sam_acc);
}

// stage 2: one thread per image turns the moments into {SAM, ERGAS, Q}
__global__ void quality_finish_kernel(const double* __restrict__ ws, float* __restrict__ out, int B, int C, int HW, float ratio) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const double* w = ws + (size_t)b * (QM_MAXC * 5 + 1);
    const double n = (double)HW;
    double erg = 0.0, q = 0.0;
    for (int c = 0; c < C; ++c) {
        const double sa = w[c * 5 + 0], sb = w[c * 5 + 1], saa = w[c * 5 + 2], sbb = w[c * 5 + 3], sab = w[c * 5 + 4];
        const double mse = (saa - 2.0 * sab + sbb) / n;                 // RMSE_numpy^2 of the band
        const double m1 = sa / n, m2 = sb / n;
        erg += mse / (m2 * m2);                                         // / mean(ps_band)^2, ps = the second argument
        const double d1 = (saa - n * m1 * m1) / (n - 1.0), d2 = (sbb - n * m2 * m2) / (n - 1.0), cov = (sab - n * m1 * m2) / (n - 1.0);
        q += 4.0 * cov * m1 * m2 / (d1 + d2) / (m1 * m1 + m2 * m2);
    }
    out[b * 3 + 0] = (float)(w[QM_MAXC * 5] / n * 180.0 / 3.14159265358979323846);
    out[b * 3 + 1] = (float)(100.0 * ratio * sqrt(erg / C));
    out[b * 3 + 2] = (float)(q / C);
}
}  // namespace

cudaError_t quality_metrics(const float* ref, const float* img, float* out, int B, int C, int HW, float ratio, cudaStream_t s) {
    if (C > QM_MAXC) return cudaErrorInvalidValue;
    int split = (2 * 148 + B - 1) / B, most = (HW + 4 * QM_THREADS - 1) / (4 * QM_THREADS);
    if (split > most) split = most;
    if (split < 1) split = 1;
    double* ws = nullptr;
    const size_t wbytes = (size_t)B * (QM_MAXC * 5 + 1) * sizeof(double);
    cudaError_t e = cudaMallocAsync((void**)&ws, wbytes, s);
    if (e != cudaSuccess) return e;
    cudaMemsetAsync(ws, 0, wbytes, s);
    quality_moments_kernel<<<dim3((unsigned)split, (unsigned)B), QM_THREADS, 0, s>>>(ref, img, ws, C, HW);
    quality_finish_kernel<<<(B + 127) / 128, 128, 0, s>>>(ws, out, B, C, HW, ratio);
    e = cudaGetLastError();
    cudaFreeAsync(ws, s);
    return e;
}

// Dynamic_conv attention head (core/dynamic_conv.py:6-43, 106-138, 194-218) for one sample per block: pooled[Ci] -> fc1 -> ReLU ->
// fc2 (+bias) -> softmax(. / temperature) -> alpha[K].  hidden and K are tiny: one warp per hidden unit, one thread per expert.
__global__ void __launch_bounds__(128) dynconv_head_kernel(const float* __restrict__ pooled, const float* __restrict__ fc1, const float* __restrict__ fc2,
                                                            const float* __restrict__ fc2_bias, float* __restrict__ alpha, int Ci, int hidden, int K,
                                                            float temperature) {
    extern __shared__ float sh[];            // hidden activations, then K logits
    const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* p = pooled + (size_t)b * Ci;
    for (int j = warp; j < hidden; j += 4) {
        float a = 0.f;
        for (int c = lane; c < Ci; c += 32) a = fmaf(fc1[(size_t)j * Ci + c], p[c], a);
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) sh[j] = fmaxf(a, 0.f);
    }
    __syncthreads();
    if (threadIdx.x < K) {
        float a = fc2_bias ? fc2_bias[threadIdx.x] : 0.f;
        for (int j = 0; j < hidden; ++j) a = fmaf(fc2[(size_t)threadIdx.x * hidden + j], sh[j], a);
        sh[hidden + threadIdx.x] = a / temperature;
    }
    __syncthreads();
    if (threadIdx.x < K) {
        float mx = -INFINITY, sum = 0.f;
        for (int k = 0; k < K; ++k) mx = fmaxf(mx, sh[hidden + k]);
        for (int k = 0; k < K; ++k) sum += expf(sh[hidden + k] - mx);
        alpha[(size_t)b * K + threadIdx.x] = expf(sh[hidden + threadIdx.x] - mx) / sum;
    }
}

// aggregate_weight / aggregate_bias (:251-257) written straight into the layouts the conv kernels read: per sample
// pf [taps][Ci][Co] fp32 (CUDA-core kernel) or pb [Co][taps*Ci] bf16 (K-major B operand of the tcgen05 kernel).  The experts come
// as [K][Co][Ci][kt] with kt = k^ndim taps that sit in the centre of the k^3 stencil (1-D / 2-D layers), zero elsewhere.
__global__ void __launch_bounds__(256) dynconv_mix_kernel(const float* __restrict__ alpha, const float* __restrict__ w, const float* __restrict__ bias,
                                                           float* __restrict__ pf, __nv_bfloat16* __restrict__ pb, float* __restrict__ ab, int K,
                                                           int Co, int Ci, int k, int ndim) {
    const int taps = k * k * k, kt = ndim == 3 ? taps : (ndim == 2 ? k * k : k);
    const size_t wn = (size_t)Co * Ci * taps, wsrc = (size_t)Co * Ci * kt;
    const int b = blockIdx.y;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float a[8];
    for (int q = 0; q < K && q < 8; ++q) a[q] = alpha[(size_t)b * K + q];
    if (i < wn) {
        const int t = (int)(i % taps);
        const int ci = (int)((i / taps) % Ci);
        const int co = (int)(i / ((size_t)taps * Ci));
        const int dz = t / (k * k), dy = (t / k) % k, dx = t % k, c = k / 2;
        int ts = -1;                                          // tap index in the expert's own stencil
        if (ndim == 3) ts = t;
        else if (ndim == 2) { if (dz == c) ts = dy * k + dx; }
        else if (dz == c && dy == c) ts = dx;
        float v = 0.f;
        if (ts >= 0)
            for (int q = 0; q < K; ++q) v = fmaf(a[q < 8 ? q : 7], w[(size_t)q * wsrc + ((size_t)co * Ci + ci) * kt + ts], v);
        if (pf) pf[(size_t)b * wn + ((size_t)t * Ci + ci) * Co + co] = v;
        if (pb) pb[(size_t)b * wn + (size_t)co * taps * Ci + (size_t)t * Ci + ci] = __float2bfloat16(v);
    }
    if (bias && ab && i < (size_t)Co) {
        float v = 0.f;
        for (int q = 0; q < K; ++q) v = fmaf(a[q < 8 ? q : 7], bias[(size_t)q * Co + i], v);
        ab[(size_t)b * Co + i] = v;
    }
}

cudaError_t dynconv_head(const float* pooled, const float* fc1, const float* fc2, const float* fc2_bias, float* alpha, int B, int Ci, int hidden, int K,
                         float temperature, cudaStream_t s) {
    if (K > 8 || K > 128) return cudaErrorInvalidValue;
    dynconv_head_kernel<<<B, 128, (hidden + K) * sizeof(float), s>>>(pooled, fc1, fc2, fc2_bias, alpha, Ci, hidden, K, temperature);
    return cudaGetLastError();
}
cudaError_t dynconv_mix(const float* alpha, const float* w, const float* bias, float* pf, __nv_bfloat16* pb, float* ab, int B, int K, int Co, int Ci,
                        int k, int ndim, cudaStream_t s) {
    const size_t wn = (size_t)Co * Ci * k * k * k;
    const dim3 grid((unsigned)((wn + 255) / 256), (unsigned)B);
    dynconv_mix_kernel<<<grid, 256, 0, s>>>(alpha, w, bias, pf, pb, ab, K, Co, Ci, k, ndim);
    return cudaGetLastError();
}

cudaError_t spatial_mean(const float* x, float* out, size_t rows, size_t S, cudaStream_t s) {
    spatial_mean_kernel<<<(unsigned)rows, 256, 0, s>>>(x, out, S);
    return cudaGetLastError();
}
cudaError_t softmax_temperature(float* x, int rows, int K, float temperature, cudaStream_t s) {
    softmax_temperature_kernel<<<(rows + 127) / 128, 128, 0, s>>>(x, rows, K, temperature);
    return cudaGetLastError();
}
cudaError_t mix_experts(const float* alpha, const float* w, float* out, int B, int K, size_t n, cudaStream_t s) {
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)B);
    mix_experts_kernel<<<grid, 256, 0, s>>>(alpha, w, out, K, n);
    return cudaGetLastError();
}

cudaError_t groupnorm_affine(const float* x, float* y, float* partial, const float* gamma, const float* beta, int gb_stride, int B,
                             int C, size_t S, int G, int split, float eps, int act, cudaStream_t s) {
    const int cpg = C / G;
    const size_t n = (size_t)cpg * S;
    if ((n & 3) == 0 && (S & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
        // smallest cluster whose slices fit in shared memory, widened until the grid covers the machine
        const size_t max_slice = 40 * 1024;            // floats per CTA (160 KB)
        int cs = 1;
        static const size_t want_slice = getenv("TMD_GN_SLICE") ? (size_t)atoi(getenv("TMD_GN_SLICE")) : 8 * 1024;   // 32 KB slices: several CTAs per SM keep enough loads in flight
        while (cs < 8 && ((n + cs - 1) / cs > want_slice || (size_t)B * G * cs < 2 * 148)) cs *= 2;
        const size_t slice = (((n + cs - 1) / cs) + 3) & ~(size_t)3;
        if (slice <= max_slice && slice >= 4) {
            static bool attr[64] = {};          // the attribute is per device
            int dev = 0;
            if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
            if (!attr[dev]) {
                const cudaError_t e = cudaFuncSetAttribute(gn_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(max_slice * sizeof(float)));
                if (e != cudaSuccess) return e;
                attr[dev] = true;
            }
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)((size_t)B * G * cs));
            cfg.blockDim = dim3(GN_THREADS);
            cfg.dynamicSmemBytes = slice * sizeof(float);
            cfg.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            return cudaLaunchKernelEx(&cfg, gn_cluster_kernel, x, y, gamma, beta, gb_stride, G, cpg, S, slice, eps, act);
        }
    }
    dim3 grid((unsigned)((size_t)B * G), (unsigned)split);
    gn_stats_kernel<<<grid, GN_THREADS, 0, s>>>(x, partial, (size_t)cpg * S, split);
    gn_apply_kernel<<<grid, GN_THREADS, 0, s>>>(x, y, partial, gamma, beta, gb_stride, G, cpg, S, split, eps, act);
    return cudaGetLastError();
}

}  // namespace tmd
