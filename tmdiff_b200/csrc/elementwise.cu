// Memory-bound kernels of the WavBEST hot path: Haar transforms, the 1->C stem, per-step
// activation of cached slices, the C->1 output projection, conditioning tables and layout
// converters.  All are coalesced / 128-bit vectorised; rooflines are HBM bandwidth.
#include "conv_op.h"
#include "kernels.h"

namespace tmd {

static inline int cdiv(size_t a, size_t b) { return (int)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------
// stem: 1x1x1 conv from a single channel + Swish + text scale  (Hyper_unet_general.py:168-171)
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void stem_expand_kernel(const float* __restrict__ x, const float* __restrict__ pan, const float* __restrict__ w,
                                   const float* __restrict__ bias, const float* __restrict__ sc, int ld_sc, T* __restrict__ out,
                                   int C, int N, int HW, size_t total4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total4) return;
    const int c4 = C / 4;
    const size_t vox = i / c4;
    const int c = (int)(i % c4) * 4;
    const int b = (int)(vox / ((size_t)N * HW));
    float v = x[vox];
    if (pan) v = pan[(size_t)b * HW + vox % HW] - v;  // cond = PAN.repeat(1,c,1,1) - MS  (:606-607)
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float u = fmaf(w[c + k], v, bias[c + k]);
        u = swish<false>(u);
        o[k] = u * sc[(size_t)b * ld_sc + c + k];
    }
    store4<T>(out + vox * C + c, o);
}


// ------------------------------------------------------------------------------------------
// bf16 fast paths of the three per-evaluation glue kernels: 8 channels (16 bytes) per thread, 32-bit index
// arithmetic, Swish through tanh.approx (as in the conv epilogue).  HBM-bound.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float swish_tanh_f(float x) {
    const float h = 0.5f * x;
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
}
__device__ __forceinline__ uint4 pack8_bf16(const float v[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ void unpack8_bf16f(const uint4& t, float v[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
__device__ __forceinline__ void ldg8(const float* p, float v[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

__global__ void __launch_bounds__(256) stem_expand_bf16_kernel(const float* __restrict__ x, const float* __restrict__ pan,
                                                               const float* __restrict__ w, const float* __restrict__ bias,
                                                               const float* __restrict__ sc, int ld_sc, __nv_bfloat16* __restrict__ out,
                                                               int C, unsigned NHW, unsigned HW, unsigned total8) {
    const unsigned i = blockIdx.x * 256u + threadIdx.x;
    if (i >= total8) return;
    const unsigned c8 = (unsigned)C >> 3;
    const unsigned vox = i / c8, c = (i % c8) * 8u;
    const unsigned b = vox / NHW;
    float v = x[vox];
    if (pan) v = pan[b * HW + vox % HW] - v;
    float ww[8], bb[8], ss[8], o[8];
    ldg8(w + c, ww); ldg8(bias + c, bb); ldg8(sc + (size_t)b * ld_sc + c, ss);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = swish_tanh_f(fmaf(ww[k], v, bb[k])) * ss[k];
    *reinterpret_cast<uint4*>(out + (size_t)vox * C + c) = pack8_bf16(o);
}

__global__ void __launch_bounds__(256) slice_activate_bf16_kernel(const __nv_bfloat16* __restrict__ src, int ld_src,
                                                                  __nv_bfloat16* __restrict__ dst, int ld_dst, const float* __restrict__ tb,
                                                                  int ld_tb, int C, unsigned vps, unsigned total8) {
    const unsigned i = blockIdx.x * 256u + threadIdx.x;
    if (i >= total8) return;
    const unsigned c8 = (unsigned)C >> 3;
    const unsigned vox = i / c8, c = (i % c8) * 8u;
    const unsigned b = vox / vps;
    float v[8], t[8];
    unpack8_bf16f(*reinterpret_cast<const uint4*>(src + (size_t)vox * ld_src + c), v);
    if (tb) {
        ldg8(tb + (size_t)b * ld_tb + c, t);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] += t[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = swish_tanh_f(v[k]);
    *reinterpret_cast<uint4*>(dst + (size_t)vox * ld_dst + c) = pack8_bf16(v);
}

// The per-step activation of all four cached skip slices in ONE launch (one per level of the U-Net): a thread takes two
// 16-byte vectors half a segment apart, so two loads are in flight per thread and three launch edges per evaluation go away.
struct SliceSeg { const __nv_bfloat16* src; __nv_bfloat16* dst; const float* tb; int ld_src, ld_dst, C; unsigned vps, total8, first; };
struct SliceSegs { SliceSeg s[4]; int n; int ld_tb; };
__global__ void __launch_bounds__(256) slice_activate_multi_kernel(const SliceSegs S) {
    const unsigned g = blockIdx.x * 256u + threadIdx.x;       // index over the first halves of all segments
    int k = 0;
#pragma unroll
    for (int j = 1; j < 4; ++j) k += (j < S.n && g >= S.s[j].first) ? 1 : 0;
    const SliceSeg& sg = S.s[k];
    const unsigned half = (sg.total8 + 1) >> 1, i0 = g - sg.first;
    if (i0 >= half) return;
    const unsigned c8 = (unsigned)sg.C >> 3;
    unsigned idx[2] = {i0, i0 + half};
    uint4 raw[2];
    float t[2][8];
    bool ok[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        ok[u] = idx[u] < sg.total8;
        if (!ok[u]) continue;
        const unsigned vox = idx[u] / c8, c = (idx[u] % c8) * 8u;
        raw[u] = *reinterpret_cast<const uint4*>(sg.src + (size_t)vox * sg.ld_src + c);
        ldg8(sg.tb + (size_t)(vox / sg.vps) * S.ld_tb + c, t[u]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        if (!ok[u]) continue;
        const unsigned vox = idx[u] / c8, c = (idx[u] % c8) * 8u;
        float v[8];
        unpack8_bf16f(raw[u], v);
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = swish_tanh_f(v[q] + t[u][q]);
        *reinterpret_cast<uint4*>(sg.dst + (size_t)vox * sg.ld_dst + c) = pack8_bf16(v);
    }
}

cudaError_t slice_activate_multi(int n, const void* const* src, const int* ld_src, void* const* dst, const int* ld_dst, const float* const* tb,
                                 int ld_tb, const int* C, int B, const size_t* vps, cudaStream_t s) {
    if (n < 1 || n > 4) return cudaErrorInvalidValue;
    SliceSegs S{};
    S.n = n; S.ld_tb = ld_tb;
    unsigned first = 0;
    for (int i = 0; i < n; ++i) {
        const size_t total8 = (size_t)B * vps[i] * (C[i] / 8);
        if (C[i] % 8 || ld_src[i] % 8 || ld_dst[i] % 8 || ld_tb % 4 || total8 >= 0x7fffffffull || !tb[i] ||
            ((reinterpret_cast<uintptr_t>(src[i]) | reinterpret_cast<uintptr_t>(dst[i]) | reinterpret_cast<uintptr_t>(tb[i])) & 15))
            return cudaErrorInvalidValue;
        S.s[i] = SliceSeg{(const __nv_bfloat16*)src[i], (__nv_bfloat16*)dst[i], tb[i], ld_src[i], ld_dst[i], C[i], (unsigned)vps[i], (unsigned)total8, first};
        first += (unsigned)((total8 + 1) / 2);
        first = (first + 255u) & ~255u;          // segments start on a CTA boundary
    }
    slice_activate_multi_kernel<<<first / 256u, 256, 0, s>>>(S);
    return cudaGetLastError();
}

// C <= 64: (C/8) lanes per voxel, one 16-byte load each, shuffle reduce
__global__ void __launch_bounds__(256) project_to_band_bf16_kernel(const __nv_bfloat16* __restrict__ a, int ld, const float* __restrict__ w,
                                                                   float* __restrict__ out, int C, unsigned nvox) {
    const unsigned c8 = (unsigned)C >> 3;                     // 1, 2, 4 or 8 lanes per voxel
    const unsigned g = blockIdx.x * 256u + threadIdx.x;
    const unsigned vox = g / c8, l = g % c8;
    float acc = 0.f;
    if (vox < nvox) {
        float v[8], ww[8];
        unpack8_bf16f(*reinterpret_cast<const uint4*>(a + (size_t)vox * ld + l * 8u), v);
        ldg8(w + l * 8u, ww);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc = fmaf(v[k], ww[k], acc);
    }
    for (unsigned o = 1; o < c8; o <<= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (vox < nvox && l == 0) out[vox] = acc;
}

cudaError_t stem_expand(int dtype, const float* x, const float* pan, const float* w, const float* bias, const float* sc,
                        int ld_sc, void* out, int C, int B, int N, int HW, cudaStream_t s) {
    const size_t total4 = (size_t)B * N * HW * (C / 4);
    if (dtype != 0 && C % 8 == 0 && ld_sc % 4 == 0 && total4 / 2 < 0xffffffffull &&
        ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(sc) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0) {
        const unsigned total8 = (unsigned)(total4 / 2);
        stem_expand_bf16_kernel<<<cdiv(total8, 256), 256, 0, s>>>(x, pan, w, bias, sc, ld_sc, (__nv_bfloat16*)out, C, (unsigned)(N * HW),
                                                                  (unsigned)HW, total8);
        return cudaGetLastError();
    }
    if (dtype == 0)
        stem_expand_kernel<float><<<cdiv(total4, 256), 256, 0, s>>>(x, pan, w, bias, sc, ld_sc, (float*)out, C, N, HW, total4);
    else
        stem_expand_kernel<__nv_bfloat16><<<cdiv(total4, 256), 256, 0, s>>>(x, pan, w, bias, sc, ld_sc, (__nv_bfloat16*)out, C, N, HW, total4);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// dst slice = swish(src slice + tb)   (the "h = x + dense1(embed); act(h)" of a ResBlock whose
// input slice is a cached, step-invariant tensor, Hyper_unet_general.py:238-242)
// ------------------------------------------------------------------------------------------
template <typename T, bool kFast>
__global__ void slice_activate_kernel(const T* __restrict__ src, int ld_src, T* __restrict__ dst, int ld_dst,
                                      const float* __restrict__ tb, int ld_tb, int C, size_t vps, size_t total4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total4) return;
    const int c4 = C / 4;
    const size_t vox = i / c4;
    const int c = (int)(i % c4) * 4;
    const int b = (int)(vox / vps);
    float v[4];
    load4<T>(src + vox * ld_src + c, v);
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = swish<kFast>(v[k] + (tb ? tb[(size_t)b * ld_tb + c + k] : 0.f));
    store4<T>(dst + vox * ld_dst + c, v);
}

cudaError_t slice_activate(int dtype, const void* src, int ld_src, void* dst, int ld_dst, const float* tb, int ld_tb, int C,
                           int B, size_t vps, cudaStream_t s) {
    const size_t total4 = (size_t)B * vps * (C / 4);
    if (dtype != 0 && C % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0 && ld_tb % 4 == 0 && total4 / 2 < 0xffffffffull &&
        ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(tb)) & 15) == 0) {
        const unsigned total8 = (unsigned)(total4 / 2);
        slice_activate_bf16_kernel<<<cdiv(total8, 256), 256, 0, s>>>((const __nv_bfloat16*)src, ld_src, (__nv_bfloat16*)dst, ld_dst, tb, ld_tb,
                                                                     C, (unsigned)vps, total8);
        return cudaGetLastError();
    }
    if (dtype == 0)
        slice_activate_kernel<float, false><<<cdiv(total4, 256), 256, 0, s>>>((const float*)src, ld_src, (float*)dst, ld_dst, tb, ld_tb, C, vps, total4);
    else
        slice_activate_kernel<__nv_bfloat16, true><<<cdiv(total4, 256), 256, 0, s>>>((const __nv_bfloat16*)src, ld_src, (__nv_bfloat16*)dst, ld_dst, tb, ld_tb, C, vps, total4);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// final.conv24: 1x1x1, C -> 1, no bias (modulation already applied by the producer)
// 8 lanes per voxel, each lane 4 channels per step, shuffle reduce.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void project_to_band_kernel(const T* __restrict__ a, int ld, const float* __restrict__ w, float* __restrict__ out, int C,
                                       size_t nvox) {
    const size_t g = ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    const size_t vox = g >> 3;
    const int lane8 = (int)(g & 7);
    float acc = 0.f;
    if (vox < nvox) {
        for (int c = lane8 * 4; c < C; c += 32) {
            float v[4];
            load4<T>(a + vox * ld + c, v);
#pragma unroll
            for (int k = 0; k < 4; ++k) acc = fmaf(v[k], w[c + k], acc);
        }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (vox < nvox && lane8 == 0) out[vox] = acc;
}

cudaError_t project_to_band(int dtype, const void* a, int ld, const float* w, float* out, int C, size_t nvox, cudaStream_t s) {
    const size_t threads = nvox * 8;
    if (dtype != 0 && (C == 8 || C == 16 || C == 32 || C == 64) && ld % 8 == 0 && nvox * (C / 8) < 0xffffffffull &&
        (reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0) {
        project_to_band_bf16_kernel<<<cdiv(nvox * (C / 8), 256), 256, 0, s>>>((const __nv_bfloat16*)a, ld, w, out, C, (unsigned)nvox);
        return cudaGetLastError();
    }
    if (dtype == 0)
        project_to_band_kernel<float><<<cdiv(threads, 256), 256, 0, s>>>((const float*)a, ld, w, out, C, nvox);
    else
        project_to_band_kernel<__nv_bfloat16><<<cdiv(threads, 256), 256, 0, s>>>((const __nv_bfloat16*)a, ld, w, out, C, nvox);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// T = IDWT(0, LH, HL, HH), channels-last: planes x [h][w][C] -> planes x [2h][2w][C]
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void idwt_highbands_kernel(const T* __restrict__ lh, const T* __restrict__ hl, const T* __restrict__ hh, T* __restrict__ out,
                                      int C, int h, int w, size_t total4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total4) return;
    const int c4 = C / 4;
    const size_t lvox = i / c4;
    const int c = (int)(i % c4) * 4;
    const int x = (int)(lvox % w), y = (int)((lvox / w) % h);
    const size_t plane = lvox / ((size_t)w * h);
    float a[4], b[4], d[4];
    load4<T>(lh + lvox * C + c, a);
    load4<T>(hl + lvox * C + c, b);
    load4<T>(hh + lvox * C + c, d);
    float o[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        o[0][k] = (a[k] + b[k] + d[k]) * 0.5f;    // (2i  ,2j  )
        o[1][k] = (-a[k] + b[k] - d[k]) * 0.5f;   // (2i  ,2j+1)
        o[2][k] = (a[k] - b[k] - d[k]) * 0.5f;    // (2i+1,2j  )
        o[3][k] = (-a[k] - b[k] + d[k]) * 0.5f;   // (2i+1,2j+1)
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const size_t ovox = (plane * (2 * h) + 2 * y + (r >> 1)) * (2 * w) + 2 * x + (r & 1);
        store4<T>(out + ovox * C + c, o[r]);
    }
}

cudaError_t idwt_highbands(int dtype, const void* lh, const void* hl, const void* hh, void* out, int C, size_t planes, int h, int w,
                           cudaStream_t s) {
    const size_t total4 = planes * h * w * (C / 4);
    if (dtype == 0)
        idwt_highbands_kernel<float><<<cdiv(total4, 256), 256, 0, s>>>((const float*)lh, (const float*)hl, (const float*)hh, (float*)out, C, h, w, total4);
    else
        idwt_highbands_kernel<__nv_bfloat16><<<cdiv(total4, 256), 256, 0, s>>>((const __nv_bfloat16*)lh, (const __nv_bfloat16*)hl, (const __nv_bfloat16*)hh, (__nv_bfloat16*)out, C, h, w, total4);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Reference-layout Haar analysis / synthesis, [BC][H][W] fp32.
// One thread: 2 input rows x 4 columns (two float4 loads) -> 2 coefficients per band (float2 stores).
// Algorithmic bytes: read 1x + write 1x of the tensor = 2*numel*4 (LL-only: 1.25x).
// ------------------------------------------------------------------------------------------
__global__ void dwt2d_kernel(const float* __restrict__ x, float* __restrict__ ll, float* __restrict__ lh, float* __restrict__ hl,
                             float* __restrict__ hh, int H, int W, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w4 = W / 4, h2 = H / 2;
    const int j = (int)(i % w4);
    const int r = (int)((i / w4) % h2);
    const size_t bc = i / ((size_t)w4 * h2);
    const float* p = x + (bc * H + 2 * r) * W + 4 * j;
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    const float4 u = __ldg(reinterpret_cast<const float4*>(p + W));
    const size_t o = (bc * h2 + r) * (W / 2) + 2 * j;
    // a=t.x b=t.y c=u.x d=u.y | a=t.z b=t.w c=u.z d=u.w
    *reinterpret_cast<float2*>(ll + o) = make_float2(((t.x + t.y) + (u.x + u.y)) * 0.5f, ((t.z + t.w) + (u.z + u.w)) * 0.5f);
    if (lh) {
        *reinterpret_cast<float2*>(lh + o) = make_float2(((t.x - t.y) + (u.x - u.y)) * 0.5f, ((t.z - t.w) + (u.z - u.w)) * 0.5f);
        *reinterpret_cast<float2*>(hl + o) = make_float2(((t.x + t.y) - (u.x + u.y)) * 0.5f, ((t.z + t.w) - (u.z + u.w)) * 0.5f);
        *reinterpret_cast<float2*>(hh + o) = make_float2(((t.x - t.y) - (u.x - u.y)) * 0.5f, ((t.z - t.w) - (u.z - u.w)) * 0.5f);
    }
}

// W % 8 == 0: 2 rows x 8 columns per thread -- four 16-byte loads in flight and 16-byte stores (the 8-byte stores of the kernel
// above write half sectors: 55 % of the copy bandwidth at the U-Net's level-0 shape, profiles/op_sweep_r02.md)
__global__ void __launch_bounds__(256) dwt2d_wide_kernel(const float* __restrict__ x, float* __restrict__ ll, float* __restrict__ lh,
                                                         float* __restrict__ hl, float* __restrict__ hh, int H, int W, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w8 = W / 8, h2 = H / 2;
    const int j = (int)(i % w8);
    const int r = (int)((i / w8) % h2);
    const size_t bc = i / ((size_t)w8 * h2);
    const float* p = x + (bc * H + 2 * r) * W + 8 * j;
    const float4 t0 = __ldg(reinterpret_cast<const float4*>(p)), t1 = __ldg(reinterpret_cast<const float4*>(p + 4));
    const float4 u0 = __ldg(reinterpret_cast<const float4*>(p + W)), u1 = __ldg(reinterpret_cast<const float4*>(p + W + 4));
    const float a[4] = {t0.x, t0.z, t1.x, t1.z}, b[4] = {t0.y, t0.w, t1.y, t1.w};
    const float c[4] = {u0.x, u0.z, u1.x, u1.z}, d[4] = {u0.y, u0.w, u1.y, u1.w};
    const size_t o = (bc * h2 + r) * (W / 2) + 4 * j;
    float4 L, LH, HL, HH;
    float* Lp = &L.x; float* LHp = &LH.x; float* HLp = &HL.x; float* HHp = &HH.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        Lp[k] = ((a[k] + b[k]) + (c[k] + d[k])) * 0.5f;
        LHp[k] = ((a[k] - b[k]) + (c[k] - d[k])) * 0.5f;
        HLp[k] = ((a[k] + b[k]) - (c[k] + d[k])) * 0.5f;
        HHp[k] = ((a[k] - b[k]) - (c[k] - d[k])) * 0.5f;
    }
    *reinterpret_cast<float4*>(ll + o) = L;
    if (lh) {
        *reinterpret_cast<float4*>(lh + o) = LH;
        *reinterpret_cast<float4*>(hl + o) = HL;
        *reinterpret_cast<float4*>(hh + o) = HH;
    }
}

__global__ void dwt2d_scalar_kernel(const float* __restrict__ x, float* __restrict__ ll, float* __restrict__ lh, float* __restrict__ hl,
                                    float* __restrict__ hh, int H, int W, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w2 = W / 2, h2 = H / 2;
    const int j = (int)(i % w2);
    const int r = (int)((i / w2) % h2);
    const size_t bc = i / ((size_t)w2 * h2);
    const float* p = x + (bc * H + 2 * r) * W + 2 * j;
    const float a = p[0], b = p[1], c = p[W], d = p[W + 1];
    ll[i] = ((a + b) + (c + d)) * 0.5f;
    if (lh) {
        lh[i] = ((a - b) + (c - d)) * 0.5f;
        hl[i] = ((a + b) - (c + d)) * 0.5f;
        hh[i] = ((a - b) - (c - d)) * 0.5f;
    }
}

cudaError_t dwt2d_nchw(const float* x, float* ll, float* lh, float* hl, float* hh, int64_t BC, int H, int W, cudaStream_t s) {
    const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (W % 8 == 0 && al16(x) && al16(ll) && (!lh || (al16(lh) && al16(hl) && al16(hh)))) {
        const size_t total = (size_t)BC * (H / 2) * (W / 8);
        dwt2d_wide_kernel<<<cdiv(total, 256), 256, 0, s>>>(x, ll, lh, hl, hh, H, W, total);
    } else if (W % 4 == 0) {
        const size_t total = (size_t)BC * (H / 2) * (W / 4);
        dwt2d_kernel<<<cdiv(total, 256), 256, 0, s>>>(x, ll, lh, hl, hh, H, W, total);
    } else {
        const size_t total = (size_t)BC * (H / 2) * (W / 2);
        dwt2d_scalar_kernel<<<cdiv(total, 256), 256, 0, s>>>(x, ll, lh, hl, hh, H, W, total);
    }
    return cudaGetLastError();
}

__global__ void idwt2d_kernel(const float* __restrict__ ll, const float* __restrict__ lh, const float* __restrict__ hl,
                              const float* __restrict__ hh, float* __restrict__ out, int h, int w, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w2 = w / 2;
    const int j = (int)(i % w2);
    const int r = (int)((i / w2) % h);
    const size_t bc = i / ((size_t)w2 * h);
    const size_t o = (bc * h + r) * w + 2 * j;
    const float2 A = __ldg(reinterpret_cast<const float2*>(ll + o)), B = __ldg(reinterpret_cast<const float2*>(lh + o));
    const float2 C = __ldg(reinterpret_cast<const float2*>(hl + o)), D = __ldg(reinterpret_cast<const float2*>(hh + o));
    float* p = out + (bc * 2 * h + 2 * r) * (2 * w) + 4 * j;
    *reinterpret_cast<float4*>(p) = make_float4(((A.x + B.x) + (C.x + D.x)) * 0.5f, ((A.x - B.x) + (C.x - D.x)) * 0.5f,
                                                ((A.y + B.y) + (C.y + D.y)) * 0.5f, ((A.y - B.y) + (C.y - D.y)) * 0.5f);
    *reinterpret_cast<float4*>(p + 2 * w) = make_float4(((A.x + B.x) - (C.x + D.x)) * 0.5f, ((A.x - B.x) - (C.x - D.x)) * 0.5f,
                                                        ((A.y + B.y) - (C.y + D.y)) * 0.5f, ((A.y - B.y) - (C.y - D.y)) * 0.5f);
}

// w % 4 == 0: four coefficients per band and thread (16-byte loads), two output rows of 8 values (four 16-byte stores)
__global__ void __launch_bounds__(256) idwt2d_wide_kernel(const float* __restrict__ ll, const float* __restrict__ lh, const float* __restrict__ hl,
                                                          const float* __restrict__ hh, float* __restrict__ out, int h, int w, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int w4 = w / 4;
    const int j = (int)(i % w4);
    const int r = (int)((i / w4) % h);
    const size_t bc = i / ((size_t)w4 * h);
    const size_t o = (bc * h + r) * w + 4 * j;
    const float4 A4 = __ldg(reinterpret_cast<const float4*>(ll + o)), B4 = __ldg(reinterpret_cast<const float4*>(lh + o));
    const float4 C4 = __ldg(reinterpret_cast<const float4*>(hl + o)), D4 = __ldg(reinterpret_cast<const float4*>(hh + o));
    const float A[4] = {A4.x, A4.y, A4.z, A4.w}, B[4] = {B4.x, B4.y, B4.z, B4.w}, C[4] = {C4.x, C4.y, C4.z, C4.w}, D[4] = {D4.x, D4.y, D4.z, D4.w};
    float top[8], bot[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        top[2 * k] = ((A[k] + B[k]) + (C[k] + D[k])) * 0.5f;
        top[2 * k + 1] = ((A[k] - B[k]) + (C[k] - D[k])) * 0.5f;
        bot[2 * k] = ((A[k] + B[k]) - (C[k] + D[k])) * 0.5f;
        bot[2 * k + 1] = ((A[k] - B[k]) - (C[k] - D[k])) * 0.5f;
    }
    float* p = out + (bc * 2 * h + 2 * r) * (2 * w) + 8 * j;
    *reinterpret_cast<float4*>(p) = make_float4(top[0], top[1], top[2], top[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(top[4], top[5], top[6], top[7]);
    *reinterpret_cast<float4*>(p + 2 * w) = make_float4(bot[0], bot[1], bot[2], bot[3]);
    *reinterpret_cast<float4*>(p + 2 * w + 4) = make_float4(bot[4], bot[5], bot[6], bot[7]);
}

__global__ void idwt2d_scalar_kernel(const float* __restrict__ ll, const float* __restrict__ lh, const float* __restrict__ hl,
                                     const float* __restrict__ hh, float* __restrict__ out, int h, int w, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int j = (int)(i % w);
    const int r = (int)((i / w) % h);
    const size_t bc = i / ((size_t)w * h);
    const float A = ll[i], B = lh[i], C = hl[i], D = hh[i];
    float* p = out + (bc * 2 * h + 2 * r) * (2 * w) + 2 * j;
    p[0] = ((A + B) + (C + D)) * 0.5f;
    p[1] = ((A - B) + (C - D)) * 0.5f;
    p[2 * w] = ((A + B) - (C + D)) * 0.5f;
    p[2 * w + 1] = ((A - B) - (C - D)) * 0.5f;
}

cudaError_t idwt2d_nchw(const float* ll, const float* lh, const float* hl, const float* hh, float* out, int64_t BC, int h, int w,
                        cudaStream_t s) {
    const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (w % 4 == 0 && al16(ll) && al16(lh) && al16(hl) && al16(hh) && al16(out)) {
        const size_t total = (size_t)BC * h * (w / 4);
        idwt2d_wide_kernel<<<cdiv(total, 256), 256, 0, s>>>(ll, lh, hl, hh, out, h, w, total);
    } else if (w % 2 == 0) {
        const size_t total = (size_t)BC * h * (w / 2);
        idwt2d_kernel<<<cdiv(total, 256), 256, 0, s>>>(ll, lh, hl, hh, out, h, w, total);
    } else {
        const size_t total = (size_t)BC * h * w;
        idwt2d_scalar_kernel<<<cdiv(total, 256), 256, 0, s>>>(ll, lh, hl, hh, out, h, w, total);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// layout converters (stand-alone conv operator only): 32x32 smem transpose
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void ncdhw_to_cl_kernel(const float* __restrict__ x, const float* __restrict__ s, T* __restrict__ out, int C, size_t nhw) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const size_t v0 = (size_t)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i;
        const size_t v = v0 + threadIdx.x;
        float val = 0.f;
        if (c < C && v < nhw) {
            val = x[((size_t)b * C + c) * nhw + v];
            if (s) val *= s[(size_t)b * C + c];
        }
        tile[i][threadIdx.x] = val;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const size_t v = v0 + i;
        const int c = c0 + threadIdx.x;
        if (c < C && v < nhw) out[((size_t)b * nhw + v) * C + c] = from_f32<T>(tile[threadIdx.x][i]);
    }
}

template <typename T>
__global__ void cl_to_ncdhw_kernel(const T* __restrict__ x, float* __restrict__ out, int C, size_t nhw) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const size_t v0 = (size_t)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const size_t v = v0 + i;
        const int c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (c < C && v < nhw) ? to_f32<T>(x[((size_t)b * nhw + v) * C + c]) : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i;
        const size_t v = v0 + threadIdx.x;
        if (c < C && v < nhw) out[((size_t)b * C + c) * nhw + v] = tile[threadIdx.x][i];
    }
}

cudaError_t ncdhw_to_cl(int dtype, const float* x, const float* s, void* out, int B, int C, size_t nhw, cudaStream_t st) {
    dim3 grid(cdiv(nhw, 32), cdiv(C, 32), B), block(32, 8);
    if (dtype == 0) ncdhw_to_cl_kernel<float><<<grid, block, 0, st>>>(x, s, (float*)out, C, nhw);
    else ncdhw_to_cl_kernel<__nv_bfloat16><<<grid, block, 0, st>>>(x, s, (__nv_bfloat16*)out, C, nhw);
    return cudaGetLastError();
}

cudaError_t cl_to_ncdhw(int dtype, const void* x, float* out, int B, int C, size_t nhw, cudaStream_t st) {
    dim3 grid(cdiv(nhw, 32), cdiv(C, 32), B), block(32, 8);
    if (dtype == 0) cl_to_ncdhw_kernel<float><<<grid, block, 0, st>>>((const float*)x, out, C, nhw);
    else cl_to_ncdhw_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)x, out, C, nhw);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// conditioning
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One CTA per sample.  Warp-per-output-row dot products (coalesced weight reads, shuffle reduce).
__global__ void time_table_kernel(const float* __restrict__ t_input, float t_scalar, const float* __restrict__ freqs, int inter_dim,
                                  const float* __restrict__ w1, const float* __restrict__ b1, const float* __restrict__ w2,
                                  const float* __restrict__ b2, const float* __restrict__ Wt, const float* __restrict__ bt,
                                  float* __restrict__ embed_out, float* __restrict__ table, int E, int rows) {
    extern __shared__ float sm[];
    float* g = sm;                 // [inter_dim]
    float* h1 = g + inter_dim;     // [E]
    float* e = h1 + E;             // [E]
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const float t = t_input ? t_input[b] : t_scalar;
    const int half = inter_dim / 2;
    for (int k = threadIdx.x; k < inter_dim; k += blockDim.x) {
        const float arg = t * freqs[k % half];
        g[k] = k < half ? cosf(arg) : sinf(arg);  // cat([cos, sin])  (:93-94)
    }
    __syncthreads();
    for (int j = warp; j < E; j += nwarp) {
        float a = 0.f;
        for (int k = lane; k < inter_dim; k += 32) a = fmaf(w1[j * inter_dim + k], g[k], a);
        a = warp_sum(a);
        if (lane == 0) h1[j] = swish<false>(a + b1[j]);
    }
    __syncthreads();
    for (int j = warp; j < E; j += nwarp) {
        float a = 0.f;
        for (int k = lane; k < E; k += 32) a = fmaf(w2[j * E + k], h1[k], a);
        a = warp_sum(a);
        if (lane == 0) {
            const float v = swish<false>(a + b2[j]);
            e[j] = v;
            if (embed_out) embed_out[(size_t)b * E + j] = v;
        }
    }
    __syncthreads();
    for (int j = warp; j < rows; j += nwarp) {
        float a = 0.f;
        for (int k = lane; k < E; k += 32) a = fmaf(Wt[(size_t)j * E + k], e[k], a);
        a = warp_sum(a);
        if (lane == 0) table[(size_t)b * rows + j] = a + bt[j];
    }
}

cudaError_t time_table(const float* t_input, float t_scalar, const float* freqs, int inter_dim, const float* w1, const float* b1, const float* w2,
                       const float* b2, const float* Wt, const float* bt, float* embed_out, float* table, int B, int E, int rows,
                       cudaStream_t s) {
    const size_t smem = (size_t)(inter_dim + 2 * E) * sizeof(float);
    time_table_kernel<<<B, 512, smem, s>>>(t_input, t_scalar, freqs, inter_dim, w1, b1, w2, b2, Wt, bt, embed_out, table, E, rows);
    return cudaGetLastError();
}

__global__ void dense_rows_kernel(const float* __restrict__ x, const float* __restrict__ W, const float* __restrict__ bias,
                                  float* __restrict__ y, int in_dim, int out_dim, int act) {
    const int r = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (j >= out_dim) return;
    float a = 0.f;
    for (int k = lane; k < in_dim; k += 32) a = fmaf(W[(size_t)j * in_dim + k], x[(size_t)r * in_dim + k], a);
    a = warp_sum(a);
    if (lane == 0) {
        if (bias) a += bias[j];
        y[(size_t)r * out_dim + j] = act == 1 ? swish<false>(a) : (act == 2 ? fmaxf(a, 0.f) : a);
    }
}

cudaError_t dense_rows(const float* x, const float* W, const float* b, float* y, int rows, int in_dim, int out_dim, int act,
                       cudaStream_t s) {
    dim3 grid(cdiv(out_dim, 8), rows);
    dense_rows_kernel<<<grid, 256, 0, s>>>(x, W, b, y, in_dim, out_dim, act);
    return cudaGetLastError();
}

__global__ void gather_rows_kernel(const float* __restrict__ table, const int* __restrict__ ids, float* __restrict__ out, int width) {
    const int b = blockIdx.y;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < width) out[(size_t)b * width + j] = table[(size_t)ids[b] * width + j];
}

__global__ void broadcast_rows_kernel(const float* __restrict__ row, float* __restrict__ out, int width) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < width) out[(size_t)blockIdx.y * width + j] = row[j];
}

cudaError_t broadcast_rows(const float* row, float* out, int B, int width, cudaStream_t s) {
    dim3 grid(cdiv(width, 256), B);
    broadcast_rows_kernel<<<grid, 256, 0, s>>>(row, out, width);
    return cudaGetLastError();
}

cudaError_t gather_rows(const float* table, const int* ids, float* out, int B, int width, cudaStream_t s) {
    dim3 grid(cdiv(width, 256), B);
    gather_rows_kernel<<<grid, 256, 0, s>>>(table, ids, out, width);
    return cudaGetLastError();
}

}  // namespace tmd
