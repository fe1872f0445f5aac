// Shared description of one fused 3-D convolution launch.
//
// Activations live in HBM channels-last: [B][N(bands)][H][W][C] with a per-voxel channel
// stride `ld` >= C, so a producer can write straight into a channel slice of a concat
// buffer (the reference materialises torch.cat, Hyper_unet_general.py:631-634).
//
// One op computes, for every output voxel and output channel,
//     v = sum_{tap,ci} W[tap][ci][co] * src[voxel+tap][ci]  + bias[co] (+ add0) (+ add1)
// and then writes up to three views of v:
//     raw_out            = v
//     act_out[i]         = swish(v + tb[i][b][co]) * sc[i][b][co]        (i = 0,1)
// which is how the reference's  "+ Dense(temb)", Swish and the text modulation of the *next*
// modulated_conv3d (== a per-(sample, input-channel) scale, Hyper_unet_general.py:68-76) are
// folded into the producer.  `mode` selects where v lands:
//     NORMAL   same resolution
//     POOL     2x2 Haar analysis in (H,W): LL/2 (= average) goes through the path above at
//              half resolution; the three un-halved high bands optionally go to hf_out
//              (WaveletUPorDown down branch, Hyper_unet_general.py:389-396)
//              (tcgen05 kernel: pool_out additionally receives the 2x2 average of v at half resolution)
//     UPSAMPLE every v is replicated to the 2x2 block of the double-resolution grid before
//              add0 (IDWT of (2h, HF) == nearest-upsample(h) + IDWT(0, HF), :383-386)
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tmd {

enum ConvMode { CONV_NORMAL = 0, CONV_POOL = 1, CONV_UPSAMPLE = 2 };

struct ConvEpilogue {
    int mode;
    const float* bias;          // [Co] or null ([B][ld_bias] when ld_bias != 0)
    int ld_bias;
    const void* add0; int ld_add0;   // at output resolution
    const void* add1; int ld_add1;   // at half the output resolution (NORMAL only), nearest-upsampled
    void* raw_out; int ld_raw;
    void* act_out[2]; int ld_act[2];
    const float* tb[2]; int ld_tb;   // [B][ld_tb] fp32 (already offset to this module's slice) or null
    const float* sc[2]; int ld_sc;   // [B][ld_sc] fp32 or null
    void* hf_out[3]; int ld_hf;      // POOL only: LH, HL, HH
    void* pool_out; int ld_pool;     // NORMAL only (tcgen05 kernel): 2x2 average of v at half resolution
    int dbg;                         // diagnostics of the tcgen05 kernel (TMD_UMMA_DEBUG), 0 in production
    // tcgen05 kernel: raw_out is written / add0 is read as IEEE half instead of bfloat16 (same bytes, 11-bit instead of 8-bit
    // significand).  For tensors that are only ever added -- the residual stream of the final block, which no MMA reads:
    // rounding it to bf16 at every block was the largest avoidable term of the bf16 mode's error.
    int raw_f16, add0_f16;
    // tcgen05 kernel: view 0 is not stored but projected onto one band plane, proj_out[voxel] += sum_co proj_w[co] * act0[co]
    // (fp32, one red.add per 16-channel slab; proj_out zeroed by the caller).  This is final.conv24 -- a 1x1x1 conv to ONE
    // channel, Hyper_unet_general.py:271-272 -- riding in the epilogue of the conv in front of it.
    float* proj_out;
    const float* proj_w;
};

struct ConvOp {
    int B, N, H, W;             // grid the convolution runs on (input resolution)
    int Ci, ld_in;
    const void* src;
    int Co;
    int taps;                   // 27 (3x3x3, pad 1) or 1
    const float* w_f32;         // [taps][Ci][Co]
    const __nv_bfloat16* w_bf16;  // [Co][taps*Ci], K-major (tcgen05 B operand)
    // Optional fused 1x1x1 convolution of a second tensor at the same resolution, accumulated into the same output
    // (tcgen05 kernel only):  v += sum_ci W2[ci][co] * src2[voxel][ci].  This is how the ResBlock skip `res_conv(x)`
    // (Hyper_unet_general.py:247-249) and the wavelet block's `Conv_2` path (:372, with the 2x2 average commuted in
    // front of the 1x1x1 conv) ride in the K loop of the block's last 3x3x3 conv instead of being separate launches.
    const void* src2; int Ci2, ld_in2;
    int src2_gap_at, src2_gap;     // the Ci2 channels of src2 skip `src2_gap` channels from channel `src2_gap_at` on (a slice of a
                                   // concat buffer left out: its contribution is step-invariant and arrives through add0)
    const __nv_bfloat16* w2_bf16;  // [Co][Ci2]
    // Per-sample weights (Dynamic_conv: a groups=B convolution, core/dynamic_conv.py:244-268): sample b reads w_f32 / w_bf16 at
    // + b * w_sample_stride elements (0: one weight set for the whole batch) and bias at + b * epi.ld_bias.
    long long w_sample_stride;
    ConvEpilogue epi;
};

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// x * sigmoid(x)  (Swish, Hyper_unet_general.py:111-113)
template <bool kFast> __device__ __forceinline__ float swish(float x) {
    if (kFast) return __fdividef(x, 1.f + __expf(-x));
    return x * (1.f / (1.f + expf(-x)));
}

template <typename T> __device__ __forceinline__ void load4(const T* p, float v[4]);
template <> __device__ __forceinline__ void load4<float>(const float* p, float v[4]) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <> __device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float v[4]) {
    uint2 t = *reinterpret_cast<const uint2*>(p);
    __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x), b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
    v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
}
template <typename T> __device__ __forceinline__ void store4(T* p, const float v[4]);
template <> __device__ __forceinline__ void store4<float>(float* p, const float v[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float v[4]) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<uint32_t*>(&a);
    t.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = t;
}

// Writes 4 consecutive output channels [co, co+4) of one output voxel.  `v` already holds
// accumulator + bias; add0/add1 are applied here.
template <typename T, bool kFast>
__device__ __forceinline__ void epilogue_store4(const ConvEpilogue& e, int b, size_t ovox, size_t hvox, int co, float v[4]) {
    if (e.add0) {
        float a[4];
        load4<T>(reinterpret_cast<const T*>(e.add0) + ovox * e.ld_add0 + co, a);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] += a[i];
    }
    if (e.add1) {
        float a[4];
        load4<T>(reinterpret_cast<const T*>(e.add1) + hvox * e.ld_add1 + co, a);
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] += a[i];
    }
    if (e.raw_out) store4<T>(reinterpret_cast<T*>(e.raw_out) + ovox * e.ld_raw + co, v);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (!e.act_out[k]) continue;
        float u[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float x = v[i];
            if (e.tb[k]) x += e.tb[k][(size_t)b * e.ld_tb + co + i];
            x = swish<kFast>(x);
            if (e.sc[k]) x *= e.sc[k][(size_t)b * e.ld_sc + co + i];
            u[i] = x;
        }
        store4<T>(reinterpret_cast<T*>(e.act_out[k]) + ovox * e.ld_act[k] + co, u);
    }
}

}  // namespace tmd
