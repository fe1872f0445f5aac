// Full-scene front / back end of the sampler (BASELINE config 4): the steps either side of the hot path as two gather / scatter
// kernels, so a scene goes from sensor DN to sampler inputs (and back) without a host round trip or intermediate tensors.
//
//   scene_to_tiles   data/LRHR_dataset.py:134-139  DN / img_scale (2047, or 1023 for GF2)
//                    data/LRHR_dataset.py:97-122   Upsample = cv2.resize(..., INTER_LINEAR): half-pixel centres, replicated border,
//                                                  horizontal pass then vertical pass in fp32
//                    data/LRHR_dataset.py:55-91    unpatch order: tile (i, j) of a gh x gw grid at index i*gw + j
//                    -> PAN tiles [n,1,t,t] and up-sampled MS tiles [n,C,t,t], the tensors tmd_precompute_condition takes
//                       (`PAN.repeat - MS` itself is formed by the stem kernel, elementwise.cu)
//   tiles_to_scene   general_sharpening_joint_random_batch_finetune.py:39-42, 136-142: clamp to [0,1], x img_scale, tiles back
//                    into the [C, Hs, Ws] cube (patch_16 order)
#include "kernels.h"

namespace tmd {

namespace {

// cv2 INTER_LINEAR source position of destination index d when enlarging by `ratio`: s = (d + 0.5)/ratio - 0.5; the left
// tap is clamped at 0 (weight 0) and the right tap at n-1.
__device__ __forceinline__ void lin_tap(int d, float inv_ratio, int n, int& i0, int& i1, float& w1) {
    float s = (d + 0.5f) * inv_ratio - 0.5f;
    int i = (int)floorf(s);
    float f = s - (float)i;
    if (i < 0) { i = 0; f = 0.f; }
    if (i >= n - 1) { i = n - 1; f = 0.f; }      // cv2: the last source column / row is replicated
    i0 = i;
    i1 = min(i + 1, n - 1);
    w1 = f;
}

// one thread per 4 consecutive x of one (tile, channel-or-pan, y) row; channel index C means the PAN plane
__global__ void __launch_bounds__(256) scene_to_tiles_kernel(const float* __restrict__ pan, const float* __restrict__ lrms,
                                                              float* __restrict__ pan_t, float* __restrict__ ms_t, int C, int Hs, int Ws,
                                                              int ratio, int tile, int first, int count, float img_scale) {
    const int t4 = tile / 4;
    const size_t total = (size_t)count * (C + 1) * tile * t4;
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const int x4 = (int)(g % t4);
    const int y = (int)((g / t4) % tile);
    const int c = (int)((g / ((size_t)t4 * tile)) % (C + 1));
    const int n = (int)(g / ((size_t)t4 * tile * (C + 1)));
    const int gw = Ws / tile, ti = (first + n) / gw, tj = (first + n) % gw;
    const int Y = ti * tile + y, X0 = tj * tile + 4 * x4;
    float o[4];
    if (c == C) {
        const float4 v = *reinterpret_cast<const float4*>(pan + (size_t)Y * Ws + X0);
        o[0] = __fdiv_rn(v.x, img_scale); o[1] = __fdiv_rn(v.y, img_scale); o[2] = __fdiv_rn(v.z, img_scale); o[3] = __fdiv_rn(v.w, img_scale);
        *reinterpret_cast<float4*>(pan_t + ((size_t)n * tile + y) * tile + 4 * x4) = make_float4(o[0], o[1], o[2], o[3]);
        return;
    }
    const int hl = Hs / ratio, wl = Ws / ratio;
    const float inv_ratio = 1.f / (float)ratio;
    int y0, y1;
    float wy;
    lin_tap(Y, inv_ratio, hl, y0, y1, wy);
    const float* r0 = lrms + ((size_t)c * hl + y0) * wl;
    const float* r1 = lrms + ((size_t)c * hl + y1) * wl;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x0, x1;
        float wx;
        lin_tap(X0 + k, inv_ratio, wl, x0, x1, wx);
        // DN scaling first (the reference scales the arrays when it loads them), then hresize, then vresize
        const float a0 = __fdiv_rn(__ldg(r0 + x0), img_scale), a1 = __fdiv_rn(__ldg(r0 + x1), img_scale);
        const float b0 = __fdiv_rn(__ldg(r1 + x0), img_scale), b1 = __fdiv_rn(__ldg(r1 + x1), img_scale);
        const float top = __fadd_rn(__fmul_rn(a0, 1.f - wx), __fmul_rn(a1, wx));
        const float bot = __fadd_rn(__fmul_rn(b0, 1.f - wx), __fmul_rn(b1, wx));
        o[k] = __fadd_rn(__fmul_rn(top, 1.f - wy), __fmul_rn(bot, wy));
    }
    *reinterpret_cast<float4*>(ms_t + (((size_t)n * C + c) * tile + y) * tile + 4 * x4) = make_float4(o[0], o[1], o[2], o[3]);
}

__global__ void __launch_bounds__(256) tiles_to_scene_kernel(const float* __restrict__ tiles, float* __restrict__ scene, int C, int Hs, int Ws,
                                                              int tile, int first, int count, float scale, int clamp) {
    const int t4 = tile / 4;
    const size_t total = (size_t)count * C * tile * t4;
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    const int x4 = (int)(g % t4);
    const int y = (int)((g / t4) % tile);
    const int c = (int)((g / ((size_t)t4 * tile)) % C);
    const int n = (int)(g / ((size_t)t4 * tile * C));
    const int gw = Ws / tile, ti = (first + n) / gw, tj = (first + n) % gw;
    float4 v = *reinterpret_cast<const float4*>(tiles + (((size_t)n * C + c) * tile + y) * tile + 4 * x4);
    float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (clamp) o[k] = fminf(fmaxf(o[k], 0.f), 1.f);
        o[k] *= scale;
    }
    *reinterpret_cast<float4*>(scene + ((size_t)c * Hs + ti * tile + y) * Ws + tj * tile + 4 * x4) = make_float4(o[0], o[1], o[2], o[3]);
}

}  // namespace

cudaError_t scene_to_tiles(const float* pan, const float* lrms, float* pan_t, float* ms_t, int C, int Hs, int Ws, int ratio, int tile, int first,
                           int count, float img_scale, cudaStream_t s) {
    const size_t total = (size_t)count * (C + 1) * tile * (tile / 4);
    if (!total) return cudaSuccess;
    scene_to_tiles_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(pan, lrms, pan_t, ms_t, C, Hs, Ws, ratio, tile, first, count, img_scale);
    return cudaGetLastError();
}

cudaError_t tiles_to_scene(const float* tiles, float* scene, int C, int Hs, int Ws, int tile, int first, int count, float img_scale, bool clamp,
                           cudaStream_t s) {
    const size_t total = (size_t)count * C * tile * (tile / 4);
    if (!total) return cudaSuccess;
    tiles_to_scene_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(tiles, scene, C, Hs, Ws, tile, first, count, img_scale, clamp ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace tmd
