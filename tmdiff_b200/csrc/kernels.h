// Launchers of the hand-written kernels (all sm_100a).  Every launcher only enqueues on `stream`.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "conv_op.h"

namespace tmd {

// ---- convolution back ends (conv_simt.cu, conv_umma.cu) ----
cudaError_t conv_simt_f32(const ConvOp& op, cudaStream_t stream);
cudaError_t conv_simt_bf16(const ConvOp& op, cudaStream_t stream);
// tcgen05/TMA implicit GEMM; returns cudaErrorNotSupported when the shape is outside its envelope
bool conv_umma_supported(const ConvOp& op);
// true when the schedule chosen for `op` skips the band planes that are pure zero padding (instead of multiplying zeros)
bool conv_umma_skips_padding_planes(const ConvOp& op);
struct UmmaPlan;  // per-op TMA descriptors + launch geometry, built once
UmmaPlan* conv_umma_plan_create(const ConvOp& op);
void conv_umma_plan_destroy(UmmaPlan*);
cudaError_t conv_umma_bf16(const ConvOp& op, const UmmaPlan* plan, cudaStream_t stream);

// ---- element-wise / glue (elementwise.cu) ----
// stem: a[vox][c] = swish(w[c]*v + bias[c]) * sc[b][c], v = x[vox] or pan[b,y,x]-ms[vox]
// (AdaptionModulateBEST.conv20 + Swish + modulation of conv21; cond = PAN.repeat - MS, :605-609)
cudaError_t stem_expand(int dtype, const float* x, const float* pan, const float* w, const float* bias, const float* sc,
                        int ld_sc, void* out, int C, int B, int N, int HW, cudaStream_t s);
// dst[vox][c] = swish(src[vox][c] + tb[b][c])  for a channel slice (per-step activation of cached tensors)
cudaError_t slice_activate(int dtype, const void* src, int ld_src, void* dst, int ld_dst, const float* tb, int ld_tb, int C,
                           int B, size_t vox_per_sample, cudaStream_t s);
// the same for up to four slices in one launch (bf16 only; every slice needs a time bias)
cudaError_t slice_activate_multi(int n, const void* const* src, const int* ld_src, void* const* dst, const int* ld_dst, const float* const* tb,
                                 int ld_tb, const int* C, int B, const size_t* vps, cudaStream_t s);
// out[vox] = sum_c w[c]*a[vox][c]   (final.conv24: 1x1x1, Co = 1, no bias)
cudaError_t project_to_band(int dtype, const void* a, int ld, const float* w, float* out, int C, size_t nvox, cudaStream_t s);
// T[hi vox][c] = IDWT(0, LH, HL, HH)  channels-last (WaveletUPorDown up branch, :383-386)
cudaError_t idwt_highbands(int dtype, const void* lh, const void* hl, const void* hh, void* out, int C, size_t planes, int h, int w,
                           cudaStream_t s);
// reference-layout Haar transforms [BC][H][W] fp32
cudaError_t dwt2d_nchw(const float* x, float* ll, float* lh, float* hl, float* hh, int64_t BC, int H, int W, cudaStream_t s);
cudaError_t idwt2d_nchw(const float* ll, const float* lh, const float* hl, const float* hh, float* out, int64_t BC, int h, int w,
                        cudaStream_t s);
// layout converters for the stand-alone conv operator: [B,C,N,H,W] fp32 <-> [B,N,H,W,C] T (optionally x *= s[b][c])
cudaError_t ncdhw_to_cl(int dtype, const float* x, const float* s, void* out, int B, int C, size_t nhw, cudaStream_t st);
cudaError_t cl_to_ncdhw(int dtype, const void* x, float* out, int B, int C, size_t nhw, cudaStream_t st);

// ---- conditioning (elementwise.cu) ----
// t_b = t_input[b] (device) or t_scalar when t_input is null.
// embed[b] = swish(L2(swish(L1(gamma(t_b)))))  then  table[b][j] = W[j].embed[b] + bias[j]
// (gamma_embedding :80-97, WavBEST.embed :529, every time-Dense :105/:346)
cudaError_t time_table(const float* t_input, float t_scalar, const float* freqs, int inter_dim, const float* w1, const float* b1, const float* w2,
                       const float* b2, const float* Wt, const float* bt, float* embed_out, float* table, int B, int E, int rows,
                       cudaStream_t s);
// y[r][j] = act(W[j].x[r] + b[j])  small dense layer (rows r), act: 0 none, 1 swish, 2 relu; b may be null
cudaError_t dense_rows(const float* x, const float* W, const float* b, float* y, int rows, int in_dim, int out_dim, int act,
                       cudaStream_t s);
cudaError_t gather_rows(const float* table, const int* ids, float* out, int B, int width, cudaStream_t s);
// out[b][:] = row[:]  (one evaluation's time table shared by the whole batch)
cudaError_t broadcast_rows(const float* row, float* out, int B, int width, cudaStream_t s);

// ---- solver (solver.cu) ----
// eps = (x - a*out)/s ; x0 = (x - s*eps)/a ; dynamic threshold (quantile 0.995, max 1) -> dst
// out_c (optional): conditional half for classifier-free guidance, eps = eps_u + g*(eps_c - eps_u); noise_type: out is eps
// ws: threshold_workspace_bytes(B) bytes of device scratch (radix-select histograms; any sample size, several CTAs per sample);
// ws_clean: the workspace was zero-filled once and has only been used by these calls since (each call leaves it clean)
// fu (optional): the dpmsolver++ singlestep update that consumes this x0, evaluated in the same pass as the clamp:
//   fu->dst = (c1*x - c2*m) [+ c3*(m1 - m)],  m / m1 == dst meaning "the fresh x0"
struct FusedUpdate {
    float* dst;
    const float* x;
    const float* m;
    const float* m1;    // null: no third term
    float c1, c2, c3;
};
size_t threshold_workspace_bytes(int B);
// writes the calling thread's tmd_last_error() text and returns 1 (for C-ABI entry points defined outside engine.cu)
int set_last_error(const char* fmt, ...);
cudaError_t x0_threshold(const float* x, const float* out, const float* out_c, float* dst, float alpha, float sigma, bool noise_type,
                         float g, int B, int n, int q_lo, float q_w, float max_val, void* ws, bool ws_clean, const FusedUpdate* fu, cudaStream_t s);
cudaError_t dynamic_threshold_inplace(float* x0, int B, int n, int q_lo, float q_w, float max_val, void* ws, cudaStream_t s);
cudaError_t solver_update(float* dst, const float* x, const float* m, const float* m1, float c1, float c2, float c3, size_t n,
                          cudaStream_t s);
cudaError_t solver_update_multistep(int form, float* dst, const float* x, const float* m, const float* m1, const float* m2, float c1,
                                    float c2, float c3, const float* k, size_t n, cudaStream_t s);

// ancestral DDPM step: out = coef1*clamp(c_recip*x - c_recipm1*eps) + coef2*x + sigma*z; z = noise or in-kernel Philox normals
cudaError_t ancestral_update(float* out, const float* x, const float* eps, const float* noise, float c_recip, float c_recipm1, float coef1,
                             float coef2, float sigma, bool clip, unsigned long long seed, unsigned long long step, size_t n,
                             cudaStream_t s);

// ---- stand-alone operators outside the sampling path (operators.cu) ----
// GroupNorm(G, C, eps) over x [B, C, S] + affine gamma/beta[b * gb_stride + c] (null: none) + optional Swish.
// partial: workspace of B*G*split*3 floats.
cudaError_t groupnorm_affine(const float* x, float* y, float* partial, const float* gamma, const float* beta, int gb_stride, int B,
                             int C, size_t S, int G, int split, float eps, int act, cudaStream_t s);
// out[bh,i,:] = softmax_j(scale * q[bh,i,:].k[bh,j,:]) v[bh,j,:];  q/out [BH,Lq,D], k/v [BH,Lk,D] fp32, D <= 2048
cudaError_t attention_f32(const float* q, const float* k, const float* v, float* out, int BH, int Lq, int Lk, int D, float scale,
                          cudaStream_t s);

// bf16 tensor-core variant (attn_umma.cu): q [BH,Lq,D], k [BH,Lk,D] token-major, vt [BH,D,lk_pitch] channel-major (lk_pitch >= Lk,
// multiple of 8), out [BH,Lq,D] fp32; D a multiple of 64
bool attention_bf16_supported(int BH, int Lq, int Lk, int D, int lk_pitch);
cudaError_t attention_bf16_general(const void* q, const void* k, const void* vt, void* out, int B, int heads, int Lq, int Lk, int D,
                                   long long q_pitch, long long k_pitch, long long v_tok_pitch, int lk_pitch, long long out_pitch, int out_bf16,
                                   float scale, cudaStream_t stream);
cudaError_t attention_bf16_strided(const void* q, const void* k, const void* vt, void* out, int B, int heads, int Lq, int Lk, int D,
                                   long long q_pitch, long long k_pitch, int lk_pitch, long long out_pitch, int out_bf16, float scale,
                                   cudaStream_t stream);
cudaError_t attention_bf16(const void* q, const void* k, const void* vt, float* out, int BH, int Lq, int Lk, int D, int lk_pitch,
                           float scale, cudaStream_t s);

// out[b] = {SAM (degrees), ERGAS, mean per-band Q index} of img[b] against ref[b]; [B, C, HW] fp32, C <= 8
cudaError_t quality_metrics(const float* ref, const float* img, float* out, int B, int C, int HW, float ratio, cudaStream_t s);
// Dynamic_conv3d pieces (core/dynamic_conv.py:194-268): channel means, expert softmax, per-sample expert mix
cudaError_t spatial_mean(const float* x, float* out, size_t rows, size_t S, cudaStream_t s);                       // out[r] = mean(x[r, :])
cudaError_t softmax_temperature(float* x, int rows, int K, float temperature, cudaStream_t s);                      // in place over K
cudaError_t mix_experts(const float* alpha, const float* w, float* out, int B, int K, size_t n, cudaStream_t s);    // out[b,:] = sum_k alpha[b,k] w[k,:]
// the attention head in one launch (fc1 -> ReLU -> fc2 (+bias) -> softmax(. / temperature)), K <= 8
cudaError_t dynconv_head(const float* pooled, const float* fc1, const float* fc2, const float* fc2_bias, float* alpha, int B, int Ci, int hidden, int K,
                         float temperature, cudaStream_t s);
// per-sample expert mix written in the conv kernels' weight layouts (pf fp32 [B][taps][Ci][Co] and / or pb bf16 [B][Co][taps*Ci]) + mixed bias
cudaError_t dynconv_mix(const float* alpha, const float* w, const float* bias, float* pf, __nv_bfloat16* pb, float* ab, int B, int K, int Co, int Ci,
                        int k, int ndim, cudaStream_t s);


// ---- scene.cu: full-scene front / back end (sensor DN <-> sampler tiles) ----
cudaError_t scene_to_tiles(const float* pan, const float* lrms, float* pan_t, float* ms_t, int C, int Hs, int Ws, int ratio, int tile, int first,
                           int count, float img_scale, cudaStream_t s);
cudaError_t tiles_to_scene(const float* tiles, float* scene, int C, int Hs, int Ws, int tile, int first, int count, float img_scale, bool clamp,
                           cudaStream_t s);

// ---- operators2.cu ----
// nn.LayerNorm(C, eps) over the rows of x [rows, C]; gamma / beta may be null
cudaError_t layernorm_rows(const float* x, float* y, const float* gamma, const float* beta, int rows, int C, float eps, cudaStream_t s);
// out[r, j] = p[r, j] * gelu(p[r, D + j]),  p [rows, 2D]  (GEGLU, core/Attention.py:69-76)
cudaError_t geglu_rows(const float* p, float* out, size_t rows, int D, cudaStream_t s);
// per image: [sum a_i | sum a_i^2 | sum b_j | sum b_j^2 | sum a_i b_j | count], optionally after the 3x3 high-pass of sobel_torch
cudaError_t band_moments(const float* a, const float* b, float* out, int B, int Ca, int Cb, int H, int W, bool highpass, cudaStream_t s);
// learned-PSF blur + every scale-th output (core/kernelAdaption.py:117-180); mode 0 softmax PSF, 1 Gaussian from a 2x2 parameter
cudaError_t psf_downsample(const float* in, float* out, const float* param, float* psf_out, int mode, int planes, int H, int W, int k, int scale,
                           cudaStream_t s);

}  // namespace tmd
