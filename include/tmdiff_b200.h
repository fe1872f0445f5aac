/*
 * tmdiff_b200 -- C ABI of the B200-native TMDiff sampling hot path.
 *
 * The reference (codgodtao/TMDiff) is pure Python/PyTorch and has no FFI of its own; the
 * entry points below are what a binding for its sampling path would bind.  Each one names
 * the reference interface it replaces (paths relative to the reference repository).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types.
 *   - every function returns 0 on success, non-zero on failure; tmd_last_error() returns a
 *     message for the last failure on the calling thread.
 *   - "dev" pointers are CUDA device pointers owned by the caller; "host" pointers are
 *     ordinary host memory.  Image tensors use the reference's layout: contiguous fp32
 *     [B, bands, H, W] (PAN: [B, 1, H, W]).
 *   - all device work is enqueued on the cudaStream_t passed as `stream` (a void* here so
 *     the header has no CUDA dependency); no call synchronises the host unless stated.
 *   - a handle is bound to the CUDA device current at tmd_create and is not thread-safe.
 */
#ifndef TMDIFF_B200_H
#define TMDIFF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tmd_engine* tmd_handle;

enum { TMD_F32 = 0, TMD_BF16 = 1 };                       /* arithmetic mode of the conv path   */
enum { TMD_QB = 0, TMD_WV3 = 1, TMD_GF2 = 2, TMD_WV2 = 3, TMD_WV4 = 4, TMD_NUM_SENSORS = 5 };

/* WavBEST(channels, embed_dim, inter_dim) -- GeneralModel/Hyper_unet_general.py:523-564 */
typedef struct {
    int32_t channels[4];   /* e.g. {32,64,128,256} (config/general_finetune.json:63-66)        */
    int32_t embed_dim;     /* 128                                                              */
    int32_t inter_dim;     /* 32 (sinusoidal time-embedding width)                             */
    int32_t max_batch;     /* workspace is sized for this many patches per call                */
    int32_t max_bands;     /* 4 (QB/GF2) or 8 (WV3)                                            */
    int32_t height, width; /* patch size, multiples of 8 (three Haar levels)                   */
    int32_t dtype;         /* TMD_F32: fp32 SIMT convs (parity mode); TMD_BF16: tcgen05 convs  */
} tmd_config;

/* One network evaluation of the sampler = DPM_Solver.model_fn(x, t) (model_type and guidance from the plan),
 * algorithm "dpmsolver++", dynamic thresholding (core/dpm_solver_pytorch.py:296-306, 430-465).
 * The scalars are the fp32 values the reference computes on the host side of each call. */
typedef struct {
    int32_t src;           /* slot holding x                                                   */
    int32_t dst;           /* slot receiving the thresholded x0 prediction                     */
    float   t_input;       /* (t - 1/N)*1000                            (:291-292)             */
    float   alpha_t;       /* marginal_alpha(t)                          (:304)                */
    float   sigma_t;       /* marginal_std(t)                                                  */
} tmd_eval;

/* One solver update, evaluated in fp32 in the reference's order of operations.
 *   form 0 (singlestep family, core/dpm_solver_pytorch.py:589-592, :648-658, :744-760):
 *       dst = (c1*x - c2*m) [+ c3*(m1 - m)]                       (m1 < 0 => no third term)
 *   form 1 (multistep_dpm_solver_second_update, :817-873, dpmsolver++ / 'dpmsolver'):
 *       D1_0 = k[0]*(m - m1);  dst = c1*x - c2*m - c3*D1_0        (k[0] = 1/r0, c3 = 0.5*(alpha_t*phi_1))
 *   form 2 (multistep_dpm_solver_third_update, :875-927):
 *       D1_0 = k[0]*(m - m1); D1_1 = k[1]*(m1 - m2); D1 = D1_0 + k[2]*(D1_0 - D1_1); D2 = k[3]*(D1_0 - D1_1)
 *       dst = c1*x - c2*m + c3*D1 - k[4]*D2                       (c3 = alpha_t*phi_2, k[4] = alpha_t*phi_3)
 * m = newest model value (model_prev_0), m1 the one before, m2 the oldest. */
typedef struct {
    int32_t dst, x, m, m1; /* slots                                                            */
    float   c1, c2, c3;
    int32_t form;          /* 0 / 1 / 2 as above                                               */
    int32_t m2;            /* slot (form 2 only)                                               */
    float   k[5];
} tmd_update;

typedef struct {
    int32_t kind;          /* 0 = eval, 1 = update                                             */
    tmd_eval   eval;
    tmd_update update;
} tmd_step;

enum { TMD_NUM_SLOTS = 8 };  /* slot 0 holds x_T on entry; the result is read from `result_slot` */

enum { TMD_MODEL_X_START = 0, TMD_MODEL_NOISE = 1 };   /* model_wrapper(model_type=...) (:296-306) */

typedef struct {
    const tmd_step* steps; /* host array                                                       */
    int32_t n_steps;
    int32_t result_slot;
    float   q_weight;      /* torch.quantile(|x0|, 0.995): rank = q*(n-1); q_lo = floor(rank), */
    int32_t q_lo;          /*   q_weight = rank - q_lo (fp32, as ATen computes them)  (:436)   */
    float   max_val;       /* thresholding_max_val (1.0)                              (:437)   */
    int32_t model_type;    /* TMD_MODEL_X_START (sample_by_dpmsolver) or TMD_MODEL_NOISE
                              (sample_by_dpmsolver_noise / _guidance, diffusion_general.py:257-335) */
    int32_t cfg;           /* 1 = classifier-free guidance (model_wrapper :336-344): the condition bound by
                              tmd_precompute_condition holds 2B patches [unconditional | conditional], x is
                              evaluated against both halves and eps = eps_u + guidance_scale*(eps_c - eps_u) */
    float   guidance_scale;
} tmd_plan;

const char* tmd_last_error(void);
int tmd_version(void);

/* networks.define_General -> WavBEST(...)                       GeneralModel/networks.py:80-85 */
int tmd_create(const tmd_config* cfg, tmd_handle* out);
int tmd_destroy(tmd_handle h);

/* nn.Module.load_state_dict for denoise_fn.*  (GeneralModel/model.py:122-134).  `key` is the
 * reference state-dict name (e.g. "down1.conv20.conv21.weight"); data is host fp32, contiguous,
 * in the reference shape.  Unknown keys are an error; call tmd_finalize_weights afterwards. */
int tmd_load_weight(tmd_handle h, const char* key, const float* host, const int64_t* shape, int32_t ndim);
/* 16 (= inter_dim/2) sinusoid frequencies of gamma_embedding (Hyper_unet_general.py:89-92), host fp32 */
int tmd_set_time_frequencies(tmd_handle h, const float* host, int32_t n);
/* WavBEST.{qb,wv3,gf2,wv2,wv4}_embeding: cached pooled CLIP vector [768], host fp32 (:566-598) */
int tmd_set_text_embedding(tmd_handle h, int32_t sensor, const float* host768);
/* repack + upload; synchronises the device */
int tmd_finalize_weights(tmd_handle h);

/* The step-invariant half of WavBEST.forward (condition branch conv1/down{1,2,3}_1, :611-618, and
 * the convH_0 high-frequency convs of the up path, :381): run once per batch of patches.
 * pan [B,1,H,W], ms [B,bands,H,W] device fp32; sensor_ids host int32[B] (all patches share `bands`). */
int tmd_precompute_condition(tmd_handle h, const float* pan_dev, const float* ms_dev, const int32_t* sensor_ids,
                             int32_t B, int32_t bands, int32_t H, int32_t W, void* stream);

/* WavBEST.forward(x_t, t_input, PAN, MS, prompt) (:600-636) for the batch bound by the last
 * tmd_precompute_condition.  x_t/out device fp32 [B,bands,H,W]; t_input device fp32[B]
 * (the reference passes a device tensor, core/dpm_solver_pytorch.py:299; no host sync here). */
int tmd_denoise(tmd_handle h, const float* x_t_dev, const float* t_input_dev, float* out_dev, void* stream);

/* DPM_Solver.sample(x_T, ...) (core/dpm_solver_pytorch.py:1081-1289) as a straight-line program of
 * evaluations and updates prepared by the host (tmdiff_b200/dpm_solver.py).  x_T / out: device fp32
 * [B,bands,H,W]; the four extents are checked against the condition bound by tmd_precompute_condition
 * (B is half of the bound patches when plan->cfg).  Patches of any size (the per-sample quantile is a
 * radix select over several CTAs).  trace_dev (optional, may be NULL): receives the raw network output
 * of every evaluation, [n_eval, B, bands, H, W] ([n_eval, 2B, ...] when plan->cfg); trace_x_dev
 * (optional): the x each evaluation was given, [n_eval, B, bands, H, W]. */
int tmd_sample(tmd_handle h, const float* x_T_dev, int32_t B, int32_t bands, int32_t H, int32_t W, const tmd_plan* plan,
               float* out_dev, float* trace_dev, float* trace_x_dev, void* stream);

/* One ancestral DDPM step after the denoiser call (GeneralDiffusion.p_sample, GeneralModel/diffusion_general.py:134-138,
 * 155-174, 196-201): x0 = c_recip*x - c_recipm1*eps, clamped to [-1,1] when clip; out = coef1*x0 + coef2*x + sigma*z.
 * noise (device, [n]) is the caller's z (the reference's torch.randn_like), or NULL: Philox4x32-10 normals are drawn inside the
 * kernel from (seed, step) and never touch HBM.  sigma = 0 on the last step (t = 0).  x, eps, out: device fp32 [n]; out may
 * alias x. */
int tmd_ancestral_update(const float* x, const float* eps, const float* noise, float* out, float c_recip, float c_recipm1,
                         float coef1, float coef2, float sigma, int32_t clip, uint64_t seed, uint64_t step, int64_t n, void* stream);

/* ---- stand-alone operators (parity tests, per-operator sweep) -------------------------------- */
/* DWT_2D("haar").forward (DWT_IDWT/DWT_IDWT_layer.py:321-334): x [B,C,H,W] -> 4 x [B,C,H/2,W/2].
 * lh/hl/hh may be NULL (LL-only variant). */
int tmd_dwt2d(const float* x, float* ll, float* lh, float* hl, float* hh, int64_t BC, int32_t H, int32_t W, void* stream);
/* IDWT_2D("haar").forward (:409-430): 4 x [B,C,h,w] -> [B,C,2h,2w] */
int tmd_idwt2d(const float* ll, const float* lh, const float* hl, const float* hh, float* out, int64_t BC,
               int32_t h, int32_t w, void* stream);
/* modulated_conv3d(x, w, s, padding=k/2) (Hyper_unet_general.py:51-77) with optional bias:
 * x [B,Ci,N,H,W], w [Co,Ci,k,k,k] (k = 1 or 3), s [B,Ci] or NULL, bias [Co] or NULL -> y [B,Co,N,H,W];
 * all device fp32 in the reference (channels-first) layout.  dtype selects the SIMT-fp32 or tcgen05-bf16 kernel. */
int tmd_conv3d_mod(const float* x, const float* w, const float* s, const float* bias, float* y, int32_t B, int32_t Ci,
                   int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t dtype, void* stream);
/* nn.Linear / a 1x1 convolution over TOKEN-MAJOR activations -- the to_q / to_k / to_v / to_out projections of CrossAttention
 * (core/Attention.py:176-183, 213), the q / k / v / proj_out convolutions of SpatialSelfAttention (:131-134) and the NIN layers
 * of AttnBlockpp (Hyper_unet_general.py:488-491) once their input is laid out [tokens, channels]:
 * y[m, :] = W x[m, :] + bias, x [M, Ci], W [Co, Ci] fp32, bias [Co] fp32 or NULL, y [M, Co].  dtype TMD_BF16: tcgen05 kernel,
 * x / y are bf16 when x_bf16 / y_bf16 are set (no layout or cast pass at all), else fp32 and cast on the way; Co of any width
 * (channel groups of 256 per launch).  dtype TMD_F32: CUDA-core kernel on fp32 operands. */
int tmd_linear(const void* x, const float* w, const float* bias, void* y, int64_t M, int32_t Ci, int32_t Co, int32_t x_bf16,
               int32_t y_bf16, int32_t dtype, void* stream);
/* DPM_Solver.dynamic_thresholding_fn (core/dpm_solver_pytorch.py:430-439) in place on x0 [B, n] */
int tmd_dynamic_threshold(float* x0, int32_t B, int32_t n, int32_t q_lo, float q_weight, float max_val, void* stream);
/* The same operation for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments and result
 * as tmd_dynamic_threshold; the radix-select scratch is a file-scope `static __device__` array of the library
 * (TMD_THRESHOLD_STATIC_MAX_BATCH samples x 10248 ints = 2 623 488 bytes, resident on every GPU that loads the library
 * until exit), so the call makes no cudaMalloc / cudaFree / pool allocation.  B > TMD_THRESHOLD_STATIC_MAX_BATCH is
 * rejected (return 1, tmd_last_error() says so); n is not limited.  All calls on a device share that array: calls that
 * overlap on the GPU (two streams, two host threads) race and give wrong thresholds -- serialise them.  The stream must
 * belong to the current device. */
#define TMD_THRESHOLD_STATIC_MAX_BATCH 64
int tmd_dynamic_threshold_static(float* x0, int32_t B, int32_t n, int32_t q_lo, float q_weight, float max_val, void* stream);

/* ---- operators the north star names that the reference's sampling path never executes;
 * device fp32, reference layouts ---------------------------------------------------------------- */
/* torch.nn.GroupNorm(G, C, eps) over x [B, C, S] (S = product of the trailing dims), then y = gamma*x_hat + beta and an
 * optional Swish.  per_sample_affine = 0: gamma/beta [C] (core/Attention.py:108-109 Normalize; AttnBlockpp.GroupNorm_0,
 * Hyper_unet_general.py:487); = 1: gamma/beta [B, C] from the style vector (AdaptiveGroupNorm, :311-331).  gamma/beta
 * may be NULL (affine=False). */
int tmd_groupnorm(const float* x, float* y, const float* gamma, const float* beta, int32_t per_sample_affine, int32_t B,
                  int32_t C, int64_t S, int32_t G, float eps, int32_t swish, void* stream);
/* out[bh,i,:] = softmax_j(scale * q[bh,i,:].k[bh,j,:]) v[bh,j,:]; q/out [BH,Lq,D], k/v [BH,Lk,D], D <= 2048.
 * The core of SpatialSelfAttention (core/Attention.py:140-159), CrossAttention (:185-214) and AttnBlockpp
 * (Hyper_unet_general.py:500-506) as one fused pass (no [Lq,Lk] matrix in HBM). */
int tmd_attention(const float* q, const float* k, const float* v, float* out, int32_t BH, int32_t Lq, int32_t Lk, int32_t D,
                  float scale, void* stream);
/* The same attention core on the tcgen05 tensor cores (bf16 operands, fp32 logits, softmax statistics and accumulation; the
 * reference's own half-precision path keeps the logits in fp32 too, core/Attention.py:194-197).  q [BH,Lq,D] and k [BH,Lk,D]
 * are token-major bf16, vt [BH,D,lk_pitch] is CHANNEL-major bf16 (the [B,C,HW] layout of the reference's 1x1 convs,
 * core/Attention.py:128-131; lk_pitch >= Lk, a multiple of 8, padding never read), out [BH,Lq,D] fp32.  D must be a multiple of
 * 64 and scale positive; anything else is an error (the fp32 entry above takes every shape). */
int tmd_attention_bf16(const void* q, const void* k, const void* vt, float* out, int32_t BH, int32_t Lq, int32_t Lk, int32_t D,
                       int32_t lk_pitch, float scale, void* stream);
/* The same kernel on the projections' own layout, so that the heads need not be split and merged by copies
 * (`rearrange(t, 'b n (h d) -> (b h) n d')` and back, core/Attention.py:189-191, 211): q [B,Lq,q_pitch] and k [B,Lk,k_pitch] hold
 * head h in columns [h*D, (h+1)*D) (bf16; pitches >= heads*D, multiples of 8), vt [B*heads,D,lk_pitch] as above, out
 * [B,Lq,out_pitch] with head h in the same columns, fp32 or bf16 (out_bf16).  heads = 1, pitches = D is tmd_attention_bf16. */
int tmd_attention_bf16_strided(const void* q, const void* k, const void* vt, void* out, int32_t B, int32_t heads, int32_t Lq,
                               int32_t Lk, int32_t D, int64_t q_pitch, int64_t k_pitch, int32_t lk_pitch, int64_t out_pitch,
                               int32_t out_bf16, float scale, void* stream);
/* ... and with v token-major as well, [B,Lk,v_pitch] like k (the value projection's own output: the kernel reads it as the
 * MN-major B operand of P V), so a module lays nothing out between its projections and the core. */
int tmd_attention_bf16_tokens(const void* q, const void* k, const void* v, void* out, int32_t B, int32_t heads, int32_t Lq,
                              int32_t Lk, int32_t D, int64_t q_pitch, int64_t k_pitch, int64_t v_pitch, int64_t out_pitch,
                              int32_t out_bf16, float scale, void* stream);
/* Dynamic_conv3d.forward (core/dynamic_conv.py:244-268; groups=1, stride 1, 'same' padding, k in {1,3}):
 * x [B,Ci,N,H,W], weight [K,Co,Ci,k,k,k], bias [K,Co] or NULL, fc1 [hidden,Ci], fc2 [K,hidden] (attention3d's 1x1x1 convs,
 * :194-218) -> y [B,Co,N,H,W]. */
int tmd_dynamic_conv3d(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2, float* y,
                       int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t K,
                       int32_t hidden, float temperature, int32_t dtype, void* stream);

/* Full-reference quality of fused images, one fused reduction per image (no host gather): out[b] = {SAM in degrees
 * (SAM_numpy, core/metrics.py:88-111), ERGAS (ERGAS_numpy, :221-241, ratio = 0.25), mean per-band universal quality
 * index (UIQC_numpy / QIndex_numpy, :244-284)} of img[b] against ref[b]; ref/img device fp32 [B, C, H, W], C <= 8. */
int tmd_quality_metrics(const float* ref, const float* img, float* out, int32_t B, int32_t C, int32_t H, int32_t W, float ratio,
                        void* stream);

/* Dynamic_conv1d / Dynamic_conv2d / Dynamic_conv3d.forward in one entry (core/dynamic_conv.py:45-104, 140-190, 221-268): ndim = 1, 2
 * or 3 trailing dims carry the k^ndim kernel (x viewed as [B,Ci,N,H,W] with the unused leading extents = 1, weight
 * [K,Co,Ci,k^ndim]); fc2_bias [K] is the bias attention1d / attention2d put on fc2 (:18, :108; attention3d has none -> NULL). */
int tmd_dynamic_conv(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2,
                     const float* fc2_bias, float* y, int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k,
                     int32_t ndim, int32_t K, int32_t hidden, float temperature, int32_t dtype, void* stream);
/* The same with the caller's scratch block (no allocation inside the call): `workspace` is device memory of at least
 * tmd_dynamic_conv_workspace_bytes(...) bytes, 256-byte aligned.  All three entries run one groups=B convolution as ONE launch of
 * the conv kernel with per-sample weight rows: spatial mean, attention head, expert mix, layout in, conv, layout out. */
int64_t tmd_dynamic_conv_workspace_bytes(int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t K, int32_t dtype);
int tmd_dynamic_conv_ws(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2,
                        const float* fc2_bias, float* y, int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k,
                        int32_t ndim, int32_t K, int32_t hidden, float temperature, int32_t dtype, void* workspace, int64_t workspace_bytes,
                        void* stream);
/* torch.nn.LayerNorm(C, eps) over the rows of x [rows, C] (BasicTransformerBlock.norm1-3, core/Attention.py:282-284);
 * gamma / beta [C] or NULL. */
int tmd_layernorm(const float* x, float* y, const float* gamma, const float* beta, int32_t rows, int32_t C, float eps, void* stream);
/* GEGLU gate (core/Attention.py:69-76): out[r, j] = proj[r, j] * gelu(proj[r, D + j]), proj [rows, 2D] (exact erf GELU). */
int tmd_geglu(const float* proj, float* out, int64_t rows, int32_t D, void* stream);
/* Band-pair moments of image pairs, one fused reduction per image: a [B,Ca,H,W], b [B,Cb,H,W] (Ca, Cb <= 8) ->
 * out [B, 2Ca + 2Cb + Ca*Cb + 1] = [sum a_i | sum a_i^2 | sum b_j | sum b_j^2 | sum a_i*b_j (i major) | pixel count].
 * highpass != 0: both images first go through the 3x3 filter of sobel_torch (core/metrics.py:411-422; valid region).
 * Everything QIndex_torch / D_lambda_torch / D_s_torch / SCC_torch (:425-503) and cc (utils/util.py:375-386) need. */
int tmd_band_moments(const float* a, const float* b, float* out, int32_t B, int32_t Ca, int32_t Cb, int32_t H, int32_t W,
                     int32_t highpass, void* stream);
/* EntireKernel / GaussianKernel.forward (core/kernelAdaption.py:117-180): depthwise valid cross-correlation of x [B,C,H,W]
 * with one learned k x k PSF, keeping every scale-th output -> y [B,C,(H-k)/scale+1,(W-k)/scale+1].  mode 0: PSF =
 * softmax(param[k*k]) (EntireKernel, k = scale_factor); mode 1: the Gaussian of kernel_generator (:12-47, shift 'center')
 * from the 2x2 parameter param[4].  psf_out [k*k] (or NULL) receives the PSF (the reference keeps it as self.psf). */
int tmd_psf_downsample(const float* x, float* y, const float* param, float* psf_out, int32_t mode, int32_t B, int32_t C, int32_t H,
                       int32_t W, int32_t k, int32_t scale, void* stream);

/* ---- full-scene front / back end (BASELINE config 4): the steps either side of the sampler, device to device ------------
 * Front end: tiles [first, first + count) (row-major tile order of unpatch_16, data/LRHR_dataset.py:55-91, any grid) of a PAN
 * scene [Hs, Ws] and its low-resolution MS cube [C, Hs/ratio, Ws/ratio], both sensor DN, become the sampler's inputs:
 * pan_tiles [count,1,tile,tile] = DN / img_scale (2047, or 1023 for GF2, :134-139) and ms_tiles [count,C,tile,tile] = the scaled
 * MS up-sampled by `ratio` as cv2.resize(INTER_LINEAR) does (Upsample, :97-122).  tile must be a multiple of 4. */
int tmd_scene_to_tiles(const float* pan_dn, const float* lrms_dn, float* pan_tiles, float* ms_tiles, int32_t C, int32_t Hs, int32_t Ws,
                       int32_t ratio, int32_t tile, int32_t first, int32_t count, float img_scale, void* stream);
/* Back end: fused tiles [count,C,tile,tile] written to their places in the scene cube [C,Hs,Ws] (patch_16 order), clamped to
 * [0,1] when `clamp` and multiplied by img_scale (normlization(min_max=(0,1)) * img_scale,
 * general_sharpening_joint_random_batch_finetune.py:39-42, 136-142). */
int tmd_tiles_to_scene(const float* tiles, float* scene, int32_t C, int32_t Hs, int32_t Ws, int32_t tile, int32_t first, int32_t count,
                       float img_scale, int32_t clamp, void* stream);

/* Debug aid (no reference counterpart): with TMD_ARENA_REDZONE=1 in the environment when the condition is bound, every tensor
 * of the handle's workspace is followed by a 4 KB canary-filled red zone; this call synchronises the device, checks all of them
 * and fails (naming the first tensor whose red zone was overwritten) if a kernel stored out of bounds. */
int tmd_debug_check_redzones(tmd_handle h, int32_t* n_checked, int32_t* n_bad);

/* number of kernel launches issued by this handle since creation (bench.py's gpu_launches) */
int64_t tmd_launch_count(tmd_handle h);

/* ---- measurement (no reference counterpart: the reference only prints time.time() deltas,
 * general_sharpening_joint_random_batch_finetune.py:131,146-147) -------------------------------
 * Between tmd_profile_begin and tmd_profile_end every kernel this handle launches is bracketed by
 * a CUDA event pair on the launching stream.  tmd_profile_end synchronises the device and returns,
 * per kernel kind, the summed device time and the ALGORITHMIC flops / bytes of those launches
 * (2*MACs; each operand and result tensor counted once) -- the numerator of bench.py's roofline. */
enum { TMD_K_CONV3 = 0,        /* 3x3x3 implicit-GEMM convolutions (tensor-bound)                 */
       TMD_K_CONV1 = 1,        /* 1x1x1 convolutions (HBM-bound)                                  */
       TMD_K_ELEMENTWISE = 2,  /* stems, Haar synthesis of cached bands, slice activations, 32->1 */
       TMD_K_SOLVER = 3,       /* x0 conversion + dynamic threshold, dpmsolver++ updates          */
       TMD_K_COND = 4,         /* time-embedding MLP + per-layer bias table                       */
       TMD_NUM_KINDS = 5 };
typedef struct {
    double ms, flops, bytes;
    int64_t launches;
    double flops_executed;   /* FLOPs the kernels issued: the tcgen05 conv skips the zero-padding band planes that `flops` counts */
} tmd_profile_kind;
typedef struct { tmd_profile_kind kind[TMD_NUM_KINDS]; } tmd_profile;
int tmd_profile_begin(tmd_handle h);
/* csv_path (optional, may be NULL): also write one line per launch (label, ms, flops, bytes) */
int tmd_profile_end(tmd_handle h, tmd_profile* out, const char* csv_path);

#ifdef __cplusplus
}
#endif
#endif /* TMDIFF_B200_H */
