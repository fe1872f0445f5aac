// tcgen05 / TMA implicit-GEMM 3-D convolution for sm_100a (bf16 operands, fp32 accumulation in TMEM).
//
// Replaces the nn.Conv3d / F.conv3d(groups=B) call sites of the reference denoiser
// (GeneralModel/Hyper_unet_general.py:74, :161-164, :224-231, :344-361) in bf16 mode.
//
// GEMM view:  M = output voxels (128 per accumulator), N = Co (one MMA covers all output
// channels), K = taps x Ci.  Activations are channels-last bf16 [B][N][H][W][C], so the A operand
// of one tap is a 5-D TMA box of the input shifted by that tap, with the zero padding supplied by
// TMA's out-of-bounds fill.  Weights are [Co][taps*Ci] (K-major) and arrive as 2-D TMA tiles.
//
// Two schedules share the kernel:
//   * REUSE (3x3x3, H >= 16, W >= 8): the CTA owns an output tile of NACC band-planes x 16 x 8
//     voxels (NACC accumulators of M=128).  Per (dx, 32-channel chunk) ONE haloed box
//     [NACC+2][18][8] voxels is loaded; the 9 (dz,dy) taps x NACC planes are then UMMA descriptor
//     offsets into that box (8 voxels = one swizzle atom, so every offset is atom-aligned).  This
//     cuts L2->smem activation traffic ~5x against loading each tap separately, which is what
//     bounds a 3x3x3 conv with 32..128 output channels (L2 feeds ~42 B/cycle/SM).
//   * SIMPLE (1x1x1, small planes): one box of 128 voxels [TB][TN][TH][TW] per (tap, chunk).
//
// Warp roles (512 threads, persistent CTAs, one per SM):
//   warp 0  : TMA producer for activation boxes     (ring of SA stages; one elected lane issues)
//   warp 1  : TMA producer for weight tiles         (ring of SB stages)
//   warp 2  : tcgen05.mma issuer; two accumulator sets in TMEM
//   warp 3  : TMEM allocation / release; REUSE kernels: helper -- decodes the next tile and stages its per-sample epilogue
//             vectors (bias, time biases, modulation scales) in shared memory one tile ahead
//   warps 4-15 : epilogue, three warps per TMEM lane quarter: tcgen05.ld (next slab read ahead) -> bias / residual adds /
//             Haar pooling or 2x upsampling / time-bias + Swish + text-scale views / 32->1 projection -> bf16 or half stores
// The epilogue's shape (which views, residuals, mode) is a template word (Epi<EF>): the kernels of the sampling path carry no
// run-time branches on it; the generic instantiation (EF = -1) serves the operator API and the TMD_UMMA_DEBUG role sweeps.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <set>
#include <utility>

#include "conv_op.h"
#include "kernels.h"

#ifndef TMD_UMMA_PART
#define TMD_UMMA_PART 0
#endif

namespace tmd {

namespace {

#ifndef TMD_EPI_PER_QUARTER
#define TMD_EPI_PER_QUARTER 3
#endif
constexpr int kEpiPerQuarter = TMD_EPI_PER_QUARTER;      // epilogue warps per TMEM lane quarter
constexpr int kThreads = 128 + 128 * kEpiPerQuarter;   // 4 role warps + 16 epilogue warps
constexpr int kMaxSA = 4, kMaxSB = 16, kMaxAcc = 4;
constexpr int kMaxCo = 256;
constexpr int kCstPerThread = (5 * kMaxCo + 128 * kEpiPerQuarter - 1) / (128 * kEpiPerQuarter);

struct UmmaParams {
    CUtensorMap tmA, tmB;
    CUtensorMap tmA2, tmB2;     // fused 1x1x1 second source (ConvOp::src2): nkc2 extra K chunks after the taps
    int nkc2;
    int c2_gap_at, c2_gap;      // second source: channel coordinate of chunk kc = kc*KC (+ c2_gap from c2_gap_at on)
    uint32_t a2_piece_bytes;    // REUSE: one un-haloed band plane [TH][TW][KC] of the second source
    ConvEpilogue epi;
    int B, N, H, W, Ci, Co, taps;   // Co: output channels of ONE tile (= the layer's Co / co_split)
    uint32_t fd[5][2];          // {multiplier, shift} of the exact divisions by co_split, tiles_x, tiles_y, tiles_n and Co (decode_tile, staging)
    int co_split;               // REUSE: channel groups the layer's Co is split into (a tile = voxels x planes x one group)
    int w_rows_per_sample;      // per-sample weights (ConvOp::w_sample_stride): rows of the weight map per sample, else 0
    int mode;                   // 0 = SIMPLE, 1 = REUSE
    int KC, nkc;                // channels per chunk, chunks per tap
    int TW, TH, TN, TB;         // output tile (voxels per accumulator = 128)
    int NACC;                   // accumulators (band planes) per tile
    int tiles_x, tiles_y, tiles_n, tiles_b, num_tiles;
    int SA, SB;
    int acc_stages, acc_shift;  // accumulator sets in TMEM (2 or 4: 4 when 4*NACC*Co fits the 512 columns) and log2 of it
    int b_taps;                 // REUSE: (dy,dx) taps per weight stage (1, 3 or 9): fewer, larger stages amortise the per-stage handshake
    uint32_t a_stage_bytes, a_stage_stride, b_tile_bytes, b_tile_stride;
    uint32_t row_bytes, layout_type, sbo16;
    uint32_t tmem_cols, idesc;
    int a_planes, a_ysplit;     // REUSE: the haloed box is fetched as a_planes x a_ysplit TMA instructions (more requests in flight)
    int pl_off;                 // REUSE: 1 when a tile spans the whole band axis (both halo planes are padding and get no slot in the stage)
    uint32_t a_piece_bytes;
    long long* trace;           // diagnostics (TMD_UMMA_TRACE=1): per-role cycle counters of CTA 0
    int dbg;                    // diagnostics (TMD_UMMA_DEBUG): 1 no MMA, 2 no activation loads, 4 no weight loads, 8 no epilogue stores
};

// ------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Spin on the phase parity; a wait that lasts ~4 s means a broken pipeline: trap instead of hanging the GPU.
// (A suspend-time hint on try_wait, meant to take the probes off the shared-memory data pipe, was measured: -4 % -- the parked
// warps wake up later than the spinning ones.)
#ifndef TMD_PRODUCER_SLEEP
#define TMD_PRODUCER_SLEEP 0
#endif
template <int kSleepNs = 0>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t spins = 0;; ++spins) {
        if (kSleepNs && spins) __nanosleep(kSleepNs);
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
        if ((spins & 0xfff) == 0xfff) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 8000000000LL) __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float v[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// The same load split into issue and wait, so the next slab's TMEM read overlaps the arithmetic on the current one.  The wait
// names the registers as in/out operands: nothing that uses them can be scheduled above it.
__device__ __forceinline__ void tc_ld16_issue(uint32_t taddr, uint32_t r[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tc_ld16_wait(uint32_t r[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]),
                   "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}

// K-major shared-memory matrix descriptors (tcgen05 "smem descriptor") are assembled in the MMA loop:
// start address >> 4 in bits [0,14), LBO = 1 at [16,30), SBO (8-row swizzle atom pitch) >> 4 at [32,46),
// version 1 at [46,48), swizzle mode at [61,64).

// ------------------------------------------------------------------------------------------
// epilogue helpers (8 consecutive output channels = one 16-byte bf16 vector)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void load8_bf16(const __nv_bfloat16* p, float v[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
__device__ __forceinline__ void store8_bf16(__nv_bfloat16* p, const float v[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ void load8_f32(const float* p, float v[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// 256-bit global accesses (sm_100): one lane moves 16 bf16 channels = one full 32-byte sector per instruction.
// The epilogue's accesses are one voxel per lane, i.e. scattered with the channel pitch, so each instruction costs one
// L1 wavefront per distinct line; the L1 data pipe is shared with the tensor core's shared-memory operand reads, which
// makes wavefronts (not bytes) the budget: 32-byte pieces halve them against 16-byte ones.
__device__ __forceinline__ void st_global_v8(void* p, const uint32_t w[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
                 "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}
__device__ __forceinline__ void ld_global_v8(const void* p, uint32_t w[8]) {
    asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(p));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}
// Packed fp32 pairs (sm_100 FADD2 / FMUL2 / FFMA2): same IEEE results as the scalar instructions at half the issue slots,
// which is what the epilogue warps are short of.
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ void add_bf16x16(float v[16], const uint32_t w[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint64_t r = add2(pk2(v[2 * i], v[2 * i + 1]), pk2(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u)));
        upk2(r, v[2 * i], v[2 * i + 1]);
    }
}
__device__ __forceinline__ void add_bf16x16_scalar(float v[16], const uint32_t w[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] += __low2float(h);
        v[2 * i + 1] += __high2float(h);
    }
}
__device__ __forceinline__ void store16_bf16(__nv_bfloat16* p, const float v[16]) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
    st_global_v8(p, w);
}
// IEEE-half twins for add-only tensors (ConvEpilogue::raw_f16 / add0_f16); values beyond the half range saturate
__device__ __forceinline__ void store16_f16(void* p, const float v[16]) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
    st_global_v8(p, w);
}
__device__ __forceinline__ void add_f16x16(float v[16], const uint32_t w[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        v[2 * i] += f.x;
        v[2 * i + 1] += f.y;
    }
}

// Shape of the epilogue as a compile-time word.  The epilogue of a tile is a short, dependent instruction stream run by twelve
// warps; with the shape in the launch parameters every `if (e.add0)` was a constant-bank load, a compare and a branch the warp
// had to resolve before going on (~15 per 16-channel slab, ~1000 of the ~1500 cycles a slab took at Co = 32:
// profiles/dbg_sweep_r02a.txt, umma_timeline_r02*.txt).  EF >= 0: the shape is this template argument (the kernels of the
// sampling path); EF < 0: the generic kernel reads it from the parameters (operator API, diagnostics).
enum : int { EF_POOL = 1, EF_UP = 2, EF_RAW = 4, EF_ACT0 = 8, EF_ACT1 = 16, EF_ADD0 = 32, EF_ADD1 = 64, EF_POOLOUT = 128, EF_HF = 256,
             EF_RAWF16 = 512, EF_ADD0F16 = 1024, EF_PROJ = 2048 };
template <int EF> struct Epi {
    static __device__ __forceinline__ int mode(const ConvEpilogue& e) { return EF < 0 ? e.mode : (EF & 3); }
    static __device__ __forceinline__ bool raw(const ConvEpilogue& e) { return EF < 0 ? e.raw_out != nullptr : (EF & EF_RAW) != 0; }
    static __device__ __forceinline__ bool act(const ConvEpilogue& e, int k) {      // view 0 is also computed when it is projected instead of stored
        return EF < 0 ? (e.act_out[k] != nullptr || (k == 0 && e.proj_out != nullptr)) : (EF & (EF_ACT0 << k)) != 0;
    }
    static __device__ __forceinline__ bool add0(const ConvEpilogue& e) { return EF < 0 ? e.add0 != nullptr : (EF & EF_ADD0) != 0; }
    static __device__ __forceinline__ bool add1(const ConvEpilogue& e) { return EF < 0 ? e.add1 != nullptr : (EF & EF_ADD1) != 0; }
    static __device__ __forceinline__ bool pool_out(const ConvEpilogue& e) { return EF < 0 ? e.pool_out != nullptr : (EF & EF_POOLOUT) != 0; }
    static __device__ __forceinline__ bool hf(const ConvEpilogue& e) { return EF < 0 ? e.hf_out[0] != nullptr : (EF & EF_HF) != 0; }
    static __device__ __forceinline__ bool raw_f16(const ConvEpilogue& e) { return EF < 0 ? e.raw_f16 != 0 : (EF & EF_RAWF16) != 0; }
    static __device__ __forceinline__ bool add0_f16(const ConvEpilogue& e) { return EF < 0 ? e.add0_f16 != 0 : (EF & EF_ADD0F16) != 0; }
    static __device__ __forceinline__ bool proj(const ConvEpilogue& e) { return EF < 0 ? e.proj_out != nullptr : (EF & EF_PROJ) != 0; }
    static __device__ __forceinline__ int dbg(const ConvEpilogue& e, int bit) { return EF < 0 ? (e.dbg & bit) : 0; }   // TMD_UMMA_DEBUG: generic kernel only
};

// Residual operand of one 16-channel slab, fetched ahead of the accumulator (NORMAL mode).
struct Resid { uint32_t a0[8], a1[8]; };      // add0 at the output resolution, add1 at half of it (both bf16 / half x 16)
template <int EF>
__device__ __forceinline__ void load_resid(const ConvEpilogue& e, size_t ovox, size_t hvox, int co, bool valid, Resid& r) {
    typedef __nv_bfloat16 T;
    if (Epi<EF>::add0(e)) {
#pragma unroll
        for (int i = 0; i < 8; ++i) r.a0[i] = 0u;
        if (valid && !Epi<EF>::dbg(e, 32)) ld_global_v8(reinterpret_cast<const T*>(e.add0) + ovox * e.ld_add0 + co, r.a0);
    }
    if (EF >= 0 && Epi<EF>::add1(e)) {        // generic kernel: add1 is read where it is used (registers)
#pragma unroll
        for (int i = 0; i < 8; ++i) r.a1[i] = 0u;
        if (valid) ld_global_v8(reinterpret_cast<const T*>(e.add1) + hvox * e.ld_add1 + co, r.a1);
    }
}

// Swish through the hardware tanh: x*sigmoid(x) = h + h*tanh(h), h = x/2 (one MUFU op; its 2^-11 relative
// error is below the bf16 rounding of the stored activation).
__device__ __forceinline__ float swish_tanh(float x) {
#ifdef TMD_SWISH_EXACT
    return __fdividef(x, 1.f + __expf(-x));      // diagnostic variant: ex2 + rcp (two MUFU ops)
#else
    const float h = 0.5f * x;
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
#endif
}

// v = accumulator + bias for channels [co, co+16) of one output voxel: residual adds, raw view and the activated
// views (same contract as epilogue_store4 in conv_op.h).
template <int EF, bool kPre>
__device__ __forceinline__ void add_residuals16(const ConvEpilogue& e, size_t ovox, size_t hvox, int co, float v[16], const uint32_t* pre0 = nullptr,
                                                const uint32_t* pre1 = nullptr) {
    typedef __nv_bfloat16 T;
    if (Epi<EF>::add0(e)) {
        uint32_t w[8];
        if (!kPre) ld_global_v8(reinterpret_cast<const T*>(e.add0) + ovox * e.ld_add0 + co, w);
        const uint32_t* src = kPre ? pre0 : w;
        if (Epi<EF>::add0_f16(e)) add_f16x16(v, src);
        else add_bf16x16(v, src);
    }
    if (Epi<EF>::add1(e)) {
        uint32_t w[8];
        const bool pre = kPre && EF >= 0;
        if (!pre) ld_global_v8(reinterpret_cast<const T*>(e.add1) + hvox * e.ld_add1 + co, w);
        add_bf16x16(v, pre ? pre1 : w);
    }
}

// kS: the per-sample vectors come from the tile's staged copy in shared memory, cs = [bias | tb0 | tb1 | sc0 | sc1][Co]
// (tb halved; absent tb reads as 0, absent sc as 1), instead of global memory.
template <int EF, bool kS>
__device__ __forceinline__ void store_views16(const ConvEpilogue& e, int b, size_t ovox, int co, const float v[16], const float* cs = nullptr, int Co = 0,
                                              int cb = 0) {
    typedef __nv_bfloat16 T;
    const int cg = co + cb;          // channel in the tensors (co: channel of this tile's group, the index into cs)
    if (Epi<EF>::raw(e) && !Epi<EF>::dbg(e, 64)) {
        if (Epi<EF>::raw_f16(e)) store16_f16(reinterpret_cast<T*>(e.raw_out) + ovox * e.ld_raw + cg, v);
        else store16_bf16(reinterpret_cast<T*>(e.raw_out) + ovox * e.ld_raw + cg, v);
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (!Epi<EF>::act(e, k)) continue;
        uint32_t w[8];
        float dot = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float t[8], s[8], u[8];
            if (kS) {
                // cs holds tb/2: h = (v + tb)/2 is one FFMA, swish = h + h*tanh(h) another, then the modulation scale.
                // (Reading only the vectors a layer has, with the bias folded into the time bias where no un-activated view
                // needs it, saves a third of these broadcast loads but measured -2.5 %: the uniform branches cost more.)
                const ulonglong2* pt = reinterpret_cast<const ulonglong2*>(cs + (1 + k) * Co + co + 8 * h);
                const ulonglong2* ps = reinterpret_cast<const ulonglong2*>(cs + (3 + k) * Co + co + 8 * h);
                const ulonglong2 t0 = pt[0], t1 = pt[1], s0 = ps[0], s1 = ps[1];
                const uint64_t tt[4] = {t0.x, t0.y, t1.x, t1.y}, ss[4] = {s0.x, s0.y, s1.x, s1.y};
                const uint64_t half2 = pk2(0.5f, 0.5f);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint64_t x = fma2(pk2(v[8 * h + 2 * i], v[8 * h + 2 * i + 1]), half2, tt[i]);
                    if (!Epi<EF>::dbg(e, 16)) {
                        float a, b2, ta, tb2;
                        upk2(x, a, b2);
#ifdef TMD_SWISH_EXACT
                        x = pk2(swish_tanh(2.f * a), swish_tanh(2.f * b2));
#else
                        asm("tanh.approx.f32 %0, %1;" : "=f"(ta) : "f"(a));
                        asm("tanh.approx.f32 %0, %1;" : "=f"(tb2) : "f"(b2));
                        x = fma2(x, pk2(ta, tb2), x);
#endif
                    } else x = add2(x, x);
                    x = mul2(x, ss[i]);
                    upk2(x, u[2 * i], u[2 * i + 1]);
                }
            } else {
                if (e.tb[k]) load8_f32(e.tb[k] + (size_t)b * e.ld_tb + cg + 8 * h, t);
                if (e.sc[k]) load8_f32(e.sc[k] + (size_t)b * e.ld_sc + cg + 8 * h, s);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float x = v[8 * h + i];
                    if (e.tb[k]) x += t[i];
                    if (!Epi<EF>::dbg(e, 16)) x = swish_tanh(x);
                    if (e.sc[k]) x *= s[i];
                    u[i] = x;
                }
            }
            if (k == 0 && Epi<EF>::proj(e)) {
                float pw[8];
                load8_f32(e.proj_w + cg + 8 * h, pw);
#pragma unroll
                for (int i = 0; i < 8; ++i) dot = fmaf(pw[i], u[i], dot);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) w[4 * h + i] = pack_bf16x2(u[2 * i], u[2 * i + 1]);
        }
        if (k == 0 && Epi<EF>::proj(e)) {
            atomicAdd(e.proj_out + ovox, dot);       // two slabs of 16 channels per voxel at Co = 32: a + b == b + a, so still deterministic
            continue;
        }
        if (!Epi<EF>::dbg(e, 128)) st_global_v8(reinterpret_cast<T*>(e.act_out[k]) + ovox * e.ld_act[k] + cg, w);
    }
}

// One 16-channel slab of one accumulator row.  All 32 lanes of the warp call this together (the
// POOL mode exchanges rows through shuffles); `valid` guards the stores only.  ovox / hvox are the row's voxel index at
// the convolution's resolution and at half of it.
template <int EF, bool kS>
__device__ __forceinline__ void epilogue16(const UmmaParams& P, float acc[16], int co, int b, int y, int x, size_t ovox, size_t hvox, bool valid,
                                           const Resid& pre, const float* cs, int cb) {
    const ConvEpilogue& e = P.epi;
    const int cg = co + cb;
    float bias[16];
    if (kS) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 t = reinterpret_cast<const float4*>(cs + co)[i];
            bias[4 * i] = t.x; bias[4 * i + 1] = t.y; bias[4 * i + 2] = t.z; bias[4 * i + 3] = t.w;
        }
    } else if (e.bias) {
        load8_f32(e.bias + (size_t)b * e.ld_bias + cg, bias);
        load8_f32(e.bias + (size_t)b * e.ld_bias + cg + 8, bias + 8);
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) bias[i] = 0.f;
    }
    if (Epi<EF>::mode(e) == CONV_NORMAL) {
        const bool pooled = Epi<EF>::pool_out(e);            // warp-uniform
        if (!valid && !pooled) return;
#pragma unroll
        for (int i = 0; i < 16; i += 2) upk2(add2(pk2(acc[i], acc[i + 1]), pk2(bias[i], bias[i + 1])), acc[i], acc[i + 1]);
        if (valid) add_residuals16<EF, true>(e, ovox, hvox, cg, acc, pre.a0, pre.a1);
        if (pooled) {
            // 2x2 average of v for the consumer's fused Conv_2 path; a quad is wholly inside or outside the tensor
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
                float o[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const float L = acc[i + j] + __shfl_xor_sync(0xffffffffu, acc[i + j], 1);
                    o[j] = (L + __shfl_xor_sync(0xffffffffu, L, P.TW)) * 0.25f;
                }
                w[i >> 1] = pack_bf16x2(o[0], o[1]);
            }
            if (valid && !(x & 1) && !(y & 1))
                st_global_v8(reinterpret_cast<__nv_bfloat16*>(e.pool_out) + hvox * e.ld_pool + cg, w);
        }
        if (valid) store_views16<EF, kS>(e, b, ovox, co, acc, cs, P.Co, cb);
    } else if (Epi<EF>::mode(e) == CONV_POOL) {
        // 2x2 Haar analysis across lanes: x-neighbour = lane^1, y-neighbour = lane^TW
        const bool writer = valid && !(x & 1) && !(y & 1);
        const bool want_hf = Epi<EF>::hf(e);
        float ll[16];
        uint32_t lh[8], hl[8], hh[8];
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            float o[2][3];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float a = acc[i + j];
                const float t = __shfl_xor_sync(0xffffffffu, a, 1);
                const float L = a + t, D = a - t;                           // a+b, a-b   (on the even-x lane)
                const float u = __shfl_xor_sync(0xffffffffu, L, P.TW), w = __shfl_xor_sync(0xffffffffu, D, P.TW);
                ll[i + j] = (L + u) * 0.25f + bias[i + j];                   // LL/2 == 2x2 average (:396)
                o[j][0] = (D + w) * 0.5f;                                    // LH: (a-b)+(c-d)
                o[j][1] = (L - u) * 0.5f;                                    // HL: (a+b)-(c+d)
                o[j][2] = (D - w) * 0.5f;                                    // HH: (a-b)-(c-d)
            }
            if (want_hf) {
                lh[i >> 1] = pack_bf16x2(o[0][0], o[1][0]);
                hl[i >> 1] = pack_bf16x2(o[0][1], o[1][1]);
                hh[i >> 1] = pack_bf16x2(o[0][2], o[1][2]);
            }
        }
        if (!writer) return;
        if (want_hf) {
            typedef __nv_bfloat16 T;
            st_global_v8(reinterpret_cast<T*>(e.hf_out[0]) + hvox * e.ld_hf + cg, lh);
            st_global_v8(reinterpret_cast<T*>(e.hf_out[1]) + hvox * e.ld_hf + cg, hl);
            st_global_v8(reinterpret_cast<T*>(e.hf_out[2]) + hvox * e.ld_hf + cg, hh);
        }
        add_residuals16<EF, false>(e, hvox, 0, cg, ll);
        store_views16<EF, kS>(e, b, hvox, co, ll, cs, P.Co, cb);
    } else {  // CONV_UPSAMPLE: replicate to the 2x2 block of the double-resolution grid, then add0 / views there
        if (!valid) return;
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] += bias[i];
        const size_t up0 = 4 * ovox - 2 * (size_t)x;         // voxel (2y, 2x) of the double-resolution plane
        // the four residual rows are requested before any arithmetic (the loads are volatile asm: left inside the loop below,
        // each would wait for the previous row's stores to issue and expose its whole latency)
        uint32_t res[EF >= 0 ? 4 : 1][8];
        if (EF >= 0 && Epi<EF>::add0(e)) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
                ld_global_v8(reinterpret_cast<const __nv_bfloat16*>(e.add0) + (up0 + (size_t)(r >> 1) * (2 * P.W) + (r & 1)) * e.ld_add0 + cg, res[EF >= 0 ? r : 0]);
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const size_t uvox = up0 + (size_t)(r >> 1) * (2 * P.W) + (r & 1);
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = acc[i];
            if (EF >= 0) add_residuals16<EF, true>(e, uvox, 0, cg, v, res[EF >= 0 ? r : 0], nullptr);
            else add_residuals16<EF, false>(e, uvox, 0, cg, v);
            store_views16<EF, kS>(e, b, uvox, co, v, cs, P.Co, cb);
        }
    }
}

// (back-off in the producers' waits -- nanosleep 32 / 200 ns, try_wait suspend hints 0.5-2 us -- was measured: 0 to -6 %)
#define PRODUCER_WAIT(bar, ph) mbar_wait<TMD_PRODUCER_SLEEP>(bar, ph)
struct TileCoord { int b0, n0, y0, x0, ch; };
// REUSE kernels: what the epilogue warps need to know about a tile, prepared one tile ahead by the helper warp
struct TileInfo {
    int b0, n0, y0, x0, cb;          // sample, first band plane, tile origin, first channel of the tile's channel group
    int pad;
    long long vox_base, hvox_base;   // voxel index of the tile origin at the conv's resolution and at half of it
};
// n / d for n < 2^31 with the host's {multiplier, shift} (round-up method): a 32-bit multiply-high, an add and a shift instead of
// the ~40-instruction runtime division -- the epilogue warps decode two tiles per tile and are short of issue slots.
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const uint32_t f[2]) { return (__umulhi(n, f[0]) + n) >> f[1]; }
__device__ __forceinline__ TileCoord decode_tile(const UmmaParams& P, int tile) {
    TileCoord c;
    // the channel groups of one voxel tile are neighbours in the tile order: they run at the same time on different SMs
    // and share the activation boxes through L2
    uint32_t t = (uint32_t)tile, q;
    q = fdiv(t, P.fd[0]); c.ch = (int)(t - q * P.co_split); t = q;
    q = fdiv(t, P.fd[1]); c.x0 = (int)(t - q * P.tiles_x) * P.TW; t = q;
    q = fdiv(t, P.fd[2]); c.y0 = (int)(t - q * P.tiles_y) * P.TH; t = q;
    q = fdiv(t, P.fd[3]); c.n0 = (int)(t - q * P.tiles_n) * (P.mode ? P.NACC : P.TN); t = q;
    c.b0 = (int)t * P.TB;
    return c;
}

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
// Warp-uniform control flow throughout: every lane of a role's warp runs the loops and one elected
// lane issues the TMA / tcgen05 instructions, so their operands stay in uniform registers (a
// divergent `lane == 0` branch makes ptxas wrap every UTCHMMA in an election loop).
// Issue order of the input planes q = -1..NACC of a haloed box for the merged-dz MMAs, and how many
// leading entries of that order overwrite (their column spans tile the NACC accumulators exactly).
template <int NACC> __device__ __forceinline__ constexpr int merge_order(int i) {
    if (NACC == 4) { constexpr int o[6] = {1, 4, -1, 0, 2, 3}; return o[i]; }      // q=1 -> planes 2,1,0 ; q=4 -> plane 3
    if (NACC == 2) { constexpr int o[4] = {0, -1, 1, 2}; return o[i]; }            // q=0 -> planes 1,0
    if (NACC == 8) { constexpr int o[10] = {1, 4, 7, -1, 0, 2, 3, 5, 6, 8}; return o[i]; }   // 2,1,0 ; 5,4,3 ; 7,6 (q=7: dz 1,2)
    return i - 1;                                                                  // NACC == 1: q=-1 -> plane 0
}
template <int NACC> __device__ __forceinline__ constexpr int merge_fresh() { return NACC == 4 ? 2 : (NACC == 8 ? 3 : 1); }

__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred;
}

template <int MODE, int NACC, int KMMA, bool MERGE, int EF>
__global__ void __launch_bounds__(kThreads, 1) conv_umma_kernel(const __grid_constant__ UmmaParams P) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kMaxSA + 2 * kMaxSB + 2 * kMaxAcc];
    __shared__ uint32_t tmem_base_slot;
    __shared__ __align__(16) float cst[MODE ? 2 : 1][MODE ? 5 * kMaxCo : 4];   // REUSE: per-tile epilogue vectors (see the helper warp)
    __shared__ __align__(16) TileInfo tinfo[2];

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // swizzle atoms need 1024-byte alignment
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + P.SA * P.a_stage_stride;
    const uint32_t bar0 = smem_u32(bars);
    auto a_full = [&](int s) { return bar0 + 8u * s; };
    auto a_empty = [&](int s) { return bar0 + 8u * (kMaxSA + s); };
    auto b_full = [&](int s) { return bar0 + 8u * (2 * kMaxSA + s); };
    auto b_empty = [&](int s) { return bar0 + 8u * (2 * kMaxSA + kMaxSB + s); };
    auto acc_full = [&](int s) { return bar0 + 8u * (2 * kMaxSA + 2 * kMaxSB + s); };
    auto acc_empty = [&](int s) { return bar0 + 8u * (2 * kMaxSA + 2 * kMaxSB + kMaxAcc + s); };

    if (threadIdx.x == 0) {
        for (int s = 0; s < P.SA; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < P.SB; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
        for (int s = 0; s < P.acc_stages; ++s) { mbar_init(acc_full(s), 1); mbar_init(acc_empty(s), 4 * kEpiPerQuarter); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 3) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "r"(P.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    // Programmatic dependent launch: everything above (barriers, TMEM) touched no global memory, so under PDL it overlaps the
    // tail of the previous kernel of the stream; from here on every warp reads or writes tensors of earlier kernels.
    // launch_dependents right away: the next kernel's CTAs may be placed as soon as this grid's CTAs leave their SMs.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    constexpr int KC = KMMA * 16;
    const int ngroups = MODE ? 1 : P.taps;              // activation boxes per K chunk (REUSE: ONE haloed box serves all 27 taps)
    const int b_taps = MODE ? P.b_taps : 1;             // (dy,dx) taps held by one weight stage
    const int b_stages_per_a = MODE ? 9 / b_taps : 1;   // weight stages consumed per activation stage (REUSE: the 9 (dy,dx) taps)
    const int nkc = P.nkc;

    if (warp == 0) {
        // ================= activation boxes =================
        uint32_t s = 0, ph = 1;                            // ring slot, parity of its "empty" phase
        for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x) {
            const TileCoord tc = decode_tile(P, tile);
            // band planes outside [0, N) are zero padding: neither loaded nor multiplied (see the MMA warp)
            int planes_in = 0;
            if (MODE) {
#pragma unroll
                for (int pl = 0; pl < NACC + 2; ++pl) planes_in += (unsigned)(tc.n0 - 1 + pl) < (unsigned)P.N;
            }
            for (int g = 0; g < ngroups; ++g) {
                int cx = tc.x0, cy = tc.y0, cn = tc.n0;
                if (MODE) { cx -= 1; cy -= 1; cn -= 1; }
                else if (P.taps == 27) {
                    cx += g % 3 - 1; cy += (g / 3) % 3 - 1; cn += g / 9 - 1;
                    if (cn + P.TN <= 0 || cn >= P.N) continue;            // the whole box is padding
                }
                for (int kc = 0; kc < nkc; ++kc) {
                    PRODUCER_WAIT(a_empty(s), ph);
                    if (elect_one()) {
                        if (P.dbg & 2) mbar_arrive(a_full(s));
                        else {
                            mbar_expect_tx(a_full(s), MODE ? P.a_piece_bytes * (uint32_t)planes_in : P.a_stage_bytes);
                            uint32_t dst = a_base + s * P.a_stage_stride;
                            if (MODE) {
#pragma unroll
                                for (int pl = 0; pl < NACC + 2; ++pl)
                                    if ((unsigned)(cn + pl) < (unsigned)P.N)
                                        tma_load_5d(dst + (uint32_t)(pl - P.pl_off) * P.a_piece_bytes, &P.tmA, a_full(s), kc * KC, cx, cy, cn + pl, tc.b0);
                            } else {
                                tma_load_5d(dst, &P.tmA, a_full(s), kc * KC, cx, cy, cn, tc.b0);
                            }
                        }
                    }
                    __syncwarp();
                    if (++s == (uint32_t)P.SA) { s = 0; ph ^= 1; }
                }
            }
            if (P.nkc2) {   // second source: un-haloed planes of the tile, centre tap only
                int planes2 = 0;
#pragma unroll
                for (int pl = 0; pl < NACC; ++pl) planes2 += (tc.n0 + pl) < P.N;
                for (int kc = 0; kc < P.nkc2; ++kc) {
                    PRODUCER_WAIT(a_empty(s), ph);
                    const int c2 = kc * KC + (kc * KC >= P.c2_gap_at ? P.c2_gap : 0);
                    if (elect_one()) {
                        uint32_t dst = a_base + s * P.a_stage_stride;
                        if (MODE) {
                            mbar_expect_tx(a_full(s), P.a2_piece_bytes * (uint32_t)planes2);
#pragma unroll
                            for (int pl = 0; pl < NACC; ++pl, dst += P.a2_piece_bytes)
                                if (tc.n0 + pl < P.N) tma_load_5d(dst, &P.tmA2, a_full(s), c2, tc.x0, tc.y0, tc.n0 + pl, tc.b0);
                        } else {
                            mbar_expect_tx(a_full(s), P.a_stage_bytes);
                            tma_load_5d(dst, &P.tmA2, a_full(s), c2, tc.x0, tc.y0, tc.n0, tc.b0);
                        }
                    }
                    __syncwarp();
                    if (++s == (uint32_t)P.SA) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= weight tiles =================
        uint32_t s = 0, ph = 1;
        for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x) {
            const TileCoord tc = decode_tile(P, tile);
            for (int g = 0; g < ngroups; ++g) {
                if (!MODE && P.taps == 27) {
                    const int cn = tc.n0 + g / 9 - 1;
                    if (cn + P.TN <= 0 || cn >= P.N) continue;
                }
                for (int kc = 0; kc < nkc; ++kc)
#pragma unroll 1
                    for (int ts = 0; ts < b_stages_per_a; ++ts) {
                        PRODUCER_WAIT(b_empty(s), ph);
                        if (elect_one()) {
                            if (P.dbg & 4) mbar_arrive(b_full(s));
                            else {
                                uint32_t dst = b_base + s * P.b_tile_stride;
                                if (MODE) {
                                    mbar_expect_tx(b_full(s), 3 * P.b_tile_bytes * (uint32_t)b_taps);
                                    for (int t = ts * b_taps; t < (ts + 1) * b_taps; ++t)    // t = dy*3+dx
#pragma unroll
                                        for (int dz = 0; dz < 3; ++dz, dst += P.b_tile_bytes)   // tap = dz*9+t; rows [dz*Co, (dz+1)*Co) of the tap's block
                                            tma_load_2d(dst, &P.tmB, b_full(s), (dz * 9 + t) * P.Ci + kc * KC, tc.ch * P.Co + tc.b0 * P.w_rows_per_sample);
                                } else {
                                    mbar_expect_tx(b_full(s), P.b_tile_bytes);
                                    tma_load_2d(dst, &P.tmB, b_full(s), g * P.Ci + kc * KC, tc.ch * P.Co + tc.b0 * P.w_rows_per_sample);
                                }
                            }
                        }
                        __syncwarp();
                        if (++s == (uint32_t)P.SB) { s = 0; ph ^= 1; }
                    }
            }
            for (int kc = 0; kc < P.nkc2; ++kc) {
                PRODUCER_WAIT(b_empty(s), ph);
                if (elect_one()) {
                    mbar_expect_tx(b_full(s), P.b_tile_bytes);
                    tma_load_2d(b_base + s * P.b_tile_stride, &P.tmB2, b_full(s), kc * KC, tc.ch * P.Co);
                }
                __syncwarp();
                if (++s == (uint32_t)P.SB) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 2) {
        // ================= MMA issue =================
        // Descriptors differ only in their 14-bit start-address field, so the loops below just add
        // 16-byte-unit offsets to the low word.
        const uint32_t desc_hi = P.sbo16 | (1u << 14) | (P.layout_type << 29);                  // weights, SIMPLE boxes, second source
        // REUSE: the haloed box is [planes][TH+2 lines][TW+2 voxels]; an 8-voxel line segment is one 8-row group, the
        // groups of a view are one line pitch apart (SBO), and a tap is a start-address offset of dy lines + dx voxels.
        // The swizzle is a function of the absolute shared-memory address for TMA and UMMA alike, so views that start
        // inside a swizzle atom read back exactly what TMA wrote (profiles/micro/shifted_view.cu).
        const uint32_t line16 = ((uint32_t)(P.TW + 2) * P.row_bytes) >> 4;
        const uint32_t desc_hi_a = MODE ? (line16 | (1u << 14) | (P.layout_type << 29)) : desc_hi;
        const uint32_t a_lo0 = ((a_base & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t b_lo0 = ((b_base & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t a_stride16 = P.a_stage_stride >> 4, b_stride16 = P.b_tile_stride >> 4;
        const uint32_t plane16 = (uint32_t)(P.TH + 2) * line16;                      // one haloed band plane
        const uint32_t vox16 = P.row_bytes >> 4;                                     // one voxel
        const int pl_off = P.pl_off;
        const uint32_t idesc = P.idesc;
        const uint32_t idesc_base = P.idesc & ~(0x3Fu << 17);
        const uint32_t Co = P.Co;
        const uint32_t co16 = (Co * P.row_bytes) >> 4;                                 // one tap's weight rows
        const bool do_mma = !(P.dbg & 1);
        uint32_t sa = 0, pha = 0, sb = 0, phb = 0, lt = 0;
        long long tw_acc = 0, tw_a = 0, tw_b = 0, t_issue = 0, t_all0 = clock64(), tq = 0;
        const bool tr = P.trace && blockIdx.x == 0;
        for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, ++lt) {
            const uint32_t ab = lt & (uint32_t)(P.acc_stages - 1);             // accumulator set (2 or 4 of them)
            if (tr) tq = clock64();
            mbar_wait(acc_empty(ab), ((lt >> P.acc_shift) & 1) ^ 1);
            if (tr) tw_acc += clock64() - tq;
            if (tr && lane == 0 && lt >= 4 && lt < 8) P.trace[200 + 2 * (lt - 4)] = clock64();          // timeline: MMA issue of this tile starts
            tc_fence_after();
            const uint32_t d0 = tmem_base + ab * NACC * Co;
            uint32_t fresh = 0;                                   // 0 for the first MMA into each accumulator of this tile
            const int n0 = decode_tile(P, tile).n0;
            const unsigned NB = (unsigned)P.N;                    // band planes outside [0, NB) are zero padding: skipped
            for (int g = 0; g < ngroups; ++g) {
                if (!MODE && P.taps == 27) {
                    const int cn = n0 + g / 9 - 1;
                    if (cn + P.TN <= 0 || cn >= P.N) continue;
                }
                for (int kc = 0; kc < nkc; ++kc) {
                    if (tr) tq = clock64();
                    mbar_wait(a_full(sa), pha);
                    if (tr) tw_a += clock64() - tq;
                    tc_fence_after();
                    const uint32_t a_lo = a_lo0 + sa * a_stride16;
#pragma unroll 1
                    for (int ts = 0; ts < b_stages_per_a; ++ts) {
                        if (tr) tq = clock64();
                        mbar_wait(b_full(sb), phb);
                        if (tr) { tw_b += clock64() - tq; tq = clock64(); }
                        tc_fence_after();
                        if (elect_one()) {
#pragma unroll 1
                          for (int tt = 0; tt < b_taps; ++tt) {
                            const int t = ts * b_taps + tt;                    // REUSE: t = dy*3 + dx
                            const uint32_t fr = fresh | (uint32_t)(tt > 0);    // 0: first MMAs of the tile (overwrite)
                            const uint32_t b_lo = b_lo0 + sb * b_stride16 + (uint32_t)tt * 3u * co16;
                            const uint32_t tap16 = (uint32_t)(t / 3) * line16 + (uint32_t)(t % 3) * vox16;
                            if (do_mma) {
                                if (!MODE) {
#pragma unroll
                                    for (int k = 0; k < KMMA; ++k)
                                        tc_mma_bf16(d0, ((uint64_t)desc_hi_a << 32) | (a_lo + k * 2), ((uint64_t)desc_hi << 32) | (b_lo + k * 2), idesc,
                                                    k == 0 ? fr : 1u);
                                } else if (MERGE) {
                                    // One MMA per input plane q (box plane q+1) and K=16 slice: its A view feeds the output planes
                                    // q+1, q, q-1 through the taps dz = 0, 1, 2, whose weights are consecutive row blocks of the
                                    // stage and whose accumulators are consecutive TMEM column blocks (planes in decreasing order).
                                    // The first tap of a tile has to overwrite: `init` tracks which accumulators already hold
                                    // this tile's data; an MMA whose span is all-fresh overwrites, all-initialised accumulates,
                                    // and a mixed span (first tap only) is issued plane by plane.
                                    uint32_t init = fr ? 0xffffffffu : 0u;
#pragma unroll
                                    for (int i = 0; i < NACC + 2; ++i) {
                                        const int q = merge_order<NACC>(i);                  // -1 .. NACC
                                        if ((unsigned)(n0 + q) >= NB) continue;
                                        const int dz_lo = q + 1 > NACC - 1 ? q + 1 - (NACC - 1) : 0;   // p = q+1-dz <= NACC-1
                                        const int dz_hi = q + 1 < 2 ? q + 1 : 2;                         // p >= 0
                                        const int nblk = dz_hi - dz_lo + 1;
                                        const int p_top = q + 1 - dz_lo;
                                        const uint32_t mask = ((1u << nblk) - 1u) << (p_top - nblk + 1);
                                        const uint32_t have = init & mask;
                                        init |= mask;
                                        const uint32_t a_view = a_lo + (uint32_t)(q + 1 - pl_off) * plane16 + tap16;
                                        if (have == mask || have == 0u) {
                                            const uint32_t id = idesc_base | ((uint32_t)((nblk * Co) >> 3) << 17);
                                            const uint32_t b_view = b_lo + (uint32_t)dz_lo * co16;
                                            const uint32_t d = d0 + (uint32_t)(NACC - 1 - p_top) * Co;
#pragma unroll
                                            for (int k = 0; k < KMMA; ++k)
                                                tc_mma_bf16(d, ((uint64_t)desc_hi_a << 32) | (a_view + k * 2), ((uint64_t)desc_hi << 32) | (b_view + k * 2), id,
                                                            (k == 0 && have == 0u) ? 0u : 1u);
                                        } else {
#pragma unroll
                                            for (int j = 0; j < nblk; ++j) {                 // plane p_top - j through dz_lo + j
                                                const uint32_t b_view = b_lo + (uint32_t)(dz_lo + j) * co16;
                                                const uint32_t d = d0 + (uint32_t)(NACC - 1 - (p_top - j)) * Co;
                                                const uint32_t acc_on = (have >> (p_top - j)) & 1u;
#pragma unroll
                                                for (int k = 0; k < KMMA; ++k)
                                                    tc_mma_bf16(d, ((uint64_t)desc_hi_a << 32) | (a_view + k * 2), ((uint64_t)desc_hi << 32) | (b_view + k * 2), idesc,
                                                                k == 0 ? acc_on : 1u);
                                            }
                                        }
                                    }
                                } else {
                                    // one MMA per (output plane pp, dz); the first valid dz of each plane overwrites its accumulator
#pragma unroll
                                    for (int dz = 0; dz < 3; ++dz)
#pragma unroll
                                        for (int pp = 0; pp < NACC; ++pp) {
                                            if ((unsigned)(n0 + pp + dz - 1) >= NB) continue;
                                            const int dz_first = (n0 + pp == 0) ? 1 : 0;
                                            const uint32_t a_view = a_lo + (uint32_t)(pp + dz - pl_off) * plane16 + tap16;
                                            const uint32_t b_view = b_lo + (uint32_t)dz * co16;
                                            const uint32_t d = d0 + (uint32_t)(NACC - 1 - pp) * Co;
#pragma unroll
                                            for (int k = 0; k < KMMA; ++k)
                                                tc_mma_bf16(d, ((uint64_t)desc_hi_a << 32) | (a_view + k * 2), ((uint64_t)desc_hi << 32) | (b_view + k * 2), idesc,
                                                            (k == 0 && dz == dz_first) ? fr : 1u);
                                        }
                                }
                            }
                          }
                            tc_commit(b_empty(sb));              // weight stage free once these MMAs retire
                        }
                        __syncwarp();
                        fresh = 1;
                        if (++sb == (uint32_t)P.SB) { sb = 0; phb ^= 1; }
                        if (tr) t_issue += clock64() - tq;
                    }
                    if (elect_one()) tc_commit(a_empty(sa));
                    __syncwarp();
                    if (++sa == (uint32_t)P.SA) { sa = 0; pha ^= 1; }
                }
            }
            for (int kc = 0; kc < P.nkc2; ++kc) {   // fused 1x1x1 second source: one MMA group per output plane, K = KC
                mbar_wait(a_full(sa), pha);
                mbar_wait(b_full(sb), phb);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t a_lo = a_lo0 + sa * a_stride16, b_lo = b_lo0 + sb * b_stride16;
                    const uint32_t plane2_16 = P.a2_piece_bytes >> 4;
                    if (do_mma) {
#pragma unroll
                        for (int pp = 0; pp < NACC; ++pp) {
                            if (MODE && (unsigned)(n0 + pp) >= NB) continue;
                            const uint32_t a_view = a_lo + (uint32_t)pp * plane2_16;
                            const uint32_t d = d0 + (uint32_t)(NACC - 1 - pp) * Co;
#pragma unroll
                            for (int k = 0; k < KMMA; ++k)
                                tc_mma_bf16(d, ((uint64_t)desc_hi << 32) | (a_view + k * 2), ((uint64_t)desc_hi << 32) | (b_lo + k * 2), idesc, 1u);
                        }
                    }
                    tc_commit(b_empty(sb));
                    tc_commit(a_empty(sa));
                }
                __syncwarp();
                if (++sb == (uint32_t)P.SB) { sb = 0; phb ^= 1; }
                if (++sa == (uint32_t)P.SA) { sa = 0; pha ^= 1; }
            }
            if (elect_one()) tc_commit(acc_full(ab));
            __syncwarp();
            if (tr && lane == 0 && lt >= 4 && lt < 8) P.trace[201 + 2 * (lt - 4)] = clock64();          // timeline: last MMA issued + commit
        }
        if (tr && lane == 0) {
            P.trace[0] = clock64() - t_all0; P.trace[1] = tw_acc; P.trace[2] = tw_a; P.trace[3] = tw_b; P.trace[4] = t_issue;
            P.trace[6] = lt;
        }
    } else if (warp == 3) {
        if (MODE) {
            // ================= helper: tile records + per-sample epilogue vectors, one tile ahead =================
            // Joins the epilogue warps' per-tile named barrier: after barrier lt nobody reads buffer (lt + 1) & 1 any more (it
            // served tile lt - 1), so the record of tile lt + 1 is written there and published by barrier lt + 1.
            const ConvEpilogue& e = P.epi;
            const int n_cst = 5 * P.Co;
            auto prepare = [&](int tile, int buf) {
                const TileCoord tc = decode_tile(P, tile);
                const int cb = tc.ch * P.Co;
                float* dst = &cst[buf][0];
#pragma unroll 4
                for (int idx = lane; idx < n_cst; idx += 32) {
                    const int v = (int)fdiv((uint32_t)idx, P.fd[4]), c = idx - v * P.Co + cb;
                    float val;
                    if (v == 0) val = e.bias ? __ldg(e.bias + (size_t)tc.b0 * e.ld_bias + c) : 0.f;
                    else if (v < 3) val = e.tb[v - 1] ? 0.5f * __ldg(e.tb[v - 1] + (size_t)tc.b0 * e.ld_tb + c) : 0.f;   // stored halved (store_views16)
                    else val = e.sc[v - 3] ? __ldg(e.sc[v - 3] + (size_t)tc.b0 * e.ld_sc + c) : 1.f;
                    dst[idx] = val;
                }
                if (lane == 0) {
                    TileInfo ti;
                    ti.b0 = tc.b0; ti.n0 = tc.n0; ti.y0 = tc.y0; ti.x0 = tc.x0; ti.cb = cb; ti.pad = 0;
                    const long long bn0 = (long long)tc.b0 * P.N + tc.n0;
                    ti.vox_base = (bn0 * P.H + tc.y0) * P.W + tc.x0;
                    ti.hvox_base = (bn0 * (P.H / 2) + tc.y0 / 2) * (P.W / 2) + tc.x0 / 2;
                    tinfo[buf] = ti;
                }
                // (Requesting the tile's residual rows into L2 from here, one tile ahead, was measured: 870 -> 842 patches/s,
                // profiles/ab_r02l_l2pf.txt -- the rows are L2 hits already; what the epilogue waits for is its own request queue.)
            };
            uint32_t lt = 0;
            prepare(blockIdx.x, 0);
            for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, ++lt) {
                asm volatile("bar.sync 1, %0;" ::"n"(128 * kEpiPerQuarter + 32) : "memory");
                const int nt = tile + gridDim.x;
                if (nt < P.num_tiles) prepare(nt, (lt + 1) & 1);
            }
        }
    } else if (warp >= 4) {
        // ================= epilogue =================
        const int q = warp & 3;                           // TMEM lane quarter this warp may read
        const int sub = (warp - 4) >> 2;                  // which of the quarter's warps
        const int r = q * 32 + lane;                      // accumulator row = voxel of the tile
        uint32_t lt = 0;
        long long tw_full = 0, tw_ld = 0, tw_bar = 0, tw_sync = 0, t_all0 = clock64(), tq = 0;
        const bool tr = P.trace && blockIdx.x == 0 && lane == 0;      // lane 0 of every epilogue warp of CTA 0
        const size_t plane_vox = (size_t)P.H * P.W, hplane_vox = (size_t)(P.H / 2) * (P.W / 2);
        // REUSE kernels: a tile lies in one sample; its coordinates and per-sample epilogue vectors [bias | tb0 | tb1 | sc0 | sc1][Co]
        // are prepared in shared memory one tile ahead by the helper warp (warp 3), so the epilogue warps neither decode tiles
        // nor wait for global loads of those vectors (their per-tile fixed cost was ~4000 of ~9000 cycles at Co = 32,
        // profiles/umma_timeline_r02c.txt).
        // the row's place in its tile: REUSE (x, y) of the 16 x 8 plane tile; SIMPLE (x, y, band plane, sample) of the box
        int x_in, y_in, n_in = 0, b_in = 0;
        if (MODE) { x_in = r % P.TW; y_in = r / P.TW; }
        else {
            int t = r;
            x_in = t % P.TW; t /= P.TW;
            y_in = t % P.TH; t /= P.TH;
            n_in = t % P.TN; b_in = t / P.TN;
        }
        const int lane_vox = y_in * P.W + x_in, lane_hvox = (y_in >> 1) * (P.W >> 1) + (x_in >> 1);     // REUSE: offsets from the tile origin
        constexpr int kBarCount = 128 * kEpiPerQuarter + (MODE ? 32 : 0);
        const int nslab = P.Co >> 4, nitems = NACC * nslab;
        const bool normal = Epi<EF>::mode(P.epi) == CONV_NORMAL && (Epi<EF>::add0(P.epi) || Epi<EF>::add1(P.epi));    // residual read-ahead
        int n_stamp = 0;
        auto stamp = [&]() { if (tr && warp == 4 && lt >= 4 && lt < 8 && n_stamp < 24) P.trace[96 + 24 * (lt - 4) + n_stamp++] = clock64(); };
        for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, ++lt) {
            n_stamp = 0;
            stamp();                                     // 0: loop top
            const int ab = lt & (uint32_t)(P.acc_stages - 1);
            // one warp probes the accumulator barrier for all twelve (every probe is a wavefront on the shared-memory data pipe
            // the tensor core reads through: +2.6 %); the named barrier hands the observation on (and publishes the helper
            // warp's tile record), tcgen05.fence::after_thread_sync orders the TMEM loads behind it
            if (tr) tq = clock64();
            if (warp == 4) mbar_wait(acc_full(ab), (lt >> P.acc_shift) & 1);
            if (tr) tw_full += clock64() - tq;
            stamp();                                     // 1: accumulator full
            asm volatile("bar.sync 1, %0;" ::"n"(kBarCount) : "memory");
            tc_fence_after();
            int b, y, x, n0, cb;
            size_t vox0, hvox0;
            const float* cs = nullptr;
            if (MODE) {
                const TileInfo& ti = tinfo[lt & 1];
                b = ti.b0; n0 = ti.n0; cb = ti.cb;
                y = y_in + ti.y0; x = x_in + ti.x0;
                vox0 = (size_t)ti.vox_base + lane_vox;
                hvox0 = (size_t)ti.hvox_base + lane_hvox;
                cs = &cst[lt & 1][0];
            } else {
                const TileCoord tc = decode_tile(P, tile);
                b = tc.b0 + b_in; n0 = tc.n0 + n_in; cb = tc.ch * P.Co;
                y = y_in + tc.y0; x = x_in + tc.x0;
                const size_t bn0 = (size_t)b * P.N + n0;
                vox0 = (bn0 * P.H + y) * P.W + x;
                hvox0 = (bn0 * (P.H / 2) + y / 2) * (P.W / 2) + x / 2;
            }
            // work items = (plane, 16-channel slab); the warps of a lane quarter take items round-robin, fetch the residual
            // operands of their next item and start the TMEM read of the next slab before working on the current one
            const bool in_tile = b < P.B && b_in < P.TB && y < P.H && x < P.W && !Epi<EF>::dbg(P.epi, 8);
            int pl = 0, sl = sub;                           // item = (plane pl, slab sl)
            while (sl >= nslab) { sl -= nslab; ++pl; }
            const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * NACC) * P.Co;
            Resid cur, nxt;
            uint32_t ra[16], rb[16];
            // (Requesting the residual rows of three work items together instead of one item ahead was measured: the first row
            // still takes ~3500 cycles to arrive and the tile takes as long as before, profiles/umma_timeline_r02l.txt -- the rows
            // queue behind the stores of the twelve epilogue warps in the L1 request path, which handles one 128-byte line of a
            // warp's access per ~2 cycles, profiles/store_patterns_r02.txt.)
            if (pl < NACC) {
                if (MODE) tc_ld16_issue(t_lane + (uint32_t)(NACC - 1 - pl) * P.Co + sl * 16, ra);
                if (normal) load_resid<EF>(P.epi, vox0 + (size_t)pl * plane_vox, hvox0 + (size_t)pl * hplane_vox, cb + sl * 16, in_tile && n0 + pl < P.N, cur);
            }
            stamp();                                     // 2: barrier + first TMEM read / residual prefetch issued
            for (int it = sub; it < nitems; it += kEpiPerQuarter) {
                stamp();                                 // item: start
                int pl2 = pl, sl2 = sl + kEpiPerQuarter;
                while (sl2 >= nslab) { sl2 -= nslab; ++pl2; }
                if (tr) tq = clock64();
                float acc[16];
                if (MODE) {
                    tc_ld16_wait(ra);
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc[i] = __uint_as_float(ra[i]);
                } else {
                    tc_ld16(t_lane + (uint32_t)(NACC - 1 - pl) * P.Co + sl * 16, acc);                  // SIMPLE kernels: no read-ahead (registers)
                }
                if (tr) { tw_ld += clock64() - tq; tq = clock64(); }
                if (pl2 < NACC) {
                    if (MODE) tc_ld16_issue(t_lane + (uint32_t)(NACC - 1 - pl2) * P.Co + sl2 * 16, rb);
                    if (normal) load_resid<EF>(P.epi, vox0 + (size_t)pl2 * plane_vox, hvox0 + (size_t)pl2 * hplane_vox, cb + sl2 * 16, in_tile && n0 + pl2 < P.N, nxt);
                }
                const bool valid = in_tile && n0 + pl < P.N;
                stamp();                                 // item: accumulator slab in registers
                epilogue16<EF, MODE != 0>(P, acc, sl * 16, b, y, x, vox0 + (size_t)pl * plane_vox, hvox0 + (size_t)pl * hplane_vox, valid, cur, cs, cb);
                if (tr) tw_bar += clock64() - tq;
                stamp();                                 // item: done
                cur = nxt;
                if (MODE) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) ra[i] = rb[i];
                }
                pl = pl2; sl = sl2;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty(ab));
            stamp();                                     // accumulator released
        }
        if (tr) {
            if (warp == 4) { P.trace[8] = clock64() - t_all0; P.trace[9] = tw_full; P.trace[10] = tw_ld; P.trace[11] = tw_bar; }
            long long* w = P.trace + 16 + 4 * (warp - 4);
            w[0] = clock64() - t_all0; w[1] = tw_sync; w[2] = tw_ld; w[3] = tw_bar;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 3) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(P.tmem_cols) : "memory");
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side: geometry, tensor maps, launch
// ------------------------------------------------------------------------------------------
struct UmmaPlan {
    UmmaParams p;
    int grid;
    size_t smem;
};

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            return nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

int floor_pow2(int v) {
    int p = 1;
    while (p * 2 <= v) p *= 2;
    return p;
}

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

bool umma_disabled() {
    static const bool off = []() {
        const char* e = getenv("TMD_DISABLE_UMMA");
        return e && e[0] == '1';
    }();
    return off;
}

// per device: a process may hold engines on several GPUs (the Python layer rebuilds the engine when the device changes)
constexpr int kMaxDevices = 64;
int sm_count() {
    static int n[kMaxDevices] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) dev = 0;
    if (!n[dev]) {
        int v = 148;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        n[dev] = v;
    }
    return n[dev];
}

}  // namespace

#if TMD_UMMA_PART == 0
bool conv_umma_supported(const ConvOp& op) {
    if (umma_disabled()) return false;
    if (op.taps != 27 && op.taps != 1) return false;
    if (op.Ci % 32 || op.Co % 16 || op.Co < 16 || op.Co > 256) return false;
    if (op.ld_in % 8 || (reinterpret_cast<uintptr_t>(op.src) & 15)) return false;
    const ConvEpilogue& e = op.epi;
    // 32-byte vector accesses in the epilogue
    if ((e.raw_out && e.ld_raw % 16) || (e.add0 && e.ld_add0 % 16) || (e.add1 && e.ld_add1 % 16) || (e.hf_out[0] && e.ld_hf % 16)) return false;
    for (int k = 0; k < 2; ++k)
        if (e.act_out[k] && e.ld_act[k] % 16) return false;
    {
        const void* wide[] = {e.add0, e.add1, e.raw_out, e.act_out[0], e.act_out[1], e.hf_out[0], e.hf_out[1], e.hf_out[2]};
        for (const void* q : wide)
            if (reinterpret_cast<uintptr_t>(q) & 31) return false;
    }
    if ((e.tb[0] || e.tb[1]) && e.ld_tb % 4) return false;
    if ((e.sc[0] || e.sc[1]) && e.ld_sc % 4) return false;
    const void* ptrs[] = {e.bias, e.add0, e.add1, e.raw_out, e.act_out[0], e.act_out[1], e.tb[0], e.tb[1], e.sc[0], e.sc[1],
                          e.hf_out[0], e.hf_out[1], e.hf_out[2], op.w_bf16};
    for (const void* q : ptrs)
        if (reinterpret_cast<uintptr_t>(q) & 15) return false;
    if ((e.mode == CONV_POOL || e.pool_out) && (op.H % 2 || op.W % 2)) return false;
    if (e.pool_out && (e.mode != CONV_NORMAL || e.ld_pool % 16 || (reinterpret_cast<uintptr_t>(e.pool_out) & 31))) return false;
    if (e.proj_out && (e.mode != CONV_NORMAL || !e.proj_w || (reinterpret_cast<uintptr_t>(e.proj_w) & 15) || e.act_out[0])) return false;
    if (op.src2) {
        const int kc = (op.taps == 27 && op.H >= 16 && op.W >= 8) ? 32 : ((op.Ci % 64 == 0) ? 64 : 32);   // KC of the plan below
        if (op.src2_gap && (op.src2_gap_at % kc || op.src2_gap % kc)) return false;
        if (op.Ci2 % kc || op.ld_in2 % 8 || (reinterpret_cast<uintptr_t>(op.src2) & 15) || (reinterpret_cast<uintptr_t>(op.w2_bf16) & 15)) return false;
    }
    if (op.W < 2 || op.H < 2) return false;
    if (op.w_sample_stride && (op.w_sample_stride != (long long)op.taps * op.Ci * op.Co || op.src2)) return false;
    if (e.bias && e.ld_bias % 4) return false;
    return true;
}

bool conv_umma_skips_padding_planes(const ConvOp& op) {
    if (op.taps != 27) return false;
    if (op.H >= 16 && op.W >= 8) return true;                       // REUSE: planes outside [0, N) are neither loaded nor issued
    // SIMPLE: only when a box holds one band plane (the whole-padding boxes are skipped)
    const int TW = floor_pow2(op.W < 16 ? op.W : 16), TH = floor_pow2(op.H < 128 / TW ? op.H : 128 / TW);
    const int TN = floor_pow2(op.N < 128 / (TW * TH) ? op.N : 128 / (TW * TH));
    return TN == 1 || op.B >= 128 / (TW * TH);
}

UmmaPlan* conv_umma_plan_create(const ConvOp& op) {
    EncodeTiledFn enc = encode_fn();
    if (!enc) return nullptr;
    auto* plan = new UmmaPlan();
    UmmaParams& P = plan->p;
    memset(&P, 0, sizeof(P));
    P.epi = op.epi;
    P.B = op.B; P.N = op.N; P.H = op.H; P.W = op.W; P.Ci = op.Ci; P.Co = op.Co; P.taps = op.taps;
    P.co_split = 1;
    P.mode = (op.taps == 27 && op.H >= 16 && op.W >= 8) ? 1 : 0;
    cuuint32_t boxA[5];
    if (P.mode) {
        P.KC = 32;
        P.TW = 8; P.TH = 16; P.TN = 1; P.TB = 1;
        // accumulators per tile: two sets must fit the 512 TMEM columns, the haloed box must leave room for
        // two stages in shared memory, and NACC divides the band count (no plane is computed for nothing)
        // Co = 256: a tile computes half of the output channels for two band planes instead of all of them for one, so a
        // weight stage (the dominant L2 -> shared-memory stream of these layers) feeds twice the MMA work
        if (op.Co == 256 && op.N % 2 == 0 && env_int("TMD_UMMA_COSPLIT", 1)) { P.co_split = 2; P.Co = 128; }
        int cap = 256 / P.Co;
        if (cap > 8) cap = 8;
        P.NACC = 1;
        for (int c = 2; c <= cap; c *= 2)
            if (op.N % c == 0) P.NACC = c;
        // Small batches (BASELINE config 1 is ONE patch): with the widest tile there are fewer tiles than SMs (32 for one 8-band
        // 64x64 patch at 8 planes per tile), and a persistent grid is only as fast as its busiest CTA: trade planes per tile for
        // tiles until the machine is covered.
        if (env_int("TMD_UMMA_SMALLB", 1)) {
            const long long per_plane_group = (long long)((op.W + 7) / 8) * ((op.H + 15) / 16) * op.B * P.co_split;
            while (P.NACC > 1 && 2 * per_plane_group * ((op.N + P.NACC - 1) / P.NACC) <= sm_count()) P.NACC /= 2;   // less than half the SMs busy
        }
        P.pl_off = (P.NACC == op.N) ? 1 : 0;
        P.a_planes = P.NACC + 2 - 2 * P.pl_off;
        P.a_ysplit = 1;
        boxA[0] = P.KC; boxA[1] = P.TW + 2; boxA[2] = P.TH + 2; boxA[3] = 1; boxA[4] = 1;   // one TMA instruction per haloed plane
        P.tiles_n = (op.N + P.NACC - 1) / P.NACC;
    } else {
        P.KC = (op.Ci % 64 == 0) ? 64 : 32;
        P.TW = floor_pow2(op.W < 16 ? op.W : 16);
        P.TH = floor_pow2(op.H < 128 / P.TW ? op.H : 128 / P.TW);
        P.TN = floor_pow2(op.N < 128 / (P.TW * P.TH) ? op.N : 128 / (P.TW * P.TH));
        // 3x3x3 on small planes: one band plane per box (batch folded into M instead), so the dz taps that fall on the
        // zero padding of the band axis are skipped outright
        if (op.taps == 27 && op.B >= 128 / (P.TW * P.TH)) P.TN = 1;
        P.TB = 128 / (P.TW * P.TH * P.TN);
        if (op.w_sample_stride) P.TB = 1;              // per-sample weights: a tile may not span samples (rows beyond the box are not stored)
        P.NACC = 1;
        boxA[0] = P.KC; boxA[1] = P.TW; boxA[2] = P.TH; boxA[3] = P.TN; boxA[4] = P.TB;
        P.tiles_n = (op.N + P.TN - 1) / P.TN;
    }
    P.nkc = op.Ci / P.KC;
    P.tiles_x = (op.W + P.TW - 1) / P.TW;
    P.tiles_y = (op.H + P.TH - 1) / P.TH;
    P.tiles_b = (op.B + P.TB - 1) / P.TB;
    P.num_tiles = P.tiles_x * P.tiles_y * P.tiles_n * P.tiles_b * P.co_split;
    {
        const uint32_t ds[5] = {(uint32_t)P.co_split, (uint32_t)P.tiles_x, (uint32_t)P.tiles_y, (uint32_t)P.tiles_n, (uint32_t)P.Co};
        for (int i = 0; i < 5; ++i) {          // round-up magic numbers: n / d == (umulhi(n, m) + n) >> s for n < 2^31
            uint32_t sh = 0;
            while ((1u << sh) < ds[i]) ++sh;
            P.fd[i][0] = (uint32_t)((((uint64_t)1 << 32) * (((uint64_t)1 << sh) - ds[i])) / ds[i] + 1);
            P.fd[i][1] = sh;
        }
    }
    P.row_bytes = P.KC * 2;
    P.layout_type = P.KC == 64 ? 2u : 4u;           // SWIZZLE_128B : SWIZZLE_64B
    P.sbo16 = (8 * P.row_bytes) >> 4;               // 8-row swizzle atoms are contiguous
    P.a_piece_bytes = boxA[1] * boxA[2] * boxA[3] * boxA[4] * P.row_bytes;
    P.a_stage_bytes = P.mode ? P.a_piece_bytes * P.a_planes * P.a_ysplit : P.a_piece_bytes;
    P.dbg = env_int("TMD_UMMA_DEBUG", 0);
    P.epi.dbg = P.dbg;
    P.a_stage_stride = (P.a_stage_bytes + 1023u) & ~1023u;
    P.b_tile_bytes = P.Co * P.row_bytes;
    // REUSE: a weight stage holds the 3 dz taps of b_taps (dy,dx) taps.  Small Co: three or all nine taps per stage, so that a
    // stage carries enough MMA work to hide the issue warp's per-stage wait / commit round (Co <= 32: 9 taps +1 % over 3;
    // the ring-depth loop below falls back to fewer taps where shared memory is short).
    P.b_taps = 1;
    if (P.mode) {
        P.b_taps = env_int("TMD_UMMA_BTAPS", P.Co <= 32 ? 9 : (P.Co <= 64 ? 3 : 1));
        if (P.b_taps != 1 && P.b_taps != 3 && P.b_taps != 9) P.b_taps = 1;
    }
    // accumulator sets: two overlap the epilogue of tile i with the MMAs of tile i+1.  Four (possible when 4*NACC*Co <= 512,
    // TMD_UMMA_ACC4=1) were measured for the Co = 32 layers: no change -- the two roles do not lose time to jitter but to the
    // shared L1 / shared-memory data pipe
    P.acc_stages = (4 * P.NACC * P.Co <= 512 && env_int("TMD_UMMA_ACC4", 0)) ? 4 : 2;
    P.acc_shift = P.acc_stages == 4 ? 2 : 1;
    int cols = P.acc_stages * P.NACC * P.Co;
    P.tmem_cols = 32;
    while ((int)P.tmem_cols < cols) P.tmem_cols *= 2;
    // instruction descriptor: D=f32, A=B=bf16, both K-major, N>>3 at [17,23), M>>4 at [24,29)
    P.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(P.Co >> 3) << 17) | ((128u >> 4) << 24);

    // ring depths from the shared-memory budget
    const size_t budget = 215 * 1024;              // 227 KB less the kernel's static shared memory (barriers, epilogue vectors)
    for (;; P.b_taps /= 3) {
        P.b_tile_stride = ((P.mode ? 3 * P.b_taps : 1) * P.b_tile_bytes + 1023u) & ~1023u;
        // activation stages first (their TMA round trip is the long one), at least four weight stages behind them
        const int want_sb = P.mode ? (P.b_taps == 1 ? 6 : (P.b_taps == 3 ? 4 : 2)) : 4;
        // REUSE with a fused second source: its chunks share the activation ring with the haloed box, and with two stages every
        // chunk waits a full TMA round trip for its slot (profiles/umma_trace_r02a.txt: the MMA warp of final.conv20.conv21 waited for
        // loads 60 % of the time) -- three stages where shared memory allows (fewer taps per weight stage make the room)
        const int want_sa = (P.mode && op.src2) ? 3 : 2;
        P.SA = P.mode ? env_int("TMD_UMMA_SA", want_sa) : 4;
        while (P.SA > 2 && (size_t)P.SA * P.a_stage_stride + 4 * (size_t)P.b_tile_stride > budget) --P.SA;
        while (P.SA > 1 && (size_t)P.SA * P.a_stage_stride + 2 * (size_t)P.b_tile_stride > budget) --P.SA;
        size_t left = budget - (size_t)P.SA * P.a_stage_stride;
        P.SB = (int)(left / P.b_tile_stride);
        if (P.SB > want_sb) P.SB = want_sb;
        if (P.SB > kMaxSB) P.SB = kMaxSB;
        if (P.b_taps == 1 || (P.SA >= want_sa && P.SB >= (P.b_taps == 3 ? 3 : 2))) break;   // else: fewer taps per stage
    }
    if (P.SB < 2 || P.SA < 1) { delete plan; return nullptr; }
    plan->smem = (size_t)P.SA * P.a_stage_stride + (size_t)P.SB * P.b_tile_stride + 1024;

    const CUtensorMapSwizzle sw = P.KC == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    {   // activations [B][N][H][W][ld] bf16, channels innermost
        cuuint64_t dims[5] = {(cuuint64_t)op.Ci, (cuuint64_t)op.W, (cuuint64_t)op.H, (cuuint64_t)op.N, (cuuint64_t)op.B};
        const cuuint64_t e = 2ull * op.ld_in;
        cuuint64_t strides[4] = {e, e * op.W, e * op.W * op.H, e * op.W * op.H * op.N};
        cuuint32_t estr[5] = {1, 1, 1, 1, 1};
        if (enc(&P.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(op.src), dims, strides, boxA, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            delete plan;
            return nullptr;
        }
    }
    {   // weights [Co][taps*Ci] bf16
        P.w_rows_per_sample = op.w_sample_stride ? op.Co : 0;
        cuuint64_t dims[2] = {(cuuint64_t)op.taps * op.Ci, (cuuint64_t)op.Co * (op.w_sample_stride ? op.B : 1)};
        cuuint64_t strides[1] = {2ull * op.taps * op.Ci};
        cuuint32_t box[2] = {(cuuint32_t)P.KC, (cuuint32_t)P.Co};
        cuuint32_t estr[2] = {1, 1};
        if (enc(&P.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(op.w_bf16), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            delete plan;
            return nullptr;
        }
    }
    if (op.src2) {
        P.nkc2 = op.Ci2 / P.KC;
        P.c2_gap_at = op.src2_gap ? op.src2_gap_at : (1 << 30);
        P.c2_gap = op.src2_gap;
        cuuint32_t box2[5] = {boxA[0], boxA[1], boxA[2], boxA[3], boxA[4]};
        if (P.mode) { box2[1] = P.TW; box2[2] = P.TH; }   // no halo: the 1x1x1 conv only reads the tile's own voxels
        P.a2_piece_bytes = box2[1] * box2[2] * box2[3] * box2[4] * P.row_bytes;
        cuuint64_t dims[5] = {(cuuint64_t)(op.Ci2 + op.src2_gap), (cuuint64_t)op.W, (cuuint64_t)op.H, (cuuint64_t)op.N, (cuuint64_t)op.B};
        const cuuint64_t e = 2ull * op.ld_in2;
        cuuint64_t strides[4] = {e, e * op.W, e * op.W * op.H, e * op.W * op.H * op.N};
        cuuint32_t estr[5] = {1, 1, 1, 1, 1};
        cuuint64_t dimsB[2] = {(cuuint64_t)op.Ci2, (cuuint64_t)op.Co};
        cuuint64_t stridesB[1] = {2ull * op.Ci2};
        cuuint32_t boxB[2] = {(cuuint32_t)P.KC, (cuuint32_t)P.Co};
        if (enc(&P.tmA2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(op.src2), dims, strides, box2, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
            enc(&P.tmB2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(op.w2_bf16), dimsB, stridesB, boxB, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            delete plan;
            return nullptr;
        }
    }
    plan->grid = P.num_tiles < sm_count() ? P.num_tiles : sm_count();
    return plan;
}

void conv_umma_plan_destroy(UmmaPlan* p) { delete p; }
#endif

namespace {
typedef void (*KernelFn)(const UmmaParams);
// the three dz taps that share an activation view are issued as one MMA when N = 3*Co fits an instruction
bool merged(const UmmaParams& p) { return p.mode && p.NACC > 1 && 3 * p.Co <= 256; }

// Epilogue shapes with a kernel of their own (see Epi<EF>): every shape the sampling path gives the 4- and 8-accumulator REUSE
// kernels, i.e. the level-0 / level-1 layers with 32 or 64 output channels, whose time is the epilogue's (profiles/
// umma_trace_r02a.txt).  Everything else -- the 128-channel and 8x8 layers, whose epilogue waits for the MMAs, the operator API,
// TMD_UMMA_DEBUG runs -- takes the generic kernels.
constexpr int kSpecEF[] = {
    EF_ACT0,                                                     // ResBlock conv20
    EF_RAW,                                                      // convH_0, last condition-branch Conv_1
    EF_RAW | EF_ACT0,                                            // stems / Conv_1 of the condition branch, up-path conv21 (+res_conv)
    EF_RAW | EF_ACT0 | EF_ACT1,                                  // conv2.conv21, main-branch Conv_1 (+Conv_2)
    EF_RAW | EF_ACT0 | EF_POOLOUT,                               // down-path conv21 (+res_conv, +pooled copy)
    EF_POOL | EF_ACT0,                                           // wavelet-down Conv_0
    EF_POOL | EF_ACT0 | EF_HF,                                   // ... of the condition branch (high bands kept)
    EF_UP | EF_ACT0 | EF_ADD0,                                   // wavelet-up Conv_0
    EF_RAW | EF_ACT0 | EF_ADD0 | EF_ADD1,                        // wavelet-up Conv_1
    EF_RAW | EF_RAWF16 | EF_ACT0,                                // final.conv20.conv21
    EF_RAW | EF_RAWF16 | EF_ACT0 | EF_ADD0 | EF_ADD0F16,         // final.conv21/22.conv21
    EF_ACT0 | EF_ADD0 | EF_ADD0F16,                              // final.conv23.conv21
    EF_ACT0 | EF_ADD0 | EF_ADD0F16 | EF_PROJ,                    // ... with final.conv24 (32 -> 1) in its epilogue
    EF_ACT0 | EF_ADD0 | EF_PROJ,
    EF_RAW | EF_ACT0 | EF_ADD0 | EF_ADD0F16,                     // up-path conv21 (+ the step-variant two thirds of res_conv; the rest arrives as add0)
    EF_RAW | EF_ACT0 | EF_ADD0,                                  // identity-skip ResBlock without the half-precision stream
    EF_ACT0 | EF_ADD0,
};
constexpr int kNumSpecEF = (int)(sizeof(kSpecEF) / sizeof(kSpecEF[0]));

int epi_flags(const ConvEpilogue& e) {
    int f = e.mode & 3;
    if (e.raw_out) f |= EF_RAW;
    if (e.act_out[0] || e.proj_out) f |= EF_ACT0;
    if (e.proj_out) f |= EF_PROJ;
    if (e.act_out[1]) f |= EF_ACT1;
    if (e.add0) f |= EF_ADD0;
    if (e.add1) f |= EF_ADD1;
    if (e.pool_out) f |= EF_POOLOUT;
    if (e.hf_out[0]) f |= EF_HF;
    if (e.raw_f16) f |= EF_RAWF16;
    if (e.add0_f16) f |= EF_ADD0F16;
    return f;
}

template <int MODE, int NACC, int KMMA, bool MERGE, int... I>
KernelFn find_special(int ef, std::integer_sequence<int, I...>) {
    KernelFn f = nullptr;
    (void)std::initializer_list<int>{(ef == kSpecEF[I] ? (f = conv_umma_kernel<MODE, NACC, KMMA, MERGE, kSpecEF[I]>, 0) : 0)...};
    return f;
}
}  // namespace

// One translation unit would take minutes to compile ~35 instantiations of the kernel: the library builds this file three
// times (tmdiff_b200/build.py), TMD_UMMA_PART = 0 host side + generic kernels, 1-4 the specialised kernels of one schedule each.
void* conv_umma_special_nacc4(int ef);
void* conv_umma_special_nacc8(int ef);
void* conv_umma_special_nacc2(int ef);      // two accumulators, one MMA per (plane, dz): the 128-channel layers
void* conv_umma_special_simple64(int ef);   // SIMPLE schedule, 64-channel chunks: the 8x8 level and the 1x1x1 convs
void* conv_umma_special_simple32(int ef);   // SIMPLE schedule, 32-channel chunks (Ci = 32 or 96)
#if TMD_UMMA_PART == 1
void* conv_umma_special_nacc4(int ef) { return (void*)find_special<1, 4, 2, true>(ef, std::make_integer_sequence<int, kNumSpecEF>()); }
#elif TMD_UMMA_PART == 2
void* conv_umma_special_nacc8(int ef) { return (void*)find_special<1, 8, 2, true>(ef, std::make_integer_sequence<int, kNumSpecEF>()); }
#elif TMD_UMMA_PART == 3
void* conv_umma_special_nacc2(int ef) { return (void*)find_special<1, 2, 2, false>(ef, std::make_integer_sequence<int, kNumSpecEF>()); }
#elif TMD_UMMA_PART == 4
void* conv_umma_special_simple64(int ef) { return (void*)find_special<0, 1, 4, false>(ef, std::make_integer_sequence<int, kNumSpecEF>()); }
#elif TMD_UMMA_PART == 5
void* conv_umma_special_simple32(int ef) { return (void*)find_special<0, 1, 2, false>(ef, std::make_integer_sequence<int, kNumSpecEF>()); }
#endif

#if TMD_UMMA_PART == 0
namespace {
KernelFn pick_kernel(const UmmaParams& p) {
    // TMD_UMMA_DEBUG / TMD_UMMA_GENERIC=1: the generic kernels (the debug bits live only there)
    if (!p.dbg && !env_int("TMD_UMMA_GENERIC", 0)) {
        void* f = nullptr;
        if (p.mode && p.NACC == 4) f = conv_umma_special_nacc4(epi_flags(p.epi));
        else if (p.mode && p.NACC == 8) f = conv_umma_special_nacc8(epi_flags(p.epi));
        else if (p.mode && p.NACC == 2 && !merged(p)) f = conv_umma_special_nacc2(epi_flags(p.epi));
        else if (!p.mode && p.KC == 64) f = conv_umma_special_simple64(epi_flags(p.epi));
        else if (!p.mode && p.KC == 32) f = conv_umma_special_simple32(epi_flags(p.epi));
        if (f) return (KernelFn)f;
    }
    if (p.mode) {
        if (p.NACC == 8) return conv_umma_kernel<1, 8, 2, true, -1>;
        if (p.NACC == 4) return conv_umma_kernel<1, 4, 2, true, -1>;
        if (p.NACC == 2) return merged(p) ? conv_umma_kernel<1, 2, 2, true, -1> : conv_umma_kernel<1, 2, 2, false, -1>;
        return conv_umma_kernel<1, 1, 2, false, -1>;
    }
    return p.KC == 64 ? conv_umma_kernel<0, 1, 4, false, -1> : conv_umma_kernel<0, 1, 2, false, -1>;
}
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device, per-function setting: done once per (device, kernel)
cudaError_t allow_big_smem(KernelFn f) {
    static std::set<std::pair<int, KernelFn>> done;
    int dev = 0;
    cudaGetDevice(&dev);
    if (done.count({dev, f})) return cudaSuccess;
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, f);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - (int)fa.sharedSizeBytes);   // static + dynamic <= 227 KB
    if (e != cudaSuccess) return e;
    done.insert({dev, f});
    return cudaSuccess;
}
}  // namespace

cudaError_t conv_umma_bf16(const ConvOp& op, const UmmaPlan* plan, cudaStream_t stream) {
    (void)op;
    KernelFn kern = pick_kernel(plan->p);
    {
        const cudaError_t e = allow_big_smem(kern);
        if (e != cudaSuccess) return e;
    }
    if (env_int("TMD_UMMA_TRACE", 0)) {
        UmmaParams p = plan->p;
        cudaMalloc((void**)&p.trace, 256 * sizeof(long long));
        cudaMemset(p.trace, 0, 256 * sizeof(long long));
        kern<<<plan->grid, kThreads, plan->smem, stream>>>(p);
        cudaDeviceSynchronize();
        long long t[256];
        cudaMemcpy(t, p.trace, sizeof(t), cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        const long long mmas = t[6] * (p.mode ? 27 : p.taps) * p.nkc * p.NACC * (p.KC / 16);
        fprintf(stderr, "[umma trace] epi=%d%s%s%s%s%s ", p.epi.mode, p.epi.raw_out ? "+raw" : "", p.epi.act_out[0] ? "+act0" : "", p.epi.act_out[1] ? "+act1" : "",
                p.epi.add0 ? "+add0" : "", p.epi.add1 ? "+add1" : "");
        fprintf(stderr, "Ci=%d Co=%d %dx%dx%d mode=%d NACC=%d tiles=%d grid=%d | mma warp: total %lld cyc, wait acc %lld, wait A %lld, "
                "wait B %lld, issue %lld (%.1f cyc/mma), tiles %lld | epilogue: total %lld, wait full %lld, tmem ld %lld, compute+store %lld\n",
                p.Ci, p.Co, p.N, p.H, p.W, p.mode, p.NACC, p.num_tiles, plan->grid, t[0], t[1], t[2], t[3], t[4],
                mmas ? (double)t[4] / mmas : 0.0, t[6], t[8], t[9], t[10], t[11]);
        if (env_int("TMD_UMMA_TRACE", 0) > 1) {
            fprintf(stderr, "    epilogue warps (total / barrier wait / tmem ld / compute+store):");
            for (int w = 0; w < 4 * kEpiPerQuarter; ++w) fprintf(stderr, " [w%d %lld %lld %lld %lld]", w + 4, t[16 + 4 * w], t[17 + 4 * w], t[18 + 4 * w], t[19 + 4 * w]);
            fprintf(stderr, "\n");
            // timeline of local tiles 4..7 (cycles relative to the MMA start of tile 4): the MMA warp's [issue start, commit] and
            // warp 4's stamps
            const long long z = t[200];
            for (int i = 0; i < 4; ++i) {
                fprintf(stderr, "    tile %d: mma [%lld, %lld] epi", 4 + i, t[200 + 2 * i] - z, t[201 + 2 * i] - z);
                for (int k = 0; k < 24 && t[96 + 24 * i + k]; ++k) fprintf(stderr, " %lld", t[96 + 24 * i + k] - z);
                fprintf(stderr, "\n");
            }
        }
        return cudaGetLastError();
    }
    static const bool pdl = env_int("TMD_PDL", 0) != 0;     // measured twice on one box: 855 / 850 patches/s with, 862 / 855 without (profiles/ab_r02e_pdl.txt)
    if (pdl) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)plan->grid);
        cfg.blockDim = dim3((unsigned)kThreads);
        cfg.dynamicSmemBytes = plan->smem;
        cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        return cudaLaunchKernelEx(&cfg, kern, plan->p);
    }
    kern<<<plan->grid, kThreads, plan->smem, stream>>>(plan->p);
    return cudaGetLastError();
}

#endif  // TMD_UMMA_PART == 0

}  // namespace tmd
