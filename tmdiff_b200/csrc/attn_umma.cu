// tcgen05 / TMA fused softmax attention for sm_100a (bf16 operands, fp32 logits / statistics / accumulation).
//
// out[bh, i, :] = softmax_j(scale * q[bh,i,:] . k[bh,j,:]) v[bh,j,:] -- the core of SpatialSelfAttention
// (core/Attention.py:140-159: one head as wide as the channel count), CrossAttention (:185-214: 8 heads x 64, 77 text
// keys) and AttnBlockpp (GeneralModel/Hyper_unet_general.py:500-506) in bf16 mode.  The fp32 CUDA-core kernel in
// operators.cu stays the fp32 parity mode.
//
// One CTA owns 128 queries x DV <= 256 output channels of one (batch, head) and walks the keys in blocks of 128:
//   S = Q K^T        128 x 128 fp32 in TMEM, two buffers (the logits of block j+1 are issued before the softmax of block j)
//   P = exp2(S - m)  by the 128 softmax threads (one query row per thread = one TMEM lane: no shuffles), bf16, written to
//                    shared memory in the 128-byte-swizzled K-major layout a TMA load would have produced
//   O += P V         128 x DV fp32 in TMEM; the running maximum is only raised when a block exceeds it by 2^8, so the
//                    accumulator is rescaled (tcgen05.ld / st) rarely
// Operands: q, k token-major [BH, L, D] (K-major for S), v channel-major [BH, D, LkPad] (K-major for P V: exactly the
// [B, C, HW] layout the reference's 1x1 convs produce).  Heads wider than 256 split their OUTPUT channels over
// gridDim.y CTAs, each recomputing the logits; Q stays resident in shared memory when D <= 256.
//
// Warp roles (384 threads): 0 = TMA producer of K (and Q) chunks, 1 = TMA producer of V blocks, 2 = MMA issue,
// 3 = TMEM allocation, 4-7 and 8-11 = two softmax / rescale / output warpgroups (keys 0-63 / 64-127 of every block).  Keys that fit ONE block (Lk <= 128) take the SINGLE variant below.
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "kernels.h"

namespace tmd {

namespace {

constexpr int kStagesMax = 4;
constexpr uint32_t kTileBytes = 128 * 128;      // [128 rows][64 bf16]: one swizzled K-major operand tile

struct AttnParams {
    CUtensorMap tmQ, tmK, tmV;
    void* out;                    // fp32, or bf16 when out_bf16; row (b, query) at b*Lq*out_pitch + query*out_pitch, head h at column h*D
    int v_tok;                    // v is token-major [B, Lk, heads*D] (tmV is a 4-D map like tmQ / tmK): MN-major B operand of P V
    int out_bf16, heads;          // heads: q / k / out hold `heads` heads side by side in a row ([B, L, heads*D]); bh = b*heads + h
    long long out_pitch;          // elements
    int BH, Lq, Lk, D, DV;
    int nkc, nkb;                 // 64-wide chunks of D; 128-key blocks
    int q_resident, NS, NVB;
    int kch;                      // 64-wide K chunks per ring stage (one barrier round trip per stage: fewer, larger stages)
    int ndv;                      // SINGLE: DV-wide output chunks one CTA walks (the other mode spreads them over gridDim.y)
    uint32_t ab_bytes;            // shared memory of the Q / K region (SINGLE: P is written over it once the logits are complete)
    uint32_t tmem_cols;
    long long* trace;             // diagnostics (TMD_ATTN_TRACE=1): cycle counters of CTA 0's first softmax warp / MMA thread
    uint32_t stage_bytes, v_bytes;
    uint32_t idesc_qk, idesc_pv;
    float scale_log2;
};

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
// A wait that lasts seconds means a broken pipeline: trap instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t spins = 0;; ++spins) {
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
            else if (now - t0 > 4000000000LL) __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// 16 accumulator values * inv -> 64 bytes of fp32 or 32 bytes of bf16 at p
__device__ __forceinline__ void store_out16(char* p, const float v[16], float inv, bool bf16) {
    if (bf16) {
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i] * inv, v[2 * i + 1] * inv);
            w[i] = *reinterpret_cast<const uint32_t*>(&h);
        }
        reinterpret_cast<uint4*>(p)[0] = make_uint4(w[0], w[1], w[2], w[3]);
        reinterpret_cast<uint4*>(p)[1] = make_uint4(w[4], w[5], w[6], w[7]);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            reinterpret_cast<float4*>(p)[i] = make_float4(v[4 * i] * inv, v[4 * i + 1] * inv, v[4 * i + 2] * inv, v[4 * i + 3] * inv);
    }
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
// 32 columns of this warp's 32 TMEM lanes; the caller waits (tc_wait_ld) once for a batch of these
__device__ __forceinline__ void tc_ld32_nowait(uint32_t taddr, float* v) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_st16(uint32_t taddr, const float v[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
          "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
          "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
          "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
        : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// K-major, 128-byte-swizzled shared-memory matrix descriptor: start address >> 4 at [0,14), LBO = 1 at [16,30),
// SBO = 1024 B (one 8-row swizzle atom) at [32,46), version 1 at [46,48), SWIZZLE_128B (2) at [61,64).
__device__ __forceinline__ uint64_t sdesc(uint32_t lo) {
    const uint32_t hi = (1024u >> 4) | (1u << 14) | (2u << 29);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint32_t sdesc_lo(uint32_t addr) { return ((addr & 0x3FFFFu) >> 4) | (1u << 16); }

// SINGLE = the keys fit one 128-key block (Lk <= 128: the 77 text tokens of cross attention, the 8x8 level).  Then nothing is
// carried from block to block, which removes every reason for a big footprint: P is written over the Q / K tiles (dead once the
// logits are complete), the output accumulators sit on top of the logits in TMEM (dead once P is written), one CTA walks ALL
// output chunks of a wide head against the same P (no logit recomputation), and two CTAs share an SM so that the loads of one
// overlap the softmax / stores of the other -- these shapes are HBM-bound.
template <bool SINGLE>
__global__ void __launch_bounds__(SINGLE ? 192 : 384, SINGLE ? 4 : 1) attn_umma_kernel(const __grid_constant__ AttnParams P) {
    // SINGLE runs 192 threads (one producer warp for all three operands, MMA issue and TMEM allocation in one warp, softmax warps
    // 2-5 = TMEM lane quarters 2, 3, 0, 1) at <= 85 registers, so that four CTAs fit an SM
    constexpr int W_K = 0, W_V = SINGLE ? 0 : 1, W_MMA = SINGLE ? 1 : 2, W_ALLOC = SINGLE ? 1 : 3, W_SOFTMAX = SINGLE ? 2 : 4;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kStagesMax + 13];
    __shared__ uint32_t tmem_base_slot;
    __shared__ float xch[2][SINGLE ? 1 : 128];      // long sequences: row statistics exchanged between the two softmax warpgroups

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // swizzle atoms need 1024-byte alignment
    const uint32_t q_base = smem_base;
    const uint32_t st_base = q_base + (P.q_resident ? (uint32_t)P.nkc * kTileBytes : 0u);
    const uint32_t p_base = SINGLE ? q_base : q_base + P.ab_bytes;
    const uint32_t v_base = q_base + P.ab_bytes + (SINGLE ? 0u : 2 * kTileBytes);
    const uint32_t bar0 = smem_u32(bars);
    auto k_full = [&](int s) { return bar0 + 8u * s; };
    auto k_empty = [&](int s) { return bar0 + 8u * (kStagesMax + s); };
    const uint32_t q_full = bar0 + 8u * (2 * kStagesMax);
    auto v_full = [&](int s) { return bar0 + 8u * (2 * kStagesMax + 1 + s); };
    auto v_empty = [&](int s) { return bar0 + 8u * (2 * kStagesMax + 3 + s); };
    auto s_full = [&](int s) { return bar0 + 8u * (2 * kStagesMax + 5 + s); };
    const uint32_t p_full = bar0 + 8u * (2 * kStagesMax + 7);
    const uint32_t pv_done = bar0 + 8u * (2 * kStagesMax + 8);
    auto o_full = [&](int s) { return bar0 + 8u * (2 * kStagesMax + 9 + s); };
    auto o_empty = [&](int s) { return bar0 + 8u * (2 * kStagesMax + 11 + s); };

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStagesMax; ++s) { mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); }
        mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1); mbar_init(s_full(s), 1);
            mbar_init(o_full(s), 1); mbar_init(o_empty(s), 128);
        }
        mbar_init(p_full, SINGLE ? 128 : 256);
        mbar_init(pv_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == W_ALLOC) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "r"(P.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_slot;
    const uint32_t tmem_o = tmem_base + (SINGLE ? 0u : 256u);

    const int q0 = blockIdx.x * 128, dv0 = SINGLE ? 0 : blockIdx.y * P.DV, bh = blockIdx.z;
    const int hh = bh % P.heads, bb = bh / P.heads;      // head and batch of this (batch, head) index
    const int nkc = P.nkc, nkb = P.nkb;                  // SINGLE: one block, or two for a wide head (Lk <= 256)
    const int nvload = SINGLE ? P.ndv : nkb;             // V tiles this CTA consumes

    if (warp == W_K) {
        if (lane == 0) {
            // ================= K (and Q) chunks =================
            if (P.q_resident) {
                mbar_expect_tx(q_full, (uint32_t)nkc * kTileBytes);
                for (int kc = 0; kc < nkc; ++kc) tma_load_4d(q_base + (uint32_t)kc * kTileBytes, &P.tmQ, q_full, kc * 64, hh, q0, bb);
            }
            uint32_t s = 0, ph = 1;
            const int kch = P.kch;
            // a stage holds kch K chunks; when Q is streamed, kch (K chunk, Q chunk) pairs
            const uint32_t pitch = P.q_resident ? kTileBytes : 2 * kTileBytes;
            for (int j = 0; j < nkb; ++j)
                for (int kc = 0; kc < nkc; kc += kch) {
                    mbar_wait(k_empty(s), ph);
                    mbar_expect_tx(k_full(s), P.stage_bytes);
                    const uint32_t dst = st_base + s * P.stage_bytes;
                    for (int c = 0; c < kch; ++c) {
                        tma_load_4d(dst + (uint32_t)c * pitch, &P.tmK, k_full(s), (kc + c) * 64, hh, j * 128, bb);
                        if (!P.q_resident) tma_load_4d(dst + (uint32_t)c * pitch + kTileBytes, &P.tmQ, k_full(s), (kc + c) * 64, hh, q0, bb);
                    }
                    if (++s == (uint32_t)P.NS) { s = 0; ph ^= 1; }
                }
        }
    }
    if (warp == W_V) {
        if (lane == 0) {
            // ================= V tiles: [DV channels][128 keys] as two 64-key swizzle columns =================
            const uint32_t half = (uint32_t)P.DV * 128u;
            for (int it = 0; it < nvload; ++it) {
                const int vb = it % P.NVB, use = it / P.NVB;
                const int key0 = SINGLE ? 0 : it * 128, dv = SINGLE ? it * P.DV : dv0;
                mbar_wait(v_empty(vb), (use & 1) ^ 1);
                if (P.v_tok) {
                    // token-major v: [128 keys][64 channels] boxes (rows = keys, 128 bytes each), DV / 64 of them per key block; keys
                    // past the end are zero-filled by TMA and multiplied by P = 0
                    mbar_expect_tx(v_full(vb), P.v_bytes);
                    const uint32_t dstv = v_base + (uint32_t)vb * P.v_bytes;
                    const int nblk = SINGLE ? nkb : 1;
                    for (int kb = 0; kb < nblk; ++kb)
                        for (int t = 0; t < P.DV / 64; ++t)
                            tma_load_4d(dstv + (uint32_t)kb * ((uint32_t)P.DV * 256u) + (uint32_t)t * kTileBytes, &P.tmV, v_full(vb), dv + 64 * t, hh,
                                        key0 + 128 * kb, bb);
                    continue;
                }
                // 64-key columns of this tile that hold keys: one past the end is neither loaded nor multiplied
                int ncol = (P.Lk - key0 + 63) >> 6;
                const int maxcol = SINGLE ? 2 * nkb : 2;
                if (ncol > maxcol) ncol = maxcol;
                mbar_expect_tx(v_full(vb), (uint32_t)ncol * half);
                const uint32_t dst = v_base + (uint32_t)vb * P.v_bytes;
                for (int t = 0; t < ncol; ++t) tma_load_3d(dst + (uint32_t)t * half, &P.tmV, v_full(vb), key0 + 64 * t, dv, bh);
            }
        }
    }
    if (warp == W_MMA) {
        if (lane == 0) {
            // ================= MMA issue =================
            const uint32_t q_lo0 = sdesc_lo(q_base), st_lo0 = sdesc_lo(st_base), p_lo0 = sdesc_lo(p_base), v_lo0 = sdesc_lo(v_base);
            const uint32_t tile16 = kTileBytes >> 4, stage16 = P.stage_bytes >> 4, half16 = ((uint32_t)P.DV * 128u) >> 4;
            // token-major v (MN-major operand): LBO = distance of the 64-channel boxes (16 KB), a 16-key step is 16 rows = 2 KB
            const uint32_t v_lo0_mn = ((v_base & 0x3FFFFu) >> 4) | ((kTileBytes >> 4) << 16), blk16 = ((uint32_t)P.DV * 256u) >> 4;
            // (one formula for both layouts, its constants chosen once: this thread's instruction stream is the critical path)
            const uint32_t vd_base = P.v_tok ? v_lo0_mn : v_lo0, vd_shift = P.v_tok ? 3u : 2u, vd_mask = P.v_tok ? 7u : 3u;
            const uint32_t vd_blk = P.v_tok ? blk16 : half16, vd_step = P.v_tok ? 128u : 2u, vd_buf = P.v_bytes >> 4;
            auto v_desc = [&](uint32_t vb, uint32_t k) -> uint64_t {
                return sdesc(vd_base + vb * vd_buf + (k >> vd_shift) * vd_blk + (k & vd_mask) * vd_step);
            };
            uint32_t ks = 0, kph = 0;
            if (P.q_resident) mbar_wait(q_full, 0);
            auto issue_qk = [&](int j) {
                const uint32_t d = tmem_base + (uint32_t)(j & 1) * 128u;
                // one wait / commit round trip per STAGE: at one 64-wide chunk (4 MMAs = 256 tensor cycles) per stage the round trip
                // (~250 cycles) kept the pipe half idle -- a fixed ~1700 cycles per key block on top of the MMA time
                for (int kc = 0; kc < nkc; kc += P.kch) {
                    mbar_wait(k_full(ks), kph);
                    tc_fence_after();
                    for (int c = 0; c < P.kch; ++c) {
                        const uint32_t b_lo = st_lo0 + ks * stage16 + (uint32_t)c * (P.q_resident ? tile16 : 2 * tile16);
                        const uint32_t a_lo = P.q_resident ? q_lo0 + (uint32_t)(kc + c) * tile16 : b_lo + tile16;
#pragma unroll
                        for (int k = 0; k < 4; ++k) tc_mma_bf16(d, sdesc(a_lo + k * 2), sdesc(b_lo + k * 2), P.idesc_qk, (uint32_t)(kc | c | k));
                    }
                    tc_commit(k_empty(ks));
                    if (++ks == (uint32_t)P.NS) { ks = 0; kph ^= 1; }
                }
                tc_commit(s_full(j & 1));
            };
            issue_qk(0);
            if (SINGLE) {
                if (nkb > 1) issue_qk(1);
                mbar_wait(p_full, 0);
                const int ksteps = (P.Lk + 15) >> 4;
                for (int c = 0; c < P.ndv; ++c) {
                    const int vb = c % P.NVB, use = c / P.NVB, ob = c & 1;
                    mbar_wait(v_full(vb), use & 1);
                    if (c >= 2) mbar_wait(o_empty(ob), ((c >> 1) - 1) & 1);
                    tc_fence_after();
                    for (int k = 0; k < ksteps; ++k) {
                        const uint32_t blk = k >> 2, kk = k & 3;
                        tc_mma_bf16(tmem_o + (uint32_t)ob * P.DV, sdesc(p_lo0 + blk * tile16 + kk * 2), v_desc((uint32_t)vb, (uint32_t)k), P.idesc_pv,
                                    (uint32_t)k);
                    }
                    tc_commit(v_empty(vb));
                    tc_commit(o_full(ob));
                }
            } else {
                for (int j = 0; j < nkb; ++j) {
                    // the logits of the next block go out first: their buffer was last read by softmax(j-1), which had signalled
                    // p_full(j-1) before PV(j-1) was issued
                    if (j + 1 < nkb) issue_qk(j + 1);
                    const int vb = j % P.NVB, use = j / P.NVB;
                    mbar_wait(v_full(vb), use & 1);
                    mbar_wait(p_full, j & 1);
                    tc_fence_after();
                    const int keys = P.Lk - j * 128;
                    const int ksteps = keys >= 128 ? 8 : (keys + 15) >> 4;
                    for (int k = 0; k < ksteps; ++k) {
                        const uint32_t blk = k >> 2, kk = k & 3;
                        tc_mma_bf16(tmem_o, sdesc(p_lo0 + blk * tile16 + kk * 2), v_desc((uint32_t)vb, (uint32_t)k), P.idesc_pv, (uint32_t)(j | k));
                    }
                    tc_commit(v_empty(vb));
                    tc_commit(pv_done);
                }
            }
        }
    }
    if (warp >= W_SOFTMAX) {
        // ================= softmax, accumulator rescale, output =================
        const int qtr = warp & 3;
        const int r = qtr * 32 + lane;                                   // query row of this thread = TMEM lane
        const uint32_t lane_addr = (uint32_t)(qtr * 32) << 16;
        const float sl2 = P.scale_log2;
        const uint32_t p_row = p_base + (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);
        const bool valid = q0 + r < P.Lq;
        const uint32_t esz = P.out_bf16 ? 2u : 4u;
        char* orow = reinterpret_cast<char*>(P.out) +
                     (((size_t)bb * P.Lq + (size_t)(q0 + r)) * (size_t)P.out_pitch + (size_t)hh * P.D + dv0) * esz;
        if (SINGLE) {
            // the logits of all keys are at hand: exact softmax in two passes over 64-column pieces (few registers: four CTAs per SM)
            mbar_wait(s_full(0), 0);
            if (nkb > 1) mbar_wait(s_full(1), 0);
            tc_fence_after();
            const uint32_t s_addr = tmem_base + lane_addr;
            const int nvalid = P.Lk, nhalf = (nvalid + 63) >> 6;
            float mb = -INFINITY;
            for (int h = 0; h < nhalf; ++h) {
                float sv[64];
                tc_ld32_nowait(s_addr + h * 64, sv);
                tc_ld32_nowait(s_addr + h * 64 + 32, sv + 32);
                tc_wait_ld();
                float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                for (int i = 0; i < 64; ++i)
                    if (h * 64 + i < nvalid) mx[i & 3] = fmaxf(mx[i & 3], sv[i]);
                mb = fmaxf(mb, fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])));
            }
            const float m = mb * sl2;                                        // scale > 0
            float l = 0.f;
            for (int h = 0; h < nhalf; ++h) {
                float sv[64];
                tc_ld32_nowait(s_addr + h * 64, sv);
                tc_ld32_nowait(s_addr + h * 64 + 32, sv + 32);
                tc_wait_ld();
                // every thread's logit reads precede ITS writes below, but P overwrites the Q / K tiles, which only the completed
                // logit MMAs (s_full) read: no further ordering is needed
                float ls[2] = {0.f, 0.f};
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    uint32_t w[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int col = c * 8 + 2 * i;
                        float p0 = ex2(fmaf(sv[col], sl2, -m)), p1 = ex2(fmaf(sv[col + 1], sl2, -m));
                        if (h * 64 + col >= nvalid) p0 = 0.f;
                        if (h * 64 + col + 1 >= nvalid) p1 = 0.f;
                        w[i] = pack_bf16x2(p0, p1);
                        ls[0] += __uint_as_float(w[i] << 16);
                        ls[1] += __uint_as_float(w[i] & 0xffff0000u);
                    }
                    const uint32_t addr = p_row + (uint32_t)h * kTileBytes + ((((uint32_t)c) ^ sw) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
                }
                l += ls[0] + ls[1];
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_fence_before();
            mbar_arrive(p_full);
            const float inv = 1.f / l;
            if (P.ndv == 1) {
                // the whole head in one chunk: a thread's output row is one contiguous D*4-byte piece of global memory.  Written
                // straight from registers it costs one L1 wavefront per lane and instruction (rows are D*4 bytes apart), which is
                // what bounded the 77-key cross-attention shape; instead the row is staged in shared memory (over the dead Q / K /
                // P / V tiles, pitch D*4 + 16 bytes: conflict-free 16-byte stores) and leaves as ONE bulk copy per row.
                mbar_wait(o_full(0), 0);
                tc_fence_after();
                const uint32_t srow = q_base + (uint32_t)r * ((uint32_t)P.DV * esz + 16u);
                for (int cc = 0; cc < P.DV / 16; ++cc) {
                    float v[16];
                    tc_ld16(tmem_o + lane_addr + cc * 16, v);
                    if (P.out_bf16) {
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            uint32_t w[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) w[j] = pack_bf16x2(v[8 * i + 2 * j] * inv, v[8 * i + 2 * j + 1] * inv);
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(srow + (uint32_t)cc * 32u + (uint32_t)i * 16u), "r"(w[0]), "r"(w[1]),
                                         "r"(w[2]), "r"(w[3])
                                         : "memory");
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(srow + (uint32_t)cc * 64u + (uint32_t)i * 16u), "f"(v[4 * i] * inv),
                                         "f"(v[4 * i + 1] * inv), "f"(v[4 * i + 2] * inv), "f"(v[4 * i + 3] * inv)
                                         : "memory");
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                if (valid)
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(orow), "r"(srow), "r"((uint32_t)P.DV * esz) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            } else
            for (int c = 0; c < P.ndv; ++c) {
                const int ob = c & 1;
                mbar_wait(o_full(ob), (c >> 1) & 1);
                tc_fence_after();
                for (int cc = 0; cc < P.DV / 16; ++cc) {
                    float v[16];
                    tc_ld16(tmem_o + lane_addr + (uint32_t)ob * P.DV + cc * 16, v);
                    if (valid) store_out16(orow + (size_t)(c * P.DV + cc * 16) * esz, v, inv, P.out_bf16 != 0);
                }
                tc_fence_before();
                mbar_arrive(o_empty(ob));
            }
        } else {
        // Two softmax warpgroups share every key block: warps 4-7 take keys 0-63 of the block, warps 8-11 keys 64-127 of the SAME
        // query rows (a TMEM lane quarter is open to every warp with that index mod 4).  The block's exp2 / pack work -- which
        // bounded head widths <= 128 with one warpgroup (TMD_ATTN_TRACE: 1610 of 2700 cycles per block) -- halves per thread; the
        // partner threads of a row exchange their partial row maxima through shared memory (one named barrier per block), keep
        // partial row sums that are only added at the end, write one 64-key P tile each and rescale / store one half of the
        // output channels each.
        const int wg = (warp - W_SOFTMAX) >> 2;
        float m_run = -INFINITY, l_run = 0.f;
        const bool tr = P.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == W_SOFTMAX;
        long long t_s = 0, t_ld = 0, t_math = 0, t_pv = 0, t_p = 0, t0 = clock64(), tq = 0;
        const int hdv = P.DV >> 1;                                           // output channels this warpgroup rescales and stores
        for (int j = 0; j < nkb; ++j) {
            if (tr) tq = clock64();
            mbar_wait(s_full(j & 1), (j >> 1) & 1);
            tc_fence_after();
            if (tr) { const long long n = clock64(); t_s += n - tq; tq = n; }
            const uint32_t s_addr = tmem_base + lane_addr + (uint32_t)(j & 1) * 128u + (uint32_t)wg * 64u;
            const int nvalid = P.Lk - j * 128 - wg * 64;                     // valid keys of this half (may be <= 0 in the last block)
            float sv[64];
            tc_ld32_nowait(s_addr, sv);
            tc_ld32_nowait(s_addr + 32, sv + 32);
            tc_wait_ld();
            if (tr) { const long long n = clock64(); t_ld += n - tq; tq = n; }
            if (nvalid < 64) {
#pragma unroll
                for (int i = 0; i < 64; ++i)
                    if (i >= nvalid) sv[i] = -INFINITY;
            }
            float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
                mx[0] = fmaxf(mx[0], sv[i]); mx[1] = fmaxf(mx[1], sv[i + 1]); mx[2] = fmaxf(mx[2], sv[i + 2]); mx[3] = fmaxf(mx[3], sv[i + 3]);
            }
            const float mpart = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
            // (one 1 KB exchange buffer, hence the second barrier: shared memory is full at D = 256)
            xch[wg][r] = mpart;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            const float mb = fmaxf(mpart, xch[wg ^ 1][r]) * sl2;             // scale > 0; key 0 of the block is valid, so mb is finite
            asm volatile("bar.sync 1, 256;" ::: "memory");
            // raise the reference maximum only when some row of the warp outgrew it by 2^8 (P stays below 256 otherwise); the
            // partner warp sees the same maxima and the same history, i.e. takes the same decision
            const bool raise = __any_sync(0xffffffffu, mb > m_run + 8.f);
            float alpha = 1.f;
            if (raise) {
                const float m_new = fmaxf(m_run, mb);
                alpha = ex2(m_run - m_new);                                  // 0 on the first block
                l_run *= alpha;
                m_run = m_new;
            }
            // P = exp2(S * scale * log2(e) - m), rounded to bf16; masked keys (-inf) give exactly 0.  The row sum is taken over
            // the ROUNDED probabilities the tensor core multiplies, so that sum_j P_ij / l_i = 1 holds for what is actually
            // accumulated (constant values come back unchanged).
            uint32_t pk[32];
            float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float p0 = ex2(fmaf(sv[2 * i], sl2, -m_run)), p1 = ex2(fmaf(sv[2 * i + 1], sl2, -m_run));
                const uint32_t w = pack_bf16x2(p0, p1);
                pk[i] = w;
                ls[i & 1] += __uint_as_float(w << 16);
                ls[2 + (i & 1)] += __uint_as_float(w & 0xffff0000u);
            }
            l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
            if (tr) { const long long n = clock64(); t_math += n - tq; tq = n; }
            if (j > 0) {
                mbar_wait(pv_done, (j - 1) & 1);                             // P is single-buffered; the accumulator is at rest
                tc_fence_after();
                if (raise) {
                    for (int c = 0; c < hdv / 16; ++c) {
                        float v[16];
                        const uint32_t a = tmem_o + lane_addr + (uint32_t)(wg * hdv + c * 16);
                        tc_ld16(a, v);
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] *= alpha;
                        tc_st16(a, v);
                    }
                    tc_wait_st();
                }
            }
            if (tr) { const long long n = clock64(); t_pv += n - tq; tq = n; }
            // this warpgroup's 64-key P tile: 16-byte chunk c of row r sits at chunk (c ^ (r & 7)) of its 128-byte row (SWIZZLE_128B)
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t addr = p_row + (uint32_t)wg * kTileBytes + ((((uint32_t)c) ^ sw) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[4 * c]), "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]),
                             "r"(pk[4 * c + 3])
                             : "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> the tensor core's async-proxy reads
            tc_fence_before();
            mbar_arrive(p_full);
            if (tr) t_p += clock64() - tq;
        }
        if (tr && lane == 0) {
            P.trace[0] = clock64() - t0; P.trace[1] = t_s; P.trace[2] = t_ld; P.trace[3] = t_math; P.trace[4] = t_pv; P.trace[5] = t_p;
            P.trace[6] = nkb;
        }
        // the row sum is the sum of the two warpgroups' partial sums (same reference maximum on both sides)
        xch[wg][r] = l_run;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float inv = 1.f / (l_run + xch[wg ^ 1][r]);
        mbar_wait(pv_done, (nkb - 1) & 1);
        tc_fence_after();
        for (int c = 0; c < hdv / 16; ++c) {
            float v[16];
            tc_ld16(tmem_o + lane_addr + (uint32_t)(wg * hdv + c * 16), v);
            if (valid) store_out16(orow + (size_t)(wg * hdv + c * 16) * esz, v, inv, P.out_bf16 != 0);
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == W_ALLOC) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(P.tmem_cols) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn attn_encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            return nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// bf16 [batch][rows][inner] with a row pitch of `pitch` elements, boxes of [box_rows][64]
bool make_map(CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint64_t batch, uint64_t pitch, uint32_t box_rows) {
    EncodeTiledFn enc = attn_encode_fn();
    if (!enc) return false;
    const cuuint64_t dims[3] = {inner, rows, batch};
    const cuuint64_t strides[2] = {pitch * 2, pitch * rows * 2};
    const cuuint32_t box[3] = {64, box_rows, 1}, es[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// bf16 [batch][rows][heads][inner]: `heads` heads of `inner` channels side by side in a row of `pitch` elements; boxes of
// [128 rows][one head][64] (coordinates: channel, head, row, batch -- strides in ascending order)
bool make_map4(CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint64_t heads, uint64_t batch, uint64_t pitch) {
    EncodeTiledFn enc = attn_encode_fn();
    if (!enc) return false;
    const cuuint64_t dims[4] = {inner, heads, rows, batch};
    const cuuint64_t strides[3] = {inner * 2, pitch * 2, pitch * rows * 2};
    const cuuint32_t box[4] = {64, 1, 128, 1}, es[4] = {1, 1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

bool attention_bf16_supported(int BH, int Lq, int Lk, int D, int lk_pitch) {
    return BH > 0 && BH <= 65535 && Lq > 0 && Lk > 0 && D >= 64 && D % 64 == 0 && D <= 4096 && lk_pitch >= Lk && lk_pitch % 8 == 0;
}

cudaError_t attention_bf16(const void* q, const void* k, const void* vt, float* out, int BH, int Lq, int Lk, int D, int lk_pitch,
                           float scale, cudaStream_t stream) {
    return attention_bf16_strided(q, k, vt, out, BH, 1, Lq, Lk, D, D, D, lk_pitch, D, 0, scale, stream);
}

// q [B, Lq, q_pitch] and k [B, Lk, k_pitch] with head h in columns [h*D, (h+1)*D); vt [B*heads, D, lk_pitch]; out [B, Lq, out_pitch]
// (fp32, or bf16 when out_bf16) with head h in the same columns: the projections' own token-major layout on both sides.
cudaError_t attention_bf16_strided(const void* q, const void* k, const void* vt, void* out, int B, int heads, int Lq, int Lk, int D,
                                   long long q_pitch, long long k_pitch, int lk_pitch, long long out_pitch, int out_bf16, float scale,
                                   cudaStream_t stream) {
    return attention_bf16_general(q, k, vt, out, B, heads, Lq, Lk, D, q_pitch, k_pitch, 0, lk_pitch, out_pitch, out_bf16, scale, stream);
}

// v_tok_pitch > 0: v is token-major [B, Lk, v_tok_pitch] like k (head h in columns [h*D, (h+1)*D)) and lk_pitch is ignored;
// v_tok_pitch == 0: v is channel-major [B*heads, D, lk_pitch].
cudaError_t attention_bf16_general(const void* q, const void* k, const void* vt, void* out, int B, int heads, int Lq, int Lk, int D,
                                   long long q_pitch, long long k_pitch, long long v_tok_pitch, int lk_pitch, long long out_pitch, int out_bf16,
                                   float scale, cudaStream_t stream) {
    if (B <= 0 || heads <= 0 || (long long)B * heads > 65535) return cudaErrorInvalidValue;
    const int BH = B * heads;
    if (v_tok_pitch) {
        if (v_tok_pitch < (long long)heads * D || v_tok_pitch % 8) return cudaErrorInvalidValue;
        lk_pitch = (Lk + 7) / 8 * 8;          // only for the shape check below
    }
    if (!attention_bf16_supported(BH, Lq, Lk, D, lk_pitch)) return cudaErrorInvalidValue;
    if (q_pitch < (long long)heads * D || k_pitch < (long long)heads * D || out_pitch < (long long)heads * D || q_pitch % 8 || k_pitch % 8 ||
        out_pitch % 8)
        return cudaErrorInvalidValue;
    if ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(vt) | reinterpret_cast<uintptr_t>(out)) & 15)
        return cudaErrorMisalignedAddress;
    // cudaFuncSetAttribute is per device: one record per device of the process
    static int smem_limit_dev[64][2] = {};
    int dev = 0;
    {
        const cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    }
    int* smem_limit = smem_limit_dev[dev];
    if (!smem_limit[0]) {
        const void* fns[2] = {(const void*)attn_umma_kernel<false>, (const void*)attn_umma_kernel<true>};
        for (int i = 0; i < 2; ++i) {
            cudaFuncAttributes fa;
            cudaError_t e = cudaFuncGetAttributes(&fa, fns[i]);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(fns[i], cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - (int)fa.sharedSizeBytes);
            if (e != cudaSuccess) return e;
            smem_limit[i] = 227 * 1024 - (int)fa.sharedSizeBytes;
        }
    }
    const bool single = Lk <= 128 || (Lk <= 256 && D > 256);     // all logits of a query tile fit TMEM next to nothing else
    AttnParams P;
    memset(&P, 0, sizeof(P));
    P.out = out;
    P.out_bf16 = out_bf16 ? 1 : 0; P.heads = heads; P.out_pitch = out_pitch;
    P.v_tok = v_tok_pitch ? 1 : 0;
    P.BH = BH; P.Lq = Lq; P.Lk = Lk; P.D = D;
    P.nkc = D / 64;
    P.nkb = (Lk + 127) / 128;
    P.q_resident = D <= 256;
    P.kch = 1;
    P.stage_bytes = P.q_resident ? kTileBytes : 2 * kTileBytes;
    size_t smem;
    if (single) {
        // one key block: every output chunk against the same P, accumulators on top of the logits, P on top of Q / K
        P.DV = D <= 256 ? D : ((D % 128 == 0 && P.nkb == 1) ? 128 : 64);
        P.ndv = D / P.DV;
        P.NVB = 1;
        P.NS = P.q_resident ? P.nkc : 2;
        int cols = (P.ndv > 1 ? 2 : 1) * P.DV;
        if (cols < 128 * P.nkb) cols = 128 * P.nkb;
        P.tmem_cols = cols <= 128 ? 128 : 256;
        P.v_bytes = (uint32_t)P.DV * 256u * (uint32_t)P.nkb;
        P.ab_bytes = (P.q_resident ? (uint32_t)P.nkc * kTileBytes : 0u) + (uint32_t)P.NS * P.stage_bytes;
        if (P.ab_bytes < 2 * kTileBytes * (uint32_t)P.nkb) P.ab_bytes = 2 * kTileBytes * (uint32_t)P.nkb;
        smem = 1024 + (size_t)P.ab_bytes + (size_t)P.NVB * P.v_bytes;
        // co-resident CTAs must find their TMEM columns: at most 512 / tmem_cols of them fit by shared memory
        const size_t floor_bytes = P.tmem_cols == 128 ? 46 * 1024 : 80 * 1024;
        if (smem < floor_bytes) smem = floor_bytes;
    } else {
        P.DV = D <= 256 ? D : (D % 256 == 0 ? 256 : (D % 128 == 0 ? 128 : 64));
        P.ndv = 1;
        if (P.q_resident) {
            P.kch = P.nkc <= 3 ? P.nkc : (P.nkc % 2 == 0 ? 2 : 1);
            P.stage_bytes = (uint32_t)P.kch * kTileBytes;
            P.NS = 4 / P.kch > 2 ? 4 / P.kch : 2;
        } else if (P.nkc % 2 == 0) {
            P.kch = 2;
            P.stage_bytes = 4 * kTileBytes;
            P.NS = 2;
        } else {
            P.NS = 3;
        }
        P.v_bytes = (uint32_t)P.DV * 256u;
        P.NVB = P.DV <= 128 ? 2 : 1;
        P.tmem_cols = 512;
        auto need = [&]() {
            P.ab_bytes = (P.q_resident ? (uint32_t)P.nkc * kTileBytes : 0u) + (uint32_t)P.NS * P.stage_bytes;
            return 1024 + (size_t)P.ab_bytes + 2 * kTileBytes + (size_t)P.NVB * P.v_bytes;
        };
        while (P.NS > 2 && need() > (size_t)smem_limit[0]) --P.NS;
        smem = need();
        // every CTA allocates all 512 TMEM columns: keep it at one CTA per SM
        if (smem < 120 * 1024) smem = 120 * 1024;
    }
    if (smem > (size_t)smem_limit[single ? 1 : 0]) return cudaErrorInvalidConfiguration;
    // D = f32, A = B = bf16, both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
    P.idesc_qk = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
    P.idesc_pv = (1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)P.DV >> 3) << 17) | ((128u >> 4) << 24);
    if (P.v_tok) P.idesc_pv |= 1u << 16;          // B operand MN-major (channels contiguous)
    P.scale_log2 = scale * 1.4426950408889634f;
    if (!make_map4(&P.tmQ, q, D, Lq, heads, B, q_pitch) || !make_map4(&P.tmK, k, D, Lk, heads, B, k_pitch) ||
        !(P.v_tok ? make_map4(&P.tmV, vt, D, Lk, heads, B, v_tok_pitch) : make_map(&P.tmV, vt, Lk, D, BH, lk_pitch, (uint32_t)P.DV)))
        return cudaErrorInvalidValue;
    if (single) {
        const dim3 grid((Lq + 127) / 128, 1, BH);
        attn_umma_kernel<true><<<grid, 192, smem, stream>>>(P);
    } else {
        const dim3 grid((Lq + 127) / 128, D / P.DV, BH);
        const char* te = getenv("TMD_ATTN_TRACE");
        if (te && te[0] == '1') {
            cudaMalloc((void**)&P.trace, 8 * sizeof(long long));
            cudaMemset(P.trace, 0, 8 * sizeof(long long));
            attn_umma_kernel<false><<<grid, 384, smem, stream>>>(P);
            cudaDeviceSynchronize();
            long long t[8];
            cudaMemcpy(t, P.trace, sizeof(t), cudaMemcpyDeviceToHost);
            cudaFree(P.trace);
            const double n = t[6] ? (double)t[6] : 1.0;
            fprintf(stderr, "[attn trace] Lq=%d Lk=%d D=%d | softmax warp per key block: total %.0f cyc = wait S %.0f + tmem ld %.0f + max/exp/pack %.0f + "
                    "wait PV %.0f + P store/fence/arrive %.0f\n", Lq, Lk, D, t[0] / n, t[1] / n, t[2] / n, t[3] / n, t[4] / n, t[5] / n);
            return cudaGetLastError();
        }
        attn_umma_kernel<false><<<grid, 384, smem, stream>>>(P);
    }
    return cudaGetLastError();
}

}  // namespace tmd
