// K4/K6: implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 + TMEM), fed by TMA.
//
// Replaces the Conv2d / ConvTranspose2d library calls (cuDNN under torch) that the reference reaches through
// nnUNetPredictor (reference ts2d/core/inference/nnu.py:149-166, prediction_worker.py:209).
//
// GEMM view:  D[M = 128 output pixels][N = BN output channels] += A[M][K] * B[N][K],  K = taps * Cin.
//   A  NHWC activations. One K-chunk = (filter tap, KC input channels). The 128 pixels of a tile are a
//      (tn images) x (th rows) x (tw cols) box, so a tap is the same box shifted by (r-pad, s-pad): ONE 4-D TMA
//      box load per K-chunk; TMA zero-fills out-of-bounds coordinates, which *is* the conv zero padding.
//      Strided convs use a 5-D "parity" view of the same buffer (w = 2*w2 + px folded into the channel axis,
//      h = 2*h2 + py as its own axis; each axis independently, so (2,1) / (1,2) strides work too), so every tap is again a dense box.  The box lands in shared memory as
//      128 rows x KC channels with the 128-byte (KC=64) or 64-byte (KC=32) swizzle = the canonical K-major
//      UMMA operand layout.
//   B  weights repacked [tap][Cout][Cin] (K-major), 3-D TMA box (KC, BN, 1).
//   D  fp32 accumulators in TMEM, double buffered (2 x BN columns): the epilogue of tile i overlaps the MMAs
//      of tile i+1.
// Transposed conv (kernel == stride): a 1x1 GEMM whose N axis is (tap, Cout): all taps of a pixel come out of ONE pass
// over A; the epilogue scatters column block (tap, c0) to output pixel (y*my + tap/mx, x*mx + tap%mx).
// Warp roles (384 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM allocator + MMA issuer (whole warp
// convergent, one elected lane issues), warps 2-3 idle, warps 4-11 = epilogue, two warpgroups (TMEM lane quarter =
// warp % 4): tcgen05.ld -> +bias -> round to bf16/fp16 -> InstanceNorm sum / sum-of-squares of the ROUNDED values ->
// 32-byte global stores.  BN == 32: the groups alternate tiles; BN >= 64: each group takes half of the tile's columns.
// Statistics: per-thread running sums (BN <= 64) or per-tile warp transpose-reductions into per-lane sums (BN >= 128),
// flushed with double atomics when the (image, channel block) of the CTA's tile sequence changes.
// Persistent grid: one CTA per SM, static CONTIGUOUS partition of the tile sequence (channel block slowest, then image,
// y, x): a CTA touches few images (few flushes, spread in time) and neighbouring CTAs share weights through L2.
#pragma once
#include "common.cuh"

namespace ts2d {

struct ConvTcParams {
    int N, Hout, Wout, Cout;     // GEMM-side output grid (for transposed conv: the INPUT grid)
    int cin_chunks;              // Cin / KC
    int split_chunks;            // K-chunks [0, split_chunks) of a tap come from tmA, the rest from tmA2 (decoder convs: the up-sampled tensor and the
                                 // encoder skip are two dense tensors instead of one concatenated buffer); == cin_chunks when there is one input
    int taps, kw;                // taps = kh*kw (1 for 1x1 / transposed conv slices)
    int pad_h, pad_w;
    int stride2;                 // 0: stride 1 (4-D map); else the 5-D parity map, bit 0 = stride 2 along w, bit 1 = stride 2 along h
    int in_ps;                   // input pixel stride in elements (parity view: px * in_ps selects the odd column)
    int tw, th, tn;              // tile box: tw*th*tn == 128
    int tiles_w, tiles_h, tiles_n;
    int n_tiles_n;               // Cout / BN
    int tap_cout;                // transposed conv: output channels per tap (Cout = taps_out * tap_cout); else == Cout
    // output addressing: pixel (n, y*my + dy, x*mx + dx) of an image Hfull x Wfull with pixel stride out_ps elements
    int my, mx, Hfull, Wfull;
    long long out_ps;
    void *out;
    const float *bias;           // [Cout] or null
    double *stats;               // [N][Cout][2] or null
    // split-K (tiny levels: a handful of M tiles, K = 9*512 .. 9*1024): tile index = (output tile, K slice), slice ks covers
    // k-chunks [ks*k_chunks/ksplit, (ks+1)*k_chunks/ksplit); the epilogue stores the fp32 accumulators to
    // partial[ks][(n*Hout + y)*Wout + x][Cout] (no bias / rounding / statistics: splitk_finish_kernel does those in a fixed order)
    int ksplit;                  // 0 / 1: off
    float *partial;
    long long *dbg;              // TS2D_TC_DBG: per-CTA clock stamps (8 per CTA), see run_conv
    // mc = 1: launched as clusters of 2 CTAs that walk neighbouring M tiles of the same channel block in lock step; each CTA loads its own
    // activation tile and HALF of the weight tile, multicast into both CTAs' shared memory.  The TMA unit issues one request per 64/128-byte
    // box row (~1.7 cycles each, measured): 128 + BN/2 rows per k-chunk instead of 128 + BN.  (tmB box = BN/2 rows.)
    int mc;
};
constexpr int kTcThreads = 384;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// `make SPIN_LIMIT=n` (-DTS2D_MBAR_SPIN_LIMIT=n): a wait that fails n try_wait rounds prints the barrier and traps, so a wrong
// expect-tx byte count or a lost arrival is an error code (cudaErrorLaunchFailure) instead of a hung GPU.  Off in the product build.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
#ifdef TS2D_MBAR_SPIN_LIMIT
    for (long long spin = 0;; ++spin) {
        uint32_t ok;
        asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (ok) return;
        if (spin > (long long)(TS2D_MBAR_SPIN_LIMIT)) {
            printf("ts2d: mbarrier wait timed out: block %d thread %d barrier smem+%u parity %u\n", (int)blockIdx.x, (int)threadIdx.x, smem_u32(bar), parity);
            __trap();
        }
    }
#endif
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
#ifndef TS2D_PARK_NS
#define TS2D_PARK_NS 64
#endif
// Same wait for roles that are NOT latency critical (epilogue warps waiting for an accumulator, the TMA producer waiting for a free
// slot): try_wait with a suspend-time hint parks the warp in hardware instead of spinning.  A spinning warp takes issue slots from the
// warps that share its scheduler - measured on conv_halo_kernel<FUSE>: the transform warps (2, 3) share schedulers with four spinning
// epilogue warps and ran at ~0.3 instructions per cycle (profiles/r02e_halo_bench_fuse.txt).
__device__ __forceinline__ void mbar_wait_parked(uint64_t *bar, uint32_t parity) {
#ifdef TS2D_MBAR_SPIN_LIMIT
    mbar_wait(bar, parity);
#else
    uint32_t ok;
    asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    while (!ok) {
        __nanosleep(TS2D_PARK_NS);
        asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
#endif
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_5d(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile("cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
// ---- warp-uniform TMA issue (same reason as the warp-uniform MMA issue below) ----
// The producer warp runs its loop CONVERGENTLY: all 32 lanes compute the (uniform) coordinates, which then live in uniform registers, and
// only the copy itself is predicated on one elected lane.  Issued from inside `if (lane == 0)` every operand is a per-thread value and each
// UTMALDG costs an ELECT + 6 x R2UR.BROADCAST "waterfall" loop; measured on conv_tc2_kernel, whose k-chunks are bound by how fast the
// producer can issue: 256 -> 256 channels at 64^2 0.064 -> 0.056 ms for one extra branch in that loop (profiles/r02t_layers_exp.txt).
__device__ __forceinline__ void mbar_expect_tx_pred(uint32_t leader, uint64_t *bar, uint32_t bytes) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %0, 0;\n@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %2;\n}" ::"r"(leader), "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_3d_pred(uint32_t leader, uint32_t dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %0, 0;\n@q cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%1], [%2, {%4, %5, %6}], [%3];\n}"
                 ::"r"(leader), "r"(dst), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_4d_pred(uint32_t leader, uint32_t dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, int c3) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %0, 0;\n@q cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%1], [%2, {%4, %5, %6, %7}], [%3];\n}"
                 ::"r"(leader), "r"(dst), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_5d_pred(uint32_t leader, uint32_t dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %0, 0;\n@q cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%1], [%2, {%4, %5, %6, %7, %8}], [%3];\n}"
                 ::"r"(leader), "r"(dst), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
__device__ __forceinline__ uint32_t elect_one();
__device__ __forceinline__ void tma_load_3d_mc(void *dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4, %5}], [%2], %6;"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "h"(mask) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tcgen05_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// ---- warp-uniform MMA issue -------------------------------------------------------------------------------------
// The MMA warp runs its loop CONVERGENTLY (all 32 lanes wait on the barriers and compute the descriptors, which
// the compiler then keeps in uniform registers); only the tcgen05 instructions are predicated on one elected lane.
// Issuing from inside an `if (lane == 0)` region instead makes every descriptor a per-thread value and costs a
// R2UR "waterfall" loop plus ~20 scalar instructions per MMA (measured: ~100 cycles per N=32 MMA that needs 16).
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}" : "=r"(pred));
    return pred;
}
// descriptor = {lo, hi}: lo = start address >> 4 (14 bits) | LBO field, hi = SBO | version | swizzle mode (constant per operand)
__device__ __forceinline__ void tcgen05_mma_f16_lohi(uint32_t leader, uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                     uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, q;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 q, %0, 0;\n"
        "setp.ne.b32 p, %7, 0;\n"
        "mov.b64 da, {%2, %3};\n"
        "mov.b64 db, {%4, %5};\n"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%1], da, db, %6, p;\n"
        "}\n" ::"r"(leader), "r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tcgen05_commit_pred(uint32_t leader, uint64_t *bar) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %0, 0;\n"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%1];\n"
        "}\n" ::"r"(leader), "r"(smem_u32(bar)) : "memory");
}
// same, arriving on the barrier at this offset in every CTA of `mask` (frees the slot for both producers of a multicast pair)
__device__ __forceinline__ void tcgen05_commit_mc_pred(uint32_t leader, uint64_t *bar, uint16_t mask) {
    asm volatile(
        "{\n"
        ".reg .pred q;\n"
        "setp.ne.b32 q, %0, 0;\n"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%1], %2;\n"
        "}\n" ::"r"(leader), "r"(smem_u32(bar)), "h"(mask) : "memory");
}
template <int KC> __device__ __forceinline__ uint32_t desc_lo(uint32_t saddr) { return ((saddr >> 4) & 0x3FFFu) | (1u << 16); }
template <int KC> __device__ __forceinline__ constexpr uint32_t desc_hi(uint32_t sbo_bytes) {
    return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | ((KC == 64 ? 2u : 4u) << 29);
}

__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout), swizzle by row width.
template <int KC>
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    constexpr uint64_t layout = (KC == 64) ? 2ull /*SWIZZLE_128B*/ : 4ull /*SWIZZLE_64B*/;
    constexpr uint64_t sbo = (KC == 64) ? 1024 : 512;   // bytes between 8-row groups
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                 // LBO (unused for swizzled K-major)
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;                 // descriptor version (sm_100)
    d |= layout << 61;
    return d;
}
// instruction descriptor: D=F32, A/B = F16(0)/BF16(1), both K-major, N>>3 at [17,23), M>>4 at [24,29)
template <typename DT>
__device__ __forceinline__ uint32_t make_idesc(int n) {
    const uint32_t fmt = sizeof(DT) == 2 && std::is_same<DT, __nv_bfloat16>::value ? 1u : 0u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}

template <int BN> struct TcCfg {
    static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
};
template <int BN, int KC> constexpr int tc_stage_bytes() { return (128 + BN) * KC * 2; }
template <int BN, int KC> constexpr int tc_num_stages() {
    return (200 * 1024 / tc_stage_bytes<BN, KC>()) > 8 ? 8 : (200 * 1024 / tc_stage_bytes<BN, KC>());
}
template <int BN, int KC> constexpr int tc_smem_bytes() { return tc_num_stages<BN, KC>() * tc_stage_bytes<BN, KC>() + 1024 /*align*/ + 256 /*barriers*/; }

template <typename DT> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
template <typename DT> __device__ __forceinline__ float2 unpack2(uint32_t u);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162 *>(&u)); }
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t u) { return __half22float2(*reinterpret_cast<__half2 *>(&u)); }

// ---- packed fp32x2 arithmetic (FADD2 / FFMA2 on sm_100): two lanes per issue slot in the epilogues ----
__device__ __forceinline__ uint64_t f32x2_pack(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ uint64_t f32x2_pack_bits(uint32_t lo, uint32_t hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi)); return r; }
__device__ __forceinline__ float f32x2_lo(uint64_t v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float f32x2_hi(uint64_t v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ uint64_t f32x2_add(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t f32x2_mul(uint64_t a, uint64_t b) { uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t f32x2_fma(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// One 32-column accumulator chunk of a thread's pixel row: v = acc + bias (fp32x2), pk = v rounded to DT (what is stored),
// and - when `stats` - the thread's running InstanceNorm sums t1 += v, t2 += v*v.  The statistics are taken from the fp32
// values BEFORE rounding: that is what InstanceNorm sees in the fp32 reference, and it saves the unpack of the rounded pairs
// (the epilogue is issue-slot bound: 64 instructions per chunk instead of 144).  bias32: 32 floats, 16-byte aligned, or null.
template <typename DT>
__device__ __forceinline__ void epi_chunk(const uint32_t (&r)[32], const float *bias32, uint32_t (&pk)[16], uint64_t (&t1)[16], uint64_t (&t2)[16],
                                          bool stats) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        uint64_t v0 = f32x2_pack_bits(r[4 * k], r[4 * k + 1]), v1 = f32x2_pack_bits(r[4 * k + 2], r[4 * k + 3]);
        if (bias32) {
            const float4 b = *reinterpret_cast<const float4 *>(bias32 + 4 * k);       // warp-uniform address: broadcast
            v0 = f32x2_add(v0, f32x2_pack(b.x, b.y));
            v1 = f32x2_add(v1, f32x2_pack(b.z, b.w));
        }
        pk[2 * k] = pack2<DT>(f32x2_lo(v0), f32x2_hi(v0));
        pk[2 * k + 1] = pack2<DT>(f32x2_lo(v1), f32x2_hi(v1));
        if (stats) {
            t1[2 * k] = f32x2_add(t1[2 * k], v0);
            t1[2 * k + 1] = f32x2_add(t1[2 * k + 1], v1);
            t2[2 * k] = f32x2_fma(v0, v0, t2[2 * k]);
            t2[2 * k + 1] = f32x2_fma(v1, v1, t2[2 * k + 1]);
        }
    }
}
// same without accumulation: v[16] = the 32 fp32 values (acc + bias) as pairs
template <typename DT>
__device__ __forceinline__ void epi_vals(const uint32_t (&r)[32], const float *bias32, uint32_t (&pk)[16], uint64_t (&v)[16]) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        uint64_t v0 = f32x2_pack_bits(r[4 * k], r[4 * k + 1]), v1 = f32x2_pack_bits(r[4 * k + 2], r[4 * k + 3]);
        if (bias32) {
            const float4 b = *reinterpret_cast<const float4 *>(bias32 + 4 * k);
            v0 = f32x2_add(v0, f32x2_pack(b.x, b.y));
            v1 = f32x2_add(v1, f32x2_pack(b.z, b.w));
        }
        pk[2 * k] = pack2<DT>(f32x2_lo(v0), f32x2_hi(v0));
        pk[2 * k + 1] = pack2<DT>(f32x2_lo(v1), f32x2_hi(v1));
        v[2 * k] = v0;
        v[2 * k + 1] = v1;
    }
}
// per-thread fp32x2 sums -> lane L gets the warp total of channel L (and the accumulators are cleared)
__device__ __forceinline__ float warp_transpose_sum32(float (&x)[32], int lane);
__device__ __forceinline__ void epi_flush(uint64_t (&t1)[16], uint64_t (&t2)[16], int lane, float &s1, float &s2) {
    float a[32];
#pragma unroll
    for (int j = 0; j < 16; ++j) { a[2 * j] = f32x2_lo(t1[j]); a[2 * j + 1] = f32x2_hi(t1[j]); t1[j] = 0ull; }
    s1 = warp_transpose_sum32(a, lane);
#pragma unroll
    for (int j = 0; j < 16; ++j) { a[2 * j] = f32x2_lo(t2[j]); a[2 * j + 1] = f32x2_hi(t2[j]); t2[j] = 0ull; }
    s2 = warp_transpose_sum32(a, lane);
}

// lane L ends up with the sum over the warp of x[L]; 31 shuffles
__device__ __forceinline__ float warp_transpose_sum32(float (&x)[32], int lane) {
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        const bool upper = (lane & step) != 0;
#pragma unroll
        for (int i = 0; i < step; ++i) {
            const float send = upper ? x[i] : x[i + step];
            const float keep = upper ? x[i + step] : x[i];
            x[i] = keep + __shfl_xor_sync(0xffffffffu, send, step);
        }
    }
    return x[0];
}


template <int BN, int KC, typename DT>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB, const ConvTcParams p) {
    constexpr int kStages = tc_num_stages<BN, KC>();
    constexpr int kABytes = 128 * KC * 2;
    constexpr int kStageBytes = tc_stage_bytes<BN, KC>();
    constexpr int kTmemCols = TcCfg<BN>::kTmemCols;
    constexpr bool kAlt = BN == 32;                          // epilogue groups alternate tiles (else: split the columns)
    constexpr int kMyChunks = kAlt ? 1 : BN / 64;            // 32-column chunks per epilogue thread and tile
    constexpr bool kThreadAcc = kMyChunks == 1;              // per-thread statistics accumulators (64 registers)

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes);
    uint64_t *empty_bar = full_bar + kStages;
    uint64_t *tmem_full = empty_bar + kStages;     // [2]
    uint64_t *tmem_empty = tmem_full + 2;          // [2]
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    long long *dbg = p.dbg ? p.dbg + 8 * blockIdx.x : nullptr;
    const long long t_start = dbg ? clock64() : 0;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA2) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], p.mc ? 2 : 1); }      // mc: both CTAs' MMAs free a slot
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], kAlt ? 128 : 256); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    if (p.mc) cluster_sync_all();                   // the peer's barriers exist before anything is multicast to them
    const uint32_t rank = p.mc ? cluster_ctarank() : 0u;
    const uint32_t tmem_base = *tmem_ptr;
    griddep_launch_dependents();
    griddep_wait();                                                     // activations / statistics of the previous launches are complete
    if (dbg && threadIdx.x == 0) dbg[0] = clock64() - t_start;          // prologue (barriers, TMEM allocation) + wait for the predecessor

    const int m_tiles = p.tiles_n * p.tiles_h * p.tiles_w;
    const int ksplit = p.ksplit > 1 ? p.ksplit : 1;
    // the tile sequence is walked in units of `mstep` M tiles (mc: a pair of CTAs takes M tiles 2j and 2j + 1; m_tiles is even then)
    const int mstep = p.mc ? 2 : 1;
    const int mseq = m_tiles / mstep;
    const int total_tiles = mseq * p.n_tiles_n * ksplit;
    const int k_chunks = p.taps * p.cin_chunks;
    const int k_per = k_chunks / ksplit;          // k-chunks per tile (the host guarantees divisibility)
    // tile t -> channel block nt = t / mseq (slowest), m tile = (t % mseq) * mstep + rank = (image group, y tile, x tile) with x fastest
    const int nparts = gridDim.x / mstep, part = blockIdx.x / mstep;
    const int t_begin = (int)((long long)total_tiles * part / nparts);
    const int t_end = (int)((long long)total_tiles * (part + 1) / nparts);

    if (warp == 0) {
        // ===================== TMA producer (whole warp convergent, one elected lane issues; no divisions in the k loop) =====================
        const uint32_t leader = elect_one();
        const uint32_t s0 = smem_u32(smem);
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = t_begin; tile < t_end; ++tile) {
            const int ot = tile / ksplit, ks = tile - ot * ksplit;      // (output tile, K slice), slice fastest
            const int nt = ot / mseq, mt = (ot - nt * mseq) * mstep + (int)rank;
            const int xt = mt % p.tiles_w, yt = (mt / p.tiles_w) % p.tiles_h, nn = mt / (p.tiles_w * p.tiles_h);
            const int x0 = xt * p.tw, y0 = yt * p.th, n0 = nn * p.tn;
            const int kc0 = ks * k_per;
            int tap = kc0 / p.cin_chunks, cc = kc0 - tap * p.cin_chunks;
            int r = tap / p.kw, s = tap - r * p.kw;
            for (int kc = 0; kc < k_per; ++kc) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                const uint32_t sa = s0 + (uint32_t)(stage * kStageBytes);
                const uint32_t sb = sa + kABytes;
                mbar_expect_tx_pred(leader, &full_bar[stage], kStageBytes);
                if (!p.stride2) {
                    const bool second = cc >= p.split_chunks;                       // K-chunks of the second input tensor (decoder: the skip)
                    tma_load_4d_pred(leader, sa, second ? &tmA2 : &tmA, &full_bar[stage], (second ? cc - p.split_chunks : cc) * KC, x0 + s - p.pad_w,
                                     y0 + r - p.pad_h, n0);
                } else {
                    // input row sh*oy + r - pad = sh*h2 + py ; input col sw*ox + s - pad = sw*w2 + px
                    const int dr = r - p.pad_h, ds = s - p.pad_w;
                    const int py = (p.stride2 & 2) ? (dr & 1) : 0, px = (p.stride2 & 1) ? (ds & 1) : 0;
                    const int hoff = (p.stride2 & 2) ? (dr - py) / 2 : dr, woff = (p.stride2 & 1) ? (ds - px) / 2 : ds;
                    tma_load_5d_pred(leader, sa, &tmA, &full_bar[stage], px * p.in_ps + cc * KC, x0 + woff, py, y0 + hoff, n0);
                }
                if (p.mc) {
                    if (leader) tma_load_3d_mc(smem + stage * kStageBytes + kABytes + rank * (BN / 2 * KC * 2), &tmB, &full_bar[stage], cc * KC, nt * BN + (int)rank * (BN / 2), tap, (uint16_t)3);
                } else tma_load_3d_pred(leader, sb, &tmB, &full_bar[stage], cc * KC, nt * BN, tap);
                if (++stage == kStages) { stage = 0; phase ^= 1; }
                if (dbg && lane == 0 && tile == t_begin && kc == 0) dbg[1] = clock64() - t_start;     // first loads issued
                if (++cc == p.cin_chunks) { cc = 0; ++tap; if (++s == p.kw) { s = 0; ++r; } }
            }
        }
        if (dbg && lane == 0) dbg[2] = clock64() - t_start;                                            // all loads issued
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp convergent, one elected lane issues) =====================
        const uint32_t leader = elect_one();
        const uint32_t idesc = make_idesc<DT>(BN);
        constexpr uint32_t hi = desc_hi<KC>(KC == 64 ? 1024 : 512);
        const uint32_t s0_lo = desc_lo<KC>(smem_u32(smem));
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int tile = t_begin; tile < t_end; ++tile, ++it) {
            const int buf = it & 1;
            const uint32_t use = (uint32_t)(it >> 1);
            mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
            tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
            for (int kc = 0; kc < k_per; ++kc) {
                mbar_wait(&full_bar[stage], phase);
                tcgen05_fence_after();
                if (dbg && lane == 0 && it == 0 && kc == 0) dbg[3] = clock64() - t_start;              // first operands landed
                const uint32_t a_lo = s0_lo + (uint32_t)(stage * (kStageBytes >> 4));
                const uint32_t b_lo = a_lo + (uint32_t)(kABytes >> 4);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k)   // 16 elements = 32 bytes along K inside the swizzled row: +2 in the (addr >> 4) field
                    tcgen05_mma_f16_lohi(leader, d_tmem, a_lo + (uint32_t)(k * 2), hi, b_lo + (uint32_t)(k * 2), hi, idesc, k ? 1u : (uint32_t)(kc != 0));
                if (p.mc) tcgen05_commit_mc_pred(leader, &empty_bar[stage], (uint16_t)3);
                else tcgen05_commit_pred(leader, &empty_bar[stage]);       // frees the smem slot once these MMAs retire
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            tcgen05_commit_pred(leader, &tmem_full[buf]);             // accumulator complete
        }
        if (dbg && lane == 0) dbg[4] = clock64() - t_start;                                           // all MMAs issued
    } else if (warp >= 4) {
        // ===================== epilogue (warps 4..11, two warpgroups) =====================
        const int grp = (warp - 4) >> 2;
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;          // accumulator row = pixel index in the tile
        const int ix = row % p.tw, iy = (row / p.tw) % p.th, in_ = row / (p.tw * p.th);
        const int chunk0 = kAlt ? 0 : grp * kMyChunks;
        DT *out = reinterpret_cast<DT *>(p.out);
        const bool do_stats = p.stats != nullptr;
        const bool fast_stats = do_stats && p.tn == 1;     // a tile lies inside one image: sums can stay in registers across tiles
        uint64_t tacc1[16], tacc2[16];                                    // per-thread fp32x2 sums of this thread's pixel row (kThreadAcc)
        float racc1[kMyChunks], racc2[kMyChunks];                         // per-lane (= per-channel) sums (otherwise)
#pragma unroll
        for (int i = 0; i < 16; ++i) { tacc1[i] = 0ull; tacc2[i] = 0ull; }
#pragma unroll
        for (int i = 0; i < kMyChunks; ++i) { racc1[i] = 0.f; racc2[i] = 0.f; }
        auto flush_stats = [&](int n_img, int nt_blk) {
            if constexpr (kThreadAcc) epi_flush(tacc1, tacc2, lane, racc1[0], racc2[0]);
#pragma unroll
            for (int c = 0; c < kMyChunks; ++c) {
                double *s = p.stats + ((long long)n_img * p.Cout + nt_blk * BN + (chunk0 + c) * 32 + lane) * 2;
                atomicAdd(s, (double)racc1[c]);
                atomicAdd(s + 1, (double)racc2[c]);
                racc1[c] = 0.f; racc2[c] = 0.f;
            }
        };
        int cur_n = -1, cur_nt = -1;
        int map_nt = -1, map_c0[kMyChunks];
        long long map_off[kMyChunks];
        // tile walker (no divisions in the loop): this group's tiles are t_begin + e, e += estep
        const int estep = kAlt ? 2 : 1;
        int tile = t_begin + (kAlt ? grp : 0);
        int nt = tile / mseq;                   // (partial mode decodes its own tiles below)
        int xt, yt, nn;
        {
            const int mt = (tile - nt * mseq) * mstep + (int)rank;
            xt = mt % p.tiles_w; yt = (mt / p.tiles_w) % p.tiles_h; nn = mt / (p.tiles_w * p.tiles_h);
        }
        if (p.partial) {
            // ---- split-K: fp32 accumulators of K slice ks -> partial[ks][pixel][Cout]; few tiles per CTA, so plain divisions ----
            const long long mtot = (long long)p.N * p.Hout * p.Wout;
            for (int e = kAlt ? grp : 0, t = t_begin + e; t < t_end; t += estep, e += estep) {
                const int ot = t / ksplit, ks = t - ot * ksplit;
                const int snt = ot / m_tiles, mt = ot - snt * m_tiles;
                const int sxt = mt % p.tiles_w, syt = (mt / p.tiles_w) % p.tiles_h, snn = mt / (p.tiles_w * p.tiles_h);
                const int buf = e & 1;
                const uint32_t par = (uint32_t)(e >> 1) & 1u;
                const int x = sxt * p.tw + ix, y = syt * p.th + iy, n = snn * p.tn + in_;
                const bool valid = (x < p.Wout) && (y < p.Hout) && (n < p.N);
                mbar_wait(&tmem_full[buf], par);
                tcgen05_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN);
                float *prow = p.partial + ((long long)ks * mtot + ((long long)n * p.Hout + y) * p.Wout + x) * p.Cout + snt * BN;
#pragma unroll
                for (int c = 0; c < kMyChunks; ++c) {
                    const int chunk = chunk0 + c;
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(taddr + chunk * 32, r);
                    tmem_ld_wait();
                    if (c == kMyChunks - 1) {
                        tcgen05_fence_before();
                        mbar_arrive(&tmem_empty[buf]);
                    }
                    if (valid) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(prow + chunk * 32 + 8 * k), "r"(r[8 * k]),
                                         "r"(r[8 * k + 1]), "r"(r[8 * k + 2]), "r"(r[8 * k + 3]), "r"(r[8 * k + 4]), "r"(r[8 * k + 5]), "r"(r[8 * k + 6]),
                                         "r"(r[8 * k + 7]) : "memory");
                    }
                }
            }
        } else
        for (int e = kAlt ? grp : 0; tile < t_end; tile += estep, e += estep) {
            const int buf = e & 1;
            const uint32_t par = (uint32_t)(e >> 1) & 1u;
            const int x = xt * p.tw + ix, y = yt * p.th + iy, n = nn * p.tn + in_;
            const bool valid = (x < p.Wout) && (y < p.Hout) && (n < p.N);
            if (fast_stats && (nn != cur_n || nt != cur_nt)) {
                if (cur_n >= 0) flush_stats(cur_n, cur_nt);
                cur_n = nn;          // tn == 1 -> the tile's image
                cur_nt = nt;
            }
            mbar_wait(&tmem_full[buf], par);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN);
            DT *opix = out + (((long long)n * p.Hfull + (long long)y * p.my) * p.Wfull + (long long)x * p.mx) * p.out_ps;
#pragma unroll
            for (int c = 0; c < kMyChunks; ++c) {
                const int chunk = chunk0 + c;
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + chunk * 32, r);
                tmem_ld_wait();
                if (c == kMyChunks - 1) {
                    // all of this thread's TMEM reads for the tile are done: release the accumulator buffer
                    tcgen05_fence_before();
                    mbar_arrive(&tmem_empty[buf]);
                }
                // column block -> (tap, channel) : transposed conv scatters tap (dy, dx) to pixel (y*my + dy, x*mx + dx).  The mapping depends
                // on the channel block only: its two integer divisions are redone when nt changes, not per tile (the epilogue of the
                // 1-chunk-per-tile transposed convs is issue bound)
                if (nt != map_nt) {
#pragma unroll
                    for (int cm = 0; cm < kMyChunks; ++cm) {
                        const int gcol_m = nt * BN + (chunk0 + cm) * 32;
                        const int tap_m = gcol_m / p.tap_cout, c0_m = gcol_m - tap_m * p.tap_cout;
                        const int dy_m = tap_m / p.mx, dx_m = tap_m - dy_m * p.mx;
                        map_c0[cm] = c0_m;
                        map_off[cm] = ((long long)dy_m * p.Wfull + dx_m) * p.out_ps + c0_m;
                    }
                    map_nt = nt;
                }
                const int gcol = nt * BN + chunk * 32;
                const int c0 = map_c0[c];
                DT *dst = opix + map_off[c];
                uint32_t pk[16];
                uint64_t v[16];                  // fp32 values of this chunk (warp-reduced path only)
                const bool tstat = do_stats && kThreadAcc && fast_stats;
                if (tstat) epi_chunk<DT>(r, p.bias ? p.bias + c0 : nullptr, pk, tacc1, tacc2, valid);     // overhanging rows do not contribute
                else epi_vals<DT>(r, p.bias ? p.bias + c0 : nullptr, pk, v);
                if (valid) {
                    // 2 x 32-byte stores (STG.256): every sector is written whole by one instruction
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * k), "r"(pk[8 * k]), "r"(pk[8 * k + 1]),
                                     "r"(pk[8 * k + 2]), "r"(pk[8 * k + 3]), "r"(pk[8 * k + 4]), "r"(pk[8 * k + 5]), "r"(pk[8 * k + 6]), "r"(pk[8 * k + 7]) : "memory");
                }
                if (do_stats && !tstat) {
                    float a[32], b[32];
#pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        a[2 * k] = valid ? f32x2_lo(v[k]) : 0.f;
                        a[2 * k + 1] = valid ? f32x2_hi(v[k]) : 0.f;
                    }
#pragma unroll
                    for (int k = 0; k < 32; ++k) b[k] = a[k] * a[k];
                    if (fast_stats) {
                        racc1[c] += warp_transpose_sum32(a, lane);
                        racc2[c] += warp_transpose_sum32(b, lane);
                    } else if (valid) {
                        // several images per tile (tiny levels): one atomic pair per value
                        double *s = p.stats + ((long long)n * p.Cout + gcol) * 2;
#pragma unroll
                        for (int k = 0; k < 32; ++k) {
                            atomicAdd(s + 2 * k, (double)a[k]);
                            atomicAdd(s + 2 * k + 1, (double)b[k]);
                        }
                    }
                }
            }
            xt += estep * mstep;
            while (xt >= p.tiles_w) {
                xt -= p.tiles_w;
                if (++yt == p.tiles_h) { yt = 0; if (++nn == p.tiles_n) { nn = 0; ++nt; } }
            }
        }
        if (fast_stats && cur_n >= 0) flush_stats(cur_n, cur_nt);
        if (dbg && warp == 4 && lane == 0) dbg[5] = clock64() - t_start;                              // epilogue done
    }
    tcgen05_fence_before();
    __syncthreads();
    if (p.mc) cluster_sync_all();                   // the peer may still multicast into this CTA's shared memory / arrive on its barriers
    if (dbg && threadIdx.x == 0) dbg[6] = clock64() - t_start;
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

}  // namespace ts2d
