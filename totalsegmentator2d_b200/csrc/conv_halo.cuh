// K4h: 3x3 / stride-1 implicit-GEMM convolution with a shared-memory HALO tile and resident weights
// (tcgen05 + TMEM, fed by TMA) - the high-resolution levels of the U-Net.
//
// Why a second conv kernel: conv_tc_kernel re-loads the A operand once per filter tap (9 TMA box loads of the same
// pixels through L2).  At 32/64 channels the layer is HBM-bound by its algorithmic bytes, but 9x the input through
// L2 (measured LTS cap ~2x HBM) made it L2-bound at ~15% of HBM speed.  Here the input box of a tile is loaded ONCE:
//
//   tile     = 8 (x) x 16 (y) output pixels = the 128 rows of one UMMA (row i -> x = i & 7, y = i >> 3)
//   halo     = 10 x 18 input pixels x KC channels, one 4-D TMA box (zero fill outside the image = conv padding),
//              stored pixel-major with the 64/128-byte swizzle, i.e. pixel p of the halo is "row p" of a K-major
//              operand whose 8-row groups are 10 pixels (= one halo row) apart
//   tap(r,s) = the SAME shared memory viewed through a descriptor whose start address is advanced by
//              (r*10 + s) rows and whose stride-byte-offset is 10 rows.  The swizzle XOR is a function of the
//              absolute shared-memory address (probed on B200: tools/probe/desc_probe.cu), so a start address
//              that is 64/128-byte but not 1024-byte aligned is read correctly with base_offset = 0.
//   weights  = all 9 taps x Cin x BN of the CTA's output-channel block stay resident in shared memory for the
//              whole persistent loop (loaded once per CTA): per tile only the halo moves.
//
// Measured on B200 (tools/probe/mma_rate.cu): one M=128, K=16 tcgen05.mma costs ~45 cycles at N=32 and ~51 at N=64
// (the N/2-cycle floor only shows from N=128 on): the A operand is read from shared memory at 128 B/clk (4 KB per instruction = 32
// cycles), so these layers are bound by the NUMBER of MMAs (9 taps x Cin/16 per tile); everything else - TMA, the operand transform,
// the epilogue - must stay off that critical path.
//
// Deferred normalisation (FUSE): the input is the PRODUCER conv's raw output; its InstanceNorm + LeakyReLU are applied
// to the halo in shared memory between the TMA arrival and the MMAs (y = leaky(x*scale[n][c] + shift[n][c]), zeros
// outside the image stay zeros = the conv padding of the normalised tensor), so the normalised tensor is never
// written to or read from HBM.  Same fma / leaky arithmetic as norm_act_kernel (the transformed operand is bit-identical; the
// fused kernel's own output statistics are summed in a different grouping, see tests/test_gpu_unet.py::test_deferred_norm_matches_materialised).
//
// Warp roles.  Plain kernel, 384 threads: warps 0-7 = two EPILOGUE warpgroups, (8-9 idle), warp 10 TMA producer, warp 11 TMEM alloc +
// MMA issuer (whole warp convergent, one elected lane issues).  FUSE kernel, 512 threads: warps 0-3 = ONE epilogue warpgroup,
// warps 4-11 = eight TRANSFORM warps, (12-13 idle), warp 14 TMA producer, warp 15 MMA issuer; setmaxnreg per role (168 / 216, 96, 64).
//   transform (FUSE)  the 256 threads split the pixels of every halo stage and rewrite it in place between the TMA arrival (full) and
//                     the MMAs (ready): a thread owns one 16-byte channel block (its 8 scale / shift pairs live in registers), loads
//                     batched before the math; fp32x2 FMA, leaky = max(y, slope*y); one elected arrival per warp.  History
//                     (profiles/r02b..r02m_halo_bench_fuse.txt): the first version ran the transform on the epilogue warps (transform of
//                     tile j, then epilogue of tile j-2): the MMA warp waited for TMEM 43% of the time, 2396 cycles per tile against 1025
//                     unfused, 1790 with an EMPTY transform; generic LD/ST instead of LDS/STS cost 3x; one warp sustains ~0.4
//                     instructions per cycle on this code (tools/probe/xform_rate.cu), hence eight warps.
//   epilogue          TMEM -> +bias -> round -> per-thread InstanceNorm partial sums -> 32-byte global stores (statistics from the
//                     fp32 values, FADD2 / FFMA2: see epi_chunk).  Plain kernel: BN == 32: the two warpgroups alternate tiles (group g
//                     owns TMEM buffer g); BN == 64: both work on every tile, one half of the columns each.  FUSE kernel: the one
//                     warpgroup takes every tile and all BN / 32 chunks.
#pragma once
#include "conv_tc.cuh"

namespace ts2d {

struct ConvHaloParams {
    int N, H, W, Cout;          // images, output grid (== input grid), output channels
    int n_base;                 // first image of this launch (chunked execution over the batch)
    int cin_chunks;             // Cin / KC
    int split_chunks;           // halo chunks [0, split_chunks) come from tmA, the rest from tmA2 (see ConvTcParams); == cin_chunks for one input
    int tiles_x, tiles_y;       // ceil(W/8), ceil(H/16)
    int n_stages;               // halo ring depth (host: what fits beside the resident weights)
    long long out_ps;           // output pixel stride in elements
    int in_ps;                  // conv_halo_s2_kernel: input pixel stride in elements (selects the odd column of the parity view)
    void *out;
    const float *bias;          // [Cout] or null
    double *stats;              // [Nmax][Cout][2] or null  (sum, sum of squares of the ROUNDED outputs)
    // fused input transform (FUSE only)
    const double *in_stats;     // [Nmax][Cin][2] sum / sum of squares over the in_hw pixels of (image, channel)
    const float *in_gamma;      // [Cin] or null
    const float *in_beta;       // [Cin] or null
    int in_hw;                  // pixels per image of the producer's output
    float in_eps, in_slope;
    int Cin;
    // diagnostics (tools/probe/halo_bench.cu only; 0 in the product): knock-out bits
    //   1 = no InstanceNorm statistics, 2 = no global stores, 4 = only tap 0 is multiplied, 8 = every halo is tile 0's (L2 hits),
    //   32 = transform without the proxy fence (WRONG results, timing only), 64 = transform does no loads / math / stores
    int dbg;
    unsigned long long *dbg_out;   // [gridDim.x*gridDim.y][8] cycles: total / MMA waits on halo / MMA waits on TMEM / epilogue waits / transform waits / transform work
};

constexpr int kHaloW = 10, kHaloH = 18, kHaloPix = kHaloW * kHaloH;
constexpr int kHaloWorkers = 256;           // epilogue threads: warps 0-7
// Block shapes (see the header).  Tried and found irrelevant on B200: ordering the roles by warp id (the arbiter's preference for high warp
// ids), parking the waiting warps with try_wait suspend hints or __nanosleep - a spinning neighbour does not slow a working warp.
template <bool FUSE> constexpr int halo_threads() { return FUSE ? 512 : 384; }
template <bool FUSE> constexpr int halo_warp_tma() { return halo_threads<FUSE>() / 32 - 2; }
template <bool FUSE> constexpr int halo_warp_mma() { return halo_threads<FUSE>() / 32 - 1; }
constexpr int kHaloXformThreads = 256;      // FUSE: warps 4-11 (two warpgroups)

template <int KC> constexpr int halo_stage_bytes() { return ((kHaloPix * KC * 2 + 1023) / 1024) * 1024; }
// bytes after the halo ring: barriers (512) + bias [64] + scale/shift table [2][Cin <= kHaloMaxCin]
constexpr int kHaloMaxCin = 256;
constexpr int kHaloTailBytes = 512 + 64 * 4 + 2 * kHaloMaxCin * 4;

// explicit shared-space 16-byte accesses: the halo ring is addressed through a pointer whose alignment arithmetic hides the address
// space from the compiler, and GENERIC LD.E.128 / ST.E.128 on shared memory made the transform ~3x slower than LDS / STS
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, const uint4 &v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

template <int BN, int KC, typename DT, bool FUSE>
__global__ void __launch_bounds__(halo_threads<FUSE>(), 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB, const ConvHaloParams p) {
    static_assert(BN == 32 || BN == 64, "conv_halo_kernel: BN must be 32 or 64");
    constexpr int kRowB = KC * 2;                      // bytes per pixel row of the operand
    constexpr int kHaloBytes = kHaloPix * kRowB;       // bytes TMA delivers per halo
    constexpr int kStageBytes = halo_stage_bytes<KC>();
    constexpr int kWBlock = BN * kRowB;                // one (cin chunk, tap) weight block
    constexpr int kTmemCols = TcCfg<BN>::kTmemCols;
    constexpr int kMaxStages = 8;
    constexpr int kHaloWarpTma = halo_warp_tma<FUSE>(), kHaloWarpMma = halo_warp_mma<FUSE>();
    constexpr int kEpiWarps = FUSE ? 4 : 8;
    constexpr bool kSplit = BN > 32;                   // both worker groups on every tile (column halves) vs alternating tiles

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int w_bytes = 9 * p.cin_chunks * kWBlock;
    uint8_t *smem_w = smem;
    uint8_t *smem_a = smem + ((w_bytes + 1023) & ~1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem_a + p.n_stages * kStageBytes);
    uint64_t *empty_bar = full_bar + kMaxStages;
    uint64_t *ready_bar = empty_bar + kMaxStages;     // FUSE: halo transformed
    uint64_t *tmem_full = ready_bar + kMaxStages;     // [2]
    uint64_t *tmem_empty = tmem_full + 2;             // [2]
    uint64_t *w_bar = tmem_empty + 2;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(w_bar + 1);
    float *s_bias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(full_bar) + 512);     // [64]
    float *s_ss = s_bias + 64;                                                                   // [2][kHaloMaxCin]: scale, shift of the transform warps' current image

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nt = blockIdx.y;                         // output-channel block of this CTA

    if (warp == kHaloWarpTma && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA2) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < p.n_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&ready_bar[s], kHaloXformThreads / 32); }      // one elected arrival per transform warp
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], (!FUSE && kSplit) ? 8 : 4); }
        mbar_init(w_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kHaloWarpMma) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp < 4)
        for (int i = threadIdx.x; i < BN; i += 128) s_bias[i] = p.bias ? p.bias[blockIdx.y * BN + i] : 0.f;
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    griddep_launch_dependents();
    if (!(warp == kHaloWarpTma && lane == 0)) griddep_wait();      // (the producer thread first issues the weight loads: constants, see below)

    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int total_tiles = tiles_per_img * p.N;
    // static CONTIGUOUS partition of the tile sequence (x fastest, then y, then image) over the CTAs of this output-channel
    // block: neighbouring tiles share halo columns through L2, and a CTA touches at most two or three images, so the
    // InstanceNorm atomics of one (image, channel) come from ~1/10 of the CTAs and at different times (a round-robin
    // partition made all CTAs hit the same 64 addresses at every image boundary: ~5 us of serialised f64 atomics each).
    const int t_begin = (int)((long long)total_tiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)total_tiles * (blockIdx.x + 1) / gridDim.x);

    if (warp < kEpiWarps) {
        // ===================== epilogue warps =====================
        // plain kernel: warps 0-7 = two warpgroups; BN == 32: the groups alternate tiles (group g owns TMEM buffer g), BN == 64: both work on
        // every tile, one 32-column half each.  FUSE kernel: warps 0-3 = ONE warpgroup takes every tile and all BN / 32 chunks (it has the
        // time: the MMAs of a tile take 1000 - 4000 cycles, a chunk ~350), which leaves eight warps to the transform.
        if constexpr (FUSE) {
            if constexpr (BN == 64) asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");      // 2 x 64 statistics accumulators live in registers
            else asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
        }
        const int wt = threadIdx.x;             // 0..255 (FUSE: 0..127)
        const int grp = wt >> 7;                // epilogue warpgroup
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;
        const int ix = row & 7, iy = row >> 3;
        constexpr int kChunks = FUSE ? BN / 32 : 1;                   // 32-column chunks per thread and tile
        constexpr bool kAltTiles = !FUSE && !kSplit;                  // the two groups alternate tiles
        const int col0 = (!FUSE && kSplit) ? grp * 32 : 0;
        DT *out = reinterpret_cast<DT *>(p.out);
        const bool do_stats = p.stats != nullptr && !(p.dbg & 1);
        // InstanceNorm statistics: every thread keeps running sums of ITS pixel row for its 32 channels across the tiles of
        // one image (no cross-lane traffic per tile); the 32-lane transpose-reduction and the double atomics happen only
        // when the image changes (<= 3 flushes per CTA with the contiguous partition).
        uint64_t tacc1[kChunks][16], tacc2[kChunks][16];
#pragma unroll
        for (int c = 0; c < kChunks; ++c)
#pragma unroll
            for (int i = 0; i < 16; ++i) { tacc1[c][i] = 0ull; tacc2[c][i] = 0ull; }
        auto flush_stats = [&](int n_img) {
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                float s1, s2;
                epi_flush(tacc1[c], tacc2[c], lane, s1, s2);
                double *s = p.stats + ((long long)n_img * p.Cout + nt * BN + col0 + c * 32 + lane) * 2;
                atomicAdd(s, (double)s1);
                atomicAdd(s + 1, (double)s2);
            }
        };
        int cur_n = -1;
        long long w_epi = 0;
        // tile walker (coordinates advance with carries, no divisions in the loop)
        const int estep = kAltTiles ? 2 : 1;
        const int esx = estep % p.tiles_x, esy = (estep / p.tiles_x) % p.tiles_y, esn = estep / tiles_per_img;
        int e_tile = t_begin + (kAltTiles ? grp : 0);
        int tx = e_tile % p.tiles_x, ty = (e_tile / p.tiles_x) % p.tiles_y, n = e_tile / tiles_per_img + p.n_base;
        const int n_local = t_end - t_begin;
        for (int e = kAltTiles ? grp : 0; e < n_local; e += estep) {
            const int buf = e & 1;
            const uint32_t par = (uint32_t)(e >> 1) & 1u;
            const int x = tx * 8 + ix, y = ty * 16 + iy;
            const bool valid = (x < p.W) && (y < p.H) && !(p.dbg & 2);
            if (do_stats && n != cur_n) {
                if (cur_n >= 0) flush_stats(cur_n);
                cur_n = n;
            }
            const long long tw0 = p.dbg_out ? clock64() : 0;
            mbar_wait_parked(&tmem_full[buf], par);
            if (p.dbg_out) w_epi += clock64() - tw0;
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + col0);
            DT *dst = out + (((long long)n * p.H + y) * p.W + x) * p.out_ps + nt * BN + col0;
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(taddr + c * 32, r);
                tmem_ld_wait();
                if (c == kChunks - 1) {
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tmem_empty[buf]);        // this warp's TMEM reads of the tile are done (one arrival per warp)
                }
                uint32_t pk[16];
                epi_chunk<DT>(r, s_bias + col0 + c * 32, pk, tacc1[c], tacc2[c], do_stats && valid);
                if (valid) {
                    // 2 x 32-byte stores (STG.256): every sector is written whole by one instruction
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + c * 32 + 16 * k), "r"(pk[8 * k]), "r"(pk[8 * k + 1]),
                                     "r"(pk[8 * k + 2]), "r"(pk[8 * k + 3]), "r"(pk[8 * k + 4]), "r"(pk[8 * k + 5]), "r"(pk[8 * k + 6]), "r"(pk[8 * k + 7]) : "memory");
                }
            }
            tx += esx; if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
            ty += esy; if (ty >= p.tiles_y) { ty -= p.tiles_y; ++n; }
            n += esn;
        }
        if (do_stats && cur_n >= 0) flush_stats(cur_n);
        if (p.dbg_out && wt == 0) {
            unsigned long long *o = p.dbg_out + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 8;
            o[3] = (unsigned long long)w_epi;
        }
    } else if (FUSE && warp < kEpiWarps + kHaloXformThreads / 32) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 96;");
        // ===================== transform warps (4..11): y = leaky(x * scale[n][c] + shift[n][c]) on every halo stage, in place =====================
        // The 256 threads split the pixels of every stage (short latency per stage: the ring holds only 3 - 6 stages); a thread owns one
        // 16-byte channel block, its 8 scale / shift pairs live in registers.  Straight-line code: a batch = 4 loads, 16 independent
        // fp32x2 chains, 4 predicated stores; only border tiles test the pixel position.
        const int xt = threadIdx.x - 128;                                        // 0..255
        constexpr int kCpp = KC / 8;                                             // 16-byte chunks (8 channels) per halo pixel
        constexpr int kPixStep = kHaloXformThreads / kCpp;                       // halo pixels covered per step: 64 (KC=32) or 32 (KC=64)
        constexpr int kIters = (kHaloPix + kPixStep - 1) / kPixStep;             // steps per stage: 3 or 6 (the last one is partial)
        constexpr int kBatch = 4;
        constexpr uint32_t kXorMask = (KC == 64) ? 7u : 3u;
        const int cb = xt % kCpp;                                                // the thread's channel block (256 % kCpp == 0)
        // tile-invariant per-step data: byte offset inside a stage (absolute-address swizzle: the stage base is 1024-byte aligned, so
        // chunk index XOR ((row offset >> 7) & mask)) | halo x << 16 | halo y << 20 | bit 31: no pixel (offset 0)
        uint32_t meta[kIters];
#pragma unroll
        for (int i = 0; i < kIters; ++i) {
            const int pix = xt / kCpp + i * kPixStep;
            const uint32_t row_off = (uint32_t)(pix * kRowB);
            const uint32_t sw = (row_off >> 7) & kXorMask;
            meta[i] = pix < kHaloPix ? ((row_off + (((uint32_t)cb ^ sw) << 4)) | ((uint32_t)(pix % kHaloW) << 16) | ((uint32_t)(pix / kHaloW) << 20)) : 0x80000000u;
        }
        int xtx = t_begin % p.tiles_x, xty = (t_begin / p.tiles_x) % p.tiles_y, xn = t_begin / tiles_per_img + p.n_base;
        int xf_n = -1, xstage = 0;
        uint32_t xphase = 0;
        const uint64_t slope2 = f32x2_pack(p.in_slope, p.in_slope);
        long long x_wait = 0, x_work = 0;
        for (int tile = t_begin; tile < t_end; ++tile) {
            if (xn != xf_n) {
                // scale / shift of the producer's InstanceNorm for image xn, all Cin channels: same arithmetic as norm_act_kernel
                // (double mean / rstd, one rounding to float), once per image and CTA
                named_bar_sync(1, kHaloXformThreads);
                for (int c = xt; c < p.Cin; c += kHaloXformThreads) {
                    const double *st = p.in_stats + ((long long)xn * p.Cin + c) * 2;
                    const double mean = st[0] / (double)p.in_hw;
                    const double var = fmax(st[1] / (double)p.in_hw - mean * mean, 0.0);
                    const double rstd = 1.0 / sqrt(var + (double)p.in_eps);
                    const float ga = p.in_gamma ? __ldg(p.in_gamma + c) : 1.f, be = p.in_beta ? __ldg(p.in_beta + c) : 0.f;
                    s_ss[c] = (float)(rstd * (double)ga);
                    s_ss[kHaloMaxCin + c] = (float)((double)be - mean * rstd * (double)ga);
                }
                named_bar_sync(1, kHaloXformThreads);
                xf_n = xn;
            }
            // halo pixel (hx, hy) is image pixel (xtx*8 - 1 + hx, xty*16 - 1 + hy); outside the image TMA wrote zeros = the conv padding
            // of the NORMALISED tensor: those pixels are left alone
            const int hx_lo = 1 - xtx * 8, hx_hi = p.W + 1 - xtx * 8, hy_lo = 1 - xty * 16, hy_hi = p.H + 1 - xty * 16;
            const bool interior = hx_lo <= 0 && hx_hi >= kHaloW && hy_lo <= 0 && hy_hi >= kHaloH;       // warp-uniform: no per-pixel test
            for (int cc = 0; cc < p.cin_chunks; ++cc) {
                uint64_t sc2[4], sh2[4];
                {
                    const float4 *ps = reinterpret_cast<const float4 *>(s_ss + cc * KC + cb * 8);
                    const float4 *ph = reinterpret_cast<const float4 *>(s_ss + kHaloMaxCin + cc * KC + cb * 8);
                    const float4 a0 = ps[0], a1 = ps[1], b0 = ph[0], b1 = ph[1];
                    sc2[0] = f32x2_pack(a0.x, a0.y); sc2[1] = f32x2_pack(a0.z, a0.w); sc2[2] = f32x2_pack(a1.x, a1.y); sc2[3] = f32x2_pack(a1.z, a1.w);
                    sh2[0] = f32x2_pack(b0.x, b0.y); sh2[1] = f32x2_pack(b0.z, b0.w); sh2[2] = f32x2_pack(b1.x, b1.y); sh2[3] = f32x2_pack(b1.z, b1.w);
                }
                const long long tq0 = p.dbg_out ? clock64() : 0;
                mbar_wait_parked(&full_bar[xstage], xphase);
                const long long tq1 = p.dbg_out ? clock64() : 0;
                x_wait += tq1 - tq0;
                const uint32_t sbase = smem_u32(smem_a) + (uint32_t)(xstage * kStageBytes);
                auto run_stage = [&](auto check_tag) {
                    constexpr bool kCheck = decltype(check_tag)::value;
#pragma unroll
                    for (int b0 = 0; b0 < kIters; b0 += kBatch) {
                        uint4 v[kBatch];
                        bool ok[kBatch];
#pragma unroll
                        for (int i = 0; i < kBatch; ++i) {
                            if (b0 + i >= kIters) continue;
                            const uint32_t m = meta[b0 + i];
                            ok[i] = (b0 + i < kIters - 1) || !(m & 0x80000000u);          // compile-time true except in the last step
                            if (kCheck) {
                                const int hx = (int)((m >> 16) & 15u), hy = (int)((m >> 20) & 31u);
                                ok[i] = ok[i] && hx >= hx_lo && hx < hx_hi && hy >= hy_lo && hy < hy_hi;
                            }
#ifdef TS2D_HALO_PROBE
                            if (p.dbg & 128) { v[i] = make_uint4(m, m * 3u, m * 5u, m * 7u); continue; }      // probe: no shared-memory loads
#endif
                            v[i] = lds128(sbase + (m & 0xFFFFu));                        // (a thread without a pixel reads chunk 0: harmless)
                        }
#pragma unroll
                        for (int i = 0; i < kBatch; ++i) {
                            if (b0 + i >= kIters) continue;
                            uint32_t *w = reinterpret_cast<uint32_t *>(&v[i]);
#ifdef TS2D_HALO_PROBE
                            if (p.dbg & 256) continue;                                                             // probe: no arithmetic
#endif
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const float2 f = unpack2<DT>(w[k]);
                                // same roundings as norm_act_kernel: fma, then y > 0 ? y : y * slope (== max(y, y * slope) for 0 <= slope <= 1)
                                const uint64_t y = f32x2_fma(f32x2_pack(f.x, f.y), sc2[k], sh2[k]);
                                const uint64_t ys = f32x2_mul(y, slope2);
                                w[k] = pack2<DT>(fmaxf(f32x2_lo(y), f32x2_lo(ys)), fmaxf(f32x2_hi(y), f32x2_hi(ys)));
                            }
                        }
#pragma unroll
                        for (int i = 0; i < kBatch; ++i) {
                            if (b0 + i >= kIters) continue;
#ifdef TS2D_HALO_PROBE
                            if ((p.dbg & 128) && v[i].x != 0x12345u) continue;                                    // probe: (practically) no stores
#endif
                            if (ok[i]) sts128(sbase + (meta[b0 + i] & 0xFFFFu), v[i]);
                        }
                    }
                };
                if (!(p.dbg & 64)) {
                    if (interior) run_stage(std::false_type{});
                    else run_stage(std::true_type{});
                }
                if (!(p.dbg & 32)) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy writes -> visible to the MMA (async proxy)
                __syncwarp();
                if (lane == 0) mbar_arrive(&ready_bar[xstage]);
                if (p.dbg_out) x_work += clock64() - tq1;
                if (++xstage == p.n_stages) { xstage = 0; xphase ^= 1; }
            }
            if (++xtx == p.tiles_x) { xtx = 0; if (++xty == p.tiles_y) { xty = 0; ++xn; } }
        }
        if (p.dbg_out && xt == 0) {
            unsigned long long *o = p.dbg_out + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 8;
            o[4] = (unsigned long long)x_wait; o[5] = (unsigned long long)x_work;
        }
    } else {
        if constexpr (FUSE) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        if (warp == kHaloWarpTma) {
            // ===================== TMA producer =====================
            if (lane == 0) {
                mbar_expect_tx(w_bar, (uint32_t)w_bytes);
                for (int cc = 0; cc < p.cin_chunks; ++cc)
                    for (int tap = 0; tap < 9; ++tap)
                        tma_load_3d(smem_w + (cc * 9 + tap) * kWBlock, &tmB, w_bar, cc * KC, nt * BN, tap);
                griddep_wait();                             // the resident weights stream in while the previous launch drains
                int stage = 0;
                uint32_t phase = 0;
                for (int tile = t_begin; tile < t_end; ++tile) {
                    const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img + p.n_base;
                    for (int cc = 0; cc < p.cin_chunks; ++cc) {
                        mbar_wait_parked(&empty_bar[stage], phase ^ 1);
                        mbar_expect_tx(&full_bar[stage], kHaloBytes);
                        if (p.dbg & 8) tma_load_4d(smem_a + stage * kStageBytes, &tmA, &full_bar[stage], cc * KC, -1, -1, p.n_base);
                        else if (cc < p.split_chunks) tma_load_4d(smem_a + stage * kStageBytes, &tmA, &full_bar[stage], cc * KC, tx * 8 - 1, ty * 16 - 1, n);
                        else tma_load_4d(smem_a + stage * kStageBytes, &tmA2, &full_bar[stage], (cc - p.split_chunks) * KC, tx * 8 - 1, ty * 16 - 1, n);
                        if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        } else if (warp == kHaloWarpMma) {
            // ===================== MMA issuer (whole warp convergent, one elected lane issues; see conv_tc.cuh) =====================
            const uint32_t leader = elect_one();
            const uint32_t idesc = make_idesc<DT>(BN);
            constexpr uint32_t a_hi = desc_hi<KC>(kHaloW * kRowB);        // 8-row groups of the A view are one halo row (10 pixels) apart
            constexpr uint32_t b_hi = desc_hi<KC>(KC == 64 ? 1024 : 512);
            const uint32_t w_lo = desc_lo<KC>(smem_u32(smem_w));
            const uint32_t a0_lo = desc_lo<KC>(smem_u32(smem_a));
            mbar_wait(w_bar, 0);
            tcgen05_fence_after();
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            long long t_start = 0, w_halo = 0, w_tmem = 0;
            if (p.dbg_out) t_start = clock64();
            for (int tile = t_begin; tile < t_end; ++tile, ++it) {
                const int buf = it & 1;
                const uint32_t use = (uint32_t)(it >> 1);
                long long t0 = p.dbg_out ? clock64() : 0;
                mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
                if (p.dbg_out) w_tmem += clock64() - t0;
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
                for (int cc = 0; cc < p.cin_chunks; ++cc) {
                    t0 = p.dbg_out ? clock64() : 0;
                    mbar_wait(FUSE ? &ready_bar[stage] : &full_bar[stage], phase);
                    if (p.dbg_out) w_halo += clock64() - t0;
                    tcgen05_fence_after();
                    const uint32_t a_lo = a0_lo + (uint32_t)(stage * (kStageBytes >> 4));
                    const uint32_t b_lo = w_lo + (uint32_t)(cc * 9 * (kWBlock >> 4));
    #pragma unroll
                    for (int tap = 0; tap < 9; ++tap) {
                        if ((p.dbg & 4) && tap) continue;
                        const int r = tap / 3, s = tap % 3;
    #pragma unroll
                        for (int k = 0; k < KC / 16; ++k)
                            tcgen05_mma_f16_lohi(leader, d_tmem, a_lo + (uint32_t)((((r * kHaloW + s) * kRowB) >> 4) + k * 2), a_hi,
                                                 b_lo + (uint32_t)(((tap * kWBlock) >> 4) + k * 2), b_hi, idesc, (tap | k) ? 1u : (uint32_t)(cc != 0));
                    }
                    tcgen05_commit_pred(leader, &empty_bar[stage]);
                    if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
                }
                tcgen05_commit_pred(leader, &tmem_full[buf]);
            }
            if (p.dbg_out && leader) {
                unsigned long long *o = p.dbg_out + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 8;
                o[0] = (unsigned long long)(clock64() - t_start); o[1] = (unsigned long long)w_halo; o[2] = (unsigned long long)w_tmem;
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == kHaloWarpMma) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

}  // namespace ts2d
