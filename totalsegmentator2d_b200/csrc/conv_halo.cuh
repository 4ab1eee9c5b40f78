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
// Optional fused InstanceNorm-apply + LeakyReLU of the PRODUCER layer on the A operand ("deferred norm"): four
// transform warps rewrite the halo in place (y = leaky(x*scale[n][c] + shift[n][c]), 0 outside the image) between
// the TMA arrival and the MMAs, so the normalised tensor is never written to or read from HBM.
//
// Warp roles: warp 0 TMA producer, warp 1 TMEM alloc + MMA issuer, warps 2-9 epilogue (two groups of four, one per TMEM
// accumulator buffer: TMEM -> +bias -> round -> InstanceNorm sum/sumsq -> global), warps 10-13 halo transform (only when FUSE).
#pragma once
#include "conv_tc.cuh"

namespace ts2d {

struct ConvHaloParams {
    int N, H, W, Cout;          // images, output grid (== input grid), output channels
    int n_base;                 // first image of this launch (chunked execution over the batch)
    int cin_chunks;             // Cin / KC
    int tiles_x, tiles_y;       // ceil(W/8), ceil(H/16)
    int n_stages;               // halo ring depth (host: what fits beside the resident weights)
    long long out_ps;           // output pixel stride in elements
    void *out;
    const float *bias;          // [Cout] or null
    double *stats;              // [Nmax][Cout][2] or null  (sum, sum of squares of the ROUNDED outputs)
    // fused input transform (FUSE only): per (image, input channel) scale / shift, LeakyReLU slope per channel
    const float *in_scale;      // [Nmax][Cin]
    const float *in_shift;      // [Nmax][Cin]
    const float *in_slope;      // [Cin]   (1.0 = no activation, e.g. the transposed-conv half of a concat buffer)
    int Cin;
    // diagnostics (tools/probe/halo_bench.cu only; 0 in the product): knock-out bits
    //   1 = no InstanceNorm statistics, 2 = no global stores, 4 = only tap 0 is multiplied, 8 = every halo is tile 0's (L2 hits)
    int dbg;
    unsigned long long *dbg_out;   // [gridDim.x*gridDim.y][4]: cycles total / MMA-issuer waits on halo / on TMEM / epilogue waits
};

constexpr int kHaloW = 10, kHaloH = 18, kHaloPix = kHaloW * kHaloH;

template <int KC> constexpr int halo_stage_bytes() { return ((kHaloPix * KC * 2 + 1023) / 1024) * 1024; }

template <int KC>
__device__ __forceinline__ uint64_t make_smem_desc_sbo(uint32_t saddr, uint32_t sbo_bytes) {
    constexpr uint64_t layout = (KC == 64) ? 2ull /*SWIZZLE_128B*/ : 4ull /*SWIZZLE_64B*/;
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= layout << 61;
    return d;
}

template <int BN, int KC, typename DT, bool FUSE>
__global__ void __launch_bounds__(FUSE ? 448 : 320, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ConvHaloParams p) {
    constexpr int kRowB = KC * 2;                      // bytes per pixel row of the operand
    constexpr int kHaloBytes = kHaloPix * kRowB;       // bytes TMA delivers per halo
    constexpr int kStageBytes = halo_stage_bytes<KC>();
    constexpr int kWBlock = BN * kRowB;                // one (tap, cin chunk) weight block
    constexpr int kTmemCols = TcCfg<BN>::kTmemCols;
    constexpr int kMaxStages = 8;

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
    float *s_bias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(full_bar) + 256);     // [BN], 16-byte aligned

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nt = blockIdx.y;                         // output-channel block of this CTA

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < p.n_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); mbar_init(&ready_bar[s], 128); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], BN > 32 ? 256 : 128); }
        mbar_init(w_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp >= 2 && warp < 6)
        for (int i = threadIdx.x - 64; i < BN; i += 128) s_bias[i] = p.bias ? p.bias[blockIdx.y * BN + i] : 0.f;
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int total_tiles = tiles_per_img * p.N;
    // static CONTIGUOUS partition of the tile sequence (x fastest, then y, then image) over the CTAs of this output-channel
    // block: neighbouring tiles share halo columns through L2, and a CTA touches at most two or three images, so the
    // InstanceNorm atomics of one (image, channel) come from ~1/10 of the CTAs and at different times (a round-robin
    // partition made all CTAs hit the same 64 addresses at every image boundary: ~5 us of serialised f64 atomics each).
    const int t_begin = (int)((long long)total_tiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)total_tiles * (blockIdx.x + 1) / gridDim.x);

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            mbar_expect_tx(w_bar, (uint32_t)w_bytes);
            for (int tap = 0; tap < 9; ++tap)
                for (int cc = 0; cc < p.cin_chunks; ++cc)
                    tma_load_3d(smem_w + (cc * 9 + tap) * kWBlock, &tmB, w_bar, cc * KC, nt * BN, tap);
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = t_begin; tile < t_end; ++tile) {
                const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img + p.n_base;
                for (int cc = 0; cc < p.cin_chunks; ++cc) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], kHaloBytes);
                    if (p.dbg & 8) tma_load_4d(smem_a + stage * kStageBytes, &tmA, &full_bar[stage], cc * KC, -1, -1, p.n_base);
                    else tma_load_4d(smem_a + stage * kStageBytes, &tmA, &full_bar[stage], cc * KC, tx * 8 - 1, ty * 16 - 1, n);
                    if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
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
            unsigned long long *o = p.dbg_out + (size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 4;
            o[0] = (unsigned long long)(clock64() - t_start); o[1] = (unsigned long long)w_halo; o[2] = (unsigned long long)w_tmem;
        }
    } else if (warp < 10) {
        // ===================== epilogue: two groups of 4 warps (warps 2..5, warps 6..9), 32 accumulator columns per warp =====================
        // One warp per TMEM lane quarter is latency bound (~900 cycles per 32 columns), so the work is spread over 8 warps:
        //   BN == 32: the groups alternate tiles (group g owns TMEM buffer g), the epilogue of tile i overlaps that of tile i-1;
        //   BN == 64: both groups work on every tile, each on one half of the columns (register budget: 2*32 statistics
        //             accumulators per thread either way; 10 warps are allocated as 12, i.e. 168 registers per thread).
        constexpr bool kSplit = BN > 32;
        static_assert(BN == 32 || BN == 64, "conv_halo_kernel: BN must be 32 or 64");
        const int grp = (warp - 2) >> 2;
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;
        const int ix = row & 7, iy = row >> 3;
        const int col0 = kSplit ? grp * 32 : 0;
        DT *out = reinterpret_cast<DT *>(p.out);
        const bool do_stats = p.stats != nullptr && !(p.dbg & 1);
        // InstanceNorm statistics: every thread keeps running sums of ITS pixel row for its 32 channels across the tiles of
        // one image (no cross-lane traffic per tile); the 32-lane transpose-reduction and the double atomics happen only
        // when the image changes (a CTA walks the images in order: <= N flushes).
        float tacc1[32], tacc2[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) { tacc1[i] = 0.f; tacc2[i] = 0.f; }
        auto flush_stats = [&](int n_img) {
            float b[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) b[j] = tacc2[j];
            const float s1 = warp_transpose_sum32(tacc1, lane), s2 = warp_transpose_sum32(b, lane);
#pragma unroll
            for (int j = 0; j < 32; ++j) { tacc1[j] = 0.f; tacc2[j] = 0.f; }
            double *s = p.stats + ((long long)n_img * p.Cout + nt * BN + col0 + lane) * 2;
            atomicAdd(s, (double)s1);
            atomicAdd(s + 1, (double)s2);
        };
        float bias_r[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) bias_r[j] = s_bias[col0 + j];
        int cur_n = -1;
        long long w_epi = 0;
        // tile walker: coordinates advance with carries, no divisions in the loop
        const int step = kSplit ? 1 : 2;
        const int sx = step % p.tiles_x, sy = (step / p.tiles_x) % p.tiles_y, sn = step / tiles_per_img;
        int tile = t_begin + (kSplit ? 0 : grp);
        int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img + p.n_base;
        for (uint32_t k = 0; tile < t_end; tile += step, ++k) {
            const int buf = kSplit ? (int)(k & 1) : grp;
            const uint32_t par = kSplit ? ((k >> 1) & 1) : (k & 1);
            const int x = tx * 8 + ix, y = ty * 16 + iy;
            const bool valid = (x < p.W) && (y < p.H) && !(p.dbg & 2);
            if (do_stats && n != cur_n) {
                if (cur_n >= 0) flush_stats(cur_n);
                cur_n = n;
            }
            const long long tw0 = p.dbg_out ? clock64() : 0;
            mbar_wait(&tmem_full[buf], par);
            if (p.dbg_out) w_epi += clock64() - tw0;
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + col0);
            DT *dst = out + (((long long)n * p.H + y) * p.W + x) * p.out_ps + nt * BN + col0;
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr, r);
            tmem_ld_wait();
            tcgen05_fence_before();
            mbar_arrive(&tmem_empty[buf]);        // this thread's TMEM reads of the tile are done
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = pack2<DT>(__uint_as_float(r[2 * j]) + bias_r[2 * j], __uint_as_float(r[2 * j + 1]) + bias_r[2 * j + 1]);
            if (valid) {
                // 2 x 32-byte stores (STG.256): every sector is written whole by one instruction
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * j), "r"(pk[8 * j]), "r"(pk[8 * j + 1]),
                                 "r"(pk[8 * j + 2]), "r"(pk[8 * j + 3]), "r"(pk[8 * j + 4]), "r"(pk[8 * j + 5]), "r"(pk[8 * j + 6]), "r"(pk[8 * j + 7]) : "memory");
                if (do_stats) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float2 f = unpack2<DT>(pk[j]);          // statistics of the ROUNDED values (what the consumer normalises)
                        tacc1[2 * j] += f.x;
                        tacc1[2 * j + 1] += f.y;
                        tacc2[2 * j] = fmaf(f.x, f.x, tacc2[2 * j]);
                        tacc2[2 * j + 1] = fmaf(f.y, f.y, tacc2[2 * j + 1]);
                    }
                }
            }
            tx += sx; if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
            ty += sy; if (ty >= p.tiles_y) { ty -= p.tiles_y; ++n; }
            n += sn;
        }
        if (do_stats && cur_n >= 0) flush_stats(cur_n);
        if (p.dbg_out && threadIdx.x == 64) p.dbg_out[(size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 4 + 3] = (unsigned long long)w_epi;
    } else if (FUSE) {
        // ===================== halo transform (warps 6..9): producer's InstanceNorm-apply + LeakyReLU =====================
        // thread t handles 16-byte chunks (8 channels) of the halo: chunk g -> pixel g / (KC/8), channel block g % (KC/8)
        const int t = threadIdx.x - 320;
        constexpr int kChunksPerPix = KC / 8;                 // 16-byte chunks (8 channels) per halo pixel
        constexpr int kPixStep = 128 / kChunksPerPix;         // 128 % kChunksPerPix == 0: a thread keeps ONE channel block
        constexpr uint32_t kXorMask = (KC == 64) ? 7u : 3u;
        const int cb = t % kChunksPerPix, pix0 = t / kChunksPerPix;
        float scs[8], shs[8], sls[8];
        int cur_n = -1, cur_cc = -1;
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = t_begin; tile < t_end; ++tile) {
            const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img + p.n_base;
            for (int cc = 0; cc < p.cin_chunks; ++cc) {
                if (n != cur_n || cc != cur_cc) {
                    const long long o = (long long)n * p.Cin + cc * KC + cb * 8;
#pragma unroll
                    for (int j = 0; j < 8; ++j) { scs[j] = __ldg(p.in_scale + o + j); shs[j] = __ldg(p.in_shift + o + j); sls[j] = __ldg(p.in_slope + cc * KC + cb * 8 + j); }
                    cur_n = n; cur_cc = cc;
                }
                mbar_wait(&full_bar[stage], phase);
                uint8_t *base = smem_a + stage * kStageBytes;
                const uint32_t base_u32 = smem_u32(base);
                for (int pix = pix0; pix < kHaloPix; pix += kPixStep) {
                    const int hx = pix % kHaloW, hy = pix / kHaloW;
                    const int gx = tx * 8 - 1 + hx, gy = ty * 16 - 1 + hy;
                    // outside the image TMA wrote zeros = the conv padding of the NORMALISED tensor: leave them
                    if (gx < 0 || gx >= p.W || gy < 0 || gy >= p.H) continue;
                    // absolute-address swizzle: 16-byte chunk index XOR ((byte address >> 7) & mask)
                    const uint32_t row_off = (uint32_t)(pix * kRowB);
                    const uint32_t sw = ((base_u32 + row_off) >> 7) & kXorMask;
                    uint4 *ptr = reinterpret_cast<uint4 *>(base + row_off + (((uint32_t)cb ^ sw) << 4));
                    uint4 v = *ptr;
                    uint32_t *w = reinterpret_cast<uint32_t *>(&v);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack2<DT>(w[j]);
                        float y0 = fmaf(f.x, scs[2 * j], shs[2 * j]), y1 = fmaf(f.y, scs[2 * j + 1], shs[2 * j + 1]);
                        y0 = fmaxf(y0, y0 * sls[2 * j]);              // LeakyReLU for 0 <= slope <= 1 (slope 1 = identity)
                        y1 = fmaxf(y1, y1 * sls[2 * j + 1]);
                        w[j] = pack2<DT>(y0, y1);
                    }
                    *ptr = v;
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy writes -> visible to the MMA (async proxy)
                mbar_arrive(&ready_bar[stage]);
                if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

}  // namespace ts2d
