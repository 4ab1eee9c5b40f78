// K4s: 3x3 / STRIDE-2 implicit-GEMM convolution with parity-plane halos and resident weights (tcgen05 + TMEM, fed by TMA):
// the first strided convolution of the encoder (enc1.0 of the CS-512 plan: 512^2 x 32 -> 256^2 x 64).
//
// Why: conv_tc_kernel loads one 128-pixel box per filter tap.  On this layer a box row is 64 bytes (32 channels), the TMA unit issues
// one request per box row (~1.7 cycles each), and 9 taps x (128 + 64) rows = 2900 cycles of fill stand against ~1000 cycles of MMAs:
// 0.154 ms where the tensor pipe needs 0.025 and HBM 0.046 (profiles/r02q_layers.txt).  Here the input of a tile is loaded ONCE:
//
//   tile     = 8 (x) x 16 (y) OUTPUT pixels = the 128 rows of one UMMA (row i -> x = i & 7, y = i >> 3)
//   input    = rows 2y-1 .. 2y+1, columns 2x-1 .. 2x+1: in the parity view of the NHWC tensor (w = 2*w2 + px, h = 2*h2 + py, the same
//              5-D tensor map conv_tc_kernel uses) that is four PLANES (py, px), each a dense 9 (w2) x 17 (h2) pixel box starting at
//              (w2, h2) = (8*tx - 1, 16*ty - 1); TMA zero-fills what lies outside the image = the conv padding
//   tap(r,s) = plane (py, px) = ((r-1) & 1, (s-1) & 1) viewed through a descriptor whose start address is advanced by
//              ((1 + hoff) * 9 + (1 + woff)) pixel rows (hoff = -1 for r = 0, woff = -1 for s = 0, else 0) and whose 8-row groups are
//              9 pixels apart - the same shifted-descriptor trick as conv_halo_kernel (the swizzle XOR is a function of the absolute
//              shared-memory address; every plane starts 1024-byte aligned)
//   weights  = all 9 taps x Cin x BN of the CTA's output-channel block stay resident in shared memory.
// 4 x 153 = 612 box rows per tile and K-chunk instead of 9 x 128 = 1152, and each input pixel crosses L2 -> SM once instead of 2.25 times.
//
// Warp roles (384 threads) as conv_halo_kernel: warps 0-7 epilogue (two warpgroups; BN == 32: alternate tiles, BN == 64: column halves),
// warp 10 TMA producer, warp 11 TMEM allocator + MMA issuer.  Epilogue arithmetic = conv_halo_kernel's (bias, rounding, fp32 InstanceNorm
// statistics in per-thread accumulators, 32-byte stores).
#pragma once
#include "conv_halo.cuh"

namespace ts2d {

constexpr int kS2W = 9, kS2H = 17, kS2Pix = kS2W * kS2H;       // one parity plane of the input halo
template <int KC> constexpr int s2_plane_bytes() { return ((kS2Pix * KC * 2 + 1023) / 1024) * 1024; }
template <int KC> constexpr int s2_stage_bytes() { return 4 * s2_plane_bytes<KC>(); }

template <int BN, int KC, typename DT>
__global__ void __launch_bounds__(384, 1)
conv_halo_s2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ConvHaloParams p) {
    static_assert(BN == 32 || BN == 64, "conv_halo_s2_kernel: BN must be 32 or 64");
    constexpr int kRowB = KC * 2;
    constexpr int kPlaneBytes = s2_plane_bytes<KC>();
    constexpr int kStageBytes = s2_stage_bytes<KC>();
    constexpr int kTmaBytes = 4 * kS2Pix * kRowB;      // bytes TMA delivers per stage
    constexpr int kWBlock = BN * kRowB;
    constexpr int kTmemCols = TcCfg<BN>::kTmemCols;
    constexpr int kMaxStages = 8;
    constexpr bool kSplit = BN > 32;
    constexpr int kWarpTma = 10, kWarpMma = 11;

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int w_bytes = 9 * p.cin_chunks * kWBlock;
    uint8_t *smem_w = smem;
    uint8_t *smem_a = smem + ((w_bytes + 1023) & ~1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem_a + p.n_stages * kStageBytes);
    uint64_t *empty_bar = full_bar + kMaxStages;
    uint64_t *tmem_full = empty_bar + kMaxStages;     // [2]
    uint64_t *tmem_empty = tmem_full + 2;             // [2]
    uint64_t *w_bar = tmem_empty + 2;
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(w_bar + 1);
    float *s_bias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(full_bar) + 512);     // [64]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nt = blockIdx.y;                         // output-channel block of this CTA

    if (warp == kWarpTma && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < p.n_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], kSplit ? 256 : 128); }
        mbar_init(w_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kWarpMma) {
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
    if (!(warp == kWarpTma && lane == 0)) griddep_wait();      // (the producer thread first issues the weight loads: constants)

    // p.H / p.W = OUTPUT grid (input = 2H x 2W); static contiguous partition of the tile sequence as in conv_halo_kernel
    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int total_tiles = tiles_per_img * p.N;
    const int t_begin = (int)((long long)total_tiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)total_tiles * (blockIdx.x + 1) / gridDim.x);

    if (warp < 8) {
        // ===================== epilogue warps =====================
        const int wt = threadIdx.x;
        const int grp = wt >> 7;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int ix = row & 7, iy = row >> 3;
        const int col0 = kSplit ? grp * 32 : 0;
        DT *out = reinterpret_cast<DT *>(p.out);
        const bool do_stats = p.stats != nullptr;
        uint64_t tacc1[16], tacc2[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { tacc1[i] = 0ull; tacc2[i] = 0ull; }
        auto flush_stats = [&](int n_img) {
            float s1, s2;
            epi_flush(tacc1, tacc2, lane, s1, s2);
            double *s = p.stats + ((long long)n_img * p.Cout + nt * BN + col0 + lane) * 2;
            atomicAdd(s, (double)s1);
            atomicAdd(s + 1, (double)s2);
        };
        int cur_n = -1;
        const int estep = kSplit ? 1 : 2;
        const int esx = estep % p.tiles_x, esy = (estep / p.tiles_x) % p.tiles_y, esn = estep / tiles_per_img;
        int e_tile = t_begin + (kSplit ? 0 : grp);
        int tx = e_tile % p.tiles_x, ty = (e_tile / p.tiles_x) % p.tiles_y, n = e_tile / tiles_per_img + p.n_base;
        const int n_local = t_end - t_begin;
        for (int e = kSplit ? 0 : grp; e < n_local; e += estep) {
            const int buf = e & 1;
            const uint32_t par = (uint32_t)(e >> 1) & 1u;
            const int x = tx * 8 + ix, y = ty * 16 + iy;
            const bool valid = (x < p.W) && (y < p.H);
            if (do_stats && n != cur_n) {
                if (cur_n >= 0) flush_stats(cur_n);
                cur_n = n;
            }
            mbar_wait_parked(&tmem_full[buf], par);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BN + col0);
            DT *dst = out + (((long long)n * p.H + y) * p.W + x) * p.out_ps + nt * BN + col0;
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr, r);
            tmem_ld_wait();
            tcgen05_fence_before();
            mbar_arrive(&tmem_empty[buf]);
            uint32_t pk[16];
            epi_chunk<DT>(r, s_bias + col0, pk, tacc1, tacc2, do_stats && valid);
            if (valid) {
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * k), "r"(pk[8 * k]), "r"(pk[8 * k + 1]),
                                 "r"(pk[8 * k + 2]), "r"(pk[8 * k + 3]), "r"(pk[8 * k + 4]), "r"(pk[8 * k + 5]), "r"(pk[8 * k + 6]), "r"(pk[8 * k + 7]) : "memory");
            }
            tx += esx; if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
            ty += esy; if (ty >= p.tiles_y) { ty -= p.tiles_y; ++n; }
            n += esn;
        }
        if (do_stats && cur_n >= 0) flush_stats(cur_n);
    } else if (warp == kWarpTma) {
        // ===================== TMA producer: four parity planes per (tile, K-chunk) =====================
        if (lane == 0) {
            mbar_expect_tx(w_bar, (uint32_t)w_bytes);
            for (int cc = 0; cc < p.cin_chunks; ++cc)
                for (int tap = 0; tap < 9; ++tap)
                    tma_load_3d(smem_w + (cc * 9 + tap) * kWBlock, &tmB, w_bar, cc * KC, nt * BN, tap);
            griddep_wait();
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = t_begin; tile < t_end; ++tile) {
                const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img + p.n_base;
                for (int cc = 0; cc < p.cin_chunks; ++cc) {
                    mbar_wait_parked(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], kTmaBytes);
                    uint8_t *sa = smem_a + stage * kStageBytes;
#pragma unroll
                    for (int pl = 0; pl < 4; ++pl) {
                        const int py = pl >> 1, px = pl & 1;
                        // parity view coordinates: (px * pixel stride + channel, w2, py, h2, n)
                        tma_load_5d(sa + pl * kPlaneBytes, &tmA, &full_bar[stage], px * p.in_ps + cc * KC, tx * 8 - 1, py, ty * 16 - 1, n);
                    }
                    if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == kWarpMma) {
        // ===================== MMA issuer (whole warp convergent, one elected lane issues; see conv_tc.cuh) =====================
        const uint32_t leader = elect_one();
        const uint32_t idesc = make_idesc<DT>(BN);
        constexpr uint32_t a_hi = desc_hi<KC>(kS2W * kRowB);          // 8-row groups of the A view are one plane row (9 pixels) apart
        constexpr uint32_t b_hi = desc_hi<KC>(KC == 64 ? 1024 : 512);
        const uint32_t w_lo = desc_lo<KC>(smem_u32(smem_w));
        const uint32_t a0_lo = desc_lo<KC>(smem_u32(smem_a));
        mbar_wait(w_bar, 0);
        tcgen05_fence_after();
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int tile = t_begin; tile < t_end; ++tile, ++it) {
            const int buf = it & 1;
            const uint32_t use = (uint32_t)(it >> 1);
            mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
            tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
            for (int cc = 0; cc < p.cin_chunks; ++cc) {
                mbar_wait(&full_bar[stage], phase);
                tcgen05_fence_after();
                const uint32_t a_lo = a0_lo + (uint32_t)(stage * (kStageBytes >> 4));
                const uint32_t b_lo = w_lo + (uint32_t)(cc * 9 * (kWBlock >> 4));
#pragma unroll
                for (int tap = 0; tap < 9; ++tap) {
                    const int r = tap / 3, s = tap % 3;
                    const int dr = r - 1, ds = s - 1;
                    const int py = dr & 1, px = ds & 1;                       // input row 2y + dr = 2*(y + hoff) + py
                    const int hoff = (dr - py) / 2, woff = (ds - px) / 2;     // -1 or 0
                    const int plane_off = (py * 2 + px) * kPlaneBytes + ((1 + hoff) * kS2W + (1 + woff)) * kRowB;
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)
                        tcgen05_mma_f16_lohi(leader, d_tmem, a_lo + (uint32_t)((plane_off >> 4) + k * 2), a_hi,
                                             b_lo + (uint32_t)(((tap * kWBlock) >> 4) + k * 2), b_hi, idesc, (tap | k) ? 1u : (uint32_t)(cc != 0));
                }
                tcgen05_commit_pred(leader, &empty_bar[stage]);
                if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
            }
            tcgen05_commit_pred(leader, &tmem_full[buf]);
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == kWarpMma) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

}  // namespace ts2d
