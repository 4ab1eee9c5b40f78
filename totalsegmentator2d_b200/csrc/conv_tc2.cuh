// K4b: conv_tc_kernel with TWO 128-pixel tiles per weight block (M = 256 per B operand), for the 3x3 convolutions of the
// 32^2 - 128^2 levels with 128 / 256 / 512 output channels (stride 1 and stride 2).
//
// Why (clock stamps in profiles/r01g_tiny_levels_clock_stamps.txt + the row-rate probe): the TMA unit of an SM issues one request per
// 128-byte box row at ~1.7 cycles each.  conv_tc_kernel moves 128 (A) + BN (B) rows per k-chunk for KC/16 MMAs: at BN = 128 that is
// 435 cycles of fill for 268 cycles of tensor work, at BN = 256 653 for 524 - the tensor pipe waits for operands.  Sharing every weight
// block between two pixel tiles halves the B rows per MMA:
//     BN = 256:  256 (A0, A1) + 256 (B) rows = 870 cycles of fill for 8 x 131 = 1048 cycles of MMA   -> tensor bound
//     BN = 128:  256 + 128 rows = 653 cycles for 8 x 67 = 536                                          -> 0.82 of tensor bound
// and a CTA owns 2 tiles at once, which also removes the 1.3-wave tail of the 32^2 level (192 tiles on 148 CTAs -> 96 pairs).
// TMEM: 2 tiles x BN fp32 columns; double buffered at BN = 128 (4 x 128 = 512 columns), single buffered at BN = 256 (the epilogue
// of a pair, ~3 k cycles, is exposed once per >= 36 k-chunks x 1048 cycles).
//
// Same GEMM view, operand layouts, descriptors, epilogue arithmetic (bias, rounding, fp32 InstanceNorm statistics) as conv_tc_kernel;
// restrictions (host-checked): tiles lie inside one image (tn == 1), an even number of pixel tiles, no split-K, no transposed conv.
#pragma once
#include "conv_tc.cuh"

namespace ts2d {

template <int BN, int KC> constexpr int tc2_stage_bytes() { return (256 + BN) * KC * 2; }
template <int BN, int KC> constexpr int tc2_num_stages() { return (200 * 1024 / tc2_stage_bytes<BN, KC>()) > 8 ? 8 : (200 * 1024 / tc2_stage_bytes<BN, KC>()); }
template <int BN, int KC> constexpr int tc2_smem_bytes() { return tc2_num_stages<BN, KC>() * tc2_stage_bytes<BN, KC>() + 1024 + 256; }

template <int BN, int KC, typename DT>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB, const ConvTcParams p) {
    static_assert(BN == 128 || BN == 256, "conv_tc2_kernel: BN must be 128 or 256");
    constexpr int kStages = tc2_num_stages<BN, KC>();
    constexpr int kABytes = 128 * KC * 2;                    // one pixel tile
    constexpr int kStageBytes = tc2_stage_bytes<BN, KC>();
    constexpr int kBufs = (4 * BN <= 512) ? 2 : 1;           // accumulator buffers of 2 x BN columns
    constexpr int kTmemCols = 512;
    constexpr int kMyChunks = BN / 64;                       // 32-column chunks per epilogue thread and tile (the two groups split the columns)

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes);
    uint64_t *empty_bar = full_bar + kStages;
    uint64_t *tmem_full = empty_bar + kStages;     // [kBufs]
    uint64_t *tmem_empty = tmem_full + 2;          // [kBufs]
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA2) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], 256); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    griddep_launch_dependents();
    griddep_wait();

    const int m_tiles = p.tiles_n * p.tiles_h * p.tiles_w;       // even (host-checked)
    const int mpairs = m_tiles >> 1;
    const int total_units = mpairs * p.n_tiles_n;                // unit = (channel block, pair of pixel tiles), channel block slowest
    const int k_chunks = p.taps * p.cin_chunks;
    const int u_begin = (int)((long long)total_units * blockIdx.x / gridDim.x);
    const int u_end = (int)((long long)total_units * (blockIdx.x + 1) / gridDim.x);

    if (warp == 0) {
        // ===================== TMA producer (whole warp convergent, one elected lane issues; no divisions in the k loop) =====================
        const uint32_t leader = elect_one();
        const uint32_t s0 = smem_u32(smem);
        int stage = 0;
        uint32_t phase = 0;
        for (int u = u_begin; u < u_end; ++u) {
            const int nt = u / mpairs, pi = u - nt * mpairs;
            int x0[2], y0[2], n0[2];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int mt = 2 * pi + t;
                x0[t] = (mt % p.tiles_w) * p.tw; y0[t] = ((mt / p.tiles_w) % p.tiles_h) * p.th; n0[t] = mt / (p.tiles_w * p.tiles_h);
            }
            int r = 0, s = 0;
            for (int tap = 0; tap < p.taps; ++tap) {
                const int dr = r - p.pad_h, ds = s - p.pad_w;
                const int py = (p.stride2 & 2) ? (dr & 1) : 0, px = (p.stride2 & 1) ? (ds & 1) : 0;
                const int hoff = (p.stride2 & 2) ? (dr - py) / 2 : dr, woff = (p.stride2 & 1) ? (ds - px) / 2 : ds;
                for (int cc = 0; cc < p.cin_chunks; ++cc) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    const uint32_t sa = s0 + (uint32_t)(stage * kStageBytes);
                    mbar_expect_tx_pred(leader, &full_bar[stage], kStageBytes);
                    const bool second = cc >= p.split_chunks;                       // K-chunks of the second input tensor (decoder: the skip)
                    const CUtensorMap *tm = second ? &tmA2 : &tmA;
                    const int c0 = (second ? cc - p.split_chunks : cc) * KC;
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        if (!p.stride2) tma_load_4d_pred(leader, sa + t * kABytes, tm, &full_bar[stage], c0, x0[t] + woff, y0[t] + hoff, n0[t]);
                        else tma_load_5d_pred(leader, sa + t * kABytes, &tmA, &full_bar[stage], px * p.in_ps + cc * KC, x0[t] + woff, py, y0[t] + hoff, n0[t]);
                    }
                    tma_load_3d_pred(leader, sa + 2 * kABytes, &tmB, &full_bar[stage], cc * KC, nt * BN, tap);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                if (++s == p.kw) { s = 0; ++r; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp convergent, one elected lane issues) =====================
        const uint32_t leader = elect_one();
        const uint32_t idesc = make_idesc<DT>(BN);
        constexpr uint32_t hi = desc_hi<KC>(KC == 64 ? 1024 : 512);
        const uint32_t s0_lo = desc_lo<KC>(smem_u32(smem));
        int stage = 0;
        uint32_t phase = 0;
        int it = 0;
        for (int u = u_begin; u < u_end; ++u, ++it) {
            const int buf = kBufs == 2 ? (it & 1) : 0;
            const uint32_t use = kBufs == 2 ? (uint32_t)(it >> 1) : (uint32_t)it;
            mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
            tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(buf * 2 * BN);
            for (int kc = 0; kc < k_chunks; ++kc) {
                mbar_wait(&full_bar[stage], phase);
                tcgen05_fence_after();
                const uint32_t a_lo = s0_lo + (uint32_t)(stage * (kStageBytes >> 4));
                const uint32_t b_lo = a_lo + (uint32_t)((2 * kABytes) >> 4);
#pragma unroll
                for (int t = 0; t < 2; ++t)
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k)
                        tcgen05_mma_f16_lohi(leader, d_tmem + (uint32_t)(t * BN), a_lo + (uint32_t)(t * (kABytes >> 4) + k * 2), hi, b_lo + (uint32_t)(k * 2), hi,
                                             idesc, k ? 1u : (uint32_t)(kc != 0));
                tcgen05_commit_pred(leader, &empty_bar[stage]);
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            tcgen05_commit_pred(leader, &tmem_full[buf]);
        }
    } else if (warp >= 4) {
        // ===================== epilogue (warps 4..11): group g takes column half g of both tiles =====================
        const int grp = (warp - 4) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int ix = row % p.tw, iy = row / p.tw;                  // tn == 1: the 128 rows are th x tw pixels of one image
        const int chunk0 = grp * kMyChunks;
        DT *out = reinterpret_cast<DT *>(p.out);
        const bool do_stats = p.stats != nullptr;
        float racc1[kMyChunks], racc2[kMyChunks];                    // per-lane (= per-channel) InstanceNorm sums of the current (image, channel block)
#pragma unroll
        for (int i = 0; i < kMyChunks; ++i) { racc1[i] = 0.f; racc2[i] = 0.f; }
        int cur_n = -1, cur_nt = -1;
        auto flush_stats = [&]() {
#pragma unroll
            for (int c = 0; c < kMyChunks; ++c) {
                double *s = p.stats + ((long long)cur_n * p.Cout + cur_nt * BN + (chunk0 + c) * 32 + lane) * 2;
                atomicAdd(s, (double)racc1[c]);
                atomicAdd(s + 1, (double)racc2[c]);
                racc1[c] = 0.f; racc2[c] = 0.f;
            }
        };
        int it = 0;
        for (int u = u_begin; u < u_end; ++u, ++it) {
            const int nt = u / mpairs, pi = u - nt * mpairs;
            const int buf = kBufs == 2 ? (it & 1) : 0;
            const uint32_t par = (kBufs == 2 ? (uint32_t)(it >> 1) : (uint32_t)it) & 1u;
            mbar_wait(&tmem_full[buf], par);
            tcgen05_fence_after();
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int mt = 2 * pi + t;
                const int x = (mt % p.tiles_w) * p.tw + ix, y = ((mt / p.tiles_w) % p.tiles_h) * p.th + iy, n = mt / (p.tiles_w * p.tiles_h);
                const bool valid = (x < p.Wout) && (y < p.Hout) && (n < p.N);
                if (do_stats && (n != cur_n || nt != cur_nt)) {
                    if (cur_n >= 0) flush_stats();
                    cur_n = n; cur_nt = nt;
                }
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * 2 * BN + t * BN);
                DT *opix = out + (((long long)n * p.Hout + y) * p.Wout + x) * p.out_ps + nt * BN;
#pragma unroll
                for (int c = 0; c < kMyChunks; ++c) {
                    const int chunk = chunk0 + c;
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(taddr + chunk * 32, r);
                    tmem_ld_wait();
                    if (t == 1 && c == kMyChunks - 1) {
                        tcgen05_fence_before();
                        mbar_arrive(&tmem_empty[buf]);       // this thread's TMEM reads of the pair are done
                    }
                    uint32_t pk[16];
                    uint64_t v[16];
                    epi_vals<DT>(r, p.bias ? p.bias + nt * BN + chunk * 32 : nullptr, pk, v);
                    if (valid) {
#pragma unroll
                        for (int k = 0; k < 2; ++k)
                            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(opix + chunk * 32 + 16 * k), "r"(pk[8 * k]), "r"(pk[8 * k + 1]),
                                         "r"(pk[8 * k + 2]), "r"(pk[8 * k + 3]), "r"(pk[8 * k + 4]), "r"(pk[8 * k + 5]), "r"(pk[8 * k + 6]), "r"(pk[8 * k + 7]) : "memory");
                    }
                    if (do_stats) {
                        float a[32], b[32];
#pragma unroll
                        for (int k = 0; k < 16; ++k) {
                            a[2 * k] = valid ? f32x2_lo(v[k]) : 0.f;
                            a[2 * k + 1] = valid ? f32x2_hi(v[k]) : 0.f;
                        }
#pragma unroll
                        for (int k = 0; k < 32; ++k) b[k] = a[k] * a[k];
                        racc1[c] += warp_transpose_sum32(a, lane);
                        racc2[c] += warp_transpose_sum32(b, lane);
                    }
                }
            }
        }
        if (do_stats && cur_n >= 0) flush_stats();
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

}  // namespace ts2d
