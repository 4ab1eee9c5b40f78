// K4f: first U-Net conv (3x3, stride 1, Cin = 2) on the tensor cores.
//
// With two input channels the layer has K = 9 taps x 2 = 18: far too thin for a TMA-fed implicit GEMM (a 16-channel
// pad would read 8x the input), and on the CUDA cores it is FMA-bound (conv_first_kernel: 0.20 ms for 12 tiles of
// 512^2 against 0.03 ms for its HBM bytes).  Here the worker threads build the im2col operand themselves:
//
//   tile   = 8 (x) x 16 (y) output pixels = the 128 rows of one UMMA (row i -> x = i & 7, y = i >> 3, as in conv_halo)
//   A row  = the 3 x 3 x 2 neighbourhood of the pixel as K = r*6 + s*2 + c (the three x-neighbours of a row are 12
//            contiguous bytes of the NHWC input), zero-padded to K = 32: one 64-byte K-major row, written with the
//            64-byte swizzle (16-byte chunk index ^ ((row >> 1) & 3)) that TMA would have produced
//   B      = weights [32 couts][K = 32] built once per CTA from the fp32 [tap][Cin][Cout] copy, same layout
//   MMA    = 2 x tcgen05.mma M128 N32 K16 per tile, fp32 accumulators in TMEM
//
// Two worker warpgroups per CTA alternate tiles; each runs build -> bar.sync -> (elected lane) MMA + commit -> mbarrier
// wait -> tcgen05.ld (bias already inside: two extra K columns) -> round -> InstanceNorm statistics -> 2 x 32-byte stores, with the 9 input words of its
// NEXT tile already in flight.  Several CTAs per SM hide the chain's latency; the kernel is bound by the output write.
#pragma once
#include "conv_halo.cuh"

namespace ts2d {

struct FirstTcParams {
    const void *in; int N, H, W;       // dense NHWC input [N][H][W][2], 16-bit
    const float *w;                    // [tap][2][Cout] fp32
    const float *bias;                 // [Cout] or null
    void *out; int Cout;               // dense NHWC output [N][H][W][Cout]
    double *stats;                     // [N][Cout][2] or null
    int tiles_x, tiles_y;
};
constexpr int kFtcThreads = 256;
constexpr int kFtcSmemBytes = 1024 /*align*/ + 2 * 8192 /*A, one stage per group*/ + 2048 /*B*/ + 256 /*barriers, tmem ptr, bias*/;

template <typename DT>
__global__ void __launch_bounds__(kFtcThreads, 2) conv_first_tc_kernel(const FirstTcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *s_a = smem;                       // [2][128 rows][64 B]
    uint8_t *s_b = smem + 2 * 8192;            // [32 rows][64 B]
    uint64_t *mma_bar = reinterpret_cast<uint64_t *>(s_b + 2048);    // [2]
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(mma_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = warp >> 2, q = warp & 3;   // worker warpgroup, TMEM lane quarter
    const int nt = blockIdx.y;                 // block of 32 output channels

    if (threadIdx.x == 0) {
        mbar_init(&mma_bar[0], 1);
        mbar_init(&mma_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // B[n][k]: k = tap*2 + c < 18 from the fp32 weights; k = 18, 19 carry the bias as a rounded value plus its rounding
    // residual (the matching A columns are 1.0), so the accumulators leave the tensor core with the bias already added
    for (int i = threadIdx.x; i < 32 * 32; i += kFtcThreads) {
        const int n = i >> 5, k = i & 31;
        float v = 0.f;
        if (k < 18) v = p.w[(long long)k * p.Cout + nt * 32 + n];
        else if (k < 20 && p.bias) {
            const float b = p.bias[nt * 32 + n];
            const float hi = to_f32<DT>(from_f32<DT>(b));
            v = k == 18 ? hi : b - hi;
        }
        const uint32_t off = (uint32_t)(n * 64 + ((((k >> 3) ^ ((n >> 1) & 3))) << 4) + (k & 7) * 2);
        *reinterpret_cast<DT *>(s_b + off) = from_f32<DT>(v);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    griddep_launch_dependents();
    griddep_wait();          // the weights / bias above are constants; the tiles and the output buffer belong to earlier launches

    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int total_tiles = tiles_per_img * p.N;
    // contiguous partition of the tile sequence over the CTAs (same reasons as conv_halo_kernel); the two groups alternate
    const int t_begin = (int)((long long)total_tiles * blockIdx.x / gridDim.x);
    const int t_end = (int)((long long)total_tiles * (blockIdx.x + 1) / gridDim.x);

    const int row = q * 32 + lane;
    const int ix = row & 7, iy = row >> 3;
    uint8_t *a_row = s_a + grp * 8192 + row * 64;
    const uint32_t sw = (uint32_t)((row >> 1) & 3);
    const uint32_t idesc = make_idesc<DT>(32);
    constexpr uint32_t ab_hi = desc_hi<32>(512);
    const uint32_t a_lo = desc_lo<32>(smem_u32(s_a + grp * 8192));
    const uint32_t b_lo = desc_lo<32>(smem_u32(s_b));
    const uint32_t d_tmem = tmem_base + (uint32_t)(grp * 32);
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(grp * 32);
    const uint32_t kOnes = pack2<DT>(1.f, 1.f);
    const uint32_t *in = reinterpret_cast<const uint32_t *>(p.in);      // one 32-bit word = the 2 channels of a pixel
    DT *out = reinterpret_cast<DT *>(p.out);
    const bool do_stats = p.stats != nullptr;

    uint64_t tacc1[16], tacc2[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { tacc1[i] = 0ull; tacc2[i] = 0ull; }
    auto flush_stats = [&](int n_img) {
        float s1, s2;
        epi_flush(tacc1, tacc2, lane, s1, s2);
        double *s = p.stats + ((long long)n_img * p.Cout + nt * 32 + lane) * 2;
        atomicAdd(s, (double)s1);
        atomicAdd(s + 1, (double)s2);
    };
    // the 3x3 neighbourhood of this thread's pixel in tile (tx, ty) of image n (zeros outside the image = the conv padding);
    // tiles that do not touch the image border (warp-uniform test) take 9 unconditional loads off one base pointer
    auto gather = [&](int tx, int ty, int n, uint32_t (&w)[9]) {
        const int x = tx * 8 + ix, y = ty * 16 + iy;
        const uint32_t *base = in + ((long long)n * p.H + (y - 1)) * p.W + (x - 1);
        if (tx > 0 && ty > 0 && tx * 8 + 9 <= p.W && ty * 16 + 17 <= p.H) {
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int s = 0; s < 3; ++s) w[r * 3 + s] = __ldg(base + r * p.W + s);
        } else {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int yy = y + r - 1;
                const bool yok = yy >= 0 && yy < p.H;
#pragma unroll
                for (int s = 0; s < 3; ++s) {
                    const int xx = x + s - 1;
                    w[r * 3 + s] = (yok && xx >= 0 && xx < p.W) ? __ldg(base + r * p.W + s) : 0u;
                }
            }
        }
    };
    // tile walker: (tx, ty, n) of the tile this group works on advance by 2 with carries (no divisions in the loop)
    auto advance2 = [&](int &tx, int &ty, int &n) {
        tx += 2;
        while (tx >= p.tiles_x) { tx -= p.tiles_x; if (++ty == p.tiles_y) { ty = 0; ++n; } }
    };

    int cur_n = -1;
    uint32_t w[9];
    int tile = t_begin + grp;
    int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, n = tile / tiles_per_img;
    int gx = tx, gy = ty, gn = n;                    // tile whose input words are being fetched
    if (tile < t_end) gather(gx, gy, gn, w);
    for (uint32_t it = 0; tile < t_end; tile += 2, ++it) {
        // ---- im2col row -> shared memory (64-byte swizzle) ----
        *reinterpret_cast<uint4 *>(a_row + ((0u ^ sw) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<uint4 *>(a_row + ((1u ^ sw) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
        *reinterpret_cast<uint4 *>(a_row + ((2u ^ sw) << 4)) = make_uint4(w[8], kOnes, 0u, 0u);        // k = 18, 19: 1.0 (bias columns)
        *reinterpret_cast<uint4 *>(a_row + ((3u ^ sw) << 4)) = make_uint4(0u, 0u, 0u, 0u);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the MMA (async proxy)
        tcgen05_fence_before();                                            // this thread's TMEM reads of the previous tile are done
        named_bar_sync(1 + grp, 128);
        if (q == 0) {
            tcgen05_fence_after();
            const uint32_t leader = elect_one();
            tcgen05_mma_f16_lohi(leader, d_tmem, a_lo, ab_hi, b_lo, ab_hi, idesc, 0u);
            tcgen05_mma_f16_lohi(leader, d_tmem, a_lo + 2u, ab_hi, b_lo + 2u, ab_hi, idesc, 1u);
            tcgen05_commit_pred(leader, &mma_bar[grp]);
        }
        const int x = tx * 8 + ix, y = ty * 16 + iy;
        const bool valid = x < p.W && y < p.H;
        const int n_cur = n;
        if (do_stats && n != cur_n) {
            if (cur_n >= 0) flush_stats(cur_n);
            cur_n = n;
        }
        advance2(gx, gy, gn);
        if (tile + 2 < t_end) gather(gx, gy, gn, w);     // next tile's input words in flight during the MMA and the epilogue
        mbar_wait(&mma_bar[grp], it & 1u);
        tcgen05_fence_after();
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr, r);
        tmem_ld_wait();
        uint32_t pk[16];
        epi_chunk<DT>(r, nullptr, pk, tacc1, tacc2, do_stats && valid);
        if (valid) {
            DT *dst = out + (((long long)n_cur * p.H + y) * p.W + x) * p.Cout + nt * 32;
#pragma unroll
            for (int k = 0; k < 2; ++k)
                asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 16 * k), "r"(pk[8 * k]), "r"(pk[8 * k + 1]),
                             "r"(pk[8 * k + 2]), "r"(pk[8 * k + 3]), "r"(pk[8 * k + 4]), "r"(pk[8 * k + 5]), "r"(pk[8 * k + 6]), "r"(pk[8 * k + 7]) : "memory");
        }
        tx = gx; ty = gy; n = gn;
    }
    if (do_stats && cur_n >= 0) flush_stats(cur_n);
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
    }
}

}  // namespace ts2d
