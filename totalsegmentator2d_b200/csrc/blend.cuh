// K8/K9 (16-bit feature path): mirror-TTA average + Gaussian-weighted tile blend + 1x1 segmentation head + threshold / paste,
// one launch per group model.  Replaces blend_head_kernel for fp16 / bf16 features (that kernel stays for fp32 and as the
// parity reference): same arithmetic order per output value up to fp32 re-association of the head's dot product.
//
// Reference: nnUNetPredictor.predict_sliding_window_return_logits + _internal_maybe_mirror_and_predict (mirror average,
// Gaussian accumulate / normalise) and export_prediction_from_logits (sigmoid > 0.5 == logit > 0, paste at the crop box) as
// reached from reference ts2d/core/inference/prediction_worker.py:209,215-221.
//
// Why a second kernel (ncu of blend_head_kernel, profiles/r02_blend_ncu.txt): one thread per pixel read its 64-byte feature
// vector as four 16-byte loads at a 64-byte lane stride with L1 bypass, so every 32-byte sector crossed the L2 -> SM path twice,
// and the 32 x heads matvec + 16 fp32-pair accumulators + 16 staged 16-byte loads per thread left 128-thread blocks at ~25 %
// occupancy.  Here:
//   phase 1  F/8 adjacent lanes own one pixel (8 channels each): a warp's load instruction covers whole 128-byte lines; every
//            lane keeps 8 fp32 accumulators; InstanceNorm {scale, shift} of all T*M images sit in shared memory;
//   phase 2  the blended fp32 features go through shared memory ([pixel][F + 4] floats, conflict-free 16-byte rows) and thread
//            (pixel, head group) computes ceil(heads / G) logits with the head weights as warp-broadcast 16-byte loads;
//   phase 3  mask bytes leave through shared memory as consecutive bytes of each pixel's channel window.
// The grid covers the WHOLE output image: pixels outside the crop box get zeros in this model's channel window, so the
// 60 MB memset of the merged mask per study is gone (mask_covers_image = 1).
#pragma once
#include "conv_tc.cuh"
#include "unet_kernels.cuh"

namespace ts2d {

constexpr int kBlend2Threads = 256;

template <typename T, int F, int CAP, int MM /* mirror mask: 0, 1 (flip rows), 2 (flip columns), 3 (all four variants) */>
__global__ void __launch_bounds__(kBlend2Threads, F <= 32 ? 4 : 2) blend_head_v2_kernel(const BlendParams p, const __grid_constant__ TileListT<CAP> tl, int full_h, int full_w, int ss_smem) {
    static_assert(F % 8 == 0 && F <= 64, "blend_head_v2_kernel: F must be 8, 16, 32 or 64");
    constexpr int LPP = F / 8;                       // lanes per pixel
    constexpr int PPB = kBlend2Threads / LPP;        // pixels per block
    constexpr int G = LPP;                           // head groups in phase 2 (PPB * G == 256)
    constexpr int FS = F + 4;                        // padded feature row (floats): 16-byte aligned, conflict-free float4 reads
    constexpr int M = MM == 3 ? 4 : (MM ? 2 : 1);    // mirror variants per tile (compile time: no branches in the tile loop)
    extern __shared__ __align__(16) float s_mem[];
    // layout: head_w [heads][F] | head_b [heads] (padded to 4) | feat [PPB][FS] | ss [T*M][F][2] (scale, shift; only when feat_ss) |
    //         pixel base offsets [PPB] (long long) | mask bytes [PPB][heads] | any flags [heads] (int)
    const int heads = p.heads;
    const int hb_pad = (heads + 3) & ~3;
    float *s_w = s_mem;
    float *s_b = s_w + heads * F;
    float *s_feat = s_b + hb_pad;
    float *s_ss = s_feat + PPB * FS;
    const int n_img = p.T * M;
    long long *s_base = reinterpret_cast<long long *>(s_ss + ((p.feat_ss && ss_smem) ? n_img * F * 2 : 0));
    uint8_t *s_mask = reinterpret_cast<uint8_t *>(s_base + PPB);
    int *s_any = reinterpret_cast<int *>(s_mask + ((PPB * heads + 15) & ~15));

    const int tid = threadIdx.x;
    for (int i = tid; i < heads * F; i += kBlend2Threads) s_w[i] = p.head_w[i];
    for (int i = tid; i < hb_pad; i += kBlend2Threads) s_b[i] = (i < heads && p.head_b) ? p.head_b[i] : 0.f;
    for (int i = tid; i < heads; i += kBlend2Threads) s_any[i] = 0;
    if (p.feat_ss && ss_smem) {
        // norm_finalize_kernel pair layout {sc[2j], sc[2j+1], sh[2j], sh[2j+1]} -> [img][c] scale, [img][F + c] shift
        for (int i = tid; i < n_img * F; i += kBlend2Threads) {
            const int img = i / F, c = i - img * F;
            const float *src = p.feat_ss + (long long)img * F * 2 + (c >> 1) * 4 + (c & 1);
            s_ss[img * 2 * F + c] = src[0];
            s_ss[img * 2 * F + F + c] = src[2];
        }
    }
    __syncthreads();

    // PERSISTENT blocks: the prologue above (head weights + InstanceNorm table into shared memory) was 9 % of the executed instructions
    // and 15 % of the stall samples when every 64-pixel group had its own block (profiles/r02_blend_ncu.txt); now a block walks
    // groups blockIdx.x, blockIdx.x + gridDim.x, ...
    const int pl = tid / LPP, q = tid % LPP;                    // local pixel, channel block (8 channels)
    const long long npix_full = (long long)full_h * full_w;
    const long long n_groups = (npix_full + PPB - 1) / PPB;
    const T *feat = reinterpret_cast<const T *>(p.feat);
    const float inv_m = 1.0f / (float)M;
    const float slope = p.slope;
    const uint64_t slope2 = f32x2_pack(slope, slope);
    const bool slope01 = slope >= 0.f && slope <= 1.f;          // LeakyReLU(y) == max(y, slope * y)
    const int img_stride = p.Ph * p.Pw * F;                     // elements of one tile image (fits 32 bits: 512 x 512 x 64 = 2^24)
    const int hp = tid % PPB, hg = tid / PPB;                   // phase 2: pixel, head group (a warp shares hg: weight loads broadcast)
    const int hpg = (heads + G - 1) / G;                        // heads per group
    const int c_lo = hg * hpg, c_hi = min(heads, c_lo + hpg);

    for (long long group = blockIdx.x; group < n_groups; group += gridDim.x) {
        // ---------------- phase 1: blend ----------------
        const long long gp = group * PPB + pl;                  // pixel of the full output image (mask space) or of the crop (no mask)
        const bool in_img = gp < npix_full;
        const int fy = in_img ? (int)(gp / full_w) : 0, fx = in_img ? (int)(gp - (long long)fy * full_w) : 0;
        // crop coordinates (the network's grid): the mask window starts at (h0, w0)
        const int y = fy - (p.mask ? p.h0 : 0), x = fx - (p.mask ? p.w0 : 0);
        const bool active = in_img && y >= 0 && y < p.h && x >= 0 && x < p.w;
        const int yp = y + p.pad_top, xp = x + p.pad_left;
        uint64_t acc2[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) acc2[k] = 0ull;
        float wsum = 0.f;
        if (active) {
            for (int t = 0; t < p.T; ++t) {
                const int ly = yp - tl.yx[t][0], lx = xp - tl.yx[t][1];
                if (ly < 0 || ly >= p.Ph || lx < 0 || lx >= p.Pw) continue;
                const float g = __ldg(p.gauss + ly * p.Pw + lx);
                wsum += g;
                const float gm = g * inv_m;
                const uint64_t gm2 = f32x2_pack(gm, gm);
                uint4 v[M];
                const T *tbase = feat + (long long)(t * M) * img_stride + q * 8;
                // the four mirror variants of a tile read (row, col) in {ly, Ph-1-ly} x {lx, Pw-1-lx}: 32-bit offsets inside the tile's M images
                const int r0 = ly * p.Pw, r1 = (p.Ph - 1 - ly) * p.Pw, c0 = lx, c1 = p.Pw - 1 - lx;
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const bool fh = MM == 3 ? (m & 1) != 0 : (MM == 1 && m == 1);
                    const bool fw = MM == 3 ? (m & 2) != 0 : (MM == 2 && m == 1);
                    v[m] = __ldg(reinterpret_cast<const uint4 *>(tbase + (m * img_stride + ((fh ? r1 : r0) + (fw ? c1 : c0)) * F)));
                }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const uint32_t *w = reinterpret_cast<const uint32_t *>(&v[m]);
                    uint64_t f2[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) { const float2 e = unpack2<T>(w[j]); f2[j] = f32x2_pack(e.x, e.y); }
                    if (p.feat_ss) {
                        uint64_t sc2[4], sh2[4];
                        if (ss_smem) {
                            const float4 *sc = reinterpret_cast<const float4 *>(s_ss + (t * M + m) * 2 * F + q * 8);
                            const float4 *sh = reinterpret_cast<const float4 *>(s_ss + (t * M + m) * 2 * F + F + q * 8);
                            const float4 a0 = sc[0], a1 = sc[1], b0 = sh[0], b1 = sh[1];
                            sc2[0] = f32x2_pack(a0.x, a0.y); sc2[1] = f32x2_pack(a0.z, a0.w); sc2[2] = f32x2_pack(a1.x, a1.y); sc2[3] = f32x2_pack(a1.z, a1.w);
                            sh2[0] = f32x2_pack(b0.x, b0.y); sh2[1] = f32x2_pack(b0.z, b0.w); sh2[2] = f32x2_pack(b1.x, b1.y); sh2[3] = f32x2_pack(b1.z, b1.w);
                        } else {
                            // too many tiles for shared memory: the pair layout {sc, sc, sh, sh} straight from global memory (L1 hits)
                            const float4 *ss = reinterpret_cast<const float4 *>(p.feat_ss) + (long long)(t * M + m) * (F / 2) + q * 4;
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const float4 c4 = __ldg(ss + j);
                                sc2[j] = f32x2_pack(c4.x, c4.y); sh2[j] = f32x2_pack(c4.z, c4.w);
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            // same roundings as norm_act_kernel: fma, then LeakyReLU
                            const uint64_t yv = f32x2_fma(f2[j], sc2[j], sh2[j]);
                            if (slope01) {
                                const uint64_t ys = f32x2_mul(yv, slope2);
                                f2[j] = f32x2_pack(fmaxf(f32x2_lo(yv), f32x2_lo(ys)), fmaxf(f32x2_hi(yv), f32x2_hi(ys)));
                            } else {
                                const float a = f32x2_lo(yv), b = f32x2_hi(yv);
                                f2[j] = f32x2_pack(a > 0.f ? a : a * slope, b > 0.f ? b : b * slope);
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc2[j] = f32x2_fma(gm2, f2[j], acc2[j]);
                }
            }
        }
        {
            const float inv = active ? 1.0f / wsum : 0.f;
            float4 *dst = reinterpret_cast<float4 *>(s_feat + pl * FS + q * 8);
            dst[0] = make_float4(f32x2_lo(acc2[0]) * inv, f32x2_hi(acc2[0]) * inv, f32x2_lo(acc2[1]) * inv, f32x2_hi(acc2[1]) * inv);
            dst[1] = make_float4(f32x2_lo(acc2[2]) * inv, f32x2_hi(acc2[2]) * inv, f32x2_lo(acc2[3]) * inv, f32x2_hi(acc2[3]) * inv);
        }
        if (q == 0) s_base[pl] = in_img ? (active ? 1LL : 0LL) | (((long long)fy * p.W0 + fx) * p.ch_total + p.ch_off) << 1 : -1LL;
        __syncthreads();

        // ---------------- phase 2: head ----------------
        const long long hbase = s_base[hp];
        const bool h_in = hbase >= 0, h_act = h_in && (hbase & 1);
        {
            // fp32x2 dot products: the feature row and the weight row are read as pairs (f[2k], f[2k+1]) x (w[2k], w[2k+1]) into two
            // pair accumulators = 4 partial sums (k mod 4, the association of the fp32 reference head kernel)
            uint64_t fv2[F / 2];
            const float4 *fr = reinterpret_cast<const float4 *>(s_feat + hp * FS);
#pragma unroll
            for (int k = 0; k < F / 4; ++k) { const float4 t4 = fr[k]; fv2[2 * k] = f32x2_pack(t4.x, t4.y); fv2[2 * k + 1] = f32x2_pack(t4.z, t4.w); }
            // crop-space coordinates of this thread's pixel (logits output)
            const long long gp2 = group * PPB + hp;
            const int fy2 = (int)(gp2 / full_w), fx2 = (int)(gp2 - (long long)fy2 * full_w);
            const int y2 = fy2 - (p.mask ? p.h0 : 0), x2 = fx2 - (p.mask ? p.w0 : 0);
            for (int c = c_lo; c < c_hi; ++c) {
                const float4 *hw = reinterpret_cast<const float4 *>(s_w + c * F);
                uint64_t d01 = 0ull, d23 = 0ull;
#pragma unroll
                for (int k = 0; k < F / 4; ++k) {
                    const float4 wq = hw[k];
                    d01 = f32x2_fma(f32x2_pack(wq.x, wq.y), fv2[2 * k], d01);
                    d23 = f32x2_fma(f32x2_pack(wq.z, wq.w), fv2[2 * k + 1], d23);
                }
                const float v = s_b[c] + ((f32x2_lo(d01) + f32x2_hi(d01)) + (f32x2_lo(d23) + f32x2_hi(d23)));
                const bool on = h_act && v > 0.f;
                if (p.logits && h_act) p.logits[((long long)c * p.h + y2) * p.w + x2] = v;
                if (p.mask) s_mask[hp * heads + c] = on ? 1 : 0;
                if (p.present) {
                    const unsigned any = __ballot_sync(0xffffffffu, on);
                    if (any && (tid & 31) == 0) s_any[c] = 1;
                }
            }
        }
        __syncthreads();
        // ---------------- phase 3: mask bytes ----------------
        if (p.mask) {
            const int total = PPB * heads;
            for (int i = tid; i < total; i += kBlend2Threads) {
                const int pp = i / heads, c = i - pp * heads;
                const long long b = s_base[pp];
                if (b >= 0) p.mask[(b >> 1) + c] = s_mask[i];
            }
        }
        __syncthreads();                                        // s_feat / s_base / s_mask are rewritten by the next group
    }
    if (p.present)
        for (int c = tid; c < heads; c += kBlend2Threads)
            if (s_any[c] && __ldcg(p.present + p.ch_off + c) == 0) p.present[p.ch_off + c] = 1;
}

template <int F> constexpr size_t blend2_smem_bytes(int heads, int n_img, bool ss) {
    return (size_t)(heads * F + ((heads + 3) & ~3) + (kBlend2Threads / (F / 8)) * (F + 4) + (ss ? n_img * F * 2 : 0)) * sizeof(float) +
           (size_t)(kBlend2Threads / (F / 8)) * sizeof(long long) + (((size_t)(kBlend2Threads / (F / 8)) * heads + 15) & ~(size_t)15) + (size_t)heads * sizeof(int);
}

}  // namespace ts2d
