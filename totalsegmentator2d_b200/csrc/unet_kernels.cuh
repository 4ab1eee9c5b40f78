// CUDA-core kernels of the U-Net engine: direct convolution (parity lock / first layer / odd shapes),
// InstanceNorm-apply + LeakyReLU, and the fused mirror-average + Gaussian blend + 1x1 head + threshold kernel.
//
// Reference: dynamic_network_architectures ConvDropoutNormReLU / UNetDecoder as reached from
// ts2d/core/inference/nnu.py:149-166; blending = nnUNetPredictor._internal_predict_sliding_window_return_logits
// and export = export_prediction_from_logits (prediction_worker.py:209,215-221).
#pragma once
#include "common.cuh"

namespace ts2d {

// ------------------------------------------------------------------------------------------------------
// Direct convolution, NHWC, fp32 accumulate.  Block = 128 threads = 32 pixels (lanes) x 4 warps of 8 output
// channels; each block walks `pix_iters` groups of 32 pixels so that the InstanceNorm statistics leave the
// block as 2*8 double atomics per warp.  Weights [tap][Cin][Cout] fp32 (warp-uniform loads).
// Also used for transposed conv (k == stride): one launch per (dy,dx) with a 1x1 weight slice and the
// scatter (y*my+dy, x*mx+dx).
// ------------------------------------------------------------------------------------------------------
struct DirectConvParams {
    const void *in; long long in_ps; int Hin, Win, Cin;
    const void *in2; long long in2_ps; int c_split;     // channels [c_split, Cin) come from in2 (decoder convs: up-sampled tensor + skip); c_split == Cin: one input
    const float *w; const float *bias;
    void *out; long long out_ps; int Hout, Wout, Cout;
    int kh, kw, sh, sw, ph, pw;
    int my, mx, dy, dx, Hfull, Wfull;
    double *stats;     // [N][Cout][2] or null
    int pix_iters;
};

template <typename T>
__global__ void __launch_bounds__(128) conv_direct_kernel(const DirectConvParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.z;
    const int co0 = blockIdx.y * 32 + warp * 8;
    const int npix = p.Hout * p.Wout;
    const T *in = reinterpret_cast<const T *>(p.in) + (long long)n * p.Hin * p.Win * p.in_ps;
    const T *in2 = p.in2 ? reinterpret_cast<const T *>(p.in2) + (long long)n * p.Hin * p.Win * p.in2_ps : in;
    T *out = reinterpret_cast<T *>(p.out);
    float ssum[8], ssq[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { ssum[j] = 0.f; ssq[j] = 0.f; }
    for (int it = 0; it < p.pix_iters; ++it) {
        const int pix = (blockIdx.x * p.pix_iters + it) * 32 + lane;
        const bool valid = pix < npix;
        const int oy = valid ? pix / p.Wout : 0, ox = valid ? pix % p.Wout : 0;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = (p.bias && co0 + j < p.Cout) ? __ldg(&p.bias[co0 + j]) : 0.f;
        if (co0 < p.Cout) {
            for (int r = 0; r < p.kh; ++r) {
                const int iy = oy * p.sh + r - p.ph;
                for (int s = 0; s < p.kw; ++s) {
                    const int ix = ox * p.sw + s - p.pw;
                    const bool inb = valid && iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win;
                    const T *ip = in + ((long long)iy * p.Win + ix) * p.in_ps;
                    const T *ip2 = in2 + ((long long)iy * p.Win + ix) * p.in2_ps - p.c_split;
                    const float *wp = p.w + (long long)(r * p.kw + s) * p.Cin * p.Cout + co0;
                    for (int ci = 0; ci < p.Cin; ++ci) {
                        const float x = inb ? to_f32<T>(ci < p.c_split ? ip[ci] : ip2[ci]) : 0.f;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float wv = (co0 + j < p.Cout) ? __ldg(wp + (long long)ci * p.Cout + j) : 0.f;
                            acc[j] = fmaf(x, wv, acc[j]);
                        }
                    }
                }
            }
        }
        if (valid) {
            T *op = out + (((long long)n * p.Hfull + (long long)oy * p.my + p.dy) * p.Wfull + (long long)ox * p.mx + p.dx) * p.out_ps + co0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (co0 + j < p.Cout) {
                    const T r = from_f32<T>(acc[j]);
                    op[j] = r;
                    const float rf = to_f32<T>(r);
                    ssum[j] += rf; ssq[j] += rf * rf;
                }
            }
        }
    }
    if (p.stats) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float a = warp_sum(ssum[j]), b = warp_sum(ssq[j]);
            if (lane == 0 && co0 + j < p.Cout) {
                double *s = p.stats + ((long long)n * p.Cout + co0 + j) * 2;
                atomicAdd(s, (double)a);
                atomicAdd(s + 1, (double)b);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// First layer of the network: 3x3 / stride 1 conv with a handful of input channels (CIN <= 4: the 'mean'/'max'
// projection channels, reference ts2d/tool.py:145-160) and a 16-bit activation type.  K = 9*CIN is far too short
// for an MMA tile, and the layer is bound by writing its output (64 B per pixel at 32 channels), so it runs on the
// CUDA cores with a register tile of 4 pixels x 16 output channels per thread (64 fp32 accumulators; each weight
// fetched from shared memory feeds 4 FMAs).  Block = 256 threads = (256/CG) pixel quads x CG channel groups
// (CG = Cout/16), covering 64 x rows pixels per iteration, 4 iterations; InstanceNorm sum / sum-of-squares of the
// ROUNDED outputs leave the block as one double atomic pair per channel.
// Summation order per output = bias, then taps (r, s) outer, input channel inner: identical to conv_direct_kernel.
// ------------------------------------------------------------------------------------------------------
struct FirstConvParams {
    const void *in; int H, W;          // dense NHWC input [N][H][W][CIN]
    const float *w;                    // [tap][CIN][Cout] fp32
    const float *bias;                 // [Cout] or null
    void *out; int Cout;               // dense NHWC output [N][H][W][Cout]
    double *stats;                     // [N][Cout][2] or null
};
constexpr int kFirstConvIters = 4;

template <typename T, int CIN>
__global__ void __launch_bounds__(256) conv_first_kernel(const FirstConvParams p) {
    extern __shared__ float s_first[];           // w [9*CIN][Cout], bias [Cout], reduction scratch [8 warps][Cout][2]
    const int Cout = p.Cout, CG = Cout >> 4;
    float *s_w = s_first, *s_b = s_first + 9 * CIN * Cout, *s_red = s_b + Cout;
    for (int i = threadIdx.x; i < 9 * CIN * Cout; i += 256) s_w[i] = p.w[i];
    for (int i = threadIdx.x; i < Cout; i += 256) s_b[i] = p.bias ? p.bias[i] : 0.f;
    __syncthreads();
    const int cg = threadIdx.x % CG, quad = threadIdx.x / CG;
    const int rows = (256 / CG) / 16;            // quad rows per iteration
    const int qx = quad & 15, qy = quad >> 4;
    const int n = blockIdx.z;
    const int x0 = blockIdx.x * 64 + qx * 4;
    const T *in = reinterpret_cast<const T *>(p.in) + (long long)n * p.H * p.W * CIN;
    T *out = reinterpret_cast<T *>(p.out) + (long long)n * p.H * p.W * Cout + cg * 16;
    float ssum[16], ssq[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { ssum[k] = 0.f; ssq[k] = 0.f; }
    for (int it = 0; it < kFirstConvIters; ++it) {
        const int y = (blockIdx.y * kFirstConvIters + it) * rows + qy;
        if (y >= p.H || x0 >= p.W) continue;
        float xin[3][6][CIN];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int iy = y + r - 1;
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                const int ix = x0 + c - 1;
                const bool inb = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) xin[r][c][ci] = inb ? to_f32<T>(in[((long long)iy * p.W + ix) * CIN + ci]) : 0.f;
            }
        }
        float acc[4][16];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[j][k] = s_b[cg * 16 + k];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int s = 0; s < 3; ++s)
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) {
                    const float4 *wp = reinterpret_cast<const float4 *>(s_w + ((r * 3 + s) * CIN + ci) * Cout + cg * 16);
                    float wv[16];
#pragma unroll
                    for (int q = 0; q < 4; ++q) { const float4 t = wp[q]; wv[4 * q] = t.x; wv[4 * q + 1] = t.y; wv[4 * q + 2] = t.z; wv[4 * q + 3] = t.w; }
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int k = 0; k < 16; ++k) acc[j][k] = fmaf(xin[r][j + s][ci], wv[k], acc[j][k]);
                }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (x0 + j >= p.W) continue;
            alignas(16) T rv[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                rv[k] = from_f32<T>(acc[j][k]);
                const float rf = to_f32<T>(rv[k]);
                ssum[k] += rf; ssq[k] += rf * rf;
            }
            uint4 *dst = reinterpret_cast<uint4 *>(out + ((long long)y * p.W + x0 + j) * Cout);
            dst[0] = *reinterpret_cast<const uint4 *>(&rv[0]);
            dst[1] = *reinterpret_cast<const uint4 *>(&rv[8]);
        }
    }
    if (p.stats) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            for (int o = 16; o >= CG; o >>= 1) {           // lanes that share this channel group
                ssum[k] += __shfl_xor_sync(0xffffffffu, ssum[k], o);
                ssq[k] += __shfl_xor_sync(0xffffffffu, ssq[k], o);
            }
            if (lane < CG) { s_red[(warp * Cout + cg * 16 + k) * 2] = ssum[k]; s_red[(warp * Cout + cg * 16 + k) * 2 + 1] = ssq[k]; }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < Cout * 2; i += 256) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += (double)s_red[w * Cout * 2 + i];
            atomicAdd(p.stats + (long long)n * Cout * 2 + i, t);
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// InstanceNorm2d(affine, biased variance over H*W per (n,c)) + LeakyReLU, from the epilogue statistics.
// raw: dense [N][HW][C]; out: [N][HW] pixels with stride out_ps (may be a channel slice of a concat buffer).
// ------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) norm_act_kernel(const T *raw, const double *stats, const float *gamma, const float *beta, T *out,
                                                       long long out_ps, int HW, int C, float eps, float slope, int pix_per_block) {
    extern __shared__ float s_ss[];    // scale[C], shift[C]
    const int n = blockIdx.y;
    griddep_launch_dependents();
    griddep_wait();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const double mean = stats[((long long)n * C + c) * 2] / (double)HW;
        const double var = fmax(stats[((long long)n * C + c) * 2 + 1] / (double)HW - mean * mean, 0.0);
        const double rstd = 1.0 / sqrt(var + (double)eps);
        const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
        const float sc = (float)(rstd * (double)g);
        s_ss[c] = sc;
        s_ss[C + c] = (float)((double)b - mean * rstd * (double)g);
    }
    __syncthreads();
    const long long p0 = (long long)blockIdx.x * pix_per_block;
    const long long p1 = min((long long)HW, p0 + pix_per_block);
    const T *src = raw + ((long long)n * HW) * C;
    T *dst = out + ((long long)n * HW) * out_ps;
    if constexpr (sizeof(T) == 2) {
        if ((C & 7) == 0 && (out_ps & 7) == 0 && (256 % (C >> 3)) == 0) {
            // a thread keeps ONE block of 8 channels (256 is a multiple of C/8): scale / shift live in registers, the pixel
            // index advances by a constant, and 4 independent 16-byte loads are in flight per thread (no divisions in the loop)
            const int cv = C >> 3;
            const int cb = threadIdx.x % cv, pstep = 256 / cv;
            float sc[8], sh[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { sc[j] = s_ss[cb * 8 + j]; sh[j] = s_ss[C + cb * 8 + j]; }
            long long pix = p0 + threadIdx.x / cv;
            const T *s = src + pix * C + cb * 8;
            T *d = dst + pix * out_ps + cb * 8;
            const long long s_step = (long long)pstep * C, d_step = (long long)pstep * out_ps;
            auto apply = [&](uint4 &v) {
                uint32_t *w = reinterpret_cast<uint32_t *>(&v);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = unpack2<T>(w[k]);
                    float y0 = fmaf(f.x, sc[2 * k], sh[2 * k]), y1 = fmaf(f.y, sc[2 * k + 1], sh[2 * k + 1]);
                    y0 = y0 > 0.f ? y0 : y0 * slope;
                    y1 = y1 > 0.f ? y1 : y1 * slope;
                    w[k] = pack2<T>(y0, y1);
                }
            };
            for (; pix + 7 * pstep < p1; pix += 8 * pstep, s += 8 * s_step, d += 8 * d_step) {
                uint4 v[8];                       // 8 independent 16-byte loads in flight per thread (the projection kernel's depth)
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = ld_stream_16(s + u * s_step);
#pragma unroll
                for (int u = 0; u < 8; ++u) apply(v[u]);
#pragma unroll
                for (int u = 0; u < 8; ++u) *reinterpret_cast<uint4 *>(d + u * d_step) = v[u];
            }
            for (; pix + 3 * pstep < p1; pix += 4 * pstep, s += 4 * s_step, d += 4 * d_step) {
                uint4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = ld_stream_16(s + u * s_step);
#pragma unroll
                for (int u = 0; u < 4; ++u) apply(v[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) *reinterpret_cast<uint4 *>(d + u * d_step) = v[u];
            }
            for (; pix < p1; pix += pstep, s += s_step, d += d_step) {
                uint4 v = ld_stream_16(s);
                apply(v);
                *reinterpret_cast<uint4 *>(d) = v;
            }
            return;
        }
        if ((C & 7) == 0 && (out_ps & 7) == 0) {
            const int cv = C >> 3;
            for (long long i = p0 * cv + threadIdx.x; i < p1 * cv; i += blockDim.x) {
                const long long pix = i / cv;
                const int c0 = (int)(i % cv) * 8;
                uint4 v = *reinterpret_cast<const uint4 *>(src + pix * C + c0);
                T *e = reinterpret_cast<T *>(&v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float y = fmaf(to_f32<T>(e[j]), s_ss[c0 + j], s_ss[C + c0 + j]);
                    y = y > 0.f ? y : y * slope;
                    e[j] = from_f32<T>(y);
                }
                *reinterpret_cast<uint4 *>(dst + pix * out_ps + c0) = v;
            }
            return;
        }
    }
    for (long long i = p0 * C + threadIdx.x; i < p1 * C; i += blockDim.x) {
        const long long pix = i / C;
        const int c = (int)(i % C);
        float y = fmaf(to_f32<T>(src[i]), s_ss[c], s_ss[C + c]);
        y = y > 0.f ? y : y * slope;
        dst[pix * out_ps + c] = from_f32<T>(y);
    }
}

// scale / shift of InstanceNorm per (image, channel) from the epilogue statistics (C even; pair layout, see below);
// same arithmetic as norm_act_kernel.  Used when the last decoder conv's normalisation is applied by blend_head_kernel.
__global__ void __launch_bounds__(256) norm_finalize_kernel(const double *stats, const float *gamma, const float *beta, float *ss, int NC, int C,
                                                            int HW, float eps) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NC) return;
    const int c = i % C;
    const double mean = stats[(long long)i * 2] / (double)HW;
    const double var = fmax(stats[(long long)i * 2 + 1] / (double)HW - mean * mean, 0.0);
    const double rstd = 1.0 / sqrt(var + (double)eps);
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    // layout per channel PAIR: {scale[2j], scale[2j+1], shift[2j], shift[2j+1]} - one 16-byte load gives the two fp32x2 operands
    float *o = ss + (long long)(i >> 1) * 4 + (i & 1);
    o[0] = (float)(rstd * (double)g);
    o[2] = (float)((double)b - mean * rstd * (double)g);
}

// ------------------------------------------------------------------------------------------------------
// Tiny levels (H*W <= 256 per image): finish of a split-K convolution.  conv_tc_kernel left `ksplit` fp32 partial sums per
// output value; one block per (32-channel block, image) adds them in slice order (deterministic), adds the bias, takes the
// InstanceNorm statistics of the fp32 values, rounds to the storage type (what the unsplit epilogue stores as "raw") and
// applies normalisation + LeakyReLU with the arithmetic of norm_act_kernel.  Replaces epilogue statistics + norm_act_kernel.
// lane = channel, warp w takes pixels w, w + 8, ...   partial: [ksplit][N*HW][C] fp32;  out: [N][HW] pixels, stride out_ps.
// ------------------------------------------------------------------------------------------------------
// S = slices, J = pixels per thread (H*W <= 8*J); everything unrolled so that the S*J loads of a thread are independent and in flight together
template <typename T, int S, int J>
__global__ void __launch_bounds__(256, 1) splitk_finish_kernel(const float *partial, const float *bias, const float *gamma, const float *beta, T *out,
                                                               long long out_ps, int HW, int C, long long mtot, float eps, float slope) {
    __shared__ float s_red[2][8][32];
    __shared__ float s_ss[2][32];
    const int n = blockIdx.y, lane = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 32 + lane;
    const float b = bias ? bias[c] : 0.f;
    griddep_launch_dependents();
    griddep_wait();
    const float *base = partial + (long long)n * HW * C + c;
    const long long slice = mtot * C;
    float v[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int pix = min(w + 8 * j, HW - 1);            // clamped: no branch between the loads
        const float *src = base + (long long)pix * C;
        float a = src[0];
#pragma unroll
        for (int ks = 1; ks < S; ++ks) a += src[ks * slice];
        v[j] = a + b;
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < J; ++j)
        if (w + 8 * j < HW) { s1 += v[j]; s2 = fmaf(v[j], v[j], s2); }
    s_red[0][w][lane] = s1;
    s_red[1][w][lane] = s2;
    __syncthreads();
    if (w == 0) {
        double t1 = 0.0, t2 = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { t1 += (double)s_red[0][k][lane]; t2 += (double)s_red[1][k][lane]; }
        const double mean = t1 / (double)HW;
        const double var = fmax(t2 / (double)HW - mean * mean, 0.0);
        const double rstd = 1.0 / sqrt(var + (double)eps);
        const float g = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
        s_ss[0][lane] = (float)(rstd * (double)g);
        s_ss[1][lane] = (float)((double)be - mean * rstd * (double)g);
    }
    __syncthreads();
    const float sc = s_ss[0][lane], sh = s_ss[1][lane];
    T *dst = out + ((long long)n * HW) * out_ps + c;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int pix = w + 8 * j;
        if (pix < HW) {
            float y = fmaf(to_f32<T>(from_f32<T>(v[j])), sc, sh);
            y = y > 0.f ? y : y * slope;
            dst[(long long)pix * out_ps] = from_f32<T>(y);
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Fused: mirror-TTA average + Gaussian-weighted tile blend + 1x1 segmentation head + (logits | threshold+paste).
// The head is linear, so  sum_t g_t (W f_t + b) / sum_t g_t  ==  W (sum_t g_t f_t / sum_t g_t) + b : the F-channel
// decoder features are blended and the head applied once per output pixel.  One thread per output pixel.
// feat: [T*M][Ph][Pw][F] normalised+activated features of the last decoder conv.
// ------------------------------------------------------------------------------------------------------
struct BlendParams {
    const void *feat; int T, M, mirror_mask; int Ph, Pw;
    int pad_top, pad_left, h, w;
    const float *gauss;        // [Ph][Pw]
    const float *head_w;       // [heads][F]
    const float *head_b;       // [heads]
    int heads;
    float *logits;             // [heads][h][w] or null
    uint8_t *mask;             // [H0][W0][ch_total] or null
    int W0, h0, w0, ch_off, ch_total;
    uint8_t *present;          // [ch_total] or null
    const float *feat_ss;      // null: feat is normalised + activated; else feat is the raw conv output and {scale, shift}
    float slope;               //       (pair layout of norm_finalize_kernel) + LeakyReLU(slope) are applied here in fp32
};

template <typename T, int F, int CAP>
__global__ void __launch_bounds__(128) blend_head_kernel(const BlendParams p, const __grid_constant__ TileListT<CAP> tl) {
    extern __shared__ __align__(16) float s_head[];            // head_w [heads][F], head_b [heads], then the block's mask bytes [128][heads]
    uint8_t *s_mask = reinterpret_cast<uint8_t *>(s_head + p.heads * F + p.heads);
    int *s_any = reinterpret_cast<int *>(reinterpret_cast<long long *>(s_head) + ((p.heads * F + p.heads) * 4 + 128 * p.heads + 7) / 8 + 128);   // [heads]
    for (int i = threadIdx.x; i < p.heads; i += blockDim.x) s_any[i] = 0;
    for (int i = threadIdx.x; i < p.heads * F; i += blockDim.x) s_head[i] = p.head_w[i];
    for (int i = threadIdx.x; i < p.heads; i += blockDim.x) s_head[p.heads * F + i] = p.head_b ? p.head_b[i] : 0.f;
    __syncthreads();
    const long long npix = (long long)p.h * p.w;
    const long long pix0 = (long long)blockIdx.x * blockDim.x;
    const long long pix = pix0 + threadIdx.x;
    const bool active = pix < npix;
    const int y = active ? (int)(pix / p.w) : 0, x = active ? (int)(pix % p.w) : 0;
    const int yp = y + p.pad_top, xp = x + p.pad_left;
    float acc[F];
    uint64_t acc2[F / 2];                          // 16-bit fast path: fp32 pairs
#pragma unroll
    for (int k = 0; k < F; ++k) acc[k] = 0.f;
#pragma unroll
    for (int k = 0; k < F / 2; ++k) acc2[k] = 0ull;
    const uint64_t slope2 = f32x2_pack(p.slope, p.slope);
    const bool slope01 = p.slope >= 0.f && p.slope <= 1.f;       // LeakyReLU(y) == max(y, slope*y)
    float wsum = 0.f;
    const T *feat = reinterpret_cast<const T *>(p.feat);
    const float inv_m = 1.0f / (float)p.M;
    for (int t = 0; t < p.T && active; ++t) {
        const int ly = yp - tl.yx[t][0], lx = xp - tl.yx[t][1];
        if (ly < 0 || ly >= p.Ph || lx < 0 || lx >= p.Pw) continue;
        const float g = p.gauss[ly * p.Pw + lx];
        wsum += g;
        const float gm = g * inv_m;
        if constexpr (sizeof(T) == 2 && (F % 8) == 0) {
            // all mirror variants of this tile are loaded before any arithmetic (F/8 x M independent 16-byte loads in flight);
            // normalisation, LeakyReLU and the weighted accumulation run on fp32 PAIRS (FFMA2 / FMUL2)
            uint4 v[4][F / 8];
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                if (m >= p.M) break;
                bool fh = false, fw = false;
                if (p.mirror_mask == 3) { fh = (m & 1) != 0; fw = (m & 2) != 0; }
                else if (p.mirror_mask == 1) fh = (m == 1);
                else if (p.mirror_mask == 2) fw = (m == 1);
                const int sy = fh ? p.Ph - 1 - ly : ly, sx = fw ? p.Pw - 1 - lx : lx;
                const T *f = feat + (((long long)(t * p.M + m) * p.Ph + sy) * p.Pw + sx) * F;
#pragma unroll
                for (int k = 0; k < F / 8; ++k) v[m][k] = ld_stream_16(f + 8 * k);
            }
            const uint64_t gm2 = f32x2_pack(gm, gm);
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                if (m >= p.M) break;
                const float4 *ss = reinterpret_cast<const float4 *>(p.feat_ss) + (long long)(t * p.M + m) * (F / 2);    // warp-uniform broadcast loads
#pragma unroll
                for (int k = 0; k < F / 8; ++k) {
                    const uint32_t *w = reinterpret_cast<const uint32_t *>(&v[m][k]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int kp = k * 4 + j;
                        uint64_t x2;
                        if constexpr (std::is_same<T, __nv_bfloat16>::value) x2 = f32x2_pack_bits(w[j] << 16, w[j] & 0xffff0000u);
                        else { const float2 e = unpack2<T>(w[j]); x2 = f32x2_pack(e.x, e.y); }
                        if (p.feat_ss) {
                            const float4 c = __ldg(ss + kp);
                            const uint64_t y2 = f32x2_fma(x2, f32x2_pack(c.x, c.y), f32x2_pack(c.z, c.w));
                            const uint64_t z2 = f32x2_mul(y2, slope2);
                            const float y0 = f32x2_lo(y2), y1 = f32x2_hi(y2), z0 = f32x2_lo(z2), z1 = f32x2_hi(z2);
                            x2 = slope01 ? f32x2_pack(fmaxf(y0, z0), fmaxf(y1, z1)) : f32x2_pack(y0 > 0.f ? y0 : z0, y1 > 0.f ? y1 : z1);
                        }
                        acc2[kp] = f32x2_fma(gm2, x2, acc2[kp]);
                    }
                }
            }
        } else {
            for (int m = 0; m < p.M; ++m) {
                bool fh = false, fw = false;
                if (p.mirror_mask == 3) { fh = (m == 1 || m == 3); fw = (m == 2 || m == 3); }
                else if (p.mirror_mask == 1) fh = (m == 1);
                else if (p.mirror_mask == 2) fw = (m == 1);
                const int sy = fh ? p.Ph - 1 - ly : ly, sx = fw ? p.Pw - 1 - lx : lx;
                const T *f = feat + (((long long)(t * p.M + m) * p.Ph + sy) * p.Pw + sx) * F;
                const float *ss = p.feat_ss + (long long)(t * p.M + m) * F * 2;
#pragma unroll
                for (int k = 0; k < F; ++k) {
                    float v = to_f32<T>(f[k]);
                    if (p.feat_ss) {
                        const float sc = __ldg(ss + (k >> 1) * 4 + (k & 1)), sh = __ldg(ss + (k >> 1) * 4 + 2 + (k & 1));
                        v = fmaf(v, sc, sh);
                        v = v > 0.f ? v : v * p.slope;
                    }
                    acc[k] = fmaf(gm, v, acc[k]);
                }
            }
        }
    }
    const float inv = 1.0f / wsum;
    if constexpr (sizeof(T) == 2 && (F % 8) == 0) {
        const uint64_t inv2 = f32x2_pack(inv, inv);
#pragma unroll
        for (int k = 0; k < F / 2; ++k) acc2[k] = f32x2_mul(acc2[k], inv2);
    } else {
#pragma unroll
        for (int k = 0; k < F; ++k) acc[k] *= inv;
    }
    const int lane = threadIdx.x & 31;
    for (int c = 0; c < p.heads; ++c) {
        float v = s_head[p.heads * F + c];
        if constexpr (sizeof(T) == 2 && (F % 8) == 0) {
            const float4 *hw = reinterpret_cast<const float4 *>(s_head + c * F);       // warp-uniform 16-byte broadcast loads
            uint64_t d2 = 0ull;
#pragma unroll
            for (int k = 0; k < F / 4; ++k) {
                const float4 q = hw[k];
                d2 = f32x2_fma(f32x2_pack(q.x, q.y), acc2[2 * k], d2);
                d2 = f32x2_fma(f32x2_pack(q.z, q.w), acc2[2 * k + 1], d2);
            }
            v += f32x2_lo(d2) + f32x2_hi(d2);
        } else {
#pragma unroll
            for (int k = 0; k < F; ++k) v = fmaf(s_head[c * F + k], acc[k], v);
        }
        const bool on = active && v > 0.f;
        if (p.logits && active) p.logits[((long long)c * p.h + y) * p.w + x] = v;
        if (p.mask) s_mask[threadIdx.x * p.heads + c] = on ? 1 : 0;
        if (p.present) {
            // label seen by this warp -> block flag in shared memory (benign race); the global flag is written once per block
            // below, and only while it still reads 0: per-warp stores to ONE address from all SMs serialise in L2 and used
            // to cost more than the rest of the kernel
            const unsigned any = __ballot_sync(0xffffffffu, on);
            if (any && lane == 0) s_any[c] = 1;
        }
    }
    if (p.present) {
        __syncthreads();
        for (int c = threadIdx.x; c < p.heads; c += blockDim.x)
            if (s_any[c] && __ldcg(p.present + p.ch_off + c) == 0) p.present[p.ch_off + c] = 1;
    }
    if (p.mask) {
        // the block's 128 pixels x heads bytes leave through shared memory: a warp writes one pixel's channel window per
        // instruction (`heads` consecutive bytes at a pitch of ch_total); the pixel's byte offset comes from its own thread.
        // Pixels of a block are consecutive, so the first n_here are the active ones; the loop is unrolled so that the
        // shared-memory loads of 8 pixels are in flight together.
        long long *s_base = reinterpret_cast<long long *>(s_head) + ((p.heads * F + p.heads) * 4 + 128 * p.heads + 7) / 8;
        s_base[threadIdx.x] = ((long long)(p.h0 + y) * p.W0 + (p.w0 + x)) * p.ch_total + p.ch_off;
        __syncthreads();
        const int n_here = (int)min((long long)blockDim.x, npix - pix0);
        const int wrp = threadIdx.x >> 5;
#pragma unroll 8
        for (int j = 0; j < 32; ++j) {
            const int pp = wrp * 32 + j;
            if (pp < n_here)
                for (int c = lane; c < p.heads; c += 32) p.mask[s_base[pp] + c] = s_mask[pp * p.heads + c];
        }
    }
}

// per-image head (inspection): feat [N][HW][F] -> logits [N][heads][HW]
template <typename T>
__global__ void __launch_bounds__(128) tile_logits_kernel(const T *feat, int F, const float *head_w, const float *head_b, int heads, int HW, float *logits) {
    const int n = blockIdx.y;
    const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= HW) return;
    const T *f = feat + ((long long)n * HW + pix) * F;
    for (int c = 0; c < heads; ++c) {
        float v = head_b ? head_b[c] : 0.f;
        for (int k = 0; k < F; ++k) v = fmaf(head_w[c * F + k], to_f32<T>(f[k]), v);
        logits[((long long)n * heads + c) * HW + pix] = v;
    }
}

// dst[pix][j] = src[pix][perm[j]]  (channel order of combine_segmentations, reference ts2d/core/util/image.py:490-510)
// A block moves 128 pixels: 128*C contiguous bytes in (16-byte loads), permuted inside shared memory, 128*C contiguous
// bytes out (4-byte stores).  perm comes from kernel-parameter space (host order) or from device memory (present order).
struct PermList { int C; int perm[256]; };
__global__ void __launch_bounds__(128) permute_channels_kernel(const uint8_t *src, uint8_t *dst, long long n, const __grid_constant__ PermList pl,
                                                               const int *perm_dev) {
    extern __shared__ __align__(16) uint8_t s_px[];      // [128][C] then perm as bytes [C]
    const int C = pl.C;
    uint8_t *s_perm = s_px + 128 * C;
    for (int j = threadIdx.x; j < C; j += blockDim.x) s_perm[j] = (uint8_t)(perm_dev ? perm_dev[j] : pl.perm[j]);
    const long long p0 = (long long)blockIdx.x * 128;
    const int np = (int)min(128LL, n - p0);
    const long long bytes = (long long)np * C;
    const uint8_t *g = src + p0 * C;
    if (np == 128) {     // 128*C is a multiple of 16 and the block base a multiple of 128 bytes
        for (int i = threadIdx.x; i < 8 * C; i += blockDim.x) reinterpret_cast<uint4 *>(s_px)[i] = __ldg(reinterpret_cast<const uint4 *>(g) + i);
    } else {
        for (int i = threadIdx.x; i < bytes; i += blockDim.x) s_px[i] = g[i];
    }
    __syncthreads();
    uint8_t *o = dst + p0 * C;
    if (np == 128) {
        for (int w = threadIdx.x; w < 32 * C; w += blockDim.x) {
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int i = 4 * w + b, px = i / C, j = i - px * C;
                v |= (uint32_t)s_px[px * C + s_perm[j]] << (8 * b);
            }
            reinterpret_cast<uint32_t *>(o)[w] = v;
        }
    } else {
        for (int i = threadIdx.x; i < bytes; i += blockDim.x) { const int px = i / C, j = i - px * C; o[i] = s_px[px * C + s_perm[j]]; }
    }
}

// perm for the reference's merged channel order: per model, labels present in the image first (ascending), then the absent
// ones in the order get_annotation_meta meets them (reference ts2d/core/util/meta.py:254-286): that loop walks
// sitk GetMetaDataKeys(), i.e. a std::map, so segment ids come out in LEXICOGRAPHIC order of "Segment{id}_" (0, 10, 11, ..., 19,
// 1, 20, ...).  lex[] holds, per model window, the window's channels in that order (host-computed, see ts2d_merge_present_order).
// One thread per model; offsets[m]..offsets[m+1] is model m's channel window.
struct ModelOffsets { int n_models; int off[33]; int lex[256]; };
__global__ void present_order_kernel(const uint8_t *present, const __grid_constant__ ModelOffsets mo, int *perm) {
    const int m = threadIdx.x;
    if (m >= mo.n_models) return;
    int k = mo.off[m];
    for (int c = mo.off[m]; c < mo.off[m + 1]; ++c) if (present[c]) perm[k++] = c;
    for (int j = mo.off[m]; j < mo.off[m + 1]; ++j) { const int c = mo.lex[j]; if (!present[c]) perm[k++] = c; }
}

// NHWC (pixel stride ps) -> float32 NCHW (layer-wise parity tests)
template <typename T>
__global__ void __launch_bounds__(256) nhwc_to_nchw_f32_kernel(const T *in, long long ps, int C, int HW, float *out) {
    const int n = blockIdx.y;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)HW * C; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i / HW);
        const long long pix = i % HW;
        out[((long long)n * C + c) * HW + pix] = to_f32<T>(in[((long long)n * HW + pix) * ps + c]);
    }
}

}  // namespace ts2d
