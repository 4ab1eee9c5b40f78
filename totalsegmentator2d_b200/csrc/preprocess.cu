// K2/K3: nnU-Net DefaultPreprocessor.run_case_npy on the device: crop_to_nonzero bbox, per-channel statistics,
// normalisation schemes, skimage-style spline resampling; plus sliding-window tile extraction.
//
// Replaces preprocessor.run_case at reference ts2d/core/inference/prediction_worker.py:194-199 (upstream
// nnunetv2: cropping.py crop_to_nonzero, default_normalization_schemes.py, default_resampling.py ->
// skimage.transform.resize == scipy.ndimage.zoom(grid_mode=True, mode='nearest') + clip) and the padding /
// slicing / torch.flip of predict_sliding_window_return_logits (reached from prediction_worker.py:209).
// All of it is HBM-bound elementwise / small-reduction work on a few MB; statistics are kept in double.
#include "common.cuh"

#include <mutex>

namespace ts2d {

// ---------------- bbox of OR_c(img[c] != 0) ---------------------------------------------------------------------
__global__ void bbox_init_kernel(int *bbox, int H, int W) {
    bbox[0] = H; bbox[1] = -1; bbox[2] = W; bbox[3] = -1;
}
__global__ void __launch_bounds__(256) bbox_kernel(const float *img, int C, int H, int W, int *bbox) {
    __shared__ int s[4];
    if (threadIdx.x == 0) { s[0] = H; s[1] = -1; s[2] = W; s[3] = -1; }
    __syncthreads();
    const long long plane = (long long)H * W;
    int h0 = H, h1 = -1, w0 = W, w1 = -1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += (long long)gridDim.x * blockDim.x) {
        bool nz = false;
        for (int c = 0; c < C; ++c) nz |= (img[c * plane + i] != 0.0f);
        if (nz) {
            const int h = (int)(i / W), w = (int)(i % W);
            h0 = min(h0, h); h1 = max(h1, h + 1); w0 = min(w0, w); w1 = max(w1, w + 1);
        }
    }
    if (h1 >= 0) { atomicMin(&s[0], h0); atomicMax(&s[1], h1); atomicMin(&s[2], w0); atomicMax(&s[3], w1); }
    __syncthreads();
    if (threadIdx.x == 0 && s[1] >= 0) { atomicMin(&bbox[0], s[0]); atomicMax(&bbox[1], s[1]); atomicMin(&bbox[2], s[2]); atomicMax(&bbox[3], s[3]); }
}
__global__ void bbox_fix_kernel(int *bbox, int H, int W) {
    if (bbox[1] < 0) { bbox[0] = 0; bbox[1] = H; bbox[2] = 0; bbox[3] = W; }   // empty mask keeps the full extent
}

// ---------------- per-channel statistics over the bbox ------------------------------------------------------------
__device__ __forceinline__ void atomic_min_double(double *addr, double v) {
    unsigned long long *a = (unsigned long long *)addr, old = *a, assumed;
    do { assumed = old; if (__longlong_as_double((long long)assumed) <= v) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v)); } while (assumed != old);
}
__device__ __forceinline__ void atomic_max_double(double *addr, double v) {
    unsigned long long *a = (unsigned long long *)addr, old = *a, assumed;
    do { assumed = old; if (__longlong_as_double((long long)assumed) >= v) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v)); } while (assumed != old);
}
__global__ void stats_init_kernel(double *stats, int C) {
    const int c = threadIdx.x;
    if (c < C) { stats[c * 4 + 0] = 0.0; stats[c * 4 + 1] = 0.0; stats[c * 4 + 2] = 1.7976931348623157e308; stats[c * 4 + 3] = -1.7976931348623157e308; }
}
__global__ void __launch_bounds__(256) stats_kernel(const float *img, int H, int W, int h0, int h1, int w0, int w1, double *stats) {
    const int c = blockIdx.y;
    const int bw = w1 - w0;
    const long long n = (long long)(h1 - h0) * bw;
    const float *base = img + (long long)c * H * W;
    double s = 0, ss = 0, mn = 1.7976931348623157e308, mx = -1.7976931348623157e308;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int h = h0 + (int)(i / bw), w = w0 + (int)(i % bw);
        const double v = (double)base[(long long)h * W + w];
        s += v; ss += v * v; mn = fmin(mn, v); mx = fmax(mx, v);
    }
    __shared__ double sh[4][8];
    s = warp_sum(s); ss = warp_sum(ss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh[0][warp] = s; sh[1][warp] = ss; sh[2][warp] = mn; sh[3][warp] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i) { s += sh[0][i]; ss += sh[1][i]; mn = fmin(mn, sh[2][i]); mx = fmax(mx, sh[3][i]); }
        atomicAdd(&stats[c * 4 + 0], s); atomicAdd(&stats[c * 4 + 1], ss);
        atomic_min_double(&stats[c * 4 + 2], mn); atomic_max_double(&stats[c * 4 + 3], mx);
    }
}

// ---------------- normalisation -------------------------------------------------------------------------------------
struct NormParams { int scheme[8]; float p[8][4]; };
__global__ void __launch_bounds__(256) normalize_kernel(const float *img, int H, int W, int h0, int h1, int w0, int w1,
                                                        NormParams np, const double *stats, float *out) {
    const int c = blockIdx.y;
    const int bw = w1 - w0, bh = h1 - h0;
    const long long n = (long long)bh * bw;
    const int scheme = np.scheme[c];
    float a = 0.f, b = 1.f, lo = 0.f, hi = 0.f;
    if (scheme == TS2D_NORM_ZSCORE) {
        const double mean = stats[c * 4 + 0] / (double)n;
        const double var = fmax(stats[c * 4 + 1] / (double)n - mean * mean, 0.0);
        a = (float)mean; b = fmaxf((float)sqrt(var), 1e-8f);
    } else if (scheme == TS2D_NORM_CT) {
        lo = np.p[c][0]; hi = np.p[c][1]; a = np.p[c][2]; b = fmaxf(np.p[c][3], 1e-8f);
    } else if (scheme == TS2D_NORM_RESCALE01) {
        a = (float)stats[c * 4 + 2]; b = fmaxf((float)stats[c * 4 + 3] - a, 1e-8f);
    } else if (scheme == TS2D_NORM_RGB01) {
        b = 255.f;
    }
    const float *base = img + (long long)c * H * W;
    float *o = out + (long long)c * n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int h = h0 + (int)(i / bw), w = w0 + (int)(i % bw);
        float v = base[(long long)h * W + w];
        if (scheme == TS2D_NORM_CT) v = fminf(fmaxf(v, lo), hi);
        if (scheme != TS2D_NORM_NONE) v = __fdiv_rn(__fsub_rn(v, a), b);   // numpy: image -= mean; image /= std (float32)
        o[i] = v;
    }
}

// ---------------- spline resampling (scipy.ndimage.zoom order 3/1, mode='nearest', grid_mode=True, + clip) --------
constexpr int kPad = 12;   // scipy _prepad_for_spline_filter for mode 'nearest'

__global__ void __launch_bounds__(256) pad_edge_kernel(const float *in, int h, int w, double *pad, double *minmax) {
    // grid.y = channel; also reduces min/max of the channel for the post-clip
    const int c = blockIdx.y;
    const int ph = h + 2 * kPad, pw = w + 2 * kPad;
    const float *src = in + (long long)c * h * w;
    double *dst = pad + (long long)c * ph * pw;
    double mn = 1.7976931348623157e308, mx = -1.7976931348623157e308;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)ph * pw; i += (long long)gridDim.x * blockDim.x) {
        const int y = min(max((int)(i / pw) - kPad, 0), h - 1), x = min(max((int)(i % pw) - kPad, 0), w - 1);
        const double v = (double)src[(long long)y * w + x];
        dst[i] = v; mn = fmin(mn, v); mx = fmax(mx, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    if ((threadIdx.x & 31) == 0) { atomic_min_double(&minmax[c * 2], mn); atomic_max_double(&minmax[c * 2 + 1], mx); }
}
__global__ void minmax_init_kernel(double *minmax, int C) {
    if (threadIdx.x < C) { minmax[threadIdx.x * 2] = 1.7976931348623157e308; minmax[threadIdx.x * 2 + 1] = -1.7976931348623157e308; }
}

// scipy ni_splines.c apply_filter for order 3 (single pole sqrt(3)-2) with the 'reflect' boundary initialisation that
// NI_EXTEND_NEAREST uses; one thread per line, in place, lines of `n` elements `stride` apart.
__global__ void __launch_bounds__(128) spline_prefilter_kernel(double *data, int n_lines, int n, long long line_stride, long long elem_stride, long long chan_stride) {
    const int line = blockIdx.x * blockDim.x + threadIdx.x;
    if (line >= n_lines) return;
    double *c = data + (long long)blockIdx.y * chan_stride + (long long)line * line_stride;
    const double z = -0.26794919243112270647;   // sqrt(3) - 2
    const double gain = (1.0 - z) * (1.0 - 1.0 / z);
    for (int i = 0; i < n; ++i) c[i * elem_stride] *= gain;
    double z_i = z;
    const double z_n = pow(z, (double)n);
    const double c0 = c[0];
    double s = c0 + z_n * c[(long long)(n - 1) * elem_stride];
    for (int i = 1; i < n; ++i) {
        s += z_i * (c[i * elem_stride] + z_n * c[(long long)(n - 1 - i) * elem_stride]);
        z_i *= z;
    }
    c[0] = s * (z / (1.0 - z_n * z_n)) + c0;
    for (int i = 1; i < n; ++i) c[i * elem_stride] += z * c[(i - 1) * elem_stride];
    c[(long long)(n - 1) * elem_stride] *= z / (z - 1.0);
    for (int i = n - 2; i >= 0; --i) c[i * elem_stride] = z * (c[(i + 1) * elem_stride] - c[i * elem_stride]);
}

__device__ __forceinline__ void bspline3_weights(double cc, int &start, double w[4]) {
    const double f = floor(cc);
    const double y = cc - f, z = 1.0 - y;
    w[1] = (y * y * (y - 2.0) * 3.0 + 4.0) / 6.0;
    w[2] = (z * z * (z - 2.0) * 3.0 + 4.0) / 6.0;
    w[0] = z * z * z / 6.0;
    w[3] = 1.0 - w[0] - w[1] - w[2];
    start = (int)f - 1;
}
__global__ void __launch_bounds__(256) spline_eval3_kernel(const double *pad, int h, int w, float *out, int h2, int w2, const double *minmax) {
    const int c = blockIdx.y;
    const int ph = h + 2 * kPad, pw = w + 2 * kPad;
    const double *src = pad + (long long)c * ph * pw;
    const double lo = minmax[c * 2], hi = minmax[c * 2 + 1];
    const double zy = (double)h / (double)h2, zx = (double)w / (double)w2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)h2 * w2; i += (long long)gridDim.x * blockDim.x) {
        const int oy = (int)(i / w2), ox = (int)(i % w2);
        int sy, sx; double wy[4], wx[4];
        bspline3_weights(((double)oy + 0.5) * zy - 0.5 + kPad, sy, wy);
        bspline3_weights(((double)ox + 0.5) * zx - 0.5 + kPad, sx, wx);
        double acc = 0.0;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int yy = min(max(sy + a, 0), ph - 1);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int xx = min(max(sx + b, 0), pw - 1);
                acc += wy[a] * wx[b] * src[(long long)yy * pw + xx];
            }
        }
        out[(long long)c * h2 * w2 + i] = (float)fmin(fmax(acc, lo), hi);
    }
}
__device__ __forceinline__ float bilinear_nearest(const float *src, int h, int w, int oy, int ox, double zy, double zx) {
    const double cy = fmin(fmax(((double)oy + 0.5) * zy - 0.5, 0.0), (double)(h - 1));
    const double cx = fmin(fmax(((double)ox + 0.5) * zx - 0.5, 0.0), (double)(w - 1));
    const int y0 = (int)floor(cy), x0 = (int)floor(cx);
    const double ty = cy - y0, tx = cx - x0;
    const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
    const double v = (1.0 - ty) * ((1.0 - tx) * (double)src[(long long)y0 * w + x0] + tx * (double)src[(long long)y0 * w + x1]) +
                     ty * ((1.0 - tx) * (double)src[(long long)y1 * w + x0] + tx * (double)src[(long long)y1 * w + x1]);
    return (float)v;
}
__global__ void __launch_bounds__(256) linear_eval_kernel(const float *in, int h, int w, float *out, int h2, int w2) {
    const int c = blockIdx.y;
    const float *src = in + (long long)c * h * w;
    const double zy = (double)h / (double)h2, zx = (double)w / (double)w2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)h2 * w2; i += (long long)gridDim.x * blockDim.x)
        out[(long long)c * h2 * w2 + i] = bilinear_nearest(src, h, w, (int)(i / w2), (int)(i % w2), zy, zx);
}
// logits [heads][h][w] -> (order-1 resize to h2 x w2) -> > 0 -> uint8 mask [H0][W0][ch_total] at (h0, w0, ch_off)
__global__ void __launch_bounds__(256) threshold_logits_kernel(const float *lg, int heads, int h, int w, int h2, int w2, uint8_t *mask,
                                                               int W0, int h0, int w0, int ch_off, int ch_total) {
    const double zy = (double)h / (double)h2, zx = (double)w / (double)w2;
    const bool same = (h == h2 && w == w2);
    const long long n = (long long)h2 * w2 * heads;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % heads);
        const long long pix = i / heads;
        const int oy = (int)(pix / w2), ox = (int)(pix % w2);
        const float *src = lg + (long long)c * h * w;
        const float v = same ? src[(long long)oy * w + ox] : bilinear_nearest(src, h, w, oy, ox, zy, zx);
        mask[((long long)(h0 + oy) * W0 + (w0 + ox)) * ch_total + ch_off + c] = v > 0.0f ? 1 : 0;
    }
}

// ---------------- sliding-window tile extraction: pad + slice + mirror + cast + NCHW->NHWC ---------------------------
// tile origins travel in kernel-parameter space (sm_70+ accepts 32 KB of parameters): 64 tiles cover every
// production layout (512x512x2000 study = 7 tiles), the 1024-tile variant serves small-patch plans
template <typename T, int CAP>
__global__ void __launch_bounds__(256) extract_tiles_kernel(const float *img, int C, int h, int w, int pad_top, int pad_left, int Ph, int Pw,
                                                            const __grid_constant__ TileListT<CAP> tl, int M, int mirror_mask, T *out) {
    const int img_idx = blockIdx.y;                 // t*M + m
    const int t = img_idx / M, m = img_idx % M;
    // upstream axes_combinations order for allowed axes (H, W): (), (H), (W), (H, W); with one allowed axis: (), (that axis)
    bool fh = false, fw = false;
    if (mirror_mask == 3) { fh = (m == 1 || m == 3); fw = (m == 2 || m == 3); }
    else if (mirror_mask == 1) fh = (m == 1);
    else if (mirror_mask == 2) fw = (m == 1);
    const int ty = tl.yx[t][0], tx = tl.yx[t][1];
    const long long npix = (long long)Ph * Pw;
    T *dst = out + (long long)img_idx * npix * C;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < npix * C; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long pix = i / C;
        int y = (int)(pix / Pw), x = (int)(pix % Pw);
        if (fh) y = Ph - 1 - y;
        if (fw) x = Pw - 1 - x;
        const int sy = ty + y - pad_top, sx = tx + x - pad_left;      // coordinates in the un-padded image
        float v = 0.f;
        if (sy >= 0 && sy < h && sx >= 0 && sx < w) v = img[((long long)c * h + sy) * w + sx];
        dst[i] = from_f32<T>(v);
    }
}

}  // namespace ts2d

using namespace ts2d;

static inline int grid_for(long long n, int threads = 256, int cap = kNumSMs * 8) {
    return (int)std::max<long long>(1, std::min<long long>((n + threads - 1) / threads, cap));
}

extern "C" int ts2d_bbox_nonzero(const float *img_dev, int C, int H, int W, int *bbox_dev, ts2d_stream stream) {
    TS2D_REQUIRE(img_dev && bbox_dev && C > 0 && H > 0 && W > 0, "ts2d_bbox_nonzero: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    bbox_init_kernel<<<1, 1, 0, st>>>(bbox_dev, H, W);
    TS2D_LAUNCH_CHECK();
    bbox_kernel<<<grid_for((long long)H * W), 256, 0, st>>>(img_dev, C, H, W, bbox_dev);
    TS2D_LAUNCH_CHECK();
    bbox_fix_kernel<<<1, 1, 0, st>>>(bbox_dev, H, W);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_channel_stats(const float *img_dev, int C, int H, int W, int h0, int h1, int w0, int w1, double *stats_dev, ts2d_stream stream) {
    TS2D_REQUIRE(img_dev && stats_dev && C > 0 && C <= 8, "ts2d_channel_stats: bad arguments (C must be 1..8)");
    TS2D_REQUIRE(0 <= h0 && h0 < h1 && h1 <= H && 0 <= w0 && w0 < w1 && w1 <= W, "ts2d_channel_stats: bad bbox");
    cudaStream_t st = (cudaStream_t)stream;
    stats_init_kernel<<<1, 32, 0, st>>>(stats_dev, C);
    TS2D_LAUNCH_CHECK();
    dim3 grid(grid_for((long long)(h1 - h0) * (w1 - w0), 256, kNumSMs * 2), C);
    stats_kernel<<<grid, 256, 0, st>>>(img_dev, H, W, h0, h1, w0, w1, stats_dev);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_normalize(const float *img_dev, int C, int H, int W, int h0, int h1, int w0, int w1, const int *scheme_host,
                              const float *params_host, const double *stats_dev, float *out_dev, ts2d_stream stream) {
    TS2D_REQUIRE(img_dev && out_dev && scheme_host && C > 0 && C <= 8, "ts2d_normalize: bad arguments (C must be 1..8)");
    TS2D_REQUIRE(0 <= h0 && h0 < h1 && h1 <= H && 0 <= w0 && w0 < w1 && w1 <= W, "ts2d_normalize: bad bbox");
    NormParams np;
    memset(&np, 0, sizeof(np));
    for (int c = 0; c < C; ++c) {
        np.scheme[c] = scheme_host[c];
        TS2D_REQUIRE(np.scheme[c] >= 0 && np.scheme[c] <= 4, "ts2d_normalize: bad scheme %d", np.scheme[c]);
        if (np.scheme[c] == TS2D_NORM_ZSCORE || np.scheme[c] == TS2D_NORM_RESCALE01)
            TS2D_REQUIRE(stats_dev != nullptr, "ts2d_normalize: scheme needs stats_dev");
        if (np.scheme[c] == TS2D_NORM_CT) TS2D_REQUIRE(params_host != nullptr, "ts2d_normalize: CT scheme needs params_host");
        if (params_host)
            for (int k = 0; k < 4; ++k) np.p[c][k] = params_host[c * 4 + k];
    }
    dim3 grid(grid_for((long long)(h1 - h0) * (w1 - w0)), C);
    normalize_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(img_dev, H, W, h0, h1, w0, w1, np, stats_dev, out_dev);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_resample2d(const float *in_dev, int C, int h, int w, float *out_dev, int h2, int w2, int order, double *scratch_dev, ts2d_stream stream) {
    TS2D_REQUIRE(in_dev && out_dev && C > 0 && h > 0 && w > 0 && h2 > 0 && w2 > 0, "ts2d_resample2d: bad arguments");
    TS2D_REQUIRE(order == 1 || order == 3, "ts2d_resample2d: order must be 1 or 3");
    cudaStream_t st = (cudaStream_t)stream;
    if (order == 1) {
        dim3 grid(grid_for((long long)h2 * w2), C);
        linear_eval_kernel<<<grid, 256, 0, st>>>(in_dev, h, w, out_dev, h2, w2);
        TS2D_LAUNCH_CHECK();
        return TS2D_OK;
    }
    TS2D_REQUIRE(scratch_dev != nullptr, "ts2d_resample2d: order 3 needs scratch_dev (C*(h+24)*(w+24)+2*C doubles)");
    TS2D_REQUIRE(C <= 32, "ts2d_resample2d: C must be <= 32 for order 3");
    const int ph = h + 2 * kPad, pw = w + 2 * kPad;
    double *minmax = scratch_dev + (long long)C * ph * pw;
    minmax_init_kernel<<<1, 32, 0, st>>>(minmax, C);
    TS2D_LAUNCH_CHECK();
    pad_edge_kernel<<<dim3(grid_for((long long)ph * pw), C), 256, 0, st>>>(in_dev, h, w, scratch_dev, minmax);
    TS2D_LAUNCH_CHECK();
    // axis 0 (columns: lines run along H), then axis 1 (rows) - scipy spline_filter order
    spline_prefilter_kernel<<<dim3(ceil_div(pw, 128), C), 128, 0, st>>>(scratch_dev, pw, ph, 1, pw, (long long)ph * pw);
    TS2D_LAUNCH_CHECK();
    spline_prefilter_kernel<<<dim3(ceil_div(ph, 128), C), 128, 0, st>>>(scratch_dev, ph, pw, pw, 1, (long long)ph * pw);
    TS2D_LAUNCH_CHECK();
    spline_eval3_kernel<<<dim3(grid_for((long long)h2 * w2), C), 256, 0, st>>>(scratch_dev, h, w, out_dev, h2, w2, minmax);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_threshold_logits(const float *logits_dev, int heads, int h, int w, int h2, int w2, uint8_t *mask_dev, int H0, int W0,
                                     int h0, int w0, int ch_off, int ch_total, ts2d_stream stream) {
    TS2D_REQUIRE(logits_dev && mask_dev && heads > 0, "ts2d_threshold_logits: bad arguments");
    TS2D_REQUIRE(h0 >= 0 && w0 >= 0 && h0 + h2 <= H0 && w0 + w2 <= W0 && ch_off >= 0 && ch_off + heads <= ch_total,
                 "ts2d_threshold_logits: output window out of range");
    threshold_logits_kernel<<<grid_for((long long)h2 * w2 * heads), 256, 0, (cudaStream_t)stream>>>(logits_dev, heads, h, w, h2, w2, mask_dev, W0,
                                                                                                 h0, w0, ch_off, ch_total);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_extract_tiles(const float *img_dev, int C, int h, int w, int pad_top, int pad_left, int Ph, int Pw, const int *tile_yx_host,
                                  int T, int mirror_mask, int precision, void *out_dev, ts2d_stream stream) {
    TS2D_REQUIRE(img_dev && out_dev && tile_yx_host && C > 0 && T > 0 && T <= kMaxTiles, "ts2d_extract_tiles: bad arguments (T must be 1..%d)", kMaxTiles);
    TS2D_REQUIRE(mirror_mask >= 0 && mirror_mask <= 3, "ts2d_extract_tiles: bad mirror mask");
    TS2D_REQUIRE(precision >= 0 && precision <= 2, "ts2d_extract_tiles: bad precision %d", precision);
    const int M = mirror_mask == 3 ? 4 : (mirror_mask ? 2 : 1);
    dim3 grid(grid_for((long long)Ph * Pw * C, 256, kNumSMs * 4), T * M);
    cudaStream_t st = (cudaStream_t)stream;
#define TS2D_EXTRACT(CAP)                                                                                                                      \
    do {                                                                                                                                       \
        TileListT<CAP> tl;                                                                                                                     \
        fill_tile_list(tl, tile_yx_host, T);                                                                                                   \
        if (precision == TS2D_PREC_F32) extract_tiles_kernel<float, CAP><<<grid, 256, 0, st>>>(img_dev, C, h, w, pad_top, pad_left, Ph, Pw, tl, M, mirror_mask, (float *)out_dev); \
        else if (precision == TS2D_PREC_BF16) extract_tiles_kernel<__nv_bfloat16, CAP><<<grid, 256, 0, st>>>(img_dev, C, h, w, pad_top, pad_left, Ph, Pw, tl, M, mirror_mask, (__nv_bfloat16 *)out_dev); \
        else extract_tiles_kernel<__half, CAP><<<grid, 256, 0, st>>>(img_dev, C, h, w, pad_top, pad_left, Ph, Pw, tl, M, mirror_mask, (__half *)out_dev); \
    } while (0)
    if (T <= 64) TS2D_EXTRACT(64); else TS2D_EXTRACT(kMaxTiles);
#undef TS2D_EXTRACT
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

// ---------------- ts2d_resample2d_static: the order-3 resampler for callers that cannot allocate ---------------------
// ts2d_resample2d needs a double-precision work area (the edge-padded copy the spline prefilter runs over in place, plus
// two doubles per channel for the clip range) and expects the caller to hand it in.  Some callers have no way to get one:
// a plugin that runs inside a host application's callback where cudaMalloc is forbidden (it takes a context-wide lock and
// may synchronise the device), code on a stream that is otherwise kept free of allocator traffic, or a binding that only
// receives an input and an output pointer from its framework and owns no memory pool of its own.  For them the work area
// lives inside the library image: one fixed array per GPU, reserved by the driver when the library's module is loaded on
// that GPU and held until the process exits.
//
// Consequences a user must know:
//   - the limit is kStaticScratchDoubles doubles (16 MiB + 512 B): C*(h+24)*(w+24) + 2*C must fit; two channels of a
//     1000 x 1000 projection fit exactly.  Larger requests are refused (TS2D_ERR_ARG), never truncated;
//   - every call on a GPU uses the SAME array.  Two calls that can overlap on the device (different streams, or
//     different host threads on different streams) corrupt each other; put them on one stream or order them with events;
//   - the array keeps what the last call left in it.  Nothing here relies on old contents: the clip range is re-initialised
//     and the padded image is rewritten in full by every call before anything reads it;
//   - the first call on each GPU looks up the array's device address (cudaGetSymbolAddress, below).  That lookup may load
//     the module on that GPU, so make the first call per GPU outside of any stream capture; later calls only launch kernels.
constexpr long long kStaticScratchDoubles = (1ll << 21) + 64;
static __device__ __align__(256) double g_resample_scratch[kStaticScratchDoubles];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
double *g_scratch_addr[kMaxDevices] = {};   // per GPU: device address of g_resample_scratch, nullptr until looked up

// In host code `g_resample_scratch` is only a handle for the runtime's symbol table, not memory a kernel may touch: the
// address the GPU sees has to be asked for, per device, with that device current.
int static_scratch_for_current_device(double **out) {
    int dev = -1;
    TS2D_CHECK_CUDA(cudaGetDevice(&dev));
    TS2D_REQUIRE(dev >= 0 && dev < kMaxDevices, "ts2d_resample2d_static: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        double *p = nullptr;
        TS2D_CHECK_CUDA(cudaGetSymbolAddress((void **)&p, g_resample_scratch));
        TS2D_REQUIRE(p != nullptr, "ts2d_resample2d_static: no device address for the built-in work area on device %d", dev);
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return TS2D_OK;
}
}  // namespace

extern "C" int64_t ts2d_resample2d_static_capacity(void) { return kStaticScratchDoubles; }

extern "C" int ts2d_resample2d_static(const float *in_dev, int C, int h, int w, float *out_dev, int h2, int w2, int order, ts2d_stream stream) {
    TS2D_REQUIRE(in_dev && out_dev && C > 0 && h > 0 && w > 0 && h2 > 0 && w2 > 0, "ts2d_resample2d_static: bad arguments");
    TS2D_REQUIRE(order == 1 || order == 3, "ts2d_resample2d_static: order must be 1 or 3");
    cudaStream_t st = (cudaStream_t)stream;
    if (order == 1) {   // no work area at all: same launch as ts2d_resample2d
        dim3 grid(grid_for((long long)h2 * w2), C);
        linear_eval_kernel<<<grid, 256, 0, st>>>(in_dev, h, w, out_dev, h2, w2);
        TS2D_LAUNCH_CHECK();
        return TS2D_OK;
    }
    TS2D_REQUIRE(C <= 32, "ts2d_resample2d_static: C must be <= 32 for order 3");
    const int ph = h + 2 * kPad, pw = w + 2 * kPad;
    const long long need = (long long)C * ph * pw + 2ll * C;
    TS2D_REQUIRE(need <= kStaticScratchDoubles,
                 "ts2d_resample2d_static: C=%d, %d x %d needs %lld doubles of work area, the built-in one holds %lld "
                 "(C*(h+24)*(w+24)+2*C must not exceed it); use ts2d_resample2d with caller-owned scratch_dev for larger images",
                 C, h, w, need, kStaticScratchDoubles);
    double *scratch = nullptr;
    const int rc = static_scratch_for_current_device(&scratch);
    if (rc != TS2D_OK) return rc;
    double *minmax = scratch + (long long)C * ph * pw;
    minmax_init_kernel<<<1, 32, 0, st>>>(minmax, C);   // the clip range is the only part that must not carry over
    TS2D_LAUNCH_CHECK();
    pad_edge_kernel<<<dim3(grid_for((long long)ph * pw), C), 256, 0, st>>>(in_dev, h, w, scratch, minmax);   // rewrites all C*ph*pw
    TS2D_LAUNCH_CHECK();
    spline_prefilter_kernel<<<dim3(ceil_div(pw, 128), C), 128, 0, st>>>(scratch, pw, ph, 1, pw, (long long)ph * pw);
    TS2D_LAUNCH_CHECK();
    spline_prefilter_kernel<<<dim3(ceil_div(ph, 128), C), 128, 0, st>>>(scratch, ph, pw, pw, 1, (long long)ph * pw);
    TS2D_LAUNCH_CHECK();
    spline_eval3_kernel<<<dim3(grid_for((long long)h2 * w2), C), 256, 0, st>>>(scratch, h, w, out_dev, h2, w2, minmax);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
