// PNG visuals of Result.save(content='visual'): the raster arithmetic of the reference's create_visual as CUDA kernels
// (reference ts2d/core/util/image.py:383-451 create_visual, :454-481 get_auto_window, :266-290 image_vector_flatten_max,
// :293-379 resample / resample_filter / resample_uniform, ts2d/core/util/color.py:87-103 to_palette; called from
// ts2d/tool.py:272-284).  No graphics API is involved: the output is a host-side uint8 raster that the PNG encoder writes.
//
//   labels     multi-component mask [H][W][C] -> label value = 1 + highest-index non-zero component (0: none)
//              -> resample_uniform with nearest neighbour (ITK ResampleImageFilter, identity transform, default value 0)
//              -> sitk.LabelToRGB(img, 0, colormap): background -> (0, 0, 0), label L -> colormap[L % n]
//   intensity  resample_uniform with the cubic B-spline interpolator (ITK BSplineInterpolateImageFunction: recursive prefilter
//              with mirror boundaries, 4 x 4 evaluation on mirrored indices; identity when the spacing is already uniform)
//              -> window ('minmax' of the RESAMPLED image, or given) -> sitk.IntensityWindowing(lower, upper, 0, 255)
//              -> Cast(uint8) (truncation).
// Geometry of resample_filter (image.py:319-371) per axis: new_size = int(0.5 + old_size * old_sp / new_sp); the physical
// position of index int(old_size / 2) is kept at index int(new_size / 2), so output index j samples the continuous input index
//     x(j) = c_old + (j - c_new) * new_sp / old_sp,
// points outside [-0.5, size - 0.5) get the default value 0.
#include "common.cuh"

namespace ts2d {

struct VisAxis { int n_in, n_out, c_in, c_out; double ratio; };     // ratio = new_sp / old_sp

__device__ __forceinline__ double vis_src_index(const VisAxis &a, int j) { return (double)a.c_in + (double)(j - a.c_out) * a.ratio; }
__device__ __forceinline__ bool vis_inside(const VisAxis &a, double x) { return x >= -0.5 && x < (double)a.n_in - 0.5; }

struct VisPalette { int n; uint8_t rgb[257][3]; };

__global__ void __launch_bounds__(256) visual_labels_kernel(const uint8_t *mask, int C, VisAxis ay, VisAxis ax, const __grid_constant__ VisPalette pal, uint8_t *rgb) {
    const long long n = (long long)ay.n_out * ax.n_out;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int oy = (int)(i / ax.n_out), ox = (int)(i % ax.n_out);
        const double sy = vis_src_index(ay, oy), sx = vis_src_index(ax, ox);
        int label = 0;
        if (vis_inside(ay, sy) && vis_inside(ax, sx)) {
            // NearestNeighborInterpolateImageFunction: Math::RoundHalfIntegerUp
            const int iy = min(max((int)floor(sy + 0.5), 0), ay.n_in - 1), ix = min(max((int)floor(sx + 0.5), 0), ax.n_in - 1);
            const uint8_t *px = mask + ((long long)iy * ax.n_in + ix) * C;
            for (int c = C - 1; c >= 0; --c)
                if (px[c]) { label = c + 1; break; }
        }
        uint8_t r = 0, g = 0, b = 0;
        if (label != 0 && pal.n > 0) { const int k = label % pal.n; r = pal.rgb[k][0]; g = pal.rgb[k][1]; b = pal.rgb[k][2]; }
        rgb[i * 3] = r; rgb[i * 3 + 1] = g; rgb[i * 3 + 2] = b;
    }
}

// cubic B-spline prefilter with MIRROR (whole-sample symmetric) boundaries, one thread per line, in place, double precision
// (itk::BSplineDecompositionImageFilter::DataToCoefficients1D with the exact initial coefficient sums)
__global__ void __launch_bounds__(128) vis_prefilter_mirror_kernel(double *data, int n_lines, int n, long long line_stride, long long elem_stride) {
    const int line = blockIdx.x * blockDim.x + threadIdx.x;
    if (line >= n_lines || n < 2) return;
    double *c = data + (long long)line * line_stride;
    const double z = -0.26794919243112270647;   // sqrt(3) - 2
    const double gain = (1.0 - z) * (1.0 - 1.0 / z);
    for (int i = 0; i < n; ++i) c[i * elem_stride] *= gain;
    // causal initialisation (full loop): c+[0] = (sum_{k=0}^{n-1} z^k s[k] + sum_{k=1}^{n-2} z^(2n-2-k) s[k]) / (1 - z^(2n-2))
    {
        const double iz = 1.0 / z;
        double zn = z, z2n = pow(z, (double)(n - 1));
        double sum = c[0] + z2n * c[(long long)(n - 1) * elem_stride];
        z2n *= z2n * iz;
        for (int k = 1; k <= n - 2; ++k) {
            sum += (zn + z2n) * c[k * elem_stride];
            zn *= z;
            z2n *= iz;
        }
        c[0] = sum / (1.0 - zn * zn);
    }
    for (int i = 1; i < n; ++i) c[i * elem_stride] += z * c[(i - 1) * elem_stride];
    c[(long long)(n - 1) * elem_stride] = (z / (z * z - 1.0)) * (z * c[(long long)(n - 2) * elem_stride] + c[(long long)(n - 1) * elem_stride]);
    for (int i = n - 2; i >= 0; --i) c[i * elem_stride] = z * (c[(i + 1) * elem_stride] - c[i * elem_stride]);
}

__global__ void __launch_bounds__(256) vis_to_double_kernel(const float *in, double *out, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = (double)in[i];
}

__device__ __forceinline__ int vis_mirror(int i, int n) {
    if (n == 1) return 0;
    const int period = 2 * n - 2;
    i = i < 0 ? -i : i;
    i %= period;
    return i >= n ? period - i : i;
}

// out[oy][ox] (float, like the input's pixel type) ; coef == nullptr: identity geometry is assumed by the caller (plain copy)
__global__ void __launch_bounds__(256) vis_bspline_eval_kernel(const double *coef, VisAxis ay, VisAxis ax, float *out, double *minmax) {
    const long long n = (long long)ay.n_out * ax.n_out;
    double mn = 1.7976931348623157e308, mx = -1.7976931348623157e308;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int oy = (int)(i / ax.n_out), ox = (int)(i % ax.n_out);
        const double sy = vis_src_index(ay, oy), sx = vis_src_index(ax, ox);
        double v = 0.0;
        if (vis_inside(ay, sy) && vis_inside(ax, sx)) {
            double wy[4], wx[4];
            const double fy = floor(sy), fx = floor(sx);
            {
                const double t = sy - fy, u = 1.0 - t;
                wy[0] = u * u * u / 6.0; wy[1] = (t * t * (t - 2.0) * 3.0 + 4.0) / 6.0; wy[2] = (u * u * (u - 2.0) * 3.0 + 4.0) / 6.0; wy[3] = 1.0 - wy[0] - wy[1] - wy[2];
            }
            {
                const double t = sx - fx, u = 1.0 - t;
                wx[0] = u * u * u / 6.0; wx[1] = (t * t * (t - 2.0) * 3.0 + 4.0) / 6.0; wx[2] = (u * u * (u - 2.0) * 3.0 + 4.0) / 6.0; wx[3] = 1.0 - wx[0] - wx[1] - wx[2];
            }
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int yy = vis_mirror((int)fy - 1 + a, ay.n_in);
                double row = 0.0;
#pragma unroll
                for (int b = 0; b < 4; ++b) row += wx[b] * coef[(long long)yy * ax.n_in + vis_mirror((int)fx - 1 + b, ax.n_in)];
                v += wy[a] * row;
            }
        }
        const float f = (float)v;
        out[i] = f;
        mn = fmin(mn, (double)f); mx = fmax(mx, (double)f);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    if ((threadIdx.x & 31) == 0 && mn <= mx) {
        unsigned long long *a = (unsigned long long *)minmax, old = *a, assumed;
        do { assumed = old; if (__longlong_as_double((long long)assumed) <= mn) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(mn)); } while (assumed != old);
        a = (unsigned long long *)(minmax + 1); old = *a;
        do { assumed = old; if (__longlong_as_double((long long)assumed) >= mx) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(mx)); } while (assumed != old);
    }
}

__global__ void __launch_bounds__(256) vis_minmax_kernel(const float *in, long long n, double *minmax) {
    double mn = 1.7976931348623157e308, mx = -1.7976931348623157e308;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double v = (double)in[i];
        mn = fmin(mn, v); mx = fmax(mx, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    if ((threadIdx.x & 31) == 0 && mn <= mx) {
        unsigned long long *a = (unsigned long long *)minmax, old = *a, assumed;
        do { assumed = old; if (__longlong_as_double((long long)assumed) <= mn) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(mn)); } while (assumed != old);
        a = (unsigned long long *)(minmax + 1); old = *a;
        do { assumed = old; if (__longlong_as_double((long long)assumed) >= mx) break; old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(mx)); } while (assumed != old);
    }
}
__global__ void vis_minmax_init_kernel(double *minmax) { minmax[0] = 1.7976931348623157e308; minmax[1] = -1.7976931348623157e308; }

// itk::Functor::IntensityWindowingTransform<float, float> then Cast<uint8>: window taken from minmax[] when use_auto
__global__ void __launch_bounds__(256) vis_window_kernel(const float *in, long long n, double lower, double upper, int auto_lower, int auto_upper,
                                                         const double *minmax, uint8_t *out) {
    const double lo = auto_lower ? minmax[0] : lower, hi = auto_upper ? minmax[1] : upper;
    const double factor = (255.0 - 0.0) / (hi - lo), offset = 0.0 - lo * factor;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float x = in[i];
        float y;
        if ((double)x < lo) y = 0.f;
        else if ((double)x > hi) y = 255.f;
        else y = (float)((double)x * factor + offset);
        out[i] = (uint8_t)y;        // sitk.Cast(float -> uint8): truncation
    }
}

static VisAxis make_axis(int n_in, int n_out, double ratio) {
    VisAxis a;
    a.n_in = n_in; a.n_out = n_out; a.c_in = n_in / 2; a.c_out = n_out / 2; a.ratio = ratio;
    return a;
}

}  // namespace ts2d

using namespace ts2d;

extern "C" int ts2d_visual_labels(const uint8_t *mask_dev, int C, int H, int W, int H2, int W2, double ratio_y, double ratio_x,
                                  const uint8_t *palette_host, int n_colors, uint8_t *rgb_dev, ts2d_stream stream) {
    TS2D_REQUIRE(mask_dev && rgb_dev && (palette_host || n_colors == 0), "ts2d_visual_labels: null pointer");
    TS2D_REQUIRE(C >= 1 && H >= 1 && W >= 1 && H2 >= 1 && W2 >= 1, "ts2d_visual_labels: bad sizes");
    TS2D_REQUIRE(n_colors >= 0 && n_colors <= 257, "ts2d_visual_labels: the colour map holds at most 257 entries");
    VisPalette pal;
    pal.n = n_colors;
    for (int i = 0; i < n_colors; ++i)
        for (int k = 0; k < 3; ++k) pal.rgb[i][k] = palette_host[i * 3 + k];
    const long long n = (long long)H2 * W2;
    visual_labels_kernel<<<(unsigned)std::min<long long>((n + 255) / 256, (long long)kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(
        mask_dev, C, make_axis(H, H2, ratio_y), make_axis(W, W2, ratio_x), pal, rgb_dev);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_visual_intensity(const float *img_dev, int H, int W, int H2, int W2, double ratio_y, double ratio_x, int resample,
                                     double lower, double upper, int auto_lower, int auto_upper, double *scratch_dev /* H*W + 2 doubles */,
                                     float *resampled_dev /* H2*W2 floats */, uint8_t *out_dev, ts2d_stream stream) {
    TS2D_REQUIRE(img_dev && out_dev && scratch_dev && resampled_dev, "ts2d_visual_intensity: null pointer");
    TS2D_REQUIRE(H >= 1 && W >= 1 && H2 >= 1 && W2 >= 1, "ts2d_visual_intensity: bad sizes");
    TS2D_REQUIRE(resample || (H == H2 && W == W2), "ts2d_visual_intensity: sizes differ but no resampling was requested");
    cudaStream_t st = (cudaStream_t)stream;
    double *minmax = scratch_dev + (long long)H * W;
    const long long n_in = (long long)H * W, n_out = (long long)H2 * W2;
    const unsigned g_out = (unsigned)std::min<long long>((n_out + 255) / 256, (long long)kNumSMs * 8);
    vis_minmax_init_kernel<<<1, 1, 0, st>>>(minmax);
    const float *src = img_dev;
    if (resample) {
        vis_to_double_kernel<<<(unsigned)std::min<long long>((n_in + 255) / 256, (long long)kNumSMs * 8), 256, 0, st>>>(img_dev, scratch_dev, n_in);
        vis_prefilter_mirror_kernel<<<ceil_div(H, 128), 128, 0, st>>>(scratch_dev, H, W, W, 1);      // along x
        vis_prefilter_mirror_kernel<<<ceil_div(W, 128), 128, 0, st>>>(scratch_dev, W, H, 1, W);      // along y
        vis_bspline_eval_kernel<<<g_out, 256, 0, st>>>(scratch_dev, make_axis(H, H2, ratio_y), make_axis(W, W2, ratio_x), resampled_dev, minmax);
        src = resampled_dev;
    } else if (auto_lower || auto_upper) {
        vis_minmax_kernel<<<g_out, 256, 0, st>>>(img_dev, n_out, minmax);
    }
    vis_window_kernel<<<g_out, 256, 0, st>>>(src, n_out, lower, upper, auto_lower, auto_upper, minmax, out_dev);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
