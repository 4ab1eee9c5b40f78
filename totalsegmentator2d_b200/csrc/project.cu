// K1: one-pass max / mean (/min) intensity projection with DICOMOrient folded into the addressing.
//
// Replaces reference ts2d/core/util/image.py:32-43 (reorient_image -> sitk.DICOMOrient),
// :46-101 (project -> ITK Maximum/MeanProjectionImageFilter), ts2d/tool.py:181-185 (Cast to float32)
// and tool.py:160 (Compose): the volume is streamed from HBM exactly once and every requested channel
// is produced from the same registers.
//
// HBM-bound: algorithmic bytes = n0*n1*n2*sizeof(T) + n_modes*H*W*4.
//   strided kernel (reduce over n0 or n1, x contiguous): 16-byte vector loads, UNROLL independent loads in
//   flight per thread, per-column accumulators in registers, one shared-memory combine per work item;
//   persistent grid sized to a multiple of 148 SMs with a static round-robin over work items.
//   contiguous kernel (reduce over n2): one warp per line, vector loads + shuffle reduction.
// Numerics: max/min exact; mean = (double)sum / N rounded once to float32 - integer sums are exact
// (int32 per thread, int64 across threads), so the result is bit-identical to ITK's double accumulator.
#include "common.cuh"

namespace ts2d {

template <typename T> struct Acc { using sum_t = double; };
template <> struct Acc<uint8_t> { using sum_t = long long; };
template <> struct Acc<int16_t> { using sum_t = long long; };
template <> struct Acc<uint16_t> { using sum_t = long long; };
template <> struct Acc<int32_t> { using sum_t = long long; };

struct ProjParams {
    const void *vol;
    float *out;           // [n_modes][H][W]
    long long R;          // reduce count
    long long rstride;    // element stride between consecutive reduce steps
    long long O;          // number of outer lines
    long long ostride;    // element stride between outer lines
    long long X;          // contiguous extent (n2) for the strided kernel
    int H, W;
    int transposed;       // output rows come from x (the contiguous kept axis)
    int flip_h, flip_w;
    int n_modes;
    int modes[4];
};

__device__ __forceinline__ void store_out(const ProjParams &p, long long o, long long x, double sum, double mx, double mn) {
    long long h = p.transposed ? x : o;
    long long w = p.transposed ? o : x;
    if (p.flip_h) h = p.H - 1 - h;
    if (p.flip_w) w = p.W - 1 - w;
    const long long idx = h * p.W + w;
    const long long plane = (long long)p.H * p.W;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
        if (m < p.n_modes) {
            const int mode = p.modes[m];
            float v = mode == TS2D_PROJ_MEAN ? (float)(sum / (double)p.R) : (mode == TS2D_PROJ_MAX ? (float)mx : (float)mn);
            p.out[m * plane + idx] = v;
        }
    }
}

// ---------------- int16 fast path: 8 columns per thread, packed s16x2 max/min, IDP.2A sums -----------------
constexpr int kI16Threads = 256;
constexpr int kI16Unroll = 8;

__global__ void __launch_bounds__(kI16Threads, 3)
project_strided_i16_kernel(ProjParams p, int xt /*threads along x*/, int yt /*groups along the reduce axis*/, long long n_items, int xchunks) {
    extern __shared__ int smem_i[];   // [yt][xt*8] sums, then [yt][xt*4] packed max, [yt][xt*4] packed min
    const int tx = threadIdx.x % xt;
    const int ty = threadIdx.x / xt;
    const bool active = ty < yt;
    const int16_t *vol = (const int16_t *)p.vol;
    int *s_sum = smem_i;
    unsigned *s_max = (unsigned *)(smem_i + yt * xt * 8);
    unsigned *s_min = s_max + yt * xt * 4;

    for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const long long o = item / xchunks;
        const int xc = (int)(item % xchunks);
        const long long x0 = ((long long)xc * xt + tx) * 8;
        const bool inx = active && x0 < p.X;
        int sum[8];
        unsigned mx[4], mn[4];
#pragma unroll
        for (int i = 0; i < 8; ++i) sum[i] = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { mx[i] = 0x80008000u; mn[i] = 0x7fff7fffu; }
        if (inx) {
            const int16_t *base = vol + o * p.ostride + x0;
            for (long long y = ty; y < p.R; y += (long long)yt * kI16Unroll) {
                uint4 v[kI16Unroll];
#pragma unroll
                for (int j = 0; j < kI16Unroll; ++j) {
                    const long long yy = y + (long long)j * yt;
                    if (yy < p.R) {
                        const uint4 *ptr = (const uint4 *)(base + yy * p.rstride);
                        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                                     : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(ptr));
                    } else {
                        v[j] = make_uint4(0x80008000u, 0x80008000u, 0x80008000u, 0x80008000u);  // marker, handled below
                    }
                }
#pragma unroll
                for (int j = 0; j < kI16Unroll; ++j) {
                    const long long yy = y + (long long)j * yt;
                    if (yy < p.R) {
                        const unsigned w4[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            mx[q] = __vmaxs2(mx[q], w4[q]);
                            mn[q] = __vmins2(mn[q], w4[q]);
                            sum[2 * q] = __dp2a_lo((int)w4[q], 0x00000001, sum[2 * q]);
                            sum[2 * q + 1] = __dp2a_lo((int)w4[q], 0x00000100, sum[2 * q + 1]);
                        }
                    }
                }
            }
        }
        if (active) {
#pragma unroll
            for (int i = 0; i < 8; ++i) s_sum[(ty * 8 + i) * xt + tx] = sum[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) { s_max[(ty * 4 + i) * xt + tx] = mx[i]; s_min[(ty * 4 + i) * xt + tx] = mn[i]; }
        }
        __syncthreads();
        // combine: thread (tx, c) for c in 0..7 handled by the first 8*xt threads
        for (int idx = threadIdx.x; idx < xt * 8; idx += blockDim.x) {
            const int c = idx / xt, t = idx % xt;
            const long long x = ((long long)xc * xt + t) * 8 + c;
            if (x < p.X) {
                long long s = 0;
                int vmax = -32768, vmin = 32767;
                for (int g = 0; g < yt; ++g) {
                    s += s_sum[(g * 8 + c) * xt + t];
                    const unsigned pm = s_max[(g * 4 + (c >> 1)) * xt + t];
                    const unsigned pn = s_min[(g * 4 + (c >> 1)) * xt + t];
                    const int a = (c & 1) ? (int)(short)(pm >> 16) : (int)(short)(pm & 0xffffu);
                    const int b = (c & 1) ? (int)(short)(pn >> 16) : (int)(short)(pn & 0xffffu);
                    vmax = max(vmax, a);
                    vmin = min(vmin, b);
                }
                store_out(p, o, x, (double)s, (double)vmax, (double)vmin);
            }
        }
        __syncthreads();
    }
}

// ---------------- generic strided path: one column per thread -------------------------------------------------
// extrema are tracked in the voxel type when that is exact and cheaper (float: FMNMX instead of three fp64 operations per voxel);
// the sum stays in sum_t (double / 64-bit integer: mean = exact-as-ITK double sum / N)
template <typename T> struct Ext { using type = double; };
template <> struct Ext<float> { using type = float; };
template <typename T>
__global__ void __launch_bounds__(256)
project_strided_generic_kernel(ProjParams p, int yt, long long n_items, int xchunks) {
    using sum_t = typename Acc<T>::sum_t;
    using ext_t = typename Ext<T>::type;
    constexpr int kU = 8;                       // independent loads in flight per thread
    __shared__ double s_sum[256];
    __shared__ double s_max[256];
    __shared__ double s_min[256];
    const int xt = blockDim.x / yt;
    const int tx = threadIdx.x % xt, ty = threadIdx.x / xt;
    const T *vol = (const T *)p.vol;
    for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const long long o = item / xchunks;
        const long long x = (item % xchunks) * xt + tx;
        sum_t s = 0;
        ext_t mx = (ext_t)-1.7976931348623157e308, mn = (ext_t)1.7976931348623157e308;     // float: -inf / +inf
        if (x < p.X && ty < yt) {
            const T *base = vol + o * p.ostride + x;
            for (long long y = ty; y < p.R; y += (long long)yt * kU) {
                T v[kU];
#pragma unroll
                for (int j = 0; j < kU; ++j) {
                    const long long yy = y + (long long)j * yt;
                    v[j] = yy < p.R ? base[yy * p.rstride] : T(0);
                }
#pragma unroll
                for (int j = 0; j < kU; ++j) {
                    const long long yy = y + (long long)j * yt;
                    if (yy < p.R) {
                        s += (sum_t)v[j];
                        mx = v[j] > mx ? (ext_t)v[j] : mx;
                        mn = v[j] < mn ? (ext_t)v[j] : mn;
                    }
                }
            }
        }
        s_sum[threadIdx.x] = (double)s;   // exact: |s| < 2^53 for every supported integer type and size
        s_max[threadIdx.x] = (double)mx;
        s_min[threadIdx.x] = (double)mn;
        __syncthreads();
        if (ty == 0 && x < p.X) {
            double ss = 0, a = -1.7976931348623157e308, b = 1.7976931348623157e308;
            for (int g = 0; g < yt; ++g) {
                ss += s_sum[g * xt + tx];
                a = fmax(a, s_max[g * xt + tx]);
                b = fmin(b, s_min[g * xt + tx]);
            }
            store_out(p, o, x, ss, a, b);
        }
        __syncthreads();
    }
}

// ---------------- contiguous path: reduce along n2, one warp per line -----------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
project_contig_kernel(ProjParams p, long long n_lines, long long n1 /*lines per outer a*/) {
    using sum_t = typename Acc<T>::sum_t;
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const T *vol = (const T *)p.vol;
    for (long long line = warp; line < n_lines; line += nwarps) {
        const T *base = vol + line * p.R;
        sum_t s = 0;
        double mx = -1.7976931348623157e308, mn = 1.7976931348623157e308;
        for (long long x = lane; x < p.R; x += 32 * 4) {
            T v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = (x + 32 * j) < p.R ? base[x + 32 * j] : T(0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if ((x + 32 * j) < p.R) {
                    s += (sum_t)v[j];
                    mx = fmax(mx, (double)v[j]);
                    mn = fmin(mn, (double)v[j]);
                }
        }
        double sd = (double)s;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sd += __shfl_xor_sync(0xffffffffu, sd, o);
            mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        }
        if (lane == 0) store_out(p, line / n1, line % n1, sd, mx, mn);   // (a, b) = (i0, i1); "x" slot carries b
    }
}

template <typename T>
static int launch_generic(const ProjParams &p, int reduce_dim, long long n0, long long n1, cudaStream_t st) {
    if (reduce_dim == 2) {
        const long long lines = n0 * n1;
        const int blocks = (int)std::min<long long>((lines + 7) / 8, (long long)kNumSMs * 8);
        project_contig_kernel<T><<<blocks, 256, 0, st>>>(p, lines, n1);
    } else {
        int xt = 32;
        while (xt < 256 && xt < p.X) xt <<= 1;
        const int yt = 256 / xt;
        const int xchunks = ceil_div(p.X, xt);
        const long long items = p.O * xchunks;
        const int blocks = (int)std::min<long long>(items, (long long)kNumSMs * 6);
        project_strided_generic_kernel<T><<<blocks, 256, 0, st>>>(p, yt, items, xchunks);
    }
    return 0;
}

}  // namespace ts2d

using namespace ts2d;

extern "C" int ts2d_project(const void *vol_dev, int dtype, int64_t n0, int64_t n1, int64_t n2, int reduce_dim, int h_dim,
                            int flip_h, int flip_w, const int *modes_host, int n_modes, float *out_dev, ts2d_stream stream) {
    TS2D_REQUIRE(vol_dev && out_dev && modes_host, "ts2d_project: null pointer");
    TS2D_REQUIRE(n0 > 0 && n1 > 0 && n2 > 0, "ts2d_project: empty volume (%lld,%lld,%lld)", (long long)n0, (long long)n1, (long long)n2);
    TS2D_REQUIRE(reduce_dim >= 0 && reduce_dim <= 2 && h_dim >= 0 && h_dim <= 2 && h_dim != reduce_dim,
                 "ts2d_project: bad axes reduce_dim=%d h_dim=%d", reduce_dim, h_dim);
    TS2D_REQUIRE(n_modes >= 1 && n_modes <= 4, "ts2d_project: n_modes must be 1..4");
    const long long n[3] = {n0, n1, n2};
    const long long stride[3] = {n1 * n2, n2, 1};
    int kept[2], k = 0;
    for (int d = 0; d < 3; ++d)
        if (d != reduce_dim) kept[k++] = d;
    ProjParams p;
    memset(&p, 0, sizeof(p));
    p.vol = vol_dev;
    p.out = out_dev;
    p.R = n[reduce_dim];
    p.rstride = stride[reduce_dim];
    p.O = n[kept[0]];
    p.ostride = stride[kept[0]];
    p.X = n[kept[1]];
    p.transposed = (h_dim == kept[1]) ? 1 : 0;
    p.H = (int)n[h_dim];
    p.W = (int)n[p.transposed ? kept[0] : kept[1]];
    p.flip_h = flip_h ? 1 : 0;
    p.flip_w = flip_w ? 1 : 0;
    p.n_modes = n_modes;
    for (int i = 0; i < n_modes; ++i) {
        TS2D_REQUIRE(modes_host[i] >= 0 && modes_host[i] <= 2, "ts2d_project: bad mode %d", modes_host[i]);
        p.modes[i] = modes_host[i];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec_ok = dtype == TS2D_I16 && reduce_dim != 2 && (n2 % 8 == 0) && (((uintptr_t)vol_dev) % 16 == 0);
    if (vec_ok) {
        int xt = 8;
        while (xt < 64 && (long long)xt * 8 < p.X) xt <<= 1;
        int yt = kI16Threads / xt;
        while (yt > 1 && (long long)yt * kI16Unroll > p.R * 2) yt >>= 1;
        if ((p.R + yt - 1) / yt < 65536) {
            const int xchunks = ceil_div(p.X, (long long)xt * 8);
            const long long items = p.O * xchunks;
            const int blocks = (int)std::min<long long>(items, (long long)kNumSMs * 3);
            const size_t smem = (size_t)yt * xt * 16 * sizeof(int);
            project_strided_i16_kernel<<<blocks, kI16Threads, smem, st>>>(p, xt, yt, items, xchunks);
            TS2D_LAUNCH_CHECK();
            return TS2D_OK;
        }
    }
    switch (dtype) {
        case TS2D_U8: launch_generic<uint8_t>(p, reduce_dim, n0, n1, st); break;
        case TS2D_I16: launch_generic<int16_t>(p, reduce_dim, n0, n1, st); break;
        case TS2D_U16: launch_generic<uint16_t>(p, reduce_dim, n0, n1, st); break;
        case TS2D_I32: launch_generic<int32_t>(p, reduce_dim, n0, n1, st); break;
        case TS2D_F32: launch_generic<float>(p, reduce_dim, n0, n1, st); break;
        case TS2D_F64: launch_generic<double>(p, reduce_dim, n0, n1, st); break;
        default: set_error("ts2d_project: unsupported dtype %d", dtype); return TS2D_ERR_ARG;
    }
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
