// Shared helpers for libts2d_b200 (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <string>

#include "../../include/ts2d_b200.h"

namespace ts2d {

void set_error(const char *fmt, ...);
extern std::atomic<long long> g_launches;

#define TS2D_CHECK_CUDA(expr)                                                                         \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            ::ts2d::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return TS2D_ERR_CUDA;                                                                     \
        }                                                                                             \
    } while (0)

#define TS2D_REQUIRE(cond, ...)             \
    do {                                    \
        if (!(cond)) {                      \
            ::ts2d::set_error(__VA_ARGS__); \
            return TS2D_ERR_ARG;            \
        }                                   \
    } while (0)

#define TS2D_LAUNCH_CHECK()                                                                  \
    do {                                                                                     \
        cudaError_t _e = cudaGetLastError();                                                 \
        if (_e != cudaSuccess) {                                                             \
            ::ts2d::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
            return TS2D_ERR_CUDA;                                                            \
        }                                                                                    \
        ::ts2d::g_launches.fetch_add(1, std::memory_order_relaxed);                          \
    } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

constexpr int kNumSMs = 148;   // B200

// activation element <-> float
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- programmatic dependent launch (PDL) ----
// The U-Net is ~75 back-to-back launches of 10-150 us on one stream.  Kernels launched with
// cudaLaunchAttributeProgrammaticStreamSerialization may become resident while their predecessor drains (its CTAs that finished early,
// the 2-vs-3 tile imbalance of the persistent grids): launch latency, barrier init, TMEM allocation and the tensor-map fetch overlap the
// predecessor's tail.  griddep_wait() returns once every prerequisite grid has completed and its writes are visible; nothing a
// predecessor writes (or still reads) may be touched before it.  Both are no-ops for a normally launched kernel.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// 16-byte streaming load (read once: do not allocate in L1)
__device__ __forceinline__ uint4 ld_stream_16(const void *p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// sliding-window tile origins (y, x) in padded coordinates, passed by value in kernel-parameter space
constexpr int kMaxTiles = 1024;
template <int CAP> struct TileListT { int T; int yx[CAP][2]; };
template <int CAP> static inline void fill_tile_list(TileListT<CAP> &tl, const int *yx_host, int T) {
    tl.T = T;
    for (int t = 0; t < T; ++t) { tl.yx[t][0] = yx_host[2 * t]; tl.yx[t][1] = yx_host[2 * t + 1]; }
}

static inline size_t prec_size(int precision) { return precision == TS2D_PREC_F32 ? 4 : 2; }

}  // namespace ts2d
