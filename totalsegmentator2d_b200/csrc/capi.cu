// C-ABI plumbing shared by every entry point: thread-local error slot, launch counter, device check.
#include <stdarg.h>

#include "common.cuh"

namespace ts2d {
static thread_local char g_err[1024] = "";
std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

__global__ void probe_kernel(int *out) { *out = 100; }
}  // namespace ts2d

extern "C" const char *ts2d_last_error(void) { return ts2d::g_err; }
extern "C" int ts2d_abi_version(void) { return TS2D_ABI_VERSION; }
extern "C" int64_t ts2d_launch_count(void) { return ts2d::g_launches.load(); }

extern "C" int ts2d_device_check(int device) {
    cudaDeviceProp prop;
    TS2D_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        ts2d::set_error("device %d (%s) is sm_%d%d; libts2d_b200 carries sm_100a code only and has no fallback", device, prop.name, prop.major, prop.minor);
        return TS2D_ERR_DEVICE;
    }
    int cur = 0;
    TS2D_CHECK_CUDA(cudaGetDevice(&cur));
    TS2D_CHECK_CUDA(cudaSetDevice(device));
    int *d = nullptr, h = 0;
    TS2D_CHECK_CUDA(cudaMalloc(&d, sizeof(int)));
    ts2d::probe_kernel<<<1, 1>>>(d);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(&h, d, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(d);
    cudaSetDevice(cur);
    if (e != cudaSuccess || h != 100) {
        ts2d::set_error("sm_100a kernel image does not load on device %d: %s", device, cudaGetErrorString(e));
        return TS2D_ERR_DEVICE;
    }
    return TS2D_OK;
}
