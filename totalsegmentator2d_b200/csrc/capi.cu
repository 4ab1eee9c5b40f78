// C-ABI plumbing shared by every entry point: thread-local error slot, launch counter, device check.
#include <stdarg.h>

#include <mutex>
#include <thread>
#include <vector>

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

// ---- bit-packed label masks ("59.9 MB as uint8; 7.5 MB bit-packed") ---------------------------------------------
// The merged mask holds one 0/1 byte per (pixel, label).  On a host whose PCIe / memory path is shared by several GPUs the 60 MB
// download competes with the next study's 524 MB upload: ts2d_pack_mask_bits folds 8 bytes into one on the
// device (byte i of the output holds mask bytes 8i .. 8i+7, least significant bit first), the host copy is 8x smaller, and
// ts2d_unpack_mask_bits_host restores the uint8 array the API returns.
namespace ts2d {
__global__ void __launch_bounds__(256) pack_bits_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, long long n_out, long long n_in) {
    // one thread: 16 input bytes -> 2 output bytes (16-byte loads; the tail is handled byte-wise)
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // index of a 16-byte input group
    const long long in0 = i * 16;
    if (in0 >= n_in) return;
    if (in0 + 16 <= n_in) {
        const uint4 v = *reinterpret_cast<const uint4 *>(src + in0);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t bits = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // bytes are 0 / 1 (any non-zero counts): gather bit 0 of the four bytes
            uint32_t x = w[k];
            x |= x >> 4; x |= x >> 2; x |= x >> 1;        // bit 0 of every byte = OR of its bits
            x &= 0x01010101u;
            bits |= (((x * 0x01020408u) >> 24) & 0xFu) << (4 * k);      // bytes b0..b3 -> bits 0..3 (no carries: every product bit is distinct)
        }
        dst[2 * i] = (uint8_t)(bits & 0xFF);
        dst[2 * i + 1] = (uint8_t)(bits >> 8);
    } else {
        for (long long o = 2 * i; o < n_out && o < 2 * i + 2; ++o) {
            uint32_t b = 0;
            for (int k = 0; k < 8; ++k) {
                const long long j = o * 8 + k;
                if (j < n_in && src[j]) b |= 1u << k;
            }
            dst[o] = (uint8_t)b;
        }
    }
}
}  // namespace ts2d

extern "C" int ts2d_pack_mask_bits(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_bytes, ts2d_stream stream) {
    TS2D_REQUIRE(src_dev && dst_dev && n_bytes >= 0, "ts2d_pack_mask_bits: bad arguments");
    TS2D_REQUIRE(((uintptr_t)src_dev) % 16 == 0, "ts2d_pack_mask_bits: the mask must be 16-byte aligned");
    if (n_bytes == 0) return TS2D_OK;
    const long long groups = (n_bytes + 15) / 16, n_out = (n_bytes + 7) / 8;
    ts2d::pack_bits_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src_dev, dst_dev, n_out, n_bytes);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_unpack_mask_bits_host(const uint8_t *packed, uint8_t *out, int64_t n_bytes, int threads) {
    TS2D_REQUIRE(packed && out && n_bytes >= 0, "ts2d_unpack_mask_bits_host: bad arguments");
    const int64_t n_full = n_bytes / 8;                 // whole input bytes -> 8 output bytes each
    auto work = [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            // byte k of x = bit k of packed[i]: replicate, isolate one bit per byte, turn "non-zero" into 1 (no carries between bytes)
            uint64_t x = (uint64_t)packed[i] * 0x0101010101010101ULL;
            x &= 0x8040201008040201ULL;
            x = ((x + 0x7F7F7F7F7F7F7F7FULL) >> 7) & 0x0101010101010101ULL;
            memcpy(out + 8 * i, &x, 8);                 // little-endian hosts (x86-64 / aarch64)
        }
    };
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(threads, n_full / (1 << 16)));
    if (nt <= 1) work(0, n_full);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; ++t) pool.emplace_back(work, n_full * t / nt, n_full * (t + 1) / nt);
        for (auto &th : pool) th.join();
    }
    for (int64_t j = n_full * 8; j < n_bytes; ++j) out[j] = (packed[j / 8] >> (j & 7)) & 1;
    return TS2D_OK;
}

// ---- upload buffers (input ingest) ---------------------------------------------------------------------------------
// Page-locked host memory for volumes that are written once by the CPU (decoded from disk) and read once by the GPU.  write_combined = 1
// asks for cudaHostAllocWriteCombined: the memory is not snooped during the PCIe transfer (CPU reads of it are slow - it is an upload
// buffer, not a working array).
extern "C" void *ts2d_host_alloc(size_t bytes, int write_combined) {
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, bytes, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault);
    if (e != cudaSuccess) { ts2d::set_error("cudaHostAlloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e)); return nullptr; }
    return p;
}
extern "C" int ts2d_host_free(void *p) {
    if (!p) return TS2D_OK;
    cudaDeviceSynchronize();   // an asynchronous upload out of this memory may still be in flight (result deliberately ignored: at exit
                               // the context may be gone; cudaFreeHost below reports what matters)
    TS2D_CHECK_CUDA(cudaFreeHost(p));
    return TS2D_OK;
}

// ---- upload of a volume that lives in PAGEABLE host memory --------------------------------------------------------------------------
// sitk.ReadImage / read_image hand back ordinary (pageable) arrays (reference ts2d/core/util/image.py:225-238).  A cudaMemcpy out of
// such memory is staged by the driver through its own small pinned buffer with one CPU thread: ~10 GB/s, 50 ms for a 524 MB study
// (bench e2e.pageable: 19 studies/s against 75 from page-locked input).  Here the staging is done by the library: a ring of page-locked
// slots per GPU, `threads` CPU threads copy chunk i+1 into its slot while the DMA engine moves chunk i; the call returns when the source
// has been read completely (it may be freed or overwritten), the device copy completes in stream order.
namespace {
constexpr int kUploadSlots = 4;
constexpr size_t kUploadSlotBytes = 32u << 20;
struct UploadRing { char *slot[kUploadSlots] = {}; cudaEvent_t done[kUploadSlots] = {}; bool used[kUploadSlots] = {}; bool ready = false; };
std::mutex g_upload_mutex;          // one staged upload at a time per process (the ring is shared)
UploadRing g_upload_ring[64];
}  // namespace

extern "C" int ts2d_upload_pageable(void *dst_dev, const void *src_host, size_t bytes, int threads, ts2d_stream stream) {
    TS2D_REQUIRE(dst_dev && src_host, "ts2d_upload_pageable: null pointer");
    if (bytes == 0) return TS2D_OK;
    int dev = -1;
    TS2D_CHECK_CUDA(cudaGetDevice(&dev));
    TS2D_REQUIRE(dev >= 0 && dev < 64, "ts2d_upload_pageable: device ordinal %d outside 0..63", dev);
    std::lock_guard<std::mutex> lock(g_upload_mutex);
    UploadRing &ring = g_upload_ring[dev];
    if (!ring.ready) {
        for (int k = 0; k < kUploadSlots; ++k) {
            TS2D_CHECK_CUDA(cudaHostAlloc((void **)&ring.slot[k], kUploadSlotBytes, cudaHostAllocPortable));
            TS2D_CHECK_CUDA(cudaEventCreateWithFlags(&ring.done[k], cudaEventDisableTiming));
        }
        ring.ready = true;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int nt = std::max(1, std::min(threads, 32));
    const char *src = (const char *)src_host;
    size_t off = 0;
    for (int i = 0; off < bytes; ++i) {
        const int k = i % kUploadSlots;
        const size_t n = std::min(kUploadSlotBytes, bytes - off);
        if (ring.used[k]) TS2D_CHECK_CUDA(cudaEventSynchronize(ring.done[k]));      // the DMA that last read this slot (this or an earlier call)
        char *slot = ring.slot[k];
        if (nt == 1 || n < (size_t)(1u << 20)) memcpy(slot, src + off, n);
        else {
            std::vector<std::thread> pool;
            for (int t = 0; t < nt; ++t) {
                const size_t lo = n * t / nt, hi = n * (t + 1) / nt;
                pool.emplace_back([=] { memcpy(slot + lo, src + off + lo, hi - lo); });
            }
            for (auto &th : pool) th.join();
        }
        TS2D_CHECK_CUDA(cudaMemcpyAsync((char *)dst_dev + off, slot, n, cudaMemcpyHostToDevice, st));
        TS2D_CHECK_CUDA(cudaEventRecord(ring.done[k], st));
        ring.used[k] = true;
        off += n;
    }
    return TS2D_OK;
}
