// Probe: sustained cycles per tcgen05.mma (cta_group::1, kind::f16, M=128, K=16) as a function of N, of the swizzle
// mode / row pitch of the A operand and of the descriptor pattern (standard 8-row groups vs the halo kernel's
// tap-shifted views with a 10-row stride).  Operands are whatever is in shared memory; only timing matters.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o mma_rate mma_rate.cu
#include <cstdio>
#include <cstdlib>
#include "../../totalsegmentator2d_b200/csrc/conv_tc.cuh"
using namespace ts2d;
namespace ts2d { void set_error(const char *, ...) {} std::atomic<long long> g_launches{0}; }

template <int N, int KC, int MODE>   // MODE 0: same A/B descriptors every MMA; 1: halo pattern (9 taps x KC/16 k-steps, shifted A views, distinct B blocks)
__global__ void __launch_bounds__(64, 1) rate_kernel(int iters, unsigned long long *out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    if (warp == 1) {
        const uint32_t leader = elect_one();
        const uint32_t idesc = make_idesc<__nv_bfloat16>(N);
        constexpr int kRowB = KC * 2;
        constexpr uint32_t a_hi = desc_hi<KC>(MODE == 1 ? 10 * kRowB : (KC == 64 ? 1024 : 512));
        constexpr uint32_t b_hi = desc_hi<KC>(KC == 64 ? 1024 : 512);
        const uint32_t a_lo = desc_lo<KC>(smem_u32(smem));
        const uint32_t b_lo = desc_lo<KC>(smem_u32(smem + 32 * 1024));
        uint32_t phase = 0;
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
                const int r = tap / 3, s = tap % 3;
#pragma unroll
                for (int k = 0; k < KC / 16; ++k) {
                    const uint32_t ao = MODE == 1 ? (uint32_t)((((r * 10 + s) * kRowB) >> 4) + k * 2) : (uint32_t)(k * 2);
                    const uint32_t bo = MODE == 1 ? (uint32_t)(((tap * N * kRowB) >> 4) + k * 2) : (uint32_t)(k * 2);
                    tcgen05_mma_f16_lohi(leader, tmem_base, a_lo + ao, a_hi, b_lo + bo, b_hi, idesc, 1u);
                }
            }
            tcgen05_commit_pred(leader, &bar);
            mbar_wait(&bar, phase);
            phase ^= 1;
        }
        const long long t1 = clock64();
        if (leader) out[blockIdx.x] = (unsigned long long)(t1 - t0);
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
}

template <int N, int KC, int MODE> static void go(unsigned long long *d) {
    const int iters = 200, smem = 200 * 1024;
    cudaFuncSetAttribute(rate_kernel<N, KC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int rep = 0; rep < 2; ++rep) {
        rate_kernel<N, KC, MODE><<<148, 64, smem>>>(iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); exit(1); }
    }
    unsigned long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double s = 0; for (int i = 0; i < 148; ++i) s += (double)h[i];
    const double per_mma = s / 148 / iters / (9 * KC / 16);
    printf("N=%3d KC=%2d (%3d-byte rows) %s : %6.1f cycles per MMA (ideal N/2 = %d)  [batches of %d MMAs + commit/wait]\n", N, KC, KC * 2,
           MODE ? "halo views  " : "same operand", per_mma, N / 2, 9 * KC / 16);
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 8);
    go<32, 32, 0>(d); go<32, 32, 1>(d); go<32, 64, 0>(d); go<32, 64, 1>(d);
    go<64, 32, 0>(d); go<64, 64, 0>(d); go<64, 64, 1>(d);
    go<128, 64, 0>(d); go<128, 64, 1>(d); go<256, 64, 0>(d);
    return 0;
}
