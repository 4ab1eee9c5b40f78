// Micro-benchmark: how fast can ONE warp rewrite a shared-memory halo stage in place (LDS.128 -> fp32x2 fma / leaky -> STS.128), alone and
// next to warps that spin on an mbarrier (the situation of the transform warps in conv_halo_kernel<FUSE>)?
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o xform_rate xform_rate.cu && ./xform_rate
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint4 lds128(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts128(uint32_t a, const uint4 &v) { asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ uint64_t pk(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(uint64_t v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi(uint64_t v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) { uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// mode bit 0: other warps spin on an mbarrier; bit 1: no arithmetic; bit 2: no loads / stores; bit 3: scalar fp32 instead of fp32x2
__global__ void __launch_bounds__(384, 1) xform_kernel(int mode, int reps, int n_xform_warps, long long *out, float scale) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    for (int i = threadIdx.x; i < 6 * 23552 / 4; i += 384) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u + i;
    __syncthreads();
    if (warp >= 12 - n_xform_warps) {
        const int xw = warp - (12 - n_xform_warps);
        const uint32_t sbase = smem_u32(smem) + (uint32_t)(xw % 6) * 23552u;
        const int cb = lane % 8, p0 = lane / 8;
        uint64_t sc2[4], sh2[4];
        for (int k = 0; k < 4; ++k) { sc2[k] = pk(scale + k, scale - k); sh2[k] = pk(0.1f * k, -0.1f * k); }
        const uint64_t slope2 = pk(0.01f, 0.01f);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
#pragma unroll 1
            for (int s0 = 0; s0 < 45; s0 += 4) {
                uint4 v[4];
                uint32_t off[4];
                bool ok[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int pix = p0 + (s0 + i) * 4;
                    ok[i] = pix < 180;
                    const uint32_t row = (uint32_t)(ok[i] ? pix : 0) * 128u;
                    off[i] = row + ((((uint32_t)cb) ^ ((row >> 7) & 7u)) << 4);
                    if (mode & 4) v[i] = make_uint4(off[i], off[i] * 3u, off[i] * 5u, off[i] * 7u);
                    else v[i] = lds128(sbase + off[i]);
                }
                if (!(mode & 2)) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint32_t *w = reinterpret_cast<uint32_t *>(&v[i]);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            float2 f;
                            if (mode & 32) f = make_float2(__uint_as_float(w[k]), __uint_as_float(w[k] << 16));          // no real unpack
                            else if (mode & 128) f = __half22float2(*reinterpret_cast<__half2 *>(&w[k]));                // fp16 unpack
                            else f = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162 *>(&w[k]));
                            float a, b;
                            if (mode & 8) {
                                a = fmaf(f.x, lo(sc2[k]), lo(sh2[k])); b = fmaf(f.y, hi(sc2[k]), hi(sh2[k]));
                                if (!(mode & 16)) { a = fmaxf(a, a * 0.01f); b = fmaxf(b, b * 0.01f); }
                            } else {
                                const uint64_t y = fma2(pk(f.x, f.y), sc2[k], sh2[k]);
                                if (mode & 16) { a = lo(y); b = hi(y); }
                                else if (mode & 256) { a = lo(y) > 0.f ? lo(y) : lo(y) * 0.01f; b = hi(y) > 0.f ? hi(y) : hi(y) * 0.01f; }    // select-based leaky
                                else {
                                    const uint64_t ys = mul2(y, slope2);
                                    a = fmaxf(lo(y), lo(ys)); b = fmaxf(hi(y), hi(ys));
                                }
                            }
                            if (mode & 64) w[k] = __float_as_uint(a) ^ __float_as_uint(b);                                // no real pack
                            else if (mode & 128) { __half2 o = __floats2half2_rn(a, b); w[k] = *reinterpret_cast<uint32_t *>(&o); }
                            else { __nv_bfloat162 o = __floats2bfloat162_rn(a, b); w[k] = *reinterpret_cast<uint32_t *>(&o); }
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if ((mode & 4) && v[i].x != 0x12345u) continue;
                    if (ok[i]) sts128(sbase + off[i], v[i]);
                }
            }
        }
        const long long t1 = clock64();
        if (lane == 0 && xw == 0) out[blockIdx.x] = t1 - t0;
        __syncwarp();
        if (xw == 0 && lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory");
    } else if (mode & 1) {
        // spin like the MMA / epilogue warps of conv_halo_kernel do
        asm volatile(
            "{\n.reg .pred P1;\nWL:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DN;\nbra WL;\nDN:\n}\n" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    }
}

int main() {
    long long *out;
    cudaMalloc(&out, 148 * 8);
    cudaFuncSetAttribute(xform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    const int reps = 200;
    printf("one stage = 180 halo pixels x 128 B (KC = 64): 45 steps of (LDS.128, 38 instructions, STS.128) per warp\n");
    for (int nx : {1, 4, 8})
        for (int mode : {0, 4, 4 + 16, 4 + 32, 4 + 64, 4 + 16 + 32 + 64, 4 + 128, 4 + 256, 4 + 8 + 256, 4 + 8 + 256 + 128}) {
            xform_kernel<<<148, 384, 160 * 1024>>>(mode, reps, nx, out, 1.25f);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            long long h[148];
            cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0;
            for (int i = 0; i < 148; ++i) avg += (double)h[i];
            avg /= 148.0 * reps;
            printf("xform warps %d  mode %3d (%s%s%s%s%s%s%s%s): %8.0f cycles per stage per warp = %6.1f per step\n", nx, mode, (mode & 2) ? "no-math " : "",
                   (mode & 4) ? "no-lds/sts " : "", (mode & 8) ? "scalar-fp32 " : "", (mode & 16) ? "no-leaky " : "", (mode & 32) ? "no-unpack " : "",
                   (mode & 64) ? "no-pack " : "", (mode & 128) ? "fp16 " : "", (mode & 256) ? "select-leaky " : "", avg, avg / 45.0);
        }
    return 0;
}
