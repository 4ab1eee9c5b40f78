// Probe: how does a K-major SWIZZLE_128B UMMA shared-memory descriptor behave when its start address is a
// multiple of 128 B but not of 1024 B (a "tap-shifted" view into a halo tile that TMA wrote once)?
// A halo of P pixel rows x 64 channels (128 B per row) is stored exactly as TMA SWIZZLE_128B would store it
// (16-byte chunk index XOR (row & 7), from the absolute smem address).  B = identity (64 x 64), so D = A.
// For every (shift, SBO, base_offset mode) the kernel reports which (pixel, channel) each D element came from.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../totalsegmentator2d_b200/csrc/conv_tc.cuh"
using namespace ts2d;

__global__ void __launch_bounds__(128, 1) probe_kernel(int shift_rows, int sbo_bytes, int base_off_mode, int pass, int swz64, float *out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int P = 320;
    const int rowB = swz64 ? 64 : 128;                 // bytes per pixel row (KC = 32 or 64 channels)
    const int KC = rowB / 2;
    __nv_bfloat16 *A = reinterpret_cast<__nv_bfloat16 *>(smem);                 // P rows
    __nv_bfloat16 *B = reinterpret_cast<__nv_bfloat16 *>(smem + 48 * 1024);     // 64 rows (N) x KC
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + 64 * 1024);
    uint32_t *tmem_ptr = reinterpret_cast<uint32_t *>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int chunks = rowB / 16;
    const int xmask = swz64 ? 3 : 7;
    for (int i = tid; i < P * KC; i += 128) {
        const int p = i / KC, c = i % KC;
        const float v = pass == 0 ? (float)(p % 256) : (float)c;
        // absolute-address swizzle (what TMA does): phase = (byte_addr >> 7) & 7 for 128B; for 64B swizzle (addr>>7)&3
        const uint32_t row_addr = (uint32_t)(p * rowB);
        const int phase = swz64 ? ((row_addr >> 7) & 3) : ((row_addr >> 7) & 7);
        const int ch = (c / 8) ^ (phase & xmask);
        A[(p * rowB + ch * 16 + (c % 8) * 2) / 2] = __float2bfloat16(v);
    }
    for (int i = tid; i < 64 * KC; i += 128) {
        const int n = i / KC, k = i % KC;
        const uint32_t row_addr = (uint32_t)(n * rowB);
        const int phase = swz64 ? ((row_addr >> 7) & 3) : ((row_addr >> 7) & 7);
        const int ch = (k / 8) ^ (phase & xmask);
        B[(n * rowB + ch * 16 + (k % 8) * 2) / 2] = __float2bfloat16(n == k ? 1.f : 0.f);
    }
    if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "n"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    if (tid == 0) {
        const uint32_t idesc = make_idesc<__nv_bfloat16>(64);
        const uint32_t sa = smem_u32(A) + shift_rows * rowB;
        const uint32_t sb = smem_u32(B);
        uint64_t adesc = swz64 ? make_smem_desc<32>(sa) : make_smem_desc<64>(sa);
        uint64_t bdesc = swz64 ? make_smem_desc<32>(sb) : make_smem_desc<64>(sb);
        // override SBO
        adesc &= ~((uint64_t)0x3FFF << 32);
        adesc |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
        if (base_off_mode == 1) adesc |= (uint64_t)((sa >> 7) & 7) << 49;
        for (int k = 0; k < KC / 16; ++k) tcgen05_mma_f16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, k ? 1u : 0u);
        tcgen05_commit(bar);
    }
    mbar_wait(bar, 0);
    tcgen05_fence_after();
    const int row = warp * 32 + lane;
    for (int ch = 0; ch < 2; ++ch) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(warp * 32) << 16) + ch * 32, r);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[row * 64 + ch * 32 + j] = __uint_as_float(r[j]);
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
}

int main() {
    float *d;
    cudaMalloc(&d, 128 * 64 * 4);
    std::vector<float> rows(128 * 64), cols(128 * 64);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
    for (int swz64 = 0; swz64 < 2; ++swz64) {
        const int rowB = swz64 ? 64 : 128;
        const int KC = rowB / 2;
        const int sbos[3] = {8 * rowB, 10 * rowB, 18 * rowB};
        for (int s = 0; s < 3; ++s)
            for (int shift = 0; shift <= 11; ++shift)
                for (int mode = 0; mode < 2; ++mode) {
                    for (int pass = 0; pass < 2; ++pass) {
                        cudaMemset(d, 0, 128 * 64 * 4);
                        probe_kernel<<<1, 128, 80 * 1024>>>(shift, sbos[s], mode, pass, swz64, d);
                        cudaError_t e = cudaDeviceSynchronize();
                        if (e != cudaSuccess) { printf("CUDA error %s at swz64=%d sbo=%d shift=%d mode=%d\n", cudaGetErrorString(e), swz64, sbos[s], shift, mode); return 1; }
                        cudaMemcpy(pass ? cols.data() : rows.data(), d, 128 * 64 * 4, cudaMemcpyDeviceToHost);
                    }
                    // expected: D[m][n] = A[pixel(m)][n], pixel(m) = shift + (m/8)*(sbo/rowB) + m%8
                    int bad_row = 0, bad_col = 0;
                    const int step = sbos[s] / rowB;
                    for (int m = 0; m < 128; ++m)
                        for (int n = 0; n < KC; ++n) {
                            const int ep = (shift + (m / 8) * step + (m % 8)) % 256;
                            if ((int)rows[m * 64 + n] != ep) bad_row++;
                            if ((int)cols[m * 64 + n] != n) bad_col++;
                        }
                    printf("swz%s sbo=%5d shift=%2d base_off_mode=%d : bad_row=%5d bad_col=%5d", swz64 ? "64 " : "128", sbos[s], shift, mode, bad_row, bad_col);
                    if (bad_row || bad_col) {
                        printf("  | m=1: p=%d cols[0..%d by 8]=", (int)rows[1 * 64], KC);
                        for (int n = 0; n < KC; n += 8) printf("%d ", (int)cols[1 * 64 + n]);
                        printf(" m=9: p=%d", (int)rows[9 * 64]);
                    }
                    printf("\n");
                }
    }
    return 0;
}
