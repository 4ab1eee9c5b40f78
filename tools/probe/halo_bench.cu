// Stand-alone timing harness for conv_halo_kernel (no Python, no torch): runs the kernel on synthetic NHWC tensors
// for the U-Net's 3x3 / stride-1 layer shapes, with the diagnostic knock-outs of ConvHaloParams::dbg and the
// per-role wait counters, to find which role (TMA producer, MMA issuer, epilogue) bounds a layer.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo --expt-relaxed-constexpr -o halo_bench halo_bench.cu -lcuda
//   ./halo_bench
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#define TS2D_HALO_PROBE 1
#include "../../totalsegmentator2d_b200/csrc/conv_halo.cuh"
using namespace ts2d;

namespace ts2d {
void set_error(const char *, ...) {}
std::atomic<long long> g_launches{0};
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled g_encode = nullptr;

static void encode(CUtensorMap *tm, void *base, int rank, const cuuint64_t *dims, const cuuint64_t *str, const cuuint32_t *box, int kc) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, base, dims, str, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
}

__global__ void fill_kernel(__nv_bfloat16 *p, size_t n, uint32_t seed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u + seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = __float2bfloat16(((h & 0xffff) / 65536.0f - 0.5f) * 0.5f);
    }
}

struct Layer { const char *name; int N, H, W, Cin, Cout; };
static double *g_in_stats = nullptr;   // [N][Cin][2]: sum 0, sum of squares HW -> mean 0, var 1

template <int BN, int KC, bool FUSE = false>
static float run(const Layer &L, int stages, int dbg, bool timing, int reps, __nv_bfloat16 *in, __nv_bfloat16 *w, __nv_bfloat16 *out, float *bias,
                 double *stats, unsigned long long *dbg_out, double *waits /*[8] avg cycles*/) {
    CUtensorMap tmA, tmB;
    {
        cuuint64_t dims[4] = {(cuuint64_t)L.Cin, (cuuint64_t)L.W, (cuuint64_t)L.H, (cuuint64_t)L.N};
        cuuint64_t str[3] = {(cuuint64_t)L.Cin * 2, (cuuint64_t)L.Cin * 2 * L.W, (cuuint64_t)L.Cin * 2 * L.W * L.H};
        cuuint32_t box[4] = {(cuuint32_t)KC, kHaloW, kHaloH, 1};
        encode(&tmA, in, 4, dims, str, box, KC);
    }
    {
        cuuint64_t dims[3] = {(cuuint64_t)L.Cin, (cuuint64_t)L.Cout, 9};
        cuuint64_t str[2] = {(cuuint64_t)L.Cin * 2, (cuuint64_t)L.Cin * 2 * L.Cout};
        cuuint32_t box[3] = {(cuuint32_t)KC, (cuuint32_t)BN, 1};
        encode(&tmB, w, 3, dims, str, box, KC);
    }
    ConvHaloParams p;
    memset(&p, 0, sizeof(p));
    p.N = L.N; p.H = L.H; p.W = L.W; p.Cout = L.Cout; p.cin_chunks = L.Cin / KC; p.split_chunks = p.cin_chunks;
    p.tiles_x = (L.W + 7) / 8; p.tiles_y = (L.H + 15) / 16; p.n_stages = stages;
    p.out_ps = L.Cout; p.out = out; p.bias = bias; p.stats = stats; p.Cin = L.Cin;
    p.dbg = dbg; p.dbg_out = timing ? dbg_out : nullptr;
    if (FUSE) { p.in_stats = g_in_stats; p.in_gamma = nullptr; p.in_beta = nullptr; p.in_hw = L.H * L.W; p.in_eps = 1e-5f; p.in_slope = 0.01f; }
    const int wbytes = ((9 * L.Cin * BN * 2) + 1023) & ~1023;
    const int smem = 1024 + wbytes + stages * halo_stage_bytes<KC>() + kHaloTailBytes;
    CK(cudaFuncSetAttribute(conv_halo_kernel<BN, KC, __nv_bfloat16, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    const int nt = L.Cout / BN;
    const int total = p.tiles_x * p.tiles_y * p.N;
    const int gx = std::max(1, std::min(total, 148 / nt));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    std::vector<float> ms;
    for (int r = 0; r < reps + 2; ++r) {
        CK(cudaMemsetAsync(stats, 0, (size_t)L.N * L.Cout * 2 * sizeof(double)));
        CK(cudaEventRecord(e0));
        conv_halo_kernel<BN, KC, __nv_bfloat16, FUSE><<<dim3(gx, nt), halo_threads<FUSE>(), smem>>>(tmA, tmA, tmB, p);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float t; CK(cudaEventElapsedTime(&t, e0, e1));
        if (r >= 2) ms.push_back(t);
    }
    std::sort(ms.begin(), ms.end());
    if (timing && waits) {
        std::vector<unsigned long long> h((size_t)gx * nt * 8);
        CK(cudaMemcpy(h.data(), dbg_out, h.size() * 8, cudaMemcpyDeviceToHost));
        for (int k = 0; k < 8; ++k) { double s = 0; for (int c = 0; c < gx * nt; ++c) s += (double)h[(size_t)c * 8 + k]; waits[k] = s / (gx * nt); }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return ms[ms.size() / 2];
}

int main(int argc, char **argv) {
    {
        void *fp = nullptr;
        cudaDriverEntryPointQueryResult q;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
        g_encode = (PFN_encodeTiled)fp;
    }
    const Layer layers[] = {
        {"enc0.1  512^2  32->32 ", 12, 512, 512, 32, 32},
        {"dec6.0  512^2  64->32 ", 12, 512, 512, 64, 32},
        {"enc1.1  256^2  64->64 ", 12, 256, 256, 64, 64},
        {"dec5.0  256^2 128->64 ", 12, 256, 256, 128, 64},
        {"enc2.1  128^2 128->128", 12, 128, 128, 128, 128},
    };
    size_t max_in = 0, max_out = 0;
    for (const Layer &L : layers) {
        max_in = std::max(max_in, (size_t)L.N * L.H * L.W * L.Cin);
        max_out = std::max(max_out, (size_t)L.N * L.H * L.W * L.Cout);
    }
    __nv_bfloat16 *in, *out, *w;
    float *bias; double *stats; unsigned long long *dbg_out;
    CK(cudaMalloc(&in, max_in * 2)); CK(cudaMalloc(&out, max_out * 2)); CK(cudaMalloc(&w, (size_t)9 * 128 * 128 * 2));
    CK(cudaMalloc(&bias, 128 * 4)); CK(cudaMalloc(&stats, (size_t)12 * 128 * 2 * 8)); CK(cudaMalloc(&dbg_out, 148 * 8 * 8)); CK(cudaMemset(dbg_out, 0, 148 * 8 * 8));
    fill_kernel<<<1184, 256>>>(in, max_in, 1u);
    fill_kernel<<<64, 256>>>(w, (size_t)9 * 128 * 128, 7u);
    CK(cudaMemset(bias, 0, 128 * 4));
    CK(cudaDeviceSynchronize());
    {
        std::vector<double> h((size_t)12 * 128 * 2);
        for (size_t i = 0; i < h.size(); i += 2) { h[i] = 0.0; h[i + 1] = 512.0 * 512.0; }
        CK(cudaMalloc(&g_in_stats, h.size() * 8));
        CK(cudaMemcpy(g_in_stats, h.data(), h.size() * 8, cudaMemcpyHostToDevice));
    }
    const int dbgs[] = {0, 16};   // 128: transform without shared-memory loads / stores, 256: without arithmetic   // bit 16 (harness only): FUSE variant
    printf("dbg bits: 1 no stats, 2 no stores, 4 tap 0 only, 8 all halos = tile 0 (L2 hits), 16 FUSE (deferred InstanceNorm on the halo)\n");
    for (const Layer &L : layers) {
        const int kc = L.Cin % 64 == 0 ? 64 : 32;
        const int bn = L.Cout % 64 == 0 ? 64 : 32;
        int stages = 6;
        auto smem_of = [&](int st) { return 1024 + (((9 * L.Cin * bn * 2) + 1023) & ~1023) + st * (kc == 64 ? halo_stage_bytes<64>() : halo_stage_bytes<32>()) + kHaloTailBytes; };
        while (stages > 2 && smem_of(stages) > 225 * 1024) --stages;
        const double flops = 2.0 * L.N * L.H * L.W * 9.0 * L.Cin * L.Cout;
        const double bytes = 2.0 * L.N * L.H * L.W * (L.Cin + L.Cout);
        const double tiles_per_cta = (double)((L.W + 7) / 8) * ((L.H + 15) / 16) * L.N / (148 / (L.Cout / bn));
        printf("%s  BN=%d KC=%d stages=%d  tiles/CTA %.1f\n", L.name, bn, kc, stages, tiles_per_cta);
        for (int dbg : dbgs) {
            double waits[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            float ms = 0.f;
            for (int timing = 0; timing < 2; ++timing) {
                float t;
                const int d = dbg & ~16;
                if (dbg & 16) {
                    if (bn == 32 && kc == 32) t = run<32, 32, true>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                    else if (bn == 32 && kc == 64) t = run<32, 64, true>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                    else if (bn == 64 && kc == 32) t = run<64, 32, true>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                    else t = run<64, 64, true>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                } else if (bn == 32 && kc == 32) t = run<32, 32>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                else if (bn == 32 && kc == 64) t = run<32, 64>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                else if (bn == 64 && kc == 32) t = run<64, 32>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                else t = run<64, 64>(L, stages, d, timing, 5, in, w, out, bias, stats, dbg_out, waits);
                if (!timing) ms = t;
            }
            printf("  dbg=%2d  %8.4f ms  %7.1f TF/s  %6.0f GB/s | cycles/CTA %9.0f (%.0f per tile, clock %.2f GHz)  mma-wait-halo %4.1f%%  mma-wait-tmem %4.1f%%  epi-wait %4.1f%%  xform-wait %4.1f%% xform-work %4.1f%%\n",
                   dbg, ms, flops / ms / 1e9, bytes / ms / 1e6, waits[0], waits[0] / tiles_per_cta, waits[0] / (ms * 1e6), 100 * waits[1] / waits[0],
                   100 * waits[2] / waits[0], 100 * waits[3] / waits[0], 100 * waits[4] / waits[0], 100 * waits[5] / waits[0]);
        }
    }
    return 0;
}
