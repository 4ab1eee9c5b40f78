// The U-Net engine behind the C ABI: plan -> layer graph, weight repacking, workspace, TMA descriptors, forward.
//
// Mirrors what nnUNetPredictor.initialize_from_trained_model_folder builds (reference
// ts2d/core/inference/nnu.py:149-166): dynamic_network_architectures PlainConvUNet =
// encoder stages of [Conv2d k3 (first conv strided) -> InstanceNorm2d -> LeakyReLU], decoder levels of
// ConvTranspose2d(k = stride) -> cat(up, skip) -> conv blocks, 1x1 seg head (deep supervision off).
//
// Data layout in HBM (activations NHWC, element = plan precision):
//   skip[l]  [N][H_l][W_l][C_l]     encoder stage output (normalised), read by the next stage's strided conv and by the decoder
//   up[l]    [N][H_l][W_l][C_l]     transposed-conv output.  torch.cat([up, skip]) never happens: the first decoder conv of a level
//                                   takes its K-chunks from TWO tensor maps (channels [0,C_l) from up[l], [C_l,2C_l) from skip[l]).
//                                   (Round 1 interleaved the two in one [..][2*C_l] buffer; at level 0 a half pixel is 64 bytes and the
//                                   strided conv enc1.0 fetched 412 MB of DRAM for 201 MB of operand.)
//   bott     [N][H_L][W_L][C_L]     bottleneck
//   raw      conv output before normalisation (+ per-(n,c) sum / sum-of-squares in `stats`, double)
//   tmpA/B   normalised intermediates / decoder stage outputs
#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "conv_tc.cuh"
#include "conv_tc2.cuh"
#include "conv_halo.cuh"
#include "conv_halo_s2.cuh"
#include "conv_first_tc.cuh"
#include "unet_kernels.cuh"
#include "blend.cuh"

namespace ts2d {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}

struct Param {
    std::vector<float> host;
    bool have = false;
    float *dev = nullptr;
};

struct ConvLayer {
    std::string name;           // "enc{s}.{i}" / "dec{d}.{i}"
    int cin = 0, cout = 0, kh = 3, kw = 3, sh = 1, sw = 1;
    int Hin = 0, Win = 0, Hout = 0, Wout = 0;
    Param w, b, gamma, beta;    // upstream layouts: w OIHW
    float *w_direct = nullptr;  // [tap][Cin][Cout] fp32
    void *w_tc = nullptr;       // [tap][Cout][Cin] bf16/fp16
    const void *in = nullptr;   // set per forward for the first layer
    long long in_ps = 0;
    const void *in2 = nullptr;  // decoder first convs: channels [c_split, cin) come from this tensor (the encoder skip)
    long long in2_ps = 0;
    int c_split = 0;            // channels taken from `in` (== cin when there is one input)
    void *norm_out = nullptr;
    long long out_ps = 0;
    double *stats = nullptr;
    bool tc = false;
    int bn = 0, kc = 0, enc_n = 0;
    CUtensorMap tmA, tmA2, tmB;
    bool first = false;         // reads the caller's tiles: tmA encoded per forward
    bool halo = false;          // conv_halo_kernel (3x3 stride 1, halo tile + resident weights)
    int halo_bn = 0, halo_stages = 0;
    bool halo_maps = false;
    bool halo_s2 = false;       // conv_halo_s2_kernel (3x3 stride 2, parity-plane halo + resident weights)
    bool first_tc = false;      // conv_first_tc_kernel (CIN == 2, 3x3 stride 1: im2col rows built by the workers, K = 32)
    bool first_fast = false;    // conv_first_kernel (CIN <= 4, 3x3 stride 1, 16-bit activations)
    void *out_raw = nullptr;    // pre-normalisation output (ping-pong between two buffers)
    bool defer = false;         // no norm_act pass: the consumer applies InstanceNorm + LeakyReLU on its operand path
    int prod = -1;              // >= 0: the input is conv `prod`'s raw output (this layer runs conv_halo_kernel<FUSE>)
    bool materialized = false;  // per forward: norm_out holds the normalised tensor
};

struct UpLayer {
    std::string name;           // "up{d}"
    int cin = 0, cout = 0, sh = 2, sw = 2, Hin = 0, Win = 0;
    Param w, b;                 // upstream: [Cin][Cout][kh][kw]
    float *w_direct = nullptr;  // [tap][Cin][Cout]
    void *w_tc = nullptr;       // [tap][Cout][Cin]
    const void *in = nullptr;
    void *out = nullptr;
    long long out_ps = 0;
    bool tc = false;
    int bn = 0, kc = 0, enc_n = 0;
    CUtensorMap tmA, tmB;
};

struct Op { int kind; int idx; };   // 0 = conv (+norm), 1 = up

}  // namespace ts2d

using namespace ts2d;

struct ts2d_unet {
    ts2d_unet_plan plan;
    int L = 0;
    std::vector<int> H, W, C;
    std::vector<ConvLayer> convs;
    std::vector<UpLayer> ups;
    std::vector<Op> ops;
    Param head_w, head_b;
    bool finalized = false;
    size_t es = 2;
    // workspace
    uint8_t *arena = nullptr;
    size_t arena_bytes = 0;
    std::vector<void *> skip, up;
    void *bott = nullptr, *raw = nullptr, *rawB = nullptr, *tmpA = nullptr, *tmpB = nullptr;
    float *partial = nullptr;     // split-K partial sums of the tiny levels [ksplit][N*HW][Cout] fp32
    size_t partial_bytes = 0;
    float *feat_ss = nullptr;     // [N][C0][2] scale / shift of the last decoder conv when its norm is deferred to the blend kernel
    double *stats_arena = nullptr;
    size_t stats_bytes = 0;
    void *features = nullptr;     // last decoder output [N][P][P][C0]
    int last_batch = 0;
    long long launches = 0;
    std::vector<std::pair<size_t, size_t>> guards;   // TS2D_ARENA_GUARD=bytes: (offset, bytes) of the canary bands between the workspace buffers
    int direct_layers = 0;        // layers of a reduced-precision tensor-core plan that fell to conv_direct_kernel (logged at finalize)
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    int prof_used = 0;
    struct ProfEntry { std::string name; int kind; double flops, bytes; };   // kind 0 conv, 1 up, 2 norm
    std::vector<ProfEntry> prof_entries;
};

namespace ts2d {

static void choose_tile(int Hout, int Wout, int N, int &tw, int &th, int &tn) {
    double best = 1e30;
    tw = 128; th = 1; tn = 1;
    for (int cw = 128; cw >= 1; cw >>= 1) {
        if (cw > 1 && cw / 2 >= Wout) continue;          // do not take a width with > 50% overhang when a smaller one fits
        int ch = 128 / cw;
        int hh = 1;
        while (hh < ch && hh < Hout) hh <<= 1;
        ch = std::min(ch, hh);
        const int cn = 128 / (cw * ch);
        const double covered = (double)ceil_div(Wout, cw) * cw * ceil_div(Hout, ch) * ch * ceil_div(N, cn) * cn;
        const double waste = covered / ((double)Wout * Hout * N);
        if (waste < best - 1e-9) { best = waste; tw = cw; th = ch; tn = cn; }
    }
}

// output-channel block of conv_tc_kernel from a cycle model of one CTA (measured on B200, tools/probe/mma_rate.cu:
// an M=128, K=16 tcgen05.mma costs ~45 / 51 / 67 / 131 cycles at N = 32 / 64 / 128 / 256; an epilogue warp needs
// ~400 cycles per 32-column chunk): time = waves x max(MMA time, epilogue time) per tile.
static int env_int(const char *name, int dflt) { const char *v = getenv(name); return v ? atoi(v) : dflt; }
// launch with programmatic stream serialisation (see common.cuh: every kernel launched through here calls griddep_wait() before it
// touches anything an earlier launch produces).  TS2D_PDL=0 falls back to plain stream order.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x, Args &&...args) {
    static const int pdl = env_int("TS2D_PDL", 1);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (cluster_x > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = (unsigned)cluster_x; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = attr; cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    return launch_pdl_cluster(kern, grid, block, smem, st, 1, std::forward<Args>(args)...);
}
// + kTileOvh cycles per tile (TMA latency of the first chunk, tmem_full / tmem_empty round trips): decides the 1-chunk tiles of the
// full-resolution transposed conv (BN = 32: 0.125 ms, BN = 128: one pass over A).
static int choose_bn(int n_total, long long m_tiles, int k_steps) {
    static const int ovh = env_int("TS2D_TILE_OVH", 600);
    int best_bn = 32;
    double best = 1e30;
    for (int bn : {256, 128, 64, 32}) {
        if (n_total % bn) continue;
        const double mma = bn == 256 ? 131 : bn == 128 ? 67 : bn == 64 ? 51 : 45;
        const double epi = bn == 32 ? 200 : 400.0 * (bn / 64);
        const long long tiles = m_tiles * (n_total / bn);
        const double t = (double)ceil_div(tiles, kNumSMs) * (std::max(k_steps * mma, epi) + ovh) + 0.02 * k_steps * mma * (n_total / bn);   // tie-break: fewer A re-reads
        if (t < best) { best = t; best_bn = bn; }
    }
    return best_bn;
}
// Tiny levels (H*W <= 256 per image).  Measured with clock stamps inside conv_tc_kernel (TS2D_TC_DBG,
// profiles/r01g_tiny_levels_clock_stamps.txt): (1) on the 4x4 / 8x8 levels a 128-pixel tile spans several images, the per-thread
// statistics accumulators cannot be used, and the epilogue's per-value double atomics cost 20-46 k cycles per launch; (2) a tile is a
// chain of 72-144 k-chunks at ~460 cycles each on 16-96 CTAs.  "Partial" mode: K is split over `ksplit` tap groups on separate CTAs,
// the epilogue stores fp32 accumulators, and splitk_finish_kernel adds the slices in a fixed order, takes the statistics and normalises
// (one launch instead of norm_act_kernel).  enc6.0: 51.7 -> 14.8 us, enc7.0: 35.2 -> 12.7 us, dec0.0: 67.6 -> 19.0 us.
// TS2D_KSPLIT (4x4 / 8x8 levels; default 9) and TS2D_KSPLIT16 (16x16 level, one image per tile; default 3):
// 0 = epilogue statistics + norm_act_kernel, 1 = partial mode without a split, 3 / 9 = split over tap rows / taps.
constexpr int kSplitMaxHW = 256;
static int tiny_ksplit() { return env_int("TS2D_KSPLIT", 9); }

// promo: CUtensorMapL2promotion (0 none, 1 64 B, 2 128 B, 3 256 B)
static int encode_map(CUtensorMap *tm, void *base, int rank, const cuuint64_t *dims, const cuuint64_t *strides_bytes, const cuuint32_t *box,
                      int kc, size_t es, bool is_bf16, int promo = 3) {
    PFN_encodeTiled fn = get_encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return TS2D_ERR_DEVICE; }
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    (void)es;
    const CUtensorMapSwizzle sw = kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUresult r = fn(tm, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank, base, dims,
                    strides_bytes, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, (CUtensorMapL2promotion)promo,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d)", (int)r, rank); return TS2D_ERR_CUDA; }
    return TS2D_OK;
}

// activation map for a conv input: stride 1 -> (C, W, H, N); strided -> parity view (px*ps + C, W/sw, sh, H/sh, N) with
// w = sw*w2 + px (px folded into the channel axis when sw == 2) and h = sh*h2 + py; stride2 = (sw == 2) | (sh == 2) << 1
static int encode_act_map(CUtensorMap *tm, const void *base, long long ps, int Cin, int Hin, int Win, int Nmax, int stride2, int kc, int tw, int th,
                          int tn, size_t es, bool is_bf16) {
    // a box that takes only part of every pixel (half of a concat buffer, or every other pixel of the parity view) must not be
    // promoted to 256-byte L2 requests: the other half is fetched from DRAM and never used (TS2D_L2PROMO_SUB, measured on enc1.0)
    static const int promo_full = env_int("TS2D_L2PROMO", 3), promo_sub = env_int("TS2D_L2PROMO_SUB", 3);
    const int promo = (stride2 || ps > Cin) ? promo_sub : promo_full;
    if (!stride2) {
        cuuint64_t dims[4] = {(cuuint64_t)Cin, (cuuint64_t)Win, (cuuint64_t)Hin, (cuuint64_t)Nmax};
        cuuint64_t str[3] = {(cuuint64_t)(ps * es), (cuuint64_t)(ps * es * Win), (cuuint64_t)(ps * es * Win * Hin)};
        cuuint32_t box[4] = {(cuuint32_t)kc, (cuuint32_t)tw, (cuuint32_t)th, (cuuint32_t)tn};
        return encode_map(tm, const_cast<void *>(base), 4, dims, str, box, kc, es, is_bf16, promo);
    }
    const int sw = (stride2 & 1) ? 2 : 1, sh = (stride2 & 2) ? 2 : 1;
    cuuint64_t dims[5] = {(cuuint64_t)(sw == 2 ? ps + Cin : Cin), (cuuint64_t)(Win / sw), (cuuint64_t)sh, (cuuint64_t)(Hin / sh), (cuuint64_t)Nmax};
    cuuint64_t str[4] = {(cuuint64_t)(sw * ps * es), (cuuint64_t)(ps * es * Win), (cuuint64_t)(sh * ps * es * Win), (cuuint64_t)(ps * es * Win * Hin)};
    cuuint32_t box[5] = {(cuuint32_t)kc, (cuuint32_t)tw, 1, (cuuint32_t)th, (cuuint32_t)tn};
    return encode_map(tm, const_cast<void *>(base), 5, dims, str, box, kc, es, is_bf16, promo);
}

static int encode_w_map(CUtensorMap *tm, const void *base, int Cin, int Cout, int taps, int kc, int bn, size_t es, bool is_bf16) {
    cuuint64_t dims[3] = {(cuuint64_t)Cin, (cuuint64_t)Cout, (cuuint64_t)taps};
    cuuint64_t str[2] = {(cuuint64_t)(Cin * es), (cuuint64_t)(Cin * es * Cout)};
    cuuint32_t box[3] = {(cuuint32_t)kc, (cuuint32_t)bn, 1};
    static const int promo_w = env_int("TS2D_L2PROMO_W", 3);
    return encode_map(tm, const_cast<void *>(base), 3, dims, str, box, kc, es, is_bf16, promo_w);
}

// cudaFuncSetAttribute acts on the current device's copy of the kernel, so "already set" has to be remembered per device ordinal (a
// process may drive several GPUs, and engines may be finalised from several threads): one bit per device in an atomic mask, set only
// after the calls succeeded.  Two threads that race both make the (idempotent) calls.  Ordinals >= 64 are never cached.
static inline bool smem_attr_needed(const std::atomic<uint64_t> &mask, int &dev) {
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { dev = -1; return true; }
    return ((mask.load(std::memory_order_acquire) >> dev) & 1ull) == 0;
}
#define TS2D_ONCE_PER_DEVICE(...)                                                           \
    do {                                                                                    \
        static std::atomic<uint64_t> _done{0};                                              \
        int _dev = -1;                                                                      \
        if (smem_attr_needed(_done, _dev)) {                                                \
            __VA_ARGS__;                                                                    \
            if (_dev >= 0) _done.fetch_or(1ull << _dev, std::memory_order_release);         \
        }                                                                                   \
    } while (0)

template <int BN, int KC, typename DT>
static int launch_tc_inst(const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvTcParams &p, cudaStream_t st) {
    constexpr int smem = tc_smem_bytes<BN, KC>();
    TS2D_ONCE_PER_DEVICE(TS2D_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<BN, KC, DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)));
    const int m_tiles = p.tiles_n * p.tiles_h * p.tiles_w;
    const int total = m_tiles * p.n_tiles_n * std::max(1, p.ksplit);
    if (p.mc) {
        // pairs of CTAs (clusters of 2) over pairs of M tiles
        static int max_clusters = -1;
        if (max_clusters < 0) {
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3(kNumSMs & ~1); cfg.blockDim = dim3(kTcThreads); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute a;
            a.id = cudaLaunchAttributeClusterDimension; a.val.clusterDim.x = 2; a.val.clusterDim.y = 1; a.val.clusterDim.z = 1;
            cfg.attrs = &a; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, conv_tc_kernel<BN, KC, DT>, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = kNumSMs / 2; }
            max_clusters = std::min(n, kNumSMs / 2);
        }
        const int pairs = std::max(1, std::min(total / 2, max_clusters));
        TS2D_CHECK_CUDA(launch_pdl_cluster(conv_tc_kernel<BN, KC, DT>, dim3(2 * pairs), dim3(kTcThreads), smem, st, 2, tmA, tmA2, tmB, p));
        return TS2D_OK;
    }
    const int gx = std::max(1, std::min(total, kNumSMs));
    TS2D_CHECK_CUDA(launch_pdl(conv_tc_kernel<BN, KC, DT>, dim3(gx), dim3(kTcThreads), smem, st, tmA, tmA2, tmB, p));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename DT>
static int launch_tc(int bn, int kc, const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvTcParams &p, cudaStream_t st) {
#define TS2D_TC_CASE(B, K) if (bn == B && kc == K) return launch_tc_inst<B, K, DT>(tmA, tmA2, tmB, p, st);
    TS2D_TC_CASE(32, 32) TS2D_TC_CASE(64, 32) TS2D_TC_CASE(128, 32) TS2D_TC_CASE(256, 32)
    TS2D_TC_CASE(32, 64) TS2D_TC_CASE(64, 64) TS2D_TC_CASE(128, 64) TS2D_TC_CASE(256, 64)
#undef TS2D_TC_CASE
    set_error("no tensor-core conv instantiation for BN=%d KC=%d", bn, kc);
    return TS2D_ERR_UNSUPPORTED;
}

template <int BN, int KC, typename DT>
static int launch_tc2_inst(const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvTcParams &p, cudaStream_t st) {
    constexpr int smem = tc2_smem_bytes<BN, KC>();
    TS2D_ONCE_PER_DEVICE(TS2D_CHECK_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<BN, KC, DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)));
    const int units = (p.tiles_n * p.tiles_h * p.tiles_w / 2) * p.n_tiles_n;
    const int gx = std::max(1, std::min(units, kNumSMs));
    TS2D_CHECK_CUDA(launch_pdl(conv_tc2_kernel<BN, KC, DT>, dim3(gx), dim3(kTcThreads), smem, st, tmA, tmA2, tmB, p));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename DT>
static int launch_tc2(int bn, const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvTcParams &p, cudaStream_t st) {
    if (bn == 256) return launch_tc2_inst<256, 64, DT>(tmA, tmA2, tmB, p, st);
    if (bn == 128) return launch_tc2_inst<128, 64, DT>(tmA, tmA2, tmB, p, st);
    set_error("no conv_tc2_kernel instantiation for BN=%d", bn);
    return TS2D_ERR_UNSUPPORTED;
}
// conv_tc2_kernel (two pixel tiles per weight block): output-channel block from a cycle model of one CTA - per k-chunk
// max(8 MMAs, TMA rows x 1.7 cycles); the BN = 256 variant exposes its epilogue (~3 k cycles per pair).  0 = not eligible.
static int choose_bn_tc2(int cout, long long m_tiles, int k_chunks) {
    const int on = env_int("TS2D_TC2", 1), min_units = env_int("TS2D_TC2_MIN_UNITS", kNumSMs / 2);      // read per call: the tests flip them
    if (!on || (m_tiles & 1)) return 0;
    int best_bn = 0;
    double best = 1e30;
    for (int bn : {256, 128}) {
        if (cout % bn) continue;
        const long long units = (m_tiles / 2) * (cout / bn);
        if (units < min_units) continue;
        const double per = bn == 256 ? k_chunks * 1048.0 + 3000.0 : k_chunks * 653.0 + 600.0;
        const double t = (double)ceil_div(units, kNumSMs) * per;
        if (t < best) { best = t; best_bn = bn; }
    }
    return best_bn;
}

constexpr int kHaloSmemBudget = 225 * 1024;
static int halo_smem_bytes(int cin, int bn, int kc, int stages) {
    const int w = ((9 * cin * bn * 2) + 1023) & ~1023;
    const int st = kc == 64 ? halo_stage_bytes<64>() : halo_stage_bytes<32>();
    (void)bn;
    return 1024 /*align*/ + w + stages * st + kHaloTailBytes;
}
template <int BN, int KC, typename DT, bool FUSE>
static int launch_halo_inst(const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvHaloParams &p, int smem, cudaStream_t st) {
    TS2D_ONCE_PER_DEVICE(TS2D_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel<BN, KC, DT, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHaloSmemBudget)));
    const int n_tiles_n = p.Cout / BN;
    const int total = p.tiles_x * p.tiles_y * p.N;
    const int gx = std::max(1, std::min(total, kNumSMs / n_tiles_n));
    TS2D_CHECK_CUDA(launch_pdl(conv_halo_kernel<BN, KC, DT, FUSE>, dim3(gx, n_tiles_n), dim3(halo_threads<FUSE>()), smem, st, tmA, tmA2, tmB, p));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename DT>
static int launch_halo(int bn, int kc, bool fuse, const CUtensorMap &tmA, const CUtensorMap &tmA2, const CUtensorMap &tmB, const ConvHaloParams &p, int smem, cudaStream_t st) {
#define TS2D_HALO_CASE(B, K)                                                              \
    if (bn == B && kc == K) return fuse ? launch_halo_inst<B, K, DT, true>(tmA, tmA2, tmB, p, smem, st) \
                                        : launch_halo_inst<B, K, DT, false>(tmA, tmA2, tmB, p, smem, st);
    TS2D_HALO_CASE(32, 32) TS2D_HALO_CASE(64, 32) TS2D_HALO_CASE(32, 64) TS2D_HALO_CASE(64, 64)
#undef TS2D_HALO_CASE
    set_error("no halo conv instantiation for BN=%d KC=%d", bn, kc);
    return TS2D_ERR_UNSUPPORTED;
}

static int halo_s2_smem_bytes(int cin, int bn, int kc, int stages) {
    const int w = ((9 * cin * bn * 2) + 1023) & ~1023;
    const int st = kc == 64 ? s2_stage_bytes<64>() : s2_stage_bytes<32>();
    return 1024 /*align*/ + w + stages * st + kHaloTailBytes;
}
template <int BN, int KC, typename DT>
static int launch_halo_s2_inst(const CUtensorMap &tmA, const CUtensorMap &tmB, const ConvHaloParams &p, int smem, cudaStream_t st) {
    TS2D_ONCE_PER_DEVICE(TS2D_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_s2_kernel<BN, KC, DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHaloSmemBudget)));
    const int n_tiles_n = p.Cout / BN;
    const int total = p.tiles_x * p.tiles_y * p.N;
    const int gx = std::max(1, std::min(total, kNumSMs / n_tiles_n));
    TS2D_CHECK_CUDA(launch_pdl(conv_halo_s2_kernel<BN, KC, DT>, dim3(gx, n_tiles_n), dim3(384), smem, st, tmA, tmB, p));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename DT>
static int launch_halo_s2(int bn, int kc, const CUtensorMap &tmA, const CUtensorMap &tmB, const ConvHaloParams &p, int smem, cudaStream_t st) {
    if (bn == 32 && kc == 32) return launch_halo_s2_inst<32, 32, DT>(tmA, tmB, p, smem, st);
    if (bn == 64 && kc == 32) return launch_halo_s2_inst<64, 32, DT>(tmA, tmB, p, smem, st);
    if (bn == 32 && kc == 64) return launch_halo_s2_inst<32, 64, DT>(tmA, tmB, p, smem, st);
    if (bn == 64 && kc == 64) return launch_halo_s2_inst<64, 64, DT>(tmA, tmB, p, smem, st);
    set_error("no stride-2 halo conv instantiation for BN=%d KC=%d", bn, kc);
    return TS2D_ERR_UNSUPPORTED;
}
// stride-2 halo kernel: the strided convs whose parity-plane halos (4 x 9 x 17 pixels per K-chunk) and resident weights fit in shared
// memory with >= 2 ring stages - the first strided stage of a plan (enc1.0 of CS-512: 32 -> 64 channels); deeper ones use conv_tc / tc2
static bool halo_s2_config(const ConvLayer &c, int &bn, int &stages) {
    const int enable = env_int("TS2D_CONV_HALO_S2", 1), max_nt = env_int("TS2D_HALO_MAX_NT", 2);
    if (!enable || !c.tc || c.kh != 3 || c.kw != 3 || c.sh != 2 || c.sw != 2 || c.in2) return false;
    if (c.Hout < 16 || c.Wout < 8) return false;
    const double covered = (double)ceil_div(c.Wout, 8) * 8 * ceil_div(c.Hout, 16) * 16;
    if (covered > 1.3 * c.Wout * c.Hout) return false;
    for (int b : {64, 32}) {
        if (c.cout % b || c.cout / b > max_nt) continue;
        int st = 6;
        while (st >= 2 && halo_s2_smem_bytes(c.cin, b, c.kc, st) > kHaloSmemBudget) --st;
        if (st < 2) continue;
        bn = b; stages = st;
        return true;
    }
    return false;
}

// picks the output-channel block and ring depth of the halo kernel, or returns false when the layer should use conv_tc_kernel
static bool halo_config(const ConvLayer &c, int &bn, int &stages) {
    const int enable = env_int("TS2D_CONV_HALO", 1), max_nt = env_int("TS2D_HALO_MAX_NT", 2);      // read at finalize time
    if (!enable || !c.tc || c.kh != 3 || c.kw != 3 || c.sh != 1 || c.sw != 1) return false;
    if (c.Hout < 16 || c.Wout < 8) return false;
    const double covered = (double)ceil_div(c.Wout, 8) * 8 * ceil_div(c.Hout, 16) * 16;
    if (covered > 1.3 * c.Wout * c.Hout) return false;
    for (int b : {64, 32}) {
        if (c.cout % b || c.cout / b > max_nt) continue;
        int st = 6;
        while (st >= 2 && halo_smem_bytes(c.cin, b, c.kc, st) > kHaloSmemBudget) --st;
        if (st < 2) continue;
        bn = b; stages = st;
        return true;
    }
    return false;
}

template <typename T> static int launch_direct(const DirectConvParams &p, int N, cudaStream_t st) {
    const int npix = p.Hout * p.Wout;
    int iters = npix >= 32 * 64 ? 8 : 1;
    DirectConvParams q = p;
    q.pix_iters = iters;
    dim3 grid(ceil_div(npix, 32 * iters), ceil_div(p.Cout, 32), N);
    conv_direct_kernel<T><<<grid, 128, 0, st>>>(q);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename T>
static int launch_first(int cin, const FirstConvParams &p, int N, cudaStream_t st) {
    const int CG = p.Cout / 16, rows = (256 / CG) / 16;
    dim3 grid(ceil_div(p.W, 64), ceil_div(p.H, rows * kFirstConvIters), N);
    const size_t smem = (size_t)(9 * cin * p.Cout + p.Cout + 8 * p.Cout * 2) * sizeof(float);
    switch (cin) {
        case 1: conv_first_kernel<T, 1><<<grid, 256, smem, st>>>(p); break;
        case 2: conv_first_kernel<T, 2><<<grid, 256, smem, st>>>(p); break;
        case 3: conv_first_kernel<T, 3><<<grid, 256, smem, st>>>(p); break;
        case 4: conv_first_kernel<T, 4><<<grid, 256, smem, st>>>(p); break;
        default: set_error("conv_first_kernel: unsupported input channel count %d", cin); return TS2D_ERR_UNSUPPORTED;
    }
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename DT>
static int launch_first_tc(const FirstTcParams &p, cudaStream_t st) {
    TS2D_ONCE_PER_DEVICE(TS2D_CHECK_CUDA(cudaFuncSetAttribute(conv_first_tc_kernel<DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFtcSmemBytes)));
    const long long total = (long long)p.tiles_x * p.tiles_y * p.N;
    const int gx = (int)std::max(1LL, std::min(total / 2, (long long)kNumSMs * 2));
    TS2D_CHECK_CUDA(launch_pdl(conv_first_tc_kernel<DT>, dim3(gx, p.Cout / 32), dim3(kFtcThreads), kFtcSmemBytes, st, p));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
// tensor-core first conv (conv_first_tc_kernel): two 16-bit input channels, output channels in blocks of 32
static bool first_tc_eligible(const ts2d_unet_plan &pl, const ConvLayer &c) {
    const int enable = getenv("TS2D_FIRST_TC") ? atoi(getenv("TS2D_FIRST_TC")) : 1;
    return enable && pl.use_tensor_cores && pl.precision != TS2D_PREC_F32 && c.kh == 3 && c.kw == 3 && c.sh == 1 && c.sw == 1 && c.cin == 2 &&
           c.cout % 32 == 0 && c.in_ps == c.cin;
}
static bool first_fast_eligible(const ts2d_unet_plan &pl, const ConvLayer &c) {
    return pl.use_tensor_cores && pl.precision != TS2D_PREC_F32 && c.kh == 3 && c.kw == 3 && c.sh == 1 && c.sw == 1 && c.cin <= 4 &&
           (c.cout == 16 || c.cout == 32 || c.cout == 64) && c.in_ps == c.cin;
}

static int launch_direct_any(int precision, const DirectConvParams &p, int N, cudaStream_t st) {
    if (precision == TS2D_PREC_F32) return launch_direct<float>(p, N, st);
    if (precision == TS2D_PREC_BF16) return launch_direct<__nv_bfloat16>(p, N, st);
    return launch_direct<__half>(p, N, st);
}

template <typename T>
static int launch_norm(const void *raw, const double *stats, const float *g, const float *b, void *out, long long out_ps, int HW, int C, float eps,
                       float slope, int N, cudaStream_t st) {
    // <= 4 blocks per SM in total: the whole grid is resident at once (no second partial wave)
    int ppb = std::max(32, ceil_div(HW, std::max(1, (kNumSMs * 4) / std::max(1, N))));
    dim3 grid(ceil_div(HW, ppb), N);
    TS2D_CHECK_CUDA(launch_pdl(norm_act_kernel<T>, grid, dim3(256), 2 * C * sizeof(float), st, (const T *)raw, stats, g, b, (T *)out, out_ps, HW, C, eps, slope, ppb));
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
static int launch_norm_any(int precision, const void *raw, const double *stats, const float *g, const float *b, void *out, long long out_ps, int HW,
                           int C, float eps, float slope, int N, cudaStream_t st) {
    if (precision == TS2D_PREC_F32) return launch_norm<float>(raw, stats, g, b, out, out_ps, HW, C, eps, slope, N, st);
    if (precision == TS2D_PREC_BF16) return launch_norm<__nv_bfloat16>(raw, stats, g, b, out, out_ps, HW, C, eps, slope, N, st);
    return launch_norm<__half>(raw, stats, g, b, out, out_ps, HW, C, eps, slope, N, st);
}

static int upload(Param &p) {
    if (!p.have) return TS2D_OK;
    if (!p.dev) TS2D_CHECK_CUDA(cudaMalloc(&p.dev, p.host.size() * sizeof(float)));
    TS2D_CHECK_CUDA(cudaMemcpy(p.dev, p.host.data(), p.host.size() * sizeof(float), cudaMemcpyHostToDevice));
    return TS2D_OK;
}

// upstream conv weight OIHW -> direct [tap][Cin][Cout] fp32 and tc [tap][Cout][Cin] DT
static int repack_conv(const std::vector<float> &w, int cout, int cin, int taps, bool transposed_src, int precision, bool want_tc, float **w_direct,
                       void **w_tc) {
    std::vector<float> d((size_t)taps * cin * cout);
    for (int o = 0; o < cout; ++o)
        for (int i = 0; i < cin; ++i)
            for (int t = 0; t < taps; ++t) {
                // conv: [o][i][t]; transposed conv: [i][o][t]
                const float v = transposed_src ? w[((size_t)i * cout + o) * taps + t] : w[((size_t)o * cin + i) * taps + t];
                d[((size_t)t * cin + i) * cout + o] = v;
            }
    if (!*w_direct) TS2D_CHECK_CUDA(cudaMalloc(w_direct, d.size() * sizeof(float)));
    TS2D_CHECK_CUDA(cudaMemcpy(*w_direct, d.data(), d.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (want_tc) {
        std::vector<uint16_t> h((size_t)taps * cin * cout);
        for (int o = 0; o < cout; ++o)
            for (int i = 0; i < cin; ++i)
                for (int t = 0; t < taps; ++t) {
                    const float v = d[((size_t)t * cin + i) * cout + o];
                    uint16_t bits;
                    if (precision == TS2D_PREC_BF16) { __nv_bfloat16 x = __float2bfloat16_rn(v); memcpy(&bits, &x, 2); }
                    else { __half x = __float2half_rn(v); memcpy(&bits, &x, 2); }
                    h[((size_t)t * cout + o) * cin + i] = bits;
                }
        if (!*w_tc) TS2D_CHECK_CUDA(cudaMalloc(w_tc, h.size() * 2));
        TS2D_CHECK_CUDA(cudaMemcpy(*w_tc, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    }
    return TS2D_OK;
}

static bool tc_eligible(const ts2d_unet_plan &pl, int cin, int cout, int kh, int kw, int sh, int sw, int Hin, int Win, long long in_ps) {
    if (!pl.use_tensor_cores || pl.precision == TS2D_PREC_F32) return false;
    if (cin % 32 || cout % 32 || in_ps % 8) return false;
    if (!((kh == 3 && kw == 3) || (kh == 1 && kw == 1))) return false;
    // strides 1 / 2 per axis (anisotropic plans included); ts2d_unet_create guarantees Hin % sh == 0 and Win % sw == 0
    return (sh == 1 || sh == 2) && (sw == 1 || sw == 2) && Hin % sh == 0 && Win % sw == 0;
}

static void fill_tc_params(ConvTcParams &p, int N, int Hout, int Wout, int cout, int cin, int kc, int bn, int taps, int kw, int pad_h, int pad_w,
                           int stride2, long long in_ps) {
    memset(&p, 0, sizeof(p));
    p.N = N; p.Hout = Hout; p.Wout = Wout; p.Cout = cout;
    p.cin_chunks = cin / kc;
    p.split_chunks = p.cin_chunks;
    p.taps = taps; p.kw = kw; p.pad_h = pad_h; p.pad_w = pad_w;
    p.stride2 = stride2;
    p.in_ps = (int)in_ps;
    p.n_tiles_n = cout / bn;
    p.tap_cout = cout;
    p.my = 1; p.mx = 1; p.Hfull = Hout; p.Wfull = Wout;
}

}  // namespace ts2d

// =========================================================================================================
extern "C" int ts2d_unet_create(const ts2d_unet_plan *pl, ts2d_unet **out) {
    TS2D_REQUIRE(pl && out, "ts2d_unet_create: null pointer");
    TS2D_REQUIRE(pl->n_stages >= 2 && pl->n_stages <= TS2D_MAX_STAGES, "ts2d_unet_create: n_stages must be 2..%d", TS2D_MAX_STAGES);
    TS2D_REQUIRE(pl->in_channels >= 1 && pl->num_heads >= 1 && pl->max_batch >= 1, "ts2d_unet_create: bad channels/heads/batch");
    TS2D_REQUIRE(pl->precision >= 0 && pl->precision <= 2, "ts2d_unet_create: bad precision");
    TS2D_REQUIRE(pl->patch_h > 0 && pl->patch_w > 0, "ts2d_unet_create: bad patch size");
    const int L = pl->n_stages;
    ts2d_unet *u = new ts2d_unet();
    u->plan = *pl;
    u->L = L;
    u->es = prec_size(pl->precision);
    u->H.resize(L); u->W.resize(L); u->C.resize(L);
    int h = pl->patch_h, w = pl->patch_w;
    for (int s = 0; s < L; ++s) {
        const int kh = pl->kernel_sizes[s][0], kw = pl->kernel_sizes[s][1], sh = pl->strides[s][0], sw = pl->strides[s][1];
        if (!(kh % 2 == 1 && kw % 2 == 1 && sh >= 1 && sw >= 1 && pl->features[s] >= 1 && pl->n_conv_enc[s] >= 1)) {
            delete u; set_error("ts2d_unet_create: bad stage %d description", s); return TS2D_ERR_ARG;
        }
        if (h % sh || w % sw) { delete u; set_error("ts2d_unet_create: patch size is not divisible by the strides at stage %d", s); return TS2D_ERR_ARG; }
        h /= sh; w /= sw;
        u->H[s] = h; u->W[s] = w; u->C[s] = pl->features[s];
    }
    // encoder
    int cin = pl->in_channels, hin = pl->patch_h, win = pl->patch_w;
    for (int s = 0; s < L; ++s) {
        for (int i = 0; i < pl->n_conv_enc[s]; ++i) {
            ConvLayer c;
            c.name = "enc" + std::to_string(s) + "." + std::to_string(i);
            c.cin = cin; c.cout = u->C[s];
            c.kh = pl->kernel_sizes[s][0]; c.kw = pl->kernel_sizes[s][1];
            c.sh = i == 0 ? pl->strides[s][0] : 1; c.sw = i == 0 ? pl->strides[s][1] : 1;
            c.Hin = hin; c.Win = win; c.Hout = u->H[s]; c.Wout = u->W[s];
            c.first = (s == 0 && i == 0);
            u->ops.push_back({0, (int)u->convs.size()});
            u->convs.push_back(c);
            cin = u->C[s]; hin = u->H[s]; win = u->W[s];
        }
    }
    // decoder
    for (int d = 0; d < L - 1; ++d) {
        const int l = L - 2 - d;
        if (pl->n_conv_dec[d] < 1) { delete u; set_error("ts2d_unet_create: n_conv_dec[%d] < 1", d); return TS2D_ERR_ARG; }
        UpLayer up;
        up.name = "up" + std::to_string(d);
        up.cin = u->C[l + 1]; up.cout = u->C[l];
        up.sh = pl->strides[l + 1][0]; up.sw = pl->strides[l + 1][1];
        up.Hin = u->H[l + 1]; up.Win = u->W[l + 1];
        u->ops.push_back({1, (int)u->ups.size()});
        u->ups.push_back(up);
        int ci = 2 * u->C[l];
        for (int i = 0; i < pl->n_conv_dec[d]; ++i) {
            ConvLayer c;
            c.name = "dec" + std::to_string(d) + "." + std::to_string(i);
            c.cin = ci; c.cout = u->C[l];
            c.kh = pl->kernel_sizes[l][0]; c.kw = pl->kernel_sizes[l][1];
            c.Hin = c.Hout = u->H[l]; c.Win = c.Wout = u->W[l];
            u->ops.push_back({0, (int)u->convs.size()});
            u->convs.push_back(c);
            ci = u->C[l];
        }
    }
    *out = u;
    return TS2D_OK;
}

static ConvLayer *find_conv(ts2d_unet *u, const std::string &name) {
    for (auto &c : u->convs) if (c.name == name) return &c;
    return nullptr;
}

static int set_param(Param &p, const float *data, int64_t count, int64_t expect, const char *name) {
    if (count != expect) { set_error("ts2d_unet_set_weight: %s has %lld elements, expected %lld", name, (long long)count, (long long)expect); return TS2D_ERR_ARG; }
    p.host.assign(data, data + count);
    p.have = true;
    return TS2D_OK;
}

extern "C" int ts2d_unet_set_weight(ts2d_unet *u, const char *name, const float *data, int64_t count) {
    TS2D_REQUIRE(u && name && data, "ts2d_unet_set_weight: null pointer");
    std::string nm(name);
    if (nm.find("all_modules") != std::string::npos || nm.rfind("decoder.encoder.", 0) == 0) return 1;
    int s = 0, i = 0, d = 0;
    char tail[64];
    const bool is_enc = sscanf(name, "encoder.stages.%d.0.convs.%d.%63s", &s, &i, tail) == 3;
    const bool is_dec = !is_enc && sscanf(name, "decoder.stages.%d.convs.%d.%63s", &s, &i, tail) == 3;
    if (is_enc || is_dec) {
        ConvLayer *c = find_conv(u, std::string(is_enc ? "enc" : "dec") + std::to_string(s) + "." + std::to_string(i));
        if (!c) { set_error("ts2d_unet_set_weight: no layer for %s", name); return TS2D_ERR_ARG; }
        const std::string t(tail);
        if (t == "conv.weight") return set_param(c->w, data, count, (int64_t)c->cout * c->cin * c->kh * c->kw, name);
        if (t == "conv.bias") return set_param(c->b, data, count, c->cout, name);
        if (t == "norm.weight") return set_param(c->gamma, data, count, c->cout, name);
        if (t == "norm.bias") return set_param(c->beta, data, count, c->cout, name);
        return 1;
    }
    if (sscanf(name, "decoder.transpconvs.%d.%63s", &d, tail) == 2) {
        if (d < 0 || d >= (int)u->ups.size()) { set_error("ts2d_unet_set_weight: no transposed conv %d", d); return TS2D_ERR_ARG; }
        UpLayer &up = u->ups[d];
        const std::string t(tail);
        if (t == "weight") return set_param(up.w, data, count, (int64_t)up.cin * up.cout * up.sh * up.sw, name);
        if (t == "bias") return set_param(up.b, data, count, up.cout, name);
        return 1;
    }
    if (sscanf(name, "decoder.seg_layers.%d.%63s", &d, tail) == 2) {
        if (d != u->L - 2) return 1;   // deep-supervision heads are unused at inference
        const std::string t(tail);
        if (t == "weight") return set_param(u->head_w, data, count, (int64_t)u->plan.num_heads * u->C[0], name);
        if (t == "bias") return set_param(u->head_b, data, count, u->plan.num_heads, name);
        return 1;
    }
    return 1;
}

extern "C" int ts2d_unet_finalize(ts2d_unet *u) {
    TS2D_REQUIRE(u, "ts2d_unet_finalize: null handle");
    const ts2d_unet_plan &pl = u->plan;
    const int L = u->L, N = pl.max_batch;
    const size_t es = u->es;
    for (auto &c : u->convs)
        if (!c.w.have) { set_error("ts2d_unet_finalize: missing weight for %s", c.name.c_str()); return TS2D_ERR_STATE; }
    for (auto &up : u->ups)
        if (!up.w.have) { set_error("ts2d_unet_finalize: missing weight for %s", up.name.c_str()); return TS2D_ERR_STATE; }
    if (!u->head_w.have) { set_error("ts2d_unet_finalize: missing seg head weight (decoder.seg_layers.%d.weight)", L - 2); return TS2D_ERR_STATE; }
    int dev = 0;
    cudaDeviceProp prop;
    TS2D_CHECK_CUDA(cudaGetDevice(&dev));
    TS2D_CHECK_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) { set_error("ts2d_unet_finalize: device %d is sm_%d%d, this library contains sm_100a code only", dev, prop.major, prop.minor); return TS2D_ERR_DEVICE; }

    // ---- workspace ----
    // TS2D_ARENA_GUARD=n (debug; compute-sanitizer is not available on the GPU pool): n bytes of 0xA5 after every workspace buffer,
    // verified by ts2d_unet_check_guards() - an out-of-bounds store of any kernel into a neighbouring buffer shows up as a damaged band
    const size_t guard = ((size_t)std::max(0, env_int("TS2D_ARENA_GUARD", 0)) + 1023) & ~(size_t)1023;
    u->guards.clear();
    size_t off = 0;
    auto align = [&](size_t v) {
        const size_t a = (v + 1023) & ~(size_t)1023;
        if (guard) u->guards.push_back({off + a, guard});
        return a + guard;
    };
    std::vector<size_t> skip_off(L - 1), up_off(L - 1);
    for (int l = 0; l < L - 1; ++l) {
        skip_off[l] = off; off += align((size_t)N * u->H[l] * u->W[l] * u->C[l] * es);
        up_off[l] = off; off += align((size_t)N * u->H[l] * u->W[l] * u->C[l] * es);
    }
    const size_t bott_off = off; off += align((size_t)N * u->H[L - 1] * u->W[L - 1] * u->C[L - 1] * es);
    size_t max_act = 0;
    for (auto &c : u->convs) max_act = std::max(max_act, (size_t)N * c.Hout * c.Wout * c.cout * es);
    const size_t raw_off = off; off += align(max_act);
    const size_t rawb_off = off; off += align(max_act);
    const size_t ss_off = off; off += align((size_t)N * u->C[0] * 2 * sizeof(float));
    const size_t a_off = off; off += align(max_act);
    const size_t b_off = off; off += align(max_act);
    size_t max_partial = 0;
    for (auto &c : u->convs)
        if (c.Hout * c.Wout <= kSplitMaxHW && c.kh * c.kw == 9) max_partial = std::max(max_partial, (size_t)9 * N * c.Hout * c.Wout * c.cout * sizeof(float));
    const size_t partial_off = off; off += align(max_partial);
    const size_t stats_off = off;
    size_t sbytes = 0;
    auto align0 = [](size_t v) { return (v + 1023) & ~(size_t)1023; };      // the statistics block is one buffer (cleared by one memset per forward): no bands inside
    for (auto &c : u->convs) sbytes += align0((size_t)N * c.cout * 2 * sizeof(double));
    off += sbytes;
    if (guard) { u->guards.push_back({off, guard}); off += guard; }
    if (u->arena) cudaFree(u->arena);
    TS2D_CHECK_CUDA(cudaMalloc(&u->arena, off));
    TS2D_CHECK_CUDA(cudaMemset(u->arena, 0, off));
    for (auto &g : u->guards) TS2D_CHECK_CUDA(cudaMemset(u->arena + g.first, 0xA5, g.second));
    u->arena_bytes = off;
    u->skip.resize(L - 1); u->up.resize(L - 1);
    for (int l = 0; l < L - 1; ++l) { u->skip[l] = u->arena + skip_off[l]; u->up[l] = u->arena + up_off[l]; }
    u->bott = u->arena + bott_off; u->raw = u->arena + raw_off; u->rawB = u->arena + rawb_off;
    u->feat_ss = reinterpret_cast<float *>(u->arena + ss_off); u->tmpA = u->arena + a_off; u->tmpB = u->arena + b_off;
    u->partial = max_partial ? reinterpret_cast<float *>(u->arena + partial_off) : nullptr;
    u->partial_bytes = max_partial;
    u->stats_arena = reinterpret_cast<double *>(u->arena + stats_off);
    u->stats_bytes = sbytes;
    size_t so = 0;
    for (auto &c : u->convs) { c.stats = reinterpret_cast<double *>(u->arena + stats_off + so); so += align0((size_t)N * c.cout * 2 * sizeof(double)); }

    // ---- wiring ----
    size_t ci = 0;
    const void *cur = nullptr; long long cur_ps = pl.in_channels;
    for (int s = 0; s < L; ++s) {
        for (int i = 0; i < pl.n_conv_enc[s]; ++i, ++ci) {
            ConvLayer &c = u->convs[ci];
            c.in = cur; c.in_ps = cur_ps; c.in2 = nullptr; c.in2_ps = 0; c.c_split = c.cin;
            const bool last = i == pl.n_conv_enc[s] - 1;
            if (last) {
                if (s < L - 1) { c.norm_out = u->skip[s]; c.out_ps = u->C[s]; }
                else { c.norm_out = u->bott; c.out_ps = u->C[s]; }
            } else { c.norm_out = u->tmpA; c.out_ps = u->C[s]; }
            cur = c.norm_out; cur_ps = c.out_ps;
        }
    }
    const void *lres = u->bott;
    for (int d = 0; d < L - 1; ++d) {
        const int l = L - 2 - d;
        UpLayer &up = u->ups[d];
        up.in = lres; up.out = u->up[l]; up.out_ps = u->C[l];
        cur = u->up[l]; cur_ps = u->C[l];
        for (int i = 0; i < pl.n_conv_dec[d]; ++i, ++ci) {
            ConvLayer &c = u->convs[ci];
            c.in = cur; c.in_ps = cur_ps; c.in2 = nullptr; c.in2_ps = 0; c.c_split = c.cin;
            if (i == 0) { c.in2 = u->skip[l]; c.in2_ps = u->C[l]; c.c_split = u->C[l]; }      // cat([up, skip]) as two K ranges
            const bool last = i == pl.n_conv_dec[d] - 1;
            c.norm_out = last ? u->tmpB : u->tmpA; c.out_ps = u->C[l];
            cur = c.norm_out; cur_ps = c.out_ps;
        }
        lres = u->tmpB;
    }
    u->features = u->tmpB;

    // ---- parameters ----
    int rc;
    for (auto &c : u->convs) {
        c.tc = tc_eligible(pl, c.cin, c.cout, c.kh, c.kw, c.sh, c.sw, c.Hin, c.Win, c.in_ps) && c.c_split % 32 == 0 && (c.in2 == nullptr || c.in2_ps % 8 == 0);
        if ((rc = repack_conv(c.w.host, c.cout, c.cin, c.kh * c.kw, false, pl.precision, c.tc, &c.w_direct, &c.w_tc))) return rc;
        if ((rc = upload(c.b)) || (rc = upload(c.gamma)) || (rc = upload(c.beta))) return rc;
        if (c.tc) c.kc = (c.c_split % 64 == 0 && (c.cin - c.c_split) % 64 == 0) ? 64 : 32;      // a K-chunk never straddles the two inputs
        c.enc_n = 0;
        c.halo = halo_config(c, c.halo_bn, c.halo_stages);
        c.halo_s2 = !c.halo && halo_s2_config(c, c.halo_bn, c.halo_stages);
        c.halo_maps = false;
        c.first_tc = !c.tc && first_tc_eligible(pl, c);
        c.first_fast = !c.tc && !c.first_tc && first_fast_eligible(pl, c);
        c.defer = false; c.prod = -1;
    }
    // ---- deferred normalisation: producer writes raw + statistics only, the consumer normalises while loading ----
    // TS2D_FUSE_NORM: 0 = every layer runs norm_act_kernel; 1 = the last decoder conv is normalised inside the blend kernel;
    // 2 = additionally every conv_halo_kernel<FUSE> consumer (bit-identical; the transform warps rewrite the halo in shared memory);
    // 3 (default) = like 2 but only consumers with 32 output channels per CTA, the only ones where the fused kernel beats
    // conv + norm pass (profiles/r02m_halo_bench_fuse.txt: 512^2 32->32 0.168 ms against 0.111 + 0.079; 64-channel blocks break even)
    {
        const int fuse = getenv("TS2D_FUSE_NORM") ? atoi(getenv("TS2D_FUSE_NORM")) : 3;
        for (size_t i = 0; i < u->convs.size(); ++i) u->convs[i].out_raw = (i & 1) ? u->rawB : u->raw;
        size_t k = 0;
        auto stage_pass = [&](int n_convs) {
            for (int i = 0; i + 1 < n_convs; ++i) {
                ConvLayer &prod = u->convs[k + i], &cons = u->convs[k + i + 1];
                // the transform computes leaky(y) as max(y, y * slope): exact for slopes in [0, 1] only
                if (fuse >= 2 && cons.halo && pl.precision != TS2D_PREC_F32 && cons.cin <= kHaloMaxCin && pl.negative_slope >= 0.f && pl.negative_slope <= 1.f &&
                    (fuse == 2 || cons.halo_bn == 32)) {
                    prod.defer = true;
                    cons.prod = (int)(k + i);
                    cons.in = prod.out_raw; cons.in_ps = prod.cout;
                }
            }
            k += n_convs;
        };
        for (int s = 0; s < L; ++s) stage_pass(pl.n_conv_enc[s]);
        for (int d = 0; d < L - 1; ++d) stage_pass(pl.n_conv_dec[d]);
        u->convs.back().defer = fuse != 0;      // the blend kernel applies the last InstanceNorm + LeakyReLU
    }
    for (auto &up : u->ups) {
        up.tc = tc_eligible(pl, up.cin, up.cout, 1, 1, 1, 1, up.Hin, up.Win, up.cin);
        if ((rc = repack_conv(up.w.host, up.cout, up.cin, up.sh * up.sw, true, pl.precision, up.tc, &up.w_direct, &up.w_tc))) return rc;
        if ((rc = upload(up.b))) return rc;
        if (up.tc) up.kc = (up.cin % 64 == 0) ? 64 : 32;
        up.enc_n = 0;
    }
    if ((rc = upload(u->head_w)) || (rc = upload(u->head_b))) return rc;
    // a reduced-precision tensor-core plan must not fall to the CUDA-core parity kernel silently (it is 10-50x slower)
    u->direct_layers = 0;
    if (pl.use_tensor_cores && pl.precision != TS2D_PREC_F32) {
        for (auto &c : u->convs)
            if (!c.tc && !c.first_tc && !c.first_fast) {
                ++u->direct_layers;
                fprintf(stderr, "[ts2d] warning: %s (Cin %d, Cout %d, kernel %dx%d, stride %dx%d, input %dx%d) has no tcgen05 kernel: it runs on conv_direct_kernel\n",
                        c.name.c_str(), c.cin, c.cout, c.kh, c.kw, c.sh, c.sw, c.Hin, c.Win);
            }
        for (auto &up : u->ups)
            if (!up.tc) {
                ++u->direct_layers;
                fprintf(stderr, "[ts2d] warning: %s (transposed conv, Cin %d, Cout %d, stride %dx%d) has no tcgen05 kernel: it runs on conv_direct_kernel\n",
                        up.name.c_str(), up.cin, up.cout, up.sh, up.sw);
            }
    }
    u->finalized = true;
    u->last_batch = 0;
    return TS2D_OK;
}

extern "C" size_t ts2d_unet_workspace_bytes(const ts2d_unet *u) { return u ? u->arena_bytes : 0; }
extern "C" int64_t ts2d_unet_launch_count(const ts2d_unet *u) { return u ? u->launches : 0; }
extern "C" int ts2d_unet_guard_count(const ts2d_unet *u) { return u ? (int)u->guards.size() : 0; }
extern "C" int ts2d_unet_check_guards(ts2d_unet *u) {
    if (!u || !u->arena) return -1;
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    int bad = 0;
    std::vector<uint8_t> h;
    for (auto &g : u->guards) {
        h.resize(g.second);
        if (cudaMemcpy(h.data(), u->arena + g.first, g.second, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
        for (size_t i = 0; i < g.second; ++i)
            if (h[i] != 0xA5) { ++bad; fprintf(stderr, "[ts2d] guard band at workspace offset %zu damaged at byte %zu\n", g.first, i); break; }
    }
    return bad;
}
extern "C" int ts2d_unet_direct_layer_count(const ts2d_unet *u) { return u ? u->direct_layers : 0; }

extern "C" double ts2d_unet_flops(const ts2d_unet *u, int batch) {
    if (!u) return 0.0;
    double fl = 0.0;
    for (auto &c : u->convs) fl += 2.0 * c.Hout * c.Wout * c.kh * c.kw * (double)c.cin * c.cout;
    for (auto &up : u->ups) fl += 2.0 * (double)up.Hin * up.sh * up.Win * up.sw * up.cin * up.cout;
    fl += 2.0 * (double)u->plan.patch_h * u->plan.patch_w * u->C[0] * u->plan.num_heads;
    return fl * batch;
}

namespace ts2d {

static int prof_begin(ts2d_unet *u, cudaStream_t st) {
    if (!u->profiling) return TS2D_OK;
    if (u->prof_used >= (int)u->prof_events.size()) {
        cudaEvent_t a, b;
        TS2D_CHECK_CUDA(cudaEventCreate(&a));
        TS2D_CHECK_CUDA(cudaEventCreate(&b));
        u->prof_events.push_back({a, b});
    }
    TS2D_CHECK_CUDA(cudaEventRecord(u->prof_events[u->prof_used].first, st));
    return TS2D_OK;
}
static int prof_end(ts2d_unet *u, const std::string &name, int kind, double flops, double bytes, cudaStream_t st) {
    if (!u->profiling) return TS2D_OK;
    TS2D_CHECK_CUDA(cudaEventRecord(u->prof_events[u->prof_used].second, st));
    if ((int)u->prof_entries.size() <= u->prof_used) u->prof_entries.resize(u->prof_used + 1);
    u->prof_entries[u->prof_used] = {name, kind, flops, bytes};
    u->prof_used++;
    return TS2D_OK;
}

template <typename T>
static int launch_splitk_finish(int S, dim3 grid, const float *partial, const float *bias, const float *gamma, const float *beta, void *out, long long out_ps,
                                int HW, int C, long long mtot, float eps, float slope, cudaStream_t st) {
#define TS2D_SKF(SS, JJ) TS2D_CHECK_CUDA(launch_pdl(splitk_finish_kernel<T, SS, JJ>, grid, dim3(256), 0, st, partial, bias, gamma, beta, (T *)out, out_ps, HW, C, mtot, eps, slope))
    const int J = HW <= 16 ? 2 : HW <= 64 ? 8 : 32;
    if (S == 1) { if (J == 2) TS2D_SKF(1, 2); else if (J == 8) TS2D_SKF(1, 8); else TS2D_SKF(1, 32); }
    else if (S == 3) { if (J == 2) TS2D_SKF(3, 2); else if (J == 8) TS2D_SKF(3, 8); else TS2D_SKF(3, 32); }
    else if (S == 9) { if (J == 2) TS2D_SKF(9, 2); else if (J == 8) TS2D_SKF(9, 8); else TS2D_SKF(9, 32); }
    else { set_error("splitk_finish_kernel: unsupported slice count %d", S); return TS2D_ERR_UNSUPPORTED; }
#undef TS2D_SKF
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

static int run_norm(ts2d_unet *u, ConvLayer &c, int N, cudaStream_t st);

static int run_conv(ts2d_unet *u, ConvLayer &c, int N, cudaStream_t st) {
    const ts2d_unet_plan &pl = u->plan;
    const size_t es = u->es;
    const bool is_bf16 = pl.precision == TS2D_PREC_BF16;
    int rc;
    if ((rc = prof_begin(u, st))) return rc;
    if (c.halo_s2) {
        ConvHaloParams p;
        memset(&p, 0, sizeof(p));
        p.N = N; p.H = c.Hout; p.W = c.Wout; p.Cout = c.cout; p.n_base = 0;
        p.cin_chunks = c.cin / c.kc; p.split_chunks = p.cin_chunks;
        p.tiles_x = ceil_div(c.Wout, 8); p.tiles_y = ceil_div(c.Hout, 16);
        p.n_stages = c.halo_stages;
        p.out_ps = c.cout; p.in_ps = (int)c.in_ps; p.out = c.out_raw; p.bias = c.b.dev; p.stats = c.stats;
        p.Cin = c.cin;
        if (c.first || !c.halo_maps) {
            if ((rc = encode_act_map(&c.tmA, c.in, c.in_ps, c.cin, c.Hin, c.Win, pl.max_batch, 3, c.kc, kS2W, kS2H, 1, es, is_bf16))) return rc;
            if ((rc = encode_w_map(&c.tmB, c.w_tc, c.cin, c.cout, 9, c.kc, c.halo_bn, es, is_bf16))) return rc;
            c.halo_maps = true;
        }
        const int smem = halo_s2_smem_bytes(c.cin, c.halo_bn, c.kc, c.halo_stages);
        rc = is_bf16 ? launch_halo_s2<__nv_bfloat16>(c.halo_bn, c.kc, c.tmA, c.tmB, p, smem, st)
                     : launch_halo_s2<__half>(c.halo_bn, c.kc, c.tmA, c.tmB, p, smem, st);
        if (rc) return rc;
    } else if (c.halo) {
        ConvHaloParams p;
        memset(&p, 0, sizeof(p));
        p.N = N; p.H = c.Hout; p.W = c.Wout; p.Cout = c.cout; p.n_base = 0;
        p.cin_chunks = c.cin / c.kc;
        p.split_chunks = c.c_split / c.kc;
        p.tiles_x = ceil_div(c.Wout, 8); p.tiles_y = ceil_div(c.Hout, 16);
        p.n_stages = c.halo_stages;
        p.out_ps = c.cout; p.out = c.out_raw; p.bias = c.b.dev; p.stats = c.stats;
        p.Cin = c.cin;
        const bool fuse = c.prod >= 0;
        if (fuse) {
            const ConvLayer &pr = u->convs[c.prod];
            p.in_stats = pr.stats; p.in_gamma = pr.gamma.dev; p.in_beta = pr.beta.dev; p.in_hw = pr.Hout * pr.Wout;
            p.in_eps = pl.eps; p.in_slope = pl.negative_slope;
        }
        if (c.first || !c.halo_maps) {
            if ((rc = encode_act_map(&c.tmA, c.in, c.in_ps, c.c_split, c.Hin, c.Win, pl.max_batch, 0, c.kc, kHaloW, kHaloH, 1, es, is_bf16))) return rc;
            if (c.in2) { if ((rc = encode_act_map(&c.tmA2, c.in2, c.in2_ps, c.cin - c.c_split, c.Hin, c.Win, pl.max_batch, 0, c.kc, kHaloW, kHaloH, 1, es, is_bf16))) return rc; }
            else c.tmA2 = c.tmA;
            if ((rc = encode_w_map(&c.tmB, c.w_tc, c.cin, c.cout, 9, c.kc, c.halo_bn, es, is_bf16))) return rc;
            c.halo_maps = true;
        }
        const int smem = halo_smem_bytes(c.cin, c.halo_bn, c.kc, c.halo_stages);
        rc = is_bf16 ? launch_halo<__nv_bfloat16>(c.halo_bn, c.kc, fuse, c.tmA, c.tmA2, c.tmB, p, smem, st)
                     : launch_halo<__half>(c.halo_bn, c.kc, fuse, c.tmA, c.tmA2, c.tmB, p, smem, st);
        if (rc) return rc;
    } else if (c.tc) {
        ConvTcParams p;
        const int s2 = (c.sw == 2 ? 1 : 0) | (c.sh == 2 ? 2 : 0);       // parity view per strided axis
        int tw, th, tn;
        choose_tile(c.Hout, c.Wout, N, tw, th, tn);
        const long long m_tiles = (long long)ceil_div(c.Wout, tw) * ceil_div(c.Hout, th) * ceil_div(N, tn);
        int bn = choose_bn(c.cout, m_tiles, c.kh * c.kw * c.cin / 16), ksplit = 0;       // ksplit >= 1: partial mode
        const long long mtot = (long long)N * c.Hout * c.Wout;
        if (c.Hout * c.Wout <= kSplitMaxHW && c.kh * c.kw == 9 && !c.defer && u->partial) {
            const int S = tn > 1 ? tiny_ksplit() : env_int("TS2D_KSPLIT16", 3);        // 16x16 level (one image per tile): 96 CTAs x 72 chunks -> 144 x 24
            if ((S == 1 || S == 3 || S == 9) && (size_t)S * mtot * c.cout * sizeof(float) <= u->partial_bytes) {
                ksplit = S;
                const int fbn = env_int("TS2D_KSPLIT_BN", 0);
                bn = (fbn && c.cout % fbn == 0) ? fbn : choose_bn(c.cout, m_tiles * S, c.kh * c.kw * c.cin / 16 / S);
            }
        }
        const int mc_on = env_int("TS2D_MC", 0);
        int bn2 = 0;                            // > 0: conv_tc2_kernel (two pixel tiles per weight block)
        if (ksplit == 0 && tn == 1 && c.kc == 64 && c.kh * c.kw == 9 && !mc_on) bn2 = choose_bn_tc2(c.cout, m_tiles, c.kh * c.kw * c.cin / c.kc);
        if (bn2) bn = bn2;
        fill_tc_params(p, N, c.Hout, c.Wout, c.cout, c.cin, c.kc, bn, c.kh * c.kw, c.kw, (c.kh - 1) / 2, (c.kw - 1) / 2, s2, c.in_ps);
        p.tw = tw; p.th = th; p.tn = tn;
        p.tiles_w = ceil_div(c.Wout, tw); p.tiles_h = ceil_div(c.Hout, th); p.tiles_n = ceil_div(N, tn);
        p.split_chunks = c.c_split / c.kc;
        p.out_ps = c.cout; p.out = c.out_raw; p.bias = c.b.dev; p.stats = c.stats;
        if (ksplit >= 1) { p.ksplit = ksplit; p.partial = u->partial; p.stats = nullptr; p.bias = nullptr; }
        p.mc = (mc_on && ksplit == 0 && m_tiles % 2 == 0 && bn >= 32) ? 1 : 0;
        // descriptors depend on (input pointer, batch-dependent tile box, BN): cache by (N, bn)
        if (c.first || c.bn != bn * (p.mc ? -1 : 1) || c.enc_n != N) {
            if ((rc = encode_act_map(&c.tmA, c.in, c.in_ps, c.c_split, c.Hin, c.Win, pl.max_batch, s2, c.kc, tw, th, tn, es, is_bf16))) return rc;
            if (c.in2) { if ((rc = encode_act_map(&c.tmA2, c.in2, c.in2_ps, c.cin - c.c_split, c.Hin, c.Win, pl.max_batch, s2, c.kc, tw, th, tn, es, is_bf16))) return rc; }
            else c.tmA2 = c.tmA;
            if ((rc = encode_w_map(&c.tmB, c.w_tc, c.cin, c.cout, c.kh * c.kw, c.kc, p.mc ? bn / 2 : bn, es, is_bf16))) return rc;
            c.bn = bn * (p.mc ? -1 : 1);       // cache key: (BN, multicast or not)
            c.enc_n = N;
        }
        static long long *tc_dbg = nullptr;
        static const int tc_dbg_on = env_int("TS2D_TC_DBG", 0);
        for (int rep = 1; rep < tc_dbg_on; ++rep) {      // TS2D_TC_DBG = 2: every launch twice (the second finds its weights in L2)
            rc = is_bf16 ? launch_tc<__nv_bfloat16>(bn, c.kc, c.tmA, c.tmA2, c.tmB, p, st) : launch_tc<__half>(bn, c.kc, c.tmA, c.tmA2, c.tmB, p, st);
            if (rc) return rc;
        }
        if (tc_dbg_on) {
            if (!tc_dbg) TS2D_CHECK_CUDA(cudaMalloc(&tc_dbg, kNumSMs * 8 * sizeof(long long)));
            TS2D_CHECK_CUDA(cudaMemsetAsync(tc_dbg, 0, kNumSMs * 8 * sizeof(long long), st));
            p.dbg = tc_dbg;
        }
        if (bn2) rc = is_bf16 ? launch_tc2<__nv_bfloat16>(bn, c.tmA, c.tmA2, c.tmB, p, st) : launch_tc2<__half>(bn, c.tmA, c.tmA2, c.tmB, p, st);
        else rc = is_bf16 ? launch_tc<__nv_bfloat16>(bn, c.kc, c.tmA, c.tmA2, c.tmB, p, st) : launch_tc<__half>(bn, c.kc, c.tmA, c.tmA2, c.tmB, p, st);
        if (rc) return rc;
        if (tc_dbg_on) {
            // clock stamps (cycles since the CTA started): prologue, first loads issued, all loads issued, first operands landed,
            // all MMAs issued, epilogue done, teardown; CTA 0 and the maximum over CTAs
            std::vector<long long> h(kNumSMs * 8);
            TS2D_CHECK_CUDA(cudaStreamSynchronize(st));
            TS2D_CHECK_CUDA(cudaMemcpy(h.data(), tc_dbg, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
            const int total = (int)(m_tiles * (c.cout / bn) * std::max(1, ksplit)), g = std::min(total, kNumSMs) & (p.mc ? ~1 : ~0);
            long long mx[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int b = 0; b < g; ++b) for (int k = 0; k < 7; ++k) mx[k] = std::max(mx[k], h[b * 8 + k]);
            fprintf(stderr, "[tc_dbg] %-7s bn %3d kc %2d split %d grid %3d chunks/tile %3d | cta0 %lld %lld %lld %lld %lld %lld %lld | max %lld %lld %lld %lld %lld %lld %lld\n",
                    c.name.c_str(), bn, c.kc, ksplit, g, c.kh * c.kw * c.cin / c.kc / std::max(1, ksplit), h[0], h[1], h[2], h[3], h[4], h[5], h[6], mx[0], mx[1], mx[2],
                    mx[3], mx[4], mx[5], mx[6]);
        }
        if (ksplit >= 1) {
            // finish: slice sum + bias + statistics + InstanceNorm + LeakyReLU in one small launch (no norm_act pass)
            u->launches++;
            const double px = (double)mtot;
            if ((rc = prof_end(u, c.name, 0, 2.0 * px * c.kh * c.kw * c.cin * c.cout, ((double)N * c.Hin * c.Win * c.cin + px * c.cout) * es, st))) return rc;
            if ((rc = prof_begin(u, st))) return rc;
            const dim3 grid(c.cout / 32, N);
            const int HW = c.Hout * c.Wout;
            rc = is_bf16 ? launch_splitk_finish<__nv_bfloat16>(ksplit, grid, u->partial, c.b.dev, c.gamma.dev, c.beta.dev, c.norm_out, c.out_ps, HW, c.cout, mtot,
                                                               pl.eps, pl.negative_slope, st)
                         : launch_splitk_finish<__half>(ksplit, grid, u->partial, c.b.dev, c.gamma.dev, c.beta.dev, c.norm_out, c.out_ps, HW, c.cout, mtot, pl.eps,
                                                        pl.negative_slope, st);
            if (rc) return rc;
            TS2D_LAUNCH_CHECK();
            u->launches++;
            c.materialized = true;
            return prof_end(u, c.name + ".norm", 2, 0.0, (double)ksplit * px * c.cout * 4 + px * c.cout * es, st);
        }
    } else if (c.first_tc) {
        FirstTcParams p;
        memset(&p, 0, sizeof(p));
        p.in = c.in; p.N = N; p.H = c.Hin; p.W = c.Win; p.w = c.w_direct; p.bias = c.b.dev; p.out = c.out_raw; p.Cout = c.cout; p.stats = c.stats;
        p.tiles_x = ceil_div(c.Wout, 8); p.tiles_y = ceil_div(c.Hout, 16);
        rc = is_bf16 ? launch_first_tc<__nv_bfloat16>(p, st) : launch_first_tc<__half>(p, st);
        if (rc) return rc;
    } else if (c.first_fast) {
        FirstConvParams p;
        memset(&p, 0, sizeof(p));
        p.in = c.in; p.H = c.Hin; p.W = c.Win; p.w = c.w_direct; p.bias = c.b.dev; p.out = c.out_raw; p.Cout = c.cout; p.stats = c.stats;
        rc = is_bf16 ? launch_first<__nv_bfloat16>(c.cin, p, N, st) : launch_first<__half>(c.cin, p, N, st);
        if (rc) return rc;
    } else {
        DirectConvParams p;
        memset(&p, 0, sizeof(p));
        p.in = c.in; p.in_ps = c.in_ps; p.Hin = c.Hin; p.Win = c.Win; p.Cin = c.cin;
        p.in2 = c.in2; p.in2_ps = c.in2_ps; p.c_split = c.c_split;
        p.w = c.w_direct; p.bias = c.b.dev;
        p.out = c.out_raw; p.out_ps = c.cout; p.Hout = c.Hout; p.Wout = c.Wout; p.Cout = c.cout;
        p.kh = c.kh; p.kw = c.kw; p.sh = c.sh; p.sw = c.sw; p.ph = (c.kh - 1) / 2; p.pw = (c.kw - 1) / 2;
        p.my = 1; p.mx = 1; p.dy = 0; p.dx = 0; p.Hfull = c.Hout; p.Wfull = c.Wout;
        p.stats = c.stats;
        if ((rc = launch_direct_any(pl.precision, p, N, st))) return rc;
    }
    u->launches++;
    {
        const double px = (double)N * c.Hout * c.Wout;
        if ((rc = prof_end(u, c.name, 0, 2.0 * px * c.kh * c.kw * c.cin * c.cout,
                           ((double)N * c.Hin * c.Win * c.cin + px * c.cout) * es, st))) return rc;
    }
    c.materialized = false;
    if (c.defer) return TS2D_OK;
    return run_norm(u, c, N, st);
}

// InstanceNorm-apply + LeakyReLU of conv `c` into its norm_out (skipped for deferred layers unless someone needs the tensor)
static int run_norm(ts2d_unet *u, ConvLayer &c, int N, cudaStream_t st) {
    const ts2d_unet_plan &pl = u->plan;
    const size_t es = u->es;
    int rc;
    if ((rc = prof_begin(u, st))) return rc;
    rc = launch_norm_any(pl.precision, c.out_raw, c.stats, c.gamma.dev, c.beta.dev, c.norm_out, c.out_ps, c.Hout * c.Wout, c.cout, pl.eps,
                         pl.negative_slope, N, st);
    if (rc) return rc;
    u->launches++;
    c.materialized = true;
    if ((rc = prof_end(u, c.name + ".norm", 2, 0.0, 2.0 * N * c.Hout * c.Wout * c.cout * es, st))) return rc;
    return TS2D_OK;
}

static int run_up(ts2d_unet *u, UpLayer &up, int N, cudaStream_t st) {
    const ts2d_unet_plan &pl = u->plan;
    const size_t es = u->es;
    const bool is_bf16 = pl.precision == TS2D_PREC_BF16;
    int rc;
    if ((rc = prof_begin(u, st))) return rc;
    if (up.tc) {
        ConvTcParams p;
        int tw, th, tn;
        choose_tile(up.Hin, up.Win, N, tw, th, tn);
        const long long m_tiles = (long long)ceil_div(up.Win, tw) * ceil_div(up.Hin, th) * ceil_div(N, tn);
        const int taps = up.sh * up.sw;
        // GEMM N axis = (tap, Cout): all taps of a pixel in one pass over A; the epilogue scatters column block (tap, c0)
        const int n_total = taps * up.cout;
        const int bn = choose_bn(n_total, m_tiles, up.cin / 16);
        fill_tc_params(p, N, up.Hin, up.Win, n_total, up.cin, up.kc, bn, 1, 1, 0, 0, 0, up.cin);
        p.tap_cout = up.cout;
        p.tw = tw; p.th = th; p.tn = tn;
        p.tiles_w = ceil_div(up.Win, tw); p.tiles_h = ceil_div(up.Hin, th); p.tiles_n = ceil_div(N, tn);
        p.my = up.sh; p.mx = up.sw; p.Hfull = up.Hin * up.sh; p.Wfull = up.Win * up.sw;
        p.out_ps = up.out_ps; p.out = up.out; p.bias = up.b.dev; p.stats = nullptr;
        const int mc_on = env_int("TS2D_MC", 0);
        p.mc = (mc_on && m_tiles % 2 == 0) ? 1 : 0;
        if (up.bn != bn * (p.mc ? -1 : 1) || up.enc_n != N) {
            if ((rc = encode_act_map(&up.tmA, up.in, up.cin, up.cin, up.Hin, up.Win, pl.max_batch, 0, up.kc, tw, th, tn, es, is_bf16))) return rc;
            if ((rc = encode_w_map(&up.tmB, up.w_tc, up.cin, n_total, 1, up.kc, p.mc ? bn / 2 : bn, es, is_bf16))) return rc;      // [tap][Cout][Cin] viewed as [taps*Cout][Cin]
            up.bn = bn * (p.mc ? -1 : 1);
            up.enc_n = N;
        }
        rc = is_bf16 ? launch_tc<__nv_bfloat16>(bn, up.kc, up.tmA, up.tmA, up.tmB, p, st) : launch_tc<__half>(bn, up.kc, up.tmA, up.tmA, up.tmB, p, st);
        if (rc) return rc;
        u->launches++;
    } else {
        for (int dy = 0; dy < up.sh; ++dy)
            for (int dx = 0; dx < up.sw; ++dx) {
                DirectConvParams p;
                memset(&p, 0, sizeof(p));
                p.in = up.in; p.in_ps = up.cin; p.Hin = up.Hin; p.Win = up.Win; p.Cin = up.cin; p.c_split = up.cin;
                p.w = up.w_direct + (size_t)(dy * up.sw + dx) * up.cin * up.cout; p.bias = up.b.dev;
                p.out = up.out; p.out_ps = up.out_ps; p.Hout = up.Hin; p.Wout = up.Win; p.Cout = up.cout;
                p.kh = 1; p.kw = 1; p.sh = 1; p.sw = 1; p.ph = 0; p.pw = 0;
                p.my = up.sh; p.mx = up.sw; p.dy = dy; p.dx = dx; p.Hfull = up.Hin * up.sh; p.Wfull = up.Win * up.sw;
                p.stats = nullptr;
                if ((rc = launch_direct_any(pl.precision, p, N, st))) return rc;
                u->launches++;
            }
    }
    {
        const double px = (double)N * up.Hin * up.Win;
        if ((rc = prof_end(u, up.name, 1, 2.0 * px * up.sh * up.sw * up.cin * up.cout,
                           (px * up.cin + px * up.sh * up.sw * up.cout) * es, st))) return rc;
    }
    return TS2D_OK;
}

static int forward_impl(ts2d_unet *u, const void *x, int batch, const char *stop_layer, float *debug_out, cudaStream_t st) {
    if (!u->finalized) { set_error("ts2d_unet_forward: engine not finalised"); return TS2D_ERR_STATE; }
    if (batch < 1 || batch > u->plan.max_batch) { set_error("ts2d_unet_forward: batch %d outside 1..%d", batch, u->plan.max_batch); return TS2D_ERR_STATE; }
    if (((uintptr_t)x) % 16) { set_error("ts2d_unet_forward: input tiles must be 16-byte aligned"); return TS2D_ERR_ARG; }
    TS2D_CHECK_CUDA(cudaMemsetAsync(u->stats_arena, 0, u->stats_bytes, st));
    u->prof_used = 0;
    u->convs[0].in = x;
    int rc;
    for (const Op &op : u->ops) {
        if (op.kind == 0) {
            ConvLayer &c = u->convs[op.idx];
            if ((rc = run_conv(u, c, batch, st))) return rc;
            if (stop_layer && c.name == stop_layer) {
                if (!c.materialized && (rc = run_norm(u, c, batch, st))) return rc;
                const int HW = c.Hout * c.Wout;
                dim3 grid(std::min(ceil_div((long long)HW * c.cout, 256), kNumSMs * 8), batch);
                if (u->plan.precision == TS2D_PREC_F32) nhwc_to_nchw_f32_kernel<float><<<grid, 256, 0, st>>>((const float *)c.norm_out, c.out_ps, c.cout, HW, debug_out);
                else if (u->plan.precision == TS2D_PREC_BF16) nhwc_to_nchw_f32_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16 *)c.norm_out, c.out_ps, c.cout, HW, debug_out);
                else nhwc_to_nchw_f32_kernel<__half><<<grid, 256, 0, st>>>((const __half *)c.norm_out, c.out_ps, c.cout, HW, debug_out);
                TS2D_LAUNCH_CHECK();
                u->last_batch = batch;
                return TS2D_OK;
            }
        } else {
            UpLayer &up = u->ups[op.idx];
            if ((rc = run_up(u, up, batch, st))) return rc;
            if (stop_layer && up.name == stop_layer) {
                const int HW = up.Hin * up.sh * up.Win * up.sw;
                dim3 grid(std::min(ceil_div((long long)HW * up.cout, 256), kNumSMs * 8), batch);
                if (u->plan.precision == TS2D_PREC_F32) nhwc_to_nchw_f32_kernel<float><<<grid, 256, 0, st>>>((const float *)up.out, up.out_ps, up.cout, HW, debug_out);
                else if (u->plan.precision == TS2D_PREC_BF16) nhwc_to_nchw_f32_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16 *)up.out, up.out_ps, up.cout, HW, debug_out);
                else nhwc_to_nchw_f32_kernel<__half><<<grid, 256, 0, st>>>((const __half *)up.out, up.out_ps, up.cout, HW, debug_out);
                TS2D_LAUNCH_CHECK();
                u->last_batch = batch;
                return TS2D_OK;
            }
        }
    }
    u->last_batch = batch;
    if (stop_layer) { set_error("ts2d_unet_forward_debug: unknown layer '%s'", stop_layer); return TS2D_ERR_ARG; }
    return TS2D_OK;
}

}  // namespace ts2d

extern "C" int ts2d_unet_forward(ts2d_unet *u, const void *x_dev, int batch, ts2d_stream stream) {
    TS2D_REQUIRE(u && x_dev, "ts2d_unet_forward: null pointer");
    return forward_impl(u, x_dev, batch, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int ts2d_unet_forward_debug(ts2d_unet *u, const void *x_dev, int batch, const char *layer, float *out_dev, ts2d_stream stream) {
    TS2D_REQUIRE(u && x_dev && layer && out_dev, "ts2d_unet_forward_debug: null pointer");
    return forward_impl(u, x_dev, batch, layer, out_dev, (cudaStream_t)stream);
}

extern "C" int ts2d_unet_tile_logits(ts2d_unet *u, int batch, float *logits_dev, ts2d_stream stream) {
    TS2D_REQUIRE(u && logits_dev, "ts2d_unet_tile_logits: null pointer");
    if (!u->finalized || batch < 1 || batch > u->last_batch) { set_error("ts2d_unet_tile_logits: no forward of >= %d images to read", batch); return TS2D_ERR_STATE; }
    const int HW = u->plan.patch_h * u->plan.patch_w, F = u->C[0], heads = u->plan.num_heads;
    dim3 grid(ceil_div(HW, 128), batch);
    cudaStream_t st = (cudaStream_t)stream;
    if (!u->convs.back().materialized) {
        int rc = run_norm(u, u->convs.back(), u->last_batch, st);
        if (rc) return rc;
    }
    if (u->plan.precision == TS2D_PREC_F32) tile_logits_kernel<float><<<grid, 128, 0, st>>>((const float *)u->features, F, u->head_w.dev, u->head_b.dev, heads, HW, logits_dev);
    else if (u->plan.precision == TS2D_PREC_BF16) tile_logits_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>((const __nv_bfloat16 *)u->features, F, u->head_w.dev, u->head_b.dev, heads, HW, logits_dev);
    else tile_logits_kernel<__half><<<grid, 128, 0, st>>>((const __half *)u->features, F, u->head_w.dev, u->head_b.dev, heads, HW, logits_dev);
    TS2D_LAUNCH_CHECK();
    u->launches++;
    return TS2D_OK;
}

template <typename T, int CAP>
static int launch_blend(const BlendParams &p, const int *tile_yx_host, int F, cudaStream_t st) {
    const long long npix = (long long)p.h * p.w;
    const int grid = ceil_div(npix, 128);
    const size_t smem = (size_t)(p.heads * F + p.heads) * sizeof(float) + (size_t)128 * p.heads + 16 + 128 * sizeof(long long) + (size_t)p.heads * sizeof(int);
    TileListT<CAP> tl;
    fill_tile_list(tl, tile_yx_host, p.T);
    switch (F) {
        case 8: blend_head_kernel<T, 8, CAP><<<grid, 128, smem, st>>>(p, tl); break;
        case 16: blend_head_kernel<T, 16, CAP><<<grid, 128, smem, st>>>(p, tl); break;
        case 32: blend_head_kernel<T, 32, CAP><<<grid, 128, smem, st>>>(p, tl); break;
        case 64: blend_head_kernel<T, 64, CAP><<<grid, 128, smem, st>>>(p, tl); break;
        default: set_error("ts2d_unet_blend_head: features_per_stage[0] must be 8, 16, 32 or 64 (got %d)", F); return TS2D_ERR_UNSUPPORTED;
    }
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
// 16-bit features: blend_head_v2_kernel (blend.cuh).  The grid covers the whole output image when a mask is written.
template <typename T, int F, int CAP>
static int launch_blend2_inst(const BlendParams &p, const int *tile_yx_host, int H0, cudaStream_t st) {
    constexpr int PPB = kBlend2Threads / (F / 8);
    const int n_img = p.T * p.M;
    bool ss_smem = p.feat_ss != nullptr;
    size_t smem = blend2_smem_bytes<F>(p.heads, n_img, ss_smem);
    if (smem > 200 * 1024) { ss_smem = false; smem = blend2_smem_bytes<F>(p.heads, n_img, false); }
    if (smem > 48 * 1024)
        TS2D_ONCE_PER_DEVICE(
            TS2D_CHECK_CUDA(cudaFuncSetAttribute(blend_head_v2_kernel<T, F, CAP, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            TS2D_CHECK_CUDA(cudaFuncSetAttribute(blend_head_v2_kernel<T, F, CAP, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            TS2D_CHECK_CUDA(cudaFuncSetAttribute(blend_head_v2_kernel<T, F, CAP, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            TS2D_CHECK_CUDA(cudaFuncSetAttribute(blend_head_v2_kernel<T, F, CAP, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    const int full_h = p.mask ? H0 : p.h, full_w = p.mask ? p.W0 : p.w;
    const long long npix = (long long)full_h * full_w;
    TileListT<CAP> tl;
    fill_tile_list(tl, tile_yx_host, p.T);
    // persistent blocks: four 256-thread blocks per SM walk the pixel groups (F <= 32; two per SM at F = 64: launch bounds of the kernel)
    const unsigned grid = (unsigned)std::min<long long>(ceil_div(npix, PPB), (long long)kNumSMs * (F <= 32 ? 4 : 2));
    switch (p.mirror_mask) {
        case 0: blend_head_v2_kernel<T, F, CAP, 0><<<grid, kBlend2Threads, smem, st>>>(p, tl, full_h, full_w, ss_smem ? 1 : 0); break;
        case 1: blend_head_v2_kernel<T, F, CAP, 1><<<grid, kBlend2Threads, smem, st>>>(p, tl, full_h, full_w, ss_smem ? 1 : 0); break;
        case 2: blend_head_v2_kernel<T, F, CAP, 2><<<grid, kBlend2Threads, smem, st>>>(p, tl, full_h, full_w, ss_smem ? 1 : 0); break;
        default: blend_head_v2_kernel<T, F, CAP, 3><<<grid, kBlend2Threads, smem, st>>>(p, tl, full_h, full_w, ss_smem ? 1 : 0); break;
    }
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}
template <typename T, int CAP>
static int launch_blend2(const BlendParams &p, const int *tile_yx_host, int F, int H0, cudaStream_t st) {
    switch (F) {
        case 8: return launch_blend2_inst<T, 8, CAP>(p, tile_yx_host, H0, st);
        case 16: return launch_blend2_inst<T, 16, CAP>(p, tile_yx_host, H0, st);
        case 32: return launch_blend2_inst<T, 32, CAP>(p, tile_yx_host, H0, st);
        case 64: return launch_blend2_inst<T, 64, CAP>(p, tile_yx_host, H0, st);
        default: set_error("ts2d_unet_blend_head: features_per_stage[0] must be 8, 16, 32 or 64 (got %d)", F); return TS2D_ERR_UNSUPPORTED;
    }
}
static bool blend_v2_enabled() { return env_int("TS2D_BLEND_V2", 1) != 0; }      // read per call: the tests flip it
template <typename T>
static int launch_blend_any(const BlendParams &p, const int *tile_yx_host, int F, int H0, cudaStream_t st) {
    if constexpr (sizeof(T) == 2) {
        if (blend_v2_enabled())
            return p.T <= 64 ? launch_blend2<T, 64>(p, tile_yx_host, F, H0, st) : launch_blend2<T, kMaxTiles>(p, tile_yx_host, F, H0, st);
    }
    return p.T <= 64 ? launch_blend<T, 64>(p, tile_yx_host, F, st) : launch_blend<T, kMaxTiles>(p, tile_yx_host, F, st);
}

extern "C" int ts2d_unet_blend_head(ts2d_unet *u, const int *tile_yx_host, int T, int M, int mirror_mask, int pad_top, int pad_left, int h, int w,
                                    const float *gaussian_dev, float *logits_dev, uint8_t *mask_dev, int H0, int W0, int h0, int w0, int ch_off,
                                    int ch_total, uint8_t *present_dev, ts2d_stream stream) {
    TS2D_REQUIRE(u && tile_yx_host && gaussian_dev, "ts2d_unet_blend_head: null pointer");
    TS2D_REQUIRE(T >= 1 && T <= kMaxTiles, "ts2d_unet_blend_head: T must be 1..%d", kMaxTiles);
    TS2D_REQUIRE(M == (mirror_mask == 3 ? 4 : (mirror_mask ? 2 : 1)), "ts2d_unet_blend_head: M does not match the mirror mask");
    TS2D_REQUIRE(logits_dev || mask_dev, "ts2d_unet_blend_head: no output requested");
    if (!u->finalized || T * M > u->last_batch) { set_error("ts2d_unet_blend_head: the last forward ran %d images, %d needed", u->last_batch, T * M); return TS2D_ERR_STATE; }
    if (mask_dev)
        TS2D_REQUIRE(h0 >= 0 && w0 >= 0 && h0 + h <= H0 && w0 + w <= W0 && ch_off >= 0 && ch_off + u->plan.num_heads <= ch_total,
                     "ts2d_unet_blend_head: mask window out of range");
    BlendParams p;
    memset(&p, 0, sizeof(p));
    p.feat = u->features; p.T = T; p.M = M; p.mirror_mask = mirror_mask; p.Ph = u->plan.patch_h; p.Pw = u->plan.patch_w;
    p.pad_top = pad_top; p.pad_left = pad_left; p.h = h; p.w = w;
    p.gauss = gaussian_dev; p.head_w = u->head_w.dev; p.head_b = u->head_b.dev; p.heads = u->plan.num_heads;
    p.logits = logits_dev; p.mask = mask_dev; p.W0 = W0; p.h0 = h0; p.w0 = w0; p.ch_off = ch_off; p.ch_total = ch_total;
    p.present = present_dev;
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    ConvLayer &lastc = u->convs.back();
    if (!lastc.materialized) {
        // the last decoder conv left raw output + statistics: finalise {scale, shift} per (image, channel), normalise inside the blend
        const int NC = u->last_batch * lastc.cout;
        norm_finalize_kernel<<<ceil_div(NC, 256), 256, 0, st>>>(lastc.stats, lastc.gamma.dev, lastc.beta.dev, u->feat_ss, NC, lastc.cout,
                                                                 lastc.Hout * lastc.Wout, u->plan.eps);
        TS2D_LAUNCH_CHECK();
        u->launches++;
        p.feat = lastc.out_raw; p.feat_ss = u->feat_ss; p.slope = u->plan.negative_slope;
    }
    if (u->plan.precision == TS2D_PREC_F32) rc = launch_blend_any<float>(p, tile_yx_host, u->C[0], H0, st);
    else if (u->plan.precision == TS2D_PREC_BF16) rc = launch_blend_any<__nv_bfloat16>(p, tile_yx_host, u->C[0], H0, st);
    else rc = launch_blend_any<__half>(p, tile_yx_host, u->C[0], H0, st);
    if (rc) return rc;
    u->launches++;
    return TS2D_OK;
}

static int launch_permute(const uint8_t *src, uint8_t *dst, int64_t n_pix, const PermList &pl, const int *perm_dev, cudaStream_t st) {
    if (n_pix == 0) return TS2D_OK;
    const int grid = (int)((n_pix + 127) / 128);
    permute_channels_kernel<<<grid, 128, 128 * pl.C + pl.C + 16, st>>>(src, dst, n_pix, pl, perm_dev);
    TS2D_LAUNCH_CHECK();
    return TS2D_OK;
}

extern "C" int ts2d_permute_channels(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_pix, int C, const int *perm_host, ts2d_stream stream) {
    TS2D_REQUIRE(src_dev && dst_dev && perm_host && src_dev != dst_dev, "ts2d_permute_channels: bad pointers");
    TS2D_REQUIRE(C >= 1 && C <= 256 && n_pix >= 0, "ts2d_permute_channels: C must be 1..256");
    PermList pl;
    pl.C = C;
    for (int j = 0; j < C; ++j) {
        TS2D_REQUIRE(perm_host[j] >= 0 && perm_host[j] < C, "ts2d_permute_channels: perm[%d] out of range", j);
        pl.perm[j] = perm_host[j];
    }
    return launch_permute(src_dev, dst_dev, n_pix, pl, nullptr, (cudaStream_t)stream);
}

extern "C" int ts2d_merge_present_order(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_pix, int C, const uint8_t *present_dev,
                                        const int *offsets_host, int n_models, int *perm_dev, ts2d_stream stream) {
    TS2D_REQUIRE(src_dev && dst_dev && present_dev && offsets_host && perm_dev && src_dev != dst_dev, "ts2d_merge_present_order: bad pointers");
    TS2D_REQUIRE(C >= 1 && C <= 256 && n_pix >= 0 && n_models >= 1 && n_models <= 32, "ts2d_merge_present_order: C must be 1..256, models 1..32");
    TS2D_REQUIRE(offsets_host[0] == 0 && offsets_host[n_models] == C, "ts2d_merge_present_order: offsets must cover [0, C)");
    ModelOffsets mo;
    mo.n_models = n_models;
    for (int m = 0; m <= n_models; ++m) {
        TS2D_REQUIRE(m == 0 || offsets_host[m] >= offsets_host[m - 1], "ts2d_merge_present_order: offsets must be ascending");
        mo.off[m] = offsets_host[m];
    }
    for (int m = 0; m < n_models; ++m) {
        // segment ids of model m in the order of the keys "Segment{id}_" in a std::map (what sitk GetMetaDataKeys() returns)
        const int lo = offsets_host[m], n = offsets_host[m + 1] - lo;
        std::vector<std::pair<std::string, int>> keys;
        for (int i = 0; i < n; ++i) keys.push_back({"Segment" + std::to_string(i) + "_", i});
        std::sort(keys.begin(), keys.end());
        for (int i = 0; i < n; ++i) mo.lex[lo + i] = lo + keys[i].second;
    }
    cudaStream_t st = (cudaStream_t)stream;
    present_order_kernel<<<1, 32, 0, st>>>(present_dev, mo, perm_dev);
    TS2D_LAUNCH_CHECK();
    PermList pl;
    pl.C = C;
    return launch_permute(src_dev, dst_dev, n_pix, pl, perm_dev, st);
}

extern "C" int ts2d_unet_blend_covers_image(const ts2d_unet *u) {
    // 1: ts2d_unet_blend_head writes this model's channel window of EVERY pixel of the H0 x W0 mask (zeros outside the crop box),
    // so the caller does not have to clear the mask first
    if (!u) return 0;
    const int F = u->C.empty() ? 0 : u->C[0];
    return (u->plan.precision != TS2D_PREC_F32 && blend_v2_enabled() && (F == 8 || F == 16 || F == 32 || F == 64)) ? 1 : 0;
}

extern "C" int ts2d_unet_resize(ts2d_unet *u, int max_batch) {
    TS2D_REQUIRE(u && max_batch >= 1, "ts2d_unet_resize: bad arguments");
    if (u->finalized && u->plan.max_batch == max_batch) return TS2D_OK;
    u->plan.max_batch = max_batch;
    return ts2d_unet_finalize(u);
}

extern "C" int ts2d_unet_set_profiling(ts2d_unet *u, int enable) {
    TS2D_REQUIRE(u, "ts2d_unet_set_profiling: null handle");
    u->profiling = enable != 0;
    u->prof_used = 0;
    return TS2D_OK;
}

static double prof_sum_ms(ts2d_unet *u, bool all_kinds) {
    if (!u || !u->profiling) return -1.0;
    double total = 0.0;
    for (int i = 0; i < u->prof_used; ++i) {
        if (!all_kinds && u->prof_entries[i].kind == 2) continue;     // norm-apply / split-K finish passes are not conv time
        if (cudaEventSynchronize(u->prof_events[i].second) != cudaSuccess) return -1.0;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, u->prof_events[i].first, u->prof_events[i].second) != cudaSuccess) return -1.0;
        total += ms;
    }
    return total;
}
extern "C" double ts2d_unet_last_conv_ms(ts2d_unet *u) { return prof_sum_ms(u, false); }
extern "C" double ts2d_unet_last_forward_ms(ts2d_unet *u) { return prof_sum_ms(u, true); }

extern "C" int ts2d_unet_profile_count(const ts2d_unet *u) { return (u && u->profiling) ? u->prof_used : 0; }

extern "C" int ts2d_unet_profile_entry(ts2d_unet *u, int i, char *name_out, int name_cap, int *kind, double *ms, double *flops, double *bytes) {
    TS2D_REQUIRE(u && u->profiling && i >= 0 && i < u->prof_used, "ts2d_unet_profile_entry: no such entry");
    TS2D_CHECK_CUDA(cudaEventSynchronize(u->prof_events[i].second));
    float t = 0.f;
    TS2D_CHECK_CUDA(cudaEventElapsedTime(&t, u->prof_events[i].first, u->prof_events[i].second));
    const auto &e = u->prof_entries[i];
    if (name_out && name_cap > 0) { strncpy(name_out, e.name.c_str(), name_cap - 1); name_out[name_cap - 1] = 0; }
    if (kind) *kind = e.kind;
    if (ms) *ms = t;
    if (flops) *flops = e.flops;
    if (bytes) *bytes = e.bytes;
    return TS2D_OK;
}

extern "C" int ts2d_unet_destroy(ts2d_unet *u) {
    if (!u) return TS2D_OK;
    auto fr = [](void *p) { if (p) cudaFree(p); };
    for (auto &c : u->convs) { fr(c.w_direct); fr(c.w_tc); fr(c.b.dev); fr(c.gamma.dev); fr(c.beta.dev); }
    for (auto &up : u->ups) { fr(up.w_direct); fr(up.w_tc); fr(up.b.dev); }
    fr(u->head_w.dev); fr(u->head_b.dev); fr(u->arena);
    for (auto &e : u->prof_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    delete u;
    return TS2D_OK;
}
