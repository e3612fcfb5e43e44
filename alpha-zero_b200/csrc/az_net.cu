// Policy/value network inference on sm_100a: the evaluator of the self-play hot path.
//
// Replaces `model(torch.from_numpy(canonical_board).unsqueeze(0))` + the NumPy softmax of
// ai/mcts.py:46-49 for ALL pending leaves at once.  The network is the reference's Net
// (ai/nn.py:63-84): conv patch embedding + cls token + positional embedding, `depth`
// pre-LN transformer blocks (nn.MultiheadAttention + GELU MLP), final LayerNorm, policy and
// value heads on the cls token.  Dense projections run on tcgen05 tensor cores (az_gemm.cuh,
// bf16 in / fp32 accumulate); everything else is fused CUDA-core kernels in this file.
// Only the cls row feeds the heads (ai/nn.py:81), so in the LAST block queries, the output
// projection and the MLP are evaluated for the cls row only -- an exact pruning.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <string>
#include <vector>
#include "az_gemm.cuh"
#include "az_gemm2.cuh"
#include "az_host.h"
#include "../../include/az_b200.h"

#define AZ_CUDA_OK(expr)                                                                       \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess) return az_fail(AZ_ECUDA, std::string(#expr ": ") + cudaGetErrorString(e__)); \
    } while (0)

typedef __nv_bfloat16 bf16;

// ---- TMA descriptors ------------------------------------------------------------------
typedef CUresult (*az_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                       const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                       CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static az_encode_tiled_fn az_get_encode() {
    static az_encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<az_encode_tiled_fn>(p);
    }
    return fn;
}

// bf16 matrix [rows][cols] with row stride `ld` elements, K-major; box = 64 cols x box_rows,
// 128-byte swizzle (what az_gemm_kernel's UMMA descriptors expect).
static int az_make_tmap(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    az_encode_tiled_fn enc = az_get_encode();
    if (!enc) return az_fail(AZ_ECUDA, "cuTensorMapEncodeTiled entry point not found");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * sizeof(bf16)};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return az_fail(AZ_ECUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
    return AZ_OK;
}

static int g_num_sms = 0;

static int az_pick_bn(int N) { return N % 256 == 0 ? 256 : (N % 192 == 0 ? 192 : (N % 128 == 0 ? 128 : (N % 64 == 0 ? 64 : 0))); }

// Column-tile width for a GEMM with few row tiles (the cls-row GEMMs: M = boards): the widest tile
// is not the best one when it leaves SMs idle.  cost ~ waves x bytes staged per k-block.
static int az_pick_bn_small_m(int N, int max_rows) {
    if (const char* e = getenv("AZ_SMALL_M_ROWS")) {          // experiments: choose the tile for this many live rows instead of the capacity
        const int r = atoi(e);
        if (r > 0 && r < max_rows) max_rows = r;
    }
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int cands[3] = {256, 128, 64};
    int best = 0;
    long long best_cost = 0;
    for (int bn : cands) {
        if (N % bn) continue;
        const long long tiles = (long long)((max_rows + 127) / 128) * (N / bn);
        const long long cost = ((tiles + g_num_sms - 1) / g_num_sms) * (128 + bn);
        if (!best || cost < best_cost) { best = bn; best_cost = cost; }
    }
    return best ? best : az_pick_bn(N);
}

// CTA-pair tiles (az_gemm2.cuh): 256 x bn per pair of SMs, each CTA stages 128 A rows + bn/2 W rows per k-block.
// Returned as bn + AZ_PAIR so that one integer describes the tile through the launch path.
#define AZ_PAIR 1000
static int az_pick_bn_pair(int N, int max_rows) {
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int cands[2] = {256, 128};
    int best = 0;
    long long best_cost = 0;
    for (int bn : cands) {
        if (N % bn) continue;
        const long long tiles = (long long)((max_rows + 255) / 256) * (N / bn);
        const long long pairs = g_num_sms / 2;
        const long long cost = ((tiles + pairs - 1) / pairs) * (128 + bn / 2);
        if (!best || cost < best_cost) { best = bn; best_cost = cost; }
    }
    return best ? best + AZ_PAIR : 0;
}
static bool az_pair_enabled() { return !getenv("AZ_NO_PAIR"); }
static uint32_t az_w_box_rows(int bn) { return bn >= AZ_PAIR ? (uint32_t)(bn - AZ_PAIR) / 2 : (uint32_t)bn; }
static int az_bn_width(int bn) { return bn >= AZ_PAIR ? bn - AZ_PAIR : bn; }

// Launch with programmatic stream serialization: the kernel may begin (prologue only -- it calls
// griddepcontrol.wait before touching its inputs) while its predecessor in the stream drains.
// AZ_NO_PDL=1 falls back to fully serialised launches.
template <typename... KArgs, typename... Args>
static cudaError_t az_launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t s, Args&&... args) {
    static const bool pdl = !getenv("AZ_NO_PDL");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// Output tensor map of a GEMM whose bf16 tiles leave through shared memory + TMA stores (az_gemm.cuh): box = 64 columns x 32
// rows (one epilogue warp), rows beyond max_rows are clipped by the hardware.  Encoded maps are cached per (pointer, shape).
template <int BN, int EPI>
static int az_gemm_out_map(AzGemmArgs* a, CUtensorMap* to, int max_rows) {
    static const bool off = getenv("AZ_NO_TMA_STORE") != nullptr;
    a->stage_out = 0;
    if (off || !(BN == 128 || BN == 256) || EPI == AZ_EPI_HEAD || !a->out || a->ldo % 8 || (reinterpret_cast<uintptr_t>(a->out) & 15) || max_rows < 1)
        return AZ_OK;
    struct Key {
        const void* p; int rows, n, ld;
        bool operator<(const Key& o) const { return std::tie(p, rows, n, ld) < std::tie(o.p, o.rows, o.n, o.ld); }
    };
    static std::map<Key, CUtensorMap> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    const Key k{a->out, max_rows, a->N, a->ldo};
    auto it = cache.find(k);
    if (it == cache.end()) {
        CUtensorMap m;
        int rc = az_make_tmap(&m, a->out, (uint64_t)max_rows, (uint64_t)a->N, (uint64_t)a->ldo, 32);
        if (rc) return rc;
        if (cache.size() > 256) cache.clear();
        it = cache.emplace(k, m).first;
    }
    *to = it->second;
    a->stage_out = 1;
    return AZ_OK;
}

template <int BN, int EPI>
static int az_launch_gemm_t(const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& a, int max_rows, cudaStream_t s) {
    constexpr int smem = AzGemmCfg<BN>::SMEM + (EPI == AZ_EPI_HEAD ? AZ_HEAD_STAGE_BYTES : 0);
    static bool attr_done = false;
    if (!attr_done) {
        AZ_CUDA_OK(cudaFuncSetAttribute(az_gemm_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int tiles = ((max_rows + 127) / 128) * (a.N / BN);
    const int grid = tiles < g_num_sms ? (tiles > 0 ? tiles : 1) : g_num_sms;
    AzGemmArgs a2 = a;
    CUtensorMap to = ta;                                       // placeholder when the tile does not go through TMA stores
    int rc_o = az_gemm_out_map<BN, EPI>(&a2, &to, max_rows);
    if (rc_o) return rc_o;
    AZ_CUDA_OK(az_launch_pdl(az_gemm_kernel<BN, EPI>, grid, AZ_GEMM_THREADS, smem, s, ta, tb, to, a2));
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

template <int BN, int EPI>
static int az_launch_gemm2_t(const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& a, int max_rows, cudaStream_t s) {
    constexpr int smem = AzGemm2Cfg<BN>::SMEM;
    static bool attr_done = false;
    if (!attr_done) {
        AZ_CUDA_OK(cudaFuncSetAttribute(az_gemm2_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_done = true;
    }
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int tiles = ((max_rows + 255) / 256) * (a.N / BN);
    const int pairs = std::max(1, std::min(tiles, g_num_sms / 2));
    static const bool pdl = !getenv("AZ_NO_PDL");
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs));
    cfg.blockDim = dim3(AZ_GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    AzGemmArgs a2 = a;
    CUtensorMap to = ta;
    int rc_o = az_gemm_out_map<BN, EPI>(&a2, &to, max_rows);
    if (rc_o) return rc_o;
    AZ_CUDA_OK(cudaLaunchKernelEx(&cfg, az_gemm2_kernel<BN, EPI>, ta, tb, to, a2));
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

// A: bf16 [rows][K] (ld = lda), W: bf16 [N][K]; max_rows bounds the grid when M is device-side.
static int az_launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& a, int epi, int max_rows,
                          cudaStream_t s, int bn_forced = 0) {
    if (bn_forced >= AZ_PAIR) {                                // CTA-pair tile: the W tensor map has a box of bn/2 rows
        const int bn2 = bn_forced - AZ_PAIR;
        if ((bn2 != 128 && bn2 != 256) || a.N % bn2 || a.K % 64 || epi == AZ_EPI_HEAD || a.out_f32)
            return az_fail(AZ_EINVAL, "gemm: unsupported CTA-pair tile");
        if ((a.stats_in && (!a.colsum || a.stats_parts < 1 || a.stats_parts > AZ_STATS_LD / 2)) || (a.stats_out && 2 * (a.N / bn2) > AZ_STATS_LD / 2))
            return az_fail(AZ_EINVAL, "gemm: bad LayerNorm-fold arguments");
#define AZ_G2(BN_)                                                                             \
    if (bn2 == BN_) {                                                                          \
        if (epi == AZ_EPI_BIAS) return az_launch_gemm2_t<BN_, AZ_EPI_BIAS>(ta, tb, a, max_rows, s);  \
        if (epi == AZ_EPI_GELU) return az_launch_gemm2_t<BN_, AZ_EPI_GELU>(ta, tb, a, max_rows, s);  \
        return az_launch_gemm2_t<BN_, AZ_EPI_RESID>(ta, tb, a, max_rows, s);                   \
    }
        AZ_G2(256) AZ_G2(128)
#undef AZ_G2
    }
    const int bn = bn_forced ? bn_forced : az_pick_bn(a.N);
    if (!bn || a.K % 64) return az_fail(AZ_EINVAL, "gemm: N and K must be multiples of 64");
    if (epi == AZ_EPI_HEAD) {
        if (bn != 64 || a.N != 64 || !a.stats_in || !a.colsum || !a.priors || !a.values || a.head_A < 1 || a.head_A > 63)
            return az_fail(AZ_EINVAL, "gemm: head epilogue needs N = 64, row statistics and output pointers");
        return az_launch_gemm_t<64, AZ_EPI_HEAD>(ta, tb, a, max_rows, s);
    }
    if ((a.stats_in && (!a.colsum || a.stats_parts < 1 || a.stats_parts > AZ_STATS_LD / 2)) || (a.stats_out && 2 * (a.N / bn) > AZ_STATS_LD / 2))
        return az_fail(AZ_EINVAL, "gemm: bad LayerNorm-fold arguments");
#define AZ_G(BN_)                                                                              \
    if (bn == BN_) {                                                                           \
        if (epi == AZ_EPI_BIAS) return az_launch_gemm_t<BN_, AZ_EPI_BIAS>(ta, tb, a, max_rows, s);   \
        if (epi == AZ_EPI_GELU) return az_launch_gemm_t<BN_, AZ_EPI_GELU>(ta, tb, a, max_rows, s);   \
        return az_launch_gemm_t<BN_, AZ_EPI_RESID>(ta, tb, a, max_rows, s);                    \
    }
    AZ_G(256) AZ_G(192) AZ_G(128) AZ_G(64)
#undef AZ_G
    return az_fail(AZ_EINVAL, "gemm: unsupported tile");
}

// Stand-alone GEMM entry point (parity tests of the tensor-core kernel against torch).
extern "C" int az_gemm_bf16(const void* a_dev, int lda, const void* w_dev, const float* bias_dev, const void* resid_dev,
                            int ldr, void* out_dev, int ldo, int M, int N, int K, int epi, void* stream) {
    if (!a_dev || !w_dev || !bias_dev || !out_dev || M <= 0) return az_fail(AZ_EINVAL, "az_gemm_bf16: bad arguments");
    if (epi == AZ_EPI_RESID && !resid_dev) return az_fail(AZ_EINVAL, "az_gemm_bf16: residual missing");
    int bn = az_pick_bn(N);
    if (!bn || K % 64 || lda % 8 || ldo % 8) return az_fail(AZ_EINVAL, "az_gemm_bf16: N,K multiples of 64; ld multiples of 8");
    if (M >= 512 && N % 128 == 0 && az_pair_enabled()) bn = (N % 256 == 0 ? 256 : 128) + AZ_PAIR;   // CTA-pair kernel
    if (const char* e = getenv("AZ_GEMM_BN")) {                 // experiments (tools/gemm_bench.py): force the tile width
        const int f = atoi(e), w = f >= AZ_PAIR ? f - AZ_PAIR : f;
        if (w > 0 && N % w == 0 && (f < AZ_PAIR ? (w == 64 || w == 128 || w == 192 || w == 256) : (w == 128 || w == 256))) bn = f;
    }
    CUtensorMap ta, tb;
    int rc;
    if ((rc = az_make_tmap(&ta, a_dev, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 128))) return rc;
    if ((rc = az_make_tmap(&tb, w_dev, (uint64_t)N, (uint64_t)K, (uint64_t)K, az_w_box_rows(bn)))) return rc;
    AzGemmArgs a{};
    a.bias = bias_dev;
    a.resid = reinterpret_cast<const bf16*>(resid_dev);
    a.out = reinterpret_cast<bf16*>(out_dev);
    a.ldo = ldo;
    a.ldr = ldr;
    a.N = N;
    a.K = K;
    a.m_fixed = M;
    a.m_count = nullptr;
    if (const char* e = getenv("AZ_GEMM_DBG")) a.dbg = atoi(e);
    return az_launch_gemm(ta, tb, a, epi, M, (cudaStream_t)stream, bn);
}

// ======================================================================================
// The network object
// ======================================================================================
#include "az_net_kernels.cuh"
#include "az_embed_mma.cuh"
#include "az_front.cuh"
#include "az_front2.cuh"
#include "az_attn.cuh"
#include "az_tail.cuh"
#include "az_chain.cuh"

struct AzBlockW {
    float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
    bf16 *w_in = nullptr, *w_out = nullptr, *w1 = nullptr, *w2 = nullptr;
    float *b_in = nullptr, *b_out = nullptr, *b1 = nullptr, *b2 = nullptr;
    CUtensorMap tm_in, tm_q, tm_kv, tm_out, tm_w1, tm_w2;
    CUtensorMap tm_out_c, tm_w1_c, tm_w2_c;     // same weights, tile width chosen for the cls-row GEMMs
    int bn_out_c = 0, bn_w1_c = 0, bn_w2_c = 0;
    // norm2 folded into mlp1 for the cls rows (az_gemm.cuh): w1 * diag(gain), bias + w1 beta, row sums
    bf16* w1_ln = nullptr;
    float *b1_ln = nullptr, *cs1 = nullptr;
    CUtensorMap tm_w1_ln_c;
};

struct az_net {
    int rows = 0, cols = 0, C = 0, patch = 0, D = 0, H = 0, L = 0, A = 0, T = 0;
    int chunk = 0;                       // boards per pass through the activation buffers
    int max_batch = 0;
    bool finalized = false;
    std::vector<void*> allocs;
    std::map<std::string, std::vector<float>> host;   // raw fp32 parameters until finalize
    std::vector<float> host_folded_w_in, host_folded_b_in;   // block 0 in_proj with norm1's affine folded in
    // embedding
    bf16* wt_embed = nullptr;
    float* posb = nullptr;
    bf16* w_emb_mma = nullptr;           // [D][kpad]: conv taps | pos-emb rows (tensor-core embedding)
    int emb_kc = 0, emb_kblocks = 0;     // 0 kblocks: shape not supported, CUDA-core embedding is used
    CUtensorMap tm_emb, tm_x_st, tm_xn_st;   // embedding weights; x / xn as TMA-store targets (box 64 x 32)
    bool emb_tma_out = false;
    float *ones = nullptr, *zeros = nullptr;
    // depth-1 nets: cls query folded into the projection (V + 64 score columns), see az_cls_attention_folded_kernel
    bool folded = false;
    int n2 = 0;
    bf16* w_vs = nullptr;
    float *b_vs = nullptr, *sc = nullptr;
    CUtensorMap tm_vs;
    // depth-1 nets on boards of <= 63 cells: embedding .. cls attention in one kernel (az_front.cuh)
    bool front = false;
    bool front2 = false;                 // reassociated value path (az_front2.cuh): U vectors + out-proj over K = H*128
    bf16 *w_out2 = nullptr, *uc = nullptr;
    float* b_out2 = nullptr;
    CUtensorMap tm_wout2, tm_uc;
    int bn_out2 = 0;
    std::vector<double> host_wemb;       // [kpad][D] rows: conv taps, then pos-emb + bias rows (fp64 copy for the folds)
    int host_wemb_kpad = 0;
    bf16 *f_wst = nullptr, *f_wv = nullptr;
    float *f_bias_v = nullptr, *b_out_x0 = nullptr;
    CUtensorMap tm_fwst, tm_fwv;
    AzFrontArgs fa{};
    std::vector<AzBlockW> blk;
    float *lnf_g = nullptr, *lnf_b = nullptr, *w_pol = nullptr, *b_pol = nullptr, *w_val = nullptr, *b_val = nullptr;
    // heads on tensor cores when 32 < A + 1 <= 64: [policy rows ; value row ; zero rows] as one [64][D] bf16 weight
    bf16* w_head = nullptr;
    float *b_head = nullptr, *head_out = nullptr;
    CUtensorMap tm_whead;
    // cls-row tail without LayerNorm kernels: row statistics travel between the GEMM epilogues,
    // the final norm + heads + softmax are the epilogue of one N = 64 GEMM (A + 1 <= 64)
    bool tail_fold = false, head_fold = false;
    bool attn_mma = false;               // all-token attention of the full blocks on tcgen05 (az_attn.cuh)
    CUtensorMap tm_qkv;
    bool fused_tail = false;             // out-proj .. heads of the cls rows in one cluster kernel (az_tail.cuh), small batches
    bool chain = false;                  // ... as one persistent kernel running the four GEMMs back to back (az_chain.cuh)
    unsigned int* chain_sync = nullptr;
    CUtensorMap tm_t_u, tm_t_wo, tm_t_w1, tm_t_w2;
    float *stats1 = nullptr, *stats2 = nullptr;
    bf16* w_head_ln = nullptr;
    float *b_head_ln = nullptr, *cs_head = nullptr;
    CUtensorMap tm_whead_ln, tm_xc, tm_xc2;
    bool head_gemm = false;
    // activations (one chunk)
    bf16 *x = nullptr, *xn = nullptr, *qkv = nullptr, *ao = nullptr, *hbuf = nullptr;
    bf16 *qc = nullptr, *oc = nullptr, *xc = nullptr, *xcn = nullptr, *hc = nullptr, *xc2 = nullptr;
    CUtensorMap tm_xn, tm_ao, tm_hbuf, tm_xn_cls, tm_oc, tm_xcn, tm_hc;
    // scratch for the stand-alone forward (planes in, logits out)
    uint64_t* in_words = nullptr;
    uint8_t* in_player = nullptr;
    float* in_planes = nullptr;          // host planes staged on the device (az_net_forward_planes), allocated on first use
    float *out_priors = nullptr, *out_values = nullptr, *out_logits = nullptr;
    unsigned long long launches = 0;     // kernels launched by forward passes
    unsigned long long weights_version = 0;   // changes with every az_net_finalize: tag of the forests' evaluation caches
    // optional per-kernel timing of eager forward passes (bench.py roofline numbers)
    bool prof_on = false;
    struct ProfRec { int kind; cudaEvent_t a, b; };
    std::vector<ProfRec> prof_pending;
    double prof_ms[16] = {0};
    long long prof_n[16] = {0};
    double prof_units[16] = {0};         // flops (GEMMs) or bytes (others) of the profiled launches
};

enum { AZ_K_EMBED = 0, AZ_K_GEMM_KV, AZ_K_GEMM_CLS, AZ_K_CLS_ATT, AZ_K_LN, AZ_K_HEAD, AZ_K_GEMM_FULL, AZ_K_ATT_FULL, AZ_K_FRONT, AZ_K_TAIL, AZ_K_COUNT };
static const char* const AZ_K_NAMES[AZ_K_COUNT] = {"embed_ln", "gemm_kv", "gemm_cls_rows", "cls_attention", "layernorm",
                                                   "head_softmax", "gemm_full_block", "attention_full", "front_fused", "tail_fused"};

struct AzProfScope {
    az_net* n; cudaStream_t s; int idx = -1;
    AzProfScope(az_net* n_, int kind, cudaStream_t s_) : n(n_), s(s_) {
        if (!n->prof_on) return;
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(s, &st);
        if (st != cudaStreamCaptureStatusNone) return;
        az_net::ProfRec r{kind, nullptr, nullptr};
        cudaEventCreate(&r.a); cudaEventCreate(&r.b);
        cudaEventRecord(r.a, s);
        n->prof_pending.push_back(r);
        idx = (int)n->prof_pending.size() - 1;
    }
    ~AzProfScope() { if (idx >= 0) cudaEventRecord(n->prof_pending[idx].b, s); }
};

template <typename T>
static int az_net_alloc(az_net* n, T** p, size_t count) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, count * sizeof(T) + 256);
    if (e != cudaSuccess) return az_fail(AZ_ENOMEM, std::string("net cudaMalloc: ") + cudaGetErrorString(e));
    n->allocs.push_back(q);
    *p = reinterpret_cast<T*>(q);
    return AZ_OK;
}

static int az_upload_f32(az_net* n, float** dst, const std::vector<float>& v) {
    int rc = az_net_alloc(n, dst, v.size());
    if (rc) return rc;
    AZ_CUDA_OK(cudaMemcpy(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
    return AZ_OK;
}

static int az_upload_bf16(az_net* n, bf16** dst, const float* v, size_t count) {
    std::vector<bf16> tmp(count);
    for (size_t i = 0; i < count; ++i) tmp[i] = __float2bfloat16(v[i]);
    int rc = az_net_alloc(n, dst, count);
    if (rc) return rc;
    AZ_CUDA_OK(cudaMemcpy(*dst, tmp.data(), count * 2, cudaMemcpyHostToDevice));
    return AZ_OK;
}

extern "C" int az_net_create(int rows, int cols, int planes, int patch, int embed_dim, int num_heads, int depth,
                             int action_dim, int max_batch, az_net** out) {
    if (!out || rows <= 0 || cols <= 0 || planes < 2 || planes > 3 || patch <= 0 || !(patch & 1) || depth < 1 ||
        action_dim <= 0 || max_batch <= 0)
        return az_fail(AZ_EINVAL, "az_net_create: bad arguments");
    if (embed_dim % 64 || embed_dim > 1024) return az_fail(AZ_EINVAL, "az_net_create: embed_dim must be a multiple of 64, <= 1024");
    if (num_heads <= 0 || embed_dim % num_heads || embed_dim / num_heads > 128 || (embed_dim / num_heads) % 8)
        return az_fail(AZ_EINVAL, "az_net_create: head dim must divide embed_dim, be a multiple of 8 and <= 128");
    if (rows * cols + 1 > 256) return az_fail(AZ_EINVAL, "az_net_create: at most 255 board cells");
    az_net* n = new az_net();
    n->rows = rows; n->cols = cols; n->C = planes; n->patch = patch; n->D = embed_dim; n->H = num_heads;
    n->L = depth; n->A = action_dim; n->T = rows * cols + 1;
    n->max_batch = max_batch;
    n->chunk = max_batch < 4096 ? max_batch : 4096;
    n->blk.resize(depth);
    *out = n;
    return AZ_OK;
}

extern "C" int az_net_invalidate_graphs(az_net* n);
extern "C" void az_net_destroy(az_net* n) {
    if (!n) return;
    az_net_invalidate_graphs(n);
    for (void* p : n->allocs) cudaFree(p);
    delete n;
}

// Parameters arrive under their state_dict names (ai/nn.py module tree), fp32, host memory.
extern "C" int az_net_set_param(az_net* n, const char* name, const float* data_host, long long numel) {
    if (!n || !name || !data_host || numel <= 0) return az_fail(AZ_EINVAL, "az_net_set_param: bad arguments");
    n->host[name] = std::vector<float>(data_host, data_host + numel);
    n->finalized = false;
    return AZ_OK;
}

static const std::vector<float>* az_param(az_net* n, const std::string& name, size_t numel, std::string* err) {
    auto it = n->host.find(name);
    if (it == n->host.end()) { *err = "missing parameter " + name; return nullptr; }
    if (it->second.size() != numel) { *err = "parameter " + name + " has wrong size"; return nullptr; }
    return &it->second;
}

// Converts the fp32 state_dict into device layouts, allocates one chunk of activations and
// encodes the TMA descriptors.  Call again after az_net_set_param to refresh the weights.
extern "C" int az_net_finalize(az_net* n) {
    if (!n) return az_fail(AZ_EINVAL, "az_net_finalize: null");
    static unsigned long long g_weights_version = 0;
    n->weights_version = 0x4E45540000000000ull + (++g_weights_version);
    az_net_invalidate_graphs(n);
    for (void* p : n->allocs) cudaFree(p);
    n->allocs.clear();
    n->in_planes = nullptr;
    const int D = n->D, T = n->T, C = n->C, P = n->patch, A = n->A, cells = n->rows * n->cols;
    std::string err;
    int rc;
#define AZ_P(var, name, numel)                                      \
    const std::vector<float>* var = az_param(n, name, numel, &err); \
    if (!var) return az_fail(AZ_EINVAL, "az_net_finalize: " + err);
    // ---- embedding: conv weight [D][C][P][P] -> channel-last [C*P*P][D]; pos/bias tables
    AZ_P(cw, "embedding.patch_embed.patch_embed.weight", (size_t)D * C * P * P)
    AZ_P(cb, "embedding.patch_embed.patch_embed.bias", (size_t)D)
    AZ_P(cls, "embedding.cls_token", (size_t)D)
    AZ_P(pos, "embedding.pos_embedding", (size_t)T * D)
    {
        std::vector<float> wt((size_t)C * P * P * D);
        for (int d = 0; d < D; ++d)
            for (int k = 0; k < C * P * P; ++k) wt[(size_t)k * D + d] = (*cw)[(size_t)d * C * P * P + k];
        if ((rc = az_upload_bf16(n, &n->wt_embed, wt.data(), wt.size()))) return rc;
        // the kernel sums bf16-rounded weights: build the tables from the same rounded values
        std::vector<float> posb((size_t)2 * T * D);
        for (int stm = 0; stm < 2; ++stm)
            for (int t = 0; t < T; ++t)
                for (int d = 0; d < D; ++d) {
                    float v = (*pos)[(size_t)t * D + d];
                    if (t == 0) v += (*cls)[d];
                    else {
                        v += (*cb)[d];
                        if (stm == 1 && C == 3) {          // plane 2 is all ones: every in-bounds tap contributes
                            const int r = (t - 1) / n->cols, c = (t - 1) % n->cols;
                            for (int kr = 0; kr < P; ++kr)
                                for (int kc = 0; kc < P; ++kc) {
                                    const int rr = r + kr - P / 2, cc = c + kc - P / 2;
                                    if (rr < 0 || rr >= n->rows || cc < 0 || cc >= n->cols) continue;
                                    v += __bfloat162float(__float2bfloat16(wt[(size_t)((2 * P + kr) * P + kc) * D + d]));
                                }
                        }
                    }
                    posb[((size_t)stm * T + t) * D + d] = v;
                }
        if ((rc = az_upload_f32(n, &n->posb, posb))) return rc;
        // tensor-core embedding: A = [taps | one-hot(t)], Wemb = [conv ; pos + bias], K padded to 64/128
        const int kc = C * P * P;
        n->emb_kc = kc;
        n->emb_kblocks = 0;
        n->host_wemb.clear();
        n->host_wemb_kpad = 0;
        if (kc + T <= 128) {
            const int kpad = (kc + T + 63) / 64 * 64;
            n->host_wemb.assign((size_t)kpad * D, 0.0);
            for (int d = 0; d < D; ++d) {
                for (int k = 0; k < kc; ++k) n->host_wemb[(size_t)k * D + d] = (*cw)[(size_t)d * kc + k];
                for (int t = 0; t < T; ++t)
                    n->host_wemb[(size_t)(kc + t) * D + d] = (double)(*pos)[(size_t)t * D + d] + (t == 0 ? (*cls)[d] : (*cb)[d]);
            }
            n->host_wemb_kpad = kpad;
        }
        const bool d_ok = (D <= 256 && D % 16 == 0) || D == 512;
        if (kc + T <= 128 && d_ok && !getenv("AZ_EMBED_SIMT")) {
            const int kb = (kc + T + 63) / 64, kpad = kb * 64;
            std::vector<float> we((size_t)D * kpad, 0.f);
            for (int d = 0; d < D; ++d) {
                for (int k = 0; k < kc; ++k) we[(size_t)d * kpad + k] = (*cw)[(size_t)d * kc + k];
                for (int t = 0; t < T; ++t)
                    we[(size_t)d * kpad + kc + t] = (*pos)[(size_t)t * D + d] + (t == 0 ? (*cls)[d] : (*cb)[d]);
            }
            if ((rc = az_upload_bf16(n, &n->w_emb_mma, we.data(), we.size()))) return rc;
            if ((rc = az_make_tmap(&n->tm_emb, n->w_emb_mma, D, kpad, kpad, D > 256 ? 256 : D))) return rc;
            n->emb_kblocks = kb;
        }
        std::vector<float> one((size_t)D, 1.f), zero((size_t)D, 0.f);
        if ((rc = az_upload_f32(n, &n->ones, one)) || (rc = az_upload_f32(n, &n->zeros, zero))) return rc;
    }
    // ---- blocks
    for (int l = 0; l < n->L; ++l) {
        AzBlockW& w = n->blk[l];
        const std::string p = "blocks." + std::to_string(l) + ".";
        AZ_P(g1, p + "norm1.weight", (size_t)D) AZ_P(be1, p + "norm1.bias", (size_t)D)
        AZ_P(g2, p + "norm2.weight", (size_t)D) AZ_P(be2, p + "norm2.bias", (size_t)D)
        AZ_P(wi, p + "attn.in_proj_weight", (size_t)3 * D * D) AZ_P(bi, p + "attn.in_proj_bias", (size_t)3 * D)
        AZ_P(wo, p + "attn.out_proj.weight", (size_t)D * D) AZ_P(bo, p + "attn.out_proj.bias", (size_t)D)
        AZ_P(w1, p + "mlp.0.weight", (size_t)4 * D * D) AZ_P(bb1, p + "mlp.0.bias", (size_t)4 * D)
        AZ_P(w2, p + "mlp.3.weight", (size_t)4 * D * D) AZ_P(bb2, p + "mlp.3.bias", (size_t)D)
        std::vector<float> wi_f(*wi), bi_f(*bi);
        if (l == 0) {        // norm1 of block 0 runs inside the embedding kernel without its affine:
                             // W' = W diag(gamma), b' = b + W beta  (exact in real arithmetic)
            for (int r = 0; r < 3 * D; ++r) {
                double acc = 0.0;
                for (int k = 0; k < D; ++k) {
                    acc += (double)(*wi)[(size_t)r * D + k] * (*be1)[k];
                    wi_f[(size_t)r * D + k] = (*wi)[(size_t)r * D + k] * (*g1)[k];
                }
                bi_f[r] = (*bi)[r] + (float)acc;
            }
        }
        if (l == 0) { n->host_folded_w_in = wi_f; n->host_folded_b_in = bi_f; }
        if ((rc = az_upload_f32(n, &w.ln1_g, *g1)) || (rc = az_upload_f32(n, &w.ln1_b, *be1)) ||
            (rc = az_upload_f32(n, &w.ln2_g, *g2)) || (rc = az_upload_f32(n, &w.ln2_b, *be2)) ||
            (rc = az_upload_f32(n, &w.b_in, bi_f)) || (rc = az_upload_f32(n, &w.b_out, *bo)) ||
            (rc = az_upload_f32(n, &w.b1, *bb1)) || (rc = az_upload_f32(n, &w.b2, *bb2)) ||
            (rc = az_upload_bf16(n, &w.w_in, wi_f.data(), wi_f.size())) || (rc = az_upload_bf16(n, &w.w_out, wo->data(), wo->size())) ||
            (rc = az_upload_bf16(n, &w.w1, w1->data(), w1->size())) || (rc = az_upload_bf16(n, &w.w2, w2->data(), w2->size())))
            return rc;
        if ((rc = az_make_tmap(&w.tm_in, w.w_in, 3 * D, D, D, az_pick_bn(3 * D))) ||
            (rc = az_make_tmap(&w.tm_q, w.w_in, D, D, D, az_pick_bn(D))) ||
            (rc = az_make_tmap(&w.tm_kv, w.w_in + (size_t)D * D, 2 * D, D, D, az_pick_bn(2 * D))) ||
            (rc = az_make_tmap(&w.tm_out, w.w_out, D, D, D, az_pick_bn(D))) ||
            (rc = az_make_tmap(&w.tm_w1, w.w1, 4 * D, D, D, az_pick_bn(4 * D))) ||
            (rc = az_make_tmap(&w.tm_w2, w.w2, D, 4 * D, 4 * D, az_pick_bn(D))))
            return rc;
        if (l == n->L - 1) {                                  // norm2 folded into mlp1 (cls rows of the last block)
            auto r16 = [](float v) { return __bfloat162float(__float2bfloat16(v)); };
            std::vector<float> wf((size_t)4 * D * D), bfold((size_t)4 * D), cs((size_t)4 * D);
            for (int r = 0; r < 4 * D; ++r) {
                double acc = (*bb1)[r], sum = 0.0;
                for (int k = 0; k < D; ++k) {
                    const float wv = (*w1)[(size_t)r * D + k] * (*g2)[k];
                    wf[(size_t)r * D + k] = wv;
                    sum += (double)r16(wv);                   // what the tensor cores will multiply by
                    acc += (double)(*w1)[(size_t)r * D + k] * (*be2)[k];
                }
                bfold[r] = (float)acc;
                cs[r] = (float)sum;
            }
            if ((rc = az_upload_bf16(n, &w.w1_ln, wf.data(), wf.size())) || (rc = az_upload_f32(n, &w.b1_ln, bfold)) ||
                (rc = az_upload_f32(n, &w.cs1, cs)))
                return rc;
        }
    }
    // ---- depth 1: fold the (constant) cls query into the K projection
    n->folded = false;
    {
        const int H = n->H, hd = D / H;
        const bool pow2 = (hd & (hd - 1)) == 0;
        if (n->L == 1 && pow2 && hd >= 8 && H <= 64 && !getenv("AZ_NO_FOLD")) {
            const std::vector<float>& wi = n->host_folded_w_in;
            const std::vector<float>& bi = n->host_folded_b_in;
            auto r16 = [](float v) { return __bfloat162float(__float2bfloat16(v)); };
            std::vector<double> x0(D), xh(D), q(D);
            double mean = 0, var = 0;
            for (int d = 0; d < D; ++d) {
                const float raw = (*pos)[d] + (*cls)[d];
                x0[d] = n->emb_kblocks ? r16(raw) : raw;    // the tensor-core embedding carries pos/cls in bf16
                mean += x0[d];
            }
            mean /= D;
            for (int d = 0; d < D; ++d) var += (x0[d] - mean) * (x0[d] - mean);
            const double rstd = 1.0 / sqrt(var / D + 1e-5);
            for (int d = 0; d < D; ++d) xh[d] = r16((float)((x0[d] - mean) * rstd));
            for (int r = 0; r < D; ++r) {
                double acc = bi[r];
                for (int k = 0; k < D; ++k) acc += (double)r16(wi[(size_t)r * D + k]) * xh[k];
                q[r] = acc;
            }
            const int N2 = D + 64;
            const double scale = 1.0 / sqrt((double)hd);
            std::vector<float> w2((size_t)N2 * D, 0.f), b2((size_t)N2, 0.f);
            for (int r = 0; r < D; ++r) {
                for (int k = 0; k < D; ++k) w2[(size_t)r * D + k] = wi[(size_t)(2 * D + r) * D + k];
                b2[r] = bi[2 * D + r];
            }
            for (int h = 0; h < H; ++h) {
                double c = 0;
                for (int d = 0; d < hd; ++d) c += q[h * hd + d] * bi[D + h * hd + d];
                b2[D + h] = (float)(c * scale);
                for (int k = 0; k < D; ++k) {
                    double u = 0;
                    for (int d = 0; d < hd; ++d) u += q[h * hd + d] * (double)r16(wi[(size_t)(D + h * hd + d) * D + k]);
                    w2[(size_t)(D + h) * D + k] = (float)(u * scale);
                }
            }
            if ((rc = az_upload_bf16(n, &n->w_vs, w2.data(), w2.size())) || (rc = az_upload_f32(n, &n->b_vs, b2))) return rc;
            if ((rc = az_make_tmap(&n->tm_vs, n->w_vs, N2, D, D, az_pick_bn(N2)))) return rc;
            n->n2 = N2;
            n->folded = true;
        }
    }
    // ---- depth 1, small boards: fold the embedding into the projection as well (az_front.cuh)
    n->front = false;
    {
        const int H = n->H, kc = n->emb_kc, kused = kc + T;
        if (n->folded && n->host_wemb_kpad && cells <= 63 && H <= 8 && kused + 8 <= 128 && P <= 5 && D % 128 == 0 && D <= 512 &&
            !getenv("AZ_NO_FRONT")) {
            const int kpad = n->host_wemb_kpad, KB = kpad / 64;
            const std::vector<float>& wi = n->host_folded_w_in;
            const std::vector<float>& bi = n->host_folded_b_in;
            // Wc: every K-row of Wemb centred over D
            std::vector<double> wc(n->host_wemb);
            for (int k = 0; k < kused; ++k) {
                double m = 0;
                for (int d = 0; d < D; ++d) m += wc[(size_t)k * D + d];
                m /= D;
                for (int d = 0; d < D; ++d) wc[(size_t)k * D + d] -= m;
            }
            // the cls query (a constant of the model) and the folded score rows Ws [H][D], fp64
            std::vector<double> x0(D), q(D), ws((size_t)H * D, 0.0);
            {
                double mean = 0, var = 0;
                for (int d = 0; d < D; ++d) { x0[d] = n->host_wemb[(size_t)kc * D + d]; mean += x0[d]; }
                mean /= D;
                for (int d = 0; d < D; ++d) var += (x0[d] - mean) * (x0[d] - mean);
                const double rstd = 1.0 / sqrt(var / D + 1e-5);
                for (int r = 0; r < D; ++r) {
                    double acc = bi[r];
                    for (int k = 0; k < D; ++k) acc += (double)wi[(size_t)r * D + k] * ((x0[k] - mean) * rstd);
                    q[r] = acc;
                }
                const int hd = D / H;
                const double scale = 1.0 / sqrt((double)hd);
                for (int h = 0; h < H; ++h)
                    for (int d = 0; d < hd; ++d) {
                        const double qq = q[h * hd + d] * scale;
                        const float* wk = &wi[(size_t)(D + h * hd + d) * D];
                        for (int k = 0; k < D; ++k) ws[(size_t)h * D + k] += qq * wk[k];
                    }
            }
            const int NS = (8 + kused + 31) / 32 * 32;
            std::vector<float> wst((size_t)NS * kpad, 0.f), wv2((size_t)D * kpad, 0.f), bv(D), box(D);
            for (int h = 0; h < H; ++h)                                  // rows 0..7: scores
                for (int k = 0; k < kused; ++k) {
                    double acc = 0;
                    for (int d = 0; d < D; ++d) acc += ws[(size_t)h * D + d] * wc[(size_t)k * D + d];
                    wst[(size_t)h * kpad + k] = (float)acc;
                }
            {   // rows 8..: L^T, with G = Wc Wc^T = L L^T (Cholesky; a vanishing pivot leaves a zero column,
                // which is exact for a positive semi-definite G), so that |L^T a|^2 = a^T G a >= 0
                std::vector<double> G((size_t)kused * kused), L((size_t)kused * kused, 0.0);
                double trace = 0;
                for (int i = 0; i < kused; ++i)
                    for (int k = 0; k <= i; ++k) {
                        double acc = 0;
                        for (int d = 0; d < D; ++d) acc += wc[(size_t)i * D + d] * wc[(size_t)k * D + d];
                        G[(size_t)i * kused + k] = G[(size_t)k * kused + i] = acc;
                        if (i == k) trace += acc;
                    }
                const double tol = 1e-12 * (trace > 0 ? trace : 1.0);
                for (int j = 0; j < kused; ++j) {
                    double dj = G[(size_t)j * kused + j];
                    for (int k = 0; k < j; ++k) dj -= L[(size_t)j * kused + k] * L[(size_t)j * kused + k];
                    if (dj <= tol) continue;
                    const double ljj = sqrt(dj);
                    L[(size_t)j * kused + j] = ljj;
                    for (int i = j + 1; i < kused; ++i) {
                        double v = G[(size_t)i * kused + j];
                        for (int k = 0; k < j; ++k) v -= L[(size_t)i * kused + k] * L[(size_t)j * kused + k];
                        L[(size_t)i * kused + j] = v / ljj;
                    }
                }
                for (int j = 0; j < kused; ++j)                          // output column 8 + j = sum_k L[k][j] a_k
                    for (int k = j; k < kused; ++k) wst[(size_t)(8 + j) * kpad + k] = (float)L[(size_t)k * kused + j];
            }
            for (int f = 0; f < D; ++f) {                                // V rows with the embedding folded in
                const float* wr = &wi[(size_t)(2 * D + f) * D];
                for (int k = 0; k < kused; ++k) {
                    double acc = 0;
                    for (int d = 0; d < D; ++d) acc += (double)wr[d] * wc[(size_t)k * D + d];
                    wv2[(size_t)f * kpad + k] = (float)acc;
                }
                bv[f] = bi[2 * D + f];
            }
            const std::vector<float>* bo = az_param(n, "blocks.0.attn.out_proj.bias", (size_t)D, &err);
            if (!bo) return az_fail(AZ_EINVAL, "az_net_finalize: " + err);
            for (int d = 0; d < D; ++d) box[d] = (*bo)[d] + (float)x0[d];   // residual of the cls row is a constant
            if ((rc = az_upload_bf16(n, &n->f_wst, wst.data(), wst.size())) || (rc = az_upload_bf16(n, &n->f_wv, wv2.data(), wv2.size())) ||
                (rc = az_upload_f32(n, &n->f_bias_v, bv)) || (rc = az_upload_f32(n, &n->b_out_x0, box)))
                return rc;
            if ((rc = az_make_tmap(&n->tm_fwst, n->f_wst, NS, kpad, kpad, NS)) || (rc = az_make_tmap(&n->tm_fwv, n->f_wv, D, kpad, kpad, 128)))
                return rc;
            AzFrontArgs& fa = n->fa;
            fa = AzFrontArgs{};
            fa.rows = n->rows; fa.cols = n->cols; fa.planes = C; fa.patch = P; fa.T = T; fa.kc = kc; fa.D = D; fa.H = H;
            fa.TS = (T + 7) / 8 * 8;
            fa.nb = std::min(128 / fa.TS, 128 / H);
            fa.NTP = (fa.nb * fa.TS + 15) / 16 * 16;
            fa.NS = NS;
            fa.KB = KB;
            fa.ksteps = (kused + 15) / 16;
            fa.nslots = std::min(std::min(4, D / 128), 384 / fa.NTP);
            // Slots must not rotate between the two value-issuing warps (a warp a whole use ahead of the other
            // would alias the single parity bit of v_empty): even slot counts keep slot parity == block parity.
            if (fa.nslots == 3 && !getenv("AZ_FRONT_SLOTS3")) fa.nslots = 2;
            fa.bias_v = n->f_bias_v;
            int tpp = 32;
            while (tpp * fa.nb * H > 128) tpp >>= 1;
            n->front = fa.nb >= 1 && fa.nslots >= 1 && tpp >= 1 && (fa.TS + tpp - 1) / tpp <= 8 && az_front_smem(fa) <= 227 * 1024;
            // Reassociated value path: out = sum_h (Wo_h Wv2_h) u_h + (Wo bv + bo + x0), u_h = A^T w_h (az_front2.cuh)
            fa.NP = (fa.nb * H + 15) / 16 * 16;
            n->front2 = n->front && KB == 2 && kc <= 80 && fa.NP <= 128 && az_front2_smem(fa) <= 227 * 1024 && !getenv("AZ_NO_FRONT2");
            if (n->front2) {
                const std::vector<float>* wo = az_param(n, "blocks.0.attn.out_proj.weight", (size_t)D * D, &err);
                if (!wo) return az_fail(AZ_EINVAL, "az_net_finalize: " + err);
                const int hd = D / H, K2 = H * 128;
                // wv2 in fp64 (the bf16 rounding happens once, on the product)
                std::vector<double> wv2d((size_t)D * 128, 0.0);
                for (int f = 0; f < D; ++f) {
                    const float* wr = &wi[(size_t)(2 * D + f) * D];
                    for (int k = 0; k < kused; ++k) {
                        double acc = 0;
                        for (int d = 0; d < D; ++d) acc += (double)wr[d] * wc[(size_t)k * D + d];
                        wv2d[(size_t)f * 128 + k] = acc;
                    }
                }
                std::vector<float> w2((size_t)D * K2, 0.f), b2(D);
                for (int d = 0; d < D; ++d) {
                    const float* wor = &(*wo)[(size_t)d * D];
                    for (int h = 0; h < H; ++h)
                        for (int k = 0; k < kused; ++k) {
                            double acc = 0;
                            for (int j = 0; j < hd; ++j) acc += (double)wor[h * hd + j] * wv2d[(size_t)(h * hd + j) * 128 + k];
                            w2[(size_t)d * K2 + h * 128 + k] = (float)acc;
                        }
                    double acc = box[d];
                    for (int f = 0; f < D; ++f) acc += (double)wor[f] * bv[f];
                    b2[d] = (float)acc;
                }
                if ((rc = az_upload_bf16(n, &n->w_out2, w2.data(), w2.size())) || (rc = az_upload_f32(n, &n->b_out2, b2))) return rc;
            }
        }
    }
    // ---- heads
    AZ_P(gf, "norm.weight", (size_t)D) AZ_P(bf, "norm.bias", (size_t)D)
    AZ_P(wp, "policy_head.weight", (size_t)A * D) AZ_P(bp, "policy_head.bias", (size_t)A)
    AZ_P(wv, "value_head.weight", (size_t)D) AZ_P(bv, "value_head.bias", (size_t)1)
#undef AZ_P
    if ((rc = az_upload_f32(n, &n->lnf_g, *gf)) || (rc = az_upload_f32(n, &n->lnf_b, *bf)) ||
        (rc = az_upload_f32(n, &n->w_pol, *wp)) || (rc = az_upload_f32(n, &n->b_pol, *bp)) ||
        (rc = az_upload_f32(n, &n->w_val, *wv)) || (rc = az_upload_f32(n, &n->b_val, *bv)))
        return rc;
    // ---- activations for one chunk of boards
    // the fused front end keeps every per-token quantity on chip: only per-board rows remain
    n->chunk = n->front ? std::min(n->max_batch, 65536) : std::min(n->max_batch, 4096);
    const size_t B = (size_t)n->chunk, M = n->front ? 128 : B * T;
    const bool deep = n->L > 1;
    if ((rc = az_net_alloc(n, &n->x, M * D)) || (rc = az_net_alloc(n, &n->xn, M * D)) ||
        (rc = az_net_alloc(n, &n->qkv, M * (deep ? 3 : 2) * D)) ||
        (rc = az_net_alloc(n, &n->qc, B * D)) || (rc = az_net_alloc(n, &n->oc, B * D)) ||
        (rc = az_net_alloc(n, &n->xc, B * D)) || (rc = az_net_alloc(n, &n->xcn, B * D)) ||
        (rc = az_net_alloc(n, &n->hc, B * 4 * D)) || (rc = az_net_alloc(n, &n->xc2, B * D)))
        return rc;
    if (n->folded && (rc = az_net_alloc(n, &n->sc, M * 64))) return rc;
    if (deep && ((rc = az_net_alloc(n, &n->ao, M * D)) || (rc = az_net_alloc(n, &n->hbuf, M * 4 * D)))) return rc;
    const int nw = cells <= 64 ? 1 : AZ_MAX_WORDS;
    if ((rc = az_net_alloc(n, &n->in_words, B * 2 * nw)) || (rc = az_net_alloc(n, &n->in_player, B)) ||
        (rc = az_net_alloc(n, &n->out_priors, B * A)) || (rc = az_net_alloc(n, &n->out_values, B)) ||
        (rc = az_net_alloc(n, &n->out_logits, B * A)))
        return rc;
    AZ_CUDA_OK(cudaMemset(n->x, 0, M * D * 2));
    AZ_CUDA_OK(cudaMemset(n->xn, 0, M * D * 2));
    if ((rc = az_make_tmap(&n->tm_xn, n->xn, M, D, D, 128)) ||
        (rc = az_make_tmap(&n->tm_xn_cls, n->xn, n->front ? 1 : B, D, (uint64_t)T * D, 128)) ||
        (rc = az_make_tmap(&n->tm_oc, n->oc, B, D, D, 128)) || (rc = az_make_tmap(&n->tm_xcn, n->xcn, B, D, D, 128)) ||
        (rc = az_make_tmap(&n->tm_hc, n->hc, B, 4 * D, 4 * D, 128)))
        return rc;
    if (deep && ((rc = az_make_tmap(&n->tm_ao, n->ao, M, D, D, 128)) || (rc = az_make_tmap(&n->tm_hbuf, n->hbuf, M, 4 * D, 4 * D, 128))))
        return rc;
    n->tm_x_st = n->tm_xn;                                    // placeholders unless the embedding kernel stores through TMA
    n->tm_xn_st = n->tm_xn;
    if (!n->front && n->emb_kblocks && D % 64 == 0 && !getenv("AZ_NO_TMA_STORE") &&
        az_embed_mma_smem(D, n->emb_kblocks, T, n->patch * n->patch, true) <= 200 * 1024) {
        if ((rc = az_make_tmap(&n->tm_x_st, n->x, M, D, D, 32)) || (rc = az_make_tmap(&n->tm_xn_st, n->xn, M, D, D, 32))) return rc;
        n->emb_tma_out = true;
    }
    if (deep && az_attn_supported(T, D, n->H) && !getenv("AZ_NO_ATTN_MMA")) {
        // key / value rows past a board's last token are read (and meet probability 0): they have to be finite from the start
        AZ_CUDA_OK(cudaMemset(n->qkv, 0, M * 3 * D * sizeof(bf16)));
        if ((rc = az_make_tmap(&n->tm_qkv, n->qkv, M, 3 * D, 3 * D, 64))) return rc;
        n->attn_mma = true;
    }
    // CTA-pair kernels for the cls-row GEMMs: measured (tools/gemm_bench.py) 7 % faster on the K = 4D GEMM at 65536 rows, not
    // faster at 4096 rows (one tile per CTA: start-up latency, not L2 traffic, bounds those); AZ_PAIR_MIN_ROWS overrides
    const int pair_min = getenv("AZ_PAIR_MIN_ROWS") ? atoi(getenv("AZ_PAIR_MIN_ROWS")) : 16384;
    const bool pair_ok = az_pair_enabled() && n->chunk >= pair_min && D % 128 == 0;
    if (n->front2) {
        const int K2 = n->H * 128;
        n->bn_out2 = pair_ok ? az_pick_bn_pair(D, n->chunk) : az_pick_bn_small_m(D, n->chunk);
        if ((rc = az_net_alloc(n, &n->uc, (B + 128) * K2)) || (rc = az_make_tmap(&n->tm_uc, n->uc, B, K2, K2, 128)) ||
            (rc = az_make_tmap(&n->tm_wout2, n->w_out2, D, K2, K2, az_w_box_rows(n->bn_out2))))
            return rc;
    }
    {   // cls-row GEMMs of the last block: M = boards of one chunk, tile width chosen for that M
        AzBlockW& w = n->blk[n->L - 1];
        w.bn_out_c = pair_ok ? az_pick_bn_pair(D, n->chunk) : az_pick_bn_small_m(D, n->chunk);
        w.bn_w1_c = pair_ok ? az_pick_bn_pair(4 * D, n->chunk) : az_pick_bn_small_m(4 * D, n->chunk);
        w.bn_w2_c = pair_ok ? az_pick_bn_pair(D, n->chunk) : az_pick_bn_small_m(D, n->chunk);
        if ((rc = az_make_tmap(&w.tm_out_c, w.w_out, D, D, D, az_w_box_rows(w.bn_out_c))) ||
            (rc = az_make_tmap(&w.tm_w1_c, w.w1, 4 * D, D, D, az_w_box_rows(w.bn_w1_c))) ||
            (rc = az_make_tmap(&w.tm_w2_c, w.w2, D, 4 * D, 4 * D, az_w_box_rows(w.bn_w2_c))))
            return rc;
    }
    n->tail_fold = !getenv("AZ_NO_TAIL_FOLD");
    n->head_fold = n->tail_fold && A + 1 <= 64;
    if (n->tail_fold) {
        AzBlockW& w = n->blk[n->L - 1];
        if ((rc = az_make_tmap(&w.tm_w1_ln_c, w.w1_ln, 4 * D, D, D, az_w_box_rows(w.bn_w1_c))) ||
            (rc = az_net_alloc(n, &n->stats1, B * AZ_STATS_LD)) || (rc = az_net_alloc(n, &n->stats2, B * AZ_STATS_LD)) ||
            (rc = az_make_tmap(&n->tm_xc, n->xc, B, D, D, 128)) || (rc = az_make_tmap(&n->tm_xc2, n->xc2, B, D, D, 128)))
            return rc;
    }
    if (n->head_fold) {                                       // final norm folded into [policy ; value] rows
        auto r16 = [](float v) { return __bfloat162float(__float2bfloat16(v)); };
        std::vector<float> wh((size_t)64 * D, 0.f), bh(64, 0.f), cs(64, 0.f);
        for (int o = 0; o <= A; ++o) {
            const float* src = o < A ? wp->data() + (size_t)o * D : wv->data();
            double acc = o < A ? (*bp)[o] : (*bv)[0], sum = 0.0;
            for (int d = 0; d < D; ++d) {
                const float v = src[d] * (*gf)[d];
                wh[(size_t)o * D + d] = v;
                sum += (double)r16(v);
                acc += (double)src[d] * (*bf)[d];
            }
            bh[o] = (float)acc;
            cs[o] = (float)sum;
        }
        if ((rc = az_upload_bf16(n, &n->w_head_ln, wh.data(), wh.size())) || (rc = az_upload_f32(n, &n->b_head_ln, bh)) ||
            (rc = az_upload_f32(n, &n->cs_head, cs)) || (rc = az_make_tmap(&n->tm_whead_ln, n->w_head_ln, 64, D, D, 64)))
            return rc;
    }
    {   // fused cls-row tail (az_tail.cuh): one cluster kernel instead of four GEMM launches for evaluation batches of up to
        // AZ_TAIL_MAX_ROWS boards.  Opt-in (AZ_FUSED_TAIL=1): measured on B200 it is 58 us per pass against 70 us for the chain
        // timed kernel by kernel, but inside the search graph (PDL, back to back) the chain is the faster one: Connect4
        // 66.6 M simulations/s with the chain, 64.1 M with this kernel.
        const int tail_max = getenv("AZ_TAIL_MAX_ROWS") ? atoi(getenv("AZ_TAIL_MAX_ROWS")) : 4096;
        const bool want = getenv("AZ_FUSED_TAIL") && atoi(getenv("AZ_FUSED_TAIL")) != 0;
        if (want && n->front2 && n->tail_fold && n->head_fold && D % 128 == 0 && D <= 512 && n->chunk <= tail_max) {
            AzBlockW& w = n->blk[n->L - 1];
            const int K2 = n->H * 128;
            if ((rc = az_make_tmap(&n->tm_t_u, n->uc, B, K2, K2, 64)) || (rc = az_make_tmap(&n->tm_t_wo, n->w_out2, D, K2, K2, 128)) ||
                (rc = az_make_tmap(&n->tm_t_w1, w.w1_ln, 4 * D, D, D, 128)) || (rc = az_make_tmap(&n->tm_t_w2, w.w2, D, 4 * D, 4 * D, 128)))
                return rc;
            n->fused_tail = true;
        }
    }
    {   // the four cls-row GEMMs as one persistent kernel (az_chain.cuh) for evaluation batches of up to AZ_CHAIN_MAX_ROWS boards.
        // Opt-in (AZ_CHAIN=1): bit-identical to the four launches and 64 us against 70 us when the kernels are timed one by one,
        // but slower inside the search graph (Connect4 62.9 M simulations/s against 66.8 M): a layer still costs ~16 us without
        // its launch and prologue, and the grid-wide barriers give up the overlap programmatic dependent launch provides.
        AzBlockW& w = n->blk[n->L - 1];
        const int chain_max = getenv("AZ_CHAIN_MAX_ROWS") ? atoi(getenv("AZ_CHAIN_MAX_ROWS")) : 8192;
        auto ok_bn = [](int bn) { return bn == 128 || bn == 256; };
        const bool want = getenv("AZ_CHAIN") && atoi(getenv("AZ_CHAIN")) != 0;
        if (want && n->front2 && n->tail_fold && n->head_fold && !n->fused_tail && n->chunk <= chain_max && ok_bn(n->bn_out2) && ok_bn(w.bn_w1_c) &&
            ok_bn(w.bn_w2_c)) {
            if ((rc = az_net_alloc(n, &n->chain_sync, 2))) return rc;
            AZ_CUDA_OK(cudaMemset(n->chain_sync, 0, 2 * sizeof(unsigned int)));
            n->chain = true;
        }
    }
    n->head_gemm = false;
    if (!n->head_fold && A + 1 > 32 && A + 1 <= 64 && !getenv("AZ_NO_HEAD_GEMM")) {   // heads as one N = 64 tensor-core GEMM
        std::vector<float> wh((size_t)64 * D, 0.f), bh(64, 0.f);
        for (int o = 0; o < A; ++o) {
            for (int d = 0; d < D; ++d) wh[(size_t)o * D + d] = (*wp)[(size_t)o * D + d];
            bh[o] = (*bp)[o];
        }
        for (int d = 0; d < D; ++d) wh[(size_t)A * D + d] = (*wv)[d];
        bh[A] = (*bv)[0];
        if ((rc = az_upload_bf16(n, &n->w_head, wh.data(), wh.size())) || (rc = az_upload_f32(n, &n->b_head, bh)) ||
            (rc = az_net_alloc(n, &n->head_out, B * 64)) || (rc = az_make_tmap(&n->tm_whead, n->w_head, 64, D, D, 64)))
            return rc;
        n->head_gemm = true;
    }
    AZ_CUDA_OK(cudaDeviceSynchronize());
    n->finalized = true;
    return AZ_OK;
}

// ---- forward pass over one chunk of boards ------------------------------------------------
static AzGemmArgs az_gargs(const AzBatch& b, int rows_per_item, const float* bias, const bf16* resid, int ldr, bf16* out,
                           int ldo, int N, int K) {
    AzGemmArgs a{};
    a.bias = bias; a.resid = resid; a.ldr = ldr; a.out = out; a.ldo = ldo; a.N = N; a.K = K;
    a.m_fixed = b.fixed; a.m_count = b.count; a.m_item_offset = b.offset; a.m_item_cap = b.cap;
    a.m_rows_per_item = rows_per_item;
    if (!b.count) {                                         // host-known batch: rows directly
        int items = b.fixed - b.offset;
        items = items < 0 ? 0 : (items > b.cap ? b.cap : items);
        a.m_fixed = items * rows_per_item;
    }
    return a;
}

template <typename F>
static void az_ln_dispatch(int D, F&& f) {
    if (D <= 256) f(std::integral_constant<int, 1>());
    else if (D <= 512) f(std::integral_constant<int, 2>());
    else f(std::integral_constant<int, 4>());
}

static int az_net_forward_chunk(az_net* n, AzBatch batch, const uint64_t* words, const uint8_t* players, int nw,
                                float* priors, float* values, float* logits, cudaStream_t s) {
    const int D = n->D, T = n->T, H = n->H, L = n->L, A = n->A;
    const int cap = batch.cap;                               // grid bound (items)
    const long long max_rows = (long long)cap * T;
    int rc;
    auto layernorm = [&](int rows_per_item, const bf16* in, int ld_in, bf16* out, int ld_out, const float* g, const float* b) {
        AzProfScope ps(n, AZ_K_LN, s);
        const long long rows = (long long)cap * rows_per_item;
        const int grid = (int)((rows + 7) / 8);
        az_ln_dispatch(D, [&](auto nv) {
            az_layernorm_kernel<decltype(nv)::value><<<grid, 256, 0, s>>>(batch, rows_per_item, in, ld_in, out, ld_out, D, g, b);
        });
        ++n->launches;
    };
    auto gemm = [&](int kind, const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& ga, int epi, int mr) {
        AzProfScope ps(n, kind, s);
        return az_launch_gemm(ta, tb, ga, epi, mr, s);
    };
    static const bool l2_hints = !getenv("AZ_NO_L2_HINTS");
    static const bool w_early = getenv("AZ_W_EARLY") != nullptr;   // measured: no gain (41.8 vs 41.8 M sims/s on Connect4), off by default
    auto gemm_bn = [&](int kind, const CUtensorMap& ta, const CUtensorMap& tb, AzGemmArgs ga, int epi, int mr, int bn) {
        AzProfScope ps(n, kind, s);
        ga.l2_hints = l2_hints ? 1 : 0;
        ga.w_early = w_early ? 1 : 0;                          // the leaf counter was written by the advance kernel, >= 2 kernels upstream
        return az_launch_gemm(ta, tb, ga, epi, mr, s, bn);
    };
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    // Net.norm + heads + softmax (ai/nn.py:78-83, ai/mcts.py:49) on the cls rows xc_final [cap][D]
    auto heads = [&](const bf16* xc_final) -> int {
        AzHeadArgs ha{};
        ha.batch = batch; ha.xc = xc_final; ha.D = D; ha.A = A; ha.ln_g = n->lnf_g; ha.ln_b = n->lnf_b;
        ha.w_pol = n->w_pol; ha.b_pol = n->b_pol; ha.w_val = n->w_val; ha.b_val = n->b_val;
        ha.priors = priors; ha.values = values; ha.logits = logits;
        if (n->head_fold) {                                    // one GEMM: raw rows x folded head weights, norm + softmax in the epilogue
            AzGemmArgs ga = az_gargs(batch, 1, n->b_head_ln, nullptr, 0, nullptr, 64, 64, D);
            ga.stats_in = n->stats2;
            ga.stats_parts = 2 * (D / az_bn_width(n->blk[L - 1].bn_w2_c));
            ga.colsum = n->cs_head;
            ga.ln_inv_k = 1.0f / (float)D;
            ga.ln_eps = 1e-5f;
            ga.head_A = A;
            ga.priors = priors; ga.values = values; ga.logits = logits;
            int rc2 = gemm_bn(AZ_K_HEAD, xc_final == n->xc2 ? n->tm_xc2 : n->tm_xc, n->tm_whead_ln, ga, AZ_EPI_HEAD, cap, 64);
            if (rc2) return rc2;
            ++n->launches;
            return AZ_OK;
        }
        if (n->head_gemm) {                                    // LayerNorm -> [cap][D] x [64][D]^T on tensor cores -> softmax
            layernorm(1, xc_final, D, n->xcn, D, n->lnf_g, n->lnf_b);
            AzGemmArgs ga = az_gargs(batch, 1, n->b_head, nullptr, 0, n->xcn, 64, 64, D);
            ga.out_f32 = n->head_out;
            ga.f32_col0 = 0;
            int rc2 = gemm_bn(AZ_K_HEAD, n->tm_xcn, n->tm_whead, ga, AZ_EPI_BIAS, cap, 64);
            if (rc2) return rc2;
            {
                AzProfScope ps(n, AZ_K_HEAD, s);
                az_head_softmax_kernel<<<(cap + 7) / 8, 256, 0, s>>>(batch, n->head_out, A, priors, values, logits);
            }
            AZ_CUDA_OK(cudaGetLastError());
            n->launches += 2;
            return AZ_OK;
        }
        if (A + 1 <= 32 && (size_t)(A + 1) * D * 4 <= 96 * 1024) {   // one warp per board, weights in shared memory
            const size_t smem = (size_t)(A + 1) * D * sizeof(float);
            const int grid = std::max(1, std::min((cap + 7) / 8, g_num_sms * 4));
            AzProfScope ps(n, AZ_K_HEAD, s);
            az_ln_dispatch(D, [&](auto nv) {
                constexpr int NV = decltype(nv)::value;
                if (smem > 48 * 1024) cudaFuncSetAttribute(az_head_warp_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                az_head_warp_kernel<NV><<<grid, 256, smem, s>>>(ha);
            });
        } else {
            AzProfScope ps(n, AZ_K_HEAD, s);
            az_head_kernel<<<cap, 128, (size_t)(D + A + 8) * sizeof(float), s>>>(ha);
        }
        AZ_CUDA_OK(cudaGetLastError());
        ++n->launches;
        return AZ_OK;
    };
    // out-proj -> norm2 -> mlp (ai/nn.py:44-50) on the cls rows: oc [cap][D] -> xc (residual stream) -> xc2.
    // With tail_fold the out-proj epilogue emits the row statistics and mlp1 applies norm2 in its own
    // epilogue (az_gemm.cuh); mlp2 emits the statistics the head GEMM needs.
    auto cls_tail = [&](AzBlockW& w, AzGemmArgs out_args, int out_epi, bool from_u = false) -> int {
        int rc2;
        const CUtensorMap& tm_oa = from_u ? n->tm_uc : n->tm_oc;         // out-proj input: U vectors (front2) or oc
        const CUtensorMap& tm_ow = from_u ? n->tm_wout2 : w.tm_out_c;
        const int bn_out = from_u ? n->bn_out2 : w.bn_out_c;
        if (n->tail_fold) {
            out_args.stats_out = n->stats1;
            if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, tm_oa, tm_ow, out_args, out_epi, cap, bn_out))) return rc2;
            AzGemmArgs g1 = az_gargs(batch, 1, w.b1_ln, nullptr, 0, n->hc, 4 * D, 4 * D, D);
            g1.stats_in = n->stats1;
            g1.stats_parts = 2 * (D / az_bn_width(bn_out));
            g1.colsum = w.cs1;
            g1.ln_inv_k = 1.0f / (float)D;
            g1.ln_eps = 1e-5f;
            if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, n->tm_xc, w.tm_w1_ln_c, g1, AZ_EPI_GELU, cap, w.bn_w1_c))) return rc2;
            AzGemmArgs g2 = az_gargs(batch, 1, w.b2, n->xc, D, n->xc2, D, D, 4 * D);
            if (n->head_fold) g2.stats_out = n->stats2;
            if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, n->tm_hc, w.tm_w2_c, g2, AZ_EPI_RESID, cap, w.bn_w2_c))) return rc2;
            n->launches += 3;
            return AZ_OK;
        }
        if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, tm_oa, tm_ow, out_args, out_epi, cap, bn_out))) return rc2;
        layernorm(1, n->xc, D, n->xcn, D, w.ln2_g, w.ln2_b);
        if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, n->tm_xcn, w.tm_w1_c, az_gargs(batch, 1, w.b1, nullptr, 0, n->hc, 4 * D, 4 * D, D), AZ_EPI_GELU, cap, w.bn_w1_c))) return rc2;
        if ((rc2 = gemm_bn(AZ_K_GEMM_CLS, n->tm_hc, w.tm_w2_c, az_gargs(batch, 1, w.b2, n->xc, D, n->xc2, D, D, 4 * D), AZ_EPI_RESID, cap, w.bn_w2_c))) return rc2;
        n->launches += 3;
        return AZ_OK;
    };
    if (n->front) {
        // embedding + LN1 + V/score projection + cls attention, all on chip (az_front.cuh)
        AzBlockW& w = n->blk[0];
        AzFrontArgs fa = n->fa;
        fa.batch = batch; fa.leaf_words = words; fa.leaf_player = players; fa.oc = n->oc; fa.uc = n->uc;
        const size_t smem = n->front2 ? az_front2_smem(fa) : az_front_smem(fa);
        const int tiles = (cap + fa.nb - 1) / fa.nb;
        const int grid = tiles < g_num_sms ? (tiles > 0 ? tiles : 1) : g_num_sms;
        {
            AzProfScope ps(n, AZ_K_FRONT, s);
#define AZ_FRONT(TS8_)                                                                                                     \
    case TS8_: {                                                                                                           \
        static bool done = false;                                                                                          \
        if (!done) {                                                                                                       \
            AZ_CUDA_OK(cudaFuncSetAttribute(az_front_kernel<TS8_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)); \
            AZ_CUDA_OK(cudaFuncSetAttribute(az_front2_kernel<TS8_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)); \
            done = true;                                                                                                   \
        }                                                                                                                  \
        if (n->front2) AZ_CUDA_OK(az_launch_pdl(az_front2_kernel<TS8_>, grid, AZ_FRONT_THREADS, smem, s, n->tm_fwst, fa)); \
        else AZ_CUDA_OK(az_launch_pdl(az_front_kernel<TS8_>, grid, AZ_FRONT_THREADS, smem, s, n->tm_fwst, n->tm_fwv, fa)); \
    } break;
            switch (fa.TS / 8) {
                AZ_FRONT(1) AZ_FRONT(2) AZ_FRONT(3) AZ_FRONT(4) AZ_FRONT(5) AZ_FRONT(6) AZ_FRONT(7) AZ_FRONT(8)
                default: return az_fail(AZ_ESTATE, "front kernel: unsupported token stride");
            }
#undef AZ_FRONT
        }
        AZ_CUDA_OK(cudaGetLastError());
        ++n->launches;
        if (n->fused_tail) {
            AzTailArgs ta{};
            ta.batch = batch; ta.D = D; ta.K2 = n->H * 128; ta.A = A; ta.C = D / 128;
            ta.b_out = n->b_out2; ta.b1 = w.b1_ln; ta.cs1 = w.cs1; ta.b2 = w.b2; ta.b_head = n->b_head_ln; ta.cs_head = n->cs_head;
            ta.priors = priors; ta.values = values; ta.logits = logits; ta.ln_eps = 1e-5f;
            static bool attr_done = false;
            if (!attr_done) {
                AZ_CUDA_OK(cudaFuncSetAttribute(az_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AZ_TAIL_SMEM));
                attr_done = true;
            }
            const int tiles = (cap + AZ_TAIL_ROWS - 1) / AZ_TAIL_ROWS;
            const int clusters = std::max(1, std::min(tiles, g_num_sms / ta.C));
            static const bool pdl = !getenv("AZ_NO_PDL");
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(clusters * ta.C));
            cfg.blockDim = dim3(AZ_TAIL_THREADS);
            cfg.dynamicSmemBytes = AZ_TAIL_SMEM;
            cfg.stream = s;
            cudaLaunchAttribute attr[2];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned)ta.C;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = pdl ? 2 : 1;
            {
                AzProfScope ps(n, AZ_K_TAIL, s);
                AZ_CUDA_OK(cudaLaunchKernelEx(&cfg, az_tail_kernel, n->tm_t_u, n->tm_t_wo, n->tm_t_w1, n->tm_t_w2, n->tm_whead_ln, ta));
            }
            AZ_CUDA_OK(cudaGetLastError());
            ++n->launches;
            return AZ_OK;
        }
        if (n->chain) {
            AzChainMaps cm;
            AzChainArgs ca{};
            // the same arguments cls_tail() and heads() give the four separate launches
            ca.g[0] = az_gargs(batch, 1, n->b_out2, nullptr, 0, n->xc, D, D, n->H * 128);
            ca.g[0].stats_out = n->stats1;
            ca.g[1] = az_gargs(batch, 1, w.b1_ln, nullptr, 0, n->hc, 4 * D, 4 * D, D);
            ca.g[1].stats_in = n->stats1;
            ca.g[1].stats_parts = 2 * (D / n->bn_out2);
            ca.g[1].colsum = w.cs1;
            ca.g[1].ln_inv_k = 1.0f / (float)D;
            ca.g[1].ln_eps = 1e-5f;
            ca.g[2] = az_gargs(batch, 1, w.b2, n->xc, D, n->xc2, D, D, 4 * D);
            ca.g[2].stats_out = n->stats2;
            ca.g[3] = az_gargs(batch, 1, n->b_head_ln, nullptr, 0, nullptr, 64, 64, D);
            ca.g[3].stats_in = n->stats2;
            ca.g[3].stats_parts = 2 * (D / w.bn_w2_c);
            ca.g[3].colsum = n->cs_head;
            ca.g[3].ln_inv_k = 1.0f / (float)D;
            ca.g[3].ln_eps = 1e-5f;
            ca.g[3].head_A = A;
            ca.g[3].priors = priors; ca.g[3].values = values; ca.g[3].logits = logits;
            ca.bn[0] = n->bn_out2; ca.bn[1] = w.bn_w1_c; ca.bn[2] = w.bn_w2_c; ca.bn[3] = 64;
            ca.sync = n->chain_sync;
            cm.a[0] = n->tm_uc; cm.a[1] = n->tm_xc; cm.a[2] = n->tm_hc; cm.a[3] = n->tm_xc2;
            cm.b[0] = n->tm_wout2; cm.b[1] = w.tm_w1_ln_c; cm.b[2] = w.tm_w2_c; cm.b[3] = n->tm_whead_ln;
            int tiles_max = 1;
            for (int l = 0; l < 3; ++l) {
                ca.g[l].l2_hints = 1;
                cm.o[l] = cm.a[l];
                if ((rc = az_gemm_out_map<256, AZ_EPI_BIAS>(&ca.g[l], &cm.o[l], cap))) return rc;       // box 64 x 32 whatever the tile width
                tiles_max = std::max(tiles_max, ((cap + 127) / 128) * (ca.g[l].N / ca.bn[l]));
            }
            cm.o[3] = cm.a[3];
            static bool attr_done = false;
            if (!attr_done) {
                AZ_CUDA_OK(cudaFuncSetAttribute(az_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AZ_CHAIN_SMEM));
                attr_done = true;
            }
            const int grid = std::min(tiles_max, g_num_sms);
            {
                AzProfScope ps(n, AZ_K_TAIL, s);
                AZ_CUDA_OK(az_launch_pdl(az_chain_kernel, grid, AZ_GEMM_THREADS, (size_t)AZ_CHAIN_SMEM, s, cm, ca));
            }
            AZ_CUDA_OK(cudaGetLastError());
            ++n->launches;
            return AZ_OK;
        }
        if (n->front2) {
            if ((rc = cls_tail(w, az_gargs(batch, 1, n->b_out2, nullptr, 0, n->xc, D, D, n->H * 128), AZ_EPI_BIAS, true))) return rc;
            return heads(n->xc2);
        }
        if ((rc = cls_tail(w, az_gargs(batch, 1, n->b_out_x0, nullptr, 0, n->xc, D, D, D), AZ_EPI_BIAS))) return rc;
        return heads(n->xc2);
    }
    // 1. boards -> tokens -> LN1 of block 0 (affine folded into block 0's in_proj at load time)
    if (n->emb_kblocks) {
        AzProfScope ps(n, AZ_K_EMBED, s);
        AzEmbedMmaArgs ma{};
        ma.batch = batch; ma.leaf_words = words; ma.leaf_player = players; ma.nw = nw;
        ma.rows = n->rows; ma.cols = n->cols; ma.planes = n->C; ma.patch = n->patch; ma.D = D; ma.T = T;
        ma.kc = n->emb_kc; ma.kblocks = n->emb_kblocks; ma.x = n->x; ma.xn = n->xn; ma.write_all_x = L > 1;
        ma.tma_out = n->emb_tma_out ? 1 : 0;
        const size_t smem = az_embed_mma_smem(D, n->emb_kblocks, T, n->patch * n->patch, n->emb_tma_out);
        const int tiles = (int)((max_rows + 127) / 128);
        const int grid = tiles < g_num_sms ? (tiles > 0 ? tiles : 1) : g_num_sms;
#define AZ_EMB(COLS)                                                                                                  \
    {                                                                                                                 \
        static bool done = false;                                                                                     \
        if (!done) {                                                                                                  \
            AZ_CUDA_OK(cudaFuncSetAttribute(az_embed_mma_kernel<COLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
            done = true;                                                                                              \
        }                                                                                                             \
        az_embed_mma_kernel<COLS><<<grid, 288, smem, s>>>(n->tm_emb, n->tm_x_st, n->tm_xn_st, ma);                     \
    }
        if (D <= 64) AZ_EMB(64) else if (D <= 128) AZ_EMB(128) else if (D <= 256) AZ_EMB(256) else AZ_EMB(512)
#undef AZ_EMB
        AZ_CUDA_OK(cudaGetLastError());
        ++n->launches;
    } else {
        AzEmbedArgs ea{};
        ea.batch = batch; ea.leaf_words = words; ea.leaf_player = players; ea.nw = nw;
        ea.rows = n->rows; ea.cols = n->cols; ea.planes = n->C; ea.patch = n->patch; ea.D = D; ea.T = T;
        ea.wt = n->wt_embed; ea.posb = n->posb; ea.ln_g = n->ones; ea.ln_b = n->zeros;
        ea.x = n->x; ea.xn = n->xn; ea.write_all_x = L > 1;
        AzProfScope ps(n, AZ_K_EMBED, s);
        const int grid = (int)((max_rows + 7) / 8);
        az_ln_dispatch(D, [&](auto nv) { az_embed_ln_kernel<decltype(nv)::value><<<grid, 256, 0, s>>>(ea); });
        AZ_CUDA_OK(cudaGetLastError());
        ++n->launches;
    }
    for (int l = 0; l < L - 1; ++l) {                        // full blocks (ai/nn.py:52-61)
        AzBlockW& w = n->blk[l];
        if (l > 0) layernorm(T, n->x, D, n->xn, D, w.ln1_g, w.ln1_b);
        if ((rc = gemm(AZ_K_GEMM_FULL, n->tm_xn, w.tm_in, az_gargs(batch, T, w.b_in, nullptr, 0, n->qkv, 3 * D, 3 * D, D), AZ_EPI_BIAS, (int)max_rows))) return rc;
        const int hd = D / H;
        if (n->attn_mma) {
            AzAttnArgs aa{};
            aa.batch = batch; aa.out = n->ao; aa.T = T; aa.D = D; aa.H = H; aa.hd = hd;
            aa.pack = T <= 64 ? 1 : 0;
            aa.KT = aa.pack ? 2 : (T + 63) / 64; aa.Tp = (T + 15) & ~15; aa.NS = aa.pack ? 128 : aa.Tp;
            aa.groups = D / 64; aa.hpg = 64 / hd; aa.hpg_log2 = hd == 64 ? 0 : (hd == 32 ? 1 : 2); aa.nq = (T + 127) / 128;
            aa.nbuf = aa.KT <= 2 ? 2 : 1; aa.nstg = aa.pack ? 3 : aa.nbuf;
            aa.sl2e = 1.4426950408889634f / sqrtf((float)hd);
            const size_t smem = az_attn_smem(aa);
            static bool attr_done = false;
            if (!attr_done) {
                AZ_CUDA_OK(cudaFuncSetAttribute(az_attn_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                attr_done = true;
            }
            const long long work = (long long)(aa.pack ? (cap + 1) / 2 : cap) * aa.groups;
            const int grid = (int)std::min<long long>(work > 0 ? work : 1, (long long)g_num_sms);     // one CTA per SM (TMEM, shared memory)
            AzProfScope ps(n, AZ_K_ATT_FULL, s);
            az_attn_mma_kernel<<<grid, AZ_ATTN_THREADS, smem, s>>>(n->tm_qkv, aa);
        } else {
            const size_t smem = ((size_t)2 * T * (hd + 1) + 4 * T) * sizeof(float);
            if (smem > 48 * 1024) AZ_CUDA_OK(cudaFuncSetAttribute(az_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            AzProfScope ps(n, AZ_K_ATT_FULL, s);
            az_attention_kernel<<<cap * H, 128, smem, s>>>(batch, n->qkv, n->ao, T, D, H);
        }
        AZ_CUDA_OK(cudaGetLastError());
        if ((rc = gemm(AZ_K_GEMM_FULL, n->tm_ao, w.tm_out, az_gargs(batch, T, w.b_out, n->x, D, n->x, D, D, D), AZ_EPI_RESID, (int)max_rows))) return rc;
        layernorm(T, n->x, D, n->xn, D, w.ln2_g, w.ln2_b);
        if ((rc = gemm(AZ_K_GEMM_FULL, n->tm_xn, w.tm_w1, az_gargs(batch, T, w.b1, nullptr, 0, n->hbuf, 4 * D, 4 * D, D), AZ_EPI_GELU, (int)max_rows))) return rc;
        if ((rc = gemm(AZ_K_GEMM_FULL, n->tm_hbuf, w.tm_w2, az_gargs(batch, T, w.b2, n->x, D, n->x, D, D, 4 * D), AZ_EPI_RESID, (int)max_rows))) return rc;
        n->launches += 5;
    }
    {                                                        // last block, cls-pruned
        AzBlockW& w = n->blk[L - 1];
        if (L > 1) layernorm(T, n->x, D, n->xn, D, w.ln1_g, w.ln1_b);
        if (n->folded) {
            // V of every token + the 64 folded score columns (fp32) in one projection
            AzGemmArgs ga = az_gargs(batch, T, n->b_vs, nullptr, 0, n->qkv, n->n2, n->n2, D);
            ga.out_f32 = n->sc;
            ga.f32_col0 = D;
            if ((rc = gemm(AZ_K_GEMM_KV, n->tm_xn, n->tm_vs, ga, AZ_EPI_BIAS, (int)max_rows))) return rc;
            {
                AzProfScope ps(n, AZ_K_CLS_ATT, s);
                az_cls_attention_folded_kernel<<<(cap * H + 3) / 4, 128, 0, s>>>(batch, n->sc, n->qkv, n->n2, n->oc, T, D, H);
            }
            AZ_CUDA_OK(cudaGetLastError());
            n->launches -= 1;                                 // one kernel fewer than the generic path
        } else {
        // K,V of every token; Q of the cls token only (rows of xn with stride T*D)
        if ((rc = gemm(AZ_K_GEMM_KV, n->tm_xn, w.tm_kv, az_gargs(batch, T, w.b_in + D, nullptr, 0, n->qkv, 2 * D, 2 * D, D), AZ_EPI_BIAS, (int)max_rows))) return rc;
        if ((rc = gemm(AZ_K_GEMM_CLS, n->tm_xn_cls, w.tm_q, az_gargs(batch, 1, w.b_in, nullptr, 0, n->qc, D, D, D), AZ_EPI_BIAS, cap))) return rc;
        {
            AzProfScope ps(n, AZ_K_CLS_ATT, s);
            const int hd = D / H;
            static const bool old_kernel = getenv("AZ_NO_CLS_ATT2") != nullptr;
            const int grid = (cap * H + 3) / 4;
            if (!old_kernel && hd == 32 && T <= 64) az_cls_attention2_kernel<32, 2><<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
            else if (!old_kernel && hd == 64 && T <= 64) az_cls_attention2_kernel<64, 2><<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
            else if (!old_kernel && hd == 16 && T <= 64) az_cls_attention2_kernel<16, 2><<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
            else if (!old_kernel && hd == 32 && T <= 128) az_cls_attention2_kernel<32, 4><<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
            else if (!old_kernel && hd == 16 && T <= 256) az_cls_attention2_kernel<16, 8><<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
            else az_cls_attention_kernel<<<grid, 128, 0, s>>>(batch, n->qc, n->qkv, n->oc, T, D, H);
        }
        AZ_CUDA_OK(cudaGetLastError());
        }
        if ((rc = cls_tail(w, az_gargs(batch, 1, w.b_out, n->x, T * D, n->xc, D, D, D), AZ_EPI_RESID))) return rc;
        n->launches += 3;
    }
    return heads(n->xc2);
}

// Evaluate the forest's pending leaves (the `model(...)` call of ai/mcts.py:46 for the whole
// batch): priors/values land in the forest's evaluation buffers.  Asynchronous, capturable.
extern "C" int az_net_eval_leaves(az_net* n, az_forest* f, void* stream) {
    if (!n || !f) return az_fail(AZ_EINVAL, "az_net_eval_leaves: null");
    if (!n->finalized) return az_fail(AZ_ESTATE, "az_net_eval_leaves: call az_net_finalize first");
    AzForestDev* d = az_forest_dev(f);
    if (d->g.rows != n->rows || d->g.cols != n->cols || d->g.planes != n->C || d->g.action_dim != n->A)
        return az_fail(AZ_EINVAL, "az_net_eval_leaves: net and forest disagree on the game");
    if (int rc = az_forest_check_eval_tag(f, n->weights_version)) return rc;
    for (int off = 0; off < d->max_leaves; off += n->chunk) {
        AzBatch b{d->n_leaves, 0, off, std::min(n->chunk, d->max_leaves - off)};
        int rc = az_net_forward_chunk(n, b, d->leaf_words, d->leaf_player, d->g.nw, d->priors, d->values, nullptr, (cudaStream_t)stream);
        if (rc) return rc;
    }
    return AZ_OK;
}

__global__ void az_planes_to_leaves_kernel(const float* __restrict__ planes, int n, int C, int cells, int nw,
                                           uint64_t* words, uint8_t* players) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * 2 * nw) return;
    const int b = t / (2 * nw), r = t % (2 * nw), side = r / nw, wi = r % nw;
    const float* p = planes + ((size_t)b * C + side) * cells;
    uint64_t v = 0;
    for (int i = 0; i < 64; ++i) {
        const int cell = wi * 64 + i;
        if (cell < cells && p[cell] == 1.0f) v |= 1ull << i;
    }
    words[t] = v;
    if (r == 0) players[b] = (C == 3 && planes[((size_t)b * C + 2) * cells] == 1.0f) ? 1 : 0;
}

// Net.forward on host planes (ai/nn.py:74-84): float32 [B][C][rows][cols] canonical boards in,
// logits [B][A], values [B] and softmax priors [B][A] out (host).  Synchronous.
extern "C" int az_net_forward_planes(az_net* n, const float* planes_host, int B, float* logits_host, float* values_host,
                                     float* priors_host, void* stream) {
    if (!n || !planes_host || B <= 0) return az_fail(AZ_EINVAL, "az_net_forward_planes: bad arguments");
    if (!n->finalized) return az_fail(AZ_ESTATE, "az_net_forward_planes: call az_net_finalize first");
    cudaStream_t s = (cudaStream_t)stream;
    const int cells = n->rows * n->cols, nw = cells <= 64 ? 1 : AZ_MAX_WORDS;
    // staging for the float planes: owned by the net (allocated at the first call, freed with it) -- a cudaMalloc /
    // cudaFree pair per call synchronises the device and cost more than the forward pass of a small batch
    if (!n->in_planes) {
        int rc = az_net_alloc(n, &n->in_planes, (size_t)n->chunk * n->C * cells);
        if (rc) return rc;
    }
    float* pl_dev = n->in_planes;
    for (int off = 0; off < B; off += n->chunk) {
        const int nb = std::min(n->chunk, B - off);
        AZ_CUDA_OK(cudaMemcpyAsync(pl_dev, planes_host + (size_t)off * n->C * cells, (size_t)nb * n->C * cells * 4, cudaMemcpyHostToDevice, s));
        az_planes_to_leaves_kernel<<<(nb * 2 * nw + 127) / 128, 128, 0, s>>>(pl_dev, nb, n->C, cells, nw, n->in_words, n->in_player);
        AzBatch b{nullptr, nb, 0, nb};
        int rc = az_net_forward_chunk(n, b, n->in_words, n->in_player, nw, n->out_priors, n->out_values, n->out_logits, s);
        if (rc) return rc;
        if (logits_host) AZ_CUDA_OK(cudaMemcpyAsync(logits_host + (size_t)off * n->A, n->out_logits, (size_t)nb * n->A * 4, cudaMemcpyDeviceToHost, s));
        if (priors_host) AZ_CUDA_OK(cudaMemcpyAsync(priors_host + (size_t)off * n->A, n->out_priors, (size_t)nb * n->A * 4, cudaMemcpyDeviceToHost, s));
        if (values_host) AZ_CUDA_OK(cudaMemcpyAsync(values_host + off, n->out_values, (size_t)nb * 4, cudaMemcpyDeviceToHost, s));
        AZ_CUDA_OK(cudaStreamSynchronize(s));
    }
    return AZ_OK;
}

// Output width of the all-token projection of the last block (the dominant GEMM): D + 64 when the
// cls query is folded (depth 1), else 2D (K and V).
extern "C" int az_net_proj_width(az_net* n) { return n ? (n->folded ? n->n2 : 2 * n->D) : 0; }

extern "C" int az_net_path_flags(az_net* n) {
    if (!n) return 0;
    return (n->front ? 1 : 0) | (n->front2 ? 2 : 0) | (n->blk[n->L - 1].bn_w1_c >= AZ_PAIR ? 4 : 0) | (n->tail_fold ? 8 : 0) | (n->head_fold ? 16 : 0) |
           (n->attn_mma ? 32 : 0) | (n->fused_tail ? 64 : 0) | (getenv("AZ_NO_TMA_STORE") ? 0 : 128) | (n->chain ? 256 : 0);
}

extern "C" long long az_net_launch_count(az_net* n) { return n ? (long long)n->launches : 0; }

// ======================================================================================
// Whole search with the built-in net: the loop of ai/mcts.py:16 for every tree at once.
// One step = az_forest_advance (apply answers + select) followed by the net on the new
// leaves; the step is captured once into a CUDA graph and replayed (no per-node host work).
// ======================================================================================
struct AzStepGraph {
    unsigned long long kernels = 0;     // kernels in one captured graph
    int steps = 1;                      // search steps per graph
    unsigned long long forest_uid = 0;  // az_forest_uid of the forest whose buffers the graph bakes in (never reused, unlike the pointer)
    int iters = -1, use_noise = -1, cache_on = -1;
    unsigned long long last_use = 0;
    cudaGraphExec_t exec = nullptr;
};
// A captured step bakes in AzForestDev by value (device pointers, n_trees, cap, vl), so a graph belongs to ONE forest:
// graphs are keyed by the forest's unique id (a destroyed forest's address is routinely handed out again by `new`),
// and az_forest_destroy drops every graph of that forest (az_net_forget_forest).  A net keeps up to AZ_MAX_GRAPHS graphs
// (e.g. two arena engines or a warm-up forest next to the main one) and evicts the least recently used.
#define AZ_MAX_GRAPHS 4
static std::map<az_net*, std::vector<AzStepGraph>> g_graphs;
static unsigned long long g_graph_clock = 0;
static cudaStream_t g_own_stream = nullptr;

static int az_step_eager(az_net* n, az_forest* f, cudaStream_t s) {
    int rc = az_forest_advance(f, s);
    if (rc) return rc;
    n->launches += az_forest_dev(f)->cache.cap ? 2 : 1;      // the advance kernel (+ the cache publish kernel in front of it)
    return az_net_eval_leaves(n, f, s);
}

void az_net_forget_forest(unsigned long long forest_uid) {
    for (auto& kv : g_graphs) {
        auto& v = kv.second;
        for (size_t i = 0; i < v.size();) {
            if (v[i].forest_uid == forest_uid) {
                if (v[i].exec) cudaGraphExecDestroy(v[i].exec);
                v.erase(v.begin() + (long)i);
            } else {
                ++i;
            }
        }
    }
}

extern "C" int az_forest_search_net(az_forest* f, az_net* n, int iters, void* stream) {
    if (!f || !n) return az_fail(AZ_EINVAL, "az_forest_search_net: null");
    if (!n->finalized) return az_fail(AZ_ESTATE, "az_forest_search_net: call az_net_finalize first");
    cudaStream_t s = (cudaStream_t)stream;
    if (!s) {                                               // graphs cannot be captured on the legacy stream
        if (!g_own_stream) AZ_CUDA_OK(cudaStreamCreateWithFlags(&g_own_stream, cudaStreamNonBlocking));
        AZ_CUDA_OK(cudaStreamSynchronize(0));
        s = g_own_stream;
    }
    int rc = az_forest_set_eval_tag(f, n->weights_version, s);   // a weight upload clears the cache (main.py:55-57)
    if (rc) return rc;
    if ((rc = az_forest_begin_move(f, iters, s))) return rc;
    AzForestDev* d = az_forest_dev(f);
    const unsigned long long uid = az_forest_uid(f);
    const int cache_on = d->cache.cap ? 1 : 0;
    std::vector<AzStepGraph>& gv = g_graphs[n];
    AzStepGraph* gp = nullptr;
    for (auto& c : gv)
        if (c.forest_uid == uid && c.iters == iters && c.use_noise == d->use_noise && c.cache_on == cache_on && c.exec) gp = &c;
    int steps = 0;
    if (!gp) {
        if (gv.size() >= AZ_MAX_GRAPHS) {                   // evict the least recently used graph of this net
            size_t lru = 0;
            for (size_t i = 1; i < gv.size(); ++i) if (gv[i].last_use < gv[lru].last_use) lru = i;
            if (gv[lru].exec) cudaGraphExecDestroy(gv[lru].exec);
            gv.erase(gv.begin() + (long)lru);
        }
        AzStepGraph g;
        if ((rc = az_step_eager(n, f, s))) return rc;      // warm-up outside capture (sets func attributes)
        ++steps;
        cudaGraph_t graph = nullptr;
        AZ_CUDA_OK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        const unsigned long long before = n->launches;
        // AZ_GRAPH_STEPS > 1 captures several steps per graph (steps beyond the last simulation find every tree finished and
        // an empty evaluation batch).  Measured on Connect4 4096 x 800: 39.5 / 39.4 / 39.4 / 39.2 M sims/s at 1 / 3 / 9 / 27
        // steps per graph -- the gap between graph launches is not what the step waits for; the default stays 1.
        g.steps = 1;
        if (const char* e = getenv("AZ_GRAPH_STEPS")) g.steps = std::max(1, std::min(64, atoi(e)));
        for (int i = 0; i < g.steps && !rc; ++i) rc = az_step_eager(n, f, s);
        g.kernels = n->launches - before;
        cudaError_t ce = cudaStreamEndCapture(s, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (ce != cudaSuccess) return az_fail(AZ_ECUDA, std::string("graph capture: ") + cudaGetErrorString(ce));
        ce = cudaGraphInstantiate(&g.exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) return az_fail(AZ_ECUDA, std::string("graph instantiate: ") + cudaGetErrorString(ce));
        g.forest_uid = uid; g.iters = iters; g.use_noise = d->use_noise; g.cache_on = cache_on;
        gv.push_back(g);
        gp = &gv.back();
    }
    AzStepGraph& g = *gp;
    g.last_use = ++g_graph_clock;
    int done = 0;
    for (;;) {
        const int want = az_search_burst(d, iters, steps);
        const int burst = std::max(1, (want + g.steps - 1) / g.steps);
        for (int i = 0; i < burst; ++i) {
            AZ_CUDA_OK(cudaGraphLaunch(g.exec, s));
            n->launches += g.kernels;
            steps += g.steps;
        }
        if ((rc = az_forest_counts(f, nullptr, &done, s))) return rc;
        if (done >= d->n_trees) break;
        if (steps > 2 * iters + 16) return az_fail(AZ_ESTATE, "az_forest_search_net: did not converge");
    }
    return AZ_OK;
}

// Continuous self-play with the built-in net (az_selfplay.cu: az_selfplay_service): one CUDA graph holds `K` search steps
// (advance + evaluator) followed by the service kernels; it is replayed until every game is over.  No tree waits for
// another one: a slot whose search is finished moves and starts its next search at the next service point.
extern "C" int az_selfplay_run_async_net(az_selfplay* sp, az_net* n, int iters, long long max_steps_arg, int resume, void* stream) {
    if (!sp || !n || iters < 1) return az_fail(AZ_EINVAL, "az_selfplay_run_async_net: bad arguments");
    if (!n->finalized) return az_fail(AZ_ESTATE, "az_selfplay_run_async_net: call az_net_finalize first");
    az_forest* f = az_selfplay_forest(sp);
    AzForestDev* d = az_forest_dev(f);
    if (iters > d->iters && iters > 0) { /* begin_move below checks against the forest's max_iters */ }
    cudaStream_t s = (cudaStream_t)stream;
    if (!s) {
        if (!g_own_stream) AZ_CUDA_OK(cudaStreamCreateWithFlags(&g_own_stream, cudaStreamNonBlocking));
        AZ_CUDA_OK(cudaStreamSynchronize(0));
        s = g_own_stream;
    }
    int rc;
    if (!resume) {                                                       // resume: the slots continue the searches a budgeted call left
        if ((rc = az_selfplay_prepare(sp, s))) return rc;                // root noise of the first plies (and use_noise before capture)
        if ((rc = az_forest_set_eval_tag(f, n->weights_version, s))) return rc;
        if ((rc = az_forest_begin_move(f, iters, s))) return rc;
    }
    if ((rc = az_step_eager(n, f, s))) return rc;                        // warm-up outside capture (function attributes)
    if ((rc = az_selfplay_service(sp, s))) return rc;
    int K = 4;                                                           // search steps per service point
    if (const char* e = getenv("AZ_ASYNC_STEPS")) K = std::max(1, std::min(64, atoi(e)));
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    AZ_CUDA_OK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    const unsigned long long before = n->launches;
    for (int i = 0; i < K && !rc; ++i) rc = az_step_eager(n, f, s);
    if (!rc) rc = az_selfplay_service(sp, s);
    const unsigned long long per_graph = n->launches - before;
    cudaError_t ce = cudaStreamEndCapture(s, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (ce != cudaSuccess) return az_fail(AZ_ECUDA, std::string("async graph capture: ") + cudaGetErrorString(ce));
    ce = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess) return az_fail(AZ_ECUDA, std::string("async graph instantiate: ") + cudaGetErrorString(ce));
    long long steps = 1 + K;
    int live = 1;
    const long long max_steps = (long long)d->g.state_dim * (iters + 2) * 64 + 1024;   // per slot generation; the loop also ends on live == 0
    long long since_tick = steps, guard = 0;
    while (live > 0) {
        int burst = std::max(1, 64 / K);                                 // ~64 steps between two looks at the counters
        if (max_steps_arg > 0) burst = (int)std::max<long long>(1, std::min<long long>(burst, (max_steps_arg - steps + K - 1) / K));
        for (int i = 0; i < burst; ++i) {
            ce = cudaGraphLaunch(exec, s);
            if (ce != cudaSuccess) { cudaGraphExecDestroy(exec); return az_fail(AZ_ECUDA, std::string("async graph launch: ") + cudaGetErrorString(ce)); }
            n->launches += per_graph;
            steps += K;
            since_tick += K;
            if (since_tick >= iters) { az_forest_cache_tick(f, s); since_tick = 0; }
        }
        if (max_steps_arg > 0 && steps >= max_steps_arg) break;          // budgeted run (benchmarks): the caller resumes or reads what is there
        if ((rc = az_selfplay_status(sp, nullptr, nullptr, &live, nullptr, s))) { cudaGraphExecDestroy(exec); return rc; }
        if (++guard > max_steps) { cudaGraphExecDestroy(exec); return az_fail(AZ_ESTATE, "az_selfplay_run_async_net: did not converge"); }
    }
    cudaGraphExecDestroy(exec);
    return (int)std::min<long long>(steps, 0x7fffffff);
}

extern "C" int az_net_invalidate_graphs(az_net* n) {
    auto it = g_graphs.find(n);
    if (it != g_graphs.end()) {
        for (auto& g : it->second)
            if (g.exec) cudaGraphExecDestroy(g.exec);
        g_graphs.erase(it);
    }
    return AZ_OK;
}

// ---- per-kernel timing of eager forward passes (CUDA events on the launching stream) -------
extern "C" int az_net_profile(az_net* n, int enable) {
    if (!n) return az_fail(AZ_EINVAL, "az_net_profile: null");
    if (enable == 2 || enable == 3) {                          // resume / pause, totals kept
        n->prof_on = enable == 2;
        return AZ_OK;
    }
    n->prof_on = enable != 0;
    if (enable) {
        for (int k = 0; k < 16; ++k) { n->prof_ms[k] = 0; n->prof_n[k] = 0; }
    }
    return AZ_OK;
}

// Synchronises, folds pending event pairs into totals.  names_out: AZ_K_COUNT C strings.
extern "C" int az_net_profile_read(az_net* n, int max_kinds, const char** names_out, double* ms_out, long long* count_out) {
    if (!n) return az_fail(AZ_EINVAL, "az_net_profile_read: null");
    AZ_CUDA_OK(cudaDeviceSynchronize());
    for (auto& r : n->prof_pending) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { n->prof_ms[r.kind] += ms; n->prof_n[r.kind] += 1; }
        cudaEventDestroy(r.a); cudaEventDestroy(r.b);
    }
    n->prof_pending.clear();
    const int k = max_kinds < AZ_K_COUNT ? max_kinds : AZ_K_COUNT;
    for (int i = 0; i < k; ++i) {
        if (names_out) names_out[i] = AZ_K_NAMES[i];
        if (ms_out) ms_out[i] = n->prof_ms[i];
        if (count_out) count_out[i] = n->prof_n[i];
    }
    return k;
}

#ifdef AZ_FRONT_TRACE
extern "C" __attribute__((visibility("default"))) int az_debug_front_trace(unsigned long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, g_front_trace, sizeof(g_front_trace));
}
#endif
