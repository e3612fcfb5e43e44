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
#include <string>
#include <vector>
#include "az_gemm.cuh"
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

static int az_pick_bn(int N) { return N % 256 == 0 ? 256 : (N % 128 == 0 ? 128 : (N % 64 == 0 ? 64 : 0)); }

template <int BN, int EPI>
static int az_launch_gemm_t(const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& a, int max_rows, cudaStream_t s) {
    static bool attr_done = false;
    if (!attr_done) {
        AZ_CUDA_OK(cudaFuncSetAttribute(az_gemm_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, AzGemmCfg<BN>::SMEM));
        attr_done = true;
    }
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int tiles = ((max_rows + 127) / 128) * (a.N / BN);
    const int grid = tiles < g_num_sms ? (tiles > 0 ? tiles : 1) : g_num_sms;
    az_gemm_kernel<BN, EPI><<<grid, 192, AzGemmCfg<BN>::SMEM, s>>>(ta, tb, a);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

// A: bf16 [rows][K] (ld = lda), W: bf16 [N][K]; max_rows bounds the grid when M is device-side.
static int az_launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const AzGemmArgs& a, int epi, int max_rows,
                          cudaStream_t s) {
    const int bn = az_pick_bn(a.N);
    if (!bn || a.K % 64) return az_fail(AZ_EINVAL, "gemm: N and K must be multiples of 64");
#define AZ_G(BN_)                                                                              \
    if (bn == BN_) {                                                                           \
        if (epi == AZ_EPI_BIAS) return az_launch_gemm_t<BN_, AZ_EPI_BIAS>(ta, tb, a, max_rows, s);   \
        if (epi == AZ_EPI_GELU) return az_launch_gemm_t<BN_, AZ_EPI_GELU>(ta, tb, a, max_rows, s);   \
        return az_launch_gemm_t<BN_, AZ_EPI_RESID>(ta, tb, a, max_rows, s);                    \
    }
    AZ_G(256) AZ_G(128) AZ_G(64)
#undef AZ_G
    return az_fail(AZ_EINVAL, "gemm: unsupported tile");
}

// Stand-alone GEMM entry point (parity tests of the tensor-core kernel against torch).
extern "C" int az_gemm_bf16(const void* a_dev, int lda, const void* w_dev, const float* bias_dev, const void* resid_dev,
                            int ldr, void* out_dev, int ldo, int M, int N, int K, int epi, void* stream) {
    if (!a_dev || !w_dev || !bias_dev || !out_dev || M <= 0) return az_fail(AZ_EINVAL, "az_gemm_bf16: bad arguments");
    if (epi == AZ_EPI_RESID && !resid_dev) return az_fail(AZ_EINVAL, "az_gemm_bf16: residual missing");
    const int bn = az_pick_bn(N);
    if (!bn || K % 64 || lda % 8 || ldo % 8) return az_fail(AZ_EINVAL, "az_gemm_bf16: N,K multiples of 64; ld multiples of 8");
    CUtensorMap ta, tb;
    int rc;
    if ((rc = az_make_tmap(&ta, a_dev, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 128))) return rc;
    if ((rc = az_make_tmap(&tb, w_dev, (uint64_t)N, (uint64_t)K, (uint64_t)K, (uint32_t)bn))) return rc;
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
    return az_launch_gemm(ta, tb, a, epi, M, (cudaStream_t)stream);
}
