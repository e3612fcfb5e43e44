// Persistent, warp-specialised bf16 GEMM for sm_100a:  out[M,N] = epi(A[M,K] * W[N,K]^T + bias).
//
// This is the dense contraction of the policy/value net (reference: the nn.Linear /
// nn.MultiheadAttention projections of ai/nn.py:42-50,71-72) on the 5th-gen tensor cores:
//   warp 0   TMA producer: cp.async.bulk.tensor 128B-swizzled K-major tiles of A and W into a
//            STAGES-deep shared-memory ring, completion on mbarriers;
//   warp 1   MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) with the fp32
//            accumulator in TMEM; tcgen05.commit releases ring slots / publishes the tile;
//   warps 2-5 epilogue: tcgen05.ld the accumulator (each warp its 32-lane quarter), add
//            bias (+ exact-erf GELU | + residual), convert to bf16, 64-byte row stores.
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
// M may come from a device-side counter (the evaluation batch of the forest is only known
// on the device), so the kernel is CUDA-graph capturable with a fixed grid.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "az_ptx.cuh"

#define AZ_EPI_BIAS 0
#define AZ_EPI_GELU 1
#define AZ_EPI_RESID 2

struct AzGemmArgs {
    const float* bias;          // [N] fp32
    const __nv_bfloat16* resid; // [M][ldr] (AZ_EPI_RESID)
    __nv_bfloat16* out;         // [M][ldo]
    int ldo, ldr;
    int N, K;
    int m_fixed;                // rows when m_count == nullptr
    const int* m_count;         // device counter of items (leaves) ...
    int m_item_offset;          // ... this launch handles items [offset, offset + cap)
    int m_item_cap;
    int m_rows_per_item;        // rows per item (tokens per board, or 1)
    float* out_f32;             // optional: columns >= f32_col0 are written here as fp32 [M][64] instead
    int f32_col0;
};

__device__ __forceinline__ int az_gemm_rows(const AzGemmArgs& a) {
    if (!a.m_count) return a.m_fixed;
    int items = *a.m_count - a.m_item_offset;
    items = items < 0 ? 0 : (items > a.m_item_cap ? a.m_item_cap : items);
    return items * a.m_rows_per_item;
}

template <int BN>
struct AzGemmCfg {
    static constexpr int BM = 128, BK = 64;
    static constexpr int A_BYTES = BM * BK * 2;              // 16 KB
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGES = BN >= 192 ? 4 : 6;
    static constexpr int TMEM_COLS = 2 * BN <= 32 ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
    static constexpr int ACC_STRIDE = TMEM_COLS / 2;          // TMEM columns between the two accumulator stages
    static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + 256 + 1024;   // + barriers + align slack
};

__device__ __forceinline__ float az_gelu_erf(float x) {      // nn.GELU() default: exact erf form
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

template <int BN, int EPI>
__global__ void __launch_bounds__(192, 1)
az_gemm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, AzGemmArgs args) {
    using Cfg = AzGemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    unsigned char* s_a = smem;
    unsigned char* s_b = smem + STAGES * Cfg::A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES));
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M = az_gemm_rows(args);
    const int m_tiles = (M + Cfg::BM - 1) / Cfg::BM;
    const int n_tiles = args.N / BN;
    const int total = m_tiles * n_tiles;
    const int kblocks = args.K / Cfg::BK;

    if (warp == 0 && lane == 0) {
        azptx::prefetch_tmap(&tm_a);
        azptx::prefetch_tmap(&tm_b);
        for (int i = 0; i < STAGES; ++i) {
            azptx::mbar_init(&full[i], 1);
            azptx::mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&tfull[i], 1);
            azptx::mbar_init(&tempty[i], 128);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 1) azptx::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {                                      // ---- TMA producer
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
                const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
                for (int kb = 0; kb < kblocks; ++kb) {
                    azptx::mbar_wait(&empty[stage], phase ^ 1);
                    azptx::mbar_arrive_expect_tx(&full[stage], Cfg::A_BYTES + Cfg::B_BYTES);
                    azptx::tma_load_2d(s_a + stage * Cfg::A_BYTES, &tm_a, &full[stage], kb * Cfg::BK, m_blk * Cfg::BM);
                    azptx::tma_load_2d(s_b + stage * Cfg::B_BYTES, &tm_b, &full[stage], kb * Cfg::BK, n_blk * BN);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {                                   // ---- MMA issuer
        constexpr uint32_t idesc = azptx::umma_idesc_bf16(Cfg::BM, BN);
        int stage = 0, as = 0;
        uint32_t phase = 0, aphase = 0;
        for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
            azptx::mbar_wait(&tempty[as], aphase ^ 1);        // epilogue drained this accumulator
            azptx::tc_fence_after();
            for (int kb = 0; kb < kblocks; ++kb) {
                azptx::mbar_wait(&full[stage], phase);
                azptx::tc_fence_after();
                if (lane == 0) {
                    const uint64_t da = azptx::umma_desc_k128(azptx::smem_u32(s_a + stage * Cfg::A_BYTES));
                    const uint64_t db = azptx::umma_desc_k128(azptx::smem_u32(s_b + stage * Cfg::B_BYTES));
#pragma unroll
                    for (int k = 0; k < Cfg::BK / 16; ++k)    // +32 bytes (>>4 = 2) per K=16 step
                        azptx::umma_bf16(tmem_base + as * Cfg::ACC_STRIDE, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    azptx::umma_commit(&empty[stage]);
                    if (kb == kblocks - 1) azptx::umma_commit(&tfull[as]);
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
    } else {                                                  // ---- epilogue (warps 2..5)
        const int q = warp & 3;                               // TMEM lane quarter this warp may read
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
            const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
            azptx::mbar_wait(&tfull[as], aphase);
            azptx::tc_fence_after();
            const int row = m_blk * Cfg::BM + q * 32 + lane;
            const bool live = row < M;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t r[32];
                azptx::tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * Cfg::ACC_STRIDE + c * 32, r);
                azptx::tmem_ld_wait();
                const int col0 = n_blk * BN + c * 32;
                if (live && args.out_f32 && col0 >= args.f32_col0) {      // fp32 side output (attention scores)
                    const float4* b4 = reinterpret_cast<const float4*>(args.bias + col0);
                    float4* op = reinterpret_cast<float4*>(args.out_f32 + (size_t)row * 64 + (col0 - args.f32_col0));
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 bb = __ldg(b4 + i);
                        op[i] = make_float4(__uint_as_float(r[4 * i + 0]) + bb.x, __uint_as_float(r[4 * i + 1]) + bb.y,
                                            __uint_as_float(r[4 * i + 2]) + bb.z, __uint_as_float(r[4 * i + 3]) + bb.w);
                    }
                } else if (live) {
                    const float4* b4 = reinterpret_cast<const float4*>(args.bias + col0);
                    uint4 rv[4];
                    if (EPI == AZ_EPI_RESID) {
                        const uint4* rp = reinterpret_cast<const uint4*>(args.resid + (size_t)row * args.ldr + col0);
#pragma unroll
                        for (int i = 0; i < 4; ++i) rv[i] = rp[i];
                    }
                    uint4 ov[4];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 bb = __ldg(b4 + i);
                        float v0 = __uint_as_float(r[4 * i + 0]) + bb.x;
                        float v1 = __uint_as_float(r[4 * i + 1]) + bb.y;
                        float v2 = __uint_as_float(r[4 * i + 2]) + bb.z;
                        float v3 = __uint_as_float(r[4 * i + 3]) + bb.w;
                        if (EPI == AZ_EPI_GELU) {
                            v0 = az_gelu_erf(v0); v1 = az_gelu_erf(v1); v2 = az_gelu_erf(v2); v3 = az_gelu_erf(v3);
                        }
                        if (EPI == AZ_EPI_RESID) {
                            const uint32_t* rw = reinterpret_cast<const uint32_t*>(&rv[i >> 1]) + (i & 1) * 2;
                            const __nv_bfloat162 r01 = *reinterpret_cast<const __nv_bfloat162*>(&rw[0]);
                            const __nv_bfloat162 r23 = *reinterpret_cast<const __nv_bfloat162*>(&rw[1]);
                            v0 += __bfloat162float(r01.x); v1 += __bfloat162float(r01.y);
                            v2 += __bfloat162float(r23.x); v3 += __bfloat162float(r23.y);
                        }
                        __nv_bfloat162 o01 = __floats2bfloat162_rn(v0, v1);
                        __nv_bfloat162 o23 = __floats2bfloat162_rn(v2, v3);
                        uint32_t* ow = reinterpret_cast<uint32_t*>(&ov[i >> 1]) + (i & 1) * 2;
                        ow[0] = *reinterpret_cast<uint32_t*>(&o01);
                        ow[1] = *reinterpret_cast<uint32_t*>(&o23);
                    }
                    uint4* op = reinterpret_cast<uint4*>(args.out + (size_t)row * args.ldo + col0);
#pragma unroll
                    for (int i = 0; i < 4; ++i) op[i] = ov[i];
                }
            }
            azptx::tc_fence_before();
            azptx::mbar_arrive(&tempty[as]);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) azptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}
