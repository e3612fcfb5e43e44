// The cls-row tail of small evaluation batches as ONE persistent kernel: out-proj -> mlp1 -> mlp2 -> heads, the same four
// GEMMs as the chain of az_gemm_kernel launches (same tiles, same epilogues, same operand order: bit-identical results), run
// back to back by one resident grid with a grid-wide barrier between layers instead of four launches.
//
// Why: at the ~1-2 k rows an evaluation pass of Connect4 has left after the evaluation cache, a GEMM launch is 2-4 us of MMA
// work inside ~12 us of launch + prologue (barrier init, TMEM allocation, tensor-map fetch) + first-load latency + drain +
// dependency resolution (9.5 of 18 us); four of them are a quarter of the pass.  Here the prologue is paid
// once, the operand ring and the two TMEM accumulators run on across layers, and a layer boundary costs one atomic + spin.
// (A different fusion -- az_tail.cuh: feature-split clusters, activations kept on chip -- was built first and is slower
// than the chain; this one keeps the tiling that is already at speed and removes only the launches.)
// STATUS: bit-identical to the four launches (tests/test_gpu_net.py::test_chain_kernel_is_bit_identical_to_separate_launches)
// and OPT-IN (AZ_CHAIN=1).  Measured on B200 at 1792 boards: 65.6 us against 70 us for the four launches timed one by one, but
// Connect4 self-play runs at 62.9 M simulations/s with it against 66.8 M without: a layer still takes ~16 us with its launch
// and prologue gone (first-load latency, a 16-32 step K loop on one tile per CTA, an epilogue whose bias / residual loads
// start only when the accumulator is ready), and a grid-wide barrier cannot overlap a layer's tail with its successor's
// head the way programmatic dependent launch does.
//
// Layer l: out[M, N_l] = epi_l(A_l[M, K_l] . W_l[N_l, K_l]^T), tiles of 128 rows x bn_l columns, roles as in az_gemm.cuh
// (warp 0 TMA producer, warp 1 MMA issue, warps 2-9 epilogue).  A_{l+1} is what layer l wrote to global memory (TMA stores),
// so between layers: every epilogue warp waits for its bulk stores, all threads __threadfence, one thread per CTA arrives on
// a global counter and spins until all CTAs have (the grid is at most one CTA per SM and every CTA is resident: 226 KB of
// shared memory each, launched after the predecessor has drained), then fence.proxy.async before the next layer's TMA loads.
// The last CTA to leave resets the counters for the next launch (the launches of one stream are ordered).
#pragma once
#include "az_gemm.cuh"

#define AZ_CHAIN_LAYERS 4
#define AZ_CHAIN_STAGES 4
#define AZ_CHAIN_B_SLOT 32768            // W tile slot: up to 256 rows x 64 k
#define AZ_CHAIN_SMEM (AZ_CHAIN_STAGES * (16384 + AZ_CHAIN_B_SLOT) + AZ_BAR_BYTES + AZ_OUT_STAGE_BYTES + 1024)

struct AzChainMaps {
    CUtensorMap a[AZ_CHAIN_LAYERS], b[AZ_CHAIN_LAYERS], o[AZ_CHAIN_LAYERS];
};
struct AzChainArgs {
    AzGemmArgs g[AZ_CHAIN_LAYERS];       // layer 0 bias, 1 GELU (+ LayerNorm fold), 2 residual, 3 heads
    int bn[AZ_CHAIN_LAYERS];             // 128 | 256 for layers 0-2, 64 for the heads
    unsigned int* sync;                  // [2]: layer-barrier arrivals, CTAs that have left
};

__device__ __forceinline__ unsigned int az_ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int BN, int EPI>
__device__ __forceinline__ void az_chain_epilogue(const AzGemmArgs& args, const CUtensorMap* tm_o, unsigned char* stg_base, uint32_t tacc, int m_blk,
                                                  int n_blk, int q, int half, int warp, int lane, int M) {
    const int row = m_blk * 128 + q * 32 + lane;
    const bool live = row < M;
    float mean = 0.f, rstd = 1.f;
    if (args.stats_in && live) {                              // LayerNorm statistics of this A row (written by an earlier layer)
        const float2* sp = reinterpret_cast<const float2*>(args.stats_in + (size_t)row * AZ_STATS_LD);
        float s1 = 0.f, s2 = 0.f;
        for (int p = 0; p < args.stats_parts; ++p) {
            const float2 v = __ldcg(sp + p);
            s1 += v.x;
            s2 += v.y;
        }
        mean = s1 * args.ln_inv_k;
        rstd = rsqrtf(fmaxf(s2 * args.ln_inv_k - mean * mean, 0.f) + args.ln_eps);
    }
    const float nmr = -mean * rstd;
    if constexpr (EPI == AZ_EPI_HEAD) {
        if (warp < 6) az_gemm_epilogue_head(args, tacc, reinterpret_cast<float*>(stg_base) + q * 32 * 64, m_blk * 128 + q * 32, lane, live, M, rstd, nmr);
    } else {
        unsigned char* stg = args.stage_out ? stg_base + (warp - 2) * 4096 : nullptr;
        az_gemm_epilogue_row<BN, EPI>(args, tacc, row, live, n_blk, half, rstd, nmr, stg, tm_o, m_blk * 128 + q * 32);
    }
}

__global__ void __launch_bounds__(AZ_GEMM_THREADS, 1) az_chain_kernel(const __grid_constant__ AzChainMaps maps, const __grid_constant__ AzChainArgs ca) {
    constexpr int STAGES = AZ_CHAIN_STAGES, A_BYTES = 16384, ACC_STRIDE = 256;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    unsigned char* s_a = smem;
    unsigned char* s_b = smem + STAGES * A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + AZ_CHAIN_B_SLOT));
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    unsigned char* stg_base = smem + STAGES * (A_BYTES + AZ_CHAIN_B_SLOT) + AZ_BAR_BYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        for (int l = 0; l < AZ_CHAIN_LAYERS; ++l) {
            azptx::prefetch_tmap(&maps.a[l]);
            azptx::prefetch_tmap(&maps.b[l]);
        }
        for (int i = 0; i < STAGES; ++i) {
            azptx::mbar_init(&full[i], 1);
            azptx::mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&tfull[i], 1);
            azptx::mbar_init(&tempty[i], 256);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 1) azptx::tmem_alloc<512>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    azptx::griddep_wait();
    azptx::griddep_launch();
    const int M = az_gemm_rows(ca.g[0]);
    const int m_tiles = (M + 127) / 128;

    // ring / accumulator positions run on across the layers (each role keeps its own copy)
    int stage = 0, as = 0;
    uint32_t phase = 0, aphase = 0;
    for (int l = 0; l < AZ_CHAIN_LAYERS; ++l) {
        const AzGemmArgs& args = ca.g[l];
        const int bn = ca.bn[l];
        const int n_tiles = args.N / bn, kblocks = args.K >> 6;
        const int total = m_tiles * n_tiles;
        if (warp == 0) {
            if (lane == 0) {                                  // ---- TMA producer
                const uint64_t pol_a = azptx::l2_policy_evict_first(), pol_w = azptx::l2_policy_evict_last();
                const uint32_t bytes = (uint32_t)(A_BYTES + bn * 128);
                for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
                    const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
                    for (int kb = 0; kb < kblocks; ++kb) {
                        azptx::mbar_wait(&empty[stage], phase ^ 1);
                        azptx::mbar_arrive_expect_tx(&full[stage], bytes);
                        azptx::tma_load_2d_hint(s_a + stage * A_BYTES, &maps.a[l], &full[stage], kb * 64, m_blk * 128, pol_a);
                        azptx::tma_load_2d_hint(s_b + stage * AZ_CHAIN_B_SLOT, &maps.b[l], &full[stage], kb * 64, n_blk * bn, pol_w);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
        } else if (warp == 1) {                               // ---- MMA issuer
            const uint32_t idesc = azptx::umma_idesc_bf16(128, (uint32_t)bn);
            for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
                azptx::mbar_wait(&tempty[as], aphase ^ 1);
                azptx::tc_fence_after();
                for (int kb = 0; kb < kblocks; ++kb) {
                    azptx::mbar_wait(&full[stage], phase);
                    azptx::tc_fence_after();
                    if (lane == 0) {
                        const uint64_t da = azptx::umma_desc_k128(azptx::smem_u32(s_a + stage * A_BYTES));
                        const uint64_t db = azptx::umma_desc_k128(azptx::smem_u32(s_b + stage * AZ_CHAIN_B_SLOT));
#pragma unroll
                        for (int k = 0; k < 4; ++k) azptx::umma_bf16(tmem_base + as * ACC_STRIDE, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                        azptx::umma_commit(&empty[stage]);
                        if (kb == kblocks - 1) azptx::umma_commit(&tfull[as]);
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                as ^= 1;
                if (as == 0) aphase ^= 1;
            }
        } else {                                              // ---- epilogue (warps 2..9)
            const int q = warp & 3, half = (warp - 2) >> 2;
            for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
                const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
                azptx::mbar_wait(&tfull[as], aphase);
                azptx::tc_fence_after();
                const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + as * ACC_STRIDE;
                const CUtensorMap* tm_o = &maps.o[l];
                if (l == 0) {
                    if (bn == 256) az_chain_epilogue<256, AZ_EPI_BIAS>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                    else az_chain_epilogue<128, AZ_EPI_BIAS>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                } else if (l == 1) {
                    if (bn == 256) az_chain_epilogue<256, AZ_EPI_GELU>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                    else az_chain_epilogue<128, AZ_EPI_GELU>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                } else if (l == 2) {
                    if (bn == 256) az_chain_epilogue<256, AZ_EPI_RESID>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                    else az_chain_epilogue<128, AZ_EPI_RESID>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                } else {
                    az_chain_epilogue<64, AZ_EPI_HEAD>(args, tm_o, stg_base, tacc, m_blk, n_blk, q, half, warp, lane, M);
                }
                azptx::tc_fence_before();
                azptx::mbar_arrive(&tempty[as]);
                as ^= 1;
                if (as == 0) aphase ^= 1;
            }
            if (lane == 0) azptx::bulk_wait0();               // this warp's TMA stores of the layer are complete
            __threadfence();
        }
        if (l + 1 < AZ_CHAIN_LAYERS) {                        // ---- grid-wide layer boundary
            __syncthreads();
            if (threadIdx.x == 0) {
                __threadfence();
                atomicAdd(ca.sync, 1u);
                const unsigned int target = (unsigned int)(l + 1) * gridDim.x;
                unsigned int spins = 0;
                while (az_ld_acquire_gpu(ca.sync) < target) {
                    if (++spins > 400000000u) {
                        printf("az_chain: layer barrier timeout (block %d layer %d)\n", blockIdx.x, l);
                        __trap();
                    }
                }
                __threadfence();
            }
            __syncthreads();
            azptx::fence_proxy_async_global();                // the previous layer's global writes, then this thread's TMA loads
            // every role carries the ring / accumulator position of ITS OWN loop: the counts are equal by construction
        }
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int left = atomicAdd(ca.sync + 1, 1u);
        if (left == gridDim.x - 1) {                           // last CTA out: every CTA is past every barrier
            ca.sync[0] = 0u;
            ca.sync[1] = 0u;
            __threadfence();
        }
    }
    if (warp == 1) azptx::tmem_dealloc<512>(tmem_base);
}
