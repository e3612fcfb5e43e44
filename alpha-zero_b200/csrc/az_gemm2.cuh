// CTA-pair variant of az_gemm_kernel: two SMs of one TPC compute one 256 x BN output tile with
// tcgen05.mma.cta_group::2.
//
// Why: the cls-row GEMMs have M = boards (4096): few row tiles, so every CTA streams its whole operand
// panel through shared memory and the kernel is bound by the chip-wide L2 -> SM rate (~6300 B/clk), not by
// the tensor pipe.  A pair shares the W operand: CTA r loads its own 128 A rows and HALF of the W rows
// (BN/2), the MMA of the leader CTA reads both halves and writes 128 accumulator rows into each CTA's
// TMEM.  L2 traffic per output tile drops from (128 + BN) to (128 + BN/2) rows per CTA.
//
//   warp 0  TMA producer (both CTAs; every load signals the LEADER's full barrier)
//   warp 1  MMA issuer (leader CTA only); commits are multicast to both CTAs' barriers
//   warps 2-9 epilogue (both CTAs, each on its own 128 accumulator rows); the peer's epilogue releases
//           the accumulator stage on the leader's barrier through the cluster shared window
// Launched with cluster dimension 2 (+ programmatic stream serialization).
#pragma once
#include "az_gemm.cuh"

namespace azptx {
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory object in CTA `rank` of the cluster (shared::cluster window)
__device__ __forceinline__ uint32_t mapa(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// tile load into THIS CTA's shared memory, completion bytes on a barrier that may live in the peer CTA
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const void* tmap, uint32_t bar_cluster_addr, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst) {   // whole warp, same warp in both CTAs
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols));
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at the same offset in both CTAs once all MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3)
                 : "memory");
}
}  // namespace azptx

template <int BN>
struct AzGemm2Cfg {
    static constexpr int BM = 128, BK = 64;                   // per CTA; the pair tile is 256 x BN
    static constexpr int A_BYTES = BM * BK * 2;               // 16 KB
    static constexpr int B_BYTES = (BN / 2) * BK * 2;         // this CTA's half of the W tile
    static constexpr int STAGES = BN >= 256 ? 6 : 8;
    static constexpr int TMEM_COLS = 2 * BN <= 256 ? 256 : 512;
    static constexpr int ACC_STRIDE = TMEM_COLS / 2;
    static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + AZ_BAR_BYTES + AZ_OUT_STAGE_BYTES + 1024;
};

template <int BN, int EPI>
__global__ void __launch_bounds__(AZ_GEMM_THREADS, 1)
az_gemm2_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, const __grid_constant__ CUtensorMap tm_o,
                AzGemmArgs args) {
    using Cfg = AzGemm2Cfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ unsigned char smem_dyn[];
    // identical offsets in both CTAs: the MMA applies the leader's descriptors to the peer's shared memory
    unsigned char* smem = smem_dyn + ((1024u - (azptx::smem_u32(smem_dyn) & 1023u)) & 1023u);
    unsigned char* s_a = smem;
    unsigned char* s_b = smem + STAGES * Cfg::A_BYTES;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES));
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = azptx::cluster_ctarank();
    const bool leader = rank == 0;
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int n_tiles = args.N / BN;
    const int kblocks = args.K / Cfg::BK;

    if (warp == 0 && lane == 0) {
        azptx::prefetch_tmap(&tm_a);
        azptx::prefetch_tmap(&tm_b);
        for (int i = 0; i < STAGES; ++i) {
            azptx::mbar_init(&full[i], 1);                     // leader's: one expect_tx arrival + the bytes of both CTAs
            azptx::mbar_init(&empty[i], 1);                    // one multicast commit
        }
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&tfull[i], 1);                    // one multicast commit
            azptx::mbar_init(&tempty[i], 16);                  // leader's: one lane of each of the 8 epilogue warps of both CTAs
        }
        azptx::fence_barrier_init();
    }
    if (warp == 1) azptx::tmem_alloc_pair<Cfg::TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::cluster_sync();                                      // the peer's barriers exist before anything signals them
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    azptx::griddep_wait();
    azptx::griddep_launch();
    const int M = az_gemm_rows(args);
    const int m_tiles = (M + 2 * Cfg::BM - 1) / (2 * Cfg::BM);
    const int total = m_tiles * n_tiles;

    if (warp == 0) {
        if (lane == 0) {                                        // ---- TMA producer (both CTAs)
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = pair; tile < total; tile += n_pairs) {
                const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
                for (int kb = 0; kb < kblocks; ++kb) {
                    azptx::mbar_wait(&empty[stage], phase ^ 1);
                    const uint32_t lfull = azptx::mapa(azptx::smem_u32(&full[stage]), 0);
                    if (leader) azptx::mbar_arrive_expect_tx(&full[stage], 2 * (Cfg::A_BYTES + Cfg::B_BYTES));
                    azptx::tma_load_2d_pair(s_a + stage * Cfg::A_BYTES, &tm_a, lfull, kb * Cfg::BK, m_blk * 2 * Cfg::BM + (int)rank * Cfg::BM);
                    azptx::tma_load_2d_pair(s_b + stage * Cfg::B_BYTES, &tm_b, lfull, kb * Cfg::BK, n_blk * BN + (int)rank * (BN / 2));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader) {                                           // ---- MMA issuer (leader CTA)
            constexpr uint32_t idesc = azptx::umma_idesc_bf16(2 * Cfg::BM, BN);
            int stage = 0, as = 0;
            uint32_t phase = 0, aphase = 0;
            for (int tile = pair; tile < total; tile += n_pairs) {
                azptx::mbar_wait(&tempty[as], aphase ^ 1);      // both CTAs' epilogues drained this accumulator
                azptx::tc_fence_after();
                for (int kb = 0; kb < kblocks; ++kb) {
                    azptx::mbar_wait(&full[stage], phase);
                    azptx::tc_fence_after();
                    if (lane == 0) {
                        const uint64_t da = azptx::umma_desc_k128(azptx::smem_u32(s_a + stage * Cfg::A_BYTES));
                        const uint64_t db = azptx::umma_desc_k128(azptx::smem_u32(s_b + stage * Cfg::B_BYTES));
#pragma unroll
                        for (int k = 0; k < Cfg::BK / 16; ++k)
                            azptx::umma_bf16_pair(tmem_base + as * Cfg::ACC_STRIDE, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                        azptx::umma_commit_pair(&empty[stage]);
                        if (kb == kblocks - 1) azptx::umma_commit_pair(&tfull[as]);
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                as ^= 1;
                if (as == 0) aphase ^= 1;
            }
        }
    } else {                                                    // ---- epilogue (warps 2..9, both CTAs)
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        int as = 0;
        uint32_t aphase = 0;
        const uint32_t ltempty0 = azptx::mapa(azptx::smem_u32(&tempty[0]), 0);
        for (int tile = pair; tile < total; tile += n_pairs) {
            const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
            azptx::mbar_wait(&tfull[as], aphase);
            azptx::tc_fence_after();
            const int row = m_blk * 2 * Cfg::BM + (int)rank * Cfg::BM + q * 32 + lane;
            const bool live = row < M;
            float mean = 0.f, rstd = 1.f;
            if (args.stats_in && live) {
                const float2* sp = reinterpret_cast<const float2*>(args.stats_in + (size_t)row * AZ_STATS_LD);
                float s1 = 0.f, s2 = 0.f;
                for (int p = 0; p < args.stats_parts; ++p) {
                    const float2 v = sp[p];
                    s1 += v.x;
                    s2 += v.y;
                }
                mean = s1 * args.ln_inv_k;
                rstd = rsqrtf(fmaxf(s2 * args.ln_inv_k - mean * mean, 0.f) + args.ln_eps);
            }
            const float nmr = -mean * rstd;
            const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + as * Cfg::ACC_STRIDE;
            unsigned char* stg = args.stage_out ? smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES) + AZ_BAR_BYTES + (warp - 2) * 4096 : nullptr;
            az_gemm_epilogue_row<BN, EPI>(args, tacc, row, live, n_blk, half, rstd, nmr, stg, &tm_o,
                                          m_blk * 2 * Cfg::BM + (int)rank * Cfg::BM + q * 32);
            azptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) azptx::mbar_arrive_cluster(ltempty0 + as * 8);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
        if (lane == 0) azptx::bulk_wait0();
    }
    azptx::tc_fence_before();
    __syncthreads();
    azptx::cluster_sync();                                      // nobody leaves while the peer may still signal or read it
    if (warp == 1) azptx::tmem_dealloc_pair<Cfg::TMEM_COLS>(tmem_base);
}
