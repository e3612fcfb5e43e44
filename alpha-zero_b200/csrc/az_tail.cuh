// Fused cls-row tail for small evaluation batches: out-proj -> norm2 -> mlp1 (GELU) -> mlp2 (+ residual) -> final norm ->
// policy / value heads -> softmax / tanh in ONE kernel (reference: ai/nn.py:44-50 block tail, :78-83 Net.norm + heads,
// ai/mcts.py:49 softmax), replacing four dependent GEMM launches of az_gemm.cuh whose cost at a few hundred to a few
// thousand rows is launch + prologue + barrier chain, not arithmetic.
// STATUS: correct (tests/test_gpu_net.py::test_fused_tail_matches_gemm_chain) and OPT-IN (AZ_FUSED_TAIL=1).  Measured on B200 at
// 1792 boards: 58 us against 70 us for the four launches timed one by one, but the search graph runs those four back to back
// under programmatic dependent launch in ~50 us, and Connect4 self-play is 64.1 M simulations/s with this kernel against 66.6 M
// without.  What holds it at 58 us (tools/build_tail_trace.sh): 328 dependent MMA steps per CTA at ~100 cycles (33 k cycles),
// 22 k cycles of waiting for the 5-stage weight ring (every tile of 64 rows streams all 5.2 MB of weights through four SMs,
// twice the L2 -> SM traffic of 128-row GEMM tiles), and ~35 k cycles of cluster steps and reductions after the last MMA.
//
// Orientation: the activations are the N operand.  A cluster of C = D / 128 CTAs owns a tile of 64 board rows; every MMA is
// M = 128 output FEATURES (weights = A operand, K-major tiles streamed by TMA) x N = 64 rows.  `tcgen05.mma` costs the same
// ~66 cycles per K = 16 step for any N <= 128 (tools/umma_rate.cu), so what matters is the number of 128-feature x 16
// slabs a CTA has to issue: the features of every layer are split over the cluster.
//   S1  xc^T[128 own features][64]   = Wout2[own] . U^T            K = heads * 128     bias -> bf16, bulk-copied into every peer (DSMEM)
//   S2  h^T[512 own hidden][64]      = W1'[own] . xc^T             K = D               norm2 folded (rstd, mean x colsum), GELU -> bf16 (local)
//   S3  part[D out features][64]     = W2[:, own hidden] . h^T     K = 512 (own)       partial sums dumped to shared memory (fp32); the owner of a
//       xc2^T[128 own features][64]  = sum of partials + bias + xc                     feature PULLS them (ld.shared::cluster), rank order -> bf16 (local)
//   S4  plog[64 heads][64]           = Whead'[:, own features] . xc2^T   K = 128 (own) partial logits bulk-copied to CTA 0, final norm folded, softmax / tanh
// Remote traffic is bulk copies or loads / stores in which the 32 lanes of a warp touch 512 contiguous bytes: per-thread
// remote stores of 16 bytes at a 128-byte stride (first version) cost 6-8 k cycles per exchange.
// The epilogue thread owns one feature (TMEM lane) and 32 rows (TMEM columns) and writes activations transposed,
// [feature][64 rows] = one 128-byte swizzled line per feature: exactly the MN-major B operand of the next layer.
// Row statistics of the two LayerNorms are taken from the bf16 values the next layer reads (as in az_gemm.cuh), summed in a
// fixed order, so a row's result does not depend on where in the batch it sits.
//
//   warp 0  TMA producer (weights + U tiles, 3-stage ring)      warp 1  MMA issue      warps 2-9  epilogue (quarter x column half)
// Cluster-wide steps are mbarriers with one remote arrive per CTA (the producer / MMA warps never join a barrier.cluster).
#pragma once
#include "az_gemm2.cuh"
#include "az_front2.cuh"

#define AZ_TAIL_ROWS 64
#define AZ_TAIL_STAGES 5
#define AZ_TAIL_THREADS 320
#define AZ_TAIL_STAGE_BYTES 16384                   // one weight tile (128 features x 64 k) or one U tile (64 rows x 64 k, half used)
#define AZ_TAIL_XCT 0                               // 64 KB  xc^T; later partial-sum dumps 0, 1; in CTA 0 the peers' head partials
#define AZ_TAIL_HT 65536                            // 64 KB  h^T (two halves); later dump 2 | xc2^T | head-partial staging
#define AZ_TAIL_X2T (65536 + 32768)
#define AZ_TAIL_HPS (65536 + 49152)
#define AZ_TAIL_RING 131072
#define AZ_TAIL_MISC (131072 + AZ_TAIL_STAGES * AZ_TAIL_STAGE_BYTES)
#define AZ_TAIL_SMEM (AZ_TAIL_MISC + 8192 + 1024)

// -DAZ_TAIL_TRACE: cumulative phase clocks of block 0 (MMA thread, first epilogue thread), printed by the kernel
#ifdef AZ_TAIL_TRACE
#define AZ_TT_DECL long long tt_t[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long tt_c = clock64();
#define AZ_TT(k) { const long long n_ = clock64(); tt_t[k] += n_ - tt_c; tt_c = n_; }
#else
#define AZ_TT_DECL
#define AZ_TT(k)
#endif

struct AzTailArgs {
    AzBatch batch;
    int D, K2, A, C;                                   // C = D / 128 = cluster size
    const float *b_out, *b1, *cs1, *b2, *b_head, *cs_head;
    float *priors, *values, *logits;
    float ln_eps;
};

namespace azptx {
__device__ __forceinline__ void st_cluster_v4(uint32_t cluster_addr, uint4 v) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_cluster_v2f(uint32_t cluster_addr, float a, float b) {
    asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(cluster_addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ uint4 ld_cluster_v4(uint32_t cluster_addr) {
    uint4 v;
    asm volatile("ld.shared::cluster.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(cluster_addr) : "memory");
    return v;
}
// bulk copy of this CTA's shared memory into a peer's, completion (bytes) on a barrier of the peer
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t dst_cluster_addr, uint32_t src_cta_addr, uint32_t bytes, uint32_t bar_cluster_addr) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_cluster_addr), "r"(src_cta_addr), "r"(bytes), "r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void fence_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait_cluster(bar, parity)) {
        if (++spins > 20000000u) {
            printf("az_tail: cluster mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
            __trap();
        }
    }
}
}  // namespace azptx

// One step of the cluster: every epilogue thread of every CTA has done its part (incl. remote stores) before anyone goes on.
__device__ __forceinline__ void az_tail_step(uint64_t* bar, uint32_t parity, int C, int te) {
    azptx::fence_cluster();
    azptx::bar_sync_named(3, 256);
    if (te == 0) {
        const uint32_t a = azptx::smem_u32(bar);
        for (int p = 0; p < C; ++p) azptx::mbar_arrive_cluster(azptx::mapa(a, (uint32_t)p));
    }
    azptx::mbar_wait_cluster(bar, parity);
}

// byte offset of rows [8 * chunk, 8 * chunk + 8) of feature line f in a [feature][64 rows] bf16 tile (SWIZZLE_128B)
__device__ __forceinline__ uint32_t az_tail_line(int f, int chunk) { return (uint32_t)f * 128u + ((((uint32_t)chunk) ^ ((uint32_t)f & 7u)) << 4); }

// (sum, sum of squares) over features [f0, f0 + nf) of row r of a transposed bf16 tile
__device__ __forceinline__ float2 az_tail_row_stats(const unsigned char* tile, int f0, int nf, int r) {
    float s1 = 0.f, s2 = 0.f;
    for (int f = f0; f < f0 + nf; ++f) {
        const float v = __bfloat162float(*reinterpret_cast<const bf16*>(tile + az_tail_line(f, r >> 3) + (r & 7) * 2));
        s1 += v;
        s2 = fmaf(v, v, s2);
    }
    return make_float2(s1, s2);
}


// fp32 exchange tile of one (warp, 16-byte chunk i): the 32 lanes of a warp touch 512 contiguous bytes
__device__ __forceinline__ uint32_t az_tail_xoff(int w8, int i, int lane) { return (uint32_t)((w8 * 8 + i) * 512 + lane * 16); }

__global__ void __launch_bounds__(AZ_TAIL_THREADS, 1)
az_tail_kernel(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_wo, const __grid_constant__ CUtensorMap tm_w1,
               const __grid_constant__ CUtensorMap tm_w2, const __grid_constant__ CUtensorMap tm_wh, AzTailArgs a) {
    extern __shared__ unsigned char az_tail_smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)az_tail_smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* xcT = smem + AZ_TAIL_XCT;
    unsigned char* hT = smem + AZ_TAIL_HT;
    unsigned char* x2T = smem + AZ_TAIL_X2T;
    unsigned char* hps = smem + AZ_TAIL_HPS;                   // this CTA's partial head outputs, staged for the bulk copy to CTA 0
    float* s_logit = reinterpret_cast<float*>(hT);             // CTA 0, after the partial sums there have been pulled
    unsigned char* ring = smem + AZ_TAIL_RING;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AZ_TAIL_MISC);
    uint64_t *full = bars, *empty = bars + 5, *s1_full = bars + 10, *s2_full = bars + 11, *s3_full = bars + 15, *head_full = bars + 16,
             *h_ready = bars + 17, *x2_ready = bars + 19, *xc_full = bars + 20, *xc_ack = bars + 21, *hp_full = bars + 22, *tile_done = bars + 23,
             *cl = bars + 24;                                  // cl[3]: cluster steps (dumps ready, pulls done, end of tile)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 28);
    float* s_part = reinterpret_cast<float*>(smem + AZ_TAIL_MISC + 256);            // [4][64][2]
    float* s_rstd = s_part + 512;                                                    // [64]
    float* s_nmr = s_rstd + 64;                                                      // [64]  -mean * rstd
    float* s_stat = s_nmr + 64;                                                      // [4 ranks][64][2]  final-norm partials (CTA 0)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int C = a.C, D = a.D;
    const int c = (int)azptx::cluster_ctarank();
    const int cluster_id = (int)blockIdx.x / C, n_clusters = (int)gridDim.x / C;

    if (threadIdx.x == 0) {
        azptx::prefetch_tmap(&tm_u);
        azptx::prefetch_tmap(&tm_wo);
        azptx::prefetch_tmap(&tm_w1);
        azptx::prefetch_tmap(&tm_w2);
        azptx::prefetch_tmap(&tm_wh);
        for (int i = 0; i < AZ_TAIL_STAGES; ++i) {
            azptx::mbar_init(&full[i], 1);
            azptx::mbar_init(&empty[i], 1);
        }
        azptx::mbar_init(s1_full, 1);
        for (int i = 0; i < 4; ++i) azptx::mbar_init(&s2_full[i], 1);
        azptx::mbar_init(s3_full, 1);
        azptx::mbar_init(head_full, 1);
        azptx::mbar_init(&h_ready[0], 256);
        azptx::mbar_init(&h_ready[1], 256);
        azptx::mbar_init(x2_ready, 256);
        azptx::mbar_init(xc_full, 1);
        azptx::mbar_init(xc_ack, (uint32_t)(C > 1 ? C - 1 : 1));
        azptx::mbar_init(hp_full, 1);
        azptx::mbar_init(tile_done, 256);
        for (int i = 0; i < 3; ++i) azptx::mbar_init(&cl[i], (uint32_t)C);
        azptx::fence_barrier_init();
    }
    if (warp == 1) azptx::tmem_alloc<512>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (C > 1) azptx::cluster_sync();                          // every CTA's barriers exist before anyone arrives remotely
    azptx::griddep_wait();
    azptx::griddep_launch();

    const int M = az_batch_items(a.batch);
    const int tiles = (M + AZ_TAIL_ROWS - 1) / AZ_TAIL_ROWS;
    const int kb1 = a.K2 >> 6, kb2 = D >> 6;
    constexpr uint32_t idesc_k = azptx::umma_idesc_bf16(128, 64);                    // B K-major (U)
    constexpr uint32_t idesc_mn = azptx::umma_idesc_bf16(128, 64) | (1u << 16);      // B MN-major (transposed activations)

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------------------------------------------
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            const uint64_t pol_w = azptx::l2_policy_evict_last(), pol_a = azptx::l2_policy_evict_first();
            auto load = [&](const CUtensorMap* tm, int k0, int r0, uint32_t bytes, uint64_t pol) {
                azptx::mbar_wait(&empty[stage], phase ^ 1);
                azptx::mbar_arrive_expect_tx(&full[stage], bytes);
                azptx::tma_load_2d_hint(ring + (size_t)stage * AZ_TAIL_STAGE_BYTES, tm, &full[stage], k0, r0, pol);
                if (++stage == AZ_TAIL_STAGES) { stage = 0; phase ^= 1; }
            };
            for (int tile = cluster_id; tile < tiles; tile += n_clusters) {
                for (int kb = 0; kb < kb1; ++kb) {                                    // S1: a weight stage, then a U stage
                    load(&tm_wo, kb * 64, c * 128, 16384, pol_w);
                    load(&tm_u, kb * 64, tile * AZ_TAIL_ROWS, 8192, pol_a);
                }
                for (int mt = 0; mt < 4; ++mt)                                         // S2
                    for (int kb = 0; kb < kb2; ++kb) load(&tm_w1, kb * 64, c * 512 + mt * 128, 16384, pol_w);
                for (int h = 0; h < 2; ++h)                                            // S3
                    for (int mt = 0; mt < C; ++mt)
                        for (int kb = 0; kb < 4; ++kb) load(&tm_w2, c * 512 + h * 256 + kb * 64, mt * 128, 16384, pol_w);
                for (int kb = 0; kb < 2; ++kb) load(&tm_wh, c * 128 + kb * 64, 0, 8192, pol_w);   // S4 (64 head rows)
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------ MMA issue -----------------------------------------------------------------------
        if (lane == 0) {
            int stage = 0, it = 0;
            uint32_t phase = 0;
            AZ_TT_DECL
            auto next = [&]() { if (++stage == AZ_TAIL_STAGES) { stage = 0; phase ^= 1; } };
            // four K = 16 steps of one ring stage against a transposed activation block of 64 features
            auto kblock_mn = [&](uint32_t acc, uint32_t act, bool first) {
                AZ_TT(0)
                azptx::mbar_wait(&full[stage], phase);
                AZ_TT(1)
                azptx::tc_fence_after();
                const uint64_t dw = azptx::umma_desc_k128(azptx::smem_u32(ring + (size_t)stage * AZ_TAIL_STAGE_BYTES));
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    azptx::umma_bf16(acc, dw + 2 * k, azptx::umma_desc_mn128(act + (uint32_t)k * 2048, 8192), idesc_mn, !(first && k == 0));
                azptx::umma_commit(&empty[stage]);
                next();
            };
            for (int tile = cluster_id; tile < tiles; tile += n_clusters, ++it) {
                const uint32_t par = (uint32_t)(it & 1);
                if (it > 0) azptx::mbar_wait(tile_done, par ^ 1);                     // the previous tile's TMEM reads are over
                azptx::tc_fence_after();
                for (int kb = 0; kb < kb1; ++kb) {                                    // S1
                    AZ_TT(0)
                    azptx::mbar_wait(&full[stage], phase);
                    const int sw_stage = stage;
                    next();
                    azptx::mbar_wait(&full[stage], phase);
                    AZ_TT(1)
                    azptx::tc_fence_after();
                    const uint64_t dw = azptx::umma_desc_k128(azptx::smem_u32(ring + (size_t)sw_stage * AZ_TAIL_STAGE_BYTES));
                    const uint64_t du = azptx::umma_desc_k128(azptx::smem_u32(ring + (size_t)stage * AZ_TAIL_STAGE_BYTES));
#pragma unroll
                    for (int k = 0; k < 4; ++k) azptx::umma_bf16(tmem_base, dw + 2 * k, du + 2 * k, idesc_k, (kb | k) != 0);
                    azptx::umma_commit(&empty[sw_stage]);
                    azptx::umma_commit(&empty[stage]);
                    next();
                }
                azptx::umma_commit(s1_full);
                AZ_TT(0)
                azptx::mbar_wait(xc_full, par);                                       // xc^T: own slice written, the peers' slices have landed
                AZ_TT(2)
                azptx::tc_fence_after();
                for (int mt = 0; mt < 4; ++mt) {                                      // S2
                    for (int kb = 0; kb < kb2; ++kb) kblock_mn(tmem_base + (uint32_t)(mt * 64), azptx::smem_u32(xcT) + (uint32_t)kb * 8192, kb == 0);
                    azptx::umma_commit(&s2_full[mt]);
                }
                for (int h = 0; h < 2; ++h) {                                         // S3
                    AZ_TT(0)
                    azptx::mbar_wait(&h_ready[h], par);
                    AZ_TT(3 + h)
                    azptx::tc_fence_after();
                    for (int mt = 0; mt < C; ++mt)
                        for (int kb = 0; kb < 4; ++kb)
                            kblock_mn(tmem_base + 256u + (uint32_t)(mt * 64), azptx::smem_u32(hT) + (uint32_t)(h * 32768 + kb * 8192), h == 0 && kb == 0);
                }
                azptx::umma_commit(s3_full);
                AZ_TT(0)
                azptx::mbar_wait(x2_ready, par);                                      // S4
                AZ_TT(5)
                azptx::tc_fence_after();
                for (int kb = 0; kb < 2; ++kb) kblock_mn(tmem_base, azptx::smem_u32(x2T) + (uint32_t)kb * 8192, kb == 0);
                azptx::umma_commit(head_full);
            }
#ifdef AZ_TAIL_TRACE
            if (blockIdx.x == 0) printf("tail mma thread: issue %lld  ring wait %lld  wait xc^T %lld  wait h0 %lld  wait h1 %lld  wait xc2^T %lld\n", tt_t[0], tt_t[1],
                                        tt_t[2], tt_t[3], tt_t[4], tt_t[5]);
#endif
        }
        __syncwarp();
    } else {
        // ------------------------------ epilogue: thread = (feature lane fl, rows col0 .. col0 + 31) -----------------------
        const int te = threadIdx.x - 64;
        const int w8 = warp - 2;
        const int q = warp & 3, ch = (warp - 2) >> 2;
        const int fl = q * 32 + lane, col0 = ch * 32;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)col0;
        const int sr = te & 63, sp = te >> 6;                                         // statistics: row, feature quarter
        const uint32_t xct_u32 = azptx::smem_u32(xcT), ht_u32 = azptx::smem_u32(hT);
        int it = 0;
        AZ_TT_DECL
        for (int tile = cluster_id; tile < tiles; tile += n_clusters, ++it) {
            const uint32_t par = (uint32_t)(it & 1);
            const int row0 = tile * AZ_TAIL_ROWS;
            AZ_TT(0)
            uint32_t xr[16];                                                          // this thread's xc values (bf16 pairs): the residual of S3
            // ---- S1 epilogue: bias, bf16 -> own slice of xc^T; one bulk copy per peer
            {
                azptx::mbar_wait(s1_full, par);
                AZ_TT(1)
                azptx::tc_fence_after();
                uint32_t v[32];
                azptx::tmem_ld32(lane_addr, v);
                azptx::tmem_ld_wait();
                const int feat = c * 128 + fl;
                const float bo = __ldg(a.b_out + feat);
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const __nv_bfloat162 p = __floats2bfloat162_rn(__uint_as_float(v[2 * e]) + bo, __uint_as_float(v[2 * e + 1]) + bo);
                    xr[e] = *reinterpret_cast<const uint32_t*>(&p);
                }
#pragma unroll
                for (int qd = 0; qd < 4; ++qd)
                    *reinterpret_cast<uint4*>(xcT + az_tail_line(feat, ch * 4 + qd)) = make_uint4(xr[4 * qd], xr[4 * qd + 1], xr[4 * qd + 2], xr[4 * qd + 3]);
                azptx::fence_proxy_async_all();
                azptx::tc_fence_before();
                azptx::bar_sync_named(2, 256);
                if (te == 0) {
                    if (C > 1) azptx::mbar_arrive_expect_tx(xc_full, (uint32_t)(C - 1) * 16384u); else azptx::mbar_arrive(xc_full);
                    const uint32_t src = xct_u32 + (uint32_t)c * 16384u, bar = azptx::smem_u32(xc_full);
                    for (int p = 0; p < C; ++p)
                        if (p != c) azptx::bulk_copy_to_peer(azptx::mapa(src, (uint32_t)p), src, 16384u, azptx::mapa(bar, (uint32_t)p));
                }
            }
            AZ_TT(2)
            azptx::mbar_wait(xc_full, par);
            if (te == 0 && C > 1) {                                                   // tell the peers their copy into this CTA has been seen (their source may be reused)
                const uint32_t ack = azptx::smem_u32(xc_ack);
                for (int p = 0; p < C; ++p)
                    if (p != c) azptx::mbar_arrive_cluster(azptx::mapa(ack, (uint32_t)p));
            }
            AZ_TT(3)
            // ---- norm2 statistics of the 64 rows over all D features (bf16 values, fixed order)
            {
                const float2 st = az_tail_row_stats(xcT, sp * (D >> 2), D >> 2, sr);
                s_part[(sp * 64 + sr) * 2] = st.x;
                s_part[(sp * 64 + sr) * 2 + 1] = st.y;
                azptx::bar_sync_named(2, 256);
                if (te < 64) {
                    float s1 = 0.f, s2 = 0.f;
                    for (int p = 0; p < 4; ++p) {
                        s1 += s_part[(p * 64 + te) * 2];
                        s2 += s_part[(p * 64 + te) * 2 + 1];
                    }
                    const float mean = s1 / (float)D;
                    const float rstd = rsqrtf(fmaxf(s2 / (float)D - mean * mean, 0.f) + a.ln_eps);
                    s_rstd[te] = rstd;
                    s_nmr[te] = -mean * rstd;
                }
                azptx::bar_sync_named(2, 256);
            }
            AZ_TT(4)
            // ---- S2 epilogues: norm2 fold, GELU, bf16 -> h^T (local)
            for (int mt = 0; mt < 4; ++mt) {
                AZ_TT(6)
                azptx::mbar_wait(&s2_full[mt], par);
                AZ_TT(5)
                azptx::tc_fence_after();
                uint32_t v[32];
                azptx::tmem_ld32(lane_addr + (uint32_t)(mt * 64), v);
                azptx::tmem_ld_wait();
                const int hf = c * 512 + mt * 128 + fl;
                const float cs = __ldg(a.cs1 + hf), bb = __ldg(a.b1 + hf);
                const int lf = mt * 128 + fl;
#pragma unroll
                for (int qd = 0; qd < 4; ++qd) {
                    uint32_t w[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int r0 = col0 + qd * 8 + 2 * e;
                        const float g0 = az_gelu(fmaf(s_rstd[r0], __uint_as_float(v[qd * 8 + 2 * e]), fmaf(s_nmr[r0], cs, bb)));
                        const float g1 = az_gelu(fmaf(s_rstd[r0 + 1], __uint_as_float(v[qd * 8 + 2 * e + 1]), fmaf(s_nmr[r0 + 1], cs, bb)));
                        const __nv_bfloat162 p = __floats2bfloat162_rn(g0, g1);
                        w[e] = *reinterpret_cast<const uint32_t*>(&p);
                    }
                    *reinterpret_cast<uint4*>(hT + az_tail_line(lf, ch * 4 + qd)) = make_uint4(w[0], w[1], w[2], w[3]);
                }
                if (mt & 1) {
                    azptx::fence_proxy_async();
                    azptx::tc_fence_before();
                    azptx::mbar_arrive(&h_ready[mt >> 1]);
                }
            }
            AZ_TT(6)
            azptx::mbar_wait(s3_full, par);
            if (C > 1) azptx::mbar_wait(xc_ack, par);                                 // every peer has received this CTA's xc^T slice: xc^T may be overwritten
            AZ_TT(7)
            azptx::tc_fence_after();
            // ---- partial sums of the other CTAs' output features -> dumps in this CTA's shared memory (fp32, warp-contiguous)
            for (int mt = 0; mt < C; ++mt) {
                if (mt == c) continue;
                uint32_t v[32];
                azptx::tmem_ld32(lane_addr + 256u + (uint32_t)(mt * 64), v);
                azptx::tmem_ld_wait();
                const int j = mt < c ? mt : mt - 1;
                unsigned char* dump = j < 2 ? xcT + (size_t)j * 32768 : hT;
#pragma unroll
                for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(dump + az_tail_xoff(w8, i, lane)) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            azptx::tc_fence_before();
            AZ_TT(8)
            az_tail_step(&cl[0], par, C, te);                                         // every CTA's dumps are complete
            AZ_TT(3)
            // ---- S3 epilogue: own partial + the peers' (pulled, rank order) + bias + residual -> xc2^T (local)
            {
                uint32_t v[32];
                azptx::tmem_ld32(lane_addr + 256u + (uint32_t)(c * 64), v);
                uint4 pv[3][8];                                                       // slot s = peer rank s (below this CTA) or s + 1 (above)
#pragma unroll
                for (int s3 = 0; s3 < 3; ++s3) {
                    if (s3 < C - 1) {
                        const int p = s3 < c ? s3 : s3 + 1;
                        const int j = c < p ? c : c - 1;
                        const uint32_t src = azptx::mapa(j < 2 ? xct_u32 + (uint32_t)j * 32768u : ht_u32, (uint32_t)p);
#pragma unroll
                        for (int i = 0; i < 8; ++i) pv[s3][i] = azptx::ld_cluster_v4(src + az_tail_xoff(w8, i, lane));
                    }
                }
                azptx::tmem_ld_wait();
                float acc[32];
#pragma unroll
                for (int e = 0; e < 32; ++e) acc[e] = 0.f;
#pragma unroll
                for (int pos = 0; pos < 4; ++pos) {                                   // rank order: peers below, own, peers above
                    if (pos < 3 && pos < c && pos < C - 1) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const uint4 t = pv[pos < 3 ? pos : 0][i];
                            acc[4 * i] += __uint_as_float(t.x); acc[4 * i + 1] += __uint_as_float(t.y);
                            acc[4 * i + 2] += __uint_as_float(t.z); acc[4 * i + 3] += __uint_as_float(t.w);
                        }
                    }
                }
#pragma unroll
                for (int e = 0; e < 32; ++e) acc[e] += __uint_as_float(v[e]);
#pragma unroll
                for (int pos = 0; pos < 3; ++pos) {
                    if (pos >= c && pos < C - 1) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const uint4 t = pv[pos][i];
                            acc[4 * i] += __uint_as_float(t.x); acc[4 * i + 1] += __uint_as_float(t.y);
                            acc[4 * i + 2] += __uint_as_float(t.z); acc[4 * i + 3] += __uint_as_float(t.w);
                        }
                    }
                }
                const float b2 = __ldg(a.b2 + c * 128 + fl);
#pragma unroll
                for (int qd = 0; qd < 4; ++qd) {
                    uint32_t w[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const __nv_bfloat162 xv = *reinterpret_cast<const __nv_bfloat162*>(&xr[qd * 4 + e]);
                        const __nv_bfloat162 p = __floats2bfloat162_rn(acc[qd * 8 + 2 * e] + b2 + __low2float(xv), acc[qd * 8 + 2 * e + 1] + b2 + __high2float(xv));
                        w[e] = *reinterpret_cast<const uint32_t*>(&p);
                    }
                    *reinterpret_cast<uint4*>(x2T + az_tail_line(fl, ch * 4 + qd)) = make_uint4(w[0], w[1], w[2], w[3]);
                }
                azptx::fence_proxy_async();
                azptx::tc_fence_before();
                azptx::mbar_arrive(x2_ready);
            }
            // ---- final-norm partial statistics of this CTA's 128 features -> CTA 0
            {
                azptx::bar_sync_named(2, 256);
                const float2 st = az_tail_row_stats(x2T, sp * 32, 32, sr);
                s_part[(sp * 64 + sr) * 2] = st.x;
                s_part[(sp * 64 + sr) * 2 + 1] = st.y;
                azptx::bar_sync_named(2, 256);
                if (te < 64) {
                    float s1 = 0.f, s2 = 0.f;
                    for (int p = 0; p < 4; ++p) {
                        s1 += s_part[(p * 64 + te) * 2];
                        s2 += s_part[(p * 64 + te) * 2 + 1];
                    }
                    azptx::st_cluster_v2f(azptx::mapa(azptx::smem_u32(s_stat + (c * 64 + te) * 2), 0u), s1, s2);
                }
            }
            // ---- this CTA's partial head outputs (lanes 0-63 of the accumulator): registers (CTA 0) / staged for CTA 0
            uint32_t hp[32];
            AZ_TT(9)
            azptx::mbar_wait(head_full, par);
            AZ_TT(10)
            azptx::tc_fence_after();
            if (q < 2) {
                azptx::tmem_ld32(lane_addr, hp);
                azptx::tmem_ld_wait();
                if (c != 0) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(hps + az_tail_xoff(ch * 2 + q, i, lane)) = make_uint4(hp[4 * i], hp[4 * i + 1], hp[4 * i + 2], hp[4 * i + 3]);
                    azptx::fence_proxy_async_all();
                }
            }
            azptx::tc_fence_before();
            AZ_TT(9)
            az_tail_step(&cl[1], par, C, te);                                         // every CTA has pulled: the dumps (CTA 0's xc^T region) are free
            AZ_TT(3)
            if (te == 0 && C > 1) {
                if (c == 0) azptx::mbar_arrive_expect_tx(hp_full, (uint32_t)(C - 1) * 16384u);
                else azptx::bulk_copy_to_peer(azptx::mapa(xct_u32 + (uint32_t)(c - 1) * 16384u, 0u), azptx::smem_u32(hps), 16384u, azptx::mapa(azptx::smem_u32(hp_full), 0u));
            }
            // ---- CTA 0: final norm fold, logits, softmax / tanh
            if (c == 0) {
                if (te < 64) {
                    float s1 = 0.f, s2 = 0.f;
                    for (int p = 0; p < C; ++p) {
                        s1 += s_stat[(p * 64 + te) * 2];
                        s2 += s_stat[(p * 64 + te) * 2 + 1];
                    }
                    const float mean = s1 / (float)D;
                    const float rstd = rsqrtf(fmaxf(s2 / (float)D - mean * mean, 0.f) + a.ln_eps);
                    s_rstd[te] = rstd;
                    s_nmr[te] = -mean * rstd;
                }
                if (C > 1) azptx::mbar_wait(hp_full, par);
                azptx::bar_sync_named(2, 256);
                if (q < 2) {
                    float acc[32];
#pragma unroll
                    for (int e = 0; e < 32; ++e) acc[e] = __uint_as_float(hp[e]);
                    for (int p = 1; p < C; ++p) {
                        const unsigned char* src = xcT + (size_t)(p - 1) * 16384;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 t = *reinterpret_cast<const float4*>(src + az_tail_xoff(ch * 2 + q, i, lane));
                            acc[4 * i] += t.x; acc[4 * i + 1] += t.y; acc[4 * i + 2] += t.z; acc[4 * i + 3] += t.w;
                        }
                    }
                    const float cs = __ldg(a.cs_head + fl), bb = __ldg(a.b_head + fl);
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const int r = col0 + e;
                        s_logit[r * 64 + (fl ^ r)] = fmaf(s_rstd[r], acc[e], fmaf(s_nmr[r], cs, bb));
                    }
                }
                azptx::bar_sync_named(2, 256);
                if (te < 64 && row0 + te < M) {
                    const int r = te, A = a.A;
                    const size_t g = (size_t)(a.batch.offset + row0 + r);
                    float mx = -3.0e38f;
                    for (int o = 0; o < A; ++o) mx = fmaxf(mx, s_logit[r * 64 + (o ^ r)]);
                    float sum = 0.f;
                    for (int o = 0; o < A; ++o) {
                        const float l = s_logit[r * 64 + (o ^ r)];
                        if (a.logits) a.logits[g * A + o] = l;
                        sum += expf(l - mx);
                    }
                    const float inv = 1.0f / sum;
                    for (int o = 0; o < A; ++o) a.priors[g * A + o] = expf(s_logit[r * 64 + (o ^ r)] - mx) * inv;
                    a.values[g] = tanhf(s_logit[r * 64 + (A ^ r)]);
                }
            }
            azptx::mbar_arrive(tile_done);
            AZ_TT(11)
            if (tile + n_clusters < tiles) az_tail_step(&cl[2], par, C, te);          // more tiles: nobody overwrites a peer's buffers before it is done
            AZ_TT(3)
        }
#ifdef AZ_TAIL_TRACE
        if (blockIdx.x == 0 && te == 0)
            printf("tail epilogue: wait S1 %lld  S1 epi %lld  cluster steps %lld  norm2 stats %lld  wait S2 %lld  S2 epi %lld  wait S3 %lld  dump %lld  pull+S3 epi+stats %lld  wait head %lld  final %lld\n",
                   tt_t[1], tt_t[2], tt_t[3], tt_t[4], tt_t[5], tt_t[6], tt_t[7], tt_t[8], tt_t[9], tt_t[10], tt_t[11]);
#endif
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (C > 1) azptx::cluster_sync();                          // nobody leaves while a peer may still read or write its shared memory
    if (warp == 1) azptx::tmem_dealloc<512>(tmem_base);
}
