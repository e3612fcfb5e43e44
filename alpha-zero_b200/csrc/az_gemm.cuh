// Persistent, warp-specialised bf16 GEMM for sm_100a:  out[M,N] = epi(A[M,K] * W[N,K]^T + bias).
//
// This is the dense contraction of the policy/value net (reference: the nn.Linear /
// nn.MultiheadAttention projections of ai/nn.py:42-50,71-72) on the 5th-gen tensor cores:
//   warp 0   TMA producer: cp.async.bulk.tensor 128B-swizzled K-major tiles of A and W into a
//            STAGES-deep shared-memory ring, completion on mbarriers;
//   warp 1   MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) with the fp32
//            accumulator in TMEM; tcgen05.commit releases ring slots / publishes the tile;
//   warps 2-9 epilogue: tcgen05.ld the accumulator (each warp its 32-lane quarter and one half of the
//            tile's columns), add bias (+ erf GELU | + residual), convert to bf16; tiles of 128 / 256 columns
//            stage 32 rows x 64 columns per warp in shared memory (128-byte swizzle, conflict-free 16-byte
//            writes) and store them with cp.async.bulk.tensor (TMA store: one instruction instead of 128
//            16-byte global stores at a row stride); narrower tiles and fp32 side outputs store directly.
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
//
// LayerNorm around a GEMM never runs as its own kernel on the cls-row path.  The GEMM that PRODUCES a
// row also emits partial (sum, sum of squares) of its bf16 outputs (stats_out); the GEMM that CONSUMES
// LayerNorm(row) runs on the raw row with the gain folded into the weights and applies
//     out = rstd * (acc - mean * colsum) + bias',   colsum[n] = sum_k W'[n][k],  bias' = bias + W beta
// in its epilogue (stats_in).  AZ_EPI_HEAD (BN = 64, one thread owns all head outputs of a row)
// finishes with the softmax over the policy logits and tanh of the value (ai/nn.py:78-83, ai/mcts.py:49).
// M may come from a device-side counter (the evaluation batch of the forest is only known
// on the device), so the kernel is CUDA-graph capturable with a fixed grid.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "az_ptx.cuh"

#define AZ_EPI_BIAS 0
#define AZ_EPI_GELU 1
#define AZ_EPI_RESID 2
#define AZ_EPI_HEAD 3
#define AZ_STATS_LD 64          // floats per row of a statistics buffer: up to 32 (sum, sum of squares) partials
#define AZ_GEMM_THREADS 320
#define AZ_HEAD_STAGE_BYTES (4 * 32 * 64 * 4)
#define AZ_OUT_STAGE_BYTES (8 * 4096)   // one 32 x 64 bf16 tile per epilogue warp
#define AZ_BAR_BYTES 1024               // barriers + TMEM slot, keeps what follows 1024-byte aligned

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
    // LayerNorm folded around the GEMM (see the header comment)
    float* stats_out;           // [M][AZ_STATS_LD]: partial (sum, sum sq) per (N tile, column half) of the bf16 outputs
    const float* stats_in;      // partials of the A rows, stats_parts pairs valid
    const float* colsum;        // [N]
    int stats_parts;
    float ln_inv_k, ln_eps;
    // AZ_EPI_HEAD: softmax over columns [0, head_A) -> priors, tanh(column head_A) -> values
    int head_A;
    float *priors, *values, *logits;
    int w_early;                // the first W tiles are requested BEFORE griddepcontrol.wait (weights do not depend on the predecessor;
                                // only where the row counter was written at least two kernels upstream)
    int l2_hints;               // A loads evict_first, W loads evict_last (the cls-row chain inside the search step)
    int stage_out;              // 1: bf16 tiles leave through shared memory + TMA stores (tm_o describes `out`, box 64 columns x 32 rows)
    int dbg;                    // experiments (tools/gemm_bench.py): bit 0 no MMA, bit 1 no W loads, bit 2 no A loads (results are garbage)
};

__device__ __forceinline__ int az_gemm_rows(const AzGemmArgs& a) {
    if (!a.m_count) return a.m_fixed;
    int items = *a.m_count - a.m_item_offset;
    items = items < 0 ? 0 : (items > a.m_item_cap ? a.m_item_cap : items);
    return items * a.m_rows_per_item;
}

// The epilogue of the GELU GEMM is SFU-bound with the erf form (2 MUFU per element: 4 SFU lanes per SM sub-partition
// = 16 cycles per warp and element pair against a K = 512 mainloop of ~4.4 k cycles per 128 x 256 tile).  The tanh form
//     gelu(x) ~ 0.5 x (1 + tanh(0.7978845608 (x + 0.044715 x^3)))
// needs ONE MUFU (tanh.approx, |error| ~ 5e-4) and 6 instructions.  It differs from nn.GELU()'s erf form by <= 4.7e-4 absolute
// (reached near |x| = 2.3, relative 2e-4) -- below the bf16 rounding of the stored activation (2^-9 relative) at every
// magnitude where |gelu| >= 0.25, but up to 12 bf16 ulps in the shallow negative lobe (gelu ~ -0.01 near x = -2.7).
// Measured: mlp1 at 65536 rows 185 -> 173 us, at 4096 rows 16.3 -> 15.3 us, nothing visible in the step time -- so the
// erf polynomial (below half a bf16 ulp wherever |gelu| >= 1e-2) stays the default; -DAZ_GELU_TANH selects this form.
__device__ __forceinline__ float az_gelu_tanh(float x) {
    const float x2 = x * x;
    const float inner = fmaf(x2, 0.044715f * 0.7978845608028654f, 0.7978845608028654f);
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(x * inner));
    const float hx = 0.5f * x;
    return fmaf(hx, t, hx);
}
#ifdef AZ_GELU_TANH
#define az_gelu az_gelu_tanh
#else
#define az_gelu az_gelu_erf
#endif

template <int BN>
struct AzGemmCfg {
    static constexpr int BM = 128, BK = 64;
    static constexpr int A_BYTES = BM * BK * 2;              // 16 KB
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGES = BN >= 192 ? 4 : 6;
    static constexpr int TMEM_COLS = 2 * BN <= 32 ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
    static constexpr int ACC_STRIDE = TMEM_COLS / 2;          // TMEM columns between the two accumulator stages
    static constexpr int SMEM = STAGES * (A_BYTES + B_BYTES) + AZ_BAR_BYTES + (BN == 128 || BN == 256 ? AZ_OUT_STAGE_BYTES : 0) + 1024;   // + align slack
};

// nn.GELU() default (erf form): 0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz & Stegun 7.1.25
// (|error| <= 2.5e-5, far below the bf16 output resolution):  z = |x|/sqrt2, t = 1/(1 + 0.47047 z),
// E = (a1 t + a2 t^2 + a3 t^3) exp(-z^2) = 1 - erf(z)  =>  gelu(x) = max(x, 0) - 0.5 |x| E.
__device__ __forceinline__ float az_gelu_erf(float x) {
    const float ax = fabsf(x);
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.47047f * 0.70710678118654752440f, 1.0f)));
    const float p = t * fmaf(t, fmaf(t, 0.5f * 0.7478556f, 0.5f * -0.0958798f), 0.5f * 0.3480242f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * (-0.5f * 1.4426950408889634f)));
    return fmaf(-ax, p * e, fmaxf(x, 0.0f));
}

// Epilogue of one accumulator tile for one thread (= one output row): BN/2 columns starting at
// column half*BN/2 of the tile, bias (+ LayerNorm fold) (+ GELU | + residual), bf16 row stores,
// optional row statistics.  Shared by the single-CTA and the CTA-pair kernels.
// stg: this warp's 4 KB staging tile (nullptr: direct stores); row_w0: first row of the warp's 32 rows.
template <int BN, int EPI>
__device__ __forceinline__ void az_gemm_epilogue_row(const AzGemmArgs& args, uint32_t tacc, int row, bool live, int n_blk, int half,
                                                     float rstd, float nmr, unsigned char* stg = nullptr, const CUtensorMap* tm_o = nullptr,
                                                     int row_w0 = 0) {
    const int lane_ = threadIdx.x & 31;
    constexpr int CPW = BN / 64;                               // 32-column chunks per warp
    float st1 = 0.f, st2 = 0.f;
#ifndef AZ_NO_EPI_PREFETCH                                          // the next chunk's tcgen05.ld runs under this chunk's arithmetic
    uint32_t rn[32];
    azptx::tmem_ld32(tacc + half * CPW * 32, rn);
#endif
#pragma unroll 1
    for (int cc = 0; cc < CPW; ++cc) {
        const int c = half * CPW + cc;
        uint32_t r[32];
#ifndef AZ_NO_EPI_PREFETCH
        azptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = rn[i];
        if (cc + 1 < CPW) azptx::tmem_ld32(tacc + (c + 1) * 32, rn);
#else
        azptx::tmem_ld32(tacc + c * 32, r);
        azptx::tmem_ld_wait();
#endif
        const int col0 = n_blk * BN + c * 32;
        const bool f32_chunk = args.out_f32 && col0 >= args.f32_col0;   // warp-uniform: the staged path below synchronises the warp
        if (f32_chunk) {                                          // fp32 side output (attention scores)
            if (!live) continue;
            const float4* b4 = reinterpret_cast<const float4*>(args.bias + col0);
            float4* op = reinterpret_cast<float4*>(args.out_f32 + (size_t)row * 64 + (col0 - args.f32_col0));
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 bb = __ldg(b4 + i);
                op[i] = make_float4(__uint_as_float(r[4 * i + 0]) + bb.x, __uint_as_float(r[4 * i + 1]) + bb.y,
                                    __uint_as_float(r[4 * i + 2]) + bb.z, __uint_as_float(r[4 * i + 3]) + bb.w);
            }
        } else if (live || stg) {
            const float4* b4 = reinterpret_cast<const float4*>(args.bias + col0);
            const float4* c4 = reinterpret_cast<const float4*>(args.colsum + col0);
            uint4 rv[4];
            if (EPI == AZ_EPI_RESID && live) {
                const uint4* rp = reinterpret_cast<const uint4*>(args.resid + (size_t)row * args.ldr + col0);
#pragma unroll
                for (int i = 0; i < 4; ++i) rv[i] = rp[i];
            }
            uint4 ov[4];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float4 bb = __ldg(b4 + i);
                if (args.stats_in) {
                    const float4 cs = __ldg(c4 + i);
                    bb.x = fmaf(nmr, cs.x, bb.x); bb.y = fmaf(nmr, cs.y, bb.y);
                    bb.z = fmaf(nmr, cs.z, bb.z); bb.w = fmaf(nmr, cs.w, bb.w);
                }
                float v0 = fmaf(rstd, __uint_as_float(r[4 * i + 0]), bb.x);
                float v1 = fmaf(rstd, __uint_as_float(r[4 * i + 1]), bb.y);
                float v2 = fmaf(rstd, __uint_as_float(r[4 * i + 2]), bb.z);
                float v3 = fmaf(rstd, __uint_as_float(r[4 * i + 3]), bb.w);
#ifdef AZ_EPI_EXPERIMENT                                          // tools/gemm_bench.py: dbg bit 4 = no GELU, bit 3 = no output stores
                if (EPI == AZ_EPI_GELU && !(args.dbg & 16)) {
#else
                if (EPI == AZ_EPI_GELU) {
#endif
                    v0 = az_gelu(v0); v1 = az_gelu(v1); v2 = az_gelu(v2); v3 = az_gelu(v3);
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
                if (args.stats_out) {             // statistics of the values the consumer will read
                    const float f0 = __bfloat162float(o01.x), f1 = __bfloat162float(o01.y);
                    const float f2 = __bfloat162float(o23.x), f3 = __bfloat162float(o23.y);
                    st1 += (f0 + f1) + (f2 + f3);
                    st2 = fmaf(f0, f0, fmaf(f1, f1, fmaf(f2, f2, fmaf(f3, f3, st2))));
                }
                uint32_t* ow = reinterpret_cast<uint32_t*>(&ov[i >> 1]) + (i & 1) * 2;
                ow[0] = *reinterpret_cast<uint32_t*>(&o01);
                ow[1] = *reinterpret_cast<uint32_t*>(&o23);
            }
#ifdef AZ_EPI_EXPERIMENT
            if ((args.dbg & 8) && ov[0].x != 0x7fc07fc1u) continue;   // value-dependent, so the arithmetic is not eliminated
#endif
            if (stg) {
                // 32 rows x 128 bytes, 16-byte chunks swizzled by the row: a 64-column pair of chunks is one TMA store
                if ((cc & 1) == 0) {
                    if (lane_ == 0) azptx::bulk_wait_read0();        // the previous store has read the tile
                    __syncwarp();
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(stg + lane_ * 128 + ((((cc & 1) * 4 + i) ^ (lane_ & 7)) << 4)) = live ? ov[i] : make_uint4(0u, 0u, 0u, 0u);
                if (cc & 1) {
                    azptx::fence_proxy_async();
                    __syncwarp();
                    if (lane_ == 0) {
                        azptx::tma_store_2d(tm_o, stg, col0 - 32, row_w0);
                        azptx::bulk_commit();
                    }
                }
            } else {
                uint4* op = reinterpret_cast<uint4*>(args.out + (size_t)row * args.ldo + col0);
#pragma unroll
                for (int i = 0; i < 4; ++i) op[i] = ov[i];
            }
        }
    }
    if (args.stats_out && live)
        *reinterpret_cast<float2*>(args.stats_out + (size_t)row * AZ_STATS_LD + (n_blk * 2 + half) * 2) = make_float2(st1, st2);
}

// Head epilogue (BN = 64: one thread owns all head outputs of its row): final-norm fold, softmax over the policy logits,
// tanh of the value (ai/nn.py:78-83, ai/mcts.py:49).  hs: 32 x 64 floats of staging for this warp; r0w: the warp's first row.
__device__ __forceinline__ void az_gemm_epilogue_head(const AzGemmArgs& args, uint32_t tacc, float* hs, int r0w, int lane, bool live, int M,
                                                      float rstd, float nmr) {
    float l[64];
    {
        uint32_t r0[32], r1[32];
        azptx::tmem_ld32(tacc, r0);
        azptx::tmem_ld32(tacc + 32, r1);
        azptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            l[i] = fmaf(rstd, __uint_as_float(r0[i]), fmaf(nmr, __ldg(args.colsum + i), __ldg(args.bias + i)));
            l[32 + i] = fmaf(rstd, __uint_as_float(r1[i]), fmaf(nmr, __ldg(args.colsum + 32 + i), __ldg(args.bias + 32 + i)));
        }
    }
    const int A = args.head_A;
    const int nlive = min(max(M - r0w, 0), 32);
    const size_t obase = (size_t)(args.m_item_offset + r0w);
    float mx = -3.0e38f, val = 0.f;
#pragma unroll
    for (int o = 0; o < 64; ++o) {
        if (o < A) mx = fmaxf(mx, l[o]);
        if (o == A) val = l[o];
    }
    if (args.logits) {
#pragma unroll
        for (int o = 0; o < 64; ++o) if (o < A && live) hs[lane * A + o] = l[o];
        __syncwarp();
        for (int i = lane; i < nlive * A; i += 32) args.logits[obase * A + i] = hs[i];
        __syncwarp();
    }
    float sum = 0.f;
#pragma unroll
    for (int o = 0; o < 64; ++o) {
        if (o < A) {
            l[o] = expf(l[o] - mx);
            sum += l[o];
        }
    }
    const float inv = 1.0f / sum;
#pragma unroll
    for (int o = 0; o < 64; ++o) if (o < A && live) hs[lane * A + o] = l[o] * inv;
    __syncwarp();
    for (int i = lane; i < nlive * A; i += 32) args.priors[obase * A + i] = hs[i];
    __syncwarp();
    if (live) args.values[obase + lane] = tanhf(val);
}

template <int BN, int EPI>
__global__ void __launch_bounds__(AZ_GEMM_THREADS, 1)
az_gemm_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, const __grid_constant__ CUtensorMap tm_o,
               AzGemmArgs args) {
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
    const int n_tiles = args.N / BN;
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
            azptx::mbar_init(&tempty[i], EPI == AZ_EPI_HEAD ? 128 : 256);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 1) azptx::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Programmatic dependent launch: everything above overlapped the predecessor's tail; its outputs
    // (the A rows, the row statistics and the device-side row count) are read only from here on.
    int pre = 0;                                              // stages whose W tile is already in flight (producer thread only)
    if (args.w_early && !args.dbg && threadIdx.x == 0) {
        const int total0 = ((az_gemm_rows(args) + Cfg::BM - 1) / Cfg::BM) * n_tiles;
        if ((int)blockIdx.x < total0) {
            const uint64_t pol_w = azptx::l2_policy_evict_last();
            const int n_blk0 = (int)blockIdx.x % n_tiles;
            pre = kblocks < STAGES ? kblocks : STAGES;
            for (int kb = 0; kb < pre; ++kb) {
                azptx::mbar_arrive_expect_tx(&full[kb], Cfg::A_BYTES + Cfg::B_BYTES);
                azptx::tma_load_2d_hint(s_b + kb * Cfg::B_BYTES, &tm_b, &full[kb], kb * Cfg::BK, n_blk0 * BN, pol_w);
            }
        }
    }
    azptx::griddep_wait();
    azptx::griddep_launch();
    const int M = az_gemm_rows(args);
    const int m_tiles = (M + Cfg::BM - 1) / Cfg::BM;
    const int total = m_tiles * n_tiles;

    if (warp == 0) {
        if (lane == 0) {                                      // ---- TMA producer
            int stage = 0;
            uint32_t phase = 0;
            const uint64_t pol_a = azptx::l2_policy_evict_first(), pol_w = azptx::l2_policy_evict_last();
            for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
                const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
                for (int kb = 0; kb < kblocks; ++kb) {
                    azptx::mbar_wait(&empty[stage], phase ^ 1);
                    if (args.dbg & 6) {
                        const uint32_t bytes = ((args.dbg & 4) ? 0 : Cfg::A_BYTES) + ((args.dbg & 2) ? 0 : Cfg::B_BYTES);
                        if (bytes) azptx::mbar_arrive_expect_tx(&full[stage], bytes); else azptx::mbar_arrive(&full[stage]);
                        if (!(args.dbg & 4)) azptx::tma_load_2d(s_a + stage * Cfg::A_BYTES, &tm_a, &full[stage], kb * Cfg::BK, m_blk * Cfg::BM);
                        if (!(args.dbg & 2)) azptx::tma_load_2d(s_b + stage * Cfg::B_BYTES, &tm_b, &full[stage], kb * Cfg::BK, n_blk * BN);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    if (pre > 0) {                                // W of this stage was requested before the dependency wait
                        --pre;
                        azptx::tma_load_2d_hint(s_a + stage * Cfg::A_BYTES, &tm_a, &full[stage], kb * Cfg::BK, m_blk * Cfg::BM, pol_a);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    azptx::mbar_arrive_expect_tx(&full[stage], Cfg::A_BYTES + Cfg::B_BYTES);
                    if (args.l2_hints) {                          // activations stream through once per step, weights come back every step
                        azptx::tma_load_2d_hint(s_a + stage * Cfg::A_BYTES, &tm_a, &full[stage], kb * Cfg::BK, m_blk * Cfg::BM, pol_a);
                        azptx::tma_load_2d_hint(s_b + stage * Cfg::B_BYTES, &tm_b, &full[stage], kb * Cfg::BK, n_blk * BN, pol_w);
                    } else {
                        azptx::tma_load_2d(s_a + stage * Cfg::A_BYTES, &tm_a, &full[stage], kb * Cfg::BK, m_blk * Cfg::BM);
                        azptx::tma_load_2d(s_b + stage * Cfg::B_BYTES, &tm_b, &full[stage], kb * Cfg::BK, n_blk * BN);
                    }
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
                if (lane == 0 && (args.dbg & 1)) {
                    azptx::mbar_arrive(&empty[stage]);
                    if (kb == kblocks - 1) azptx::mbar_arrive(&tfull[as]);
                } else if (lane == 0) {
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
    } else if (EPI != AZ_EPI_HEAD || warp < 6) {              // ---- epilogue (warps 2..9; the head epilogue uses 2..5)
        const int q = warp & 3;                               // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;                     // which half of the tile's columns
        constexpr int CPW = EPI == AZ_EPI_HEAD ? BN / 32 : BN / 64;   // 32-column chunks per warp
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
            const int m_blk = tile / n_tiles, n_blk = tile - m_blk * n_tiles;
            const int row = m_blk * Cfg::BM + q * 32 + lane;
            const bool live = row < M;
            // Everything the epilogue reads besides the accumulator is requested while the mainloop still runs: on a grid of one
            // tile per CTA (the cls-row GEMMs) these round trips to L2 would otherwise start when the accumulator is ready.
            if (EPI != AZ_EPI_HEAD) {
                const int c0 = n_blk * BN + half * (BN / 2);
                if (lane * 32 < BN / 2) {
                    azptx::prefetch_l1(args.bias + c0 + lane * 32);
                    if (args.stats_in) azptx::prefetch_l1(args.colsum + c0 + lane * 32);
                }
                if (EPI == AZ_EPI_RESID && live) {
                    for (int b = 0; b < BN; b += 128) azptx::prefetch_l1(args.resid + (size_t)row * args.ldr + c0 + b / 2);
                }
            }
            float mean = 0.f, rstd = 1.f;
            if (args.stats_in && live) {                      // LayerNorm statistics of this A row
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
            const float nmr = -mean * rstd;                   // out = rstd*acc + nmr*colsum + bias
            azptx::mbar_wait(&tfull[as], aphase);
            azptx::tc_fence_after();
            const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + as * Cfg::ACC_STRIDE;
            if constexpr (EPI == AZ_EPI_HEAD) {
                static_assert(EPI != AZ_EPI_HEAD || BN == 64, "head epilogue: one 64-column tile per row");
                float* hs = reinterpret_cast<float*>(smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES) + AZ_BAR_BYTES) + q * 32 * 64;
                az_gemm_epilogue_head(args, tacc, hs, m_blk * Cfg::BM + q * 32, lane, live, M, rstd, nmr);
            } else {
                unsigned char* stg = (BN == 128 || BN == 256) && args.stage_out
                                         ? smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES) + AZ_BAR_BYTES + (warp - 2) * 4096 : nullptr;
                az_gemm_epilogue_row<BN, EPI>(args, tacc, row, live, n_blk, half, rstd, nmr, stg, &tm_o, m_blk * Cfg::BM + q * 32);
            }
            azptx::tc_fence_before();
            azptx::mbar_arrive(&tempty[as]);
            as ^= 1;
            if (as == 0) aphase ^= 1;
        }
        if (lane == 0) azptx::bulk_wait0();                    // this warp's TMA stores have landed before the CTA retires
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) azptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}
