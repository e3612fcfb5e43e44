// Board -> token embedding + LayerNorm on the tensor cores (sm_100a).
//
// Reference: PatchEmbedding conv (ai/nn.py:9-18) + cls token + positional embedding
// (ai/nn.py:33-35) + norm1 of the first block (ai/nn.py:54).  The conv over BINARY planes is a
// GEMM  X[token][D] = A[token][K] * Wemb[D][K]^T  whose A rows are built on the fly from the
// leaf bitboards:  A = [ im2col taps of the C planes (C*p*p 0/1 values) | one-hot(token index) ]
// and  Wemb = [ conv weight ; pos-emb + conv bias (row 0: cls + pos[0]) ],  K <= 128.
// Per 128-token tile:
//   warps 4-7  build the A tile in shared memory in the 128B-swizzled K-major layout
//   warp  8    one lane issues tcgen05.mma (M128, N<=256 per instruction, accumulator = whole row in TMEM)
//   warps 0-3  read the accumulator twice (tcgen05.ld): row mean / variance, then normalise -> bf16; both outputs leave
//              through shared memory + TMA stores (32 rows x 64 columns per warp and store): per-thread 16-byte row
//              stores at a 2 D-byte stride were 4096 LSU transactions per warp and tile and made this kernel 5x its
//              HBM time
// Wemb stays resident in shared memory (TMA, once per CTA); the LayerNorm affine is folded into
// the weights of the following projections at load time, so the epilogue writes (x-mean)*rstd.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "az_net_kernels.cuh"
#include "az_ptx.cuh"

struct AzEmbedMmaArgs {
    AzBatch batch;
    const uint64_t* leaf_words;   // [items][2][nw]
    const uint8_t* leaf_player;
    int nw, rows, cols, planes, patch, D, T;
    int kc;                       // planes*patch*patch : first one-hot column
    int kblocks;                  // K padded to kblocks*64 (1 or 2)
    bf16* x;                      // residual stream (raw embedding), may be written for cls rows only
    bf16* xn;                     // (x - mean) * rstd
    int write_all_x;
    int tma_out;                  // 1: tm_x / tm_xn describe x / xn (box 64 columns x 32 rows): staged TMA stores
};

#define AZ_EMBED_STAGE_BYTES (4 * 2 * 2 * 4096)     // per epilogue warp: two buffers of (xn | x) tiles, 32 rows x 128 bytes each
static inline size_t az_embed_mma_smem(int D, int kblocks, int T, int pp, bool staged) {
    return 1024 + (size_t)kblocks * D * 128 + 2 * (size_t)kblocks * 16384 + (staged ? AZ_EMBED_STAGE_BYTES : 0) + ((size_t)T * pp + 15) / 16 * 16 + 128;
}

template <int TMEM_COLS>
__global__ void __launch_bounds__(288, 1) az_embed_mma_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                                                              const __grid_constant__ CUtensorMap tm_xn, AzEmbedMmaArgs a) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
    const int KB = a.kblocks, D = a.D, T = a.T, pp = a.patch * a.patch;
    unsigned char* s_w = smem;                                   // [KB][D rows][128 B]
    unsigned char* s_a = s_w + (size_t)KB * D * 128;             // [2 stages][KB][128 rows][128 B]
    unsigned char* s_out = s_a + 2 * (size_t)KB * 16384;         // [4 warps][2 buffers][xn | x][32 rows][128 B] staging of the TMA stores
    uint8_t* s_tap = s_out + (a.tma_out ? AZ_EMBED_STAGE_BYTES : 0);   // [T][pp] source cell of each tap, 0xFF = outside
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_tap + ((size_t)T * pp + 15) / 16 * 16);
    uint64_t* w_full = bars;          // 1
    uint64_t* a_full = bars + 1;      // 2
    uint64_t* a_empty = bars + 3;     // 2
    uint64_t* acc_full = bars + 5;    // 1
    uint64_t* acc_empty = bars + 6;   // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 7);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int items = az_batch_items(a.batch);
    const long long rows_total = (long long)items * T;
    const int tiles = (int)((rows_total + 127) / 128);
    const int n_half = D > 256 ? 256 : D;                        // N of one MMA instruction
    const int halves = D / n_half;

    for (int i = threadIdx.x; i < T * pp; i += blockDim.x) {
        const int t = i / pp, tap = i - t * pp;
        int v = 0xFF;
        if (t > 0) {
            const int cell = t - 1, r = cell / a.cols, c = cell - r * a.cols;
            const int rr = r + tap / a.patch - (a.patch >> 1), cc = c + tap % a.patch - (a.patch >> 1);
            if (rr >= 0 && rr < a.rows && cc >= 0 && cc < a.cols) v = rr * a.cols + cc;
        }
        s_tap[i] = (uint8_t)v;
    }
    if (warp == 8 && lane == 0) {
        azptx::prefetch_tmap(&tm_w);
        azptx::mbar_init(w_full, 1);
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&a_full[i], 128);
            azptx::mbar_init(&a_empty[i], 1);
        }
        azptx::mbar_init(acc_full, 1);
        azptx::mbar_init(acc_empty, 128);
        azptx::fence_barrier_init();
    }
    if (warp == 0) azptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ------------------------------ weights in, then MMA issue -------------------------------
        if (lane == 0) {
            azptx::mbar_arrive_expect_tx(w_full, (uint32_t)(KB * D * 128));
            for (int kb = 0; kb < KB; ++kb)
                for (int h = 0; h < halves; ++h)
                    azptx::tma_load_2d(s_w + ((size_t)kb * halves + h) * n_half * 128, &tm_w, w_full, kb * 64, h * n_half);
        }
        azptx::mbar_wait(w_full, 0);
        const uint32_t idesc = azptx::umma_idesc_bf16(128, (uint32_t)n_half);
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int st = it & 1;
            azptx::mbar_wait(&a_full[st], (it >> 1) & 1);
            azptx::mbar_wait(acc_empty, (it & 1) ^ 1);
            azptx::tc_fence_after();
            if (lane == 0) {
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t da = azptx::umma_desc_k128(azptx::smem_u32(s_a + ((size_t)st * KB + kb) * 16384));
                    for (int h = 0; h < halves; ++h) {
                        const uint64_t db = azptx::umma_desc_k128(azptx::smem_u32(s_w + ((size_t)kb * halves + h) * n_half * 128));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            azptx::umma_bf16(tmem_base + h * n_half, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    }
                }
                azptx::umma_commit(&a_empty[st]);
                azptx::umma_commit(acc_full);
            }
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ------------------------------ A-tile producers: one thread per token ------------------
        const int r = threadIdx.x - 128;
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int st = it & 1;
            azptx::mbar_wait(&a_empty[st], ((it >> 1) & 1) ^ 1);
            const long long row = (long long)tile * 128 + r;
            const bool live = row < rows_total;
            const int b = live ? (int)(row / T) : 0, t = live ? (int)(row - (long long)b * T) : 0;
            const int item = a.batch.offset + b;
            uint64_t w0[AZ_MAX_WORDS], w1[AZ_MAX_WORDS];
#pragma unroll
            for (int i = 0; i < AZ_MAX_WORDS; ++i) {
                w0[i] = (live && i < a.nw) ? a.leaf_words[(size_t)item * 2 * a.nw + i] : 0ull;
                w1[i] = (live && i < a.nw) ? a.leaf_words[(size_t)item * 2 * a.nw + a.nw + i] : 0ull;
            }
            const bool stm = live && a.planes == 3 && a.leaf_player[item] != 0;
            const uint8_t* tap = s_tap + t * pp;
            // the row of A as a bit mask over K (2 x 64 columns)
            uint64_t km[2] = {0ull, 0ull};
            if (live) {
                for (int ch = 0; ch < a.planes; ++ch) {
                    for (int i = 0; i < pp; ++i) {
                        const int cell = tap[i];
                        if (cell == 0xFF) continue;
                        const uint64_t bit = 1ull << (cell & 63);
                        const int wi = cell >> 6;            // selects, not indexing: keeps the words in registers
                        const uint64_t m0 = wi == 0 ? w0[0] : (wi == 1 ? w0[1] : (wi == 2 ? w0[2] : w0[3]));
                        const uint64_t m1 = wi == 0 ? w1[0] : (wi == 1 ? w1[1] : (wi == 2 ? w1[2] : w1[3]));
                        const bool on = ch == 0 ? (m0 & bit) != 0 : (ch == 1 ? (m1 & bit) != 0 : stm);
                        const int k = ch * pp + i;
                        if (on) km[k >> 6] |= 1ull << (k & 63);
                    }
                }
                const int k1 = a.kc + t;                     // one-hot token index -> pos-emb / bias row
                km[k1 >> 6] |= 1ull << (k1 & 63);
            }
            unsigned char* arow = s_a + (size_t)st * KB * 16384 + (size_t)r * 128;
            for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const uint32_t by = (uint32_t)(km[kb] >> (c * 8)) & 0xFFu;
                    uint32_t pk[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j)              // bf16 1.0 = 0x3F80 per set bit
                        pk[j] = ((by >> (2 * j)) & 1u ? 0x00003F80u : 0u) | ((by >> (2 * j + 1)) & 1u ? 0x3F800000u : 0u);
                    *reinterpret_cast<uint4*>(arow + (size_t)kb * 16384 + ((c ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
            }
            azptx::fence_proxy_async();                      // generic-proxy writes -> visible to the MMA (async proxy)
            azptx::mbar_arrive(&a_full[st]);
        }
    } else {
        // ------------------------------ epilogue: LayerNorm of the accumulator rows -------------
        const int q = warp;                                  // TMEM lane quarter
        const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16);
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            azptx::mbar_wait(acc_full, it & 1);
            azptx::tc_fence_after();
            const long long row = (long long)tile * 128 + q * 32 + lane;
            const bool live = row < rows_total;
            float s = 0.f, ss = 0.f;
            for (int c = 0; c < D / 32; ++c) {
                uint32_t v[32];
                azptx::tmem_ld32(tq + c * 32, v);
                azptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float f = __uint_as_float(v[i]);
                    s += f;
                    ss += f * f;
                }
            }
            const float mean = s / (float)D;
            const float rstd = rsqrtf(fmaxf(ss / (float)D - mean * mean, 0.f) + 1e-5f);
            const bool want_x = live && (a.write_all_x || (row % T) == 0);
            const bool stage = a.tma_out != 0;                   // warp-uniform
            const bool stage_x = stage && a.write_all_x;
            const int row_w0 = tile * 128 + q * 32;
            for (int c = 0; c < D / 32; ++c) {
                uint32_t v[32];
                azptx::tmem_ld32(tq + c * 32, v);
                azptx::tmem_ld_wait();
                uint4 o[4], xr[4];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float f0 = __uint_as_float(v[2 * i]), f1 = __uint_as_float(v[2 * i + 1]);
                    __nv_bfloat162 n2 = __floats2bfloat162_rn((f0 - mean) * rstd, (f1 - mean) * rstd);
                    __nv_bfloat162 x2 = __floats2bfloat162_rn(f0, f1);
                    reinterpret_cast<uint32_t*>(o)[i] = live ? *reinterpret_cast<uint32_t*>(&n2) : 0u;
                    reinterpret_cast<uint32_t*>(xr)[i] = live ? *reinterpret_cast<uint32_t*>(&x2) : 0u;
                }
                unsigned char* stg_n = s_out + ((size_t)q * 2 + ((c >> 1) & 1)) * 2 * 4096;   // pairs of chunks alternate between two buffers
                unsigned char* stg_x = stg_n + 4096;
                if (stage) {
                    if ((c & 1) == 0) {
                        if (lane == 0) azptx::bulk_wait_read1();     // the stores issued two pairs ago have read this buffer
                        __syncwarp();
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint32_t off = (uint32_t)lane * 128u + (((uint32_t)((c & 1) * 4 + i) ^ ((uint32_t)lane & 7u)) << 4);
                        *reinterpret_cast<uint4*>(stg_n + off) = o[i];
                        if (stage_x) *reinterpret_cast<uint4*>(stg_x + off) = xr[i];
                    }
                    if (c & 1) {
                        azptx::fence_proxy_async();
                        __syncwarp();
                        if (lane == 0) {
                            azptx::tma_store_2d(&tm_xn, stg_n, (c - 1) * 32, row_w0);
                            if (stage_x) azptx::tma_store_2d(&tm_x, stg_x, (c - 1) * 32, row_w0);
                            azptx::bulk_commit();
                        }
                    }
                    if (!stage_x && want_x) {
                        uint4* dx = reinterpret_cast<uint4*>(a.x + (size_t)row * D + c * 32);
#pragma unroll
                        for (int i = 0; i < 4; ++i) dx[i] = xr[i];
                    }
                } else if (live) {
                    uint4* dst = reinterpret_cast<uint4*>(a.xn + (size_t)row * D + c * 32);
#pragma unroll
                    for (int i = 0; i < 4; ++i) dst[i] = o[i];
                    if (want_x) {
                        uint4* dx = reinterpret_cast<uint4*>(a.x + (size_t)row * D + c * 32);
#pragma unroll
                        for (int i = 0; i < 4; ++i) dx[i] = xr[i];
                    }
                }
            }
            azptx::tc_fence_before();
            azptx::mbar_arrive(acc_empty);
        }
        if (lane == 0) azptx::bulk_wait0();                     // this warp's TMA stores are complete before the CTA retires
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<TMEM_COLS>(tmem_base);
}
