// All-token multi-head self-attention of the full (non-last) blocks on the tensor cores.
//
// Reference: nn.MultiheadAttention inside ai/nn.py:42,55 (eval mode, no masks): softmax(q k^T / sqrt(hd)) v for every
// token of a board.  The last block only needs the cls query (az_cls_attention_kernel / the fused front end); every
// block before it needs all T queries -- the `main.py:187` net (D = 256, depth 2) and any deeper one.
//
// Work item = (board, 64-column group of D): 64 / hd heads whose q / k / v columns share one 128-byte row of a TMA box.
// Unit = (item, query tile of 128 rows, head).  Per unit two tcgen05 MMAs with fp32 accumulators in TMEM:
//     S[128 q][Tp keys] = Q_h K_h^T              both operands K-major (hd contiguous), K = hd
//     O[128 q][64 cols] = P[128 q][Tp keys] V    P K-major (written by the softmax warps), V MN-major (as TMA wrote it)
// O is computed over the 64 columns of the whole group and only the hd columns of head h are kept: an N = hd MMA would
// need an MN-major operand narrower than the 64-element swizzle atom.  Rows >= T of a tile and key columns >= T are
// padding: padded S columns are masked to probability 0; padded Q rows give rows of S / O nobody reads; padded K / V rows
// hold the NEXT board's tokens (or the zero fill of TMA beyond the tensor), finite values that meet probability 0.
//
//   warps 0-3  softmax (TMEM lane = query row): max, exp2, bf16 P into the UMMA layout, row sum; then the O epilogue of
//              the unit LAG behind: TMEM -> x 1/sum -> bf16 -> global
//   warp 4     one thread: TMA loads of q / k / v (two stages when a board has <= 128 tokens) and both MMAs,
//              S(u+1) issued before O(u) so the tensor pipe works while the softmax warps hold unit u
// TMEM columns: NBUF S accumulators of SC columns, then NBUF O accumulators of 64; NBUF = 2 up to 128 tokens, else 1.
#pragma once
#include "az_ptx.cuh"
#include "az_front2.cuh"

struct AzAttnArgs {
    AzBatch batch;
    bf16* out;          // [boards * T][D]
    int T, D, H, hd;
    int KT;             // 64-row boxes per board: ceil(T / 64), 1..4
    int Tp;             // keys per MMA: T rounded up to 16
    int groups;         // D / 64
    int hpg;            // heads per group: 64 / hd
    int nq;             // query tiles of 128 rows
    int nbuf;           // S / P / O buffers (2: a unit's softmax overlaps the next unit's S MMA)
    int nstg;           // q / k / v stages
    float sl2e;         // log2(e) / sqrt(hd)
};

static inline bool az_attn_supported(int T, int D, int H) {
    const int hd = H > 0 ? D / H : 0;
    return D % 64 == 0 && (hd == 16 || hd == 32 || hd == 64) && hd * H == D && T >= 1 && T <= 256;
}
static inline size_t az_attn_smem(const AzAttnArgs& a) {
    return 1024 + (size_t)a.nstg * 3 * a.KT * 8192 + (size_t)a.nbuf * a.KT * 16384 + 256;
}
static inline int az_attn_tmem_cols(const AzAttnArgs& a) { return a.nbuf * a.KT * 64 + a.nbuf * 64 <= 256 ? 256 : 512; }

namespace azptx {
// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
}  // namespace azptx

__device__ __forceinline__ uint32_t az_attn_pack2(float lo, float hi, float& sum) {
    const __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
    sum += __low2float(p) + __high2float(p);                  // the sum of what the MMA will actually read
    return *reinterpret_cast<const uint32_t*>(&p);
}

template <int TMEM_COLS>
__global__ void __launch_bounds__(160) az_attn_mma_kernel(const __grid_constant__ CUtensorMap tm_qkv, AzAttnArgs a) {
    extern __shared__ unsigned char az_attn_smem_raw[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)az_attn_smem_raw + 1023) & ~(uintptr_t)1023);
    const int KT = a.KT, NBUF = a.nbuf, NSTG = a.nstg, LAG = a.nbuf - 1;
    const size_t stage_bytes = (size_t)3 * KT * 8192, p_bytes = (size_t)KT * 16384;
    unsigned char* s_qkv = base;                                   // [stage][q | k | v][KT * 64 rows][128 B]
    unsigned char* s_p = base + NSTG * stage_bytes;                // [buf][KT atoms][128 rows][128 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_p + NBUF * p_bytes);
    uint64_t *kv_full = bars, *kv_empty = bars + 2, *s_full = bars + 4, *p_ready = bars + 6, *o_full = bars + 8;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        azptx::prefetch_tmap(&tm_qkv);
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&kv_full[i], 1);
            azptx::mbar_init(&kv_empty[i], 1);
            azptx::mbar_init(&s_full[i], 1);
            azptx::mbar_init(&p_ready[i], 128);
            azptx::mbar_init(&o_full[i], 1);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 0) azptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int SC = KT * 64;                                        // columns of one S accumulator
    const uint32_t o_col0 = (uint32_t)(NBUF * SC);

    const int items = az_batch_items(a.batch) * a.groups;          // (board, group) pairs
    const int n_my = items > (int)blockIdx.x ? (items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int upi = a.nq * a.hpg;                                  // units per item
    const int U = n_my * upi;

    if (warp == 4) {
        if (lane == 0) {
            const uint32_t idesc_s = azptx::umma_idesc_bf16(128, (uint32_t)a.Tp);
            const uint32_t idesc_o = azptx::umma_idesc_bf16(128, 64) | (1u << 16);       // B operand (V) MN-major
            const int ksteps_s = a.hd >> 4, ksteps_o = a.Tp >> 4;
            auto load = [&](int i) {                               // q / k / v boxes of the CTA's i-th item
                const int w = (int)blockIdx.x + i * (int)gridDim.x;
                const int b = w / a.groups, g = w - b * a.groups;
                const int st = i % NSTG;
                unsigned char* dst = s_qkv + (size_t)st * stage_bytes;
                azptx::mbar_arrive_expect_tx(&kv_full[st], (uint32_t)stage_bytes);
                for (int sec = 0; sec < 3; ++sec)
                    for (int k = 0; k < KT; ++k)
                        azptx::tma_load_2d(dst + ((size_t)sec * KT + k) * 8192, &tm_qkv, &kv_full[st], sec * a.D + g * 64, b * a.T + k * 64);
            };
            for (int u = 0; u < U + LAG; ++u) {
                if (u < U) {
                    const int i = u / upi, r = u - i * upi, st = i % NSTG;
                    if (r == 0) {
                        if (NSTG == 1) {
                            if (i > 0) azptx::mbar_wait(&kv_empty[0], (uint32_t)((i - 1) & 1));
                            load(i);
                        } else if (i == 0) {
                            load(0);
                        }
                        azptx::mbar_wait(&kv_full[st], (uint32_t)((i / NSTG) & 1));
                    }
                    const int qt = r / a.hpg, j = r - qt * a.hpg, sb = u % NBUF;
                    const uint32_t sq = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes) + (uint32_t)(qt * 16384 + j * a.hd * 2);
                    const uint32_t sk = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes + (size_t)KT * 8192) + (uint32_t)(j * a.hd * 2);
                    const uint64_t dq = azptx::umma_desc_k128(sq), dk = azptx::umma_desc_k128(sk);
                    azptx::tc_fence_after();
                    for (int ks = 0; ks < ksteps_s; ++ks)
                        azptx::umma_bf16(tmem_base + (uint32_t)(sb * SC), dq + 2 * ks, dk + 2 * ks, idesc_s, ks != 0);
                    azptx::umma_commit(&s_full[sb]);
                }
                const int v = u - LAG;
                if (v >= 0) {
                    const int i = v / upi, r = v - i * upi, st = i % NSTG, pb = v % NBUF;
                    azptx::mbar_wait(&p_ready[pb], (uint32_t)((v / NBUF) & 1));
                    azptx::tc_fence_after();
                    const uint32_t sp = azptx::smem_u32(s_p + (size_t)pb * p_bytes);
                    const uint32_t sv = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes + (size_t)2 * KT * 8192);
                    for (int ks = 0; ks < ksteps_o; ++ks) {        // +16 keys: 32 bytes of a P row (next atom every 64), 2048 bytes of V
                        const uint64_t dp = azptx::umma_desc_k128(sp + (uint32_t)((ks >> 2) * 16384 + (ks & 3) * 32));
                        const uint64_t dv = azptx::umma_desc_mn128(sv + (uint32_t)ks * 2048, 8192);
                        azptx::umma_bf16(tmem_base + o_col0 + (uint32_t)(pb * 64), dp, dv, idesc_o, ks != 0);
                    }
                    azptx::umma_commit(&o_full[pb]);
                    if (r == upi - 1) azptx::umma_commit(&kv_empty[st]);
                }
                if (NSTG == 2 && u < U && u % upi == 0) {          // next item's loads, after this iteration's O MMA was issued
                    const int nx = u / upi + 1;
                    if (nx < n_my) {
                        if (nx >= 2) azptx::mbar_wait(&kv_empty[nx % 2], (uint32_t)(((nx >> 1) - 1) & 1));
                        load(nx);
                    }
                }
            }
        }
        __syncwarp();
    } else {
        const int row = threadIdx.x;                               // TMEM lane = query row of the tile
        const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
        const int T = a.T, hd = a.hd;
        const int chunks = (a.Tp + 31) >> 5;
        float inv0 = 0.f, inv1 = 0.f;                               // 1 / row sum of the units in flight (even / odd)
        for (int u = 0; u < U + LAG; ++u) {
            if (u < U) {
                const int r = u % upi, qt = r / a.hpg, sb = u % NBUF;
                azptx::mbar_wait(&s_full[sb], (uint32_t)((u / NBUF) & 1));
                azptx::tc_fence_after();
                if (qt * 128 + warp * 32 < T) {                    // warp-uniform: some row of this warp is a real query
                    const uint32_t ts = lane_addr + (uint32_t)(sb * SC);
                    float mx = -3.0e38f;
                    for (int c = 0; c < chunks; ++c) {
                        uint32_t v[32];
                        azptx::tmem_ld32(ts + c * 32, v);
                        azptx::tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 32; ++e)
                            if (c * 32 + e < T) mx = fmaxf(mx, __uint_as_float(v[e]));
                    }
                    const float off = mx * a.sl2e;
                    float sum = 0.f;
                    unsigned char* prow = s_p + (size_t)sb * p_bytes + (size_t)row * 128;
                    for (int c = 0; c < chunks; ++c) {
                        uint32_t v[32];
                        azptx::tmem_ld32(ts + c * 32, v);
                        azptx::tmem_ld_wait();
                        float e[32];
#pragma unroll
                        for (int k = 0; k < 32; ++k) {
                            float x;
                            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(x) : "f"(fmaf(__uint_as_float(v[k]), a.sl2e, -off)));
                            e[k] = (c * 32 + k < T) ? x : 0.f;
                        }
                        unsigned char* atom = prow + (size_t)(c >> 1) * 16384;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            uint4 w;
                            w.x = az_attn_pack2(e[q * 8 + 0], e[q * 8 + 1], sum);
                            w.y = az_attn_pack2(e[q * 8 + 2], e[q * 8 + 3], sum);
                            w.z = az_attn_pack2(e[q * 8 + 4], e[q * 8 + 5], sum);
                            w.w = az_attn_pack2(e[q * 8 + 6], e[q * 8 + 7], sum);
                            const uint32_t chk = (uint32_t)((c & 1) * 4 + q);
                            if (c * 32 + q * 8 < a.Tp) *reinterpret_cast<uint4*>(atom + ((chk ^ ((uint32_t)row & 7u)) << 4)) = w;
                        }
                    }
                    if (u & 1) inv1 = 1.0f / sum; else inv0 = 1.0f / sum;
                    azptx::fence_proxy_async();                    // generic-proxy writes of P -> visible to the MMA
                }
                azptx::tc_fence_before();
                azptx::mbar_arrive(&p_ready[sb]);
            }
            const int v = u - LAG;
            if (v >= 0) {
                const int i = v / upi, r = v - i * upi, qt = r / a.hpg, j = r - qt * a.hpg, ob = v % NBUF;
                if (qt * 128 + warp * 32 < T) {
                    azptx::mbar_wait(&o_full[ob], (uint32_t)((v / NBUF) & 1));
                    azptx::tc_fence_after();
                    const int w = (int)blockIdx.x + i * (int)gridDim.x;
                    const int b = w / a.groups, g = w - b * a.groups;
                    const int q = qt * 128 + row;
                    const float inv = (v & 1) ? inv1 : inv0;
                    bf16* dst = a.out + ((size_t)b * T + q) * a.D + g * 64 + j * hd;
                    const uint32_t to = lane_addr + o_col0 + (uint32_t)(ob * 64 + j * hd);
                    if (hd == 16) {
                        uint32_t o[16];
                        azptx::tmem_ld16(to, o);
                        azptx::tmem_ld_wait();
                        if (q < T) {
#pragma unroll
                            for (int c = 0; c < 2; ++c) {
                                float f[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(o[c * 8 + e]) * inv;
                                *reinterpret_cast<uint4*>(dst + c * 8) = az_pack8(f);
                            }
                        }
                    } else {
                        for (int h32 = 0; h32 < (hd >> 5); ++h32) {
                            uint32_t o[32];
                            azptx::tmem_ld32(to + h32 * 32, o);
                            azptx::tmem_ld_wait();
                            if (q < T) {
#pragma unroll
                                for (int c = 0; c < 4; ++c) {
                                    float f[8];
#pragma unroll
                                    for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(o[c * 8 + e]) * inv;
                                    *reinterpret_cast<uint4*>(dst + h32 * 32 + c * 8) = az_pack8(f);
                                }
                            }
                        }
                    }
                    azptx::tc_fence_before();
                }
            }
        }
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<TMEM_COLS>(tmem_base);
}
