// All-token multi-head self-attention of the full (non-last) blocks on the tensor cores.
//
// Reference: nn.MultiheadAttention inside ai/nn.py:42,55 (eval mode, no masks): softmax(q k^T / sqrt(hd)) v for every
// token of a board.  The last block only needs the cls query (az_cls_attention_kernel / the fused front end); every
// block before it needs all T queries -- the `main.py:187` net (D = 256, depth 2) and any deeper one.
//
// Work item = (tile of boards, 64-column group of D): 64 / hd heads whose q / k / v columns share one 128-byte row of a
// TMA box.  Boards of up to 64 tokens go TWO to a tile (rows 0-63 / 64-127: every TMEM lane carries a query instead of
// 39 % of them on a 7x7 board); longer boards one to a tile.  Unit = (item, query tile of 128 rows, head).  Per unit two
// tcgen05 MMAs with fp32 accumulators in TMEM:
//     S[128 q][NS keys] = Q_h K_h^T              both operands K-major (hd contiguous), K = hd
//     O[128 q][64 cols] = P[128 q][NS keys] V    P K-major (written by the softmax warps), V MN-major (as TMA wrote it)
// With two boards in a tile S holds the four board x board blocks and the softmax reads the diagonal ones; row r of P holds
// the probabilities over the 64 keys of ITS board, and the P V MMA takes the two boards' V boxes as the two halves of an
// N = 128 operand: the half of an O row that belongs to the row's own board is the answer, the other half is never read.
// (`tcgen05.mma` costs the issuing thread 66+ cycles per K = 16 step whatever N <= 128 is -- tools/umma_rate.cu -- so
// the count of instructions, not their width, is what this layout minimises: hd/16 for S, T/16 for O.)
// O is computed over the 64 columns of the whole group and only the hd columns of head h are kept: an N = hd MMA would
// need an MN-major operand narrower than the 64-element swizzle atom.  Rows >= T of a board and key columns >= T are
// padding: padded S columns are masked to probability 0; padded Q rows give rows of S / O nobody reads; padded K / V rows
// hold the NEXT board's tokens (or the zero fill of TMA beyond the tensor), finite values that meet probability 0.
//
//   warps 0-3 / 4-7  two softmax groups, units alternate between them (TMEM lane = query row): max, exp2, bf16 P into the
//                    UMMA layout, row sum; then the unit's O epilogue: TMEM -> x 1/sum -> bf16 -> global
//   warp 8           one thread: both MMAs, S(u+1) issued before O(u)
//   warp 9           one thread: TMA loads of q / k / v, three stages ahead (two from 65 tokens)
// Boards of more than 128 tokens (3-4 key boxes): one group, one stage (shared memory: P alone is 48-64 KB).
// TMEM columns: NG S accumulators of SC columns, then NG O accumulators of OC (packed 128 + 128, else KT x 64 + 64).
#pragma once
#include "az_ptx.cuh"
#include "az_front2.cuh"

struct AzAttnArgs {
    AzBatch batch;
    bf16* out;          // [boards * T][D]
    int T, D, H, hd;
    int pack;           // 1: two boards per 128-row tile (T <= 64)
    int KT;             // 64-row boxes per q / k / v tile: 2 when packed, else ceil(T / 64) (1..4)
    int Tp;             // keys of a board: T rounded up to 16
    int NS;             // N of the S MMA: 128 when packed, else Tp
    int groups;         // D / 64
    int hpg;            // heads per group: 64 / hd
    int hpg_log2;
    int nq;             // query tiles of 128 rows
    int nbuf;           // softmax groups = S / P / O buffers
    int nstg;           // q / k / v stages
    float sl2e;         // log2(e) / sqrt(hd)
};

static inline bool az_attn_supported(int T, int D, int H) {
    const int hd = H > 0 ? D / H : 0;
    return D % 64 == 0 && (hd == 16 || hd == 32 || hd == 64) && hd * H == D && T >= 1 && T <= 256;
}
// P tile of one softmax group: packed boards share one atom of 64 keys (each row holds its own board's keys)
__host__ __device__ static inline size_t az_attn_p_bytes(const AzAttnArgs& a) { return (size_t)(a.pack ? 1 : a.KT) * 16384; }
static inline size_t az_attn_smem(const AzAttnArgs& a) {
    return 1024 + (size_t)a.nstg * 3 * a.KT * 8192 + (size_t)a.nbuf * az_attn_p_bytes(a) + 256;
}
#define AZ_ATTN_TMEM_COLS 512       // packed: 2 x (128 S + 128 O); else nbuf x (KT x 64 + 64) <= 384: one CTA per SM
#define AZ_ATTN_THREADS 320

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

// -DAZ_ATTN_TRACE: cumulative phase clocks of block 0 (MMA thread and lane 0 of the first warp of each softmax group)
#ifdef AZ_ATTN_TRACE
#define AZ_AT_DECL long long at_t[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long at_c = clock64();
#define AZ_AT(k) { const long long n_ = clock64(); at_t[k] += n_ - at_c; at_c = n_; }
#else
#define AZ_AT_DECL
#define AZ_AT(k)
#endif

__device__ __forceinline__ uint32_t az_attn_bf2(float lo, float hi) {
    const __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&p);
}
// (A __nanosleep back-off in the MMA thread's mbarrier polls, so that it would not take issue slots from the softmax warp on
// its scheduler, was tried and dropped: no gain, mbarrier.try_wait already suspends the thread in hardware.)
template <bool MASK>
__device__ __forceinline__ void az_attn_max(const uint32_t (&v)[32], int valid, float& m0, float& m1) {
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
        if (!MASK || e < valid) m0 = fmaxf(m0, __uint_as_float(v[e]));
        if (!MASK || e + 1 < valid) m1 = fmaxf(m1, __uint_as_float(v[e + 1]));
    }
}

// 32 scores of one row -> probabilities (unnormalised, bf16) in the row's 64 bytes of the P tile.  MASK: the chunk holds
// key columns >= T (probability 0).
template <bool MASK>
__device__ __forceinline__ void az_attn_chunk(const uint32_t (&v)[32], float sl2e, float off, int valid, unsigned char* atom,
                                              uint32_t chk0, uint32_t rsw, int n16, float (&sum)[4]) {
    float e[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) {
        float x;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(x) : "f"(fmaf(__uint_as_float(v[k]), sl2e, -off)));
        e[k] = (!MASK || k < valid) ? x : 0.f;
        sum[k & 3] += e[k];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 w;
        w.x = az_attn_bf2(e[q * 8 + 0], e[q * 8 + 1]);
        w.y = az_attn_bf2(e[q * 8 + 2], e[q * 8 + 3]);
        w.z = az_attn_bf2(e[q * 8 + 4], e[q * 8 + 5]);
        w.w = az_attn_bf2(e[q * 8 + 6], e[q * 8 + 7]);
        if (q < 2 * n16) *reinterpret_cast<uint4*>(atom + (((chk0 + q) ^ rsw) << 4)) = w;    // n16: 16-key steps the P V MMA reads of this chunk
    }
}

__global__ void __launch_bounds__(AZ_ATTN_THREADS) az_attn_mma_kernel(const __grid_constant__ CUtensorMap tm_qkv, AzAttnArgs a) {
    extern __shared__ unsigned char az_attn_smem_raw[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)az_attn_smem_raw + 1023) & ~(uintptr_t)1023);
    const int KT = a.KT, NG = a.nbuf, NSTG = a.nstg, LAG = a.nbuf - 1;
    const size_t stage_bytes = (size_t)3 * KT * 8192, p_bytes = az_attn_p_bytes(a);
    unsigned char* s_qkv = base;                                   // [stage][q | k | v][KT * 64 rows][128 B]
    unsigned char* s_p = base + NSTG * stage_bytes;                // [group][atoms of 64 keys][128 rows][128 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_p + NG * p_bytes);
    uint64_t *kv_full = bars, *kv_empty = bars + 3, *s_full = bars + 6, *p_ready = bars + 8, *o_full = bars + 10;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        azptx::prefetch_tmap(&tm_qkv);
        for (int i = 0; i < 3; ++i) {
            azptx::mbar_init(&kv_full[i], 1);
            azptx::mbar_init(&kv_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&s_full[i], 1);
            azptx::mbar_init(&p_ready[i], 128);
            azptx::mbar_init(&o_full[i], 1);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 0) azptx::tmem_alloc<AZ_ATTN_TMEM_COLS>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int SC = a.pack ? 128 : KT * 64;                         // columns of one S accumulator
    const int OC = a.pack ? 128 : 64;                              // ... of one O accumulator
    const uint32_t o_col0 = (uint32_t)(NG * SC);

    const int boards = az_batch_items(a.batch);
    const int tiles = a.pack ? (boards + 1) >> 1 : boards;
    const int items = tiles * a.groups;                            // (tile, group) pairs
    const int n_my = items > (int)blockIdx.x ? (items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int upi = a.nq << a.hpg_log2;                            // units per item
    const int U = n_my * upi;

    if (warp == 9) {
        // ------------------------------ TMA: q / k / v boxes of the CTA's items, NSTG stages ahead ------------------------
        if (lane == 0) {
            for (int i = 0; i < n_my; ++i) {
                const int w = (int)blockIdx.x + i * (int)gridDim.x;
                const int t = w / a.groups, g = w - t * a.groups;
                const int st = i % NSTG, fill = i / NSTG;
                if (fill > 0) azptx::mbar_wait(&kv_empty[st], (uint32_t)((fill - 1) & 1));
                unsigned char* dst = s_qkv + (size_t)st * stage_bytes;
                azptx::mbar_arrive_expect_tx(&kv_full[st], (uint32_t)stage_bytes);
                for (int sec = 0; sec < 3; ++sec)
                    for (int k = 0; k < KT; ++k)
                        azptx::tma_load_2d(dst + ((size_t)sec * KT + k) * 8192, &tm_qkv, &kv_full[st], sec * a.D + g * 64,
                                           a.pack ? (2 * t + k) * a.T : t * a.T + k * 64);
            }
        }
        __syncwarp();
    } else if (warp == 8) {
        // ------------------------------ MMA issue -------------------------------------------------------------------------
        if (lane == 0) {
            const uint32_t idesc_s = azptx::umma_idesc_bf16(128, (uint32_t)a.NS);
            const uint32_t idesc_o = azptx::umma_idesc_bf16(128, (uint32_t)OC) | (1u << 16);     // B operand (V) MN-major
            const int ksteps_s = a.hd >> 4, ksteps_o = a.Tp >> 4;
            AZ_AT_DECL
            int i1 = 0, r1 = 0;                                    // item / unit-in-item of the S MMA being issued
            int i2 = 0, r2 = 0;                                    // ... of the O MMA
            for (int u = 0; u < U + LAG; ++u) {
                if (u < U) {
                    const int st = i1 % NSTG;
                    if (r1 == 0) {
                        AZ_AT(0)
                        azptx::mbar_wait(&kv_full[st], (uint32_t)((i1 / NSTG) & 1));
                        AZ_AT(1)
                    }
                    const int qt = r1 >> a.hpg_log2, j = r1 & (a.hpg - 1), sb = u % NG;
                    const uint32_t sq = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes) + (uint32_t)(qt * 16384 + j * a.hd * 2);
                    const uint32_t sk = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes + (size_t)KT * 8192) + (uint32_t)(j * a.hd * 2);
                    const uint64_t dq = azptx::umma_desc_k128(sq), dk = azptx::umma_desc_k128(sk);
                    azptx::tc_fence_after();
                    for (int ks = 0; ks < ksteps_s; ++ks)
                        azptx::umma_bf16(tmem_base + (uint32_t)(sb * SC), dq + 2 * ks, dk + 2 * ks, idesc_s, ks != 0);
                    azptx::umma_commit(&s_full[sb]);
                    if (++r1 == upi) { r1 = 0; ++i1; }
                    AZ_AT(2)
                }
                if (u >= LAG) {
                    const int v = u - LAG, st = i2 % NSTG, pb = v % NG;
                    azptx::mbar_wait(&p_ready[pb], (uint32_t)((v / NG) & 1));
                    AZ_AT(3)
                    azptx::tc_fence_after();
                    const uint32_t sp = azptx::smem_u32(s_p + (size_t)pb * p_bytes);
                    const uint32_t sv = azptx::smem_u32(s_qkv + (size_t)st * stage_bytes + (size_t)2 * KT * 8192);
                    // +16 keys: 32 bytes of a P row (next atom every 64 keys), 2048 bytes of V.  Packed: the two boards' V boxes
                    // are the two 64-column halves of an N = 128 operand (LBO = 8192 bytes apart); row r of P holds the
                    // probabilities of ITS board, and only the matching half of its O row is read.
                    for (int ks = 0; ks < ksteps_o; ++ks) {
                        const uint64_t dp = azptx::umma_desc_k128(sp + (uint32_t)((ks >> 2) * 16384 + (ks & 3) * 32));
                        const uint64_t dv = azptx::umma_desc_mn128(sv + (uint32_t)ks * 2048, 8192);
                        azptx::umma_bf16(tmem_base + o_col0 + (uint32_t)(pb * OC), dp, dv, idesc_o, ks != 0);
                    }
                    azptx::umma_commit(&o_full[pb]);
                    if (r2 == upi - 1) azptx::umma_commit(&kv_empty[st]);
                    if (++r2 == upi) { r2 = 0; ++i2; }
                    AZ_AT(4)
                }
            }
#ifdef AZ_ATTN_TRACE
            if (blockIdx.x == 0) printf("attn mma thread: units %d | kv_full wait %lld  S mma issue %lld  p_ready wait %lld  O mma issue %lld (loop %lld)\n",
                                        U, at_t[1], at_t[2], at_t[3], at_t[4], at_t[0]);
#endif
        }
        __syncwarp();
    } else if ((warp >> 2) < NG) {
        const int grp = warp >> 2, w4 = warp & 3;
        const int row = threadIdx.x & 127;                         // TMEM lane = query row of the tile
        const uint32_t lane_addr = tmem_base + ((uint32_t)(w4 * 32) << 16);
        const int T = a.T, hd = a.hd;
        const int blk = a.pack ? row >> 6 : 0;                     // packed: which board of the tile
        const int chunks = (a.Tp + 31) >> 5;                       // 32-key chunks of a board
        const uint32_t ts = lane_addr + (uint32_t)(grp * SC + blk * 64);           // packed: the board's own 64 x 64 block of S
        unsigned char* prow = s_p + (size_t)grp * p_bytes + (size_t)row * 128;
        const uint32_t rsw = (uint32_t)row & 7u;
        const int wrow0 = a.pack ? (w4 & 1) * 32 : w4 * 32;        // first query row of this warp inside its board / tile
        AZ_AT_DECL
        int i = 0, r = grp;                                        // item / unit-in-item of unit u
        while (r >= upi) { r -= upi; ++i; }
        for (int u = grp, n = 0; u < U; u += NG, ++n) {
            const int qt = r >> a.hpg_log2, j = r & (a.hpg - 1);
            const bool warp_on = qt * 128 + wrow0 < T;             // warp-uniform: some row of this warp is a real query
            AZ_AT(0)
            azptx::mbar_wait(&s_full[grp], (uint32_t)(n & 1));
            azptx::tc_fence_after();
            AZ_AT(1)
            float inv = 0.f;
            if (warp_on) {
                float m0 = -3.0e38f, m1 = -3.0e38f;
                float sum[4] = {0.f, 0.f, 0.f, 0.f};
                if (chunks <= 2) {                                 // up to 64 keys: the row stays in registers between the two passes
                    uint32_t v0[32], v1[32];
                    azptx::tmem_ld32(ts, v0);
                    if (chunks == 2) azptx::tmem_ld32(ts + 32, v1);
                    azptx::tmem_ld_wait();
                    if (T >= 32) az_attn_max<false>(v0, 32, m0, m1); else az_attn_max<true>(v0, T, m0, m1);
                    if (chunks == 2) {
                        if (T >= 64) az_attn_max<false>(v1, 32, m0, m1); else az_attn_max<true>(v1, T - 32, m0, m1);
                    }
                    const float off = fmaxf(m0, m1) * a.sl2e;
                    const int n16a = min(2, a.Tp >> 4), n16b = (a.Tp - 32) >> 4;
                    if (T >= 32) az_attn_chunk<false>(v0, a.sl2e, off, 32, prow, 0u, rsw, n16a, sum);
                    else az_attn_chunk<true>(v0, a.sl2e, off, T, prow, 0u, rsw, n16a, sum);
                    if (chunks == 2) {
                        if (T >= 64) az_attn_chunk<false>(v1, a.sl2e, off, 32, prow, 4u, rsw, n16b, sum);
                        else az_attn_chunk<true>(v1, a.sl2e, off, T - 32, prow, 4u, rsw, n16b, sum);
                    }
                } else {
                    for (int c = 0; c < chunks; ++c) {
                        uint32_t v[32];
                        azptx::tmem_ld32(ts + c * 32, v);
                        azptx::tmem_ld_wait();
                        const int valid = T - c * 32;
                        if (valid >= 32) az_attn_max<false>(v, 32, m0, m1); else az_attn_max<true>(v, valid, m0, m1);
                    }
                    const float off = fmaxf(m0, m1) * a.sl2e;
                    for (int c = 0; c < chunks; ++c) {
                        uint32_t v[32];
                        azptx::tmem_ld32(ts + c * 32, v);
                        azptx::tmem_ld_wait();
                        const int valid = T - c * 32;
                        const int n16 = min(2, (a.Tp - c * 32) >> 4);
                        unsigned char* atom = prow + (size_t)(c >> 1) * 16384;
                        if (valid >= 32) az_attn_chunk<false>(v, a.sl2e, off, valid, atom, (uint32_t)((c & 1) * 4), rsw, n16, sum);
                        else az_attn_chunk<true>(v, a.sl2e, off, valid, atom, (uint32_t)((c & 1) * 4), rsw, n16, sum);
                    }
                }
                inv = 1.0f / ((sum[0] + sum[1]) + (sum[2] + sum[3]));
                azptx::fence_proxy_async();                        // generic-proxy writes of P -> visible to the MMA
            }
            AZ_AT(2)
            azptx::tc_fence_before();
            azptx::mbar_arrive(&p_ready[grp]);
            AZ_AT(3)
            if (warp_on) {
                azptx::mbar_wait(&o_full[grp], (uint32_t)(n & 1));
                azptx::tc_fence_after();
                AZ_AT(4)
                const int w = (int)blockIdx.x + i * (int)gridDim.x;
                const int t = w / a.groups, g = w - t * a.groups;
                const int b = a.pack ? 2 * t + blk : t;
                const int q = a.pack ? (row & 63) : qt * 128 + row;
                const bool on = q < T && b < boards;
                bf16* dst = a.out + ((size_t)b * T + q) * a.D + g * 64 + j * hd;
                const uint32_t to = lane_addr + o_col0 + (uint32_t)(grp * OC + blk * 64 + j * hd);
                if (hd == 16) {
                    uint32_t o[16];
                    azptx::tmem_ld16(to, o);
                    azptx::tmem_ld_wait();
                    if (on) {
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
                        if (on) {
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
                AZ_AT(5)
            }
            r += NG;
            while (r >= upi) { r -= upi; ++i; }
        }
#ifdef AZ_ATTN_TRACE
        if (blockIdx.x == 0 && (threadIdx.x & 127) == 0)
            printf("attn softmax group %d: loop %lld  s_full wait %lld  softmax %lld  fence+arrive %lld  o_full wait %lld  epilogue %lld\n", grp, at_t[0], at_t[1], at_t[2],
                   at_t[3], at_t[4], at_t[5]);
#endif
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<AZ_ATTN_TMEM_COLS>(tmem_base);
}
