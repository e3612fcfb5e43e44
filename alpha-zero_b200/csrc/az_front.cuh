// Fused front end of the depth-1 policy/value net on sm_100a: leaf bitboards in, the cls
// token's attention output out -- embedding, LayerNorm 1, V / score projection and the
// cls-query attention of ai/nn.py:9-36,52-56 in ONE kernel, nothing per token leaves the SM.
//
// Algebra (exact in real arithmetic; checked in fp64 by tests/test_fold_algebra.py):
//   token row       x_t  = Wemb^T a_t          a_t in {0,1}^K = [im2col taps | one-hot(t)]
//   centred         x_t - mean(x_t) 1 = Wc^T a_t,   Wc = Wemb with every K-row centred over D
//   LayerNorm 1     xn_t = r_t (x_t - mean) (.) gamma + beta,   r_t = rsqrt(a_t^T G a_t / D + eps),
//                   G = Wc Wc^T = L L^T  (K x K Gram matrix, Cholesky factor L):  a^T G a = |L^T a|^2
//   V projection    v_t  = r_t (Wv' Wc^T) a_t + bv'        (gamma/beta folded into Wv', bv')
//   cls scores      s_t,h = r_t (Ws Wc^T)_h a_t (+ const)  (constant cls query folded into Ws)
//   attention       o_h = sum_t softmax_t(s_.,h) v_t,h
// so per token only ONE contraction over K <= 128 binary inputs is needed.  Two tcgen05 MMAs per
// tile of `nb` boards read the same bit-built operand tile A [tokens][K] (bf16 0/1, exact):
//   stats : D1[token][j]   = A  x Wst^T   (M = 128 token rows, N = NS: 8 score columns + K columns of L^T a)
//   values: D2[feature][t] = Wv2 x A^T    (M = 128 features per block, N = NTP token columns)
// D1 has tokens on TMEM lanes (the sum of squares |L^T a|^2 is thread-local); D2 has FEATURES on
// lanes, so the softmax-weighted sum over a board's tokens is thread-local as well.
//   warps 0-3 statistic + softmax | warps 4-11 value epilogue | warps 12-14 A-tile builders | warp 15 TMA + MMA issue
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "az_net_kernels.cuh"
#include "az_ptx.cuh"

struct AzFrontArgs {
    AzBatch batch;
    const uint64_t* leaf_words;   // [items][2]: mine, theirs (boards of <= 64 cells)
    const uint8_t* leaf_player;   // side to move (plane 2 of the 3-plane games)
    int rows, cols, planes, patch, T, kc, D, H;
    int nb;                       // boards per tile
    int TS;                       // token rows per board inside a tile (T rounded up to 8)
    int NTP;                      // token rows per tile = nb*TS rounded up to 16 (<= 128)
    int NS;                       // stat columns: 8 + kc + T rounded up to 32 (<= 128)
    int KB;                       // K blocks of 64 bf16 (1 or 2)
    int ksteps;                   // K = 16 MMA steps that carry non-zero columns
    int nslots;                   // TMEM accumulator slots for the value blocks
    const float* bias_v;          // [D]
    bf16* oc;                     // [cap][D] (chunk-local rows) attention output of the cls query (before out_proj)
    // az_front2.cuh (reassociated value path): U[board][head][128] instead of oc
    int NP;                       // (board, head) pairs per tile = nb*H rounded up to 16
    bf16* uc;                     // [cap][H*128]
};

static inline size_t az_front_smem(const AzFrontArgs& a) {
    const size_t tile = (size_t)a.NTP * 128;
    return 1024 + 2 * (size_t)a.KB * tile + (size_t)a.KB * a.NS * 128 + (size_t)a.KB * a.D * 128 +
           2 * 8 * 128 * 4 + 2 * 128 * 4 + 128 + 256;
}

namespace azptx {
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
// After tcgen05.wait::ld: ties the destination registers of an earlier tcgen05.ld to this point of the
// volatile-asm order, so the compiler cannot schedule their consumers above the wait.
__device__ __forceinline__ void tmem_regs_ready(uint32_t (&r)[8]) {
    asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]));
}
__device__ __forceinline__ void bar_sync_named(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
}  // namespace azptx

#ifdef AZ_FRONT_TRACE
__device__ unsigned long long g_front_trace[4][64][8];   // [role: mma, builder, S, V][tile iteration][event]
#define AZ_TR(role, ev) do { if (blockIdx.x == 0 && it < 64 && (threadIdx.x & 127) == 0) g_front_trace[role][it][ev] = clock64(); } while (0)
#define AZ_TR1(role, ev) do { if (blockIdx.x == 0 && it < 64 && lane == 0) g_front_trace[role][it][ev] = clock64(); } while (0)
#else
#define AZ_TR(role, ev)
#define AZ_TR1(role, ev)
#endif

// Threads (16 warps):
//   warps 0-3    S group: per-row statistic + scores from the stats accumulator, softmax -> weights in smem
//   warps 4-11   V group: two warps per TMEM lane quarter, each owning every other value block
//   warps 12-13  A-tile builders (64 threads, token rows r and r + 64; next tile's boards prefetched)
//   warp  15     TMA (weights, once) + MMA issue: stats and the even value blocks
//   warp  14     MMA issue: the odd value blocks (one thread cannot issue tcgen05.mma fast enough alone)
// The S group works on tile i + 1 while the V group consumes the weights of tile i (double-buffered).
#define AZ_FRONT_THREADS 512
#define AZ_FRONT_BUILDERS 64

template <int TS8>
__global__ void __launch_bounds__(AZ_FRONT_THREADS, 1) az_front_kernel(const __grid_constant__ CUtensorMap tm_wst,
                                                                       const __grid_constant__ CUtensorMap tm_wv, AzFrontArgs a) {
    extern __shared__ unsigned char smem_dyn[];
    // pointer arithmetic (not an integer round trip) keeps the shared address space: LDS/STS, not generic
    unsigned char* smem = smem_dyn + ((1024u - (azptx::smem_u32(smem_dyn) & 1023u)) & 1023u);
    const int KB = a.KB, D = a.D, T = a.T, P = a.patch, NTP = a.NTP, NS = a.NS;
    constexpr int TS = TS8 * 8;
    const int FB = D >> 7;                                       // feature blocks of 128
    const uint32_t tile_bytes = (uint32_t)NTP * 128;
    unsigned char* s_a = smem;                                   // [2 stages][KB][NTP rows][128 B]
    unsigned char* s_wst = s_a + 2 * (size_t)KB * tile_bytes;    // [KB][NS rows][128 B]
    unsigned char* s_wv = s_wst + (size_t)KB * NS * 128;         // [FB][KB][128 rows][128 B]
    float* s_w = reinterpret_cast<float*>(s_wv + (size_t)FB * KB * 16384);   // [2][8 heads][128 rows] scores -> softmax weight * r
    float* s_q = s_w + 2 * 8 * 128;                              // [2][128] sum of squares per row
    uint2* s_lut = reinterpret_cast<uint2*>(s_q + 2 * 128);      // 16 entries: nibble -> 4 bf16 (0 / 1.0)
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_lut + 16);
    uint64_t* w_full = bars;            // 1
    uint64_t* a_full = bars + 1;        // 2   (64 builder arrivals)
    uint64_t* a_empty = bars + 3;       // 2   (one commit per MMA warp: every MMA reading the stage is done)
    uint64_t* st_full = bars + 5;       // 1
    uint64_t* st_empty = bars + 6;      // 1   (128 S-group arrivals)
    uint64_t* v_full = bars + 7;        // 4
    uint64_t* v_empty = bars + 11;      // 4   (128 arrivals of the V sub-group that owns the block)
    uint64_t* sw_full = bars + 15;      // 2   (128 S-group arrivals: weights of a tile are in s_w[buf])
    uint64_t* sw_empty = bars + 17;     // 2   (256 V-group arrivals: s_w[buf] may be overwritten)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // The batch size lives in device memory and is written by the predecessor (the forest's advance
    // kernel): each role runs its own prologue, then waits for the predecessor grid and reads it.
    int items = 0, tiles = 0;
    auto wait_upstream = [&]() {
        azptx::griddep_wait();
        items = az_batch_items(a.batch);
        tiles = (items + a.nb - 1) / a.nb;
    };
    const int nslots = a.nslots;
    // Who issues the statistics MMAs: the lead value-issuing warp when every feature block owns a TMEM slot;
    // builder warp 12 when slots are scarce (then the value issuers block on slots for long stretches and the S
    // group, which paces the kernel, would wait behind them for its accumulator).
    const bool bstats = a.nslots < (a.D >> 7);

    if (threadIdx.x < 16) {
        const uint32_t n = threadIdx.x;
        s_lut[n] = make_uint2(((n & 1u) ? 0x00003F80u : 0u) | ((n & 2u) ? 0x3F800000u : 0u),
                              ((n & 4u) ? 0x00003F80u : 0u) | ((n & 8u) ? 0x3F800000u : 0u));
    }
    if (warp == 15 && lane == 0) {
        azptx::prefetch_tmap(&tm_wst);
        azptx::prefetch_tmap(&tm_wv);
        azptx::mbar_init(w_full, 1);
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&a_full[i], AZ_FRONT_BUILDERS);
            azptx::mbar_init(&a_empty[i], a.nslots < (a.D >> 7) ? 3 : 2);   // value-issuing warps (+ the statistics issuer, see below)
            azptx::mbar_init(&sw_full[i], 128);
            azptx::mbar_init(&sw_empty[i], 256);
        }
        azptx::mbar_init(st_full, 1);
        azptx::mbar_init(st_empty, 128);
        for (int i = 0; i < 4; ++i) {
            azptx::mbar_init(&v_full[i], 1);
            azptx::mbar_init(&v_empty[i], 128);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 0) azptx::tmem_alloc<512>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // TMEM columns: [0, 128) stats accumulator, value slots of NTP columns from 128 on.
    azptx::griddep_launch();                                         // the successor's CTAs may queue behind this grid

    if (warp >= 14) {
        // ------------------------------ weights in once (warp 15), then MMA issue -----------------------
        const bool lead = warp == 15;                                // stats + even value blocks; warp 14: odd value blocks
        if (lead && lane == 0) {
            azptx::mbar_arrive_expect_tx(w_full, (uint32_t)(KB * NS * 128 + FB * KB * 16384));
            for (int kb = 0; kb < KB; ++kb) azptx::tma_load_2d(s_wst + (size_t)kb * NS * 128, &tm_wst, w_full, kb * 64, 0);
            for (int fb = 0; fb < FB; ++fb)
                for (int kb = 0; kb < KB; ++kb)
                    azptx::tma_load_2d(s_wv + ((size_t)fb * KB + kb) * 16384, &tm_wv, w_full, kb * 64, fb * 128);
        }
        const uint32_t idesc_st = azptx::umma_idesc_bf16(128, (uint32_t)NS);
        const uint32_t idesc_v = azptx::umma_idesc_bf16(128, (uint32_t)NTP);
        // The issuing thread is the pacemaker of the tensor pipe: every descriptor is formed once, outside
        // the tile loop, and the K steps are fully unrolled and unpredicated (columns beyond the used K are
        // zero in both operands) -- nothing but tcgen05.mma in the hot sequence.
        uint64_t d_wst[2], d_a[2][2], d_wv[4][2];
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
            const int kbe = kb < KB ? kb : 0;
            d_wst[kb] = azptx::umma_desc_k128(azptx::smem_u32(s_wst + (size_t)kbe * NS * 128));
#pragma unroll
            for (int st = 0; st < 2; ++st)
                d_a[st][kb] = azptx::umma_desc_k128(azptx::smem_u32(s_a + (size_t)st * KB * tile_bytes + (size_t)kbe * tile_bytes));
#pragma unroll
            for (int fb = 0; fb < 4; ++fb)
                d_wv[fb][kb] = azptx::umma_desc_k128(azptx::smem_u32(s_wv + ((size_t)(fb < FB ? fb : 0) * KB + kbe) * 16384));
        }
        const int mine = lead ? 0 : 1;                               // parity of the value blocks this warp issues
        int it = 0, slot = 0, round = 0;                           // value-block counter g = round * nslots + slot
        azptx::mbar_wait(w_full, 0);
        wait_upstream();
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int st = it & 1;
            AZ_TR1(0, 0);
            azptx::mbar_wait(&a_full[st], (it >> 1) & 1);
            AZ_TR1(0, 1);
            const uint64_t da0 = st ? d_a[1][0] : d_a[0][0], da1 = st ? d_a[1][1] : d_a[0][1];
            if (lead && !bstats) {
                azptx::mbar_wait(st_empty, (it & 1) ^ 1);
                AZ_TR1(0, 2);
                azptx::tc_fence_after();
                if (azptx::elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        azptx::umma_bf16(tmem_base, da0 + 2 * ks, d_wst[0] + 2 * ks, idesc_st, ks != 0);
                    if (KB == 2) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(tmem_base, da1 + 2 * ks, d_wst[1] + 2 * ks, idesc_st, 1);
                    }
                    azptx::umma_commit(st_full);
                }
                __syncwarp();
            }
#pragma unroll
            for (int fb = 0; fb < 4; ++fb) {
                if (fb < FB) {
                    if ((fb & 1) == mine) {
                        azptx::mbar_wait(&v_empty[slot], (round & 1) ^ 1);
                        if (fb == 0) AZ_TR1(0, 3);
                        if (fb == 2) AZ_TR1(0, 4);
                        azptx::tc_fence_after();
                        if (azptx::elect_one()) {
                            const uint32_t dcol = tmem_base + 128 + slot * NTP;
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(dcol, d_wv[fb][0] + 2 * ks, da0 + 2 * ks, idesc_v, ks != 0);
                            if (KB == 2) {
#pragma unroll
                                for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(dcol, d_wv[fb][1] + 2 * ks, da1 + 2 * ks, idesc_v, 1);
                            }
                            azptx::umma_commit(&v_full[slot]);
                        }
                        __syncwarp();
                    }
                    if (++slot == nslots) { slot = 0; ++round; }
                }
            }
            if (azptx::elect_one()) azptx::umma_commit(&a_empty[st]);    // this warp's MMAs on the A stage are done
            __syncwarp();
            AZ_TR1(0, 5);
        }
    } else if (warp >= 12) {
        // ------------------------------ A-tile builders: token rows bt and bt + 64 ---------------------
        const int bt = threadIdx.x - 384;
        const int half = P >> 1, pp = P * P;
        const uint32_t colmask = (1u << a.cols) - 1u, pmask = (1u << P) - 1u;
        // geometry of this thread's (at most two) rows does not depend on the board: hoisted out of the tile loop
        int r_b[2], r_t[2], r_rs[2], r_ls[2], r_cr[2];
        uint32_t r_inb[2];
        bool r_on[2];
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {
            const int r = bt + pass * AZ_FRONT_BUILDERS;
            const int bb = r / TS, t = r - bb * TS;
            r_b[pass] = bb;
            r_t[pass] = t;
            r_on[pass] = r < NTP && bb < a.nb && t < T;
            int cr = 0, cc = 0;
            if (t > 0) { cr = (t - 1) / a.cols; cc = (t - 1) - cr * a.cols; }
            r_cr[pass] = cr;
            r_rs[pass] = cc >= half ? cc - half : 0;
            r_ls[pass] = cc >= half ? 0 : half - cc;
            uint32_t inb = 0;                                        // in-bounds taps of the window (plane 2 is all ones)
            for (int kr = 0; kr < P; ++kr) {
                const int rr = cr + kr - half;
                if (rr >= 0 && rr < a.rows) inb |= (((colmask >> r_rs[pass]) << r_ls[pass]) & pmask) << (kr * P);
            }
            r_inb[pass] = (r_on[pass] && t > 0) ? inb : 0u;
        }
        // per (row, window row): where the board row starts in the bitboard and whether it exists; window rows
        // beyond the patch or outside the board get an empty column mask, so the per-tile encode below has no
        // branches and its 20 extractions are independent of each other
        int g_sh[2][5];
        uint32_t g_cm[2][5];
        int g_ks[5];
#pragma unroll
        for (int kr = 0; kr < 5; ++kr) {
            g_ks[kr] = kr < P ? kr * P : 0;
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                const int rr = r_cr[pass] + kr - half;
                const bool ok = kr < P && rr >= 0 && rr < a.rows && r_on[pass] && r_t[pass] > 0;
                g_sh[pass][kr] = ok ? rr * a.cols : 0;
                g_cm[pass][kr] = ok ? colmask : 0u;
            }
        }
        // boards of the tile (side bitboards + side to move), fetched one tile ahead
        // Only the loads are issued here; nothing consumes the loaded registers until the next tile is
        // encoded (a compare right after the load would stall the warp for the whole memory latency).
        ulonglong2 w_cur[2], w_nxt[2];
        uint32_t s_cur[2], s_nxt[2];                                 // raw side-to-move byte, or 0xFFFFFFFF = row unused / board beyond the batch
        auto fetch = [&](int tile, ulonglong2 (&w)[2], uint32_t (&sm)[2]) {
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                const int item = tile * a.nb + r_b[pass];
                w[pass] = make_ulonglong2(0ull, 0ull);
                sm[pass] = 0xFFFFFFFFu;
                if (tile < tiles && r_on[pass] && item < items) {
                    w[pass] = *reinterpret_cast<const ulonglong2*>(a.leaf_words + (size_t)(a.batch.offset + item) * 2);
                    sm[pass] = a.leaf_player[a.batch.offset + item];
                }
            }
        };
        // The one-hot columns of a row never change and pad rows stay zero: both stages are initialised once;
        // per tile only the chunks (8 K columns = 16 bytes) that hold tap columns are rewritten.
        const int nch = (a.kc + 7) >> 3;
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {
            const int r = bt + pass * AZ_FRONT_BUILDERS;
            if (r >= NTP) continue;
            const unsigned __int128 km = r_on[pass] ? (unsigned __int128)1 << (a.kc + r_t[pass]) : (unsigned __int128)0;
            for (int st = 0; st < 2; ++st)
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t kw = kb == 0 ? (uint64_t)km : (uint64_t)(km >> 64);
                    unsigned char* arow = s_a + (size_t)st * KB * tile_bytes + (size_t)kb * tile_bytes + (size_t)r * 128;
                    for (int c = 0; c < 8; ++c) {
                        const uint32_t by = (uint32_t)(kw >> (c * 8)) & 0xFFu;
                        const uint2 lo = s_lut[by & 15u], hi = s_lut[by >> 4];
                        *reinterpret_cast<uint4*>(arow + ((c ^ (r & 7)) << 4)) = make_uint4(lo.x, lo.y, hi.x, hi.y);
                    }
                }
        }
        const bool stats_issuer = bstats && warp == 12;
        const uint32_t idesc_st = azptx::umma_idesc_bf16(128, (uint32_t)NS);
        uint64_t d_wst[2], d_as[2][2];
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
            const int kbe = kb < KB ? kb : 0;
            d_wst[kb] = azptx::umma_desc_k128(azptx::smem_u32(s_wst + (size_t)kbe * NS * 128));
#pragma unroll
            for (int st = 0; st < 2; ++st)
                d_as[st][kb] = azptx::umma_desc_k128(azptx::smem_u32(s_a + (size_t)st * KB * tile_bytes + (size_t)kbe * tile_bytes));
        }
        if (stats_issuer) azptx::mbar_wait(w_full, 0);               // statistic weights are in shared memory
        wait_upstream();                                             // boards and batch size come from the predecessor
        fetch(blockIdx.x, w_cur, s_cur);
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int st = it & 1;
            fetch(tile + gridDim.x, w_nxt, s_nxt);                   // in flight while this tile is encoded
            // K columns of a row as two 64-bit words: taps of plane 0 at bit 0, plane 1 at bit pp, plane 2
            // (in-bounds taps when side 1 is to move) at bit 2*pp, one-hot token index at bit kc + t
            uint64_t kw0[2], kw1[2];
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                const int ob = a.kc + r_t[pass];
                uint64_t k0 = ob < 64 ? 1ull << ob : 0ull, k1 = ob >= 64 ? 1ull << (ob - 64) : 0ull;
                {
                    // (boards beyond the batch were fetched as empty bitboards: their window stays empty)
                    uint32_t p0 = 0, p1 = 0;                         // taps of planes 0 / 1, bit kr*P + kcol  (P*P <= 32)
#pragma unroll
                    for (int kr = 0; kr < 5; ++kr) {
                        const uint32_t b0 = (uint32_t)(w_cur[pass].x >> g_sh[pass][kr]) & g_cm[pass][kr];
                        const uint32_t b1 = (uint32_t)(w_cur[pass].y >> g_sh[pass][kr]) & g_cm[pass][kr];
                        p0 |= (((b0 >> r_rs[pass]) << r_ls[pass]) & pmask) << g_ks[kr];
                        p1 |= (((b1 >> r_rs[pass]) << r_ls[pass]) & pmask) << g_ks[kr];
                    }
                    k0 |= (uint64_t)p0 | ((uint64_t)p1 << pp);
                    if (a.planes == 3 && s_cur[pass] != 0u && s_cur[pass] != 0xFFFFFFFFu) {   // side 1 to move; the taps may straddle bit 64
                        const uint64_t inb = r_inb[pass];
                        k0 |= inb << (2 * pp);
                        if (3 * pp > 64) k1 |= inb >> (64 - 2 * pp);
                    }
                }
                kw0[pass] = k0;
                kw1[pass] = k1;
            }
            AZ_TR(1, 0);
            azptx::mbar_wait(&a_empty[st], ((it >> 1) & 1) ^ 1);
            AZ_TR(1, 1);
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                const int r = bt + pass * AZ_FRONT_BUILDERS;
                if (r >= NTP || !r_on[pass] || r_t[pass] == 0) continue;     // pad rows and the cls row never change
                unsigned char* arow = s_a + (size_t)st * KB * tile_bytes + (size_t)r * 128;
                const int sw = r & 7;
                const int n0 = nch < 8 ? nch : 8;
#pragma unroll 4
                for (int c = 0; c < n0; ++c) {
                    const uint32_t by = (uint32_t)(kw0[pass] >> (c * 8)) & 0xFFu;
                    const uint2 lo = s_lut[by & 15u], hi = s_lut[by >> 4];
                    *reinterpret_cast<uint4*>(arow + ((c ^ sw) << 4)) = make_uint4(lo.x, lo.y, hi.x, hi.y);
                }
#pragma unroll 2
                for (int c = 0; c < nch - 8; ++c) {
                    const uint32_t by = (uint32_t)(kw1[pass] >> (c * 8)) & 0xFFu;
                    const uint2 lo = s_lut[by & 15u], hi = s_lut[by >> 4];
                    *reinterpret_cast<uint4*>(arow + (size_t)tile_bytes + ((c ^ sw) << 4)) = make_uint4(lo.x, lo.y, hi.x, hi.y);
                }
            }
            azptx::fence_proxy_async();                              // generic-proxy writes -> visible to the MMA
            azptx::mbar_arrive(&a_full[st]);
            AZ_TR(1, 2);
            if (stats_issuer) {
                azptx::mbar_wait(&a_full[st], (it >> 1) & 1);        // every builder thread has written its rows
                azptx::mbar_wait(st_empty, (it & 1) ^ 1);            // the S group has read the previous tile's statistics
                azptx::tc_fence_after();
                if (azptx::elect_one()) {
                    const uint64_t da0 = st ? d_as[1][0] : d_as[0][0], da1 = st ? d_as[1][1] : d_as[0][1];
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(tmem_base, da0 + 2 * ks, d_wst[0] + 2 * ks, idesc_st, ks != 0);
                    if (KB == 2) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(tmem_base, da1 + 2 * ks, d_wst[1] + 2 * ks, idesc_st, 1);
                    }
                    azptx::umma_commit(st_full);
                    azptx::umma_commit(&a_empty[st]);                // this thread's reads of the A stage are done
                }
                __syncwarp();
            }
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) { w_cur[pass] = w_nxt[pass]; s_cur[pass] = s_nxt[pass]; }
        }
    } else if (warp < 4) {
        // ------------------------------ S group: statistic, scores, softmax ----------------------------
        const int tid = threadIdx.x;                                 // token row
        const uint32_t tq = tmem_base + ((uint32_t)(warp * 32) << 16);
        const int H = a.H;
        const int pairs = a.nb * H;
        int tpp = 32;
        while (tpp * pairs > 128) tpp >>= 1;                         // threads per (board, head) softmax
        const int pair = tid / tpp, sub = tid - pair * tpp;
        const bool pair_on = pair < pairs;
        const int pb = pair_on ? pair / H : 0, ph = pair_on ? pair - pb * H : 0;
        const float inv_d = 1.0f / (float)D;
        const int ne = (TS + tpp - 1) / tpp;                         // tokens per softmax thread (<= 8, checked on the host)
        wait_upstream();
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            float* swb = s_w + buf * 8 * 128;
            float* sqb = s_q + buf * 128;
            AZ_TR(2, 0);
            azptx::mbar_wait(&sw_empty[buf], ((it >> 1) & 1) ^ 1);    // V group finished with this buffer
            AZ_TR(2, 1);
            // ---- |L^T a|^2 = a^T G a and raw scores from the stats accumulator
            azptx::mbar_wait(st_full, it & 1);
            AZ_TR(2, 2);
            azptx::tc_fence_after();
            float q0 = 0.f, q1 = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c * 32 < NS) {
                    uint32_t v[32];
                    azptx::tmem_ld32(tq + c * 32, v);
                    azptx::tmem_ld_wait();
                    if (c == 0) {
#pragma unroll
                        for (int h = 0; h < 8; ++h) {
                            swb[h * 128 + tid] = __uint_as_float(v[h]);
                            v[h] = 0u;                               // score columns are not part of the sum of squares
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        q0 = fmaf(__uint_as_float(v[i]), __uint_as_float(v[i]), q0);
                        q1 = fmaf(__uint_as_float(v[i + 1]), __uint_as_float(v[i + 1]), q1);
                    }
                }
            }
            azptx::tc_fence_before();
            azptx::mbar_arrive(st_empty);
            sqb[tid] = q0 + q1;
            AZ_TR(2, 3);
            azptx::bar_sync_named(1, 128);
            AZ_TR(2, 4);
            // ---- softmax over the T tokens of each (board, head); weight = p * rstd
            {
                float* sw = swb + ph * 128 + pb * TS;
                const float* sq = sqb + pb * TS;
                // branch-free: every thread reads 8 (clamped) positions first, then runs 8 independent
                // rsqrt / exp chains; invalid positions are masked by select, not skipped by a branch
                float sc[8], rs[8], qv[8], wv[8];
                bool ok[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int t = sub + i * tpp;
                    ok[i] = i < ne && t < T;
                    const int tt = ok[i] ? t : 0;
                    qv[i] = sq[tt];
                    wv[i] = sw[tt];
                }
                float mx = -1e30f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float r;
                    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(qv[i], inv_d, 1e-5f)));
                    rs[i] = ok[i] ? r : 0.f;
                    sc[i] = ok[i] ? wv[i] * r : -1e30f;
                    mx = fmaxf(mx, sc[i]);
                }
#pragma unroll
                for (int o = 16; o; o >>= 1) {
                    const float v = __shfl_xor_sync(0xffffffffu, mx, o);
                    if (o < tpp) mx = fmaxf(mx, v);
                }
                AZ_TR(2, 6);
                float sum = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float e;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((sc[i] - mx) * 1.4426950408889634f));
                    sc[i] = ok[i] ? e : 0.f;
                    sum += sc[i];
                }
#pragma unroll
                for (int o = 16; o; o >>= 1) {
                    const float v = __shfl_xor_sync(0xffffffffu, sum, o);
                    if (o < tpp) sum += v;
                }
                AZ_TR(2, 7);
                float inv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(sum));
                if (pair_on) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int t = sub + i * tpp;
                        if (i < ne && t < TS) sw[t] = sc[i] * inv * rs[i];   // pad rows (t >= T) get weight 0
                    }
                }
            }
            azptx::mbar_arrive(&sw_full[buf]);                       // release: this thread's weights are written
            AZ_TR(2, 5);
        }
    } else {
        // ------------------------------ V group: softmax-weighted sums of the value blocks -------------
        const int vw = warp - 4;                                     // 0..7
        const int grp = vw >> 2;                                     // owns blocks fb with (fb & 1) == grp
        const int tid = (vw & 3) * 32 + lane;                        // feature inside a block = TMEM lane
        const uint32_t tq = tmem_base + ((uint32_t)((vw & 3) * 32) << 16);
        const int hd = D / a.H;
        int it = 0, slot0 = 0, round0 = 0;                           // block counter at fb = 0 of the current tile
        wait_upstream();
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            const float* swb = s_w + buf * 8 * 128;
            AZ_TR(3, 0);
            azptx::mbar_wait(&sw_full[buf], (it >> 1) & 1);
            AZ_TR(3, 1);
            int slot = slot0, round = round0;
            for (int fb = 0; fb < FB; ++fb) {
                if ((fb & 1) == grp) {
                    azptx::mbar_wait(&v_full[slot], round & 1);
                    if (fb < 2) AZ_TR(3, 2); else AZ_TR(3, 4);
                    azptx::tc_fence_after();
                    const int f = fb * 128 + tid;
                    const float* wrow = swb + (f / hd) * 128;
                    const float bias = __ldg(a.bias_v + f);
                    const uint32_t tcol = tq + 128 + slot * NTP;
                    bf16* dst = a.oc + (size_t)tile * a.nb * D + f;
                    if constexpr (TS8 >= 5) {
                        // Long token strides: a board is read in two halves, the second half (and then the next
                        // board's first half) is in flight from TMEM while the previous one is accumulated.
                        constexpr int H0 = (TS8 + 1) / 2, H1 = TS8 - H0;
                        uint32_t va[H0][8], vb[H1][8];
                        auto acc_half = [&](const float* wr, auto& v, auto nh, float& a0, float& a1, float& a2, float& a3) {
                            constexpr int NH = decltype(nh)::value;
#pragma unroll
                            for (int j = 0; j < NH; ++j) {
                                const float4 w0 = *reinterpret_cast<const float4*>(wr + j * 8);
                                const float4 w1 = *reinterpret_cast<const float4*>(wr + j * 8 + 4);
                                a0 = fmaf(w0.x, __uint_as_float(v[j][0]), fmaf(w1.x, __uint_as_float(v[j][4]), a0));
                                a1 = fmaf(w0.y, __uint_as_float(v[j][1]), fmaf(w1.y, __uint_as_float(v[j][5]), a1));
                                a2 = fmaf(w0.z, __uint_as_float(v[j][2]), fmaf(w1.z, __uint_as_float(v[j][6]), a2));
                                a3 = fmaf(w0.w, __uint_as_float(v[j][3]), fmaf(w1.w, __uint_as_float(v[j][7]), a3));
                            }
                        };
#pragma unroll
                        for (int j = 0; j < H0; ++j) azptx::tmem_ld8(tcol + j * 8, va[j]);
                        azptx::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < H0; ++j) azptx::tmem_regs_ready(va[j]);
                        for (int b = 0; b < a.nb; ++b) {
#pragma unroll
                            for (int j = 0; j < H1; ++j) azptx::tmem_ld8(tcol + b * TS + (H0 + j) * 8, vb[j]);
                            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                            acc_half(wrow + b * TS, va, std::integral_constant<int, H0>(), a0, a1, a2, a3);
                            azptx::tmem_ld_wait();
#pragma unroll
                            for (int j = 0; j < H1; ++j) azptx::tmem_regs_ready(vb[j]);
                            if (b + 1 < a.nb) {
#pragma unroll
                                for (int j = 0; j < H0; ++j) azptx::tmem_ld8(tcol + (b + 1) * TS + j * 8, va[j]);
                            } else {                                 // slot fully read: hand it back to the MMA warp
                                azptx::tc_fence_before();
                                azptx::mbar_arrive(&v_empty[slot]);
                            }
                            acc_half(wrow + b * TS + H0 * 8, vb, std::integral_constant<int, H1>(), a0, a1, a2, a3);
                            if (tile * a.nb + b < items) dst[(size_t)b * D] = __float2bfloat16((a0 + a1) + (a2 + a3) + bias);
                            if (b + 1 < a.nb) {
                                azptx::tmem_ld_wait();
#pragma unroll
                                for (int j = 0; j < H0; ++j) azptx::tmem_regs_ready(va[j]);
                            }
                        }
                    } else {
                    constexpr int NBV = 8 / TS8 > 4 ? 4 : 8 / TS8;   // boards fetched per TMEM round trip
                    for (int b0 = 0; b0 < a.nb; b0 += NBV) {
                        uint32_t v[NBV][TS8][8];
#pragma unroll
                        for (int bi = 0; bi < NBV; ++bi)
                            if (b0 + bi < a.nb) {
#pragma unroll
                                for (int j = 0; j < TS8; ++j) azptx::tmem_ld8(tcol + (b0 + bi) * TS + j * 8, v[bi][j]);
                            }
                        azptx::tmem_ld_wait();
                        if (b0 + NBV >= a.nb) {                      // slot fully read: hand it back to the MMA warp
                            azptx::tc_fence_before();
                            azptx::mbar_arrive(&v_empty[slot]);
                        }
#pragma unroll
                        for (int bi = 0; bi < NBV; ++bi) {
                            const int b = b0 + bi;
                            if (b < a.nb) {
                                float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
                                for (int j = 0; j < TS8; ++j) {
                                    const float4 w0 = *reinterpret_cast<const float4*>(wrow + b * TS + j * 8);
                                    const float4 w1 = *reinterpret_cast<const float4*>(wrow + b * TS + j * 8 + 4);
                                    a0 += w0.x * __uint_as_float(v[bi][j][0]) + w1.x * __uint_as_float(v[bi][j][4]);
                                    a1 += w0.y * __uint_as_float(v[bi][j][1]) + w1.y * __uint_as_float(v[bi][j][5]);
                                    a2 += w0.z * __uint_as_float(v[bi][j][2]) + w1.z * __uint_as_float(v[bi][j][6]);
                                    a3 += w0.w * __uint_as_float(v[bi][j][3]) + w1.w * __uint_as_float(v[bi][j][7]);
                                }
                                if (tile * a.nb + b < items) dst[(size_t)b * D] = __float2bfloat16((a0 + a1) + (a2 + a3) + bias);
                            }
                        }
                    }
                    }
                }
                if (++slot == nslots) { slot = 0; ++round; }
            }
            slot0 = slot; round0 = round;
            azptx::mbar_arrive(&sw_empty[buf]);                      // this thread no longer reads s_w[buf]
            AZ_TR(3, 5);
        }
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<512>(tmem_base);
}
