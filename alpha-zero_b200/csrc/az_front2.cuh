// Fused front end, reassociated: the value projection leaves the per-token path.
//
// az_front.cuh computes  o_h = sum_t w_{h,t} (Wv2 a_t)  with one value MMA per (feature block, tile) and a
// per-token epilogue (w_{h,t} = softmax weight x rstd_t).  The sum is linear in the binary token rows:
//     o_h = Wv2_h (sum_t w_{h,t} a_t) = Wv2_h u_h ,      u_h = A^T w_h   (one K-vector per board and head)
// and the out-projection that follows is linear too:  Wo o + bo = sum_h (Wo_h Wv2_h) u_h + (Wo bv + bo).
// So this kernel emits only  U[board][head][K = 128]  (bf16) and the out-proj GEMM runs on
// [boards][H*128] x W'[D][H*128]^T with W' = Wo_h Wv2_h folded in fp64 at az_net_finalize -- the value
// projection (86 % of the front kernel's MMA work and its whole per-token TMEM epilogue) is gone.
//
// Per tile of `nb` boards (NTP <= 128 token rows), two tcgen05 MMAs read the SAME bit-built tile A[token][K]:
//   stats : D1[token][j] = A x Wst^T               A as the K-major operand   (M = 128 tokens, N = NS, K = 128)
//   U     : D2[k][pair]  = A^T x Wblk              A as the MN-major operand  (M = 128 k,      N = NP, K = NTP tokens)
// Wblk[pair = board*H + head][token] holds w_{h,t} (bf16) for the tokens of that board and zero elsewhere
// (block diagonal), written by the softmax warps straight into the UMMA operand layout.
//
//   warps 0-3 / 4-7  two statistic + softmax groups (even / odd tiles, each with its own TMEM accumulator)
//   warps 8-11       U epilogue: TMEM -> bf16 -> global
//   warps 12-13      A-tile builders (4 stages: a tile is read again by the U MMA one softmax later)
//   warp 15          TMA (statistic weights, once) + MMA issue, U MMAs trail the statistic MMAs by 2 tiles
// Reference: ai/nn.py:9-36 (embedding), :52-56 (block: norm1 + attention), restricted to the cls query.
#pragma once
#include "az_front.cuh"

#define AZ_FRONT2_STAGES 4
#define AZ_FRONT2_LAG 2
// Experiment, off: warp 9 (U epilogue group) issues the U MMAs and warp 15 only the statistic MMAs.  Measured: no gain
// (Connect4 front 23.0 -> 25.0 us, Gomoku 0.2275 -> 0.2257 ms).  The issuing thread is blocked for the ~70-95 cycles each
// small-N MMA occupies the tensor pipe, so two issuers share the same 1.13 k pipe-cycles per tile, and warp 9's issue now
// delays the drain its three sibling warps wait for.
#ifndef AZ_FRONT2_SPLIT_ISSUE
#define AZ_FRONT2_SPLIT_ISSUE 0
#endif
#define AZ_FRONT2_BUILDERS 96      // warps 12-14: one rewritten token row per thread (two when a tile has more than 96)

static inline size_t az_front2_smem(const AzFrontArgs& a) {
    const size_t tile = (size_t)a.NTP * 128;
    return 1024 + AZ_FRONT2_STAGES * 2 * tile + 2 * (size_t)a.NS * 128 + 2 * 2 * (size_t)a.NP * 128 +
           2 * 2 * 8 * 128 * 4 + 2 * 2 * 128 * 4 + 2 * (size_t)a.NP * 256 + 128 + 256;
}

namespace azptx {
// A^T of a K-major SWIZZLE_128B tile pair [2 k-halves][rows][128 B] read as an MN-major operand:
// 64 MN elements (one 128-byte row) contiguous, LBO = bytes between the two k-halves, SBO = bytes
// between groups of 8 rows (the contraction index), same swizzle.
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr, uint32_t lbo_bytes) {
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}
}  // namespace azptx

// One rewritten token row of the A tile per thread.  Only the rows of real board cells change from tile to tile
// (the cls row, pad rows and the one-hot columns are constants): active row i = cell i % (T-1) of board i / (T-1).
struct AzRowBuilder {
    bool on;
    int b, t, r, rs, ls, ob;
    uint32_t inb, pmask;
    int sh[5], ks[5];
    uint32_t cm[5];
    ulonglong2 w_cur, w_nxt;
    uint32_t s_cur, s_nxt;                                       // raw side-to-move byte, 0xFFFFFFFF = row unused / board beyond the batch
    uint64_t kw0, kw1;

    __device__ __forceinline__ void init(const AzFrontArgs& a, int ai, int TS) {
        const int T = a.T, P = a.patch, half = P >> 1;
        const uint32_t colmask = (1u << a.cols) - 1u;
        pmask = (1u << P) - 1u;
        b = ai / (T - 1);
        t = 1 + ai - b * (T - 1);
        on = ai < a.nb * (T - 1);
        r = b * TS + t;
        ob = a.kc + t;
        const int cr = (t - 1) / a.cols, cc = (t - 1) - cr * a.cols;
        rs = cc >= half ? cc - half : 0;
        ls = cc >= half ? 0 : half - cc;
        inb = 0;                                                 // in-bounds taps of the window (plane 2 is all ones)
        for (int kr = 0; kr < P; ++kr) {
            const int rr = cr + kr - half;
            if (rr >= 0 && rr < a.rows) inb |= (((colmask >> rs) << ls) & pmask) << (kr * P);
        }
        if (!on) inb = 0u;
#pragma unroll
        for (int kr = 0; kr < 5; ++kr) {                         // window rows outside the board / beyond the patch: empty mask
            const int rr = cr + kr - half;
            const bool ok = kr < P && rr >= 0 && rr < a.rows && on;
            ks[kr] = kr < P ? kr * P : 0;
            sh[kr] = ok ? rr * a.cols : 0;
            cm[kr] = ok ? colmask : 0u;
        }
    }
    // loads only: nothing consumes the registers until the next tile is encoded
    __device__ __forceinline__ void fetch(const AzFrontArgs& a, int tile, int tiles, int items, ulonglong2& w, uint32_t& sm) const {
        const int item = tile * a.nb + b;
        w = make_ulonglong2(0ull, 0ull);
        sm = 0xFFFFFFFFu;
        if (tile < tiles && on && item < items) {
            w = *reinterpret_cast<const ulonglong2*>(a.leaf_words + (size_t)(a.batch.offset + item) * 2);
            sm = a.leaf_player[a.batch.offset + item];
        }
    }
    // K columns of the row as two 64-bit words: taps of plane 0 at bit 0, plane 1 at bit pp, plane 2 (in-bounds taps
    // when side 1 is to move) at bit 2*pp, one-hot token index at bit kc + t
    __device__ __forceinline__ void encode(const AzFrontArgs& a) {
        const int pp = a.patch * a.patch;
        uint64_t k0 = ob < 64 ? 1ull << ob : 0ull, k1 = ob >= 64 ? 1ull << (ob - 64) : 0ull;
        uint32_t p0 = 0, p1 = 0;
#pragma unroll
        for (int kr = 0; kr < 5; ++kr) {
            const uint32_t b0 = (uint32_t)(w_cur.x >> sh[kr]) & cm[kr];
            const uint32_t b1 = (uint32_t)(w_cur.y >> sh[kr]) & cm[kr];
            p0 |= (((b0 >> rs) << ls) & pmask) << ks[kr];
            p1 |= (((b1 >> rs) << ls) & pmask) << ks[kr];
        }
        k0 |= (uint64_t)p0 | ((uint64_t)p1 << pp);
        if (a.planes == 3 && s_cur != 0u && s_cur != 0xFFFFFFFFu) {      // side 1 to move; the taps may straddle bit 64
            const uint64_t ib = inb;
            k0 |= ib << (2 * pp);
            if (3 * pp > 64) k1 |= ib >> (64 - 2 * pp);
        }
        kw0 = k0;
        kw1 = k1;
    }
    // every LUT read of the row first, then the stores: the compiler cannot prove that the LUT and the tile do not
    // alias, so interleaved they run as one serial load -> store chain (85 cycles per chunk)
    __device__ __forceinline__ void store(unsigned char* stage, uint32_t tile_bytes, const uint2* s_lut, int nch) const {
        if (!on) return;
        unsigned char* arow = stage + (size_t)r * 128;
        const int sw = r & 7;
        uint2 lo[10], hi[10];
#pragma unroll
        for (int c = 0; c < 10; ++c) {
            const uint32_t by = (uint32_t)((c < 8 ? kw0 : kw1) >> ((c & 7) * 8)) & 0xFFu;
            lo[c] = s_lut[by & 15u];
            hi[c] = s_lut[by >> 4];
        }
#pragma unroll
        for (int c = 0; c < 10; ++c)
            if (c < nch)
                *reinterpret_cast<uint4*>(arow + (c < 8 ? 0u : tile_bytes) + (((c & 7) ^ sw) << 4)) = make_uint4(lo[c].x, lo[c].y, hi[c].x, hi[c].y);
    }
};

template <int TS8>
__global__ void __launch_bounds__(AZ_FRONT_THREADS, 1) az_front2_kernel(const __grid_constant__ CUtensorMap tm_wst, AzFrontArgs a) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (azptx::smem_u32(smem_dyn) & 1023u)) & 1023u);
    constexpr int NA = AZ_FRONT2_STAGES, LAG = AZ_FRONT2_LAG;
    const int D = a.D, T = a.T, P = a.patch, NTP = a.NTP, NS = a.NS, NP = a.NP;
    constexpr int TS = TS8 * 8;
    const uint32_t tile_bytes = (uint32_t)NTP * 128;
    unsigned char* s_a = smem;                                        // [NA][2 k-halves][NTP rows][128 B]
    unsigned char* s_wst = s_a + (size_t)NA * 2 * tile_bytes;         // [2][NS rows][128 B]
    unsigned char* s_wb = s_wst + 2 * (size_t)NS * 128;               // [2 groups][2 token halves][NP rows][128 B]
    float* s_sc = reinterpret_cast<float*>(s_wb + 2 * 2 * (size_t)NP * 128);   // [2 groups][2][8 heads][128 rows] raw scores
    float* s_q = s_sc + 2 * 2 * 8 * 128;                              // [2 groups][2][128] sum of squares per row
    bf16* s_us = reinterpret_cast<bf16*>(s_q + 2 * 2 * 128);          // [2][NP][128] U staging: TMEM [k][pair] -> rows [pair][k]
    uint2* s_lut = reinterpret_cast<uint2*>(s_us + 2 * (size_t)NP * 128);   // 16 entries: nibble -> 4 bf16 (0 / 1.0)
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_lut + 16);
    uint64_t* w_full = bars;             // 1
    uint64_t* a_full = bars + 1;         // NA  (64 builder arrivals)
    uint64_t* a_empty = a_full + NA;     // NA  (commit: the U MMAs of the tile are done)
    uint64_t* st_full = a_empty + NA;    // 2   (commit)
    uint64_t* st_empty = st_full + 2;    // 2   (128 arrivals of the group)
    uint64_t* sw_full = st_empty + 2;    // 2   (128 arrivals: Wblk of the group's tile is written)
    uint64_t* sw_empty = sw_full + 2;    // 2   (commit: the U MMAs that read Wblk are done)
    uint64_t* u_full = sw_empty + 2;     // 2   (commit)
    uint64_t* u_empty = u_full + 2;      // 2   (128 U-epilogue arrivals)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(u_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int items = 0, tiles = 0;
    auto wait_upstream = [&]() {
        azptx::griddep_wait();
        items = az_batch_items(a.batch);
        tiles = (items + a.nb - 1) / a.nb;
    };

    if (threadIdx.x < 16) {
        const uint32_t n = threadIdx.x;
        s_lut[n] = make_uint2(((n & 1u) ? 0x00003F80u : 0u) | ((n & 2u) ? 0x3F800000u : 0u),
                              ((n & 4u) ? 0x00003F80u : 0u) | ((n & 8u) ? 0x3F800000u : 0u));
    }
    for (int i = threadIdx.x; i < 2 * 2 * NP * 8; i += AZ_FRONT_THREADS)     // Wblk: zero outside the boards' own blocks, for good
        reinterpret_cast<uint4*>(s_wb)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < NA * 2 * NTP * 8; i += AZ_FRONT_THREADS)   // A stages: pad rows and unused columns stay zero
        reinterpret_cast<uint4*>(s_a)[i] = make_uint4(0u, 0u, 0u, 0u);
    azptx::fence_proxy_async();
    if (warp == 15 && lane == 0) {
        azptx::prefetch_tmap(&tm_wst);
        azptx::mbar_init(w_full, 1);
        for (int i = 0; i < NA; ++i) {
            azptx::mbar_init(&a_full[i], a.nb * (a.T - 1) > AZ_FRONT2_BUILDERS ? AZ_FRONT2_BUILDERS + 32 : AZ_FRONT2_BUILDERS);
            azptx::mbar_init(&a_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            azptx::mbar_init(&st_full[i], 1);
            azptx::mbar_init(&st_empty[i], 128);
            azptx::mbar_init(&sw_full[i], 128);
            azptx::mbar_init(&sw_empty[i], 1);
            azptx::mbar_init(&u_full[i], 1);
            azptx::mbar_init(&u_empty[i], 128);
        }
        azptx::fence_barrier_init();
    }
    if (warp == 0) azptx::tmem_alloc<512>(tmem_slot);
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // TMEM columns: [0, 256) the two statistic accumulators, [256, 256 + 2 NP) the two U accumulators.
    azptx::griddep_launch();

    if (warp == 15) {
        // ------------------------------ statistic weights in once, then MMA issue ------------------------
        if (lane == 0) {
            azptx::mbar_arrive_expect_tx(w_full, (uint32_t)(2 * NS * 128));
            for (int kb = 0; kb < 2; ++kb) azptx::tma_load_2d(s_wst + (size_t)kb * NS * 128, &tm_wst, w_full, kb * 64, 0);
        }
        const uint32_t idesc_st = azptx::umma_idesc_bf16(128, (uint32_t)NS);
        const uint32_t idesc_u = azptx::umma_idesc_bf16(128, (uint32_t)NP) | (1u << 15);     // A operand MN-major
        const uint64_t d_wst0 = azptx::umma_desc_k128(azptx::smem_u32(s_wst));
        const uint64_t d_wst1 = azptx::umma_desc_k128(azptx::smem_u32(s_wst + (size_t)NS * 128));
        const int usteps = NTP >> 4;                                  // 16 tokens per U MMA
        azptx::mbar_wait(w_full, 0);
        wait_upstream();
        const int n_my = tiles > (int)blockIdx.x ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
        for (int j = 0; j < n_my + (AZ_FRONT2_SPLIT_ISSUE ? 0 : LAG); ++j) {
            const int it = j;
            AZ_TR1(0, 0);
            if (j < n_my) {                                           // statistics of tile j
                const int st = j % NA, g = j & 1;
                azptx::mbar_wait(&a_full[st], (j / NA) & 1);
                AZ_TR1(0, 1);
                azptx::mbar_wait(&st_empty[g], ((j >> 1) & 1) ^ 1);
                AZ_TR1(0, 2);
                azptx::tc_fence_after();
                if (azptx::elect_one()) {
                    const uint32_t sa = azptx::smem_u32(s_a + (size_t)st * 2 * tile_bytes);
                    const uint64_t da0 = azptx::umma_desc_k128(sa), da1 = azptx::umma_desc_k128(sa + tile_bytes);
                    const uint32_t dcol = tmem_base + g * 128;
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(dcol, da0 + 2 * ks, d_wst0 + 2 * ks, idesc_st, ks != 0);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) azptx::umma_bf16(dcol, da1 + 2 * ks, d_wst1 + 2 * ks, idesc_st, 1);
                    azptx::umma_commit(&st_full[g]);
                }
                __syncwarp();
            }
            const int i = j - LAG;
            if (!AZ_FRONT2_SPLIT_ISSUE && i >= 0) {                   // U of tile i: its softmax weights are in Wblk
                const int st = i % NA, g = i & 1;
                AZ_TR1(0, 3);
                azptx::mbar_wait(&sw_full[g], (i >> 1) & 1);
                AZ_TR1(0, 4);
                azptx::mbar_wait(&u_empty[g], ((i >> 1) & 1) ^ 1);
                AZ_TR1(0, 5);
                azptx::tc_fence_after();
                if (azptx::elect_one()) {
                    const uint64_t dat = azptx::umma_desc_mn128(azptx::smem_u32(s_a + (size_t)st * 2 * tile_bytes), tile_bytes);
                    const uint32_t wb = azptx::smem_u32(s_wb + (size_t)g * 2 * NP * 128);
                    const uint64_t dw0 = azptx::umma_desc_k128(wb), dw1 = azptx::umma_desc_k128(wb + NP * 128);
                    const uint32_t dcol = tmem_base + 256 + g * NP;
                    for (int ks = 0; ks < usteps; ++ks)               // +16 tokens: 2048 bytes of A, 32 bytes of a Wblk row
                        azptx::umma_bf16(dcol, dat + 128 * ks, (ks < 4 ? dw0 : dw1) + 2 * (ks & 3), idesc_u, ks != 0);
                    azptx::umma_commit(&u_full[g]);
                    azptx::umma_commit(&a_empty[st]);
                    azptx::umma_commit(&sw_empty[g]);
                }
                __syncwarp();
            }
        }
    } else if (warp >= 12 && warp < 15) {
        // ------------------------------ A-tile builders: active row bt of every tile -------------------
        // (active rows 96.. of a tile, if any, are written by warp 8 between its U epilogues)
        const int bt = threadIdx.x - 384;
        const int nch = (a.kc + 7) >> 3;
        int it = 0;
        AZ_TR(1, 5);
        for (int r = bt; r < a.nb * TS; r += AZ_FRONT2_BUILDERS) {   // one-hot columns: the stages were zero-filled before the CTA-wide sync
            const int t = r % TS;
            if (t >= T) continue;
            const int ob = a.kc + t;
            for (int st = 0; st < NA; ++st)
                *reinterpret_cast<unsigned short*>(s_a + (size_t)st * 2 * tile_bytes + (size_t)(ob >> 6) * tile_bytes + (size_t)r * 128 +
                                                   (((((uint32_t)ob & 63u) >> 3) ^ ((uint32_t)r & 7u)) << 4) + (ob & 7) * 2) = 0x3F80;
        }
        AzRowBuilder rb;
        rb.init(a, bt, TS);
        AZ_TR(1, 6);
        wait_upstream();
        rb.fetch(a, blockIdx.x, tiles, items, rb.w_cur, rb.s_cur);
        AZ_TR(1, 7);
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int st = it % NA;
            rb.fetch(a, tile + gridDim.x, tiles, items, rb.w_nxt, rb.s_nxt);     // in flight while this tile is encoded
            rb.encode(a);
            AZ_TR(1, 0);
            azptx::mbar_wait(&a_empty[st], ((it / NA) & 1) ^ 1);
            AZ_TR(1, 1);
            rb.store(s_a + (size_t)st * 2 * tile_bytes, tile_bytes, s_lut, nch);
            AZ_TR(1, 3);
            azptx::fence_proxy_async();                              // generic-proxy writes -> visible to the MMA
            AZ_TR(1, 4);
            azptx::mbar_arrive(&a_full[st]);
            AZ_TR(1, 2);
            rb.w_cur = rb.w_nxt;
            rb.s_cur = rb.s_nxt;
        }
    } else if (warp < 8) {
        // ------------------------------ S groups: statistic, scores, softmax -> Wblk --------------------
        const int g = warp >> 2;
        const int tid = threadIdx.x & 127;                           // token row
        const uint32_t tq = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + g * 128;
        const int H = a.H;
        const int pairs = a.nb * H;
        int tpp = 32;
        while (tpp * pairs > 128) tpp >>= 1;                         // threads per (board, head) softmax
        const int pair = tid / tpp, sub = tid - pair * tpp;
        const bool pair_on = pair < pairs;
        const int pb = pair_on ? pair / H : 0, ph = pair_on ? pair - pb * H : 0;
        const float inv_d = 1.0f / (float)D;
        const int ne = (TS + tpp - 1) / tpp;                         // tokens per softmax thread (<= 8, checked on the host)
        // Wblk row of this thread's (board, head): 128-byte rows of 64 tokens, 16-byte chunks swizzled by the row
        unsigned char* wrow = s_wb + (size_t)g * 2 * NP * 128 + (size_t)(pair >> 3) * 1024 + (size_t)(pair & 7) * 128;
        const uint32_t wsw = (uint32_t)(pair & 7);
        wait_upstream();
        for (int it = g;; it += 2) {
            const int tile = blockIdx.x + it * gridDim.x;
            if (tile >= tiles) break;
            const int u = it >> 1, buf = u & 1;
            float* scb = s_sc + (g * 2 + buf) * 8 * 128;
            float* sqb = s_q + (g * 2 + buf) * 128;
            AZ_TR(2, 0);
            azptx::mbar_wait(&st_full[g], u & 1);
            AZ_TR(2, 1);
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
                            scb[h * 128 + tid] = __uint_as_float(v[h]);
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
            azptx::mbar_arrive(&st_empty[g]);
            sqb[tid] = q0 + q1;
            AZ_TR(2, 2);
            azptx::bar_sync_named(1 + g, 128);
            AZ_TR(2, 3);
            // ---- softmax over the T tokens of each (board, head); weight = p * rstd
            const float* sw = scb + ph * 128 + pb * TS;
            const float* sq = sqb + pb * TS;
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
            float inv;
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(sum));
            AZ_TR(2, 4);
            azptx::mbar_wait(&sw_empty[g], (u & 1) ^ 1);             // the U MMAs of this group's previous tile have read Wblk
            if (pair_on) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int t = sub + i * tpp;
                    if (i < ne && t < TS) {                          // pad rows (t >= T) get weight 0
                        const int tok = pb * TS + t;
                        unsigned char* p = wrow + (size_t)(tok >> 6) * NP * 128 + (((((uint32_t)tok & 63u) >> 3) ^ wsw) << 4) + (tok & 7) * 2;
                        *reinterpret_cast<bf16*>(p) = __float2bfloat16(sc[i] * inv * rs[i]);
                    }
                }
            }
            azptx::fence_proxy_async();                              // generic-proxy writes -> visible to the MMA
            azptx::mbar_arrive(&sw_full[g]);
            AZ_TR(2, 5);
        }
    } else if (warp < 12) {
        // ------------------------------ U epilogue: TMEM [k][pair] -> bf16 uc[board][head][k] -----------
        const int q = warp & 3;
        const int k = q * 32 + lane;                                 // TMEM lane = K index
        const int pairs = a.nb * a.H;
        // warp 8 also writes the active rows 96.. of every tile (a tile of two 7x7 boards has 98).  It builds tile j
        // and THEN drains the accumulator of tile j - NA: both wait for the same event (the U MMAs of tile j - NA),
        // so the build never queues behind a recent MMA (with a lag of 2 the chain U MMA -> epilogue -> build ->
        // statistics MMA serialised the whole pipeline: 3.4 k cycles per tile instead of 1.5 k).
        const bool extra = a.nb * (T - 1) > AZ_FRONT2_BUILDERS;
        const bool assist = warp == 8 && extra;
#if AZ_FRONT2_SPLIT_ISSUE
        // Warp 9 issues the U MMAs of tile j - 2 (its softmax finished about one tile ago: no wait), then all four warps
        // drain tile j - 3, whose MMAs were issued one iteration earlier.  Two U accumulators: U(i) needs drain(i - 2),
        // which ran in iteration i + 1 < i + 2.
        const int ulag = 3;
        const bool uissue = warp == 9;
        const uint32_t idesc_u = azptx::umma_idesc_bf16(128, (uint32_t)NP) | (1u << 15);     // A operand MN-major
        const int usteps = NTP >> 4;                                  // 16 tokens per U MMA
#else
        const int ulag = extra ? NA : 0;
#endif
        const int nch = (a.kc + 7) >> 3;
        AzRowBuilder rb;
        rb.init(a, AZ_FRONT2_BUILDERS + lane, TS);
        wait_upstream();
        if (assist) rb.fetch(a, blockIdx.x, tiles, items, rb.w_cur, rb.s_cur);
        const int n_my = tiles > (int)blockIdx.x ? (tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
        for (int j = 0; j < n_my + ulag; ++j) {
            if (assist && j < n_my) {
                const int bst = j % NA, btile = blockIdx.x + j * gridDim.x;
                rb.fetch(a, btile + gridDim.x, tiles, items, rb.w_nxt, rb.s_nxt);
                rb.encode(a);
                azptx::mbar_wait(&a_empty[bst], ((j / NA) & 1) ^ 1);
                rb.store(s_a + (size_t)bst * 2 * tile_bytes, tile_bytes, s_lut, nch);
                azptx::fence_proxy_async();
                azptx::mbar_arrive(&a_full[bst]);
                rb.w_cur = rb.w_nxt;
                rb.s_cur = rb.s_nxt;
            }
#if AZ_FRONT2_SPLIT_ISSUE
            {
                const int i = j - 2;
                if (uissue && i >= 0 && i < n_my) {                   // U of tile i: its softmax weights are in Wblk
                    const int st = i % NA, g = i & 1;
                    azptx::mbar_wait(&sw_full[g], (i >> 1) & 1);
                    azptx::mbar_wait(&u_empty[g], ((i >> 1) & 1) ^ 1);
                    azptx::tc_fence_after();
                    if (azptx::elect_one()) {
                        const uint64_t dat = azptx::umma_desc_mn128(azptx::smem_u32(s_a + (size_t)st * 2 * tile_bytes), tile_bytes);
                        const uint32_t wb = azptx::smem_u32(s_wb + (size_t)g * 2 * NP * 128);
                        const uint64_t dw0 = azptx::umma_desc_k128(wb), dw1 = azptx::umma_desc_k128(wb + NP * 128);
                        const uint32_t dcol = tmem_base + 256 + g * NP;
                        for (int ks = 0; ks < usteps; ++ks)           // +16 tokens: 2048 bytes of A, 32 bytes of a Wblk row
                            azptx::umma_bf16(dcol, dat + 128 * ks, (ks < 4 ? dw0 : dw1) + 2 * (ks & 3), idesc_u, ks != 0);
                        azptx::umma_commit(&u_full[g]);
                        azptx::umma_commit(&a_empty[st]);
                        azptx::umma_commit(&sw_empty[g]);
                    }
                    __syncwarp();
                }
            }
#endif
            const int it = j - ulag;
            if (it < 0 || it >= n_my) continue;
            const int tile = blockIdx.x + it * gridDim.x;
            const int g = it & 1, u = it >> 1;
            AZ_TR(3, 0);
            azptx::mbar_wait(&u_full[g], u & 1);
            AZ_TR(3, 1);
            azptx::tc_fence_after();
            const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + 256 + g * NP;
            bf16* us = s_us + (size_t)(it & 1) * NP * 128;
            for (int c = 0; c < NP; c += 16) {
                uint32_t v0[8], v1[8];
                azptx::tmem_ld8(tcol + c, v0);
                azptx::tmem_ld8(tcol + c + 8, v1);
                azptx::tmem_ld_wait();
                azptx::tmem_regs_ready(v0);
                azptx::tmem_regs_ready(v1);
                if (c + 16 >= NP) {                                  // accumulator fully read: hand it back
                    azptx::tc_fence_before();
                    azptx::mbar_arrive(&u_empty[g]);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    us[(size_t)(c + i) * 128 + k] = __float2bfloat16(__uint_as_float(v0[i]));
                    us[(size_t)(c + 8 + i) * 128 + k] = __float2bfloat16(__uint_as_float(v1[i]));
                }
            }
            azptx::bar_sync_named(3, 128);
            // rows [pair][128] of a tile are contiguous in uc: 16-byte copies (rows of boards beyond the batch land in
            // the padding rows of uc and are never read)
            uint4* dst = reinterpret_cast<uint4*>(a.uc + (size_t)tile * pairs * 128);
            const uint4* src = reinterpret_cast<const uint4*>(us);
            for (int idx = k; idx < pairs * 16; idx += 128) dst[idx] = src[idx];
            AZ_TR(3, 2);
        }
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<512>(tmem_base);
}
