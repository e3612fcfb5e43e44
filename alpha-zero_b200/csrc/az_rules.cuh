// Bitboard rules of TicTacToe / Connect4 / Gomoku for sm_100a device code.
//
// Restates, on side bitboards (bit r*cols+c), what the reference computes on float planes:
//   make_move / undo_move   games/tictactoe.py:36-51  connect4.py:55-70  gomoku.py:50-63
//   check_winner            games/tictactoe.py:53-79  connect4.py:72-98  gomoku.py:65-91
//   get_valid_moves (ORDER) games/tictactoe.py:81-83  connect4.py:43-53  gomoku.py:93-106
// The Gomoku legal list is `list(set)` in the reference, i.e. CPython set slot order; the
// emulation below follows CPython 3.12 Objects/setobject.c (set_add_entry,
// set_table_resize, set_insert_clean) with the tuple hashes supplied by the host.
#pragma once
#include "az_types.h"

#ifdef __CUDA_ARCH__
#define AZ_CTZ64(x) (__ffsll((long long)(x)) - 1)
#define AZ_POPC64(x) __popcll(x)
#else
#define AZ_CTZ64(x) __builtin_ctzll(x)
#define AZ_POPC64(x) __builtin_popcountll(x)
#endif

template <int NW>
struct AzBoard {
    uint64_t w[2][NW];   // w[player][word]

    // Player and word indices are data dependent; selecting with predicates instead of indexing keeps
    // the board in registers (a dynamically indexed array is demoted to local memory).
    __host__ __device__ __forceinline__ uint64_t word(int p, int i) const {
        uint64_t v = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k)
            if (NW == 1 || k == i) v = p ? w[1][k] : w[0][k];
        return v;
    }
    __host__ __device__ __forceinline__ bool has(int p, int cell) const {
        return (word(p, NW == 1 ? 0 : (cell >> 6)) >> (cell & 63)) & 1ull;
    }
    __host__ __device__ __forceinline__ bool empty(int cell) const {
        const int i = NW == 1 ? 0 : (cell >> 6);
        return !(((word(0, i) | word(1, i)) >> (cell & 63)) & 1ull);
    }
    __host__ __device__ __forceinline__ void put(int p, int cell) {
        const uint64_t bit = 1ull << (cell & 63);
        const int i = NW == 1 ? 0 : (cell >> 6);
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (NW == 1 || k == i) {
                w[0][k] |= p ? 0ull : bit;
                w[1][k] |= p ? bit : 0ull;
            }
        }
    }
    __host__ __device__ __forceinline__ void clear(int p, int cell) {
        const uint64_t bit = 1ull << (cell & 63);
        const int i = NW == 1 ? 0 : (cell >> 6);
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (NW == 1 || k == i) {
                w[0][k] &= p ? ~0ull : ~bit;
                w[1][k] &= p ? ~bit : ~0ull;
            }
        }
    }
};

// Run of `p` stones through `cell` along +-(dr,dc), `cell` counted as 1.
template <int NW>
__host__ __device__ __forceinline__ int az_run_length(const AzGame& g, const AzBoard<NW>& b, int p,
                                                      int r0, int c0, int dr, int dc) {
    int n = 1;
#pragma unroll
    for (int s = 1; s >= -1; s -= 2) {
        int r = r0 + s * dr, c = c0 + s * dc;
        while (r >= 0 && r < g.rows && c >= 0 && c < g.cols && b.has(p, r * g.cols + c)) {
            ++n;
            r += s * dr;
            c += s * dc;
        }
    }
    return n;
}

// true iff player p, who just played `cell`, has a run >= k through it (overlines count): walking the four lines.
template <int NW>
__host__ __device__ __forceinline__ bool az_is_win_walk(const AzGame& g, const AzBoard<NW>& b, int p, int cell) {
    const int r = cell / g.cols, c = cell - r * g.cols;
    return az_run_length(g, b, p, r, c, 0, 1) >= g.k || az_run_length(g, b, p, r, c, 1, 0) >= g.k ||
           az_run_length(g, b, p, r, c, 1, 1) >= g.k || az_run_length(g, b, p, r, c, 1, -1) >= g.k;
}

// The same predicate on a one-word board by shift and mask.  Along a direction with bit shift s, bit x of
//     t = m & (m >> s) & ... & (m >> (k-1) s)
// says that k stones of p sit at x, x + s, ...; win_start drops the x whose run would wrap into the next row, and `th` keeps
// the runs that contain `cell` (start = cell - i s, i < k; a start that wrapped is not in win_start either).  Same answer as
// the walk for EVERY position, also ones that already hold a run elsewhere (the reference only looks through the last move:
// games/connect4.py:72-98 and the two other check_winner).
__host__ __device__ __forceinline__ bool az_is_win_bits(const AzGame& g, uint64_t m, int cell) {
    const uint64_t bitc = 1ull << cell;
    const int sh[4] = {1, g.cols, g.cols + 1, g.cols - 1};
    bool win = false;
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        uint64_t t = m, th = bitc;
#pragma unroll
        for (int i = 1; i < 5; ++i) {
            if (i < g.k) {
                t &= m >> (i * sh[d]);
                th |= bitc >> (i * sh[d]);
            }
        }
        win = win || (t & th & g.win_start[d]) != 0ull;
    }
    return win;
}

template <int NW>
__host__ __device__ __forceinline__ bool az_is_win(const AzGame& g, const AzBoard<NW>& b, int p, int cell) {
    if (NW == 1) return az_is_win_bits(g, p ? b.w[1][0] : b.w[0][0], cell);
    return az_is_win_walk(g, b, p, cell);
}

// ---- legal SETS for the random rollouts of vanilla mode (ai/mcts.py:62-79; the order of the moves does not matter under a
// uniform draw): TicTacToe empty cells, Connect4 the lowest empty cell of every open column, Gomoku empty 8-neighbours of
// stones (the centre on an empty board).  Two forms: the cell walk (any board size) and shift-and-mask (one-word boards).
template <int NW>
__host__ __device__ inline int az_rollout_legal_walk(const AzGame& g, const AzBoard<NW>& b, int mc, uint64_t (&legal)[NW]) {
#pragma unroll
    for (int i = 0; i < NW; ++i) legal[i] = 0ull;
    int n = 0;
    for (int cell = 0; cell < g.state_dim; ++cell) {
        if (!b.empty(cell)) continue;
        bool ok;
        if (g.gid == AZ_GAME_TICTACTOE) {
            ok = true;
        } else if (g.gid == AZ_GAME_CONNECT4) {
            ok = cell + g.cols >= g.state_dim || !b.empty(cell + g.cols);
        } else {
            ok = false;
            const int r = cell / g.cols, c = cell - r * g.cols;
            for (int dr = -1; dr <= 1 && !ok; ++dr)
                for (int dc = -1; dc <= 1; ++dc) {
                    const int rr = r + dr, cc = c + dc;
                    if ((dr || dc) && rr >= 0 && rr < g.rows && cc >= 0 && cc < g.cols && !b.empty(rr * g.cols + cc)) { ok = true; break; }
                }
            if (mc == 0) ok = cell == (g.rows / 2) * g.cols + g.cols / 2;
        }
        if (ok) {
#pragma unroll
            for (int i = 0; i < NW; ++i)
                if (NW == 1 || i == (cell >> 6)) legal[i] |= 1ull << (cell & 63);
            ++n;
        }
    }
    return n;
}

struct AzRolloutMasks {
    uint64_t board, col0, coll, floor_row;
    int centre;
};
__host__ __device__ inline AzRolloutMasks az_rollout_masks(const AzGame& g) {
    AzRolloutMasks m{0ull, 0ull, 0ull, 0ull, 0};
    const uint64_t row = g.cols >= 64 ? ~0ull : (1ull << g.cols) - 1ull;
    for (int r = 0; r < g.rows; ++r) {
        m.board |= row << (r * g.cols);
        m.col0 |= 1ull << (r * g.cols);
        m.coll |= 1ull << (r * g.cols + g.cols - 1);
    }
    m.floor_row = row << ((g.rows - 1) * g.cols);
    m.centre = (g.rows / 2) * g.cols + g.cols / 2;
    return m;
}
__host__ __device__ __forceinline__ uint64_t az_rollout_legal_bits(const AzGame& g, const AzRolloutMasks& m, uint64_t occ, int mc) {
    if (g.gid == AZ_GAME_TICTACTOE) return ~occ & m.board;
    if (g.gid == AZ_GAME_CONNECT4) return ~occ & m.board & ((occ >> g.cols) | m.floor_row);
    if (mc == 0) return 1ull << m.centre;
    const uint64_t h = (occ | ((occ << 1) & ~m.col0) | ((occ >> 1) & ~m.coll)) & m.board;   // a bit past the last cell must not come back down
    return (h | (h << g.cols) | (h >> g.cols)) & ~occ & m.board;
}

// Index into the policy vector of a move stored as a cell: games/*: get_action_idx.
__host__ __device__ __forceinline__ int az_action_index(const AzGame& g, int cell) {
    return g.gid == AZ_GAME_CONNECT4 ? cell % g.cols : cell;
}

// ---- CPython set emulation -------------------------------------------------------------
// One probe sequence of set_add_entry / set_insert_clean: home slot, up to 9 linear probes
// when they fit below the mask, then i = 5i + 1 + (perturb >>= 5).
__host__ __device__ __forceinline__ bool az_set_insert(uint8_t* tab, uint32_t mask, uint8_t x, uint64_t h,
                                                       bool check_dup) {
    uint64_t perturb = h;
    uint64_t i = h & mask;
    for (;;) {
        const int lim = (i + 9 <= mask) ? 9 : 0;
        for (int j = 0; j <= lim; ++j) {
            const uint8_t e = tab[i + j];
            if (e == 0xFF) {
                tab[i + j] = x;
                return true;
            }
            if (check_dup && e == x) return false;
        }
        perturb >>= 5;
        i = (i * 5 + 1 + perturb) & mask;
    }
}

// Legal moves (as cells) in the reference's child order, computed by ONE thread.
// tab_a / tab_b: scratch of g.set_slots bytes each (Gomoku only).  Returns the count.
template <int NW>
__host__ __device__ inline int az_legal_moves(const AzGame& g, const AzBoard<NW>& b,
                                              const uint64_t* __restrict__ cell_hash, uint8_t* tab_a,
                                              uint8_t* tab_b, uint8_t* out) {
    int n = 0;
    if (g.gid == AZ_GAME_TICTACTOE) {
        for (int cell = 0; cell < g.state_dim; ++cell)
            if (b.empty(cell)) out[n++] = (uint8_t)cell;
        return n;
    }
    if (g.gid == AZ_GAME_CONNECT4) {
        for (int c = 0; c < g.cols; ++c) {
            if (!b.empty(c)) continue;                      // top cell taken: column full
            int r = g.rows - 1;
            while (!b.empty(r * g.cols + c)) --r;           // lowest empty row (row 0 = top)
            out[n++] = (uint8_t)(r * g.cols + c);
        }
        return n;
    }
    // Gomoku: empty 8-neighbours of every stone, inserted in the reference's scan order.  Adding a
    // key the set already holds changes nothing (no fill, no resize), so only the FIRST occurrence
    // of a cell reaches the table: `blocked` = stones | cells already inserted is a bit test.
    // Pass 1 lists the first occurrences in scan order (into `out`, reused below), pass 2 replays
    // them through the table.
    // Shortcut (boards of <= 64 cells).  The table size a CPython set ends with depends only on the number of keys
    // (8 slots up to 4 keys, 32 up to 18, 128 up to 76: growth is triggered by the fill count alone, nothing is ever
    // deleted).  At the last resize the keys are reinserted into an EMPTY table and every later key is added to that
    // table, so whenever the keys' home slots `hash & mask` in the final table are pairwise distinct, no probe ever
    // collides there and every key sits in its home slot whatever the insertion history was: iteration order = keys
    // sorted by home slot.  The candidate SET is the 8-neighbour dilation of the stones, a handful of shifts; a
    // collision in the final table falls through to the full emulation below.
#ifndef AZ_NO_LEGAL_SHORTCUT
    if (NW == 1) {
        const uint64_t occ = b.w[0][0] | b.w[1][0];
        uint64_t col0 = 0ull;
        for (int r = 0; r < g.rows; ++r) col0 |= 1ull << (r * g.cols);
        const uint64_t colr = col0 << (g.cols - 1);
        const uint64_t full = g.state_dim >= 64 ? ~0ull : (1ull << g.state_dim) - 1ull;
        const uint64_t h = occ | ((occ & ~colr) << 1) | ((occ & ~col0) >> 1);
        const uint64_t cand = (h | (h << g.cols) | (h >> g.cols)) & ~occ & full;
        if (cand == 0ull) {                                   // no candidate (empty board): centre (gomoku.py:104-105)
            out[0] = (uint8_t)((g.rows / 2) * g.cols + g.cols / 2);
            return 1;
        }
        const int nk = AZ_POPC64(cand);
        const uint32_t fm = nk <= 4 ? 7u : (nk <= 18 ? 31u : 127u);
        uint64_t s0 = 0ull, s1 = 0ull;                        // occupied home slots 0..63 / 64..127
        bool clash = false;
        for (uint64_t w = cand; w; w &= w - 1) {
            const uint32_t slot = (uint32_t)cell_hash[AZ_CTZ64(w)] & fm;
            const uint64_t bit = 1ull << (slot & 63);
            if (slot < 64) { clash |= (s0 & bit) != 0ull; s0 |= bit; }
            else { clash |= (s1 & bit) != 0ull; s1 |= bit; }
        }
        if (!clash) {
            for (uint64_t w = cand; w; w &= w - 1) {
                const int x = AZ_CTZ64(w);
                const uint32_t slot = (uint32_t)cell_hash[x] & fm;
                const int pos = slot < 64 ? AZ_POPC64(s0 & ((1ull << slot) - 1ull))
                                          : AZ_POPC64(s0) + AZ_POPC64(s1 & ((1ull << (slot - 64)) - 1ull));
                out[pos] = (uint8_t)x;
            }
            return nk;
        }
    }
#endif
    uint64_t blocked[NW];
#pragma unroll
    for (int wi = 0; wi < NW; ++wi) blocked[wi] = b.w[0][wi] | b.w[1][wi];
    int ncand = 0;
    const int ddr[4] = {0, 1, 1, 1}, ddc[4] = {1, 0, 1, -1};
#pragma unroll
    for (int wi = 0; wi < NW; ++wi) {
        uint64_t occ = b.w[0][wi] | b.w[1][wi];
        while (occ) {
            const int cell = wi * 64 + AZ_CTZ64(occ);
            occ &= occ - 1;
            const int r = cell / g.cols, c = cell - r * g.cols;
#pragma unroll
            for (int d = 0; d < 4; ++d) {
#pragma unroll
                for (int s = 1; s >= -1; s -= 2) {
                    const int rr = r + s * ddr[d], cc = c + s * ddc[d];
                    if (rr < 0 || rr >= g.rows || cc < 0 || cc >= g.cols) continue;
                    const int x = rr * g.cols + cc;
                    bool taken = false;
                    if (NW == 1) {
                        taken = (blocked[0] >> x) & 1;
                        blocked[0] |= 1ull << x;
                    } else {
#pragma unroll
                        for (int wj = 0; wj < NW; ++wj) {
                            if (wj == (x >> 6)) {
                                taken = (blocked[wj] >> (x & 63)) & 1;
                                blocked[wj] |= 1ull << (x & 63);
                            }
                        }
                    }
                    if (!taken) out[ncand++] = (uint8_t)x;
                }
            }
        }
    }
    uint8_t* tab = tab_a;
    uint8_t* spare = tab_b;
    uint32_t mask = 7;
    int fill = 0;
    for (int i = 0; i < 8; ++i) tab[i] = 0xFF;
    for (int k = 0; k < ncand; ++k) {
        const uint8_t x = out[k];
        az_set_insert(tab, mask, x, cell_hash[x], false);
        ++fill;
        if ((uint32_t)fill * 5 >= mask * 3) {               // grow: smallest 2^k > 4*used
            uint32_t newsize = 8;
            while (newsize <= (uint32_t)fill * 4) newsize <<= 1;
            for (uint32_t i = 0; i < newsize; ++i) spare[i] = 0xFF;
            for (uint32_t i = 0; i <= mask; ++i)
                if (tab[i] != 0xFF) az_set_insert(spare, newsize - 1, tab[i], cell_hash[tab[i]], false);
            uint8_t* t = tab;
            tab = spare;
            spare = t;
            mask = newsize - 1;
        }
    }
    if (fill == 0) {                                        // empty board: centre (gomoku.py:104-105)
        out[0] = (uint8_t)((g.rows / 2) * g.cols + g.cols / 2);
        return 1;
    }
    for (uint32_t i = 0; i <= mask; ++i)
        if (tab[i] != 0xFF) out[n++] = tab[i];
    return n;
}
