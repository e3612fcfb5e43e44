// Device-side view of a forest of concurrent search trees (one per game) in HBM.
//
// Layout (SoA across node fields, grouped by access pattern):
//   nodes[tree*cap + i]  AzNode = {hot {N i32, Q f32, P f32, V u32} | meta {child_base, action, nchild, status} | W f64}
//                        32 B = one sector: two 16-byte loads per child in PUCT, one sector per backup / expansion
// Children of a node are contiguous (bump allocator per tree), in the reference's child
// order (= order of Game.get_valid_moves, ai/node.py:55-59).  Node 0 is the root.
#pragma once
#include <cuda_runtime.h>
#include "az_types.h"

// counters in AzForestDev::stats
#define AZ_STAT_SIMS 0        // completed simulations (iterations of ai/mcts.py:16)
#define AZ_STAT_TERMINAL 1    // ... of which ended on a terminal node
#define AZ_STAT_EXPAND 2      // expansions (= evaluations consumed)
#define AZ_STAT_CHILDREN 3    // children created
#define AZ_STAT_SELECT 4      // select steps (edges walked)
#define AZ_STAT_CHILD_READS 5 // children scored in select
#define AZ_STAT_COLLIDE 6     // virtual-loss batches cut short by a pending leaf
#define AZ_STAT_CACHE_HIT 7   // leaves answered by the evaluation cache (no evaluator row)
#define AZ_STAT_COUNT 8

// Device evaluation cache: the GPU analogue of MCTS.cache (ai/mcts.py:7,37-51) -- canonical position -> (softmax priors
// f32[A], value).  Open addressing in HBM, 8-entry windows (one 128-byte line of keys, probed by 8 lanes at once).
// A leaf that finds its position takes the cached answer instead of an evaluator row; leaves of different trees that
// reach the same new position in one launch share ONE row (the first claims the entry, the others wait for it).
// state[e]: bits 31:30 kind (0 empty, 1 locked: key being written, 2 pending: evaluation in flight, 3 ready),
//           bits 7:0 search number of the last use; entries not used for AZ_CACHE_STALE searches may be overwritten.
// Boards of <= 64 cells only (16-byte keys; the side to move of the 3-plane games rides in bit 63 of `theirs`).
#define AZ_CACHE_WINDOW 8
#define AZ_CACHE_STALE 2
#define AZ_CK_EMPTY 0u
#define AZ_CK_LOCKED 1u
#define AZ_CK_PENDING 2u
#define AZ_CK_READY 3u
struct AzCacheDev {
    uint32_t cap;             // entries, power of two; 0 = cache off
    ulonglong2* keys;         // [cap] {mine, theirs | player << 63}
    uint32_t* state;          // [cap]
    float* values;            // [cap]
    float* priors;            // [cap][A]
    uint32_t* epoch;          // device scalar: search number (az_begin_move_kernel increments it)
    uint32_t* leaf_entry;     // [max_leaves] entry claimed by the batch row, 0xffffffff = none
    int* n_prev;              // device scalar: rows of the previous batch still to be published
};

struct AzForestDev {
    AzGame g;
    int n_trees;
    int cap;          // nodes per tree
    int iters;        // simulations per move (mcts_iterations)
    int vl;           // leaves per tree per step (1 = the reference's sequential search)
    int max_depth;    // path entries per leaf
    int max_leaves;   // n_trees * vl
    int use_noise;    // Dirichlet noise mixed into the root priors (utils.py:12-27)
    int lp;           // pending leaves of a tree expanded in parallel (<= AZ_LP, <= vl)
    int ucb1;         // vanilla MCTS (model=None): UCB1 selection in f64 (utils.py:36,43), priors unused
    const double* log_tab;   // [iters + vl + 2] ln(N) computed on the host with the C library (= math.log of the reference)

    AzNode* nodes;    // [n_trees][cap]

    // per tree
    uint64_t* root_words;    // [n_trees][2*nw]  side-0 words then side-1 words
    uint8_t* root_player;    // side to move at the root
    uint16_t* root_mc;       // moves played so far
    uint8_t* active;         // tree takes part in the search
    int* n_nodes;            // bump allocator
    int* sims_done;
    uint8_t* n_pending;
    uint8_t* done_flag;

    // pending leaves [n_trees][vl]
    uint32_t* path;          // [n_trees][vl][max_depth]
    uint16_t* path_len;
    int* leaf_slot;
    uint64_t* pend_words;    // [n_trees][vl][2*nw]  leaf position (side 0, side 1)
    uint8_t* pend_player;
    uint16_t* pend_mc;
    uint32_t* pend_leaf;     // node index of the pending leaf (= last path entry)
    uint16_t* pend_action;   // cell of the move that led to the leaf (meta.action, rewritten at expansion)

    // evaluation batch (compacted)
    int* n_leaves;           // device counter, zeroed before every advance
    int* n_done;             // trees that finished all their simulations
    uint64_t* leaf_words;    // [max_leaves][2*nw]  canonical: mine, theirs
    uint8_t* leaf_player;    // side to move (plane 2 of the 3-plane games)
    int* leaf_tree;
    float* priors;           // [max_leaves][A]  exp(l)/sum(exp(l)), unmasked (ai/mcts.py:49)
    float* values;           // [max_leaves]

    // root noise
    const double* root_noise;  // [n_trees][A] or nullptr
    double* root_prior64;      // [n_trees][max_children]  f64 priors of root children

    const uint64_t* cell_hash; // CPython hash((r,c)) per cell (Gomoku child order)
    unsigned long long* stats; // AZ_STAT_*
    AzCacheDev cache;
    int max_rounds;            // vl == 1 with the cache: simulations a tree may run in one launch while its leaves hit the cache
    int no_best;               // A/B switch (AZ_NO_BEST=1): vl == 1 searches scan and score per level instead of following cached best children
};
