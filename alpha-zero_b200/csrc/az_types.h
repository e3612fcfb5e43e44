// Shared POD types of the B200 self-play engine (host + device).
#pragma once
#include <stdint.h>

#define AZ_GAME_TICTACTOE 0
#define AZ_GAME_CONNECT4 1
#define AZ_GAME_GOMOKU 2

#define AZ_MAX_CELLS 225          // 15 x 15
#define AZ_MAX_WORDS 4            // ceil(225 / 64)

// node status (meta.status)
#define AZ_ST_LEAF 0              // never expanded
#define AZ_ST_EXPANDED 1
#define AZ_ST_WIN 2               // terminal: the player who moved in has won   (ai/mcts.py:26-29)
#define AZ_ST_DRAW 3              // terminal: board full                         (ai/mcts.py:30-32)
#define AZ_ST_PENDING 4           // selected in this step, evaluation in flight (virtual-loss mode)

// Rule set of one game; the reference passes a class with these as attributes
// (games/tictactoe.py:10-12, games/connect4.py:8-10, games/gomoku.py:10-13).
struct AzGame {
    int gid;          // AZ_GAME_*
    int rows, cols;
    int k;            // run length that wins (3 / 4 / 5)
    int action_dim;   // 9 / 7 / rows*cols
    int state_dim;    // rows*cols
    int planes;       // 3 / 3 / 2 feature planes of the network input
    int nw;           // 64-bit words per side bitboard
    int max_children; // upper bound on legal moves of a node
    int set_slots;    // largest CPython-set table the Gomoku legal-move emulation can reach
    // boards of one word: cells a run of k stones may START at along (0,1), (1,0), (1,1), (1,-1) without leaving its rows
    // (column bounds only: bits beyond the board are never set), for the shift-and-mask win test of az_rules.cuh
    uint64_t win_start[4];
};

// Hot per-node record: everything PUCT reads for one child, one 16-byte load.
struct __attribute__((aligned(16))) AzHot {
    int N;            // ai/node.py:33  visit count
    float Q;          // (float)(W / N) exactly as utils.py:40 forms it, refreshed by every backup (unused while N == 0)
    float P;          // ai/node.py:36  prior (fp32, gathered from the unmasked softmax)
    uint32_t V;       // virtual-loss count (vl > 1 only)
};

// Cold per-node record, read only for the selected child.
struct __attribute__((aligned(8))) AzMeta {
    uint32_t child_base;  // local index of first child (children are contiguous)
    uint16_t action;      // cell index r*cols+c of the move that led here (prevAction)
    uint8_t nchild;
    uint8_t status;       // AZ_ST_*
};

// One node = one 32-byte sector: a child scan, a backup and an expansion each touch exactly one
// sector per node and one address stream per tree (one TLB entry instead of three).
struct __attribute__((aligned(32))) AzNode {
    AzHot hot;        // offset 0:  what PUCT scores
    AzMeta meta;      // offset 16: where the children are
    double W;         // offset 24: ai/node.py:34 value sum (viewpoint of the player who moved in), f64 in visitation order
};

// One leaf handed to the evaluator: canonical side bitboards (mine = side to move).
struct AzLeafIn {
    uint64_t mine[AZ_MAX_WORDS];
    uint64_t theirs[AZ_MAX_WORDS];
};
