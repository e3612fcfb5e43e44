// Host-side differential test of the shift-and-mask legal SETS of the vanilla-mode rollouts (az_rollout_legal_bits) against the
// cell walk (az_rollout_legal_walk) on one-word boards: random positions of every density incl. the empty board, all games.
// Built and run by tests/test_rules_bits.py (nvcc, host code only: no GPU needed).
#include <cstdio>
#include <cstdlib>
#include "az_types.h"
#include "az_rules.cuh"
static void make(int game, int rows, int cols, AzGame* out) {
    AzGame g{}; g.gid = game; g.rows = rows; g.cols = cols;
    g.k = game == 0 ? 3 : (game == 1 ? 4 : 5); g.state_dim = rows * cols; g.nw = 1;
    *out = g;
}
int main() {
    srand(2);
    long long n = 0;
    int shapes[][3] = {{0,3,3},{1,6,7},{2,5,5},{2,6,6},{2,7,7},{2,8,8}};
    for (auto& sh : shapes) {
        AzGame g; make(sh[0], sh[1], sh[2], &g);
        const AzRolloutMasks m = az_rollout_masks(g);
        for (int it = 0; it < 300000; ++it) {
            AzBoard<1> b; b.w[0][0] = b.w[1][0] = 0;
            const int cells = g.state_dim;
            const int dens = rand() % 5;                       // 0: empty board
            for (int c = 0; c < cells; ++c) {
                if (sh[0] == 1) continue;                      // Connect4 below: stones stack from the floor
                int r = rand() % 8; if (r < dens) b.w[0][0] |= 1ull << c; else if (r < 2 * dens) b.w[1][0] |= 1ull << c;
            }
            if (sh[0] == 1)
                for (int c = 0; c < g.cols; ++c) { const int hgt = rand() % (g.rows + 1); for (int k = 0; k < hgt; ++k) b.w[rand() & 1][0] |= 1ull << ((g.rows - 1 - k) * g.cols + c); }
            const int mc = __builtin_popcountll(b.w[0][0] | b.w[1][0]);
            uint64_t walk[1];
            const int nw = az_rollout_legal_walk<1>(g, b, mc, walk);
            const uint64_t bits = az_rollout_legal_bits(g, m, b.w[0][0] | b.w[1][0], mc);
            if (walk[0] != bits || nw != __builtin_popcountll(bits)) {
                printf("MISMATCH game %d %dx%d occ %llx walk %llx bits %llx\n", sh[0], sh[1], sh[2], (unsigned long long)(b.w[0][0] | b.w[1][0]), (unsigned long long)walk[0], (unsigned long long)bits);
                return 1;
            }
            ++n;
        }
    }
    printf("identical legal sets on %lld positions\n", n);
    return 0;
}
