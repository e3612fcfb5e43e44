// Host-side differential test of the shift-and-mask win test (az_is_win_bits) against the line walk (az_is_win_walk) it
// replaces on one-word boards: random positions of every density, random last move, all one-word board shapes.
// Built and run by tests/test_rules_bits.py (nvcc, host code only: no GPU needed).
#include <cstdio>
#include <cstdlib>
#include <string>
#include "az_types.h"
#include "az_rules.cuh"
int az_fail(int c, const std::string& m) { printf("fail %s\n", m.c_str()); return c; }
#define AZ_EINVAL 1
#define AZ_OK 0
int make(int game, int rows, int cols, AzGame* out) {
    AzGame g{}; g.gid = game; g.rows = rows; g.cols = cols;
    g.k = game == 0 ? 3 : (game == 1 ? 4 : 5); g.state_dim = rows * cols; g.nw = 1;
    const int dcs[4] = {1, 0, 1, -1};
    for (int d = 0; d < 4; ++d) { g.win_start[d] = 0;
        for (int r = 0; r < rows; ++r) for (int c = 0; c < cols; ++c) { const int last = c + dcs[d] * (g.k - 1); if (last >= 0 && last < cols) g.win_start[d] |= 1ull << (r * cols + c); } }
    *out = g; return 0;
}
int main() {
    srand(1);
    long long n = 0, wins = 0;
    int shapes[][3] = {{0,3,3},{1,6,7},{2,5,5},{2,6,6},{2,7,7},{2,8,8}};
    for (auto& sh : shapes) {
        AzGame g; make(sh[0], sh[1], sh[2], &g);
        for (int it = 0; it < 400000; ++it) {
            AzBoard<1> b; b.w[0][0] = b.w[1][0] = 0;
            const int cells = g.state_dim;
            const int dens = 1 + rand() % 4;
            for (int c = 0; c < cells; ++c) { int r = rand() % 8; if (r < dens) b.w[0][0] |= 1ull << c; else if (r < 2 * dens) b.w[1][0] |= 1ull << c; }
            const int cell = rand() % cells, p = rand() & 1;
            b.w[p][0] |= 1ull << cell; b.w[1 - p][0] &= ~(1ull << cell);
            const bool a = az_is_win_walk<1>(g, b, p, cell), bb = az_is_win<1>(g, b, p, cell);
            if (a != bb) { printf("MISMATCH game %d %dx%d cell %d p %d walk %d bits %d m=%llx\n", sh[0], sh[1], sh[2], cell, p, a, bb, (unsigned long long)b.w[p][0]); return 1; }
            ++n; wins += a;
        }
    }
    printf("identical on %lld positions (%lld wins)\n", n, wins);
    return 0;
}
