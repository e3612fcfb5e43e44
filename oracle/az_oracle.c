/* az_oracle.c -- CPU oracle (plain C) of the self-play hot path.  TEST INFRASTRUCTURE.
 *
 * A sequential, one-tree-at-a-time restatement of the reference's search so that parity
 * tests can check thousands of GPU trees in seconds.  It follows the same reference lines
 * as oracle/az_oracle.py and is itself pinned against tests/golden/ (reference outputs):
 *   MCTS.mcts            ai/mcts.py:11-60        Node.select/expand/backup  ai/node.py:42-74
 *   PUCT                 utils.py:29-44          Dirichlet mixing           utils.py:24-25
 *   rules                games/tictactoe.py, games/connect4.py, games/gomoku.py
 * Boards are byte arrays (0 empty, 1 side 0, 2 side 1) -- deliberately NOT the bitboards of
 * the CUDA path.  Build: see oracle/Makefile (-O2 -ffp-contract=off: no fused multiply-add,
 * the scores must round exactly as NumPy does).
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXC 225

typedef struct {
    int gid, rows, cols, k, A, planes, cells;
} Game;

static int game_init(Game* g, int gid, int rows, int cols) {
    g->gid = gid; g->rows = rows; g->cols = cols; g->cells = rows * cols;
    if (gid == 0) { g->k = 3; g->A = 9; g->planes = 3; return rows == 3 && cols == 3 ? 0 : -1; }
    if (gid == 1) { g->k = 4; g->A = 7; g->planes = 3; return rows == 6 && cols == 7 ? 0 : -1; }
    if (gid == 2) { g->k = 5; g->A = rows * cols; g->planes = 2; return rows == cols && rows <= 15 ? 0 : -1; }
    return -1;
}

/* ---- rules ------------------------------------------------------------------------- */
static int run_len(const Game* g, const uint8_t* b, int stone, int r0, int c0, int dr, int dc) {
    int n = 1;
    for (int s = 1; s >= -1; s -= 2) {
        int r = r0 + s * dr, c = c0 + s * dc;
        while (r >= 0 && r < g->rows && c >= 0 && c < g->cols && b[r * g->cols + c] == stone) {
            n++; r += s * dr; c += s * dc;
        }
    }
    return n;
}

/* check_winner: gomoku.py:65-91 etc.  1 if `player` has >= k through `cell`. */
static int is_win(const Game* g, const uint8_t* b, int player, int cell) {
    static const int D[4][2] = {{0, 1}, {1, 0}, {1, 1}, {1, -1}};
    int r = cell / g->cols, c = cell % g->cols;
    for (int d = 0; d < 4; d++)
        if (run_len(g, b, player + 1, r, c, D[d][0], D[d][1]) >= g->k) return 1;
    return 0;
}

/* CPython tuple hash of (r, c): Objects/tupleobject.c (3.8+). */
static uint64_t tuple2_hash(uint64_t r, uint64_t c) {
    const uint64_t P1 = 11400714785074694791ULL, P2 = 14029467366897019727ULL, P5 = 2870177450012600261ULL;
    uint64_t acc = P5, lane[2] = {r, c};
    for (int i = 0; i < 2; i++) {
        acc += lane[i] * P2;
        acc = (acc << 31) | (acc >> 33);
        acc *= P1;
    }
    acc += 2ULL ^ (P5 ^ 3527539ULL);
    return acc == ~0ULL ? 1546275796ULL : acc;
}

typedef struct { int16_t* slot; uint32_t mask; int fill; } PySet;

/* set_add_entry / set_insert_clean probe: Objects/setobject.c (CPython 3.12). */
static int pyset_probe(int16_t* tab, uint32_t mask, int x, uint64_t h, int check_dup) {
    uint64_t perturb = h, i = h & mask;
    for (;;) {
        int lim = (i + 9 <= mask) ? 9 : 0;
        for (int j = 0; j <= lim; j++) {
            if (tab[i + j] < 0) { tab[i + j] = (int16_t)x; return 1; }
            if (check_dup && tab[i + j] == x) return 0;
        }
        perturb >>= 5;
        i = (i * 5 + 1 + perturb) & mask;
    }
}

static void pyset_add(PySet* s, int x, const uint64_t* hashes) {
    if (!pyset_probe(s->slot, s->mask, x, hashes[x], 1)) return;
    s->fill++;
    if ((uint32_t)s->fill * 5 < s->mask * 3) return;
    uint32_t newsize = 8;
    while (newsize <= (uint32_t)s->fill * 4) newsize <<= 1;     /* set_table_resize(used*4) */
    int16_t* nt = (int16_t*)malloc(newsize * sizeof(int16_t));
    for (uint32_t i = 0; i < newsize; i++) nt[i] = -1;
    for (uint32_t i = 0; i <= s->mask; i++)
        if (s->slot[i] >= 0) pyset_probe(nt, newsize - 1, s->slot[i], hashes[s->slot[i]], 0);
    free(s->slot);
    s->slot = nt;
    s->mask = newsize - 1;
}

/* get_valid_moves in child order; returns count, cells into out[]. */
static int legal_moves(const Game* g, const uint8_t* b, const uint64_t* hashes, int* out) {
    int n = 0;
    if (g->gid == 0) {                                          /* tictactoe.py:81-83 */
        for (int i = 0; i < 9; i++) if (!b[i]) out[n++] = i;
        return n;
    }
    if (g->gid == 1) {                                          /* connect4.py:43-53 */
        for (int c = 0; c < g->cols; c++) {
            if (b[c]) continue;
            for (int r = g->rows - 1; r >= 0; r--)
                if (!b[r * g->cols + c]) { out[n++] = r * g->cols + c; break; }
        }
        return n;
    }
    static const int D[4][2] = {{0, 1}, {1, 0}, {1, 1}, {1, -1}};   /* gomoku.py:93-106 */
    PySet s;
    s.slot = (int16_t*)malloc(8 * sizeof(int16_t));
    for (int i = 0; i < 8; i++) s.slot[i] = -1;
    s.mask = 7; s.fill = 0;
    for (int r = 0; r < g->rows; r++)
        for (int c = 0; c < g->cols; c++) {
            if (!b[r * g->cols + c]) continue;
            for (int d = 0; d < 4; d++)
                for (int sg = 1; sg >= -1; sg -= 2) {
                    int rr = r + sg * D[d][0], cc = c + sg * D[d][1];
                    if (rr < 0 || rr >= g->rows || cc < 0 || cc >= g->cols) continue;
                    if (b[rr * g->cols + cc]) continue;
                    pyset_add(&s, rr * g->cols + cc, hashes);
                }
        }
    if (s.fill == 0) out[n++] = (g->rows / 2) * g->cols + g->cols / 2;
    else for (uint32_t i = 0; i <= s.mask; i++) if (s.slot[i] >= 0) out[n++] = s.slot[i];
    free(s.slot);
    return n;
}

static int action_index(const Game* g, int cell) { return g->gid == 1 ? cell % g->cols : cell; }

/* ---- evaluators ---------------------------------------------------------------------- */
static uint64_t sm64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

/* canonical words: mine (side to move) then theirs; bit r*cols+c. */
static void canon_words(const Game* g, const uint8_t* b, int player, uint64_t* mine, uint64_t* theirs) {
    int nw = (g->cells + 63) / 64;
    for (int i = 0; i < nw; i++) mine[i] = theirs[i] = 0;
    for (int i = 0; i < g->cells; i++) {
        if (b[i] == player + 1) mine[i >> 6] |= 1ULL << (i & 63);
        else if (b[i]) theirs[i >> 6] |= 1ULL << (i & 63);
    }
}

typedef struct {
    int n, nw, A, nslots;
    const uint64_t* keys;   /* [n][2*nw+1] */
    const float* priors;    /* [n][A] */
    const float* values;
    int* slots;
} Table;

static uint64_t table_hash(const uint64_t* k, int len) {
    uint64_t h = 0x1234;
    for (int i = 0; i < len; i++) h = sm64(h ^ k[i]);
    return h;
}

void* azo_table_create(int n, int nw, int A, const uint64_t* keys, const float* priors, const float* values) {
    Table* t = (Table*)malloc(sizeof(Table));
    t->n = n; t->nw = nw; t->A = A; t->keys = keys; t->priors = priors; t->values = values;
    t->nslots = 16;
    while (t->nslots < 2 * n) t->nslots <<= 1;
    t->slots = (int*)malloc(sizeof(int) * t->nslots);
    for (int i = 0; i < t->nslots; i++) t->slots[i] = -1;
    int kl = 2 * nw + 1;
    for (int i = 0; i < n; i++) {
        uint64_t h = table_hash(keys + (size_t)i * kl, kl) & (t->nslots - 1);
        while (t->slots[h] >= 0) {
            if (!memcmp(keys + (size_t)t->slots[h] * kl, keys + (size_t)i * kl, kl * 8)) break;  /* first wins */
            h = (h + 1) & (t->nslots - 1);
        }
        if (t->slots[h] < 0) t->slots[h] = i;
    }
    return t;
}

void azo_table_destroy(void* p) {
    if (!p) return;
    free(((Table*)p)->slots);
    free(p);
}

typedef struct {
    uint64_t seed; int qbits; const Table* table; long misses;
} Eval;

static void evaluate(const Game* g, Eval* ev, const uint8_t* b, int player, float* priors, float* value) {
    uint64_t key[2 * 4 + 1];
    int nw = (g->cells + 63) / 64;
    canon_words(g, b, player, key, key + nw);
    key[2 * nw] = (uint64_t)player;
    if (ev->table) {
        const Table* t = ev->table;
        int kl = 2 * nw + 1;
        if (g->planes != 3) key[2 * nw] = 0;
        uint64_t h = table_hash(key, kl) & (t->nslots - 1);
        while (t->slots[h] >= 0) {
            if (!memcmp(t->keys + (size_t)t->slots[h] * kl, key, kl * 8)) {
                memcpy(priors, t->priors + (size_t)t->slots[h] * g->A, g->A * sizeof(float));
                *value = t->values[t->slots[h]];
                return;
            }
            h = (h + 1) & (t->nslots - 1);
        }
        ev->misses++;
        for (int a = 0; a < g->A; a++) priors[a] = 0.f;
        *value = 0.f;
        return;
    }
    uint64_t h = sm64(ev->seed);
    for (int i = 0; i < nw; i++) h = sm64(h ^ key[i]);
    h = sm64(h ^ 0xA5A5A5A5A5A5A5A5ULL);
    for (int i = 0; i < nw; i++) h = sm64(h ^ key[nw + i]);
    if (g->planes == 3) h = sm64(h ^ (uint64_t)player);
    int lg = 0;
    while ((1 << lg) < g->A) lg++;
    float scale = ldexpf(1.0f, -(ev->qbits + lg));
    for (int a = 0; a < g->A; a++) {
        uint64_t x = sm64(h + (uint64_t)a + 1);
        priors[a] = (float)(((x >> 40) & ((1ULL << ev->qbits) - 1)) + 1) * scale;
    }
    uint64_t x = sm64(h + 0xFFFFULL);
    *value = (float)((int)((x >> 40) & 0xFFFF) - 32768) * (1.0f / 32768.0f);
}

/* ---- tree ------------------------------------------------------------------------------ */
typedef struct {
    int parent, first, nchild;   /* children are consecutive in the arena, in get_valid_moves order */
    int visit;                   /* node.py:33 */
    double W;                    /* node.py:34 */
    float P;                     /* node.py:36 (np.float32) */
    double P64;                  /* root children under Dirichlet noise (np.float64) */
    int cell, player, mc;        /* prevAction, currentPlayer, move_count */
    int vloss, pending;          /* virtual-loss extension (0 in the reference's sequential search) */
} ONode;

/* utils.py:29-44 'network' branch under NumPy>=2 scalar promotion. */
static float puct32(double W, float P, int n, float s) {
    float t = P * s;
    float u = t / (float)(n + 1);
    if (n == 0) return u;
    float q = (float)(W / (double)n);
    return q + u;
}

static double puct64(double W, double P, int n, double s) {
    double u = (P * s) / (double)(n + 1);
    if (n == 0) return u;
    return W / (double)n + u;
}

/* utils.py:29-44 'normal' branch (model=None, vanilla MCTS): Python floats = f64; an unvisited child scores
 * sqrt(2 ln Np / 1), not +inf.  Selected by azo_set_ucb1(1); the priors are then unused (node.py:58). */
static int g_ucb1 = 0;
void azo_set_ucb1(int on) { g_ucb1 = on; }
static double ucb1(double W, int n, double two_ln_np) {
    double e = sqrt(two_ln_np / (double)(n + 1));
    if (n == 0) return e;
    return W / (double)n + e;
}

/* node.py:42-47: first maximum in child order. */
static int select_child(const ONode* nd, int node, int noisy_root) {
    const ONode* p = &nd[node];
    int pn = p->visit + p->vloss, best = p->first;
    if (g_ucb1) {
        double s = 2 * log((double)pn), bs = 0;
        for (int c = 0; c < p->nchild; c++) {
            const ONode* ch = &nd[p->first + c];
            double sc = ucb1(ch->W - (double)ch->vloss, ch->visit + ch->vloss, s);
            if (c == 0 || sc > bs) { bs = sc; best = p->first + c; }
        }
    } else if (noisy_root) {
        double s = sqrt((double)pn), bs = 0;
        for (int c = 0; c < p->nchild; c++) {
            const ONode* ch = &nd[p->first + c];
            double sc = puct64(ch->W - (double)ch->vloss, ch->P64, ch->visit + ch->vloss, s);
            if (c == 0 || sc > bs) { bs = sc; best = p->first + c; }
        }
    } else {
        float s = (float)sqrt((double)pn), bs = 0;
        for (int c = 0; c < p->nchild; c++) {
            const ONode* ch = &nd[p->first + c];
            float sc = puct32(ch->W - (double)ch->vloss, ch->P, ch->visit + ch->vloss, s);
            if (c == 0 || sc > bs) { bs = sc; best = p->first + c; }
        }
    }
    return best;
}

/* out arrays sized for max children; stats[8] as AZ_STAT_* of the CUDA path. Returns the
 * number of root children, or <0 on error. */
int azo_search(int gid, int rows, int cols, const uint8_t* cells_in, int player0, int mc0, int iters, int vl,
               uint64_t seed, int qbits, const double* noise, void* table, int* out_cells, int* out_visits,
               double* out_W, double* out_P, int* root_visit, double* root_value, uint64_t* stats) {
    Game g;
    if (game_init(&g, gid, rows, cols)) return -1;
    if (vl < 1) vl = 1;
    int maxch = g.gid == 0 ? 9 : g.gid == 1 ? 7 : g.cells;
    ONode* nd = (ONode*)calloc((size_t)1 + (size_t)iters * maxch, sizeof(ONode));
    uint64_t hashes[MAXC];
    for (int i = 0; i < g.cells; i++) hashes[i] = tuple2_hash(i / g.cols, i % g.cols);
    uint8_t board[MAXC];
    memcpy(board, cells_in, g.cells);
    Eval ev = {seed, qbits, (const Table*)table, 0};
    int n_nodes = 1;
    nd[0].parent = -1; nd[0].player = player0; nd[0].mc = mc0; nd[0].cell = -1;
    int* path = (int*)malloc(sizeof(int) * (size_t)vl * (g.cells + 2));
    int* plen = (int*)malloc(sizeof(int) * vl);
    float* ppri = (float*)malloc(sizeof(float) * (size_t)vl * g.A);
    float* pval = (float*)malloc(sizeof(float) * vl);
    uint8_t* pboard = (uint8_t*)malloc((size_t)vl * MAXC);
    int moves[MAXC];
    uint64_t st[8] = {0};
    int sims = 0;
    const int stride = g.cells + 2;
    while (sims < iters) {
        /* phase B first (nothing is pending at loop entry): select up to vl leaves */
        /* a step also closes after vl terminal descents (bounds the work of one step; at vl == 1 the
         * order of simulations is unchanged because nothing is pending at that point) */
        int np = 0, started = sims, nterm = 0;
        while (started < iters && np < vl && nterm < vl) {
            int node = 0, len = 1;
            int* pth = path + np * stride;
            pth[0] = 0;
            while (nd[node].nchild > 0) {                         /* mcts.py:20-23 */
                st[5] += nd[node].nchild; st[4]++;
                node = select_child(nd, node, node == 0 && noise != NULL);
                pth[len++] = node;
                board[nd[node].cell] = (uint8_t)(2 - nd[node].player);   /* mover = 1 - currentPlayer */
            }
            int terminal = -1;
            if (nd[node].pending) {
                for (int i = len - 1; i >= 1; i--) board[nd[pth[i]].cell] = 0;
                st[6]++;
                break;
            }
            if (node != 0) {                                      /* mcts.py:25-32 */
                if (is_win(&g, board, 1 - nd[node].player, nd[node].cell)) terminal = 1;
                else if (nd[node].mc == g.cells) terminal = 0;
            }
            if (terminal >= 0) {
                double v = (double)terminal;
                for (int i = len - 1; i >= 0; i--) {              /* node.py:62-74 */
                    nd[pth[i]].visit++; nd[pth[i]].W += v; v = -v;
                    if (i >= 1) board[nd[pth[i]].cell] = 0;
                }
                st[0]++; st[1]++; started++; sims++; nterm++;
                continue;
            }
            plen[np] = len;
            memcpy(pboard + (size_t)np * MAXC, board, g.cells);
            evaluate(&g, &ev, board, nd[node].player, ppri + (size_t)np * g.A, &pval[np]);
            if (vl > 1) {
                nd[node].pending = 1;
                for (int i = 0; i < len; i++) nd[pth[i]].vloss++;
            }
            for (int i = len - 1; i >= 1; i--) board[nd[pth[i]].cell] = 0;
            np++; started++;
        }
        /* phase A: expand + backup in selection order */
        for (int j = 0; j < np; j++) {
            int* pth = path + j * stride;
            int len = plen[j], leaf = pth[len - 1];
            int nc = legal_moves(&g, pboard + (size_t)j * MAXC, hashes, moves);
            nd[leaf].first = n_nodes; nd[leaf].nchild = nc; nd[leaf].pending = 0;
            for (int c = 0; c < nc; c++) {                        /* node.py:50-59 */
                ONode* ch = &nd[n_nodes + c];
                memset(ch, 0, sizeof(*ch));
                ch->parent = leaf; ch->cell = moves[c]; ch->player = 1 - nd[leaf].player; ch->mc = nd[leaf].mc + 1;
                float p = ppri[(size_t)j * g.A + action_index(&g, moves[c])];
                ch->P = p;
                if (leaf == 0 && noise) {                         /* utils.py:24-25 */
                    float a = 0.75f * p;
                    ch->P64 = (double)a + 0.25 * noise[action_index(&g, moves[c])];
                    ch->P = (float)ch->P64;
                }
            }
            n_nodes += nc;
            double v = -(double)pval[j];                          /* mcts.py:56 */
            for (int i = len - 1; i >= 0; i--) {
                nd[pth[i]].visit++; nd[pth[i]].W += v; v = -v;
                if (vl > 1) nd[pth[i]].vloss--;
            }
            st[0]++; st[2]++; st[3] += nc; sims++;
        }
    }
    int nc = nd[0].nchild;
    for (int c = 0; c < nc; c++) {
        const ONode* ch = &nd[nd[0].first + c];
        if (out_cells) out_cells[c] = ch->cell;
        if (out_visits) out_visits[c] = ch->visit;
        if (out_W) out_W[c] = ch->W;
        if (out_P) out_P[c] = noise ? ch->P64 : (double)ch->P;
    }
    if (root_visit) *root_visit = nd[0].visit;
    if (root_value) *root_value = nd[0].W;
    if (stats) for (int i = 0; i < 8; i++) stats[i] += st[i];
    if (ev.misses && stats) stats[7] += (uint64_t)ev.misses;
    free(nd); free(path); free(plen); free(ppri); free(pval); free(pboard);
    return nc;
}

/* n independent searches (e.g. one per GPU tree); arrays are [n][...] with maxch columns. */
int azo_search_batch(int gid, int rows, int cols, int n, const uint8_t* cells, const int* players, const int* mcs,
                     int iters, int vl, uint64_t seed, int qbits, const double* noise, void* table, int maxch,
                     int* out_nc, int* out_cells, int* out_visits, double* out_W, double* out_P, int* root_visit,
                     double* root_value, uint64_t* stats) {
    int ncells = rows * cols, A = gid == 1 ? 7 : ncells;
    for (int t = 0; t < n; t++) {
        int rc = azo_search(gid, rows, cols, cells + (size_t)t * ncells, players[t], mcs[t], iters, vl, seed, qbits,
                            noise ? noise + (size_t)t * A : NULL, table, out_cells + (size_t)t * maxch,
                            out_visits + (size_t)t * maxch, out_W + (size_t)t * maxch, out_P + (size_t)t * maxch,
                            root_visit + t, root_value + t, stats);
        if (rc < 0) return rc;
        out_nc[t] = rc;
    }
    return 0;
}

/* rules entry points for differential tests */
int azo_legal(int gid, int rows, int cols, const uint8_t* cells, int* out) {
    Game g;
    if (game_init(&g, gid, rows, cols)) return -1;
    uint64_t hashes[MAXC];
    for (int i = 0; i < g.cells; i++) hashes[i] = tuple2_hash(i / g.cols, i % g.cols);
    return legal_moves(&g, cells, hashes, out);
}

int azo_winner(int gid, int rows, int cols, const uint8_t* cells, int player, int cell) {
    Game g;
    if (game_init(&g, gid, rows, cols)) return -2;
    return is_win(&g, cells, player, cell) ? player : -1;
}

/* PUCT score bits for the golden micro-vectors: returns selected index, scores as doubles */
int azo_puct(int n, int parent_visit, int is_f64, const double* priors, const int* visits, const double* W,
             double* out_scores) {
    int best = 0;
    double bs = 0;
    for (int c = 0; c < n; c++) {
        double sc;
        if (is_f64) sc = puct64(W[c], priors[c], visits[c], sqrt((double)parent_visit));
        else sc = (double)puct32(W[c], (float)priors[c], visits[c], (float)sqrt((double)parent_visit));
        out_scores[c] = sc;
        if (c == 0 || sc > bs) { bs = sc; best = c; }
    }
    return best;
}
