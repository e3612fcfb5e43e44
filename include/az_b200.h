/* az_b200.h -- C ABI of the B200-native self-play hot path (libaz_b200.so).
 *
 * The reference (ehgus8/alpha-zero) is pure Python and has no FFI; these entry points are
 * what a ctypes binding for its hot path binds (INTEGRATION.md shows the stub).  Each
 * group names the reference interface it replaces.  Conventions: every function returns
 * 0 on success or a negative AZ_E* code (az_last_error() has the text); no exceptions and
 * no torch types cross the boundary; `stream` is a cudaStream_t passed as void* (NULL =
 * default stream); pointers named *_host are host memory, *_dev are device memory owned
 * by the caller; calls are asynchronous on `stream` unless they return host data.
 * One host thread per handle (the reference is single-threaded: global MCTS.cache, global
 * NumPy RNG).
 */
#ifndef AZ_B200_H
#define AZ_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define AZ_API __attribute__((visibility("default")))

#define AZ_OK 0
#define AZ_EINVAL (-1)
#define AZ_ECUDA (-2)
#define AZ_ENOMEM (-3)
#define AZ_ESTATE (-4)

#define AZ_TICTACTOE 0   /* games/tictactoe.py */
#define AZ_CONNECT4 1    /* games/connect4.py  */
#define AZ_GOMOKU 2      /* games/gomoku.py    */

typedef struct az_forest az_forest;
typedef struct az_net az_net;

AZ_API const char* az_version(void);
AZ_API const char* az_last_error(void);

/* ---- forest of search trees: replaces ai/node.py:20-40 (Node objects) ----------------- */
/* rows/cols: 3x3, 6x7, or n x n Gomoku (n <= 15; the reference's class attrs, gomoku.py:10).
 * max_iters bounds mcts_iterations of later searches (node pool = 1 + max_iters*max_children
 * per tree); vl_batch = leaves per tree per step (1 = the reference's sequential search). */
AZ_API int az_forest_create(int game, int rows, int cols, int n_trees, int max_iters, int vl_batch, az_forest** out);
AZ_API void az_forest_destroy(az_forest* f);

/* Root positions.  planes_host: float32 [n_trees][C][rows][cols] exactly as the reference
 * holds boards (Game().board, games/gomoku.py:16-17); players / move_counts as int32. */
AZ_API int az_forest_set_roots_planes(az_forest* f, const float* planes_host, const int32_t* players_host,
                               const int32_t* move_counts_host, void* stream);
/* Same from side bitboards: words_host uint64 [n_trees][2][nw], bit (r*cols+c)%64 of word
 * (r*cols+c)/64; nw = 1 for boards of <= 64 cells, else 4 (az_forest_info reports it). */
AZ_API int az_forest_set_roots_words(az_forest* f, const uint64_t* words_host, const int32_t* players_host,
                              const int32_t* move_counts_host, void* stream);
/* active_host: uint8 [n_trees] (NULL = all active). */
AZ_API int az_forest_set_active(az_forest* f, const uint8_t* active_host, void* stream);
/* Dirichlet noise for the root priors, float64 [n_trees][A] playing the role of
 * np.random.dirichlet in utils.py:21; NULL turns noise off (dirichlet=False, mcts.py:11). */
AZ_API int az_forest_set_noise(az_forest* f, const double* noise_host, void* stream);

/* ---- one search: replaces MCTS.mcts, ai/mcts.py:11-60 ------------------------------------ */
/* Fresh roots (no subtree reuse, gomoku.py:134) and `iters` simulations per tree. */
AZ_API int az_forest_begin_move(az_forest* f, int iters, void* stream);
/* Apply the evaluator's answers (expand + backup, node.py:50-74) and select the next
 * leaves (node.py:42-47, utils.py:29-44).  Fills the evaluation batch. */
AZ_API int az_forest_advance(az_forest* f, void* stream);
/* Synchronises `stream`; returns the size of the evaluation batch and finished trees. */
AZ_API int az_forest_counts(az_forest* f, int* n_leaves, int* n_done, void* stream);

/* Evaluators (the `model(...)` call + softmax glue of ai/mcts.py:46-49): */
/* (a) build-defined exact hash evaluator shared with the oracle (parity runs); */
AZ_API int az_eval_hash(az_forest* f, uint64_t seed, int qbits, void* stream);
/* (b) external evaluator: canonical leaves as float32 planes [max_b][C][rows][cols] ... */
AZ_API int az_leaves_to_planes(az_forest* f, float* planes_dev, int max_b, void* stream);
/* ... and its logits [max_b][A], values [max_b] back (softmax as mcts.py:49); */
AZ_API int az_leaves_set_logits(az_forest* f, const float* logits_dev, const float* values_dev, int max_b, void* stream);
/* (c) the built-in policy/value net, az_net_* below. */

/* Vanilla MCTS (MCTS.mcts(None, ...): ai/mcts.py:57-79, utils.py:36,43).  mode 1 = UCB1 selection in f64 with
 * ln(N) from the host C library (bit-exact with math.log), children without priors; mode 0 = PUCT (default). */
AZ_API int az_forest_set_mode(az_forest* f, int mode);
/* MCTS.simulate for every pending leaf: uniformly random legal moves to the end of the game (bitboard rules on
 * the device, splitmix64 streams keyed by seed / counter / tree); writes value = +1 / -1 / 0 for the side to move. */
AZ_API int az_eval_rollout(az_forest* f, uint64_t seed, uint64_t counter, void* stream);
/* Whole vanilla search (mode 1): advance + rollouts, host loop inside the library. */
AZ_API int az_forest_search_rollout(az_forest* f, int iters, uint64_t seed, void* stream);
/* Whole search with the hash evaluator, host loop inside the library. */
AZ_API int az_forest_search_hash(az_forest* f, int iters, uint64_t seed, int qbits, void* stream);

/* ---- results: replaces utils.get_probablity_distribution_of_children (utils.py:46-55),
 *      root.value/root.visit (gomoku.py:141) and reads of root.children (test.py:44-46) ---- */
AZ_API int az_forest_root_visits(az_forest* f, int32_t* visits_host /*[n_trees][A]*/, double* q_host /*[n_trees]*/,
                          void* stream);
/* root.max_visit_child() (ai/node.py:76-81) of every tree: cell of the FIRST most-visited child in child order
 * (-1 for a root that was not searched); int32 [n_trees]. */
AZ_API int az_forest_root_best(az_forest* f, int32_t* cells_host, void* stream);
/* Children of one root in child order: cell = r*cols+c of prevAction. Returns the count. */
AZ_API int az_forest_root_children(az_forest* f, int tree, int max_children, int32_t* cells_host, int32_t* visits_host,
                            double* values_host, double* priors_host, int32_t* root_visit, double* root_value,
                            void* stream);
/* The current evaluation batch (leaves + the evaluator's answers) copied to the host: canonical
 * words uint64 [n][2][nw], side to move, owning tree, priors [n][A], values [n].  Returns n.
 * This is the log that replays a GPU search in the oracle (MCTS.cache preload, mcts.py:40-41). */
AZ_API int az_forest_read_leaves(az_forest* f, int max_n, uint64_t* words_host, uint8_t* players_host,
                                 int32_t* trees_host, float* priors_host, float* values_host, void* stream);
/* counters: [sims, terminal sims, expansions, children, selects, children scored, collisions, cache hits] */
AZ_API int az_forest_stats(az_forest* f, uint64_t* out8_host, int reset, void* stream);
AZ_API int az_forest_info(az_forest* f, int* n_trees, int* action_dim, int* planes, int* nw, int* cap);

/* ---- evaluation cache: replaces MCTS.cache (ai/mcts.py:7,37-51; cleared per model update, main.py:55-57) ------
 * canonical position -> (softmax priors, value) in an open-addressed HBM table probed by the advance kernel: a leaf
 * whose position is cached costs no evaluator row, leaves of different trees reaching the same new position in one
 * step share one row, and at vl_batch == 1 a tree keeps simulating inside a launch while its leaves hit.  Transparent
 * for a deterministic evaluator (search_stats[7] counts the hits).  On by default for boards of <= 64 cells
 * (AZ_NO_CACHE=1 turns it off); log2_entries = 0 removes it.  Entries unused for 2 searches are recycled. */
AZ_API int az_forest_cache_config(az_forest* f, int log2_entries);
AZ_API int az_forest_cache_clear(az_forest* f, void* stream);   /* also forgets which evaluator filled it */
AZ_API int az_forest_cache_info(az_forest* f, long long* entries, long long* bytes);
/* vl_batch == 1 with the cache: up to `rounds` simulations per tree and launch while its leaves hit the cache (default 1:
 * lockstep searches gain nothing from more; continuous self-play does). */
AZ_API int az_forest_set_rounds(az_forest* f, int rounds);
/* Names the evaluator whose answers may be cached from now on (any id that changes with the evaluator or its weights);
 * a different id drops the cached answers.  az_forest_search_{hash,net} do this themselves; step-by-step callers
 * (az_forest_advance + an evaluator) call it before az_forest_begin_move when they switch evaluators. */
AZ_API int az_forest_set_evaluator(az_forest* f, unsigned long long id, void* stream);

/* ---- whole self-play games on the device: replaces the per-ply driver Game.self_play (games/gomoku.py:123-164,
 *      connect4.py:117-151, tictactoe.py:99-133) and train.collect_data's game loop (train.py:54-83) -----------------
 * The forest's trees are game SLOTS: after every search one kernel records the position (side bitboards, u16 root
 * visit counts by action, f64 root Q), chooses the move -- np.random.choice over pi while ply < sample_plies
 * (node.py:83-93: f64 cumsum, searchsorted side='right') else the first most-visited child in child order
 * (node.py:76-81) --, applies it, tests win / draw, starts the next game in a finished slot and writes the next roots;
 * Dirichlet(0.03) root noise (utils.py:21) is drawn on the device.  Randomness is Philox4x32-10 keyed by (seed; global
 * game id, ply), global id of local game k = id_first + k * id_stride: trajectories do not depend on slot, GPU or rank
 * count.  n_games may exceed the forest's trees (slots are refilled). */
typedef struct az_selfplay az_selfplay;
AZ_API int az_selfplay_create(az_forest* f, int n_games, int id_first, int id_stride, uint64_t seed, int sample_plies,
                              int use_noise, az_selfplay** out);
AZ_API void az_selfplay_destroy(az_selfplay* sp);
/* adopt_roots = 0: every game starts from the empty board; 1: games 0..slots-1 continue from the forest's current roots. */
AZ_API int az_selfplay_begin(az_selfplay* sp, int adopt_roots, void* stream);
/* One ply by hand: az_selfplay_prepare (root noise), any az_forest_search_*, az_selfplay_move. */
AZ_API int az_selfplay_prepare(az_selfplay* sp, void* stream);
AZ_API int az_selfplay_move(az_selfplay* sp, void* stream);
/* Synchronises: games started / finished, live slots, positions recorded in finished games. */
AZ_API int az_selfplay_status(az_selfplay* sp, int* started, int* finished, int* live, long long* positions, void* stream);
/* Plays until every game has finished or max_plies searches were run (0 = no limit); returns the searches run. */
AZ_API int az_selfplay_run_net(az_selfplay* sp, az_net* net, int iters, int max_plies, void* stream);
AZ_API int az_selfplay_run_hash(az_selfplay* sp, int iters, uint64_t seed, int qbits, int max_plies, void* stream);
/* Continuous plies: the same games without lockstep -- a slot whose search is finished moves and starts its next search
 * while the others keep searching (nobody waits for the slowest tree of a ply).  Trajectories are identical to the
 * lockstep run (search, noise and move sampling are functions of (game id, ply)).  Returns the search steps run. */
/* max_steps > 0: return after that many search steps (benchmarks); resume != 0: continue the searches such a call left. */
AZ_API int az_selfplay_run_async_net(az_selfplay* sp, az_net* net, int iters, long long max_steps, int resume, void* stream);
AZ_API int az_selfplay_run_async_hash(az_selfplay* sp, int iters, uint64_t seed, int qbits, void* stream);
/* Trajectories, positions indexed [game][ply] with max_plies = rows*cols per game.  Device pointers (library-owned):
 * out8 = {plies i32[G], winner i32[G] (0/1, -1 draw, -2 unfinished), words u64[G][P][2][nw], visits u16[G][P][A],
 * q f64[G][P], action(cell) i16[G][P], uniform f64[G][P] (-1 = argmax ply), first recorded ply i32[G]}. */
AZ_API int az_selfplay_buffers(az_selfplay* sp, void** out8, int* n_games, int* max_plies);
/* Host copies (any pointer may be NULL).  Synchronous. */
AZ_API int az_selfplay_read(az_selfplay* sp, int32_t* plies_host, int32_t* winner_host, int32_t* first_host, uint64_t* words_host,
                            uint16_t* visits_host, double* q_host, int16_t* action_host, double* u_host, void* stream);
AZ_API int az_selfplay_read_slots(az_selfplay* sp, int32_t* slot_game_host, double* noise_host, void* stream);
AZ_API long long az_selfplay_launch_count(az_selfplay* sp);

/* ---- device-resident replay buffer: replaces train.save_data_to_buffer / rotate_data / flip_data (train.py:8-49) and
 *      ReplayBuffer.add / sample / size (replay_buffer.py:8-25,34-35) ---------------------------------------------------
 * A ring of `capacity` ENTRIES (the deque's maxlen), one per sample the reference would hold, in its order (game, ply,
 * variant: rot0, rot0-lr, rot0-tb, rot90, rot90-lr, rot90-tb, rot180, rot270; plies 0 and 1 once; Connect4: identity +
 * mirror), compact in HBM: canonical transformed side bitboards, u16 visit counts, z. */
typedef struct az_replay az_replay;
AZ_API int az_replay_create(int game, int rows, int cols, long long capacity, az_replay** out);
AZ_API void az_replay_destroy(az_replay* rb);
AZ_API long long az_replay_size(az_replay* rb);
AZ_API long long az_replay_entry_bytes(az_replay* rb);
AZ_API int az_replay_clear(az_replay* rb);
/* Appends the recorded positions of n_games finished games; device arrays in the layout of az_selfplay_buffers
 * (plies / winner / first i32[n_games], words u64[n_games][P][2][nw], visits u16[n_games][P][A], P = rows*cols), games
 * in storage order.  Returns the entries appended (< 0: error). */
AZ_API long long az_replay_add(az_replay* rb, int n_games, int max_plies, const int32_t* plies_dev, const int32_t* winner_dev,
                               const int32_t* first_dev, const uint64_t* words_dev, const uint16_t* visits_dev, void* stream);
/* ReplayBuffer.sample(B): B distinct entries drawn uniformly (a keyed permutation of (seed, counter)) expanded into the
 * caller's device buffers: states f32 [B][C][rows][cols], policies f32 [B][A] (= (float)(n_a / sum n)), z f32 [B][1];
 * indices_dev (optional, int64 [B]) receives the entry numbers (0 = oldest). */
AZ_API int az_replay_sample(az_replay* rb, int B, uint64_t seed, uint64_t counter, float* states_dev, float* policies_dev,
                            float* z_dev, long long* indices_dev, void* stream);
/* Entries start .. start+count-1 (oldest first), expanded like a sample. */
AZ_API int az_replay_read(az_replay* rb, long long start, int count, float* states_dev, float* policies_dev, float* z_dev, void* stream);
/* Source cell of output cell `cell` under augmentation variant 0..7 (host function; the transform the kernels apply). */
AZ_API int az_d4_source_cell(int rows, int cols, int variant, int cell);
/* ---- the loss of train.train (train.py:110-112) for one batch, forward + backward in one kernel: out2 = {mean_b sum_a
 *      -pi log_softmax(logits), mean_b (v - z)^2}, dlogits [B][A], dvalues [B].  All pointers device memory. */
AZ_API int az_loss_policy_value(const float* logits_dev, const float* values_dev, const float* pis_dev, const float* zs_dev, int B,
                                int A, float* dlogits_dev, float* dvalues_dev, float* out2_dev, void* stream);

/* ---- batched rules for differential tests: games/*.py get_valid_moves / check_winner ---- */
/* words_host uint64 [n][2][nw]; out_counts int32 [n]; out_cells uint8 [n][rows*cols]. */
AZ_API int az_rules_legal(int game, int rows, int cols, int n, const uint64_t* words_host, int32_t* out_counts_host,
                   uint8_t* out_cells_host);
/* az_rules_legal for callers that may not allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  It has NO counterpart in the reference.
 * words_dev uint64 [n][2][nw], out_counts_dev int32 [n], out_cells_dev uint8 [n][rows*cols] are DEVICE buffers of the
 * caller; asynchronous on `stream`; results identical to az_rules_legal.  The kernel's scratch is one static __device__
 * array of 4 MiB + 2 KiB that every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree.
 * Limit: n * 2 * (set_slots + 1) <= 4 MiB, i.e. at most 16 256 positions of 7x7 Gomoku, 4 088 of 15x15 Gomoku, 2 097 152
 * of TicTacToe / Connect4 per call; larger n returns AZ_EINVAL with the limit in az_last_error().  All calls share the
 * array (it keeps the last call's contents; nothing in it needs clearing): calls that overlap on the GPU race. */
AZ_API int az_rules_legal_noalloc(int game, int rows, int cols, int n, const uint64_t* words_dev, int32_t* out_counts_dev,
                                  uint8_t* out_cells_dev, void* stream);
/* After `player` played `cell`: winner id or -1 (check_winner). */
AZ_API int az_rules_winner(int game, int rows, int cols, int n, const uint64_t* words_host, const int32_t* players_host,
                    const int32_t* cells_host, int32_t* out_winner_host);

/* ---- policy/value network: replaces Net.forward (ai/nn.py:63-84) + softmax (ai/mcts.py:49) ---- */
/* Tensor-core GEMM building block, exposed for parity tests: out[M][N] (bf16, row stride ldo)
 * = epi(A[M][K] (bf16, row stride lda) * W[N][K]^T + bias[N] (fp32)); epi 0 = none, 1 = exact
 * GELU, 2 = + resid[M][ldr].  All pointers are device memory; N, K multiples of 64. */
AZ_API int az_gemm_bf16(const void* a_dev, int lda, const void* w_dev, const float* bias_dev, const void* resid_dev,
                        int ldr, void* out_dev, int ldo, int M, int N, int K, int epi, void* stream);

/* The reference's Net(img_size, patch_size, embed_dim, action_dim, num_heads, depth, channels)
 * (ai/nn.py:63-72); rows x cols generalises img_size to non-square boards (Connect4). */
AZ_API int az_net_create(int rows, int cols, int planes, int patch, int embed_dim, int num_heads, int depth,
                         int action_dim, int max_batch, az_net** out);
AZ_API void az_net_destroy(az_net* n);
/* One call per state_dict entry, under its PyTorch name (utils.load_model checkpoints,
 * utils.py:63-69): fp32 host data.  az_net_finalize converts them (bf16 weights, TMA maps). */
AZ_API int az_net_set_param(az_net* n, const char* name, const float* data_host, long long numel);
AZ_API int az_net_finalize(az_net* n);
/* Evaluate the forest's pending leaves in place of `model(...)` + softmax (ai/mcts.py:46-49). */
AZ_API int az_net_eval_leaves(az_net* n, az_forest* f, void* stream);
/* Net.forward on canonical float32 planes [B][C][rows][cols] (host) -> logits [B][A],
 * values [B], priors [B][A] (host; any may be NULL).  Synchronous. */
AZ_API int az_net_forward_planes(az_net* n, const float* planes_host, int B, float* logits_host, float* values_host,
                                 float* priors_host, void* stream);
AZ_API long long az_net_launch_count(az_net* n);
AZ_API int az_net_proj_width(az_net* n);   /* N of the all-token projection GEMM of the last block */
/* Which kernels this net runs (for reports): bit 0 fused front end (az_front.cuh), bit 1 its reassociated form
 * (az_front2.cuh: U vectors, out-proj over K = heads*128), bit 2 CTA-pair tiles for the cls-row GEMMs (az_gemm2.cuh),
 * bit 3 LayerNorm folded around the cls-row GEMMs, bit 4 heads + softmax as a GEMM epilogue, bit 5 all-token attention of
 * the full blocks on tcgen05 (az_attn.cuh; off: the scalar kernel, AZ_NO_ATTN_MMA=1 or an unsupported head dim), bit 6 the
 * cls-row tail fused into one cluster kernel (az_tail.cuh; opt-in, AZ_FUSED_TAIL=1), bit 7 bf16 GEMM tiles of 128 / 256
 * columns leave through shared memory + TMA stores (az_gemm.cuh; AZ_NO_TMA_STORE=1 restores per-thread row stores), bit 8
 * the four cls-row GEMMs as one persistent kernel (az_chain.cuh; opt-in, AZ_CHAIN=1). */
AZ_API int az_net_path_flags(az_net* n);
/* MCTS.mcts with the built-in net for every tree of the forest: `iters` simulations per tree,
 * one CUDA graph per step (advance + net), no per-node host work. */
AZ_API int az_forest_search_net(az_forest* f, az_net* n, int iters, void* stream);
AZ_API int az_net_invalidate_graphs(az_net* n);
/* Per-kernel CUDA-event timing of eager forward passes (roofline numbers of bench.py).
 * enable: 1 = clear totals and start, 0 = stop, 2 = resume, 3 = pause (totals kept). */
AZ_API int az_net_profile(az_net* n, int enable);
AZ_API int az_net_profile_read(az_net* n, int max_kinds, const char** names_out, double* ms_out, long long* count_out);

#ifdef __cplusplus
}
#endif
#endif
