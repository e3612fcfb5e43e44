// Whole self-play games on the device (sm_100a): the per-ply driver of the reference
//   Gomoku.self_play      games/gomoku.py:123-164   (connect4.py:117-151, tictactoe.py:99-133)
//   train.collect_data    train.py:54-83            (games one after another -> `slots` games in flight, refilled)
//   Node.sample_child     ai/node.py:83-93          np.random.choice over pi = visits / sum
//   Node.max_visit_child  ai/node.py:76-81          first maximum in CHILD order
//   add_dirichlet_noise   utils.py:12-27            Dirichlet(0.03) drawn per root
// without a host step per ply: after each search one kernel turns every slot's root statistics into the move, appends
// the position to the game's trajectory (compact: side bitboards, u16 visit counts, f64 root Q, move, the uniform that
// drew it), applies the move on bitboards, tests win / draw, starts the next game in a finished slot and writes the next
// roots; a second kernel draws the Dirichlet noise of the new roots.  All randomness is counter based -- Philox4x32-10
// keyed by (seed; global game id, ply, purpose) -- so a game's trajectory does not depend on the slot, the GPU or the
// number of ranks it was played on, and a test can replay every decision from the logged uniform.
// Compiled with -fmad=false (the f64 CDF must round like NumPy's cumsum / divide).
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "az_forest.h"
#include "az_rules.cuh"
#include "az_host.h"
#include "../../include/az_b200.h"

struct AzSelfplayDev {
    int n_games;              // games this engine plays
    int id_first, id_stride;  // global id of local game k = id_first + k * id_stride (rank + k * world)
    int max_plies;            // positions per game (state_dim)
    int sample_plies;         // plies < sample_plies draw the move from pi, later plies take the most visited child
    uint64_t seed;
    int* slot_game;           // [n_trees] local game in the slot, -1 = retired
    int* counters;            // [0] next game to start, [1] games finished, [2] live slots, [3] positions recorded
    // trajectories, [n_games][max_plies] positions
    uint64_t* t_words;        // [..][2 * nw]  the board BEFORE the move (absolute sides), boards.append(board.copy()) gomoku.py:140
    uint16_t* t_visits;       // [..][A]       root visit counts by action (pi numerators, utils.py:46-55)
    double* t_q;              // root.value / root.visit (gomoku.py:141)
    int16_t* t_action;        // cell of the move played
    double* t_u;              // the uniform that drew it, -1 when the move was the most visited child
    int* g_plies;             // [n_games] plies of the finished game (= stones on its final board), 0 = not finished
    int* g_winner;            // [n_games] 0 / 1, -1 = draw (check_winner's convention), -2 = not finished
    int* g_first;             // [n_games] first recorded ply: 0, or the move count of an adopted root (az_selfplay_begin)
    uint8_t* need_noise;      // [n_trees] continuous plies: the slot has a new root whose Dirichlet noise is not drawn yet
};

namespace {

// ---- Philox4x32-10 (Salmon et al. 2011), counter = (game id, ply, purpose, index) -----------------------------
struct AzPhilox {
    uint32_t v[4];
};
__host__ __device__ __forceinline__ AzPhilox az_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    AzPhilox r{{c0, c1, c2, c3}};
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * r.v[0];
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * r.v[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ r.v[1] ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ r.v[3] ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        r.v[0] = n0; r.v[1] = n1; r.v[2] = n2; r.v[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return r;
}
// 53-bit uniform in [0, 1), the resolution of NumPy's random_sample
__host__ __device__ __forceinline__ double az_u53(uint32_t hi, uint32_t lo) {
    return (double)((((uint64_t)hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
}
#define AZ_RNG_MOVE 0u
#define AZ_RNG_GAMMA 1u
#define AZ_RNG_BOOST 2u

__global__ void az_selfplay_begin_kernel(AzForestDev f, AzSelfplayDev sp, int adopt) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) {
        const int live = min(f.n_trees, sp.n_games);
        sp.counters[0] = live;
        sp.counters[1] = 0;
        sp.counters[2] = live;
        sp.counters[3] = 0;
    }
    for (int k = s; k < sp.n_games; k += gridDim.x * blockDim.x) {
        sp.g_plies[k] = 0;
        sp.g_winner[k] = -2;
        sp.g_first[k] = (adopt && k < f.n_trees) ? (int)f.root_mc[k] : 0;
    }
    if (s >= f.n_trees) return;
    const int live = s < sp.n_games;
    sp.slot_game[s] = live ? s : -1;
    f.active[s] = (uint8_t)live;
    if (adopt) return;                                            // games 0 .. slots-1 continue from the forest's current roots
    for (int w = 0; w < 2 * f.g.nw; ++w) f.root_words[(size_t)s * 2 * f.g.nw + w] = 0ull;
    f.root_player[s] = 0;
    f.root_mc[s] = 0;
}

// One thread per slot, after a search: record the position, choose and apply the move, finish / refill the slot.
template <int NW>
__device__ __forceinline__ void az_selfplay_move_slot(const AzForestDev& f, const AzSelfplayDev& sp, int s) {
    int k = sp.slot_game[s];
    const AzGame& g = f.g;
    const int A = g.action_dim;
    const AzNode* nodes = f.nodes + (size_t)s * f.cap;
    const AzNode root = nodes[0];
    const int ply = f.root_mc[s];
    const int player = f.root_player[s];
    AzBoard<NW> b;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        b.w[0][w] = f.root_words[(size_t)s * 2 * NW + w];
        b.w[1][w] = f.root_words[(size_t)s * 2 * NW + NW + w];
    }
    const uint32_t gid = (uint32_t)(sp.id_first + k * sp.id_stride);
    int cell;
    {
        const size_t pos = (size_t)k * sp.max_plies + ply;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            sp.t_words[pos * 2 * NW + w] = b.w[0][w];
            sp.t_words[pos * 2 * NW + NW + w] = b.w[1][w];
        }
        uint16_t* vis = sp.t_visits + pos * A;
        for (int a = 0; a < A; ++a) vis[a] = 0;
        const int nc = root.meta.status == AZ_ST_EXPANDED ? root.meta.nchild : 0;
        const uint32_t cb = root.meta.child_base;
        int sum = 0, best_n = -1, best_c = 0;
        for (int c = 0; c < nc; ++c) {                            // node.py:76-81: first maximum in child order
            const int n = nodes[cb + c].hot.N;
            vis[az_action_index(g, nodes[cb + c].meta.action)] = (uint16_t)n;
            sum += n;
            if (n > best_n) { best_n = n; best_c = c; }
        }
        sp.t_q[pos] = root.hot.N ? root.W / (double)root.hot.N : 0.0;   // gomoku.py:141
        double u = -1.0;
        int pick = best_c;
        if (ply < sp.sample_plies && sum > 0) {
            // node.py:92 np.random.choice(children by action, p = pi): cdf = cumsum(pi); cdf /= cdf[-1];
            // index = searchsorted(cdf, u, side='right') = number of entries <= u.  f64, sequential, as NumPy.
            const AzPhilox r = az_philox(sp.seed, gid, (uint32_t)ply, AZ_RNG_MOVE, 0u);
            u = az_u53(r.v[0], r.v[1]);
            const double dsum = (double)sum;
            double tot = 0.0;
            for (int a = 0; a < A; ++a) tot = __dadd_rn(tot, __ddiv_rn((double)vis[a], dsum));
            double c = 0.0;
            int idx = 0;
            for (int a = 0; a < A; ++a) {
                c = __dadd_rn(c, __ddiv_rn((double)vis[a], dsum));
                if (__ddiv_rn(c, tot) <= u) idx = a + 1;
            }
            if (idx > A - 1) idx = A - 1;
            int found = -1;
            for (int c2 = 0; c2 < nc; ++c2)
                if (az_action_index(g, nodes[cb + c2].meta.action) == idx) found = c2;
            // u landing exactly on a CDF step can name an action without visits (NumPy would return it too, and the
            // reference would then crash on a None child): keep the most visited child in that 2^-53 case
            if (found >= 0 && vis[idx] > 0) pick = found;
        }
        cell = nodes[cb + pick].meta.action;
        sp.t_action[pos] = (int16_t)cell;
        sp.t_u[pos] = u;
    }
    b.put(player, cell);                                          // Game.make_move
    const int mc = ply + 1;
    const bool won = az_is_win<NW>(g, b, player, cell);           // Game.check_winner (last move only)
    const bool over = won || mc == g.state_dim;
    int np = player ^ 1;
    if (over) {
        sp.g_plies[k] = mc;
        sp.g_winner[k] = won ? player : -1;
        atomicAdd(&sp.counters[1], 1);
        atomicAdd(&sp.counters[3], mc - sp.g_first[k]);
        const int nxt = atomicAdd(&sp.counters[0], 1);
        if (nxt < sp.n_games) {                                   // the next game takes the slot (train.py:64: for iter in range(iterations))
            k = nxt;
#pragma unroll
            for (int w = 0; w < NW; ++w) b.w[0][w] = b.w[1][w] = 0ull;
            np = 0;
            sp.slot_game[s] = k;
            f.root_mc[s] = 0;
        } else {
            sp.slot_game[s] = -1;
            f.active[s] = 0;
            atomicSub(&sp.counters[2], 1);
            f.root_mc[s] = (uint16_t)mc;
        }
    } else {
        f.root_mc[s] = (uint16_t)mc;
    }
    f.root_player[s] = (uint8_t)np;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        f.root_words[(size_t)s * 2 * NW + w] = b.w[0][w];
        f.root_words[(size_t)s * 2 * NW + NW + w] = b.w[1][w];
    }
}

// Lockstep plies: every live slot moves (the search of ALL trees has finished).
template <int NW>
__global__ void az_selfplay_move_kernel(AzForestDev f, AzSelfplayDev sp) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= f.n_trees || sp.slot_game[s] < 0) return;
    az_selfplay_move_slot<NW>(f, sp, s);
}

// Continuous plies: a slot whose search has finished (done_flag, set by az_advance_kernel) moves NOW and starts its next
// search -- fresh root, no subtree reuse (gomoku.py:134) -- while the other slots keep searching; nobody waits for the
// slowest tree of a ply.  A game's trajectory does not depend on the schedule: search, noise and move sampling are
// functions of (game id, ply).  The re-arming is what az_begin_move_kernel does for every tree at once.
template <int NW>
__global__ void az_selfplay_service_kernel(AzForestDev f, AzSelfplayDev sp) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= f.n_trees || sp.slot_game[s] < 0 || !f.done_flag[s]) return;
    az_selfplay_move_slot<NW>(f, sp, s);
    if (sp.slot_game[s] < 0) return;                              // retired: no game left for this slot
    AzNode& root = f.nodes[(size_t)s * f.cap];
    AzHot h;
    h.N = 0; h.Q = 0.f; h.P = 0.f; h.V = 0;
    AzMeta m;
    m.child_base = 0; m.action = 0; m.nchild = 0; m.status = AZ_ST_LEAF;
    root.hot = h;
    root.meta = m;
    root.W = 0.0;
    f.n_nodes[s] = 1;
    f.sims_done[s] = 0;
    f.n_pending[s] = 0;
    f.done_flag[s] = 0;
    sp.need_noise[s] = 1;
}

// Gamma(alpha) for alpha < 1 as Gamma(alpha + 1) * U^(1/alpha) (Marsaglia & Tsang 2000 for the shape >= 1 part).
// NumPy's legacy generator uses a different algorithm for the same distribution (utils.py:21): statistical parity.
__device__ double az_gamma_small(uint64_t seed, uint32_t gid, uint32_t ply, uint32_t a, double alpha) {
    const double d = alpha + 1.0 - 1.0 / 3.0;
    const double c = 1.0 / sqrt(9.0 * d);
    double gam = d;
    for (uint32_t it = 0; it < 24; ++it) {
        const AzPhilox r = az_philox(seed, gid, ply, AZ_RNG_GAMMA | (it << 8), a);
        const double u1 = ((double)r.v[0] + 0.5) * (1.0 / 4294967296.0);
        const double u2 = ((double)r.v[1] + 0.5) * (1.0 / 4294967296.0);
        const double x = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
        const double t = 1.0 + c * x;
        if (t <= 0.0) continue;
        const double v = t * t * t;
        const double u = ((double)r.v[2] + 0.5) * (1.0 / 4294967296.0);
        gam = d * v;
        if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) break;
    }
    const AzPhilox r = az_philox(seed, gid, ply, AZ_RNG_BOOST, a);
    const double ub = az_u53(r.v[0], r.v[1]) + (1.0 / 18014398509481984.0);
    return gam * exp(log(ub) / alpha);
}

// One warp per slot: noise[a] = g_a / sum g, g_a ~ Gamma(alpha): np.random.dirichlet([alpha] * A) (utils.py:21).
__global__ void az_selfplay_noise_kernel(AzForestDev f, AzSelfplayDev sp, double* noise, double alpha, int only_flagged) {
    const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (s >= f.n_trees) return;
    const int k = sp.slot_game[s];
    if (k < 0) return;
    if (only_flagged) {                                           // continuous plies: only the slots that have just moved
        if (!sp.need_noise[s]) return;
        __syncwarp();
        if (lane == 0) sp.need_noise[s] = 0;
    }
    const int A = f.g.action_dim;
    const uint32_t gid = (uint32_t)(sp.id_first + k * sp.id_stride);
    const uint32_t ply = f.root_mc[s];
    double part = 0.0;
    for (int a = lane; a < A; a += 32) {
        const double gm = az_gamma_small(sp.seed, gid, ply, (uint32_t)a, alpha);
        noise[(size_t)s * A + a] = gm;
        part += gm;
    }
    for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    __syncwarp();
    for (int a = lane; a < A; a += 32) noise[(size_t)s * A + a] = part > 0.0 ? noise[(size_t)s * A + a] / part : 1.0 / (double)A;
}

}  // namespace

// ======================================================================================
// Host side: handle + C ABI
// ======================================================================================
#define AZ_CUDA_OK(expr)                                                                       \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess) return az_fail(AZ_ECUDA, std::string(#expr ": ") + cudaGetErrorString(e__)); \
    } while (0)

struct az_selfplay {
    az_forest* f = nullptr;
    AzSelfplayDev d{};
    std::vector<void*> allocs;
    int* h_counters = nullptr;      // pinned
    int use_noise = 0;
    double alpha = 0.03;            // utils.py:21
    long long plies_run = 0;        // searches driven by az_selfplay_run_*
    long long launches = 0;         // own kernels launched
};

extern "C" void az_selfplay_destroy(az_selfplay* sp) {
    if (!sp) return;
    cudaDeviceSynchronize();
    for (void* p : sp->allocs) cudaFree(p);
    if (sp->h_counters) cudaFreeHost(sp->h_counters);
    delete sp;
}

extern "C" int az_selfplay_create(az_forest* f, int n_games, int id_first, int id_stride, uint64_t seed, int sample_plies,
                                  int use_noise, az_selfplay** out) {
    if (!f || !out || n_games <= 0 || id_stride <= 0 || id_first < 0 || sample_plies < 0)
        return az_fail(AZ_EINVAL, "az_selfplay_create: bad arguments");
    AzForestDev* fd = az_forest_dev(f);
    if (fd->iters > 65535) return az_fail(AZ_EINVAL, "az_selfplay_create: visit counts are stored as u16 (max_iters <= 65535)");
    az_selfplay* sp = new az_selfplay();
    sp->f = f;
    sp->use_noise = use_noise ? 1 : 0;
    AzSelfplayDev& d = sp->d;
    d.n_games = n_games;
    d.id_first = id_first;
    d.id_stride = id_stride;
    d.max_plies = fd->g.state_dim;
    d.sample_plies = sample_plies;
    d.seed = seed;
    const size_t npos = (size_t)n_games * d.max_plies;
    auto alloc = [&](void** p, size_t bytes) {
        if (cudaMalloc(p, bytes + 16) != cudaSuccess) return false;
        sp->allocs.push_back(*p);
        return true;
    };
    bool ok = alloc((void**)&d.slot_game, (size_t)fd->n_trees * 4) && alloc((void**)&d.counters, 64) &&
              alloc((void**)&d.t_words, npos * 2 * fd->g.nw * 8) && alloc((void**)&d.t_visits, npos * fd->g.action_dim * 2) &&
              alloc((void**)&d.t_q, npos * 8) && alloc((void**)&d.t_action, npos * 2) && alloc((void**)&d.t_u, npos * 8) &&
              alloc((void**)&d.g_plies, (size_t)n_games * 4) && alloc((void**)&d.g_winner, (size_t)n_games * 4) &&
              alloc((void**)&d.g_first, (size_t)n_games * 4) && alloc((void**)&d.need_noise, (size_t)fd->n_trees);
    if (!ok || cudaMallocHost(&sp->h_counters, 64) != cudaSuccess) {
        cudaGetLastError();
        az_selfplay_destroy(sp);
        return az_fail(AZ_ENOMEM, "az_selfplay_create: allocation failed");
    }
    cudaMemset(d.counters, 0, 64);
    cudaMemset(d.slot_game, 0xff, (size_t)fd->n_trees * 4);
    cudaMemset(d.need_noise, 0, (size_t)fd->n_trees);
    *out = sp;
    return AZ_OK;
}

// Games 0 .. slots-1 started: from empty boards, or (adopt_roots != 0) from the forest's current root positions -- synthetic
// mid-game starts for benchmarks; their earlier plies are not recorded (g_first).  Later games always start empty.
extern "C" int az_selfplay_begin(az_selfplay* sp, int adopt_roots, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_begin: null");
    AzForestDev* fd = az_forest_dev(sp->f);
    cudaStream_t s = (cudaStream_t)stream;
    const int n = std::max(fd->n_trees, 1);
    const size_t npos = (size_t)sp->d.n_games * sp->d.max_plies;   // unrecorded positions read as zeros
    AZ_CUDA_OK(cudaMemsetAsync(sp->d.t_words, 0, npos * 2 * fd->g.nw * 8, s));
    AZ_CUDA_OK(cudaMemsetAsync(sp->d.t_visits, 0, npos * fd->g.action_dim * 2, s));
    AZ_CUDA_OK(cudaMemsetAsync(sp->d.t_q, 0, npos * 8, s));
    AZ_CUDA_OK(cudaMemsetAsync(sp->d.t_action, 0, npos * 2, s));
    AZ_CUDA_OK(cudaMemsetAsync(sp->d.t_u, 0, npos * 8, s));
    az_selfplay_begin_kernel<<<(n + 255) / 256, 256, 0, s>>>(*fd, sp->d, adopt_roots ? 1 : 0);
    AZ_CUDA_OK(cudaGetLastError());
    sp->launches += 1;
    return AZ_OK;
}

// Before a search: Dirichlet noise for the current roots (noise on), drawn on the device into the forest's buffer.
extern "C" int az_selfplay_prepare(az_selfplay* sp, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_prepare: null");
    AzForestDev* fd = az_forest_dev(sp->f);
    cudaStream_t s = (cudaStream_t)stream;
    if (!sp->use_noise) return az_forest_use_device_noise(sp->f, 0);
    az_forest_use_device_noise(sp->f, 1);
    fd = az_forest_dev(sp->f);
    const int warps_per_block = 4;
    az_selfplay_noise_kernel<<<(fd->n_trees + warps_per_block - 1) / warps_per_block, 32 * warps_per_block, 0, s>>>(
        *fd, sp->d, az_forest_noise_dev(sp->f), sp->alpha, 0);
    AZ_CUDA_OK(cudaGetLastError());
    sp->launches += 1;
    return AZ_OK;
}

// After a search: one ply of every live game.
extern "C" int az_selfplay_move(az_selfplay* sp, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_move: null");
    AzForestDev* fd = az_forest_dev(sp->f);
    cudaStream_t s = (cudaStream_t)stream;
    const int n = fd->n_trees;
    if (fd->g.nw == 1) az_selfplay_move_kernel<1><<<(n + 127) / 128, 128, 0, s>>>(*fd, sp->d);
    else az_selfplay_move_kernel<AZ_MAX_WORDS><<<(n + 127) / 128, 128, 0, s>>>(*fd, sp->d);
    AZ_CUDA_OK(cudaGetLastError());
    sp->launches += 1;
    return AZ_OK;
}

// Continuous plies: finished slots move and start their next search (+ their root noise).  Asynchronous, capturable.
int az_selfplay_service(az_selfplay* sp, cudaStream_t s) {
    AzForestDev* fd = az_forest_dev(sp->f);
    const int n = fd->n_trees;
    if (fd->g.nw == 1) az_selfplay_service_kernel<1><<<(n + 127) / 128, 128, 0, s>>>(*fd, sp->d);
    else az_selfplay_service_kernel<AZ_MAX_WORDS><<<(n + 127) / 128, 128, 0, s>>>(*fd, sp->d);
    sp->launches += 1;
    if (sp->use_noise) {
        az_selfplay_noise_kernel<<<(n + 3) / 4, 128, 0, s>>>(*fd, sp->d, az_forest_noise_dev(sp->f), sp->alpha, 1);
        sp->launches += 1;
    }
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}
az_forest* az_selfplay_forest(az_selfplay* sp) { return sp->f; }
void az_selfplay_count_plies(az_selfplay* sp, long long n) { sp->plies_run += n; }

// Synchronises `stream`.  started / finished games, live slots, positions recorded in finished games.
extern "C" int az_selfplay_status(az_selfplay* sp, int* started, int* finished, int* live, long long* positions, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_status: null");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaMemcpyAsync(sp->h_counters, sp->d.counters, 16, cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    if (started) *started = std::min(sp->h_counters[0], sp->d.n_games);
    if (finished) *finished = sp->h_counters[1];
    if (live) *live = sp->h_counters[2];
    if (positions) *positions = sp->h_counters[3];
    return AZ_OK;
}

// The whole of train.collect_data's game loop: searches + moves until every game has finished (or max_plies searches
// were run, 0 = no limit).  evaluator: net != NULL -> az_forest_search_net, else the hash evaluator (seed, qbits).
static int az_selfplay_run(az_selfplay* sp, az_net* net, uint64_t hseed, int qbits, int iters, int max_plies, void* stream) {
    if (iters < 1) return az_fail(AZ_EINVAL, "az_selfplay_run: iters >= 1");
    int live = 1, done = 0;
    int rc = az_selfplay_status(sp, nullptr, nullptr, &live, nullptr, stream);
    if (rc) return rc;
    while (live > 0 && (max_plies <= 0 || done < max_plies)) {
        if ((rc = az_selfplay_prepare(sp, stream))) return rc;
        rc = net ? az_forest_search_net(sp->f, net, iters, stream) : az_forest_search_hash(sp->f, iters, hseed, qbits, stream);
        if (rc) return rc;
        if ((rc = az_selfplay_move(sp, stream))) return rc;
        if ((rc = az_selfplay_status(sp, nullptr, nullptr, &live, nullptr, stream))) return rc;
        ++done;
        ++sp->plies_run;
    }
    return done;
}

extern "C" int az_selfplay_run_net(az_selfplay* sp, az_net* net, int iters, int max_plies, void* stream) {
    if (!sp || !net) return az_fail(AZ_EINVAL, "az_selfplay_run_net: null");
    cudaStream_t s = (cudaStream_t)stream;
    static cudaStream_t own = nullptr;                          // the net search cannot capture graphs on the legacy stream
    if (!s) {
        if (!own) AZ_CUDA_OK(cudaStreamCreateWithFlags(&own, cudaStreamNonBlocking));
        AZ_CUDA_OK(cudaStreamSynchronize(0));
        s = own;
    }
    return az_selfplay_run(sp, net, 0, 0, iters, max_plies, (void*)s);
}

extern "C" int az_selfplay_run_hash(az_selfplay* sp, int iters, uint64_t seed, int qbits, int max_plies, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_run_hash: null");
    return az_selfplay_run(sp, nullptr, seed, qbits, iters, max_plies, stream);
}

// Continuous plies with the hash evaluator (parity runs): advance + evaluate + service until every game is over.  Returns
// the number of steps.  The slots are NOT in lockstep; trajectories equal those of az_selfplay_run_hash.
extern "C" int az_selfplay_run_async_hash(az_selfplay* sp, int iters, uint64_t seed, int qbits, void* stream) {
    if (!sp || iters < 1) return az_fail(AZ_EINVAL, "az_selfplay_run_async_hash: bad arguments");
    int rc, live = 1, steps = 0;
    if ((rc = az_selfplay_prepare(sp, stream))) return rc;
    if ((rc = az_forest_set_eval_tag(sp->f, az_hash_eval_tag(seed, qbits), (cudaStream_t)stream))) return rc;
    if ((rc = az_forest_begin_move(sp->f, iters, stream))) return rc;
    const long long max_steps = (long long)sp->d.n_games * sp->d.max_plies * (iters + 2) + 64;
    while (live > 0) {
        for (int i = 0; i < 32; ++i) {
            if ((rc = az_forest_advance(sp->f, stream))) return rc;
            if ((rc = az_eval_hash(sp->f, seed, qbits, stream))) return rc;
            if ((rc = az_selfplay_service(sp, (cudaStream_t)stream))) return rc;
            if (++steps % iters == 0) az_forest_cache_tick(sp->f, (cudaStream_t)stream);
        }
        if ((rc = az_selfplay_status(sp, nullptr, nullptr, &live, nullptr, stream))) return rc;
        if (steps > max_steps) return az_fail(AZ_ESTATE, "az_selfplay_run_async_hash: did not converge");
    }
    return steps;
}

// Device pointers of the trajectory store (library-owned, valid until az_selfplay_destroy), for zero-copy consumers
// (NCCL all-gather into the replay buffer): out7 = {g_plies i32[n_games], g_winner i32[n_games], t_words u64[..][2nw],
// t_visits u16[..][A], t_q f64[..], t_action i16[..], t_u f64[..], g_first i32[n_games]}, positions indexed [game][ply],
// max_plies per game.
extern "C" int az_selfplay_buffers(az_selfplay* sp, void** out8, int* n_games, int* max_plies) {
    void** out7 = out8;
    if (!sp || !out8) return az_fail(AZ_EINVAL, "az_selfplay_buffers: null");
    out8[7] = sp->d.g_first;
    out7[0] = sp->d.g_plies; out7[1] = sp->d.g_winner; out7[2] = sp->d.t_words; out7[3] = sp->d.t_visits;
    out7[4] = sp->d.t_q; out7[5] = sp->d.t_action; out7[6] = sp->d.t_u;
    if (n_games) *n_games = sp->d.n_games;
    if (max_plies) *max_plies = sp->d.max_plies;
    return AZ_OK;
}

// Host copies of the same arrays (any pointer may be NULL).  Synchronous.
extern "C" int az_selfplay_read(az_selfplay* sp, int32_t* plies_host, int32_t* winner_host, int32_t* first_host, uint64_t* words_host,
                                uint16_t* visits_host, double* q_host, int16_t* action_host, double* u_host, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_read: null");
    AzForestDev* fd = az_forest_dev(sp->f);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t ng = (size_t)sp->d.n_games, npos = ng * sp->d.max_plies;
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    if (plies_host) AZ_CUDA_OK(cudaMemcpy(plies_host, sp->d.g_plies, ng * 4, cudaMemcpyDeviceToHost));
    if (winner_host) AZ_CUDA_OK(cudaMemcpy(winner_host, sp->d.g_winner, ng * 4, cudaMemcpyDeviceToHost));
    if (first_host) AZ_CUDA_OK(cudaMemcpy(first_host, sp->d.g_first, ng * 4, cudaMemcpyDeviceToHost));
    if (words_host) AZ_CUDA_OK(cudaMemcpy(words_host, sp->d.t_words, npos * 2 * fd->g.nw * 8, cudaMemcpyDeviceToHost));
    if (visits_host) AZ_CUDA_OK(cudaMemcpy(visits_host, sp->d.t_visits, npos * fd->g.action_dim * 2, cudaMemcpyDeviceToHost));
    if (q_host) AZ_CUDA_OK(cudaMemcpy(q_host, sp->d.t_q, npos * 8, cudaMemcpyDeviceToHost));
    if (action_host) AZ_CUDA_OK(cudaMemcpy(action_host, sp->d.t_action, npos * 2, cudaMemcpyDeviceToHost));
    if (u_host) AZ_CUDA_OK(cudaMemcpy(u_host, sp->d.t_u, npos * 8, cudaMemcpyDeviceToHost));
    return AZ_OK;
}

// Slot -> local game (-1 retired) and the noise the next search will use (tests replay searches in the oracle with it).
extern "C" int az_selfplay_read_slots(az_selfplay* sp, int32_t* slot_game_host, double* noise_host, void* stream) {
    if (!sp) return az_fail(AZ_EINVAL, "az_selfplay_read_slots: null");
    AzForestDev* fd = az_forest_dev(sp->f);
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    if (slot_game_host) AZ_CUDA_OK(cudaMemcpy(slot_game_host, sp->d.slot_game, (size_t)fd->n_trees * 4, cudaMemcpyDeviceToHost));
    if (noise_host) AZ_CUDA_OK(cudaMemcpy(noise_host, az_forest_noise_dev(sp->f), (size_t)fd->n_trees * fd->g.action_dim * 8, cudaMemcpyDeviceToHost));
    return AZ_OK;
}

extern "C" long long az_selfplay_launch_count(az_selfplay* sp) { return sp ? sp->launches : 0; }
