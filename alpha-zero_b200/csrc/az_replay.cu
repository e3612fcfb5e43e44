// Device-resident replay buffer + the training step's loss (sm_100a).
//
// Replaces, without a host round trip per sample,
//   train.save_data_to_buffer + rotate_data + flip_data   train.py:8-49   (z, canonical board, 8 dihedral variants)
//   ReplayBuffer.add / sample / size                      replay_buffer.py:8-25,34-35  (deque(maxlen), uniform sample w/o replacement)
//   the loss of train.train                               train.py:101-112 (soft-target CE + MSE; L2 stays a torch reduction)
// The buffer is a ring of ENTRIES -- one per sample the reference's deque would hold, in the same order (game, ply,
// variant) -- in compact form: canonical, already transformed side bitboards, the side-to-move bit of the 3-plane
// games, the transformed root visit counts u16[A] and z as i8: 116 B per 7x7 entry instead of 788 B (f32 board + f64
// policy).  az_replay_add consumes the trajectory arrays of the self-play engine (csrc/az_selfplay.cu) as they come out
// of the NCCL all-gather; az_replay_sample expands B distinct entries into the float32 tensors ReplayBuffer.sample
// returns, straight into the caller's device buffers.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "az_types.h"
#include "az_host.h"
#include "../../include/az_b200.h"

// Source cell of output cell `cell` under dihedral variant v, in the order train.save_data_to_buffer stores them
// (train.py:40-49): 0 rot0, 1 rot0-lr, 2 rot0-tb, 3 rot90, 4 rot90-lr, 5 rot90-tb, 6 rot180, 7 rot270, with
// np.rot90(k, axes=(1,2)) (counter-clockwise) and np.flip.  Square boards; on non-square boards only 0 (identity) and 1
// (left-right mirror) are defined -- the build's Connect4 augmentation (the reference cannot augment non-square boards).
__host__ __device__ __forceinline__ int az_d4_src(int rows, int cols, int v, int cell) {
    int i = cell / cols, j = cell - i * cols;
    if (rows != cols) return v == 1 ? i * cols + (cols - 1 - j) : cell;
    const int n = rows;
    int k = 0, flip = 0;                                    // flip: 0 none, 1 lr, 2 tb (applied AFTER the rotation)
    switch (v) {
        case 1: flip = 1; break;
        case 2: flip = 2; break;
        case 3: k = 1; break;
        case 4: k = 1; flip = 1; break;
        case 5: k = 1; flip = 2; break;
        case 6: k = 2; break;
        case 7: k = 3; break;
        default: break;
    }
    if (flip == 1) j = n - 1 - j;                           // out[i][j] = rot[i][n-1-j]
    if (flip == 2) i = n - 1 - i;                           // out[i][j] = rot[n-1-i][j]
    int si = i, sj = j;                                     // rot[i][j] = in[si][sj]
    if (k == 1) { si = j; sj = n - 1 - i; }
    else if (k == 2) { si = n - 1 - i; sj = n - 1 - j; }
    else if (k == 3) { si = n - 1 - j; sj = i; }
    return si * n + sj;
}

extern "C" int az_d4_source_cell(int rows, int cols, int variant, int cell) {
    if (rows <= 0 || cols <= 0 || variant < 0 || variant > 7 || cell < 0 || cell >= rows * cols) return -1;
    return az_d4_src(rows, cols, variant, cell);
}

struct AzReplayDev {
    AzGame g;
    long long cap;            // entries (deque maxlen)
    int variants;             // samples per position from the third ply on: 8 (square, A == cells), 2 (Connect4)
    int policy_is_board;      // the policy vector is a rows x cols grid (transforms like the board); else A == cols (Connect4)
    uint64_t* words;          // [cap][2 * nw]  canonical: mine, theirs
    uint16_t* visits;         // [cap][A]
    int8_t* z;                // [cap]
    uint8_t* side;            // [cap] side to move (plane 2 of the 3-plane games)
};

namespace {

// One thread per (position, variant).  Positions are [game][ply] of the trajectory arrays; entry_base[game] = entries
// of all earlier games (exclusive scan over the games' entry counts).
template <int NW>
__global__ void az_replay_add_kernel(AzReplayDev rb, int n_games, int max_plies, const int* __restrict__ plies,
                                     const int* __restrict__ winner, const int* __restrict__ first,
                                     const uint64_t* __restrict__ t_words, const uint16_t* __restrict__ t_visits,
                                     const long long* __restrict__ entry_base, long long head, long long skip) {
    const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int V = rb.variants;
    if (t >= (long long)n_games * max_plies * V) return;
    const int v = (int)(t % V);
    const long long pos = t / V;
    const int game = (int)(pos / max_plies), ply = (int)(pos - (long long)game * max_plies);
    const int f0 = first[game], p1 = plies[game];
    if (ply < f0 || ply >= p1) return;
    const int r = ply - f0;                                     // index in the game's recorded list (train.py:33 enumerate)
    if (r < 2 && v > 0) return;                                 // plies 0 and 1 are stored once (train.py:38-39)
    const long long off = entry_base[game] + (r < 2 ? r : 2 + (long long)(r - 2) * V + v);
    if (off < skip) return;                                     // more entries than the ring holds: the oldest are dropped
    const long long e = (head + off) % rb.cap;
    const AzGame& g = rb.g;
    const int A = g.action_dim, cells = g.state_dim;
    const int side = ply & 1;                                   // ply i is played by side i % 2 (train.py:34,37)
    const int w = winner[game];
    rb.z[e] = (int8_t)(w < 0 ? 0 : (side == w ? 1 : -1));       // train.py:35: +reward for the winner's plies, -reward else; draw 0
    rb.side[e] = (uint8_t)side;
    const uint64_t* src = t_words + (size_t)pos * 2 * NW;
    uint64_t mine[NW], theirs[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) mine[i] = theirs[i] = 0ull;
    for (int c = 0; c < cells; ++c) {                           // canonical (side to move first, gomoku.py:34-40), then the variant
        const int sc = az_d4_src(g.rows, g.cols, v, c);
        const uint64_t m = (src[(side ? NW : 0) + (sc >> 6)] >> (sc & 63)) & 1ull;
        const uint64_t o = (src[(side ? 0 : NW) + (sc >> 6)] >> (sc & 63)) & 1ull;
#pragma unroll
        for (int i = 0; i < NW; ++i)
            if (NW == 1 || i == (c >> 6)) {
                mine[i] |= m << (c & 63);
                theirs[i] |= o << (c & 63);
            }
    }
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        rb.words[(size_t)e * 2 * NW + i] = mine[i];
        rb.words[(size_t)e * 2 * NW + NW + i] = theirs[i];
    }
    const uint16_t* vs = t_visits + (size_t)pos * A;
    uint16_t* vd = rb.visits + (size_t)e * A;
    for (int a = 0; a < A; ++a) {
        int sa = a;
        if (rb.policy_is_board) sa = az_d4_src(g.rows, g.cols, v, a);
        else if (v == 1) sa = A - 1 - a;                        // Connect4 mirror: policy[::-1]
        vd[a] = vs[sa];
    }
}

// Feistel permutation of [0, n) with cycle walking: i -> a keyed bijection, so the first B images are B distinct
// indices -- np.random.choice(len, B, replace=False) (replay_buffer.py:16) without a shuffle buffer.
__device__ __forceinline__ uint32_t az_mix32(uint32_t x, uint32_t k) {
    x ^= k;
    x *= 0x9E3779B1u;
    x ^= x >> 15;
    x *= 0x85EBCA77u;
    x ^= x >> 13;
    x *= 0xC2B2AE3Du;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ unsigned long long az_feistel(unsigned long long i, unsigned long long n, int half_bits, uint64_t key) {
    const uint32_t mask = half_bits >= 32 ? 0xffffffffu : ((1u << half_bits) - 1u);
    unsigned long long x = i;
    do {
        uint32_t l = (uint32_t)(x >> half_bits) & mask, r = (uint32_t)x & mask;
#pragma unroll
        for (int round = 0; round < 6; ++round) {
            const uint32_t f = az_mix32(r, (uint32_t)(key >> (round * 8)) + 0x632BE5ABu * (uint32_t)(round + 1) + (uint32_t)(key >> 32)) & mask;
            const uint32_t nl = r;
            r = l ^ f;
            l = nl;
        }
        x = ((unsigned long long)l << half_bits) | r;
    } while (x >= n);                                           // cycle walking keeps the map a bijection on [0, n)
    return x;
}

// One block per sample: entry -> float32 planes [C][rows][cols], policy f32[A] = (float)(n_a / sum n) (utils.py:46-55
// in f64, then the float32 cast of replay_buffer.py:22), z f32[1].  start >= 0: entries start + b instead of a draw.
template <int NW>
__global__ void az_replay_sample_kernel(AzReplayDev rb, long long size, long long tail, int B, uint64_t key, int half_bits,
                                        long long start, float* __restrict__ states, float* __restrict__ policies,
                                        float* __restrict__ zs, long long* __restrict__ indices_out) {
    const int b = blockIdx.x;
    if (b >= B) return;
    const long long idx = start >= 0 ? start + b : (long long)az_feistel((unsigned long long)b, (unsigned long long)size, half_bits, key);
    const long long e = (tail + idx) % rb.cap;                  // deque index idx = idx-th oldest entry
    const AzGame& g = rb.g;
    const int A = g.action_dim, cells = g.state_dim, C = g.planes;
    const uint64_t* w = rb.words + (size_t)e * 2 * NW;
    const float sidev = (float)rb.side[e];
    float* st = states + (size_t)b * C * cells;
    for (int t = threadIdx.x; t < C * cells; t += blockDim.x) {
        const int pl = t / cells, c = t - pl * cells;
        st[t] = pl < 2 ? (float)((w[pl * NW + (c >> 6)] >> (c & 63)) & 1ull) : sidev;
    }
    __shared__ int s_sum;
    if (threadIdx.x == 0) {
        int s = 0;
        for (int a = 0; a < A; ++a) s += rb.visits[(size_t)e * A + a];
        s_sum = s;
    }
    __syncthreads();
    const double tot = (double)s_sum;
    for (int a = threadIdx.x; a < A; a += blockDim.x)
        policies[(size_t)b * A + a] = s_sum ? (float)((double)rb.visits[(size_t)e * A + a] / tot) : 0.f;
    if (threadIdx.x == 0) {
        zs[b] = (float)rb.z[e];
        if (indices_out) indices_out[b] = idx;
    }
}

// train.py:110-112 for a batch, forward and backward in one pass, one warp per row:
//   policy_loss = mean_b( sum_a -pi * log_softmax(logits) ),  value_loss = mean_b (v - z)^2
//   dlogits = (softmax * sum_a pi - pi) / B,  dvalues = 2 (v - z) / B.   out2[0] += row policy loss / B, out2[1] += (v - z)^2 / B
__global__ void az_loss_kernel(const float* __restrict__ logits, const float* __restrict__ values, const float* __restrict__ pis,
                               const float* __restrict__ zs, int B, int A, float* __restrict__ dlogits, float* __restrict__ dvalues,
                               float* __restrict__ out2) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= B) return;
    const float* l = logits + (size_t)row * A;
    const float* p = pis + (size_t)row * A;
    float mx = -INFINITY;
    for (int a = lane; a < A; a += 32) mx = fmaxf(mx, l[a]);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float se = 0.f, sp = 0.f, spl = 0.f;
    for (int a = lane; a < A; a += 32) {
        se += expf(l[a] - mx);
        sp += p[a];
        spl += p[a] * (l[a] - mx);
    }
    for (int o = 16; o; o >>= 1) {
        se += __shfl_xor_sync(0xffffffffu, se, o);
        sp += __shfl_xor_sync(0xffffffffu, sp, o);
        spl += __shfl_xor_sync(0xffffffffu, spl, o);
    }
    const float lse = logf(se);
    const float inv_b = 1.0f / (float)B;
    for (int a = lane; a < A; a += 32) dlogits[(size_t)row * A + a] = (expf(l[a] - mx - lse) * sp - p[a]) * inv_b;
    if (lane == 0) {
        const float d = values[row] - zs[row];
        dvalues[row] = 2.0f * d * inv_b;
        atomicAdd(out2, (sp * lse - spl) * inv_b);              // sum_a -pi (l - mx - lse)
        atomicAdd(out2 + 1, d * d * inv_b);
    }
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

struct az_replay {
    AzReplayDev d{};
    long long head = 0;       // ring index of the next entry
    long long size = 0;       // entries held (<= cap)
    std::vector<void*> allocs;
    long long* base_dev = nullptr;
    int base_cap = 0;
};

extern "C" void az_replay_destroy(az_replay* rb) {
    if (!rb) return;
    cudaDeviceSynchronize();
    for (void* p : rb->allocs) cudaFree(p);
    if (rb->base_dev) cudaFree(rb->base_dev);
    delete rb;
}

extern "C" int az_replay_create(int game, int rows, int cols, long long capacity, az_replay** out) {
    if (!out || capacity <= 0) return az_fail(AZ_EINVAL, "az_replay_create: bad arguments");
    AzGame g;
    int rc = az_make_game(game, rows, cols, &g);
    if (rc) return rc;
    az_replay* rb = new az_replay();
    rb->d.g = g;
    rb->d.cap = capacity;
    rb->d.policy_is_board = g.action_dim == g.state_dim ? 1 : 0;
    rb->d.variants = (g.rows == g.cols && g.action_dim == g.state_dim) ? 8 : 2;       // train.py:40-49 / Connect4: identity + mirror
    auto alloc = [&](void** p, size_t bytes) {
        if (cudaMalloc(p, bytes + 16) != cudaSuccess) return false;
        rb->allocs.push_back(*p);
        return true;
    };
    if (!alloc((void**)&rb->d.words, (size_t)capacity * 2 * g.nw * 8) || !alloc((void**)&rb->d.visits, (size_t)capacity * g.action_dim * 2) ||
        !alloc((void**)&rb->d.z, (size_t)capacity) || !alloc((void**)&rb->d.side, (size_t)capacity)) {
        cudaGetLastError();
        az_replay_destroy(rb);
        return az_fail(AZ_ENOMEM, "az_replay_create: cudaMalloc failed");
    }
    *out = rb;
    return AZ_OK;
}

extern "C" long long az_replay_size(az_replay* rb) { return rb ? rb->size : 0; }
extern "C" long long az_replay_entry_bytes(az_replay* rb) {
    return rb ? (long long)(2 * rb->d.g.nw * 8 + rb->d.g.action_dim * 2 + 2) : 0;
}
extern "C" int az_replay_clear(az_replay* rb) {
    if (!rb) return az_fail(AZ_EINVAL, "az_replay_clear: null");
    rb->head = rb->size = 0;
    return AZ_OK;
}

// Appends every recorded position of n_games finished games (device arrays in the layout of az_selfplay_buffers, games
// in the order they are to be stored) as train.save_data_to_buffer would.  Returns the number of entries appended.
extern "C" long long az_replay_add(az_replay* rb, int n_games, int max_plies, const int32_t* plies_dev, const int32_t* winner_dev,
                                   const int32_t* first_dev, const uint64_t* words_dev, const uint16_t* visits_dev, void* stream) {
    if (!rb || n_games < 0 || !plies_dev || !winner_dev || !first_dev || !words_dev || !visits_dev)
        return az_fail(AZ_EINVAL, "az_replay_add: bad arguments");
    if (n_games == 0) return 0;
    if (max_plies != rb->d.g.state_dim) return az_fail(AZ_EINVAL, "az_replay_add: max_plies must be rows * cols");
    cudaStream_t s = (cudaStream_t)stream;
    std::vector<int32_t> plies(n_games), first(n_games);
    AZ_CUDA_OK(cudaMemcpyAsync(plies.data(), plies_dev, (size_t)n_games * 4, cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaMemcpyAsync(first.data(), first_dev, (size_t)n_games * 4, cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    std::vector<long long> base(n_games);
    long long total = 0;
    const int V = rb->d.variants;
    for (int k = 0; k < n_games; ++k) {
        base[k] = total;
        const int p = std::max(0, plies[k] - first[k]);
        total += std::min(p, 2) + (long long)std::max(p - 2, 0) * V;
    }
    if (total == 0) return 0;
    if (n_games > rb->base_cap) {
        if (rb->base_dev) cudaFree(rb->base_dev);
        rb->base_dev = nullptr;
        AZ_CUDA_OK(cudaMalloc(&rb->base_dev, (size_t)n_games * 8));
        rb->base_cap = n_games;
    }
    AZ_CUDA_OK(cudaMemcpyAsync(rb->base_dev, base.data(), (size_t)n_games * 8, cudaMemcpyHostToDevice, s));
    const long long skip = total > rb->d.cap ? total - rb->d.cap : 0;
    const long long work = (long long)n_games * max_plies * V;
    const int grid = (int)((work + 127) / 128);
    if (rb->d.g.nw == 1)
        az_replay_add_kernel<1><<<grid, 128, 0, s>>>(rb->d, n_games, max_plies, plies_dev, winner_dev, first_dev, words_dev, visits_dev,
                                                     rb->base_dev, rb->head, skip);
    else
        az_replay_add_kernel<AZ_MAX_WORDS><<<grid, 128, 0, s>>>(rb->d, n_games, max_plies, plies_dev, winner_dev, first_dev, words_dev,
                                                                visits_dev, rb->base_dev, rb->head, skip);
    AZ_CUDA_OK(cudaGetLastError());
    AZ_CUDA_OK(cudaStreamSynchronize(s));                       // `base` (host) was the source of an async copy
    rb->head = (rb->head + total) % rb->d.cap;
    rb->size = std::min(rb->d.cap, rb->size + total);
    return total;
}

static int az_replay_expand(az_replay* rb, int B, long long start, uint64_t key, float* states_dev, float* policies_dev, float* z_dev,
                            long long* indices_dev, cudaStream_t s) {
    int half_bits = 1;
    while (half_bits < 32 && (1ull << (2 * half_bits)) < (unsigned long long)rb->size) ++half_bits;
    const long long tail = ((rb->head - rb->size) % rb->d.cap + rb->d.cap) % rb->d.cap;
    if (rb->d.g.nw == 1)
        az_replay_sample_kernel<1><<<B, 64, 0, s>>>(rb->d, rb->size, tail, B, key, half_bits, start, states_dev, policies_dev, z_dev, indices_dev);
    else
        az_replay_sample_kernel<AZ_MAX_WORDS><<<B, 64, 0, s>>>(rb->d, rb->size, tail, B, key, half_bits, start, states_dev, policies_dev, z_dev,
                                                               indices_dev);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

// ReplayBuffer.sample (replay_buffer.py:15-25): B distinct entries, uniformly -> float32 [B][C][rows][cols], [B][A], [B][1] in
// the caller's device buffers.  The draw is a keyed permutation of (seed, counter); indices_dev (optional) gets the entry numbers.
extern "C" int az_replay_sample(az_replay* rb, int B, uint64_t seed, uint64_t counter, float* states_dev, float* policies_dev,
                                float* z_dev, long long* indices_dev, void* stream) {
    if (!rb || B <= 0 || !states_dev || !policies_dev || !z_dev) return az_fail(AZ_EINVAL, "az_replay_sample: bad arguments");
    if (B > rb->size) return az_fail(AZ_EINVAL, "az_replay_sample: cannot take a larger sample than the buffer holds (replace=False)");
    uint64_t key = seed * 0x9E3779B97F4A7C15ull + counter * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull;
    key ^= key >> 31;
    key *= 0xBF58476D1CE4E5B9ull;
    key ^= key >> 29;
    return az_replay_expand(rb, B, -1, key, states_dev, policies_dev, z_dev, indices_dev, (cudaStream_t)stream);
}

// Entries start .. start+count-1 in deque order (oldest first), expanded like a sample: parity tests, persistence.
extern "C" int az_replay_read(az_replay* rb, long long start, int count, float* states_dev, float* policies_dev, float* z_dev, void* stream) {
    if (!rb || count <= 0 || start < 0 || start + count > rb->size || !states_dev || !policies_dev || !z_dev)
        return az_fail(AZ_EINVAL, "az_replay_read: range outside the buffer");
    return az_replay_expand(rb, count, start, 0, states_dev, policies_dev, z_dev, nullptr, (cudaStream_t)stream);
}

// train.py:110-112: loss and its gradient for one batch (all device pointers; out2 = {policy_loss, value_loss}, zeroed here).
extern "C" int az_loss_policy_value(const float* logits_dev, const float* values_dev, const float* pis_dev, const float* zs_dev, int B, int A,
                                    float* dlogits_dev, float* dvalues_dev, float* out2_dev, void* stream) {
    if (!logits_dev || !values_dev || !pis_dev || !zs_dev || !dlogits_dev || !dvalues_dev || !out2_dev || B <= 0 || A <= 0)
        return az_fail(AZ_EINVAL, "az_loss_policy_value: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaMemsetAsync(out2_dev, 0, 8, s));
    az_loss_kernel<<<(B * 32 + 127) / 128, 128, 0, s>>>(logits_dev, values_dev, pis_dev, zs_dev, B, A, dlogits_dev, dvalues_dev, out2_dev);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}
