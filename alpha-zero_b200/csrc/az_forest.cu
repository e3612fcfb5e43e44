// Forest kernels: PUCT select, expand, backup for thousands of concurrent trees (sm_100a).
//
// Replaces the per-node Python of the reference's simulation loop:
//   MCTS.mcts                         ai/mcts.py:11-60
//   Node.select / expand / backup     ai/node.py:42-74
//   calcUcbOfChildrenFromParent       utils.py:29-44  ('network' branch, c_puct = 1)
//   add_dirichlet_noise (mixing only) utils.py:24-25
// One thread GROUP (8 or 32 lanes) owns one tree.  One launch of az_advance_kernel
//   A. consumes the evaluator's answers for the tree's pending leaves: generate legal moves
//      in reference order, write the child block (coalesced), back the value up the path;
//   B. runs new descents until `vl` leaves wait for evaluation (terminal leaves are backed
//      up on the spot and cost a simulation each, exactly like `continue` in mcts.py:28-32).
// Compiled with -fmad=false: every score is the reference's rounding sequence.
#include <cstdio>
#include <cstdlib>
#include "az_forest.h"
#include "az_rules.cuh"

namespace {

template <int G>
__device__ __forceinline__ unsigned az_group_mask() {
    if constexpr (G == 32) return 0xffffffffu;
    else return ((1u << (G & 31)) - 1u) << (((threadIdx.x & 31) / G) * G);
}

// fp32 PUCT of the reference under NumPy >= 2 promotion:
//   s = (float)sqrt((double)Np); t = P*s; u = t/(float)(n+1); q = (float)(W/n); ucb = q+u
// q is cached in the node by the backup that last changed W and n (az_backup_node); s is formed with
// the fp32 square root, which equals the rounded f64 root for every Np < 2^24 (53 >= 2*24+2 bits;
// checked exhaustively in tests/test_oracle_golden.py).
__device__ __forceinline__ float az_puct32(float q, float P, int n, float s) {
    const float u = __fdiv_rn(__fmul_rn(P, s), (float)(n + 1));
    if (n == 0) return u;                                     // utils.py:32-35
    return __fadd_rn(q, u);                                   // utils.py:40
}
__device__ __forceinline__ float az_q32(double W, int n) { return __double2float_rn(__ddiv_rn(W, (double)n)); }

// node.py:69-72 for one node of the path: visit += 1; value += v; (+ refresh the cached quotient)
// Views of the node array that keep the kernel text in terms of hot / meta / W.
struct AzHotView {
    AzNode* p;
    __device__ __forceinline__ AzHot& operator[](size_t i) const { return p[i].hot; }
};
struct AzMetaView {
    AzNode* p;
    __device__ __forceinline__ AzMeta& operator[](size_t i) const { return p[i].meta; }
};
struct AzMeta64View {
    AzNode* p;
    __device__ __forceinline__ unsigned long long operator[](size_t i) const { return *reinterpret_cast<const unsigned long long*>(&p[i].meta); }
};

// f64 variant: root children when Dirichlet noise made their priors np.float64.
__device__ __forceinline__ double az_puct64(double W, double P, int n, double s) {
    const double u = __ddiv_rn(__dmul_rn(P, s), (double)(n + 1));
    if (n == 0) return u;
    return __dadd_rn(__ddiv_rn(W, (double)n), u);
}

// Tree records are re-read every step, but between two advance launches the evaluator streams ~40 MB of activations and
// weights through the L2: without a hint most child scans came from DRAM (ncu: 5.7 MB DRAM read per launch against 2.2 MB of
// algorithmic bytes) and a select level cost ~1.9 k cycles.  Node loads carry an L2 evict_last policy so the streamed data
// is evicted before them.  AZ_NO_L2_KEEP (build) switches the hint off.
__device__ __forceinline__ uint64_t az_l2_keep_policy() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void az_load_node(const AzNode* nd, uint64_t pol, AzHot& h, ulonglong2& mw) {
#ifdef AZ_NO_L2_KEEP
    h = nd->hot;
    mw = *reinterpret_cast<const ulonglong2*>(&nd->meta);
#else
    uint32_t a0, a1, a2, a3;
    asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "l"(nd), "l"(pol) : "memory");
    asm volatile("ld.global.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;" : "=l"(mw.x), "=l"(mw.y) : "l"(reinterpret_cast<const char*>(nd) + 16), "l"(pol) : "memory");
    h.N = (int)a0;
    h.Q = __uint_as_float(a1);
    h.P = __uint_as_float(a2);
    h.V = a3;
#endif
}

__device__ __forceinline__ void az_store16_keep(void* p, const uint4& v, uint64_t pol) {
#ifdef AZ_NO_L2_KEEP
    *reinterpret_cast<uint4*>(p) = v;
#else
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
#endif
}
__device__ __forceinline__ void az_store8_keep(void* p, unsigned long long v, uint64_t pol) {
#ifdef AZ_NO_L2_KEEP
    *reinterpret_cast<unsigned long long*>(p) = v;
#else
    asm volatile("st.global.L2::cache_hint.u64 [%0], %1, %2;" ::"l"(p), "l"(v), "l"(pol) : "memory");
#endif
}

// Scalar loads with the same evict_last policy (per-tree state, pending-leaf records, the evaluator's answers).
#ifdef AZ_NO_L2_KEEP
template <typename T> __device__ __forceinline__ T az_ldk(const T* p, uint64_t) { return *p; }
#else
__device__ __forceinline__ uint8_t az_ldk(const uint8_t* p, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.L2::cache_hint.u8 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol) : "memory");
    return (uint8_t)v;
}
__device__ __forceinline__ uint16_t az_ldk(const uint16_t* p, uint64_t pol) {
    uint16_t v;
    asm volatile("ld.global.L2::cache_hint.u16 %0, [%1], %2;" : "=h"(v) : "l"(p), "l"(pol) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t az_ldk(const uint32_t* p, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol) : "memory");
    return v;
}
__device__ __forceinline__ int az_ldk(const int* p, uint64_t pol) { return (int)az_ldk(reinterpret_cast<const uint32_t*>(p), pol); }
__device__ __forceinline__ float az_ldk(const float* p, uint64_t pol) { return __uint_as_float(az_ldk(reinterpret_cast<const uint32_t*>(p), pol)); }
__device__ __forceinline__ uint64_t az_ldk(const uint64_t* p, uint64_t pol) {
    uint64_t v;
    asm volatile("ld.global.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol) : "memory");
    return v;
}
#endif

// node.py:69-72 for one path node: visit += 1, value += v (f64, visitation order), cached Q refreshed.
__device__ __forceinline__ void az_backup_node(AzNode* nodes, uint32_t nd, double v, bool release_vl, uint64_t pol) {
    AzHot h;
    ulonglong2 mw;
    az_load_node(&nodes[nd], pol, h, mw);
    const double W = __dadd_rn(__longlong_as_double((long long)mw.y), v);
    h.N += 1;
    h.Q = az_q32(W, h.N);
    if (release_vl) h.V -= 1;
    az_store16_keep(&nodes[nd].hot, make_uint4((uint32_t)h.N, __float_as_uint(h.Q), __float_as_uint(h.P), h.V), pol);
    az_store8_keep(&nodes[nd].W, (unsigned long long)__double_as_longlong(W), pol);
}

// Vanilla mode (model=None), utils.py:36,43 in Python floats: sqrt(2 ln Np / (n+1)) (+ W/n once visited); an
// unvisited child is NOT +inf.  two_ln = 2 * ln(Np) with ln from the host's C library (log_tab), so the score is
// the reference's bit for bit.
__device__ __forceinline__ double az_ucb1(double W, int n, double two_ln) {
    const double e = __dsqrt_rn(__ddiv_rn(two_ln, (double)(n + 1)));
    if (n == 0) return e;
    return __dadd_rn(__ddiv_rn(W, (double)n), e);
}

// First maximum in child order over the lanes of a group with the warp-reduce unit (redux.sync):
// scores become order-preserving unsigned keys (+0.0 canonicalises -0.0; a lane never holds NaN
// because `sc > best` is false for NaN), the winner is the smallest child index among the lanes
// that hold the maximum key.  2 (fp32) / 3 (fp64) reductions instead of 5 shuffle rounds.
__device__ __forceinline__ unsigned az_ord32(float f) {
    const unsigned u = __float_as_uint(f + 0.0f);
    return u ^ ((unsigned)((int)u >> 31) | 0x80000000u);
}
__device__ __forceinline__ unsigned long long az_ord64(double d) {
    const unsigned long long u = (unsigned long long)__double_as_longlong(d + 0.0);
    return u ^ ((unsigned long long)((long long)u >> 63) | 0x8000000000000000ull);
}
__device__ __forceinline__ int az_group_argmax(unsigned gmask, float best, int idx) {
    const unsigned k = az_ord32(best);
    const unsigned m = __reduce_max_sync(gmask, k);
    return (int)__reduce_min_sync(gmask, k == m ? (unsigned)idx : 0x7fffffffu);
}
__device__ __forceinline__ int az_group_argmax(unsigned gmask, double best, int idx) {
    const unsigned long long k = az_ord64(best);
    const unsigned hi = (unsigned)(k >> 32), lo = (unsigned)k;
    const unsigned mh = __reduce_max_sync(gmask, hi);
    const unsigned ml = __reduce_max_sync(gmask, hi == mh ? lo : 0u);
    return (int)__reduce_min_sync(gmask, (hi == mh && lo == ml) ? (unsigned)idx : 0x7fffffffu);
}

// ---- evaluation cache (az_forest.h: AzCacheDev) ---------------------------------------------------
__device__ __forceinline__ uint64_t az_cache_hash(uint64_t mine, uint64_t theirs) {
    uint64_t x = mine * 0x9E3779B97F4A7C15ull ^ (theirs + 0xD1B54A32D192ED03ull) * 0xBF58476D1CE4E5B9ull;
    x ^= x >> 29;
    x *= 0x94D049BB133111EBull;
    return x ^ (x >> 32);
}
__device__ __forceinline__ uint32_t az_cache_word(uint32_t kind, uint32_t epoch) { return (kind << 30) | (epoch & 0xffu); }

// Warp-cooperative lookup of one canonical position.  Returns
//   2  the entry holds the answer already (AZ_CK_READY): usable at once,
//   1  another leaf of THIS launch claimed the entry and its evaluation is in flight: the answer is in the entry by the
//      next launch (az_cache_publish_kernel runs before it),
//   0  miss; *entry = the entry claimed for this leaf (its batch row will be published there) or 0xffffffff when the
//      window had no free entry (the leaf is evaluated and simply not cached).
// Every state transition is a CAS on the entry's state word, so a reader that found a stale entry and a writer that wants
// to recycle it cannot both win; the key is written while the entry is LOCKED and only then made PENDING.
template <int G>
__device__ __noinline__ int az_cache_probe(const AzCacheDev c, uint64_t mine, uint64_t theirs, int lane, unsigned gmask, uint32_t* entry) {
    const uint32_t epoch = __ldcg(c.epoch);
    const uint32_t base = (uint32_t)az_cache_hash(mine, theirs) & (c.cap - 1u) & ~(uint32_t)(AZ_CACHE_WINDOW - 1);
    ulonglong2 k = make_ulonglong2(0ull, 0ull);
    uint32_t st = 0u;
    const bool in = lane < AZ_CACHE_WINDOW;
    if (in) {
        k = __ldcg(c.keys + base + lane);
        st = __ldcg(c.state + base + lane);
    }
    const uint32_t kind = st >> 30, age = (epoch - st) & 0xffu;
    const bool same = in && k.x == mine && k.y == theirs;
    const unsigned mb = __ballot_sync(gmask, same && (kind == AZ_CK_READY || (kind == AZ_CK_PENDING && age == 0u))) & gmask;
    if (mb) {
        const int l = (__ffs(mb) - 1) % G;
        const uint32_t e = base + (uint32_t)l;
        const uint32_t sl = __shfl_sync(gmask, st, l, G);
        *entry = e;
        if ((sl >> 30) == AZ_CK_PENDING) return 1;
        const uint32_t al = (epoch - sl) & 0xffu;
        if (al == 0u) return 2;
        int ok = 1;                                          // touch: a stale entry must be re-validated atomically
        if (lane == 0) ok = atomicCAS(c.state + e, sl, az_cache_word(AZ_CK_READY, epoch)) == sl;
        ok = __shfl_sync(gmask, ok, 0, G);
        if (ok || al < AZ_CACHE_STALE) return 2;             // a failed refresh of a fresh entry: somebody else refreshed it
        *entry = 0xffffffffu;                                // lost a stale entry to a writer: plain miss, not cached
        return 0;
    }
    unsigned cb = __ballot_sync(gmask, in && (kind == AZ_CK_EMPTY || (kind == AZ_CK_READY && age >= AZ_CACHE_STALE) ||
                                              (kind == AZ_CK_PENDING && age >= 1u))) & gmask;
    while (cb) {
        const int l = (__ffs(cb) - 1) % G;
        cb &= cb - 1;
        const uint32_t e = base + (uint32_t)l;
        const uint32_t old = __shfl_sync(gmask, st, l, G);
        int ok = 0;
        if (lane == 0) {
            ok = atomicCAS(c.state + e, old, az_cache_word(AZ_CK_LOCKED, epoch)) == old;
            if (ok) {
                __stcg(c.keys + e, make_ulonglong2(mine, theirs));
                __threadfence();
                atomicExch(c.state + e, az_cache_word(AZ_CK_PENDING, epoch));
            }
        }
        ok = __shfl_sync(gmask, ok, 0, G);
        if (ok) {
            *entry = e;
            return 0;
        }
    }
    *entry = 0xffffffffu;
    return 0;
}

// ---- vl == 1: backup + cached best child ------------------------------------------------------------------
// At vl == 1 exactly one simulation of a tree is in flight, so between two visits of a node nothing it scores changes
// except through the backup of the simulation that passed through it.  The backup therefore also computes, for EVERY
// path node, the PUCT argmax the NEXT visit of that node will take -- same inputs, same arithmetic, same first-maximum
// rule as Node.select (ai/node.py:42-47, utils.py:29-44) -- and stores it in the node's V field (unused without
// virtual loss).  The next descent is then a pointer walk: one uniform 32-byte load per level instead of a scan +
// scoring + warp reduction per level.  Work split: (1) lane i updates path node i (visit += 1, value += v,
// node.py:69-72) and publishes the record in shared memory; (2) 8-lane groups score the children of one path node
// each, four path nodes per pass, one child per lane and pass of 8 -- the scoring stays lane-parallel (issue slots, not
// memory, bound this kernel) and the child fetches of four levels share one round trip; (3) lane i stores node i.
// nd0 / nd1: path entries lane and lane + 32; `result` is credited to the last path node, alternating sign upwards;
// leaf_best: argmax among the fresh children of the last node (0 for a terminal leaf).
// Returns the root's new best child (all lanes).  *reads += children scored (statistics).
struct __attribute__((aligned(8))) AzBestRec {   // 32 bytes
    uint32_t nd;
    int N;
    float Q, P;
    double W;
    uint32_t cb;
    uint16_t nc, best;
};
#define AZ_BEST_SMEM (64 * 32)
template <int G>
__device__ __noinline__ int az_backup_best(AzNode* nodes, uint32_t nd0, uint32_t nd1, int len, double result, int leaf_best,
                                           const double* root_p64, int lane, unsigned gmask, uint64_t pol, unsigned* reads,
                                           AzBestRec* rec) {
    for (int h = 0; h * 32 < len; ++h) {                      // (1) own records
        const int i = lane + 32 * h;
        if (i < len) {
            const uint32_t nd = h ? nd1 : nd0;
            AzHot hh;
            ulonglong2 mw;
            az_load_node(&nodes[nd], pol, hh, mw);
            const AzMeta m = *reinterpret_cast<const AzMeta*>(&mw.x);
            const double v = ((len - 1 - i) & 1) ? -result : result;
            AzBestRec r;
            r.nd = nd;
            r.W = __dadd_rn(__longlong_as_double((long long)mw.y), v);
            r.N = hh.N + 1;
            r.Q = az_q32(r.W, r.N);
            r.P = hh.P;
            r.cb = m.child_base;
            r.nc = m.nchild;
            r.best = 0;
            rec[i] = r;
        }
    }
    __syncwarp(gmask);
    const int k = lane & 7, grp = lane >> 3;                  // (2) children of path nodes p0 .. p0+3, 8 lanes each
    const unsigned m8 = 0xffu << (grp * 8);
    unsigned scored = 0;
    for (int p0 = 0; p0 < len - 1; p0 += 4) {
        const int j = p0 + grp;
        if (j >= len - 1) continue;                           // group-uniform
        const AzBestRec r = rec[j];
        const uint32_t x_nd = rec[j + 1].nd;                  // the child on the path: its record in memory is stale
        const int nc = r.nc;
        int bi = 0x7fffffff;
        if (j == 0 && root_p64) {                             // noisy root: f64 priors, f64 scores (utils.py:24-25)
            const double s = __dsqrt_rn((double)r.N);
            double best = -1.0 / 0.0;
            for (int c = k; c < nc; c += 8) {
                int n;
                double W;
                if (r.cb + c == x_nd) {
                    n = rec[1].N;
                    W = rec[1].W;
                } else {
                    n = (int)az_ldk(reinterpret_cast<const uint32_t*>(&nodes[r.cb + c].hot.N), pol);
                    W = __longlong_as_double((long long)az_ldk(reinterpret_cast<const uint64_t*>(&nodes[r.cb + c].W), pol));
                }
                const double sc = az_puct64(W, root_p64[c], n, s);
                if (sc > best) { best = sc; bi = c; }
            }
            bi = az_group_argmax(m8, best, bi);
        } else {
            const float s = __fsqrt_rn((float)r.N);
            float best = -1.0f / 0.0f;
            for (int c = k; c < nc; c += 8) {
                uint32_t cn, cq, cp, unused;
                asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                             : "=r"(cn), "=r"(cq), "=r"(cp), "=r"(unused) : "l"(&nodes[r.cb + c]), "l"(pol) : "memory");
                if (r.cb + c == x_nd) {
                    cn = (uint32_t)rec[j + 1].N;
                    cq = __float_as_uint(rec[j + 1].Q);
                }
                const float sc = az_puct32(__uint_as_float(cq), __uint_as_float(cp), (int)cn, s);
                if (sc > best) { best = sc; bi = c; }
            }
            bi = az_group_argmax(m8, best, bi);
        }
        if (k == 0) {
            rec[j].best = (uint16_t)(bi == 0x7fffffff ? 0 : bi);   // all-NaN guard: max() keeps the first
            scored += (unsigned)nc;
        }
    }
    __syncwarp(gmask);
    for (int h = 0; h * 32 < len; ++h) {                      // (3) store
        const int i = lane + 32 * h;
        if (i < len) {
            const AzBestRec r = rec[i];
            const uint32_t best = i == len - 1 ? (uint32_t)leaf_best : (uint32_t)r.best;
            az_store16_keep(&nodes[r.nd].hot, make_uint4((uint32_t)r.N, __float_as_uint(r.Q), __float_as_uint(r.P), best), pol);
            az_store8_keep(&nodes[r.nd].W, (unsigned long long)__double_as_longlong(r.W), pol);
        }
    }
    *reads += __reduce_add_sync(gmask, scored);
    const int root_best = len == 1 ? leaf_best : (int)rec[0].best;
    __syncwarp(gmask);                                        // rec is reused by the next backup
    return root_best;
}

// Expansion of up to AZ_LP pending leaves of a tree runs in parallel (one lane per leaf generates
// that leaf's legal moves -- for Gomoku the sequential CPython-set emulation -- into its own scratch).
#define AZ_LP 8

#ifdef AZ_ADV_TRACE
// Per-tree phase clocks, accumulated over launches (tools/adv_trace.py): [tree][16]
__device__ unsigned long long g_adv_trace[8192][16];
__device__ unsigned int g_adv_hist[66];      // warp lifetimes of the launches since the last read: 64 buckets of 32k cycles, [64] = max, [65] = first-to-last clock span / 64
#define AZ_AT(k)                                                                    \
    do {                                                                            \
        const long long t__ = clock64();                                            \
        if (lane == 0 && tree < 8192) g_adv_trace[tree][k] += (unsigned long long)(t__ - at_prev); \
        at_prev = t__;                                                              \
    } while (0)
#define AZ_AT_COUNT(k, n) do { if (lane == 0 && tree < 8192) g_adv_trace[tree][k] += (n); } while (0)
#else
#define AZ_AT(k) do {} while (0)
#define AZ_AT_COUNT(k, n) do {} while (0)
#endif

// The legal-move listing as a real call: inlined, its loops and 64-bit masks compete for the 72 registers the kernel
// is held to (7 CTAs per SM) and the allocator spills in the select loop instead.
#ifndef AZ_LEGAL_INLINE
template <int NW>
__device__ __noinline__ int az_legal_moves_call(AzGame g, AzBoard<NW> b, const uint64_t* cell_hash, uint8_t* lm, int moves_pad) {
    return az_legal_moves<NW>(g, b, cell_hash, lm + moves_pad, lm + moves_pad + g.set_slots, lm);
}
#endif

template <int NW, int G>
__global__ void __launch_bounds__(128, 7) az_advance_kernel(AzForestDev f, int smem_per_group) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x % G;
    const int grp = threadIdx.x / G;
    const int tree = blockIdx.x * (blockDim.x / G) + grp;
    const unsigned gmask = az_group_mask<G>();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the evaluator's first kernel may queue behind this grid
    if (tree >= f.n_trees) return;
#ifdef AZ_ADV_TRACE
    long long at_prev = clock64();
    const long long at_t0 = at_prev;
#endif
    // independent loads first, decisions after: one memory round trip
    const uint64_t l2keep = az_l2_keep_policy();
    const uint8_t is_active = az_ldk(f.active + tree, l2keep), is_done = az_ldk(f.done_flag + tree, l2keep);
    int n_nodes = az_ldk(f.n_nodes + tree, l2keep);
    int sims = az_ldk(f.sims_done + tree, l2keep);
    const int npend0 = az_ldk(f.n_pending + tree, l2keep);
    const int root_player = az_ldk(f.root_player + tree, l2keep);
    const int root_mc = az_ldk(f.root_mc + tree, l2keep);
    AzBoard<NW> rootb;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        rootb.w[0][w] = az_ldk(f.root_words + (size_t)tree * 2 * NW + w, l2keep);
        rootb.w[1][w] = az_ldk(f.root_words + (size_t)tree * 2 * NW + NW + w, l2keep);
    }
    // the root's record is tracked in registers for the whole launch: every backup and every virtual
    // loss passes through it, and a reload after the backup stores would cost a round trip per descent
    AzMeta root_m;
    int root_nv, root_best;
    {
        AzHot hr;
        ulonglong2 mwr;
        az_load_node(f.nodes + (size_t)tree * f.cap, l2keep, hr, mwr);
        root_m = *reinterpret_cast<const AzMeta*>(&mwr.x);
        root_nv = hr.N + (int)hr.V;
        root_best = (int)hr.V;                                // vl == 1: V holds the cached best child (az_backup_best)
    }
    if (!is_active || is_done) return;
    AZ_AT(0);                                                 // tree scalars + root words
    AZ_AT_COUNT(15, 1);

    const AzGame& g = f.g;
    unsigned char* sm = smem_raw + (size_t)grp * smem_per_group;
    uint32_t* s_path = reinterpret_cast<uint32_t*>(sm);
    const int moves_pad = (g.max_children + 15) & ~15;
    const int lm_bytes = moves_pad + 2 * g.set_slots;           // legal-move scratch of one pending leaf
    uint8_t* s_lm = sm + 4 * f.max_depth;
    AzBestRec* s_rec = reinterpret_cast<AzBestRec*>(sm + smem_per_group - AZ_BEST_SMEM);   // vl == 1 only (az_forest_create)

    AzNode* const nodes = f.nodes + (size_t)tree * f.cap;
    const AzHotView hot{nodes};
    const AzMetaView meta{nodes};
    const AzMeta64View meta64{nodes};
    const bool use_vl = f.vl > 1;
    // one simulation in flight per tree and PUCT selection: descents follow cached best children (az_backup_best)
    const bool best_mode = (G == 32) && f.vl == 1 && !f.ucb1 && f.max_depth <= 64 && !f.no_best;
    const int A = g.action_dim;
    const int LP = f.lp;

    unsigned st_sims = 0, st_term = 0, st_exp = 0, st_child = 0, st_sel = 0, st_reads = 0, st_coll = 0, st_hit = 0;
    // An answer lives in the evaluation batch (slot < max_leaves) or in the evaluation cache (slot - max_leaves = entry).
    auto priors_of = [&](int slot) -> const float* {
        return slot < f.max_leaves ? f.priors + (size_t)slot * A : f.cache.priors + (size_t)(slot - f.max_leaves) * A;
    };
    auto value_of = [&](int slot) -> const float* {
        return slot < f.max_leaves ? f.values + slot : f.cache.values + (slot - f.max_leaves);
    };
    int npend = npend0;
    int np = 0;

  for (int round = 0;; ++round) {     // vl == 1 with the cache: a leaf answered from the cache is expanded in the same launch
    // ---------------- phase A: expand + backup the evaluated leaves -----------------------
    // Priors (lane a keeps prior a and a + 32) and the stored path (lane i keeps entries i and i + 32)
    // of leaf j + 1 are fetched while leaf j is processed: one exposed round trip per leaf (the
    // read-modify-write of the path nodes) instead of three.
    const bool pre = (G == 32) && A <= 64 && f.max_depth <= 64;
    for (int j0 = 0; j0 < npend; j0 += LP) {
        const int nj = min(LP, npend - j0);
        int my_nc = 0, my_slot = 0, my_len = 0, my_action = 0;
        uint32_t my_leaf = 0;
        float my_val = 0.f;
        AzBoard<NW> pb_;
        if (lane < nj) {                                      // lane j: scalars of leaf j0 + j
            const size_t pj = (size_t)tree * f.vl + j0 + lane;
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                pb_.w[0][w] = az_ldk(f.pend_words + pj * 2 * NW + w, l2keep);
                pb_.w[1][w] = az_ldk(f.pend_words + pj * 2 * NW + NW + w, l2keep);
            }
            my_slot = az_ldk(f.leaf_slot + pj, l2keep);
            my_len = az_ldk(f.path_len + pj, l2keep);
            my_action = az_ldk(f.pend_action + pj, l2keep);
            my_leaf = az_ldk(f.pend_leaf + pj, l2keep);
            my_val = az_ldk(value_of(my_slot), l2keep);
        }
        float pa = 0.f, pb = 0.f;
        uint32_t q0 = 0, q1 = 0;
        auto fetch = [&](int j, float& xa, float& xb, uint32_t& y0, uint32_t& y1) {
            const int slotx = __shfl_sync(gmask, my_slot, j, G);
            const float* prx = priors_of(slotx);
            const uint32_t* gp = f.path + ((size_t)tree * f.vl + j0 + j) * f.max_depth;
            xa = lane < A ? az_ldk(prx + lane, l2keep) : 0.f;
            xb = lane + 32 < A ? az_ldk(prx + lane + 32, l2keep) : 0.f;
            y0 = lane < f.max_depth ? az_ldk(gp + lane, l2keep) : 0u;
            y1 = lane + 32 < f.max_depth ? az_ldk(gp + lane + 32, l2keep) : 0u;
        };
        if (pre) fetch(0, pa, pb, q0, q1);
        if (lane < nj) {                                      // legal moves of leaf j0 + j (Gomoku: sequential set emulation)
            uint8_t* lm = s_lm + (size_t)lane * lm_bytes;
#ifndef AZ_LEGAL_INLINE
            my_nc = az_legal_moves_call<NW>(g, pb_, f.cell_hash, lm, moves_pad);
#else
            my_nc = az_legal_moves<NW>(g, pb_, f.cell_hash, lm + moves_pad, lm + moves_pad + g.set_slots, lm);
#endif
        }
        __syncwarp(gmask);                                    // move lists visible to the group
        AZ_AT(1);                                             // pending scalars + legal moves
        for (int j = 0; j < nj; ++j) {                        // in selection order: node numbering and f64 sums depend on it
            const size_t pj = (size_t)tree * f.vl + j0 + j;
            float na = 0.f, nb = 0.f;
            uint32_t r0 = 0, r1 = 0;
            if (pre && j + 1 < nj) fetch(j + 1, na, nb, r0, r1);
            const int nc = __shfl_sync(gmask, my_nc, j, G);
            const int slot = __shfl_sync(gmask, my_slot, j, G);
            const int len = __shfl_sync(gmask, my_len, j, G);
            const int action = __shfl_sync(gmask, my_action, j, G);
            const uint32_t leaf = __shfl_sync(gmask, my_leaf, j, G);
            const float val = __shfl_sync(gmask, my_val, j, G);
            const uint8_t* s_moves = s_lm + (size_t)j * lm_bytes;
            const uint32_t* gpath = f.path + pj * f.max_depth;
            const uint32_t base = (uint32_t)n_nodes;
            n_nodes += nc;
            const float* pri = priors_of(slot);
            const bool noisy_root = (leaf == 0) && f.use_noise;
            int leaf_best = 0;                                    // first maximum of the fresh children's scores = of their priors
            float lb32 = -1.0f / 0.0f;                            // (n = 0, parent visit 1: P * sqrt(1) / 1, utils.py:34)
            double lb64 = -1.0 / 0.0;
            for (int c0 = 0; c0 < nc; c0 += G) {
                const int c = c0 + lane;
                const bool valid = c < nc;
                const int cell = valid ? s_moves[c] : 0;
                const int ai = az_action_index(g, cell);
                float p;                                         // node.py:57 gather, unmasked softmax
                if (pre) {
                    const float xa = __shfl_sync(gmask, pa, ai & 31, G);
                    const float xb = A > 32 ? __shfl_sync(gmask, pb, ai & 31, G) : 0.f;
                    p = ai < 32 ? xa : xb;
                } else {
                    p = valid ? pri[ai] : 0.f;
                }
                double p64 = 0.0;
                if (valid && noisy_root)                         // utils.py:24: f32*0.75 + f64 noise*0.25
                    p64 = __dadd_rn((double)__fmul_rn(0.75f, p), __dmul_rn(0.25, f.root_noise[(size_t)tree * A + ai]));
                if (best_mode) {                                 // all lanes take part in the reductions
                    if (noisy_root) {
                        double sc = -1.0 / 0.0;
                        if (valid && p64 > sc) sc = p64;
                        const int bi = az_group_argmax(gmask, sc, valid ? c : 0x7fffffff);
                        const double top = __shfl_sync(gmask, sc, bi < 0x7fffffff ? (bi - c0) : 0, G);
                        if (bi < 0x7fffffff && top > lb64) { lb64 = top; leaf_best = bi; }
                    } else {
                        float sc = -1.0f / 0.0f;
                        if (valid && p > sc) sc = p;
                        const int bi = az_group_argmax(gmask, sc, valid ? c : 0x7fffffff);
                        const float top = __shfl_sync(gmask, sc, bi < 0x7fffffff ? (bi - c0) : 0, G);
                        if (bi < 0x7fffffff && top > lb32) { lb32 = top; leaf_best = bi; }
                    }
                }
                if (!valid) continue;
                if (noisy_root) {
                    f.root_prior64[(size_t)tree * g.max_children + c] = p64;
                    p = __double2float_rn(p64);                  // kept for reporting only
                }
                AzNode nn;
                nn.hot.N = 0;
                nn.hot.Q = 0.f;
                nn.hot.P = p;
                nn.hot.V = 0;
                nn.meta.child_base = 0;
                nn.meta.action = (uint16_t)cell;
                nn.meta.nchild = 0;
                nn.meta.status = AZ_ST_LEAF;
                nn.W = 0.0;
                uint4* dst = reinterpret_cast<uint4*>(&nodes[base + c]);   // one sector, two 16-byte stores
                az_store16_keep(dst, reinterpret_cast<const uint4*>(&nn)[0], l2keep);
                az_store16_keep(dst + 1, reinterpret_cast<const uint4*>(&nn)[1], l2keep);
            }
            {
                AzMeta m;
                m.child_base = base;
                m.action = (uint16_t)action;
                m.nchild = (uint8_t)nc;
                m.status = AZ_ST_EXPANDED;
                if (lane == 0) meta[leaf] = m;
                if (leaf == 0) root_m = m;
            }
            AZ_AT(2);                                             // child block written
            const double result = -(double)val;                   // mcts.py:56
            if (best_mode) {
                root_best = az_backup_best<G>(nodes, q0, q1, len, result, leaf_best,
                                              f.use_noise ? f.root_prior64 + (size_t)tree * g.max_children : nullptr, lane, gmask, l2keep, &st_reads, s_rec);
            } else {
                for (int i = lane; i < len; i += G) {             // node.py:69-72
                    const uint32_t nd = pre ? (i < 32 ? q0 : q1) : gpath[i];
                    az_backup_node(nodes, nd, ((len - 1 - i) & 1) ? -result : result, use_vl, l2keep);
                }
            }
            root_nv += use_vl ? 0 : 1;                            // the root: N + 1, V - 1 under virtual loss
            __syncwarp(gmask);
            AZ_AT(3);                                             // backup
            ++st_exp;
            st_child += nc;
            ++st_sims;
            ++sims;
            pa = na; pb = nb; q0 = r0; q1 = r1;
        }
    }

    // ---------------- phase B: descend until `vl` leaves are pending -----------------------
    int started = sims;
    np = 0;
    bool ready_hit = false;                                   // the last pending leaf found its answer in the cache
    // A tree whose best line ends in a terminal node burns simulation after simulation without ever
    // needing the evaluator; left unbounded, one such tree (up to `iters` descents) sets the duration
    // of the whole launch (measured: one tree 15x the 99th percentile).  The step machine therefore
    // closes a step after `vl` terminal descents as well.  At vl == 1 nothing is pending when that
    // happens, the next launch resumes the very same sequence and the reference's order is kept;
    // at vl > 1 it is part of the build-defined batching rule (oracle: az_oracle.c azo_search).
    const unsigned kMaxTerminalPerLaunch = (unsigned)max(f.vl, f.max_rounds);   // max_rounds > 1 only at vl == 1 with the cache
    AzBoard<NW> kept = rootb;
    int kept_player = 0, flushed = 0;
    uint32_t kept_entry = 0xffffffffu;                        // cache entry of this lane's pending leaf
    int kept_cached = 0;                                      // ... which already holds (or will hold) the answer: no batch row
    auto hand_off = [&](int upto) {                           // evaluation slots of pending leaves [flushed, upto)
        const int cnt = upto - flushed;
        if (cnt <= 0) return;
        const unsigned need = __ballot_sync(gmask, lane < cnt && !kept_cached) & gmask;
        const int rows = __popc(need);
        int base = 0;
        if (lane == 0 && rows) base = atomicAdd(f.n_leaves, rows);
        base = __shfl_sync(gmask, base, 0, G);
        if (lane < cnt) {
            const size_t pk = (size_t)tree * f.vl + flushed + lane;
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                f.pend_words[pk * 2 * NW + w] = kept.w[0][w];
                f.pend_words[pk * 2 * NW + NW + w] = kept.w[1][w];
            }
            if (kept_cached) {
                f.leaf_slot[pk] = f.max_leaves + (int)kept_entry;
            } else {
                const unsigned below = need & ((1u << (threadIdx.x & 31)) - 1u);
                const int slot = base + __popc(below);
                f.leaf_slot[pk] = slot;
                f.leaf_player[slot] = (uint8_t)kept_player;
                f.leaf_tree[slot] = tree;
                if (f.cache.cap) f.cache.leaf_entry[slot] = kept_entry;
#pragma unroll
                for (int w = 0; w < NW; ++w) {
                    f.leaf_words[(size_t)slot * 2 * NW + w] = kept_player ? kept.w[1][w] : kept.w[0][w];        // canonical: mine
                    f.leaf_words[(size_t)slot * 2 * NW + NW + w] = kept_player ? kept.w[0][w] : kept.w[1][w];   // theirs
                }
            }
        }
    };
    while (started < f.iters && np < f.vl && st_term < kMaxTerminalPerLaunch) {
        AzBoard<NW> b = rootb;
        int player = root_player, mc = root_mc, len = 1;
        uint32_t node = 0;
        if (lane == 0) s_path[0] = 0;
        AzMeta m = root_m;
        int pn = root_nv;
        AZ_AT(4);                                             // root record (registers)
        // One memory round trip per level: every lane fetches hot AND meta of its children, the
        // winner's meta / visit count arrive by shuffle (mcts.py:20-23, node.py:42-47).
        int walk = root_best;
        while (best_mode && m.status == AZ_ST_EXPANDED) {     // vl == 1: follow the cached best children, one uniform load per level
            node = m.child_base + (uint32_t)walk;
            AzHot h;
            ulonglong2 mw;
            az_load_node(&nodes[node], l2keep, h, mw);
            m = *reinterpret_cast<const AzMeta*>(&mw.x);
            walk = (int)h.V;
            if (lane == 0) s_path[len] = node;
            ++len;
            ++st_sel;
            b.put(player, m.action);                          // Game.make_move (mcts.py:23)
            player ^= 1;
            ++mc;
            AZ_AT(5);
            AZ_AT_COUNT(14, 1);
        }
        while (m.status == AZ_ST_EXPANDED) {
            const int nc = m.nchild;
            const uint32_t cb = m.child_base;
            int bi = 0x7fffffff, bn = 0;
            unsigned long long bm = 0ull;
            if (f.ucb1 || (node == 0 && f.use_noise)) {
                const double s = f.ucb1 ? __dmul_rn(2.0, f.log_tab[pn]) : __dsqrt_rn((double)pn);
                double best = -1.0 / 0.0;
                for (int c = lane; c < nc; c += G) {
                    AzHot h;
                    ulonglong2 mw;                                   // meta | W: same sector
                    az_load_node(&nodes[cb + c], l2keep, h, mw);
                    const unsigned long long mm = mw.x;
                    const int n = h.N + (int)h.V;
                    const double W = __longlong_as_double((long long)mw.y) - (double)h.V;
                    const double sc = f.ucb1 ? az_ucb1(W, n, s) : az_puct64(W, f.root_prior64[(size_t)tree * g.max_children + c], n, s);
                    if (sc > best) { best = sc; bi = c; bm = mm; bn = n; }
                }
                bi = az_group_argmax(gmask, best, bi);
            } else {
                const float s = __fsqrt_rn((float)pn);
                float best = -1.0f / 0.0f;
                for (int c = lane; c < nc; c += G) {
                    AzHot h;
                    ulonglong2 mw;                                   // meta | W: same sector
                    az_load_node(&nodes[cb + c], l2keep, h, mw);
                    const unsigned long long mm = mw.x;
                    const double Wc = __longlong_as_double((long long)mw.y);
                    int n = h.N;
                    float q = h.Q;
                    if (use_vl && h.V) {                              // virtual loss: n + v visits, v losses
                        n += (int)h.V;
                        q = az_q32(Wc - (double)h.V, n);
                    }
                    const float sc = az_puct32(q, h.P, n, s);
                    if (sc > best) { best = sc; bi = c; bm = mm; bn = n; }
                }
                bi = az_group_argmax(gmask, best, bi);
            }
            st_reads += nc;
            ++st_sel;
            if (bi == 0x7fffffff) {                               // all-NaN guard: max() keeps the first
                bi = 0;
                bm = meta64[cb];
                const AzHot h0 = hot[cb];
                bn = h0.N + (int)h0.V;
            } else {                                              // the lane that scored child bi holds its records
                bm = __shfl_sync(gmask, bm, bi % G, G);
                bn = __shfl_sync(gmask, bn, bi % G, G);
            }
            node = cb + bi;
            pn = bn;
            if (lane == 0) s_path[len] = node;
            ++len;
            m = *reinterpret_cast<const AzMeta*>(&bm);
            b.put(player, m.action);                          // Game.make_move (mcts.py:23)
            player ^= 1;
            ++mc;
            AZ_AT(5);                                         // one select level
            AZ_AT_COUNT(14, 1);
        }
        if (m.status == AZ_ST_PENDING) {                      // leaf already waiting: close the batch
            ++st_coll;
            break;
        }
        __syncwarp(gmask);                                    // s_path visible to all lanes
        int status = m.status;
        if (node != 0 && status == AZ_ST_LEAF) {              // mcts.py:25-32, cached in the node
            if (az_is_win<NW>(g, b, player ^ 1, m.action)) status = AZ_ST_WIN;
            else if (mc == g.state_dim) status = AZ_ST_DRAW;
            if (status != AZ_ST_LEAF && lane == 0) {
                m.status = (uint8_t)status;
                meta[node] = m;
            }
        }
        if (status == AZ_ST_WIN || status == AZ_ST_DRAW) {
            const double v = status == AZ_ST_WIN ? 1.0 : 0.0;
            if (best_mode) {
                root_best = az_backup_best<G>(nodes, lane < len ? s_path[lane] : 0u, lane + 32 < len ? s_path[lane + 32] : 0u, len, v, 0,
                                              f.use_noise ? f.root_prior64 + (size_t)tree * g.max_children : nullptr, lane, gmask, l2keep, &st_reads, s_rec);
            } else {
                for (int i = lane; i < len; i += G) az_backup_node(nodes, s_path[i], ((len - 1 - i) & 1) ? -v : v, false, l2keep);
            }
            __syncwarp(gmask);
            root_nv += 1;
            ++st_term;
            ++st_sims;
            ++started;
            ++sims;
            AZ_AT(7);                                         // terminal backup
            continue;
        }
        // leaf needs the evaluator: lane (np - flushed) keeps its board until the tree's slots are
        // reserved with ONE atomic for all leaves of this launch (hand_off below)
        const size_t pj = (size_t)tree * f.vl + np;
        uint32_t* gpath = f.path + pj * f.max_depth;
        for (int i = lane; i < len; i += G) {
            const uint32_t nd = s_path[i];
            gpath[i] = nd;
            if (use_vl) atomicAdd(&nodes[nd].hot.V, 1u);      // fire-and-forget: nobody waits for the old count
        }
        if (use_vl) root_nv += 1;
        uint32_t entry = 0xffffffffu;
        int found = 0;
        if (NW == 1 && f.cache.cap && !f.ucb1) {              // MCTS.cache lookup (ai/mcts.py:37-44) on the canonical position
            const uint64_t mine = player ? b.w[1][0] : b.w[0][0];
            uint64_t theirs = player ? b.w[0][0] : b.w[1][0];
            if (g.planes == 3) theirs |= (uint64_t)player << 63;
            found = az_cache_probe<G>(f.cache, mine, theirs, lane, gmask, &entry);
            if (found) ++st_hit;
            ready_hit = found == 2;
        }
        AZ_AT(9);                                             // terminal test + path store + cache probe
        if (lane == np - flushed) {
            kept = b;
            kept_player = player;
            kept_entry = entry;
            kept_cached = found;
        }
        if (use_vl) m.status = AZ_ST_PENDING;
        if (node == 0) root_m = m;
        if (lane == 0) {
            f.path_len[pj] = (uint16_t)len;
            f.pend_action[pj] = m.action;
            f.pend_leaf[pj] = node;
            if (use_vl) meta[node] = m;
        }
        if (np + 1 - flushed == G) {
            hand_off(np + 1);
            flushed = np + 1;
        }
        __syncwarp(gmask);
        AZ_AT(6);                                             // terminal test + leaf hand-off
        ++np;
        ++started;
    }

    hand_off(np);
    // vl == 1: the leaf's answer is in the cache -> apply it now and run the next simulation in this launch (the order
    // of a tree's simulations is untouched, so the search is the reference's; only fewer launches are needed)
    if (!(f.vl == 1 && np == 1 && ready_hit && round + 1 < f.max_rounds)) break;
    __syncwarp(gmask);
    npend = 1;
  }
    if (lane == 0) {
        f.n_nodes[tree] = n_nodes;
        f.sims_done[tree] = sims;
        f.n_pending[tree] = (uint8_t)np;
        if (np == 0 && sims >= f.iters) {
            f.done_flag[tree] = 1;
            atomicAdd(f.n_done, 1);
        }
    }
    if (f.stats && lane < 8) {
        const unsigned vals[8] = {st_sims, st_term, st_exp, st_child, st_sel, st_reads, st_coll, st_hit};
        unsigned v = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) if (lane == k) v = vals[k];
        if (v) atomicAdd(&f.stats[(size_t)tree * AZ_STAT_COUNT + lane], (unsigned long long)v);   // no reply needed
    }
    AZ_AT(8);                                                 // tail writes
#ifdef AZ_ADV_TRACE
    if (lane == 0) {
        const unsigned life = (unsigned)(clock64() - at_t0);
        atomicAdd(&g_adv_hist[min(life >> 15, 63u)], 1u);
        atomicMax(&g_adv_hist[64], life);
    }
#endif
}

}  // namespace

// ======================================================================================
// Auxiliary kernels
// ======================================================================================
namespace {

__global__ void az_begin_move_kernel(AzForestDev f) {
    const int tree = blockIdx.x * blockDim.x + threadIdx.x;
    if (tree >= f.n_trees) return;
    if (tree == 0 && f.cache.cap) *f.cache.epoch += 1u;       // a new search: entries age by one
    AzHot h;
    h.N = 0;
    h.Q = 0.f;
    h.P = 0.f;
    h.V = 0;
    f.nodes[(size_t)tree * f.cap].hot = h;
    f.nodes[(size_t)tree * f.cap].W = 0.0;
    AzMeta m;
    m.child_base = 0;
    m.action = 0;
    m.nchild = 0;
    m.status = AZ_ST_LEAF;
    f.nodes[(size_t)tree * f.cap].meta = m;
    f.n_nodes[tree] = 1;
    f.sims_done[tree] = 0;
    f.n_pending[tree] = 0;
    const bool live = f.active[tree] != 0;
    f.done_flag[tree] = live ? 0 : 1;
    if (!live) atomicAdd(f.n_done, 1);                       // counters are zeroed by the host before
}

// Runs in front of every advance launch when the cache is on (instead of the memset of the batch counter): the rows the
// evaluator answered since the last launch are copied into the entries their leaves claimed (the store of ai/mcts.py:51)
// and the entries become READY; the last block to finish zeroes the batch counter for the launch that follows.
__global__ void az_cache_publish_kernel(AzForestDev f, unsigned int* ticket) {
    const int n = min(*f.n_leaves, f.max_leaves);
    const int A = f.g.action_dim;
    const uint32_t epoch = *f.cache.epoch;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)n * (A + 1); t += (long long)gridDim.x * blockDim.x) {
        const int row = (int)(t / (A + 1));
        const int a = (int)(t - (long long)row * (A + 1));
        const uint32_t e = f.cache.leaf_entry[row];
        if (e == 0xffffffffu) continue;
        if (a < A) {
            f.cache.priors[(size_t)e * A + a] = f.priors[(size_t)row * A + a];
        } else {                                              // nobody reads the cache while this kernel runs (stream order)
            f.cache.values[e] = f.values[row];
            f.cache.state[e] = (AZ_CK_READY << 30) | (epoch & 0xffu);
        }
    }
    __shared__ unsigned int last;                             // every block has read the counter: the last one out resets it
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (last && threadIdx.x == 0) {
        *f.n_leaves = 0;
        *ticket = 0u;
    }
}

__global__ void az_cache_tick_kernel(AzForestDev f) { *f.cache.epoch += 1u; }

// splitmix64 finaliser; the build-defined hash evaluator (oracle/az_oracle.py: hash_eval_*).
__device__ __forceinline__ uint64_t az_sm64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// One thread per (leaf, action): priors exact in fp32 (integer * power of two), sum <= 1.
__global__ void az_eval_hash_kernel(AzForestDev f, uint64_t seed, int qbits, float scale) {
    const int n = *f.n_leaves;
    const int A = f.g.action_dim;
    const int nw = f.g.nw;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)n * (A + 1);
         t += (long long)gridDim.x * blockDim.x) {
        const int slot = (int)(t / (A + 1));
        const int a = (int)(t - (long long)slot * (A + 1));
        const uint64_t* w = f.leaf_words + (size_t)slot * 2 * nw;
        uint64_t h = az_sm64(seed);
        for (int i = 0; i < nw; ++i) h = az_sm64(h ^ w[i]);
        h = az_sm64(h ^ 0xA5A5A5A5A5A5A5A5ull);
        for (int i = 0; i < nw; ++i) h = az_sm64(h ^ w[nw + i]);
        if (f.g.planes == 3) h = az_sm64(h ^ (uint64_t)f.leaf_player[slot]);
        if (a < A) {
            const uint64_t x = az_sm64(h + (uint64_t)a + 1);
            const uint32_t q = (uint32_t)((x >> 40) & ((1ull << qbits) - 1)) + 1;
            f.priors[(size_t)slot * A + a] = __fmul_rn((float)q, scale);
        } else {
            const uint64_t x = az_sm64(h + 0xFFFFull);
            const int q = (int)((x >> 40) & 0xFFFF) - 32768;
            f.values[slot] = __fmul_rn((float)q, 1.0f / 32768.0f);
        }
    }
}

// MCTS.simulate (ai/mcts.py:62-79): uniformly random legal moves from the leaf to the end of the game, one thread
// per pending leaf.  The leaf is canonical (side 0 = side to move), so the value written is +1 if the side to move
// wins, -1 if it loses, 0 for a draw (the reference's `result` with the sign of mcts.py:56).  Legal SETS only --
// the reference's move order does not matter under a uniform draw: TicTacToe empty cells, Connect4 the lowest
// empty cell of every open column, Gomoku empty 8-neighbours of stones (centre on an empty board).
template <int NW>
__device__ int az_rollout(const AzGame& g, AzBoard<NW> b, uint64_t rng) {
    int mc = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) mc += __popcll(b.w[0][i]) + __popcll(b.w[1][i]);
    int player = 0;
    if (NW == 1) {
        // One-word boards: the legal set by shift and mask (az_rules.cuh; same sets, same k-th-lowest-cell draw and the same
        // random stream as the cell walk below, hence the same playouts; tests/native/legal_set_test.cu compares the two forms)
        const AzRolloutMasks m = az_rollout_masks(g);
        while (mc < g.state_dim) {
            uint64_t legal = az_rollout_legal_bits(g, m, b.w[0][0] | b.w[1][0], mc);
            const int n = __popcll(legal);
            if (n == 0) return 0;
            rng = az_sm64(rng);
            const int k = (int)((rng >> 11) % (uint64_t)n);
            for (int j = 0; j < k; ++j) legal &= legal - 1;
            const int cell = __ffsll((long long)legal) - 1;
            b.put(player, cell);
            ++mc;
            if (az_is_win(g, b, player, cell)) return player == 0 ? 1 : -1;
            player ^= 1;
        }
        return 0;
    }
    while (mc < g.state_dim) {
        uint64_t legal[NW];
        const int n = az_rollout_legal_walk<NW>(g, b, mc, legal);
        if (n == 0) return 0;
        rng = az_sm64(rng);
        int k = (int)((rng >> 11) % (uint64_t)n);
        int cell = 0;
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            uint64_t w = legal[i];
            const int pc = __popcll(w);
            if (k >= 0 && k < pc) {
                for (int j = 0; j < k; ++j) w &= w - 1;
                cell = i * 64 + __ffsll((long long)w) - 1;
                k = -1;
            } else if (k >= 0) {
                k -= pc;
            }
        }
        b.put(player, cell);
        ++mc;
        if (az_is_win(g, b, player, cell)) return player == 0 ? 1 : -1;
        player ^= 1;
    }
    return 0;
}

__global__ void az_eval_rollout_kernel(AzForestDev f, uint64_t seed, uint64_t counter) {
    const int n = *f.n_leaves;
    const int A = f.g.action_dim;
    const int nw = f.g.nw;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n; slot += gridDim.x * blockDim.x) {
        const uint64_t* w = f.leaf_words + (size_t)slot * 2 * nw;
        const uint64_t rng = az_sm64(az_sm64(seed ^ (counter * 0x9E3779B97F4A7C15ull)) + (uint64_t)f.leaf_tree[slot] * 0xD1B54A32D192ED03ull + (uint64_t)slot);
        int r;
        if (nw == 1) {
            AzBoard<1> b;
            b.w[0][0] = w[0];
            b.w[1][0] = w[1];
            r = az_rollout<1>(f.g, b, rng);
        } else {
            AzBoard<AZ_MAX_WORDS> b;
            for (int i = 0; i < AZ_MAX_WORDS; ++i) {
                b.w[0][i] = i < nw ? w[i] : 0ull;
                b.w[1][i] = i < nw ? w[nw + i] : 0ull;
            }
            r = az_rollout<AZ_MAX_WORDS>(f.g, b, rng);
        }
        f.values[slot] = (float)r;
        for (int a = 0; a < A; ++a) f.priors[(size_t)slot * A + a] = 0.f;     // node.py:58: children without priors
    }
}

// f32 planes [n][C][H][W] (the reference's board format) -> side bitboards.
__global__ void az_planes_to_words_kernel(const float* __restrict__ planes, int n, AzGame g, uint64_t* words) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * 2 * g.nw) return;
    const int tree = t / (2 * g.nw), r = t % (2 * g.nw), side = r / g.nw, wi = r % g.nw;
    const float* p = planes + ((size_t)tree * g.planes + side) * g.state_dim;
    uint64_t v = 0;
    for (int i = 0; i < 64; ++i) {
        const int cell = wi * 64 + i;
        if (cell < g.state_dim && p[cell] == 1.0f) v |= 1ull << i;
    }
    words[t] = v;
}

// Evaluation batch -> f32 canonical planes [B][C][H][W] for an external (PyTorch) evaluator:
// plane 0 = side to move, plane 1 = opponent (gomoku.py:34-40), plane 2 = side-to-move id.
__global__ void az_leaves_to_planes_kernel(AzForestDev f, float* out, int max_b) {
    const int n = min(*f.n_leaves, max_b);
    const int cells = f.g.state_dim, C = f.g.planes, nw = f.g.nw;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < (long long)n * C * cells;
         t += (long long)gridDim.x * blockDim.x) {
        const int slot = (int)(t / (C * cells));
        const int r = (int)(t - (long long)slot * C * cells);
        const int pl = r / cells, cell = r - pl * cells;
        float v;
        if (pl < 2) v = (float)((f.leaf_words[(size_t)slot * 2 * nw + pl * nw + (cell >> 6)] >> (cell & 63)) & 1ull);
        else v = (float)f.leaf_player[slot];
        out[t] = v;
    }
}

// logits [B][A], values [B] from an external evaluator -> priors exactly as ai/mcts.py:49:
// exp(l)/sum(exp(l)) in fp32 (max-subtracted for range safety; same value up to rounding).
__global__ void az_logits_to_priors_kernel(AzForestDev f, const float* __restrict__ logits,
                                           const float* __restrict__ values, int max_b) {
    const int n = min(*f.n_leaves, max_b);
    const int A = f.g.action_dim;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n) return;
    const float* l = logits + (size_t)warp * A;
    float mx = -1.0f / 0.0f;
    for (int a = lane; a < A; a += 32) mx = fmaxf(mx, l[a]);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float s = 0.f;
    for (int a = lane; a < A; a += 32) s += expf(l[a] - mx);
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    for (int a = lane; a < A; a += 32) f.priors[(size_t)warp * A + a] = expf(l[a] - mx) / s;
    if (lane == 0) f.values[warp] = values[warp];
}

// Visit counts of the root children scattered to action order (utils.py:46-55) + root Q.
__global__ void az_root_visits_kernel(AzForestDev f, int* visits, double* q) {
    const int tree = blockIdx.x;
    const int A = f.g.action_dim;
    for (int a = threadIdx.x; a < A; a += blockDim.x) visits[(size_t)tree * A + a] = 0;
    __syncthreads();
    const AzNode* nodes = f.nodes + (size_t)tree * f.cap;
    const AzMeta m = nodes[0].meta;
    if (m.status == AZ_ST_EXPANDED)
        for (int c = threadIdx.x; c < m.nchild; c += blockDim.x)
            visits[(size_t)tree * A + az_action_index(f.g, nodes[m.child_base + c].meta.action)] = nodes[m.child_base + c].hot.N;
    if (threadIdx.x == 0 && q) q[tree] = nodes[0].hot.N ? nodes[0].W / (double)nodes[0].hot.N : 0.0;   // gomoku.py:141
}

// Node.max_visit_child (ai/node.py:76-81) for every root: the FIRST most-visited child in child order, as its cell.
__global__ void az_root_best_kernel(AzForestDev f, int* cells) {
    const int tree = blockIdx.x * blockDim.x + threadIdx.x;
    if (tree >= f.n_trees) return;
    const AzNode* nodes = f.nodes + (size_t)tree * f.cap;
    const AzMeta m = nodes[0].meta;
    int best = -1, cell = -1;
    if (m.status == AZ_ST_EXPANDED)
        for (int c = 0; c < m.nchild; ++c) {
            const int n = nodes[m.child_base + c].hot.N;
            if (n > best) { best = n; cell = nodes[m.child_base + c].meta.action; }
        }
    cells[tree] = cell;
}

}  // namespace

// ======================================================================================
// Host side: handle + C ABI (include/az_b200.h)
// ======================================================================================
#include <string>
#include <vector>
#include "az_host.h"
#include "../../include/az_b200.h"

thread_local std::string g_az_error;
extern "C" int az_forest_cache_clear(az_forest* f, void* stream);

extern "C" const char* az_last_error(void) { return g_az_error.c_str(); }
extern "C" const char* az_version(void) { return "az_b200 0.1 (sm_100a)"; }

int az_fail(int code, const std::string& msg) {
    g_az_error = msg;
    return code;
}

int az_make_game(int game, int rows, int cols, AzGame* out) {
    AzGame g{};
    g.gid = game;
    g.rows = rows;
    g.cols = cols;
    if (game == AZ_GAME_TICTACTOE) {
        if (rows != 3 || cols != 3) return az_fail(AZ_EINVAL, "tictactoe is 3x3");
        g.k = 3; g.action_dim = 9; g.planes = 3; g.max_children = 9;
    } else if (game == AZ_GAME_CONNECT4) {
        if (rows != 6 || cols != 7) return az_fail(AZ_EINVAL, "connect4 is 6x7");
        g.k = 4; g.action_dim = 7; g.planes = 3; g.max_children = 7;
    } else if (game == AZ_GAME_GOMOKU) {
        if (rows != cols || rows < 5 || rows > 15) return az_fail(AZ_EINVAL, "gomoku is n x n, 5 <= n <= 15");
        g.k = 5; g.action_dim = rows * cols; g.planes = 2; g.max_children = rows * cols;
    } else {
        return az_fail(AZ_EINVAL, "unknown game id");
    }
    g.state_dim = rows * cols;
    g.nw = g.state_dim <= 64 ? 1 : AZ_MAX_WORDS;   // kernels are instantiated for 1 and 4 words
    g.set_slots = 0;
    if (game == AZ_GAME_GOMOKU) {                 // largest table a CPython set of <= cells items reaches
        int size = 8;
        for (int fill = 1; fill <= g.state_dim; ++fill)
            if (fill * 5 >= (size - 1) * 3) {
                size = 8;
                while (size <= fill * 4) size <<= 1;
            }
        g.set_slots = size;
    }
    for (int d = 0; d < 4; ++d) g.win_start[d] = 0;
    if (g.nw == 1) {
        const int dcs[4] = {1, 0, 1, -1};
        for (int d = 0; d < 4; ++d)
            for (int r = 0; r < rows; ++r)
                for (int c = 0; c < cols; ++c) {
                    const int last = c + dcs[d] * (g.k - 1);                 // column of the run's last stone
                    if (last >= 0 && last < cols) g.win_start[d] |= 1ull << (r * cols + c);
                }
    }
    *out = g;
    return AZ_OK;
}

// CPython 3.8+ tuple hash of (r, c), Objects/tupleobject.c (xxHash-derived).
uint64_t az_tuple2_hash(uint64_t r, uint64_t c) {
    const uint64_t P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P5 = 2870177450012600261ull;
    uint64_t acc = P5;
    const uint64_t lanes[2] = {r, c};
    for (int i = 0; i < 2; ++i) {
        acc += lanes[i] * P2;
        acc = (acc << 31) | (acc >> 33);
        acc *= P1;
    }
    acc += 2ull ^ (P5 ^ 3527539ull);
    if (acc == ~0ull) acc = 1546275796ull;
    return acc;
}

struct az_forest {
    AzForestDev d{};
    int max_iters = 0;
    int G = 32;
    int block = 128;
    int smem_per_group = 0;
    std::vector<void*> allocs;
    double* noise_dev = nullptr;
    int* vis_dev = nullptr;       // root_visits scratch [n_trees][A] + root Q [n_trees]
    double* q_dev = nullptr;
    int* h_counts = nullptr;      // pinned: n_leaves, n_done
    void* h_stage = nullptr;      // pinned staging for roots
    size_t h_stage_bytes = 0;
    float* planes_stage_dev = nullptr;
    int32_t* i32_stage_dev = nullptr;
    unsigned long long uid = 0;          // never reused (graphs are keyed by it, az_net.cu)
    unsigned int* cache_ticket = nullptr;
    std::vector<void*> cache_allocs;
    unsigned long long eval_tag = 0;     // identity of the evaluator whose answers the cache holds
    bool eval_tag_set = false;
};
static unsigned long long g_forest_uid = 0;

namespace {

template <typename T>
int az_dev_alloc(az_forest* f, T** p, size_t n) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T) + 16);
    if (e != cudaSuccess) return az_fail(AZ_ENOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e));
    f->allocs.push_back(q);
    *p = reinterpret_cast<T*>(q);
    return AZ_OK;
}

__global__ void az_roots_from_i32_kernel(AzForestDev f, const int32_t* players, const int32_t* mcs) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= f.n_trees) return;
    f.root_player[t] = (uint8_t)players[t];
    f.root_mc[t] = (uint16_t)mcs[t];
}

template <int NW>
__global__ void az_rules_legal_kernel(AzGame g, int n, const uint64_t* words, const uint64_t* cell_hash,
                                      int32_t* counts, uint8_t* cells, uint8_t* scratch) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    AzBoard<NW> b;
    for (int w = 0; w < NW; ++w) {
        b.w[0][w] = words[(size_t)t * 2 * NW + w];
        b.w[1][w] = words[(size_t)t * 2 * NW + NW + w];
    }
    uint8_t* ta = scratch + (size_t)t * 2 * (g.set_slots + 1);
    counts[t] = az_legal_moves<NW>(g, b, cell_hash, ta, ta + g.set_slots, cells + (size_t)t * g.state_dim);
}

template <int NW>
__global__ void az_rules_winner_kernel(AzGame g, int n, const uint64_t* words, const int32_t* players,
                                       const int32_t* cells, int32_t* out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    AzBoard<NW> b;
    for (int w = 0; w < NW; ++w) {
        b.w[0][w] = words[(size_t)t * 2 * NW + w];
        b.w[1][w] = words[(size_t)t * 2 * NW + NW + w];
    }
    out[t] = az_is_win<NW>(g, b, players[t], cells[t]) ? players[t] : -1;
}

template <int NW, int G>
int az_launch_advance(az_forest* f, cudaStream_t s) {
    const int gpb = f->block / G;
    const int grid = (f->d.n_trees + gpb - 1) / gpb;
    const size_t smem = (size_t)gpb * f->smem_per_group;
    az_advance_kernel<NW, G><<<grid, f->block, smem, s>>>(f->d, f->smem_per_group);
    return AZ_OK;
}

}  // namespace

#ifdef AZ_ADV_TRACE
extern "C" __attribute__((visibility("default"))) int az_debug_adv_trace(unsigned long long* out_host, int reset) {
    cudaError_t e = cudaMemcpyFromSymbol(out_host, g_adv_trace, sizeof(g_adv_trace));
    if (reset) {
        void* p = nullptr;
        cudaGetSymbolAddress(&p, g_adv_trace);
        cudaMemset(p, 0, sizeof(g_adv_trace));
    }
    return (int)e;
}
extern "C" __attribute__((visibility("default"))) int az_debug_adv_hist(unsigned int* out_host) {
    cudaError_t e = cudaMemcpyFromSymbol(out_host, g_adv_hist, sizeof(g_adv_hist));
    void* p = nullptr;
    cudaGetSymbolAddress(&p, g_adv_hist);
    cudaMemset(p, 0, sizeof(g_adv_hist));
    return (int)e;
}
#endif

#define AZ_CUDA_OK(expr)                                                                       \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess) return az_fail(AZ_ECUDA, std::string(#expr ": ") + cudaGetErrorString(e__)); \
    } while (0)

extern "C" int az_forest_create(int game, int rows, int cols, int n_trees, int max_iters, int vl_batch,
                                az_forest** out) {
    if (!out || n_trees <= 0 || max_iters <= 0 || vl_batch <= 0 || vl_batch > 64)
        return az_fail(AZ_EINVAL, "az_forest_create: bad arguments");
    AzGame g;
    int rc = az_make_game(game, rows, cols, &g);
    if (rc) return rc;
    az_forest* f = new az_forest();
    AzForestDev& d = f->d;
    d.g = g;
    d.n_trees = n_trees;
    d.iters = max_iters;
    d.vl = vl_batch;
    d.cap = 1 + max_iters * g.max_children;
    d.max_depth = g.state_dim + 2;
    d.max_leaves = n_trees * vl_batch;
    d.use_noise = 0;
    d.ucb1 = 0;
    d.log_tab = nullptr;
    d.cache = AzCacheDev{};
    d.max_rounds = 1;
    d.no_best = getenv("AZ_NO_BEST") ? 1 : 0;
    f->uid = ++g_forest_uid;
    f->max_iters = max_iters;
    // One warp per tree: sub-warp groups (8 lanes for the <= 9-children games) share a warp between
    // four trees whose descents diverge and serialise; measured 4x slower on Connect4.  AZ_FOREST_GROUP=8
    // re-enables them for experiments.
    f->G = 32;
    if (const char* e = getenv("AZ_FOREST_GROUP")) if (atoi(e) == 8 && g.max_children <= 9) f->G = 8;
    // legal-move scratch for up to AZ_LP pending leaves expanded in parallel, bounded by ~24 KB per tree
    const size_t lm_bytes = ((g.max_children + 15) & ~15) + 2 * (size_t)g.set_slots;
    int lp = std::min(vl_batch, AZ_LP);
    while (lp > 1 && lp * lm_bytes > 24 * 1024) --lp;
    d.lp = lp;
    size_t per_group = 4 * (size_t)d.max_depth + lp * lm_bytes;
    f->smem_per_group = (int)((per_group + 15) & ~(size_t)15);
    if (vl_batch == 1) f->smem_per_group += AZ_BEST_SMEM;     // path-node records of az_backup_best
    const size_t nodes = (size_t)n_trees * d.cap;
    const int nw = g.nw;
#define AZ_ALLOC(ptr, count)                                  \
    if ((rc = az_dev_alloc(f, &(ptr), (count))) != AZ_OK) {   \
        az_forest_destroy(f);                                 \
        return rc;                                            \
    }
    AZ_ALLOC(d.nodes, nodes);
    AZ_ALLOC(d.root_words, (size_t)n_trees * 2 * nw);
    AZ_ALLOC(d.root_player, (size_t)n_trees);
    AZ_ALLOC(d.root_mc, (size_t)n_trees);
    AZ_ALLOC(d.active, (size_t)n_trees);
    AZ_ALLOC(d.n_nodes, (size_t)n_trees);
    AZ_ALLOC(d.sims_done, (size_t)n_trees);
    AZ_ALLOC(d.n_pending, (size_t)n_trees);
    AZ_ALLOC(d.done_flag, (size_t)n_trees);
    AZ_ALLOC(d.path, (size_t)d.max_leaves * d.max_depth);
    AZ_ALLOC(d.path_len, (size_t)d.max_leaves);
    AZ_ALLOC(d.leaf_slot, (size_t)d.max_leaves);
    AZ_ALLOC(d.pend_words, (size_t)d.max_leaves * 2 * nw);
    AZ_ALLOC(d.pend_player, (size_t)d.max_leaves);
    AZ_ALLOC(d.pend_mc, (size_t)d.max_leaves);
    AZ_ALLOC(d.pend_action, (size_t)d.max_leaves);
    AZ_ALLOC(d.pend_leaf, (size_t)d.max_leaves);
    AZ_ALLOC(d.n_leaves, 4);
    d.n_done = d.n_leaves + 1;
    AZ_ALLOC(d.leaf_words, (size_t)d.max_leaves * 2 * nw);
    AZ_ALLOC(d.leaf_player, (size_t)d.max_leaves);
    AZ_ALLOC(d.leaf_tree, (size_t)d.max_leaves);
    AZ_ALLOC(d.priors, (size_t)d.max_leaves * g.action_dim);
    AZ_ALLOC(d.values, (size_t)d.max_leaves);
    AZ_ALLOC(d.root_prior64, (size_t)n_trees * g.max_children);
    AZ_ALLOC(d.stats, (size_t)n_trees * AZ_STAT_COUNT);
    AZ_ALLOC(f->noise_dev, (size_t)n_trees * g.action_dim);
    AZ_ALLOC(f->vis_dev, (size_t)n_trees * g.action_dim);
    AZ_ALLOC(f->q_dev, (size_t)n_trees);
    AZ_ALLOC(f->planes_stage_dev, (size_t)n_trees * g.planes * g.state_dim);
    AZ_ALLOC(f->i32_stage_dev, (size_t)n_trees * 2);
    uint64_t* hash_dev = nullptr;
    AZ_ALLOC(hash_dev, (size_t)g.state_dim);
    AZ_ALLOC(f->cache_ticket, 4);
#undef AZ_ALLOC
    cudaMemset(f->cache_ticket, 0, 16);
    std::vector<uint64_t> hashes(g.state_dim);
    for (int i = 0; i < g.state_dim; ++i) hashes[i] = az_tuple2_hash(i / g.cols, i % g.cols);
    cudaMemcpy(hash_dev, hashes.data(), hashes.size() * 8, cudaMemcpyHostToDevice);
    d.cell_hash = hash_dev;
    d.root_noise = nullptr;
    cudaMemset(d.active, 1, n_trees);
    cudaMemset(d.stats, 0, (size_t)n_trees * AZ_STAT_COUNT * 8);
    cudaMemset(d.done_flag, 1, n_trees);
    cudaMemset(d.n_leaves, 0, 16);
    cudaMemset(d.root_words, 0, (size_t)n_trees * 2 * nw * 8);
    cudaMemset(d.root_player, 0, n_trees);
    cudaMemset(d.root_mc, 0, (size_t)n_trees * 2);
    f->h_stage_bytes = (size_t)n_trees * (size_t)std::max<size_t>((size_t)g.planes * g.state_dim * 4, (size_t)g.action_dim * 8 + 8) + 64;
    if (cudaMallocHost(&f->h_counts, 64) != cudaSuccess || cudaMallocHost(&f->h_stage, f->h_stage_bytes) != cudaSuccess) {
        az_forest_destroy(f);
        return az_fail(AZ_ENOMEM, "cudaMallocHost failed");
    }
    const size_t smem = (size_t)(f->block / f->G) * f->smem_per_group;
    if (smem > 48 * 1024) {
        cudaError_t e;
        if (nw == 1) e = f->G == 8 ? cudaFuncSetAttribute(az_advance_kernel<1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                                   : cudaFuncSetAttribute(az_advance_kernel<1, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        else e = cudaFuncSetAttribute(az_advance_kernel<4, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { az_forest_destroy(f); return az_fail(AZ_ECUDA, "smem attribute"); }
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        az_forest_destroy(f);
        return az_fail(AZ_ECUDA, std::string("az_forest_create: ") + cudaGetErrorString(e));
    }
    // Evaluation cache on by default for boards of <= 64 cells: room for the distinct positions of ~3 searches of every
    // tree (entries older than AZ_CACHE_STALE searches are recycled), at most 2^25 entries.  AZ_NO_CACHE=1 / AZ_CACHE_LOG2=n.
    if (nw == 1 && f->G == 32 && !getenv("AZ_NO_CACHE")) {
        int lg = 10;
        const double want = 3.0 * (double)n_trees * (double)max_iters * (double)std::min(vl_batch, 2);
        while (lg < 25 && (double)((size_t)1 << lg) < want) ++lg;
        if (const char* e2 = getenv("AZ_CACHE_LOG2")) lg = std::max(4, std::min(28, atoi(e2)));
        if ((rc = az_forest_cache_config(f, lg)) != AZ_OK) {
            az_forest_destroy(f);
            return rc;
        }
    }
    *out = f;
    return AZ_OK;
}

static void az_cache_free(az_forest* f) {
    for (void* p : f->cache_allocs) cudaFree(p);
    f->cache_allocs.clear();
    f->d.cache = AzCacheDev{};
    f->d.max_rounds = 1;
}

// Evaluation cache of 2^log2_entries entries (0 = off).  Replaces any existing cache (and the step graphs that baked
// its pointers in).  Each entry costs 24 + 4 * (A + 1) bytes.
extern "C" int az_forest_cache_config(az_forest* f, int log2_entries) {
    if (!f || log2_entries < 0 || log2_entries > 28) return az_fail(AZ_EINVAL, "cache_config: log2_entries in 0..28");
    if (log2_entries && log2_entries < 4) log2_entries = 4;
    cudaDeviceSynchronize();
    az_net_forget_forest(f->uid);
    az_cache_free(f);
    if (!log2_entries) return AZ_OK;
    if (f->d.g.nw != 1 || f->G != 32) return az_fail(AZ_EINVAL, "cache_config: boards of <= 64 cells only");
    const size_t cap = (size_t)1 << log2_entries;
    const int A = f->d.g.action_dim;
    AzCacheDev c{};
    auto alloc = [&](void** p, size_t bytes) {
        if (cudaMalloc(p, bytes + 16) != cudaSuccess) return false;
        f->cache_allocs.push_back(*p);
        return true;
    };
    void *keys = nullptr, *state = nullptr, *values = nullptr, *priors = nullptr, *epoch = nullptr, *le = nullptr;
    if (!alloc(&keys, cap * 16) || !alloc(&state, cap * 4) || !alloc(&values, cap * 4) || !alloc(&priors, cap * (size_t)A * 4) ||
        !alloc(&epoch, 16) || !alloc(&le, (size_t)f->d.max_leaves * 4)) {
        az_cache_free(f);
        cudaGetLastError();
        return az_fail(AZ_ENOMEM, "cache_config: cudaMalloc failed");
    }
    c.cap = (uint32_t)cap;
    c.keys = reinterpret_cast<ulonglong2*>(keys);
    c.state = reinterpret_cast<uint32_t*>(state);
    c.values = reinterpret_cast<float*>(values);
    c.priors = reinterpret_cast<float*>(priors);
    c.epoch = reinterpret_cast<uint32_t*>(epoch);
    c.leaf_entry = reinterpret_cast<uint32_t*>(le);
    c.n_prev = nullptr;
    cudaMemset(state, 0, cap * 4);
    cudaMemset(keys, 0, cap * 16);
    cudaMemset(epoch, 0, 16);
    cudaMemset(le, 0xff, (size_t)f->d.max_leaves * 4);
    f->d.cache = c;
    // Simulations per tree and launch while leaves hit the cache (vl == 1).  Measured on Connect4 4096 x 800 (B200, round 2):
    // 46.5 / 41.0 / 32.3 / 23.1 / 15.9 M sims/s at 1 / 2 / 4 / 8 / 16 -- the slowest tree still needs ~iters launches per
    // search (a sampled move into a rarely visited child leaves nothing to reuse) while every launch lasts as long as its
    // slowest warp's R simulations -- so extra rounds do not pay; the cache's gain is the evaluator rows it removes.
    int rounds = 1;
    if (const char* e = getenv("AZ_CACHE_ROUNDS")) rounds = std::max(1, std::min(64, atoi(e)));
    f->d.max_rounds = rounds;
    if (cudaDeviceSynchronize() != cudaSuccess) return az_fail(AZ_ECUDA, "cache_config: device error");
    return AZ_OK;
}

// Forget every cached answer (main.py:55-57 clears MCTS.cache after each model update).
extern "C" int az_forest_cache_clear(az_forest* f, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "cache_clear: null");
    f->eval_tag_set = false;                                  // an empty cache belongs to whichever evaluator fills it next
    if (!f->d.cache.cap) return AZ_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaMemsetAsync(f->d.cache.state, 0, (size_t)f->d.cache.cap * 4, s) != cudaSuccess) return az_fail(AZ_ECUDA, "cache_clear: memset");
    if (cudaMemsetAsync(f->d.cache.leaf_entry, 0xff, (size_t)f->d.max_leaves * 4, s) != cudaSuccess) return az_fail(AZ_ECUDA, "cache_clear: memset");
    return AZ_OK;
}

extern "C" int az_forest_cache_info(az_forest* f, long long* entries, long long* bytes) {
    if (!f) return az_fail(AZ_EINVAL, "cache_info: null");
    const long long cap = f->d.cache.cap;
    if (entries) *entries = cap;
    if (bytes) *bytes = cap * (24 + 4 * (long long)(f->d.g.action_dim + 1));
    return AZ_OK;
}

int az_forest_set_eval_tag(az_forest* f, unsigned long long tag, cudaStream_t s) {
    if (f->eval_tag_set && f->eval_tag == tag) return AZ_OK;
    const bool had = f->eval_tag_set;
    const int rc = had ? az_forest_cache_clear(f, s) : AZ_OK;
    f->eval_tag = tag;
    f->eval_tag_set = true;
    return rc;
}
// Step-by-step searches (az_forest_advance + an evaluator call): the evaluator must be the one the cache was filled by.
int az_forest_check_eval_tag(az_forest* f, unsigned long long tag) {
    if (!f->d.cache.cap || f->d.ucb1) return AZ_OK;
    if (!f->eval_tag_set) {
        f->eval_tag = tag;
        f->eval_tag_set = true;
        return AZ_OK;
    }
    if (f->eval_tag != tag)
        return az_fail(AZ_ESTATE, "the evaluation cache holds another evaluator's answers: call az_forest_set_evaluator (or "
                                  "az_forest_cache_clear) before az_forest_begin_move when the evaluator changes");
    return AZ_OK;
}
unsigned long long az_hash_eval_tag(uint64_t seed, int qbits) {
    return 0x4841534800000000ull ^ (seed * 0x9E3779B97F4A7C15ull) ^ (unsigned long long)qbits;
}
// Public form of the tag: the caller names the evaluator whose answers may be cached from now on (any 64-bit id that
// changes when the evaluator or its weights change); a different id drops the cached answers.
extern "C" int az_forest_set_evaluator(az_forest* f, unsigned long long id, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "set_evaluator: null");
    return az_forest_set_eval_tag(f, id, (cudaStream_t)stream);
}

int az_forest_cache_tick(az_forest* f, cudaStream_t s) {
    if (!f->d.cache.cap) return AZ_OK;
    az_cache_tick_kernel<<<1, 1, 0, s>>>(f->d);
    return AZ_OK;
}

// vl == 1 with the cache: up to `rounds` simulations per tree and launch while the leaves hit the cache.  1 for lockstep
// searches (see az_forest_cache_config); continuous self-play (az_selfplay_run_async_*) gains from more, because there a
// tree that runs ahead starts its next search instead of waiting for the slowest one.
extern "C" int az_forest_set_rounds(az_forest* f, int rounds) {
    if (!f || rounds < 1 || rounds > 64) return az_fail(AZ_EINVAL, "set_rounds: 1 <= rounds <= 64");
    rounds = (f->d.cache.cap && f->d.vl == 1) ? rounds : 1;
    if (rounds != f->d.max_rounds) {
        cudaDeviceSynchronize();
        az_net_forget_forest(f->uid);                            // step graphs hold the forest's parameters by value
        f->d.max_rounds = rounds;
    }
    return AZ_OK;
}

unsigned long long az_forest_uid(az_forest* f) { return f->uid; }
// Launches to queue before the next look at the done counter.  vl == 1 with one simulation per launch: every launch
// retires exactly one simulation per live tree, iters + 1 launches finish the search.  Otherwise a launch retires up to
// `vl` (or, with the cache at vl == 1, up to max_rounds) simulations per tree: queue the least number that could
// finish, then poll.
int az_search_burst(const AzForestDev* d, int iters, int steps_done) {
    if (d->vl == 1 && d->max_rounds <= 1) return steps_done == 0 ? iters + 1 : (steps_done <= iters ? iters + 1 - steps_done : 4);
    const int per = d->vl > 1 ? d->vl : d->max_rounds;
    if (steps_done == 0) return std::max(8, (iters + per - 1) / per + 1);
    return d->vl > 1 ? 4 : 16;
}
double* az_forest_noise_dev(az_forest* f) { return f->noise_dev; }
int az_forest_use_device_noise(az_forest* f, int on) {
    f->d.use_noise = on ? 1 : 0;
    f->d.root_noise = on ? f->noise_dev : nullptr;
    return AZ_OK;
}

extern "C" void az_forest_destroy(az_forest* f) {
    if (!f) return;
    cudaDeviceSynchronize();
    az_net_forget_forest(f->uid);                              // step graphs hold this forest's device pointers
    az_cache_free(f);
    for (void* p : f->allocs) cudaFree(p);
    if (f->h_counts) cudaFreeHost(f->h_counts);
    if (f->h_stage) cudaFreeHost(f->h_stage);
    delete f;
}

AzForestDev* az_forest_dev(az_forest* f) { return &f->d; }

static int az_set_players(az_forest* f, const int32_t* players, const int32_t* mcs, cudaStream_t s) {
    const int n = f->d.n_trees;
    AZ_CUDA_OK(cudaMemcpyAsync(f->i32_stage_dev, players, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    AZ_CUDA_OK(cudaMemcpyAsync(f->i32_stage_dev + n, mcs, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    az_roots_from_i32_kernel<<<(n + 255) / 256, 256, 0, s>>>(f->d, f->i32_stage_dev, f->i32_stage_dev + n);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

extern "C" int az_forest_set_roots_planes(az_forest* f, const float* planes_host, const int32_t* players_host,
                                          const int32_t* move_counts_host, void* stream) {
    if (!f || !planes_host || !players_host || !move_counts_host) return az_fail(AZ_EINVAL, "set_roots_planes: null");
    cudaStream_t s = (cudaStream_t)stream;
    const AzGame& g = f->d.g;
    const size_t bytes = (size_t)f->d.n_trees * g.planes * g.state_dim * 4;
    AZ_CUDA_OK(cudaMemcpyAsync(f->planes_stage_dev, planes_host, bytes, cudaMemcpyHostToDevice, s));
    const int n = f->d.n_trees * 2 * g.nw;
    az_planes_to_words_kernel<<<(n + 127) / 128, 128, 0, s>>>(f->planes_stage_dev, f->d.n_trees, g, f->d.root_words);
    AZ_CUDA_OK(cudaGetLastError());
    return az_set_players(f, players_host, move_counts_host, s);
}

extern "C" int az_forest_set_roots_words(az_forest* f, const uint64_t* words_host, const int32_t* players_host,
                                         const int32_t* move_counts_host, void* stream) {
    if (!f || !words_host || !players_host || !move_counts_host) return az_fail(AZ_EINVAL, "set_roots_words: null");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaMemcpyAsync(f->d.root_words, words_host, (size_t)f->d.n_trees * 2 * f->d.g.nw * 8,
                               cudaMemcpyHostToDevice, s));
    return az_set_players(f, players_host, move_counts_host, s);
}

extern "C" int az_forest_set_active(az_forest* f, const uint8_t* active_host, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "set_active: null");
    cudaStream_t s = (cudaStream_t)stream;
    if (!active_host) AZ_CUDA_OK(cudaMemsetAsync(f->d.active, 1, f->d.n_trees, s));
    else AZ_CUDA_OK(cudaMemcpyAsync(f->d.active, active_host, f->d.n_trees, cudaMemcpyHostToDevice, s));
    return AZ_OK;
}

extern "C" int az_forest_set_noise(az_forest* f, const double* noise_host, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "set_noise: null");
    cudaStream_t s = (cudaStream_t)stream;
    if (!noise_host) {
        f->d.use_noise = 0;
        f->d.root_noise = nullptr;
        return AZ_OK;
    }
    AZ_CUDA_OK(cudaMemcpyAsync(f->noise_dev, noise_host, (size_t)f->d.n_trees * f->d.g.action_dim * 8,
                               cudaMemcpyHostToDevice, s));
    f->d.use_noise = 1;
    f->d.root_noise = f->noise_dev;
    return AZ_OK;
}

extern "C" int az_forest_begin_move(az_forest* f, int iters, void* stream) {
    if (!f || iters < 0 || iters > f->max_iters) return az_fail(AZ_EINVAL, "begin_move: iters out of range");
    f->d.iters = iters;
    AZ_CUDA_OK(cudaMemsetAsync(f->d.n_leaves, 0, 2 * sizeof(int), (cudaStream_t)stream));
    az_begin_move_kernel<<<(f->d.n_trees + 255) / 256, 256, 0, (cudaStream_t)stream>>>(f->d);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

extern "C" int az_forest_advance(az_forest* f, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "advance: null");
    cudaStream_t s = (cudaStream_t)stream;
    if (f->d.cache.cap) {
        const long long work = (long long)f->d.max_leaves * (f->d.g.action_dim + 1);
        const int grid = (int)std::max<long long>(1, std::min<long long>((work + 1023) / 1024, 148 * 4));
        az_cache_publish_kernel<<<grid, 256, 0, s>>>(f->d, f->cache_ticket);
    } else {
        AZ_CUDA_OK(cudaMemsetAsync(f->d.n_leaves, 0, sizeof(int), s));
    }
    if (f->d.g.nw == 1) {
        if (f->G == 8) az_launch_advance<1, 8>(f, s);
        else az_launch_advance<1, 32>(f, s);
    } else {
        az_launch_advance<4, 32>(f, s);
    }
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

extern "C" int az_forest_counts(az_forest* f, int* n_leaves, int* n_done, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "counts: null");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaMemcpyAsync(f->h_counts, f->d.n_leaves, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    if (n_leaves) *n_leaves = f->h_counts[0];
    if (n_done) *n_done = f->h_counts[1];
    return AZ_OK;
}

extern "C" int az_eval_hash(az_forest* f, uint64_t seed, int qbits, void* stream) {
    if (!f || qbits < 1 || qbits > 16) return az_fail(AZ_EINVAL, "eval_hash: qbits in 1..16");
    if (int rc = az_forest_check_eval_tag(f, az_hash_eval_tag(seed, qbits))) return rc;
    int lg = 0;
    while ((1 << lg) < f->d.g.action_dim) ++lg;
    const float scale = ldexpf(1.0f, -(qbits + lg));
    const long long work = (long long)f->d.max_leaves * (f->d.g.action_dim + 1);
    const int grid = (int)std::min<long long>((work + 255) / 256, 148 * 16);
    az_eval_hash_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(f->d, seed, qbits, scale);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

// Vanilla MCTS (model=None, ai/mcts.py:57-59): mode 1 = UCB1 selection, children without priors.
extern "C" int az_forest_set_mode(az_forest* f, int mode) {
    if (!f || (mode != 0 && mode != 1)) return az_fail(AZ_EINVAL, "forest_set_mode: mode 0 (PUCT) or 1 (UCB1)");
    if (mode == 1 && !f->d.log_tab) {
        const int n = f->max_iters + f->d.vl + 2;
        std::vector<double> tab((size_t)n, 0.0);
        for (int i = 1; i < n; ++i) tab[i] = std::log((double)i);          // the C library's log = Python's math.log
        double* dev = nullptr;
        int rc = az_dev_alloc(f, &dev, (size_t)n);
        if (rc) return rc;
        AZ_CUDA_OK(cudaMemcpy(dev, tab.data(), (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
        f->d.log_tab = dev;
    }
    f->d.ucb1 = mode;
    return AZ_OK;
}

// MCTS.simulate for every pending leaf (random rollouts; `counter` separates the steps of one search).
extern "C" int az_eval_rollout(az_forest* f, uint64_t seed, uint64_t counter, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "eval_rollout: null forest");
    const int grid = std::max(1, std::min((f->d.max_leaves + 63) / 64, 148 * 8));
    az_eval_rollout_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(f->d, seed, counter);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

// Whole vanilla search: UCB1 + random rollouts, host loop inside the library (as az_forest_search_hash).
extern "C" int az_forest_search_rollout(az_forest* f, int iters, uint64_t seed, void* stream) {
    if (!f || !f->d.ucb1) return az_fail(AZ_ESTATE, "search_rollout: call az_forest_set_mode(f, 1) first");
    int rc = az_forest_begin_move(f, iters, stream);
    if (rc) return rc;
    int done = 0, steps = 0;
    for (;;) {
        const int burst = steps == 0 ? (f->d.vl == 1 ? iters + 1 : std::max(8, iters / f->d.vl)) : 4;
        for (int i = 0; i < burst; ++i) {
            if ((rc = az_forest_advance(f, stream))) return rc;
            if ((rc = az_eval_rollout(f, seed, (uint64_t)steps, stream))) return rc;
            ++steps;
        }
        if ((rc = az_forest_counts(f, nullptr, &done, stream))) return rc;
        if (done >= f->d.n_trees) break;
        if (steps > 2 * iters + 16) return az_fail(AZ_ESTATE, "search_rollout: did not converge");
    }
    return AZ_OK;
}

extern "C" int az_leaves_to_planes(az_forest* f, float* planes_dev, int max_b, void* stream) {
    if (!f || !planes_dev) return az_fail(AZ_EINVAL, "leaves_to_planes: null");
    az_leaves_to_planes_kernel<<<148 * 4, 256, 0, (cudaStream_t)stream>>>(f->d, planes_dev, max_b);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

extern "C" int az_leaves_set_logits(az_forest* f, const float* logits_dev, const float* values_dev, int max_b,
                                    void* stream) {
    if (!f || !logits_dev || !values_dev) return az_fail(AZ_EINVAL, "leaves_set_logits: null");
    const int nb = std::min(max_b, f->d.max_leaves);
    az_logits_to_priors_kernel<<<(nb * 32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(f->d, logits_dev, values_dev, max_b);
    AZ_CUDA_OK(cudaGetLastError());
    return AZ_OK;
}

extern "C" int az_forest_search_hash(az_forest* f, int iters, uint64_t seed, int qbits, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "search_hash: null");
    int rc = az_forest_set_eval_tag(f, az_hash_eval_tag(seed, qbits), (cudaStream_t)stream);
    if (rc) return rc;
    if ((rc = az_forest_begin_move(f, iters, stream))) return rc;
    // vl == 1: every advance retires one simulation per live tree, so iters + 1 advances
    // finish the search (the last one only applies results); vl > 1 polls the done counter.
    int done = 0, steps = 0;
    for (;;) {
        const int burst = az_search_burst(&f->d, iters, steps);
        for (int i = 0; i < burst; ++i) {
            if ((rc = az_forest_advance(f, stream))) return rc;
            if ((rc = az_eval_hash(f, seed, qbits, stream))) return rc;
            ++steps;
        }
        if ((rc = az_forest_counts(f, nullptr, &done, stream))) return rc;
        if (done >= f->d.n_trees) break;
        if (steps > 2 * iters + 16) return az_fail(AZ_ESTATE, "search_hash: did not converge");
    }
    return AZ_OK;
}

extern "C" int az_forest_root_visits(az_forest* f, int32_t* visits_host, double* q_host, void* stream) {
    if (!f || !visits_host) return az_fail(AZ_EINVAL, "root_visits: null");
    cudaStream_t s = (cudaStream_t)stream;
    const int n = f->d.n_trees, A = f->d.g.action_dim;
    // device scratch and the pinned staging buffer are owned by the forest: a stream-ordered allocation
    // per call let the pool trim and re-map memory at the synchronize (spikes of 20-300 ms per call)
    int* h_vis = reinterpret_cast<int*>(f->h_stage);
    double* h_q = reinterpret_cast<double*>(reinterpret_cast<char*>(f->h_stage) + (((size_t)n * A * 4 + 15) & ~(size_t)15));
    az_root_visits_kernel<<<n, 64, 0, s>>>(f->d, f->vis_dev, f->q_dev);
    AZ_CUDA_OK(cudaMemcpyAsync(h_vis, f->vis_dev, (size_t)n * A * 4, cudaMemcpyDeviceToHost, s));
    if (q_host) AZ_CUDA_OK(cudaMemcpyAsync(h_q, f->q_dev, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    memcpy(visits_host, h_vis, (size_t)n * A * 4);
    if (q_host) memcpy(q_host, h_q, (size_t)n * 8);
    return AZ_OK;
}

extern "C" int az_forest_root_best(az_forest* f, int32_t* cells_host, void* stream) {
    if (!f || !cells_host) return az_fail(AZ_EINVAL, "root_best: null");
    cudaStream_t s = (cudaStream_t)stream;
    const int n = f->d.n_trees;
    az_root_best_kernel<<<(n + 127) / 128, 128, 0, s>>>(f->d, f->vis_dev);
    AZ_CUDA_OK(cudaMemcpyAsync(f->h_stage, f->vis_dev, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    memcpy(cells_host, f->h_stage, (size_t)n * 4);
    return AZ_OK;
}

extern "C" int az_forest_root_children(az_forest* f, int tree, int max_children, int32_t* cells_host,
                                       int32_t* visits_host, double* values_host, double* priors_host,
                                       int32_t* root_visit, double* root_value, void* stream) {
    if (!f || tree < 0 || tree >= f->d.n_trees) return az_fail(AZ_EINVAL, "root_children: bad tree");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    AzNode root;
    AZ_CUDA_OK(cudaMemcpy(&root, f->d.nodes + (size_t)tree * f->d.cap, sizeof(root), cudaMemcpyDeviceToHost));
    const AzMeta m = root.meta;
    if (root_visit) *root_visit = root.hot.N;
    if (root_value) *root_value = root.W;
    if (m.status != AZ_ST_EXPANDED) return 0;
    const int nc = std::min<int>(m.nchild, max_children);
    std::vector<AzNode> kids(nc);
    std::vector<double> p64(nc);
    AZ_CUDA_OK(cudaMemcpy(kids.data(), f->d.nodes + (size_t)tree * f->d.cap + m.child_base, nc * sizeof(AzNode), cudaMemcpyDeviceToHost));
    if (f->d.use_noise)
        AZ_CUDA_OK(cudaMemcpy(p64.data(), f->d.root_prior64 + (size_t)tree * f->d.g.max_children, nc * 8, cudaMemcpyDeviceToHost));
    for (int c = 0; c < nc; ++c) {
        if (cells_host) cells_host[c] = kids[c].meta.action;
        if (visits_host) visits_host[c] = kids[c].hot.N;
        if (values_host) values_host[c] = kids[c].W;
        if (priors_host) priors_host[c] = f->d.use_noise ? p64[c] : (double)kids[c].hot.P;
    }
    return m.nchild;
}

extern "C" int az_forest_read_leaves(az_forest* f, int max_n, uint64_t* words_host, uint8_t* players_host,
                                     int32_t* trees_host, float* priors_host, float* values_host, void* stream) {
    if (!f) return az_fail(AZ_EINVAL, "read_leaves: null");
    cudaStream_t s = (cudaStream_t)stream;
    int n = 0;
    int rc = az_forest_counts(f, &n, nullptr, stream);
    if (rc) return rc;
    n = std::min(n, max_n);
    const int nw = f->d.g.nw, A = f->d.g.action_dim;
    if (n > 0) {
        if (words_host) AZ_CUDA_OK(cudaMemcpyAsync(words_host, f->d.leaf_words, (size_t)n * 2 * nw * 8, cudaMemcpyDeviceToHost, s));
        if (players_host) AZ_CUDA_OK(cudaMemcpyAsync(players_host, f->d.leaf_player, (size_t)n, cudaMemcpyDeviceToHost, s));
        if (trees_host) AZ_CUDA_OK(cudaMemcpyAsync(trees_host, f->d.leaf_tree, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        if (priors_host) AZ_CUDA_OK(cudaMemcpyAsync(priors_host, f->d.priors, (size_t)n * A * 4, cudaMemcpyDeviceToHost, s));
        if (values_host) AZ_CUDA_OK(cudaMemcpyAsync(values_host, f->d.values, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        AZ_CUDA_OK(cudaStreamSynchronize(s));
    }
    return n;
}

extern "C" int az_forest_stats(az_forest* f, uint64_t* out8_host, int reset, void* stream) {
    if (!f || !out8_host) return az_fail(AZ_EINVAL, "stats: null");
    cudaStream_t s = (cudaStream_t)stream;
    AZ_CUDA_OK(cudaStreamSynchronize(s));
    std::vector<unsigned long long> all((size_t)f->d.n_trees * AZ_STAT_COUNT);
    AZ_CUDA_OK(cudaMemcpy(all.data(), f->d.stats, all.size() * 8, cudaMemcpyDeviceToHost));
    for (int k = 0; k < AZ_STAT_COUNT; ++k) out8_host[k] = 0;
    for (size_t t = 0; t < (size_t)f->d.n_trees; ++t)
        for (int k = 0; k < AZ_STAT_COUNT; ++k) out8_host[k] += all[t * AZ_STAT_COUNT + k];
    if (reset) AZ_CUDA_OK(cudaMemset(f->d.stats, 0, all.size() * 8));
    return AZ_OK;
}

extern "C" int az_forest_info(az_forest* f, int* n_trees, int* action_dim, int* planes, int* nw, int* cap) {
    if (!f) return az_fail(AZ_EINVAL, "info: null");
    if (n_trees) *n_trees = f->d.n_trees;
    if (action_dim) *action_dim = f->d.g.action_dim;
    if (planes) *planes = f->d.g.planes;
    if (nw) *nw = f->d.g.nw;
    if (cap) *cap = f->d.cap;
    return AZ_OK;
}

// ---- batched rules (differential tests) ----------------------------------------------------
extern "C" int az_rules_legal(int game, int rows, int cols, int n, const uint64_t* words_host,
                              int32_t* out_counts_host, uint8_t* out_cells_host) {
    AzGame g;
    int rc = az_make_game(game, rows, cols, &g);
    if (rc) return rc;
    uint64_t *w = nullptr, *hd = nullptr;
    int32_t* cnt = nullptr;
    uint8_t *cells = nullptr, *scratch = nullptr;
    std::vector<uint64_t> hashes(g.state_dim);
    for (int i = 0; i < g.state_dim; ++i) hashes[i] = az_tuple2_hash(i / g.cols, i % g.cols);
    AZ_CUDA_OK(cudaMalloc(&w, (size_t)n * 2 * g.nw * 8));
    AZ_CUDA_OK(cudaMalloc(&hd, (size_t)g.state_dim * 8));
    AZ_CUDA_OK(cudaMalloc(&cnt, (size_t)n * 4));
    AZ_CUDA_OK(cudaMalloc(&cells, (size_t)n * g.state_dim));
    AZ_CUDA_OK(cudaMalloc(&scratch, (size_t)n * 2 * (g.set_slots + 1)));
    AZ_CUDA_OK(cudaMemcpy(w, words_host, (size_t)n * 2 * g.nw * 8, cudaMemcpyHostToDevice));
    AZ_CUDA_OK(cudaMemcpy(hd, hashes.data(), (size_t)g.state_dim * 8, cudaMemcpyHostToDevice));
    if (g.nw == 1) az_rules_legal_kernel<1><<<(n + 63) / 64, 64>>>(g, n, w, hd, cnt, cells, scratch);
    else az_rules_legal_kernel<4><<<(n + 63) / 64, 64>>>(g, n, w, hd, cnt, cells, scratch);
    AZ_CUDA_OK(cudaGetLastError());
    AZ_CUDA_OK(cudaMemcpy(out_counts_host, cnt, (size_t)n * 4, cudaMemcpyDeviceToHost));
    AZ_CUDA_OK(cudaMemcpy(out_cells_host, cells, (size_t)n * g.state_dim, cudaMemcpyDeviceToHost));
    cudaFree(w); cudaFree(hd); cudaFree(cnt); cudaFree(cells); cudaFree(scratch);
    return AZ_OK;
}

extern "C" int az_rules_winner(int game, int rows, int cols, int n, const uint64_t* words_host,
                               const int32_t* players_host, const int32_t* cells_host, int32_t* out_winner_host) {
    AzGame g;
    int rc = az_make_game(game, rows, cols, &g);
    if (rc) return rc;
    uint64_t* w = nullptr;
    int32_t *pl = nullptr, *ce = nullptr, *out = nullptr;
    AZ_CUDA_OK(cudaMalloc(&w, (size_t)n * 2 * g.nw * 8));
    AZ_CUDA_OK(cudaMalloc(&pl, (size_t)n * 4));
    AZ_CUDA_OK(cudaMalloc(&ce, (size_t)n * 4));
    AZ_CUDA_OK(cudaMalloc(&out, (size_t)n * 4));
    AZ_CUDA_OK(cudaMemcpy(w, words_host, (size_t)n * 2 * g.nw * 8, cudaMemcpyHostToDevice));
    AZ_CUDA_OK(cudaMemcpy(pl, players_host, (size_t)n * 4, cudaMemcpyHostToDevice));
    AZ_CUDA_OK(cudaMemcpy(ce, cells_host, (size_t)n * 4, cudaMemcpyHostToDevice));
    if (g.nw == 1) az_rules_winner_kernel<1><<<(n + 127) / 128, 128>>>(g, n, w, pl, ce, out);
    else az_rules_winner_kernel<4><<<(n + 127) / 128, 128>>>(g, n, w, pl, ce, out);
    AZ_CUDA_OK(cudaGetLastError());
    AZ_CUDA_OK(cudaMemcpy(out_winner_host, out, (size_t)n * 4, cudaMemcpyDeviceToHost));
    cudaFree(w); cudaFree(pl); cudaFree(ce); cudaFree(out);
    return AZ_OK;
}
