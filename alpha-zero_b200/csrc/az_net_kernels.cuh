// CUDA-core kernels of the policy/value net (everything that is not a dense projection):
// board -> token embedding (+LayerNorm), LayerNorm, multi-head attention, cls-query
// attention of the pruned last block, final LayerNorm + heads + softmax.
// Reference: ai/nn.py:5-36 (embedding), :38-61 (block), :74-84 (heads); ai/mcts.py:49 (softmax).
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

typedef __nv_bfloat16 bf16;

// Items (boards) this launch handles: the evaluation batch size lives on the device.
struct AzBatch {
    const int* count;   // device counter (forest n_leaves) or nullptr
    int fixed;          // used when count == nullptr
    int offset, cap;    // this launch covers items [offset, offset + cap)
};
__device__ __forceinline__ int az_batch_items(const AzBatch& b) {
    int n = (b.count ? *b.count : b.fixed) - b.offset;
    return n < 0 ? 0 : (n > b.cap ? b.cap : n);
}

__device__ __forceinline__ float az_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float az_warp_max(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ void az_unpack8(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 az_pack8(const float (&f)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return u;
}

// LayerNorm of a row held as NV x 8 channels per lane (channel = (v*32+lane)*8 + e), eps 1e-5,
// biased variance (torch.nn.LayerNorm).  Lanes whose vector lies beyond D hold zeros.
template <int NV>
__device__ __forceinline__ void az_row_layernorm(float (&x)[NV][8], int D, int lane, const float* __restrict__ gamma,
                                                 const float* __restrict__ beta) {
    float s = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) s += x[v][e];
    const float mean = az_warp_sum(s) / (float)D;
    float q = 0.f;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const bool on = (v * 32 + lane) * 8 < D;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float d = on ? x[v][e] - mean : 0.f;
            q += d * d;
        }
    }
    const float rstd = rsqrtf(az_warp_sum(q) / (float)D + 1e-5f);
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int d0 = (v * 32 + lane) * 8;
        if (d0 < D) {
            const float4 g0 = *reinterpret_cast<const float4*>(gamma + d0), g1 = *reinterpret_cast<const float4*>(gamma + d0 + 4);
            const float4 b0 = *reinterpret_cast<const float4*>(beta + d0), b1 = *reinterpret_cast<const float4*>(beta + d0 + 4);
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) x[v][e] = (x[v][e] - mean) * rstd * gg[e] + bb[e];
        }
    }
}

struct AzEmbedArgs {
    AzBatch batch;
    const uint64_t* leaf_words;   // [items][2][nw] canonical: mine, theirs
    const uint8_t* leaf_player;   // side to move (plane 2 of the 3-plane games)
    int nw, rows, cols, planes, patch, D, T;
    const bf16* wt;               // [planes*patch*patch][D] conv weight, channel-last
    const float* posb;            // [2][T][D]: pos-emb + conv bias (+ plane-2 sum when stm=1); row 0 = cls + pos[0]
    const float* ln_g;            // LayerNorm 1 of block 0
    const float* ln_b;
    bf16* x;                      // [items*T][D] residual stream
    bf16* xn;                     // [items*T][D] LN1(x)
    int write_all_x;              // 0: only the cls rows of x are needed (depth == 1)
};

// One warp per token: conv patch embedding of binary planes as a sum of weight rows
// (ai/nn.py:9-18), + cls/pos (ai/nn.py:33-35), then LayerNorm 1 of the first block (:54).
template <int NV>
__global__ void __launch_bounds__(256) az_embed_ln_kernel(AzEmbedArgs a) {
    const int items = az_batch_items(a.batch);
    const int lane = threadIdx.x & 31;
    const long long token = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (token >= (long long)items * a.T) return;
    const int b = (int)(token / a.T), t = (int)(token - (long long)b * a.T);
    const int item = a.batch.offset + b;
    const int stm = a.planes == 3 ? a.leaf_player[item] : 0;
    float x[NV][8];
    const float* pb = a.posb + ((size_t)stm * a.T + t) * a.D;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int d0 = (v * 32 + lane) * 8;
        if (d0 < a.D) {
            const float4 p0 = *reinterpret_cast<const float4*>(pb + d0), p1 = *reinterpret_cast<const float4*>(pb + d0 + 4);
            x[v][0] = p0.x; x[v][1] = p0.y; x[v][2] = p0.z; x[v][3] = p0.w;
            x[v][4] = p1.x; x[v][5] = p1.y; x[v][6] = p1.z; x[v][7] = p1.w;
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) x[v][e] = 0.f;
        }
    }
    if (t > 0) {
        const int cell = t - 1, r = cell / a.cols, c = cell - r * a.cols, half = a.patch >> 1;
        const uint64_t* w = a.leaf_words + (size_t)item * 2 * a.nw;
        for (int kr = 0; kr < a.patch; ++kr) {
            const int rr = r + kr - half;
            if (rr < 0 || rr >= a.rows) continue;
            for (int kc = 0; kc < a.patch; ++kc) {
                const int cc = c + kc - half;
                if (cc < 0 || cc >= a.cols) continue;
                const int s = rr * a.cols + cc;
                const uint64_t bit = 1ull << (s & 63);
                int ch = -1;
                if (w[s >> 6] & bit) ch = 0;
                else if (w[a.nw + (s >> 6)] & bit) ch = 1;
                if (ch < 0) continue;
                const bf16* wr = a.wt + (size_t)((ch * a.patch + kr) * a.patch + kc) * a.D;
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int d0 = (v * 32 + lane) * 8;
                    if (d0 < a.D) {
                        float f[8];
                        az_unpack8(*reinterpret_cast<const uint4*>(wr + d0), f);
#pragma unroll
                        for (int e = 0; e < 8; ++e) x[v][e] += f[e];
                    }
                }
            }
        }
    }
    const size_t row = (size_t)b * a.T + t;
    if (a.write_all_x || t == 0) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int d0 = (v * 32 + lane) * 8;
            if (d0 < a.D) *reinterpret_cast<uint4*>(a.x + row * a.D + d0) = az_pack8(x[v]);
        }
    }
    az_row_layernorm<NV>(x, a.D, lane, a.ln_g, a.ln_b);
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int d0 = (v * 32 + lane) * 8;
        if (d0 < a.D) *reinterpret_cast<uint4*>(a.xn + row * a.D + d0) = az_pack8(x[v]);
    }
}

// LayerNorm of bf16 rows: in[row*ld_in .. +D) -> out[row*ld_out .. +D).  One warp per row.
template <int NV>
__global__ void __launch_bounds__(256) az_layernorm_kernel(AzBatch batch, int rows_per_item, const bf16* in, int ld_in,
                                                           bf16* out, int ld_out, int D, const float* g, const float* bt) {
    const long long rows = (long long)az_batch_items(batch) * rows_per_item;
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    float x[NV][8];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int d0 = (v * 32 + lane) * 8;
        if (d0 < D) az_unpack8(*reinterpret_cast<const uint4*>(in + (size_t)row * ld_in + d0), x[v]);
        else {
#pragma unroll
            for (int e = 0; e < 8; ++e) x[v][e] = 0.f;
        }
    }
    az_row_layernorm<NV>(x, D, lane, g, bt);
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int d0 = (v * 32 + lane) * 8;
        if (d0 < D) *reinterpret_cast<uint4*>(out + (size_t)row * ld_out + d0) = az_pack8(x[v]);
    }
}

// Full multi-head self-attention of one (board, head) per block (nn.MultiheadAttention,
// ai/nn.py:42,55; eval mode, no masks): softmax(q k^T / sqrt(hd)) v.  qkv rows hold
// [q | k | v] (3D wide).  Used by every block except the last (which is cls-pruned).
// smem: K and V of the head as fp32 [T][hd+1] each + one probability row per warp.
__global__ void __launch_bounds__(128) az_attention_kernel(AzBatch batch, const bf16* __restrict__ qkv, bf16* __restrict__ out,
                                                           int T, int D, int H) {
    extern __shared__ float sm_att[];
    const int items = az_batch_items(batch);
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    if (b >= items) return;
    const int hd = D / H, ldk = hd + 1;
    float* sk = sm_att;
    float* sv = sk + (size_t)T * ldk;
    float* sp = sv + (size_t)T * ldk;                       // [4 warps][T]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bf16* base = qkv + (size_t)b * T * 3 * D;
    for (int i = threadIdx.x; i < T * hd; i += blockDim.x) {
        const int t = i / hd, d = i - t * hd;
        sk[t * ldk + d] = __bfloat162float(base[(size_t)t * 3 * D + D + h * hd + d]);
        sv[t * ldk + d] = __bfloat162float(base[(size_t)t * 3 * D + 2 * D + h * hd + d]);
    }
    __syncthreads();
    const float scale = rsqrtf((float)hd);
    float* p = sp + warp * T;
    for (int tq = warp; tq < T; tq += 4) {
        const bf16* qp = base + (size_t)tq * 3 * D + h * hd;
        float mx = -1e30f;
        for (int tk = lane; tk < T; tk += 32) {
            float s = 0.f;
            for (int d = 0; d < hd; ++d) s += __bfloat162float(qp[d]) * sk[tk * ldk + d];
            s *= scale;
            p[tk] = s;
            mx = fmaxf(mx, s);
        }
        mx = az_warp_max(mx);
        float sum = 0.f;
        for (int tk = lane; tk < T; tk += 32) {
            const float e = __expf(p[tk] - mx);
            p[tk] = e;
            sum += e;
        }
        sum = az_warp_sum(sum);
        __syncwarp();
        const float inv = 1.0f / sum;
        for (int d = lane; d < hd; d += 32) {
            float o = 0.f;
            for (int tk = 0; tk < T; ++tk) o += p[tk] * sv[tk * ldk + d];
            out[((size_t)b * T + tq) * D + h * hd + d] = __float2bfloat16(o * inv);
        }
        __syncwarp();
    }
}

// Attention of the LAST block restricted to the cls query (only x[:,0] feeds the heads,
// ai/nn.py:81).  One warp per (board, head); kv rows hold [k | v] (2D wide), qc holds the cls
// query rows [items][D].  hd <= 128 (multiple of 8), T <= 256.
//   scores : lane = token; each lane streams its token's K slice (hd bf16, 16-byte loads)
//   softmax: warp max / sum
//   output : lane = 2 head dims; V rows are read coalesced (4 bytes per lane), 4 tokens in flight
__global__ void __launch_bounds__(128) az_cls_attention_kernel(AzBatch batch, const bf16* __restrict__ qc,
                                                               const bf16* __restrict__ kv, bf16* __restrict__ oc, int T,
                                                               int D, int H) {
    __shared__ float s_q[4][128];
    const int items = az_batch_items(batch);
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int w = blockIdx.x * (blockDim.x >> 5) + wib;
    const int b = w / H, h = w - b * H;
    if (b >= items) return;
    const int hd = D / H;
    for (int d = lane; d < hd; d += 32) s_q[wib][d] = __bfloat162float(qc[(size_t)b * D + h * hd + d]);
    __syncwarp();
    const bf16* kb = kv + (size_t)b * T * 2 * D + h * hd;
    const float scale = rsqrtf((float)hd);
    float s[8];
    float mx = -1e30f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int t = j * 32 + lane;
        s[j] = -1e30f;
        if (t < T) {
            const uint4* kr = reinterpret_cast<const uint4*>(kb + (size_t)t * 2 * D);
            float acc = 0.f;
            for (int c = 0; c < hd / 8; ++c) {
                float f[8];
                az_unpack8(__ldg(kr + c), f);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc += f[e] * s_q[wib][c * 8 + e];
            }
            s[j] = acc * scale;
            mx = fmaxf(mx, s[j]);
        }
    }
    mx = az_warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        s[j] = (j * 32 + lane < T) ? __expf(s[j] - mx) : 0.f;
        sum += s[j];
    }
    const float inv = 1.0f / az_warp_sum(sum);
    const bf16* vb = kb + D;
    float o[2][2] = {{0.f, 0.f}, {0.f, 0.f}};                // dims 2*lane(+1) and 64 + 2*lane(+1)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (j * 32 >= T) break;
        const int tn = min(32, T - j * 32);
#pragma unroll 4
        for (int tt = 0; tt < tn; ++tt) {
            const float pt = __shfl_sync(0xffffffffu, s[j], tt);
            const bf16* vr = vb + (size_t)(j * 32 + tt) * 2 * D;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int d = u * 64 + 2 * lane;
                if (d < hd) {
                    const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(vr + d));
                    o[u][0] += pt * v.x;
                    o[u][1] += pt * v.y;
                }
            }
        }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int d = u * 64 + 2 * lane;
        if (d < hd)
            *reinterpret_cast<__nv_bfloat162*>(oc + (size_t)b * D + h * hd + d) = __floats2bfloat162_rn(o[u][0] * inv, o[u][1] * inv);
    }
}

// The same attention with every load issued up front: lane t owns tokens t, t + 32, ... (at most TPL of them) and reads BOTH
// their K and V slices (HD bf16 each, 16-byte loads, independent of one another) before any arithmetic; scores, warp softmax,
// then each lane scales its V slices by its probabilities and the HD partial outputs are summed across the 32 lanes by a
// halving butterfly (HD - HD/32 shuffles: lane d ends with dims d, d + 32).  One memory round trip instead of one per four
// tokens of the V phase.  HD in {16, 32, 64}, T <= 32 * TPL.
template <int HD, int TPL>
__global__ void __launch_bounds__(128) az_cls_attention2_kernel(AzBatch batch, const bf16* __restrict__ qc, const bf16* __restrict__ kv,
                                                                bf16* __restrict__ oc, int T, int D, int H) {
    const int items = az_batch_items(batch);
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int w = blockIdx.x * (blockDim.x >> 5) + wib;
    const int b = w / H, h = w - b * H;
    if (b >= items) return;
    constexpr int C8 = HD / 8;
    const bf16* kb = kv + (size_t)b * T * 2 * D + h * HD;
    uint4 kq[TPL][C8], vq[TPL][C8];
#pragma unroll
    for (int j = 0; j < TPL; ++j) {
        const int t = j * 32 + lane;
        const uint4* kr = reinterpret_cast<const uint4*>(kb + (size_t)(t < T ? t : 0) * 2 * D);
#pragma unroll
        for (int c = 0; c < C8; ++c) {
            kq[j][c] = __ldg(kr + c);
            vq[j][c] = __ldg(kr + D / 8 + c);
        }
    }
    __shared__ float s_q2[4][HD];
    float* q = s_q2[wib];
    for (int d = lane; d < HD; d += 32) q[d] = __bfloat162float(qc[(size_t)b * D + h * HD + d]);
    __syncwarp();
    const float scale = rsqrtf((float)HD);
    float s[TPL];
    float mx = -1e30f;
#pragma unroll
    for (int j = 0; j < TPL; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int c = 0; c < C8; ++c) {
            float f[8];
            az_unpack8(kq[j][c], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc += f[e] * q[c * 8 + e];
        }
        s[j] = (j * 32 + lane < T) ? acc * scale : -1e30f;
        mx = fmaxf(mx, s[j]);
    }
    mx = az_warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < TPL; ++j) {
        s[j] = (j * 32 + lane < T) ? __expf(s[j] - mx) : 0.f;
        sum += s[j];
    }
    const float inv = 1.0f / az_warp_sum(sum);
    float o[HD];
#pragma unroll
    for (int d = 0; d < HD; ++d) o[d] = 0.f;
#pragma unroll
    for (int j = 0; j < TPL; ++j) {
#pragma unroll
        for (int c = 0; c < C8; ++c) {
            float f[8];
            az_unpack8(vq[j][c], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[c * 8 + e] = fmaf(s[j], f[e], o[c * 8 + e]);
        }
    }
    // sum over the 32 lanes, halving the values a lane carries at every step: after the step with offset `off` a lane keeps the
    // dims whose bit `off` (of dim mod 32) equals its own lane bit
#pragma unroll
    for (int off = 16, n = HD; off >= 1; off >>= 1) {
        n >>= 1;
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float send = up ? o[i] : o[i + n];            // the half this lane gives away
            const float keep = up ? o[i + n] : o[i];
            o[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
        if (n == 1 && off > 1) {                                  // HD < 32: fewer dims than lanes, finish by plain reductions
#pragma unroll
            for (int o2 = off >> 1; o2 >= 1; o2 >>= 1) o[0] += __shfl_xor_sync(0xffffffffu, o[0], o2);
            break;
        }
    }
    // which dim did this lane end up with?  step k (offset 16 >> k) splits the current range in two and an `up` lane takes the
    // upper half: dims = bit-reversed picture of the lane id, computed the same way
    if (HD >= 32) {
        int base = 0, span = HD;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            span >>= 1;
            if (lane & off) base += span;
        }
        // span == HD / 32 values left: o[0 .. span)
#pragma unroll
        for (int i = 0; i < HD / 32; ++i) oc[(size_t)b * D + h * HD + base + i] = __float2bfloat16(o[i] * inv);
    } else {
        // HD = 16: after the offsets 16, 8, 4, 2 a lane holds one dim (the last reductions summed the remaining lane pairs)
        int base = 0, span = HD;
#pragma unroll
        for (int off = 16; off >= 2; off >>= 1) {
            span >>= 1;
            if (lane & off) base += span;
        }
        if ((lane & 1) == 0) oc[(size_t)b * D + h * HD + base] = __float2bfloat16(o[0] * inv);
    }
}

struct AzHeadArgs {
    AzBatch batch;
    const bf16* xc;       // [items][D] cls rows after the last block
    int D, A;
    const float *ln_g, *ln_b;     // Net.norm
    const float *w_pol, *b_pol;   // [A][D], [A]
    const float *w_val, *b_val;   // [D], [1]
    float* priors;        // [.. + offset][A]
    float* values;
    float* logits;        // optional raw logits [.. + offset][A] (parity tests), may be null
};

// Final LayerNorm (ai/nn.py:78) + policy/value heads (:82-83) + the softmax of ai/mcts.py:49
// (max-subtracted for range safety; mathematically the same exp(l)/sum(exp(l))).
// One block of 128 threads per board; smem: D + A + 8 floats.
__global__ void __launch_bounds__(128) az_head_kernel(AzHeadArgs a) {
    extern __shared__ float sm_head[];
    const int items = az_batch_items(a.batch);
    const int b = blockIdx.x;
    if (b >= items) return;
    float* xs = sm_head;
    float* lg = xs + a.D;
    float* red = lg + a.A;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float s = 0.f;
    for (int d = threadIdx.x; d < a.D; d += blockDim.x) {
        const float v = __bfloat162float(a.xc[(size_t)b * a.D + d]);
        xs[d] = v;
        s += v;
    }
    s = az_warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    const float mean = (red[0] + red[1] + red[2] + red[3]) / (float)a.D;
    __syncthreads();
    float q = 0.f;
    for (int d = threadIdx.x; d < a.D; d += blockDim.x) {
        const float dv = xs[d] - mean;
        q += dv * dv;
    }
    q = az_warp_sum(q);
    if (lane == 0) red[warp] = q;
    __syncthreads();
    const float rstd = rsqrtf((red[0] + red[1] + red[2] + red[3]) / (float)a.D + 1e-5f);
    for (int d = threadIdx.x; d < a.D; d += blockDim.x) xs[d] = (xs[d] - mean) * rstd * a.ln_g[d] + a.ln_b[d];
    __syncthreads();
    for (int o = warp; o <= a.A; o += 4) {                   // o == A is the value head
        const float* wr = o < a.A ? a.w_pol + (size_t)o * a.D : a.w_val;
        float acc = 0.f;
        for (int d = lane; d < a.D; d += 32) acc += wr[d] * xs[d];
        acc = az_warp_sum(acc);
        if (lane == 0) {
            if (o < a.A) lg[o] = acc + a.b_pol[o];
            else a.values[a.batch.offset + b] = tanhf(acc + a.b_val[0]);
        }
    }
    __syncthreads();
    float mx = -1e30f;
    for (int i = threadIdx.x; i < a.A; i += blockDim.x) mx = fmaxf(mx, lg[i]);
    mx = az_warp_max(mx);
    if (lane == 0) red[4 + warp] = mx;
    __syncthreads();
    mx = fmaxf(fmaxf(red[4], red[5]), fmaxf(red[6], red[7]));
    float sum = 0.f;
    for (int i = threadIdx.x; i < a.A; i += blockDim.x) sum += expf(lg[i] - mx);
    sum = az_warp_sum(sum);
    __syncthreads();
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = red[0] + red[1] + red[2] + red[3];
    const size_t ob = (size_t)(a.batch.offset + b) * a.A;
    for (int i = threadIdx.x; i < a.A; i += blockDim.x) {
        a.priors[ob + i] = expf(lg[i] - mx) / sum;
        if (a.logits) a.logits[ob + i] = lg[i];
    }
}

// cls-query attention when the net has ONE block: the cls query is then a constant of the
// model (LayerNorm(cls + pos[0]) does not depend on the board), so q.K^T is folded into the
// projection itself: scores[t][h] = xn[t] . (Wk_h^T q_h)/sqrt(hd) + c_h arrive as 64 extra
// fp32 output columns of the V projection and K is never materialised.
// One warp per (board, head): softmax over the T scores, then sum_t p_t V[t, head] with
// 16-byte V loads (lane = token group x 8-dim chunk).  hd = 8 * 2^k <= 128.
__global__ void __launch_bounds__(128) az_cls_attention_folded_kernel(AzBatch batch, const float* __restrict__ sc,
                                                                      const bf16* __restrict__ v, int ldv,
                                                                      bf16* __restrict__ oc, int T, int D, int H) {
    const int items = az_batch_items(batch);
    const int lane = threadIdx.x & 31;
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int b = w / H, h = w - b * H;
    if (b >= items) return;
    const int hd = D / H, chunks = hd >> 3, groups = 32 / chunks;
    const float* sb = sc + (size_t)b * T * 64 + h;
    float s[8];
    float mx = -1e30f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int t = j * 32 + lane;
        s[j] = t < T ? __ldg(sb + (size_t)t * 64) : -1e30f;
        mx = fmaxf(mx, s[j]);
    }
    mx = az_warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        s[j] = (j * 32 + lane < T) ? __expf(s[j] - mx) : 0.f;
        sum += s[j];
    }
    const float inv = 1.0f / az_warp_sum(sum);
    const int grp = lane / chunks, c = lane - grp * chunks;
    const bf16* vb = v + (size_t)b * T * ldv + h * hd + c * 8;
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (j * 32 >= T) break;
#pragma unroll 4
        for (int base = 0; base < 32; base += groups) {
            const int tt = base + grp;                      // token within this block of 32
            const float pt = __shfl_sync(0xffffffffu, s[j], tt);
            const int t = j * 32 + tt;
            if (t < T) {
                float f[8];
                az_unpack8(__ldg(reinterpret_cast<const uint4*>(vb + (size_t)t * ldv)), f);
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] += pt * f[e];
            }
        }
    }
    for (int off = chunks; off < 32; off <<= 1)
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += __shfl_xor_sync(0xffffffffu, o[e], off);
    if (grp == 0) {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] *= inv;
        *reinterpret_cast<uint4*>(oc + (size_t)b * D + h * hd + c * 8) = az_pack8(o);
    }
}

// Final LayerNorm + heads + softmax, one WARP per board, head weights resident in shared memory
// (policy rows then the value row, fp32).  For A + 1 <= 32 outputs (TicTacToe, Connect4): lane o keeps
// output o.  Persistent: every warp strides over the boards of the batch.
template <int NV>
__global__ void __launch_bounds__(256) az_head_warp_kernel(AzHeadArgs a) {
    extern __shared__ float sm_hw[];                          // [(A + 1)][D]
    const int items = az_batch_items(a.batch);
    const int D = a.D, A = a.A;
    for (int i = threadIdx.x; i < (A + 1) * D; i += blockDim.x) sm_hw[i] = i < A * D ? a.w_pol[i] : a.w_val[i - A * D];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const float my_bias = lane < A ? a.b_pol[lane] : (lane == A ? a.b_val[0] : 0.f);
    for (int b = blockIdx.x * wpb + (threadIdx.x >> 5); b < items; b += gridDim.x * wpb) {
        float x[NV][8];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int d0 = (v * 32 + lane) * 8;
            if (d0 < D) az_unpack8(*reinterpret_cast<const uint4*>(a.xc + (size_t)b * D + d0), x[v]);
            else {
#pragma unroll
                for (int e = 0; e < 8; ++e) x[v][e] = 0.f;
            }
        }
        az_row_layernorm<NV>(x, D, lane, a.ln_g, a.ln_b);
        float mine = 0.f;
        for (int o = 0; o <= A; ++o) {
            const float* wr = sm_hw + (size_t)o * D;
            float acc = 0.f;
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const int d0 = (v * 32 + lane) * 8;
                if (d0 < D) {
                    const float4 w0 = *reinterpret_cast<const float4*>(wr + d0), w1 = *reinterpret_cast<const float4*>(wr + d0 + 4);
                    acc += x[v][0] * w0.x + x[v][1] * w0.y + x[v][2] * w0.z + x[v][3] * w0.w + x[v][4] * w1.x + x[v][5] * w1.y +
                           x[v][6] * w1.z + x[v][7] * w1.w;
                }
            }
            acc = az_warp_sum(acc);
            if (lane == o) mine = acc;
        }
        mine += my_bias;
        const float mx = az_warp_max(lane < A ? mine : -1e30f);
        const float e = lane < A ? expf(mine - mx) : 0.f;
        const float sum = az_warp_sum(e);
        const size_t ob = (size_t)(a.batch.offset + b) * A;
        if (lane < A) {
            a.priors[ob + lane] = e / sum;
            if (a.logits) a.logits[ob + lane] = mine;
        } else if (lane == A) {
            a.values[a.batch.offset + b] = tanhf(mine);
        }
    }
}

// Softmax / tanh over head outputs produced by the tensor-core head GEMM: ho [items][64] fp32 holds
// A policy logits then the value pre-activation (A + 1 <= 64).  One warp per board.
__global__ void __launch_bounds__(256) az_head_softmax_kernel(AzBatch batch, const float* __restrict__ ho, int A,
                                                              float* priors, float* values, float* logits) {
    const int items = az_batch_items(batch);
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= items) return;
    const float l0 = ho[(size_t)b * 64 + lane], l1 = ho[(size_t)b * 64 + 32 + lane];
    const bool p0 = lane < A, p1 = lane + 32 < A;
    const float mx = az_warp_max(fmaxf(p0 ? l0 : -1e30f, p1 ? l1 : -1e30f));
    const float e0 = p0 ? expf(l0 - mx) : 0.f, e1 = p1 ? expf(l1 - mx) : 0.f;
    const float inv = 1.0f / az_warp_sum(e0 + e1);
    const size_t ob = (size_t)(batch.offset + b) * A;
    if (p0) { priors[ob + lane] = e0 * inv; if (logits) logits[ob + lane] = l0; }
    if (p1) { priors[ob + lane + 32] = e1 * inv; if (logits) logits[ob + lane + 32] = l1; }
    if (lane == A) values[batch.offset + b] = tanhf(l0);
    if (lane + 32 == A) values[batch.offset + b] = tanhf(l1);
}
