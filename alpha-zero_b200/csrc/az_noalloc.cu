// az_rules_legal_noalloc: batched get_valid_moves for callers that may not allocate device memory (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same computation as az_rules_legal (csrc/az_forest.cu: games/*.py get_valid_moves in
// the reference's child order, incl. the CPython-set emulation for Gomoku), but
//   * inputs and outputs are DEVICE buffers of the caller, the call is asynchronous on `stream`;
//   * the kernel's scratch (two set-emulation tables per position + the per-cell tuple hashes) lives in ONE file-scope
//     `static __device__` array of AZ_NOALLOC_SCRATCH_BYTES: every GPU that loads the library holds it until exit;
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine.
// Every call shares the array and finds the previous call's bytes in it.  Nothing has to start at zero: the hash region
// is rewritten by every call before it is read, and az_legal_moves fills each table with 0xFF before it probes it.
// Calls that overlap on the GPU (different streams, or graphs replayed concurrently) RACE on the array: serialise them.
#include <cstdint>
#include <cstdio>
#include <string>
#include "az_types.h"
#include "az_rules.cuh"
#include "az_host.h"
#include "../../include/az_b200.h"

#define AZ_NOALLOC_HASH_BYTES 2048u                    // 256 cells x u64 (AZ_MAX_CELLS = 225)
#define AZ_NOALLOC_TABLE_BYTES (4u << 20)              // set-emulation tables of all positions of one call
#define AZ_NOALLOC_SCRATCH_BYTES (AZ_NOALLOC_HASH_BYTES + AZ_NOALLOC_TABLE_BYTES)
#define AZ_NOALLOC_MAX_DEVICES 64

// In host code this name is a host-side stand-in, NOT a device address: the address the kernels receive comes from
// cudaGetSymbolAddress (az_noalloc_scratch_ptr below), once per device.
static __device__ __align__(256) unsigned char g_az_noalloc_scratch[AZ_NOALLOC_SCRATCH_BYTES];

namespace {

// CPython 3.8+ tuple hash of (r, c) (Objects/tupleobject.c, xxHash-derived): the child order of Gomoku is the slot
// order of a CPython set of such tuples (games/gomoku.py:95,106).  Same function as az_tuple2_hash (host, az_forest.cu).
__device__ __forceinline__ uint64_t az_noalloc_tuple2_hash(uint64_t r, uint64_t c) {
    const uint64_t P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P5 = 2870177450012600261ull;
    uint64_t acc = P5;
    acc += r * P2;
    acc = (acc << 31) | (acc >> 33);
    acc *= P1;
    acc += c * P2;
    acc = (acc << 31) | (acc >> 33);
    acc *= P1;
    acc += 2ull ^ (P5 ^ 3527539ull);
    if (acc == ~0ull) acc = 1546275796ull;
    return acc;
}

__global__ void az_noalloc_hash_kernel(int cols, int cells, uint64_t* hashes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cells) hashes[i] = az_noalloc_tuple2_hash((uint64_t)(i / cols), (uint64_t)(i % cols));
}

template <int NW>
__global__ void az_noalloc_legal_kernel(AzGame g, int n, const uint64_t* __restrict__ words, const uint64_t* __restrict__ cell_hash,
                                        int32_t* __restrict__ counts, uint8_t* __restrict__ cells, uint8_t* scratch, int per_pos) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    AzBoard<NW> b;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        b.w[0][w] = words[(size_t)t * 2 * NW + w];
        b.w[1][w] = words[(size_t)t * 2 * NW + NW + w];
    }
    uint8_t* ta = scratch + (size_t)t * per_pos;
    counts[t] = az_legal_moves<NW>(g, b, cell_hash, ta, ta + g.set_slots + 1, cells + (size_t)t * g.state_dim);
}

// Device address of the scratch array on the CURRENT device, looked up once per device.
int az_noalloc_scratch_ptr(unsigned char** out) {
    static unsigned char* ptrs[AZ_NOALLOC_MAX_DEVICES] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return az_fail(AZ_ECUDA, std::string("az_rules_legal_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= AZ_NOALLOC_MAX_DEVICES) return az_fail(AZ_EINVAL, "az_rules_legal_noalloc: device index beyond 63");
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_az_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return az_fail(AZ_ECUDA, std::string("az_rules_legal_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        ptrs[dev] = reinterpret_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return AZ_OK;
}

}  // namespace

extern "C" int az_rules_legal_noalloc(int game, int rows, int cols, int n, const uint64_t* words_dev, int32_t* out_counts_dev,
                                      uint8_t* out_cells_dev, void* stream) {
    if (n < 0 || !words_dev || !out_counts_dev || !out_cells_dev) return az_fail(AZ_EINVAL, "az_rules_legal_noalloc: bad arguments");
    AzGame g;
    int rc = az_make_game(game, rows, cols, &g);
    if (rc) return rc;
    if (n == 0) return AZ_OK;
    const int per_pos = 2 * (g.set_slots + 1);                  // two tables of set_slots bytes (+1: games without a set still get distinct pointers)
    const long long max_n = (long long)AZ_NOALLOC_TABLE_BYTES / per_pos;
    if ((long long)n > max_n) {
        char msg[256];
        snprintf(msg, sizeof msg, "az_rules_legal_noalloc: n = %d positions need %lld bytes of scratch, the static array holds %u "
                 "(at most %lld positions of this game per call); split the batch or use az_rules_legal",
                 n, (long long)n * per_pos, AZ_NOALLOC_TABLE_BYTES, max_n);
        return az_fail(AZ_EINVAL, msg);
    }
    unsigned char* base = nullptr;
    if ((rc = az_noalloc_scratch_ptr(&base))) return rc;
    uint64_t* hashes = reinterpret_cast<uint64_t*>(base);
    uint8_t* tables = base + AZ_NOALLOC_HASH_BYTES;
    cudaStream_t s = (cudaStream_t)stream;
    az_noalloc_hash_kernel<<<(g.state_dim + 127) / 128, 128, 0, s>>>(g.cols, g.state_dim, hashes);
    if (g.nw == 1) az_noalloc_legal_kernel<1><<<(n + 63) / 64, 64, 0, s>>>(g, n, words_dev, hashes, out_counts_dev, out_cells_dev, tables, per_pos);
    else az_noalloc_legal_kernel<AZ_MAX_WORDS><<<(n + 63) / 64, 64, 0, s>>>(g, n, words_dev, hashes, out_counts_dev, out_cells_dev, tables, per_pos);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return az_fail(AZ_ECUDA, std::string("az_rules_legal_noalloc: launch: ") + cudaGetErrorString(e));
    return AZ_OK;
}
