// Microbenchmark: tcgen05.mma (kind::f16, bf16, SS operands, M = 128) cycles per instruction vs N.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I alpha-zero_b200/csrc -o tools/umma_rate tools/umma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "az_ptx.cuh"

__device__ __forceinline__ void ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}

// mode bit 0: extra warps 4.. hammer tcgen05.ld on columns 384..511; bit 1: every MMA reads a different A tile;
// bit 2: no wait after commit (free-running issue, wait only at the end)
__global__ void __launch_bounds__(384, 1) k(int N, int per_commit, int commits, unsigned long long* out, int mode, volatile int* stop_flag) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (azptx::smem_u32(smem_dyn) & 1023u)) & 1023u);
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) *stop_flag = 0;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3F803F80u;
    if (threadIdx.x == 0) { azptx::mbar_init(&bar, 1); azptx::fence_barrier_init(); }
    if (warp == 0) azptx::tmem_alloc<512>(&slot);
    azptx::fence_proxy_async();
    azptx::tc_fence_before();
    __syncthreads();
    azptx::tc_fence_after();
    if (warp >= 4) {
        if (mode & 1) {
            const uint32_t tq = slot + ((uint32_t)((warp & 3) * 32) << 16) + 384;
            uint32_t acc = 0;
            while (!*stop_flag) {
                uint32_t r[6][8];
#pragma unroll
                for (int j = 0; j < 6; ++j) ld8(tq + j * 8, r[j]);
                azptx::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 6; ++j)
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc ^= r[j][e];
            }
            if (acc == 0x1234567u) out[147] = acc;
        }
    } else if (warp == 1) {
        const uint32_t idesc = azptx::umma_idesc_bf16(128, (uint32_t)N);
        const uint64_t da = azptx::umma_desc_k128(azptx::smem_u32(smem));
        const uint64_t db = azptx::umma_desc_k128(azptx::smem_u32(smem + 32768));
        unsigned long long t0 = 0;
        for (int c = 0; c < commits + 1; ++c) {
            if (c == 1) t0 = clock64();
            if (lane == 0) {
                for (int i = 0; i < per_commit; ++i) {
                    const uint64_t a2 = (mode & 2) ? da + (uint64_t)(((c * per_commit + i) >> 2) % 4) * (16384 >> 4) : da;
                    azptx::umma_bf16(slot + (c & 1) * 128, a2 + 2 * (i & 3), db + 2 * (i & 3), idesc, i != 0);
                }
                azptx::umma_commit(&bar);
            }
            __syncwarp();
            if (!(mode & 4) || c == 0 || c == commits) azptx::mbar_wait(&bar, c & 1);
        }
        *stop_flag = 1;
        const unsigned long long t1 = clock64();
        if (lane == 0) out[blockIdx.x] = t1 - t0;
    }
    azptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) azptx::tmem_dealloc<512>(slot);
}

int main() {
    unsigned long long* d;
    int* flag;
    cudaMalloc(&d, 8 * 148);
    cudaMalloc(&flag, 4 * 148);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    for (int mode : {0, 1, 2, 3})
    for (int per : {8, 32}) {
        for (int N : {96, 128}) {
            const int commits = 200;
            printf("mode %d ", mode);
            k<<<1, 384, 100 * 1024>>>(N, per, commits, d, mode, flag);
            cudaError_t e = cudaDeviceSynchronize();
            unsigned long long h[148];
            cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            printf("M=128 N=%3d K=16, %2d MMAs per commit+wait: %7.1f cycles per MMA (floor %5.1f)  %s\n", N, per,
                   (double)h[0] / (commits * per), 128.0 * N / 256.0, cudaGetErrorString(e));
        }
    }
    return 0;
}
