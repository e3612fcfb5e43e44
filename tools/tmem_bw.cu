// Microbenchmark: tcgen05.ld throughput per SM (TMEM -> registers), sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tmem_bw tools/tmem_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int X>
__device__ __forceinline__ void ld(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void ld<32>(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
template <>
__device__ __forceinline__ void ld<8>(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}

// `warps` warps (multiple of 4), each issues `iters` x (NL loads of X columns) then one wait per group.
template <int X, int NL>
__global__ void k(int iters, unsigned long long* out, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t acc = 0;
    const unsigned long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        uint32_t r[NL][X];
#pragma unroll
        for (int j = 0; j < NL; ++j) ld<X>(base + ((i * NL + j) * X) % (512 - X + 1) / X * X, r[j]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int j = 0; j < NL; ++j)
#pragma unroll
            for (int e = 0; e < X; ++e) acc ^= r[j][e];
    }
    const unsigned long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot));
}

template <int X, int NL>
void run(int warps, const char* name) {
    unsigned long long* d;
    uint32_t* sink;
    cudaMalloc(&d, 8 * 148);
    cudaMalloc(&sink, 4);
    const int iters = 2000;
    k<X, NL><<<148, warps * 32>>>(iters, d, sink);
    k<X, NL><<<148, warps * 32>>>(iters, d, sink);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const double bytes = (double)warps * iters * NL * X * 32 * 4;
    printf("%-28s warps %2d : %8llu cycles, %7.1f B/cycle/SM  (%s)\n", name, warps, h[0], bytes / (double)h[0], cudaGetErrorString(e));
    cudaFree(d);
    cudaFree(sink);
}

int main() {
    for (int w : {4, 8, 12, 16}) run<32, 1>(w, "x32, 1 load per wait");
    for (int w : {4, 8, 16}) run<32, 2>(w, "x32, 2 loads per wait");
    for (int w : {4, 8, 16}) run<8, 6>(w, "x8, 6 loads per wait");
    for (int w : {4, 8, 16}) run<8, 1>(w, "x8, 1 load per wait");
    return 0;
}
