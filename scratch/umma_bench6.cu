// cta_group::2 issue rate: a CTA pair (cluster of 2) runs M = 256 UMMAs, each CTA supplies its 128 rows of A and HALF of
// B from its own shared memory.  Question for k_gn_block (N = 64, shared-memory operand bound at 48 cycles per UMMA in
// cta_group::1): does the pair get max(N/2, (4 KB + 16 * N B) / 128) = 40 cycles?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scratch/umma_bench6 scratch/umma_bench6.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../alphazero-gomoku_b200/csrc/gz_tcgen05.cuh"
using namespace gz;

__device__ __forceinline__ void umma2_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma1_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit2_elect(uint64_t *bar) {   // arrives on the leader's barrier only
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}" ::"r"(ptx::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void commit1_elect(uint64_t *bar) {
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}" ::"r"(ptx::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cta_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

template <int PAIR>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
k_bench(int N, int n_mma, long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const uint32_t rank = cta_rank();
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3C003C00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    ptx::fence_proxy_async();
    if (warp == 0) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(ptx::smem_u32(&slot)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            ptx::tmem_alloc<512>(&slot);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t idesc = ptx::make_idesc_bf16(PAIR ? 256 : 128, N);
    const uint32_t a0 = ptx::smem_u32(smem), b0 = ptx::smem_u32(smem) + 96 * 1024;
    const bool issuer = warp == 0 && (!PAIR || rank == 0);
    if (issuer) {
        const uint64_t da = ptx::make_desc(a0, 4640, 128), db = ptx::make_desc(b0, N * 16, 128);
        long long t0 = clock64();
#pragma unroll 8
        for (int j = 0; j < n_mma; ++j) {
            if (PAIR) umma2_elect(tm + (j & 1) * 256, da + (uint64_t)((j & 7) * 16), db + (uint64_t)((j & 3) * 2 * N), idesc, 1);
            else umma1_elect(tm + (j & 1) * 256, da + (uint64_t)((j & 7) * 16), db + (uint64_t)((j & 3) * 2 * N), idesc, 1);
        }
        if (PAIR) commit2_elect(&bar); else commit1_elect(&bar);
        ptx::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) cycles[0] = t1 - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (warp == 0) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
        else ptx::tmem_dealloc<512>(tm);
    }
}

template <int PAIR> void run(const char *name, long long *d) {
    cudaFuncSetAttribute(k_bench<PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    for (int N : {64, 128, 256}) {
        k_bench<PAIR><<<148, 128, 160 * 1024>>>(N, 4096, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long cyc = 0; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
        printf("%-28s N=%3d: %7.1f cycles/UMMA (%s)\n", name, N, cyc / 4096.0, cudaGetErrorString(e));
    }
}
int main() {
    long long *d; cudaMalloc(&d, 8);
    run<0>("cta_group::1, M=128", d);
    run<1>("cta_group::2, M=256 (pair)", d);
    return 0;
}
