// Issue-side variants: (0) lane-0 branch (as in the engine), (1) whole warp runs the loop, the mma is predicated
// inside the asm by elect.sync, (2) as 1 with the descriptors advanced by adds only.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../alphazero-gomoku_b200/csrc/gz_tcgen05.cuh"
using namespace gz;

__device__ __forceinline__ void umma_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit_elect(uint64_t *bar) {
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}" ::"r"(ptx::smem_u32(bar)) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(128, 1) k_bench(int N, int lbo_a, int lbo_b, int a_step, int b_step, int n_mma,
                                                  int per_commit, long long *cycles, int a_off) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3C003C00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    ptx::fence_proxy_async();
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t idesc = ptx::make_idesc_bf16(128, N);
    const uint32_t a0 = ptx::smem_u32(smem) + a_off, b0 = ptx::smem_u32(smem) + 100 * 1024;
    if (MODE == 0) {
        if (warp == 0 && lane == 0) {
            long long t0 = clock64();
            uint32_t phase = 0;
            for (int i = 0; i < n_mma; i += per_commit) {
                for (int j = 0; j < per_commit; ++j) {
                    const int m = i + j;
                    ptx::umma_bf16(tm + (m & 1) * 256, ptx::make_desc(a0 + (m & 7) * a_step, lbo_a, 128),
                                   ptx::make_desc(b0 + (m & 3) * b_step, lbo_b, 128), idesc, 1);
                }
                ptx::umma_commit(&bar);
                ptx::mbar_wait(&bar, phase);
                phase ^= 1;
            }
            long long t1 = clock64();
            if (blockIdx.x == 0) cycles[0] = t1 - t0;
        }
    } else if (warp == 0) {
        long long t0 = clock64();
        uint32_t phase = 0;
        for (int i = 0; i < n_mma; i += per_commit) {
            if (MODE == 1) {
                for (int j = 0; j < per_commit; ++j) {
                    const int m = i + j;
                    umma_elect(tm + (m & 1) * 256, ptx::make_desc(a0 + (m & 7) * a_step, lbo_a, 128),
                               ptx::make_desc(b0 + (m & 3) * b_step, lbo_b, 128), idesc, 1);
                }
            } else {
                const uint64_t da = ptx::make_desc(a0, lbo_a, 128), db = ptx::make_desc(b0, lbo_b, 128);
#pragma unroll 8
                for (int j = 0; j < per_commit; ++j)
                    umma_elect(tm + (j & 1) * 256, da + (uint64_t)((j & 7) * (a_step >> 4)),
                               db + (uint64_t)((j & 3) * (b_step >> 4)), idesc, 1);
            }
            commit_elect(&bar);
            ptx::mbar_wait(&bar, phase);
            phase ^= 1;
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) cycles[0] = t1 - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}


int main() {
    long long *d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(k_bench<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int N : {64, 128, 192, 240, 256})
        for (int a_off : {0, 16, 32, 64, 112})
            for (int lbo : {4640, 4096}) {
                k_bench<1><<<148, 128, 200 * 1024>>>(N, lbo, N * 16, 256, N * 32, 4096, 4096, d, a_off);
                cudaError_t e = cudaDeviceSynchronize();
                long long cyc = 0; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
                printf("N=%3d A offset %3d B lboA %d: %7.1f cycles/UMMA (%s)\n", N, a_off, lbo, cyc / 4096.0, cudaGetErrorString(e));
            }
    return 0;
}
