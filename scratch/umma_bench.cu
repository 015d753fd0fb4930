// Microbenchmark: issue rate of tcgen05.mma (M=128, K=16, bf16, SS mode, K-major SWIZZLE_NONE) as a
// function of N, operand strides and alignment.  One CTA per SM, operands fixed in shared memory.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../alphazero-gomoku_b200/csrc/gz_tcgen05.cuh"
using namespace gz;

__global__ void __launch_bounds__(128, 1) k_bench(int N, int lbo_a, int lbo_b, int a_off, int a_step, int b_step,
                                                  int n_mma, int per_commit, long long *cycles, int M, int sbo, unsigned long long lt) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3C003C00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    ptx::fence_proxy_async();
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    if (warp == 0 && lane == 0) {
        const uint32_t idesc = ptx::make_idesc_bf16(M, N);
        const uint32_t a0 = ptx::smem_u32(smem) + a_off, b0 = ptx::smem_u32(smem) + 100 * 1024;
        long long t0 = clock64();
        uint32_t phase = 0;
        for (int i = 0; i < n_mma; i += per_commit) {
            for (int j = 0; j < per_commit; ++j) {
                const int m = i + j;
                ptx::umma_bf16(tm + (m & 1) * 256, (ptx::make_desc(a0 + (m % 9) * a_step, lbo_a, sbo) | (lt << 61)),
                               (ptx::make_desc(b0 + (m % 4) * b_step, lbo_b, sbo) | (lt << 61)), idesc, 1);
            }
            ptx::umma_commit(&bar);
            ptx::mbar_wait(&bar, phase);
            phase ^= 1;
        }
        long long t1 = clock64();
        if (blockIdx.x == 0) cycles[0] = t1 - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}


int main() {
    long long *d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(k_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct Cfg { int M, N, lbo_a, lbo_b, a_step, b_step, sbo; unsigned long long lt; const char *name; };
    Cfg cfgs[] = {
        {128, 64, 2048, 1024, 0, 0, 128, 0, "M128 N64  none"},
        {128, 256, 2048, 4096, 0, 0, 128, 0, "M128 N256 none"},
        {128, 240, 2048, 4096, 0, 0, 128, 0, "M128 N240 none"},
        {128, 208, 2048, 4096, 0, 0, 128, 0, "M128 N208 none"},
        {64, 64, 2048, 1024, 0, 0, 128, 0, "M64  N64  none"},
        {64, 128, 2048, 2048, 0, 0, 128, 0, "M64  N128 none"},
        {64, 256, 2048, 4096, 0, 0, 128, 0, "M64  N256 none"},
        {128, 64, 16, 16, 32, 32, 1024, 2, "M128 N64  sw128"},
        {128, 128, 16, 16, 32, 32, 1024, 2, "M128 N128 sw128"},
        {128, 256, 16, 16, 32, 32, 1024, 2, "M128 N256 sw128"},
        {64, 256, 16, 16, 32, 32, 1024, 2, "M64  N256 sw128"},
        {128, 64, 16, 16, 32, 32, 512, 4, "M128 N64  sw64"},
        {128, 128, 16, 16, 32, 32, 512, 4, "M128 N128 sw64"},
        {128, 64, 16, 16, 0, 0, 256, 6, "M128 N64  sw32"},
        {128, 128, 16, 16, 0, 0, 256, 6, "M128 N128 sw32"},
    };
    for (auto &c : cfgs) {
        k_bench<<<148, 128, 200 * 1024>>>(c.N, c.lbo_a, c.lbo_b, 0, c.a_step, c.b_step, 4096, 4096, d, c.M, c.sbo, c.lt);
        cudaError_t e = cudaDeviceSynchronize();
        long long cyc = 0; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
        printf("%-28s: %7.1f cycles/UMMA (%s)\n", c.name, cyc / 4096.0, cudaGetErrorString(e));
    }
    return 0;
}
