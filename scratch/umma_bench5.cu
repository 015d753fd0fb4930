// Does epilogue traffic slow the tensor pipe?  Warp 0 issues M128 N64 K16 UMMAs back to back; R reader warps
// loop over tcgen05.ld (x16) of other TMEM columns, optionally also writing 16 B per lane to shared memory.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../alphazero-gomoku_b200/csrc/gz_tcgen05.cuh"
using namespace gz;

__global__ void __launch_bounds__(320, 1) k_bench(int N, int n_mma, int readers, int do_sts, int acc_mode, long long *cycles, float *sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3C003C00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); done = 0; }
    ptx::fence_proxy_async();
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    if (warp == 0) {
        const uint32_t idesc = ptx::make_idesc_bf16(128, N);
        const uint32_t a0 = ptx::smem_u32(smem), b0 = ptx::smem_u32(smem) + 100 * 1024;
        long long t0 = clock64();
        for (int m = 0; m < n_mma; ++m)
            ptx::umma_bf16_elect(tm + (acc_mode == 0 ? (m & 1) * 64 : acc_mode == 1 ? 0 : ((m >> 2) & 1) * 64), ptx::make_desc(a0 + (m & 7) * 256, 4640, 128),
                                 ptx::make_desc(b0 + (m & 3) * N * 32, N * 16, 128), idesc, 1);
        ptx::umma_commit_elect(&bar);
        ptx::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) cycles[0] = t1 - t0;
        done = 1;
    } else if (warp >= 2 && warp < 2 + readers) {
        float acc = 0.f;
        long long n = 0;
        while (!done) {
            float v[16];
            ptx::tmem_ld16(tm + ((uint32_t)(32 * (warp & 3)) << 16) + 256 + ((n & 15) * 16), v);
#pragma unroll
            for (int i = 0; i < 16; ++i) acc += v[i];
            if (do_sts) *reinterpret_cast<uint4 *>(smem + 150 * 1024 + ((warp * 32 + lane) * 16)) = make_uint4(__float_as_uint(v[0]), 0, 0, 0);
            ++n;
        }
        if (acc == 123.f) sink[0] = acc;
        if (blockIdx.x == 0 && lane == 0) cycles[warp] = n;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}

int main() {
    long long *d; cudaMalloc(&d, 128); float *sink; cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(k_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const char *names[] = {"alternate 2 accumulators", "one accumulator", "4 in a row then switch"};
    for (int N : {64, 128, 256})
        for (int mode : {0, 1, 2}) {
            cudaMemset(d, 0, 128);
            k_bench<<<148, 320, 200 * 1024>>>(N, 8192, 0, 0, mode, d, sink);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[16]; cudaMemcpy(h, d, 128, cudaMemcpyDeviceToHost);
            printf("N=%3d %-26s: %6.1f cycles/UMMA (%s)\n", N, names[mode], h[0] / 8192.0, cudaGetErrorString(e));
        }
    return 0;
}
