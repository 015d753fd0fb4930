// Where does row m of an M = 64 (cta_group::1) accumulator live in tensor memory?  A[m][0] = m + 1, B[0][0] = 1,
// everything else 0  =>  D[m][0] = m + 1.  All 128 lanes x columns 0..15 are read back and printed when non-zero.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "../alphazero-gomoku_b200/csrc/gz_tcgen05.cuh"
using namespace gz;

__global__ void __launch_bounds__(128, 1) k_probe(int M, float *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0u;
    __syncthreads();
    // A: K-major no-swizzle core matrices [kg 2][row 128][8]: element (m, k=0) at kg 0, row m, e 0
    __nv_bfloat16 *A = reinterpret_cast<__nv_bfloat16 *>(smem);
    for (int m = threadIdx.x; m < M; m += blockDim.x) A[m * 8] = __float2bfloat16((float)(m + 1));
    __nv_bfloat16 *B = reinterpret_cast<__nv_bfloat16 *>(smem + 32 * 1024);   // [kg 2][n 16][8]
    if (threadIdx.x == 0) { B[0] = __float2bfloat16(1.f); ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    ptx::fence_proxy_async();
    if (warp == 0) ptx::tmem_alloc<32>(&slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    if (warp == 0) {
        ptx::umma_bf16_elect(tm, ptx::make_desc(ptx::smem_u32(smem), 128 * 16, 128),
                             ptx::make_desc(ptx::smem_u32(smem) + 32 * 1024, 16 * 16, 128), ptx::make_idesc_bf16(M, 16), 0);
        ptx::umma_commit_elect(&bar);
    }
    ptx::mbar_wait(&bar, 0);
    ptx::tc_fence_after();
    float v[16];
    ptx::tmem_ld16(tm + ((uint32_t)(32 * warp) << 16), v);
    for (int j = 0; j < 16; ++j) out[(32 * warp + lane) * 16 + j] = v[j];
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<32>(tm);
}

int main() {
    float *d; cudaMalloc(&d, 128 * 16 * 4);
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int M : {128, 64}) {
        cudaMemset(d, 0, 128 * 16 * 4);
        k_probe<<<1, 128, 64 * 1024>>>(M, d);
        cudaError_t e = cudaDeviceSynchronize();
        static float h[128 * 16]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("M=%d (%s): ", M, cudaGetErrorString(e));
        for (int l = 0; l < 128; ++l) for (int j = 0; j < 16; ++j) if (h[l * 16 + j] != 0.f) printf("[lane %d col %d]=%g ", l, j, h[l * 16 + j]);
        printf("\n");
    }
    return 0;
}
