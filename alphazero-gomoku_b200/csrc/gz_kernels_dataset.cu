// gz_kernels_dataset.cu -- training samples from packed replay records: GomokuSelfPlayDataset and
// augment_sample (training.py:104-134, 38-70), SURVEY.md 8f row 2.
//
// One warp per emitted sample.  72 bytes in (two bitboards, move, mover, outcome), 2712 bytes out (675 floats,
// an int64, a float): the kernel is HBM-write bound, stores are 128 contiguous bytes per warp instruction.
// Lane l < 16 keeps word l of the record's bitboards; a cell's source position under the inverse symmetry is
// computed once per cell and serves the three planes.
#include "gz_engine.h"
#include "gz_bitboard.cuh"

namespace gz {

// destination (i, j) of the planes after k counter-clockwise quarter turns and an optional flip of axis 2
// reads source (r, c):  np.flip: out[i][j] = y[i][14-j];  np.rot90 (k=1, axes=(1,2)): y[i][j] = x[j][14-i]
__device__ __forceinline__ void source_cell(int i, int j, int k, bool flip, int &r, int &c) {
    if (flip) j = 14 - j;
    for (int t = 0; t < k; ++t) { const int ni = j, nj = 14 - i; i = ni; j = nj; }
    r = i; c = j;
}
// _transform_index (training.py:53-61): (r, c) -> (c, 14 - r) per quarter turn, then c -> 14 - c
__device__ __forceinline__ int index_as_written(int idx, int k, bool flip) {
    int r = idx / 15, c = idx % 15;
    for (int t = 0; t < k; ++t) { const int nr = c, nc = 14 - r; r = nr; c = nc; }
    if (flip) c = 14 - c;
    return r * 15 + c;
}
// where the planes' transform really sends the stone at (r, c): (14 - c, r) per quarter turn, then the flip
__device__ __forceinline__ int index_with_planes(int idx, int k, bool flip) {
    int r = idx / 15, c = idx % 15;
    for (int t = 0; t < k; ++t) { const int nr = 14 - c, nc = r; r = nr; c = nc; }
    if (flip) c = 14 - c;
    return r * 15 + c;
}

__global__ void __launch_bounds__(128) k_dataset_build(const gz_replay_record *__restrict__ records, long long n,
                                                       const int32_t *__restrict__ chosen, long long m,
                                                       const long long *__restrict__ sample_ids, long long first,
                                                       long long count, int reference_exact,
                                                       float *__restrict__ planes, long long *__restrict__ move,
                                                       float *__restrict__ value) {
    const long long o = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // output slot
    const int lane = threadIdx.x & 31;
    if (o >= count) return;
    const long long id = sample_ids ? sample_ids[o] : first + o;
    long long rec_i = id;
    int k = 0;
    bool flip = false;
    if (id >= n) {
        const long long a = id - n;
        rec_i = chosen[a >> 3];
        k = (int)((a >> 1) & 3);
        flip = (a & 1) != 0;
    }
    const gz_replay_record &rec = records[rec_i];
    const uint32_t word = lane < 16 ? rec.stones[lane >> 3][lane & 7] : 0u;
    float *dst = planes + (size_t)o * 675;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        const int cell = lane + 32 * t;
        int r = 0, c = 0;
        if (cell < kCells) source_cell(cell / 15, cell % 15, k, flip, r, c);
        const int p = 16 * r + c;
        const uint32_t wb = __shfl_sync(0xffffffffu, word, p >> 5), ww = __shfl_sync(0xffffffffu, word, 8 + (p >> 5));
        const uint32_t bl = (wb >> (p & 31)) & 1u, wh = (ww >> (p & 31)) & 1u;
        if (cell < kCells) {
            dst[cell] = (float)bl;
            dst[kCells + cell] = (float)wh;
            dst[2 * kCells + cell] = (float)(1u - bl - wh);
        }
    }
    if (lane == 0) {
        move[o] = reference_exact ? index_as_written(rec.move, k, flip) : index_with_planes(rec.move, k, flip);
        value[o] = (float)rec.outcome;
    }
}

void launch_dataset_build(const gz_replay_record *records, long long n, const int32_t *chosen, long long m,
                          const long long *sample_ids, long long first, long long count, bool reference_exact,
                          float *planes, long long *move, float *value, cudaStream_t s) {
    if (count <= 0) return;
    const long long blocks = (count * 32 + 127) / 128;
    k_dataset_build<<<(unsigned)blocks, 128, 0, s>>>(records, n, chosen, m, sample_ids, first, count,
                                                     reference_exact ? 1 : 0, planes, move, value);
}

}  // namespace gz
