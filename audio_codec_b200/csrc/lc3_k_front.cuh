// spectral front-end kernel, instantiated per frame length in lc3_k_front_*.cu (split for parallel builds)
#pragma once
#include <cuda_runtime.h>
#include "lc3_body_frame.cuh"

using namespace lc3k;

template <int N, int BITS> __global__ void __launch_bounds__(128) k_front(LC3K_Batch b)
{
    const int warp = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    front_body<N, BITS>(b, warp / b.n_tiles, warp % b.n_tiles, lane);
}

template <int N> static int front_launch(const LC3K_Batch *b, cudaStream_t s)
{
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + 3) / 4;
    if (b->bits == 16) k_front<N, 16><<<blocks, 128, 0, s>>>(*b);
    else k_front<N, 24><<<blocks, 128, 0, s>>>(*b);
    return (int)cudaGetLastError();
}
