// sm_100a kernels of the batched LC3plus encoder, part: bandwidth detector + SNS of the tile path (lc3_body_frame.cuh,
// sns_body): one lane per channel-frame, band energies in, SNS gains and indices out -- no spectrum access.
#define LC3_SAT16_PTX 1
#include <cuda_runtime.h>
#include "lc3_body_frame.cuh"

using namespace lc3k;

// Large blocks on purpose: the body is ~15 000 instructions of straight-line code that every warp runs once, so the
// kernel is bound by instruction fetch unless the warps of an SM walk through it together and share the fetches.
#ifndef SNS_WARPS
#define SNS_WARPS 12 /* measured: 24 x 1 block 0.73 ms, 12 x 2 blocks 0.67, 32 x 1 0.67, 16 x 1 (128 registers) 0.85, 8 x 3 0.77 */
#endif
#ifndef SNS_BLOCKS
#define SNS_BLOCKS 2
#endif
__global__ void __launch_bounds__(32 * SNS_WARPS, SNS_BLOCKS) k_sns(LC3K_Batch b)
{
    const int warp = blockIdx.x * SNS_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    LC3_DBG_ENTER(0, warp);
    sns_body(b, warp / b.n_tiles, warp % b.n_tiles, lane);
    LC3_DBG_LEAVE();
}
LC3_DBG_READER(sns)

extern "C" int lc3k_sns_launch(const LC3K_Batch *b, void *stream)
{
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + SNS_WARPS - 1) / SNS_WARPS;
    k_sns<<<blocks, 32 * SNS_WARPS, 0, (cudaStream_t)stream>>>(*b);
    return (int)cudaGetLastError();
}
