// spectral front end, frame lengths 60, 40, 20
#define LC3_SAT16_PTX 1 // the fused short-frame kernels hold the SNS codebook search (see sat16 in lc3_basop.cuh)
#include "lc3_k_front.cuh"

extern "C" int lc3k_front_60(const LC3K_Batch *b, void *stream) { return front_launch<60>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_40(const LC3K_Batch *b, void *stream) { return front_launch<40>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_20(const LC3K_Batch *b, void *stream) { return front_launch<20>(b, (cudaStream_t)stream); }
