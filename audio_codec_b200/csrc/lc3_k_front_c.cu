// spectral front end, frame lengths 160, 120, 80
#define LC3_SAT16_PTX 1 // the fused short-frame kernels hold the SNS codebook search (see sat16 in lc3_basop.cuh)
#include "lc3_k_front.cuh"

extern "C" int lc3k_front_160(const LC3K_Batch *b, void *stream) { return front_launch_tiled<160>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_120(const LC3K_Batch *b, void *stream) { return front_launch<120>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_80(const LC3K_Batch *b, void *stream) { return front_launch<80>(b, (cudaStream_t)stream); }
extern "C" int lc3k_tns_160(const LC3K_Batch *b, void *stream) { return tns_launch<160>(b, (cudaStream_t)stream); }
