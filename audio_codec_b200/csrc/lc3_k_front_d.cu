// spectral front end, frame lengths 60, 40, 20
#include "lc3_k_front.cuh"

extern "C" int lc3k_front_60(const LC3K_Batch *b, void *stream) { return front_launch<60>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_40(const LC3K_Batch *b, void *stream) { return front_launch<40>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_20(const LC3K_Batch *b, void *stream) { return front_launch<20>(b, (cudaStream_t)stream); }
