// spectral front end, frame lengths 480
#include "lc3_k_front.cuh"

extern "C" int lc3k_front_480(const LC3K_Batch *b, void *stream) { return front_launch_tiled<480>(b, (cudaStream_t)stream); }
LC3_DBG_READER(front)
extern "C" int lc3k_tns_480(const LC3K_Batch *b, void *stream) { return tns_launch<480>(b, (cudaStream_t)stream); }
