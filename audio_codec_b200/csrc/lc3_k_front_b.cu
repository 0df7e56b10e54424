// spectral front end, frame lengths 320, 240
#include "lc3_k_front.cuh"

extern "C" int lc3k_front_320(const LC3K_Batch *b, void *stream) { return front_launch_tiled<320>(b, (cudaStream_t)stream); }
extern "C" int lc3k_front_240(const LC3K_Batch *b, void *stream) { return front_launch_tiled<240>(b, (cudaStream_t)stream); }
extern "C" int lc3k_tns_320(const LC3K_Batch *b, void *stream) { return tns_launch<320>(b, (cudaStream_t)stream); }
extern "C" int lc3k_tns_240(const LC3K_Batch *b, void *stream) { return tns_launch<240>(b, (cudaStream_t)stream); }
