// dispatch of the spectral front end by frame length (kernels in lc3_k_front_*.cu)
#include <cuda_runtime.h>
#include "lc3_layout.h"
extern "C" {
int lc3k_front_480(const LC3K_Batch *b, void *stream);
int lc3k_front_320(const LC3K_Batch *b, void *stream);
int lc3k_front_240(const LC3K_Batch *b, void *stream);
int lc3k_front_160(const LC3K_Batch *b, void *stream);
int lc3k_front_120(const LC3K_Batch *b, void *stream);
int lc3k_front_80(const LC3K_Batch *b, void *stream);
int lc3k_front_60(const LC3K_Batch *b, void *stream);
int lc3k_front_40(const LC3K_Batch *b, void *stream);
int lc3k_front_20(const LC3K_Batch *b, void *stream);
int lc3k_tns_480(const LC3K_Batch *b, void *stream);
int lc3k_tns_320(const LC3K_Batch *b, void *stream);
int lc3k_tns_240(const LC3K_Batch *b, void *stream);
int lc3k_tns_160(const LC3K_Batch *b, void *stream);
int lc3k_sns_launch(const LC3K_Batch *b, void *stream);
// N >= 160: tile MDCT + band energies; shorter frames: the whole front end in one lane-per-frame kernel
int lc3k_launch_front(const LC3K_Batch *b, void *stream)
{
    switch (b->N) {
    case 480: return lc3k_front_480(b, stream);
    case 320: return lc3k_front_320(b, stream);
    case 240: return lc3k_front_240(b, stream);
    case 160: return lc3k_front_160(b, stream);
    case 120: return lc3k_front_120(b, stream);
    case 80: return lc3k_front_80(b, stream);
    case 60: return lc3k_front_60(b, stream);
    case 40: return lc3k_front_40(b, stream);
    case 20: return lc3k_front_20(b, stream);
    default: return (int)cudaErrorInvalidValue;
    }
}
// second and third kernel of the front end for N >= 160 (no-ops for the short frames, whose lc3k_launch_front did it all)
int lc3k_launch_sns(const LC3K_Batch *b, void *stream) { return b->N >= 160 ? lc3k_sns_launch(b, stream) : 0; }
int lc3k_launch_tns(const LC3K_Batch *b, void *stream)
{
    switch (b->N) {
    case 480: return lc3k_tns_480(b, stream);
    case 320: return lc3k_tns_320(b, stream);
    case 240: return lc3k_tns_240(b, stream);
    case 160: return lc3k_tns_160(b, stream);
    default: return 0;
    }
}
}
