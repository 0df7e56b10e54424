/* Device data layout of the batched LC3plus DECODER (fx/dec_lc3.c without concealment).
 *
 * Same conventions as lc3_layout.h: everything is tile-interleaved (32 consecutive lanes form a tile, element e of a
 * per-lane array sits at base[(tile * n_elems + e) * 32 + lane]); persistent per-channel state (what fx/setup_dec_lc3.h keeps
 * in DecSetup for the error-free path) is addressed by SLOT in a library-owned pool, per-batch work buffers by batch lane.
 */
#ifndef LC3_DEC_LAYOUT_H
#define LC3_DEC_LAYOUT_H
#include <stdint.h>
#include "lc3_layout.h"

/* per-lane configuration of a call, derived from the frame's byte count (fx/setup_dec_lc3.c:271-336 update_dec_bitrate) */
enum { DCFG_BYTES = 0, DCFG_LPC_WEIGHTING, DCFG_QGAIN_OFF, DCFG_LTPF_SCALE_IDX, DCFG_N = 4 };

/* per-slot scalar state (int32 words) */
enum {
    DST_OLA_EXP = 0,       /* stDec_ola_mem_fx_exp */
    DST_LTPF_E,            /* ltpf_mem_e */
    DST_LTPF_PITCH_INT, DST_LTPF_PITCH_FR, DST_LTPF_GAIN, DST_LTPF_ACTIVE,
    DST_LTPF_SCALE_IDX,    /* ltpf_mem_scale_fac_idx, power-on value -1 (fx/setup_dec_lc3.c:109) */
    DST32_N = 8
};

/* per-slot int16 state arrays (element offsets) */
enum {
    DS16_OLA = 0,          /* 480: overlap-add memory (frame_length - la_zeroes used) */
    DS16_LTPF_X = 480,     /* 12: ltpf_mem_x (ltpf_mem_x_len <= 10 used) */
    DS16_LTPF_Y = 492,     /* 864: ltpf_mem_y (ltpf_mem_y_len <= 861 used) */
    DST16_N = 1356
};
#define LC3D_MAX_LTPF_X 10
#define LC3D_MAX_LTPF_Y 861

/* per channel-frame outputs of the frame-parallel passes (int32 words) */
enum {
    DO_BFI = 0,            /* 0: decoded; 1: the payload is not a valid frame (zero spectrum is synthesised instead: no concealment) */
    DO_Q_EXP,              /* exponent of the shaped spectrum */
    DO_Y_E, DO_Y_S,        /* after the DCT-IV: block exponent and the up-shift of the windowing loops (fx/imdct_fx.c:84-103) */
    DO_LTPF0, DO_LTPF1, DO_LTPF2, DO_PAD, DO_N = 8
};

typedef struct {
    /* bucket geometry (fx/setup_dec_lc3.c:100-106, :169-231) */
    int fs_idx, dms, dms_idx, N, yLen, nbands, W_off, W_len, la_zeroes, ola_len, bands_off, bw_bits, ltpf_x_len, ltpf_y_len, bits /* 16 | 24 */;
    int n_lanes, n_tiles, n_frames;
    int max_bytes;             /* largest DCFG_BYTES of the batch's lanes */
    /* state pool */
    const int32_t *slot;       /* [n_lanes] */
    int32_t *st32;             /* pool, DST32_N words per slot */
    int16_t *st16;             /* pool, DST16_N words per slot */
    /* per call */
    const int32_t *cfg;        /* [n_tiles][DCFG_N][32] */
    const uint8_t *in;         /* payloads (device address) */
    const int64_t *in_off;     /* [n_lanes] byte offset of each lane's first frame payload */
    const int32_t *in_stride;  /* [n_lanes] bytes between consecutive frames of the lane */
    void *pcm;                 /* int16 / int32 output samples (device address) */
    const int64_t *pcm_off;    /* [n_lanes] element offset of each lane's first output sample; frames follow each other */
    uint8_t *bfi;              /* [n_frames][n_lanes] per-frame flag DO_BFI, or NULL */
    int frame0;                /* index of this chunk's first frame within the caller's batch */
    /* work buffers */
    int16_t *xq;               /* [n_frames][n_tiles][N][32] decoded integer spectrum */
    int32_t *spec;             /* [n_frames][n_tiles][N][32] spectrum, then the DCT-IV output */
    int32_t *dout;             /* [n_frames][n_tiles][DO_N][32] */
    int32_t *yrow;             /* [n_frames][n_tiles*32][N] DCT-IV output, ROW-MAJOR: one contiguous row per channel-frame */
    int32_t *work;             /* [n_frames][n_tiles][N][32] FFT scratch of the lane-per-frame DCT (N < 160), else NULL */
} LC3D_Batch;

#ifdef __cplusplus
extern "C" {
#endif
/* kernel launchers (lc3_k_dec.cu); stream is a cudaStream_t.  Return 0 or a cudaError_t value. */
int lc3d_launch_entropy(const LC3D_Batch *b, void *stream); /* lane per channel-frame: side information, arithmetic decoder, residual bits, spectrum up to SNS shaping */
int lc3d_launch_idct(const LC3D_Batch *b, void *stream);    /* DCT-IV of the inverse MDCT: block per tile (N >= 160) or lane per channel-frame */
int lc3d_launch_synth(const LC3D_Batch *b, void *stream);   /* warp per channel-stream, frames in order: windowing + overlap-add, LTPF, output scaling */
int lc3d_launch_reset(int32_t *st32, int16_t *st16, const int32_t *slots, int n, void *stream); /* power-on state (fx/setup_dec_lc3.c:83-167) */
#ifdef __cplusplus
}
#endif
#endif
