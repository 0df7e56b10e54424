/* Device data layout shared by the host C layer (lc3_api.c) and the CUDA kernels (lc3_kernels.cu).
 *
 * Everything the kernels touch is "tile-interleaved": 32 consecutive lanes (one warp) form a tile and
 * element e of a per-lane array lives at  base[(tile * n_elems + e) * 32 + lane].  A warp reading the
 * same element index of its 32 lanes therefore issues one fully coalesced 128-byte (int32) or 64-byte
 * (int16) request.  A lane is one channel-stream (lane-per-stream kernels) or one channel-frame
 * (lane-per-frame kernels; frame-major tiles).
 *
 * Persistent per-channel state (what fx/setup_enc_lc3.h:17-57 keeps in EncSetup) is addressed by SLOT
 * in a library-owned pool; per-batch work buffers are addressed by batch lane.
 */
#ifndef LC3_LAYOUT_H
#define LC3_LAYOUT_H
#include <stdint.h>

#define LC3K_LANES 32
#define LC3K_MAX_N 480

/* per-slot configuration derived from sample rate / frame size / bitrate (int32 words) */
enum {
    CFG_TARGET_BYTES = 0, CFG_TOTAL_BITS, CFG_TBITS_INIT, CFG_LPC_WEIGHTING, CFG_LTPF_ENABLE, CFG_QGAIN_OFF,
    CFG_ATTACK, CFG_BW_CUTOFF_BIN /* 0: no limit */, CFG_BW_INDEX,
    /* error protection (fx/setup_enc_lc3.c:155-206): CFG_EP = epmode | epmr << 4 | combined channel coding << 8 (0: off);
     * CFG_SLOT_BYTES = bytes of the channel-coded slot this lane's payload goes into (combined channel coding: the whole
     * frame on channel 0, 0 on channel 1); CFG_N_PCCW / CFG_N_PC = partial-concealment codewords / nibbles */
    CFG_EP, CFG_SLOT_BYTES, CFG_EP_DATA_BYTES /* payload bytes that go into the slot */, CFG_N_PCCW, CFG_N_PC,
    CFG_ATT_GEN /* moves on whenever set_bitrate switched the attack detector off: the pitch pass clears the detector's state when
                   it differs from ST_ATT_GEN (fx/setup_enc_lc3.c:266-275) */, CFG_N = 16
};

/* per-slot scalar state (int32 words) */
enum {
    ST_MEM50_0 = 0, ST_MEM50_1, ST_OLPA_EXP, ST_OLPA_PITCH, ST_LTPF_EXP, ST_LTPF_NC, ST_LTPF_MNC, ST_LTPF_ON,
    ST_LTPF_PITCH, ST_ATT_F0, ST_ATT_F1, ST_ATT_DET, ST_ATT_POS, ST_ATT_SCAL, ST_ATT_ACC, ST_TBITS_OFF,
    ST_MEM_TBITS, ST_MEM_SPECBITS, ST_X_EXP, ST_RESAMP_EXP, ST_ATT_GEN, ST32_N = 24
};

/* per-slot int16 state arrays (element offsets) */
enum {
    S16_MEM_OUT = 0,              /* 24: resampler output delay line */
    S16_OLPA_S12 = 24,            /* 3 (+1 pad) */
    S16_OLPA_BUF = 28,            /* 130: 6.4 kHz history */
    S16_LTPF_BUF = 158,           /* 264: 12.8 kHz history */
    ST16_N = 422
};

/* per channel-frame outputs of the per-stream time-domain pass (int32 words) */
enum { PO_LTPF_BITS = 0, PO_LTPF_IDX0, PO_LTPF_IDX1, PO_LTPF_IDX2, PO_ATTACK, PO_X_EXP, PO_N = 8 };

/* per channel-frame outputs of the spectral front end (int32 words) */
enum {
    FO_D_EXP = 0, FO_BW, FO_TNS_BITS, FO_TNS_NFILT, FO_TNS_ORDER0, FO_TNS_ORDER1, FO_X_MAX, FO_PAD,
    FO_TNS_IDX = 8,  /* 16 */
    FO_SCF = 24,     /* 7 */
    FO_N = 32
};

/* per channel-frame band data (tile MDCT path): 64 band energies + their exponent, later overwritten with the SNS gains
 * (mantissa in the low, exponent in the high half-word) */
enum { BEN_EXP = 64, BEN_N = 66 };

/* per channel-frame outputs of the rate-loop pass */
/* (RO_QGAIN: gain index of the first pass, replaced by the adjusted one by the final quantisation, which adds the last four) */
enum { RO_QGAIN = 0, RO_QGAIN_MIN, RO_NBITS1, RO_TBITS_QUANT, RO_LASTNZ, RO_LSBMODE, RO_NBITS2, RO_N = 8 };

typedef struct {
    /* bucket geometry (fx/setup_enc_lc3.c:69-127) */
    int fs_idx, dms, dms_idx, N, yLen, nbands, fs, W_off, W_len, mdct_mem_len, bands_off, nsub, ltpf_mem_in_len,
        r12k8_mem_in_len, bw_bits, bits /* 16 | 24 */;
    int n_lanes, n_tiles, n_frames;
    int max_bytes;            /* largest CFG_TARGET_BYTES of the bucket's lanes (sizes k_back's shared-memory frame buffers) */
    int frame0;               /* index of this chunk's first frame within the caller's batch */
    int stage_full;           /* 1: k_stage copies the chunk's PCM into pcm_t (24-bit input, frames shorter than 160 samples);
                                 0: history frame only -- the tile MDCT pulls its PCM rows from the caller's buffer with bulk copies */
    /* state pool */
    const int32_t *slot;      /* [n_lanes] */
    const int32_t *cfg;       /* pool, CFG_N words per slot */
    int32_t *st32;            /* pool, ST32_N words per slot */
    int16_t *st16;            /* pool, ST16_N words per slot */
    int16_t *hist16;          /* pool, LC3K_MAX_N samples per slot: last input frame (16-bit path) */
    int32_t *hist32;          /* pool, LC3K_MAX_N samples per slot: last input frame (24-bit path) */
    /* caller's buffers (device addresses) */
    const void *pcm;          /* int16 / int32 samples */
    const int64_t *pcm_off;   /* [n_lanes] element offset of each lane's first sample */
    uint8_t *out;
    const int64_t *out_off;   /* [n_lanes] byte offset of each lane's first frame payload */
    const int32_t *out_stride; /* [n_lanes] bytes between consecutive frames of the lane */
    const uint8_t *bfi;       /* [n_frames][n_lanes] "mark frame as lost" flags, or NULL */
    /* work buffers */
    void *pcm_t;              /* [n_tiles][N + n_frames*N][32] staged samples, history first */
    int32_t *pout;            /* [n_frames][n_tiles][PO_N][32] */
    int32_t *spec;            /* [n_frames][n_tiles][N][32] */
    int32_t *fout;            /* [n_frames][n_tiles][FO_N][32] */
    int32_t *en4;             /* [n_frames][n_tiles][N/4][32] */
    int32_t *rout;            /* [n_frames][n_tiles][RO_N][32] */
    int16_t *r12;             /* [n_frames][n_tiles][128][32] 12.8 kHz signal: FIR output, then high-passed in place */
    int32_t *ben;             /* [n_frames][n_tiles][BEN_N][32] band energies / SNS gains (tile MDCT path) */
    int16_t *xq;              /* [n_frames][n_tiles][N][32] quantised spectrum of the final pass */
} LC3K_Batch;

#ifdef __cplusplus
extern "C" {
#endif
/* kernel launchers (lc3_kernels.cu); stream is a cudaStream_t.  Return 0 or a cudaError_t value. */
int lc3k_launch_stage(const LC3K_Batch *b, void *stream);
int lc3k_launch_fir(const LC3K_Batch *b, void *stream);     /* 16-bit input only (no-op otherwise): resampler FIR, frame-parallel */
int lc3k_launch_scalar(const LC3K_Batch *b, void *stream);  /* 16-bit input only (no-op otherwise): 50 Hz high-pass recursion, lane per stream */
int lc3k_launch_pitch(const LC3K_Batch *b, void *stream);   /* pitch analysis + attack detector, warp per stream (24-bit: also the two above) */
int lc3k_launch_front(const LC3K_Batch *b, void *stream);   /* N >= 160: tile MDCT + band energies; N < 160: the whole spectral front end */
int lc3k_launch_sns(const LC3K_Batch *b, void *stream);     /* N >= 160 only (no-op otherwise): bandwidth detector + SNS */
int lc3k_launch_tns(const LC3K_Batch *b, void *stream);     /* N >= 160 only (no-op otherwise): shaping, TNS, quadruple energies */
int lc3k_launch_rate(const LC3K_Batch *b, void *stream);
int lc3k_launch_quant(const LC3K_Batch *b, void *stream);   /* gain adjustment + final quantisation, block per tile */
int lc3k_launch_back(const LC3K_Batch *b, void *stream);
int lc3k_launch_fec(const LC3K_Batch *b, void *stream);     /* error protection of the finished payloads, in place in the output (only when a lane has CFG_EP) */
/* set the listed state slots to their power-on values (fx/setup_enc_lc3.c:38-66) */
int lc3k_launch_reset(int32_t *st32, int16_t *st16, int16_t *hist16, int32_t *hist32, const int32_t *slots, int n, void *stream);
#ifdef __cplusplus
}
#endif
#endif
