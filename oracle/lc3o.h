/* ORACLE (test infrastructure, not product code): public interface of the plain-C restatement of
 * the ETSI LC3plus fixed-point encoder.  Only tests/, __graft_entry__.smoke() and the cpu_baseline /
 * --impl reference legs of bench.py may use anything under oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle_*.py check every stage and the whole encoder
 * bit-for-bit against the unmodified reference compiled from /root/reference (oracle/_ref, built by
 * oracle/Makefile) and against the reference's own known-answer vectors
 * (ETSI testvec md5_bin.txt EP0 lines; committed bitstream digests in tests/golden/). */
#ifndef LC3O_H
#define LC3O_H
#include <stdint.h>

enum { /* same numbering as LC3_Error, fx/lc3.h:102-126 */
    LC3O_OK = 0, LC3O_ERROR = 1, LC3O_DECODE_ERROR = 2, LC3O_NULL_ERROR = 3, LC3O_SAMPLERATE_ERROR = 4, LC3O_CHANNELS_ERROR = 5,
    LC3O_BITRATE_ERROR = 6, LC3O_FRAMEMS_ERROR = 10, LC3O_BITRATE_UNSET_ERROR = 14, LC3O_BITRATE_SET_ERROR = 15,
    LC3O_BW_WARNING = 18
};

typedef struct lc3o_enc lc3o_enc;

/* per-frame intermediates of the LAST encoded channel, for stage-level parity tests */
typedef struct {
    int32_t mdct[480];  int16_t mdct_exp;     /* after processMdct_fx */
    int32_t ener[64];   int16_t ener_exp;     /* after processPerBandEnergy_fx */
    int32_t spec[480];                         /* after shaping + TNS */
    int16_t xq[480];
    int32_t scf_idx[7];
    int16_t bw, tns_bits, ltpf_bits, qgain, lastnz, lsbMode, fac_ns, nBits, attack, pitch;
} lc3o_trace;

int lc3o_enc_size(void);
int lc3o_enc_init(lc3o_enc *e, int samplerate, int channels);
lc3o_enc *lc3o_enc_new(int samplerate, int channels);
void lc3o_enc_free(lc3o_enc *e);
int lc3o_enc_set_frame_dms(lc3o_enc *e, int dms);
int lc3o_enc_set_bitrate(lc3o_enc *e, int bitrate);
int lc3o_enc_set_bandwidth(lc3o_enc *e, int bandwidth);
int lc3o_enc_set_ep_mode(lc3o_enc *e, int epmode);          /* 0..4, fx/lc3.h:288 */
int lc3o_enc_set_ep_mode_request(lc3o_enc *e, int epmr);    /* 0..3, fx/lc3.h:300 */
int lc3o_enc_get_input_samples(const lc3o_enc *e);
int lc3o_enc_get_num_bytes(const lc3o_enc *e);
int lc3o_enc_get_delay(const lc3o_enc *e);
int lc3o_enc_get_real_bitrate(const lc3o_enc *e);
int lc3o_enc16(lc3o_enc *e, int16_t **in, uint8_t *out, int *num_bytes);
int lc3o_enc24(lc3o_enc *e, int32_t **in, uint8_t *out, int *num_bytes);
int lc3o_enc16_trace(lc3o_enc *e, int16_t **in, uint8_t *out, int *num_bytes, lc3o_trace *tr);
int lc3o_encode_stream16(lc3o_enc *e, const int16_t *pcm, long ch_stride, int n_frames, uint8_t *out);

/* ---- decoder for good frames (oracle/lc3o_dec.c; fx/dec_lc3.c without concealment and channel decoding).
 * Parity status: PINNED against the compiled reference decoder and the committed digests of its output
 * (tests/test_oracle_dec.py, tests/golden/dec_golden.json). */
typedef struct lc3o_dec lc3o_dec;
lc3o_dec *lc3o_dec_new(int samplerate, int channels);
void lc3o_dec_free(lc3o_dec *d);
int lc3o_dec_set_frame_dms(lc3o_dec *d, int dms);
int lc3o_dec_get_output_samples(const lc3o_dec *d);
int lc3o_dec_get_delay(const lc3o_dec *d);
/* LC3O_OK, or LC3O_DECODE_ERROR when the payload is not a valid frame (zero spectrum synthesised) */
int lc3o_dec16(lc3o_dec *d, const uint8_t *in, int num_bytes, int16_t **out);
int lc3o_dec24(lc3o_dec *d, const uint8_t *in, int num_bytes, int32_t **out);
#endif
