/* liblc3plus_b200 -- C ABI of the B200-native batched LC3plus encoder.
 *
 * Drop-in for the ENCODER half of the ETSI reference API (fx/lc3.h, file:line cited per function below;
 * fx/ = LC3plus_ETSI_src_v17171_20200723/src/fixed_point/ of the reference repository), plus one new entry
 * point, lc3plus_enc_batch, that encodes many independent streams and many frames per stream in one call.
 * Every function is exported under the reference's `lc3_` name and under the `lc3plus_` prefix used by
 * later ETSI releases; signatures are identical.  Output bytes are identical to the reference encoder's.
 *
 * Differences a caller must know:
 *   - The encoder block the caller allocates (lc3_enc_get_size bytes) holds configuration and a handle;
 *     the per-channel signal state lives in device memory owned by the library.  Call lc3plus_enc_release
 *     when an encoder is no longer needed (the reference has no such call because it owns no memory).
 *   - `scratch` must not be NULL (the reference returns LC3_NULL_ERROR, fx/lc3.c:250, and so does this library); it is not used.
 *   - One encoder, one bit depth: after lc3_enc16 frames, lc3_enc24 on the same encoder (or the reverse) returns LC3_ERROR
 *     instead of carrying the MDCT overlap across the switch as the reference does (fx/enc_lc3.c:35-83); re-initialise it.
 *   - There is no CPU fallback: every encode call needs a CUDA device and returns LC3_ERROR if the
 *     device path fails.
 *   - Error protection (lc3_enc_set_ep_mode, modes 1-4, and lc3_enc_set_ep_mode_request) is implemented on the
 *     device as in the reference: payload size reduced to fec_get_data_size, partial-concealment block moved to the
 *     front (modes 3/4, slots of 80 bytes and more), Reed-Solomon coded slot of lc3_enc_get_num_bytes bytes.
 */
#ifndef LC3PLUS_B200_H
#define LC3PLUS_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LC3_VERSION_INT(major, minor, micro) (((major) << 16) | ((minor) << 8) | (micro))
#define LC3_VERSION LC3_VERSION_INT(1, 4, 10)           /* fx/lc3.h:39-42 */
#define LC3_MAX_CHANNELS 2                               /* fx/lc3.h:46 */
#define LC3_MAX_SAMPLES 480                              /* fx/lc3.h:49 */
#define LC3_MAX_BYTES 870                                /* fx/lc3.h:52 */
#define LC3_ENC_MAX_SIZE 7064                            /* fx/lc3.h:55 */
#define LC3_ENC_MAX_SCRATCH_SIZE 6784                    /* fx/lc3.h:61 */

typedef enum { /* fx/lc3.h:102-126 */
    LC3_OK = 0, LC3_ERROR = 1, LC3_DECODE_ERROR = 2, LC3_NULL_ERROR = 3, LC3_SAMPLERATE_ERROR = 4,
    LC3_CHANNELS_ERROR = 5, LC3_BITRATE_ERROR = 6, LC3_NUMBYTES_ERROR = 7, LC3_PLCMODE_ERROR = 8,
    LC3_EPMODE_ERROR = 9, LC3_FRAMEMS_ERROR = 10, LC3_ALIGN_ERROR = 11, LC3_EPMR_ERROR = 12, LC3_HRMODE_ERROR = 13,
    LC3_BITRATE_UNSET_ERROR = 14, LC3_BITRATE_SET_ERROR = 15, LC3_BFI_EXT_ERROR = 16,
    LC3_WARNING = 17, LC3_BW_WARNING = 18
} LC3_Error;

typedef enum { LC3_EP_OFF = 0, LC3_EP_ZERO = 1, LC3_EP_LOW = 2, LC3_EP_MEDIUM = 3, LC3_EP_HIGH = 4 } LC3_EpMode; /* fx/lc3.h:74-82 */
typedef enum { LC3_EPMR_ZERO = 0, LC3_EPMR_LOW = 1, LC3_EPMR_MEDIUM = 2, LC3_EPMR_HIGH = 3 } LC3_EpModeRequest; /* fx/lc3.h:85-99 */

typedef struct LC3_Enc LC3_Enc;                          /* fx/lc3.h:128 */

int lc3_version(void);                                   /* fx/lc3.h:135 */
int lc3_channels_supported(int channels);                /* fx/lc3.h:142 */
int lc3_samplerate_supported(int samplerate);            /* fx/lc3.h:149 */
LC3_Error lc3_enc_init(LC3_Enc *encoder, int samplerate, int channels);                                /* fx/lc3.h:170 */
LC3_Error lc3_enc16(LC3_Enc *encoder, int16_t **input_samples, void *output_bytes, int *num_bytes, void *scratch); /* fx/lc3.h:191 */
LC3_Error lc3_enc24(LC3_Enc *encoder, int32_t **input_samples, void *output_bytes, int *num_bytes, void *scratch); /* fx/lc3.h:198 */
int lc3_enc_get_size(int samplerate, int channels);      /* fx/lc3.h:207 */
int lc3_enc_get_scratch_size(const LC3_Enc *encoder);    /* fx/lc3.h:216 */
int lc3_enc_get_input_samples(const LC3_Enc *encoder);   /* fx/lc3.h:223 */
int lc3_enc_get_real_bitrate(const LC3_Enc *encoder);    /* fx/lc3.h:231 */
int lc3_enc_get_num_bytes(const LC3_Enc *encoder);       /* fx/lc3.h:239 */
LC3_Error lc3_enc_set_bitrate(LC3_Enc *encoder, int bitrate);                                           /* fx/lc3.h:258 */
int lc3_enc_get_delay(const LC3_Enc *encoder);           /* fx/lc3.h:266 */
LC3_Error lc3_enc_set_frame_dms(LC3_Enc *encoder, int frame_ms);                                        /* fx/lc3.h:277 */
LC3_Error lc3_enc_set_ep_mode(LC3_Enc *encoder, LC3_EpMode epmode);                                     /* fx/lc3.h:288 */
LC3_Error lc3_enc_set_ep_mode_request(LC3_Enc *encoder, LC3_EpModeRequest epmr);                        /* fx/lc3.h:300 */
LC3_Error lc3_enc_set_bandwidth(LC3_Enc *encoder, int bandwidth);                                       /* fx/lc3.h:309 */

/* `lc3plus_` spellings of the same functions (north-star names) */
LC3_Error lc3plus_enc_init(LC3_Enc *encoder, int samplerate, int channels);
LC3_Error lc3plus_enc16(LC3_Enc *encoder, int16_t **input_samples, void *output_bytes, int *num_bytes, void *scratch);
LC3_Error lc3plus_enc24(LC3_Enc *encoder, int32_t **input_samples, void *output_bytes, int *num_bytes, void *scratch);
LC3_Error lc3plus_enc_set_bitrate(LC3_Enc *encoder, int bitrate);
LC3_Error lc3plus_enc_set_frame_dms(LC3_Enc *encoder, int frame_ms);
LC3_Error lc3plus_enc_set_bandwidth(LC3_Enc *encoder, int bandwidth);
int lc3plus_enc_get_size(int samplerate, int channels);
int lc3plus_enc_get_input_samples(const LC3_Enc *encoder);
int lc3plus_enc_get_num_bytes(const LC3_Enc *encoder);
int lc3plus_enc_get_real_bitrate(const LC3_Enc *encoder);
int lc3plus_enc_get_delay(const LC3_Enc *encoder);

/* NEW (no reference counterpart; replaces the caller's loop fx/codec_exe.c:332-500 around lc3_enc16/24).
 * Encodes n_frames consecutive frames of each of n_streams encoders.  Equivalent, byte for byte and in
 * its effect on every encoder's state, to calling lc3_enc16/lc3_enc24 n_frames times on each encoder.
 *
 *   pcm                 samples of stream s start at (char*)pcm + s*stream_stride_bytes; within a stream
 *                       the channels are planar: channel c occupies n_frames*N samples starting at sample
 *                       c*n_frames*N (N = lc3_enc_get_input_samples).  int16_t for bitdepth 16, int32_t
 *                       (24-bit, sign-extended) for bitdepth 24.
 *   out                 frame f of stream s is written at (uint8_t*)out + s*out_stream_stride_bytes +
 *                       f*lc3_enc_get_num_bytes(enc[s]); channel payloads are concatenated inside a frame
 *                       exactly as lc3_enc16 does (fx/enc_lc3.c:279-299).
 *   num_bytes           optional [n_streams*n_frames], stream-major.  On entry -1 marks "signal this frame as
 *                       lost" (the reference's in/out quirk, fx/lc3.c:253); on return the byte count.
 *   pcm / out may be host or device pointers (detected with cudaPointerGetAttributes); device pointers
 *   avoid the copies.  With host buffers and streams of unequal size in one call, the bytes of an output row past its own
 *   frames -- up to the length of the longest row of the batch, never past the row's stride -- are overwritten with
 *   unspecified values.  Encoders may differ in sample rate, frame size, bitrate and channel count; they are
 *   bucketed internally.  An encoder may appear only once per call.
 */
LC3_Error lc3plus_enc_batch(LC3_Enc *const *encoders, int n_streams, int n_frames, const void *pcm,
                            size_t stream_stride_bytes, int bitdepth, void *out, size_t out_stream_stride_bytes,
                            int *num_bytes);

/* NEW: the same call without the wait at its end.  It returns when the batch is queued; `out` and `num_bytes` are valid after
 * lc3plus_sync().  The caller's `pcm` (if in host memory) must stay untouched until then.  Consecutive calls on one device
 * form one continuing pipeline -- the next call's copies and first kernels overlap the previous call's last ones -- and at
 * most two calls per device are in flight: a third waits for the first to finish. */
LC3_Error lc3plus_enc_batch_async(LC3_Enc *const *encoders, int n_streams, int n_frames, const void *pcm,
                                  size_t stream_stride_bytes, int bitdepth, void *out, size_t out_stream_stride_bytes,
                                  int *num_bytes);
/* NEW: waits for everything queued by lc3plus_enc_batch_async on every device; LC3_ERROR if any of it failed. */
LC3_Error lc3plus_sync(void);
/* NEW ("one thread/stream-of-work per device"): selects the CUDA device on which the encoders that THIS
 * THREAD initialises from now on keep their state (-1: back to the default, the thread's current CUDA device).  The library
 * keeps one context per device.  lc3plus_enc_batch accepts encoders of several devices in one call when pcm and out are host
 * buffers: each device's share is queued by its own host thread and the call joins them.  With device buffers, make one
 * call per device (lc3plus_enc_batch_async lets one host thread keep several devices busy). */
LC3_Error lc3plus_set_device(int device);
/* NEW: the device an encoder lives on (-1 for an invalid handle). */
int lc3plus_enc_get_device(const LC3_Enc *encoder);

/* NEW -- no counterpart in the reference; added for callers that may not allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Encodes the first n_frames frames of n_streams FRESH mono streams (every stream starts from the power-on state of
 * lc3_enc_init; nothing is carried to the next call) that share one configuration, byte for byte what lc3_enc_init +
 * lc3_enc_set_frame_dms + lc3_enc_set_bitrate + n_frames x lc3_enc16/24 give.  The call allocates NOTHING (no cudaMalloc,
 * cudaFree, pinned or library allocation) and does no host<->device copy or synchronisation: it queues one memset and the
 * chain's kernels on `cuda_stream` (a cudaStream_t; NULL = the default stream) and returns; synchronise that stream before
 * reading `out`.  All scratch -- stream state, lane tables, the kernels' work buffers -- lives in one static 7 MiB device
 * array that every GPU loading the library holds until exit.  Limits: n_streams <= LC3PLUS_NOALLOC_MAX_STREAMS and
 * (n_streams rounded up to a multiple of 32) * n_frames <= LC3PLUS_NOALLOC_MAX_CHANNEL_FRAMES, else LC3_ERROR with the
 * reason in lc3plus_last_error().  pcm and out MUST be device pointers, laid out as for lc3plus_enc_batch (mono).
 * Every call on a GPU shares the array, which keeps the previous call's contents (the state part is cleared at the start
 * of each call): calls that overlap on the GPU -- different streams, or threads -- race; serialise them. */
#define LC3PLUS_NOALLOC_MAX_STREAMS 64
#define LC3PLUS_NOALLOC_MAX_CHANNEL_FRAMES 1024
LC3_Error lc3plus_enc_fresh_noalloc(int samplerate, int frame_dms, int bitrate, int bitdepth, int n_streams, int n_frames,
                                    const void *pcm, size_t stream_stride_bytes, void *out, size_t out_stream_stride_bytes,
                                    void *cuda_stream);

/* NEW: returns the device state of an encoder to the library's pool.  Optional: the library remembers which state belongs
 * to which encoder block by the block's address, so initialising a block again (or a new block at an address that was freed
 * without this call) hands the old state back first. */
void lc3plus_enc_release(LC3_Enc *encoder);
/* NEW: number of kernels launched by this process so far (for benchmarks' gpu_launches). */
long long lc3plus_kernel_launches(void);
/* NEW: per-kernel device timing with CUDA events on the library's launching stream.  After enable(1), every
 * batch call accumulates into ms[13] = {stage, fir, scalar, pitch, mdct (the whole front end for frames shorter than 160 samples), sns, tns, rate,
 * quant, back kernel ms (each bracketed by events
 * on the stream it was launched on; chunks overlap, so these can sum to more than the span), start-of-first-chunk to
 * end-of-last-chunk span ms, calls, chunks}; enable() also clears the accumulators. */
void lc3plus_profile_enable(int level); /* 0 off; 1 device span only (calls stay asynchronous); 2 per-kernel events too */
void lc3plus_profile_read(double *ms);                        /* summed over the devices in use */
void lc3plus_profile_read_device(int device, double *ms);    /* one device's share */
/* NEW: lc3_enc_set_bitrate(encoders[i], bitrates[i]) for i < n in one call (mid-stream switching of many streams between two
 * batches, as fx/codec_exe.c:340-349 does per frame from its switching file); returns the first error, applies the rest. */
LC3_Error lc3plus_enc_set_bitrate_many(LC3_Enc *const *encoders, int n, const int *bitrates);
/* NEW: last CUDA error string seen by the library ("" if none). */
const char *lc3plus_last_error(void);

/* ------------------------------------------------------------------------------------------------------------------
 * DECODER.  Drop-in for the decoder half of the reference API for ERROR-FREE streams: PCM identical,
 * sample for sample, to lc3_dec16 / lc3_dec24 of the reference (fx/lc3.h:334-370, fx/dec_lc3.c:15-487) as long as every
 * frame decodes.  NOT built: packet-loss and partial concealment (fx/plc_*.c, fx/pc_*.c) and the channel decoder
 * (fx/al_fec.c fec_decoder).  A frame the reference would conceal (bfi_ext != 0, num_bytes == 0, or a payload its bit-error
 * checks reject) is synthesised from a zero spectrum instead -- the stream's memories stay consistent -- and the call returns
 * LC3_DECODE_ERROR, as the reference does; from there on the PCM differs from the reference's until the memories have
 * run out (overlap: one frame; LTPF: while it stays active).  The plc_mode argument is checked and otherwise ignored. */
typedef struct LC3_Dec LC3_Dec;                          /* fx/lc3.h:129 */
typedef enum { LC3_PLC_STANDARD = 0, LC3_PLC_ADVANCED = 1 } LC3_PlcMode;                               /* fx/lc3.h:66-70 */
#define LC3_DEC_MAX_SIZE 256                             /* fx/lc3.h:58 holds the signal memories; here they live on the device */
LC3_Error lc3_dec_init(LC3_Dec *decoder, int samplerate, int channels, LC3_PlcMode plc_mode);          /* fx/lc3.h:334 */
LC3_Error lc3_dec16(LC3_Dec *decoder, void *input_bytes, int num_bytes, int16_t **output_samples, void *scratch, int bfi_ext); /* fx/lc3.h:362 */
LC3_Error lc3_dec24(LC3_Dec *decoder, void *input_bytes, int num_bytes, int32_t **output_samples, void *scratch, int bfi_ext); /* fx/lc3.h:370 */
int lc3_dec_get_size(int samplerate, int channels, LC3_PlcMode plc_mode);                              /* fx/lc3.h:381 */
int lc3_dec_get_scratch_size(const LC3_Dec *decoder);    /* fx/lc3.h:390 */
int lc3_dec_get_output_samples(const LC3_Dec *decoder);  /* fx/lc3.h:398 */
int lc3_dec_get_delay(const LC3_Dec *decoder);           /* fx/lc3.h:405 */
LC3_Error lc3_dec_set_frame_dms(LC3_Dec *decoder, int frame_ms);                                       /* fx/lc3.h:415 */
LC3_Error lc3_dec_set_ep_enabled(LC3_Dec *decoder, int ep_enabled); /* fx/lc3.h:425; LC3_EPMODE_ERROR for ep_enabled != 0 */
LC3_EpModeRequest lc3_dec_get_ep_mode_request(const LC3_Dec *decoder); /* fx/lc3.h:447; always LC3_EPMR_ZERO */
int lc3_dec_get_error_report(const LC3_Dec *decoder);    /* fx/lc3.h:457; always 0 */
int lc3_dec_get_epok_flags(const LC3_Dec *decoder);      /* fx/lc3.h:466; always 0 */
/* NEW: returns the decoder's device state to the pool.  Optional: re-initialising or abandoning the block leaks nothing. */
LC3_Error lc3_dec_free(LC3_Dec *decoder);
int lc3plus_dec_get_device(const LC3_Dec *decoder);

/* NEW (no reference counterpart; replaces the caller's loop fx/codec_exe.c:332-500 around lc3_dec16/24).  Decodes n_frames
 * consecutive frames of each of n_streams decoders: the same PCM and the same effect on every decoder's state as calling
 * lc3_dec16 / lc3_dec24 n_frames times on each decoder with bfi_ext = 0.
 *
 *   in_bytes            frame f of stream s is read at (uint8_t*)in_bytes + s*in_stream_stride_bytes + f*num_bytes[s];
 *                       channel payloads are concatenated inside a frame as lc3_enc16 writes them (fx/dec_lc3.c:455).
 *   num_bytes           [n_streams] bytes per frame of each stream (all frames of a stream in one call have one size; 0 marks
 *                       every frame of the stream as lost).  20..435 per channel (fx/defines.h:141-142), else LC3_NUMBYTES_ERROR.
 *   pcm                 samples of stream s start at (char*)pcm + s*pcm_stream_stride_bytes; channels planar: channel c
 *                       occupies n_frames*N samples from sample c*n_frames*N (N = lc3_dec_get_output_samples).  int16_t for
 *                       bitdepth 16, int32_t (24-bit) for bitdepth 24.
 *   bfi_flags           optional [n_streams*n_frames], stream-major: 1 where the frame did not decode.
 *   in_bytes / pcm may be host or device pointers (device pointers avoid the copies).  Decoders may differ in sample rate,
 *   frame duration, byte count and channel count (they are grouped internally; with host buffers and one configuration the
 *   copies of the chunks overlap the kernels) and, with host buffers, in device.  A decoder may appear only once.
 * Returns LC3_OK, or LC3_DECODE_ERROR when at least one frame did not decode (the PCM is still complete). */
LC3_Error lc3plus_dec_batch(LC3_Dec *const *decoders, int n_streams, int n_frames, const void *in_bytes, size_t in_stream_stride_bytes,
                            const int *num_bytes, void *pcm, size_t pcm_stream_stride_bytes, int bitdepth, uint8_t *bfi_flags);
void lc3plus_dec_set_chunk_frames(int channel_frames);   /* channel-frames per chunk of the decoder's passes (0: default 163840) */
long long lc3plus_dec_kernel_launches(void);
void lc3plus_dec_profile_enable(int on);                 /* CUDA-event timing */
void lc3plus_dec_profile_read(double *ms /* [4]: entropy, DCT, synthesis kernel of the first chunk of every call; device span of the calls */);
/* NEW: upper bound on channel-frames processed per internal chunk, taken literally when set here (defaults, which never cut
 * a chunk below eight frames per stream: 163840 when the PCM is already on the
 * device, 204800 when it is copied from host memory chunk by chunk); up to three chunks are in flight at once, each on
 * its own CUDA streams with its own work buffers. */
void lc3plus_set_chunk_frames(int channel_frames);

/* ---- Bitstream containers of the reference command-line tool.  Host-only helpers: they
 * never touch the device.  Little endian on every host. */
#define LC3PLUS_CONTAINER_HEADER_BYTES 18
/* NEW (replaces open_bitstream_writer's header, fx/codec_exe.c:726-751; layout ETSI/Readme.txt:308-325): writes
 * the 18-byte configuration header {0xcc1c, 18, samplerate/100, bitrate/100, channels, frame_ms*100, epmode,
 * signal_len lo, signal_len hi} to dst and returns 18; dst == NULL only returns the size.  frame_dms is the
 * frame duration in 0.1 ms units (25, 50, 100) as in lc3_enc_set_frame_dms. */
size_t lc3plus_container_header(void *dst, int samplerate, int bitrate, int channels, int frame_dms, int epmode,
                                uint32_t signal_len);
/* NEW (replaces write_bitstream_frame / write_bitstream_frame_G192, fx/codec_exe.c:799-843): packs n_frames coded
 * frames (frame f at (uint8_t*)frames + f*frame_stride_bytes, num_bytes[f] bytes, or bytes_per_frame when
 * num_bytes is NULL) into dst, each preceded by its uint16 byte count, or, with g192 != 0, as G.192 words (sync
 * 0x6B21, bit count, one 16-bit word per bit, LSB of each byte first).  Returns the bytes written; dst == NULL
 * only returns the size; 0 on a bad size. */
size_t lc3plus_container_frames(void *dst, const void *frames, int n_frames, size_t frame_stride_bytes,
                                int bytes_per_frame, const int *num_bytes, int g192);
/* NEW: header + frames into the file `path` exactly as the reference tool writes it (fx/codec_exe.c:255-260,
 * :418-437).  With g192 != 0 the header goes to cfg_path (NULL: path with ".cfg" appended,
 * fx/codec_exe.c:720-723) and `path` holds only the G.192 frames. */
LC3_Error lc3plus_write_bitstream_file(const char *path, const char *cfg_path, int g192, int samplerate, int bitrate,
                                       int channels, int frame_dms, int epmode, uint32_t signal_len,
                                       const void *frames, int n_frames, size_t frame_stride_bytes,
                                       int bytes_per_frame, const int *num_bytes);
/* NEW (replaces loopy_read64's file, fx/codec_exe.c:968-977): reads the int64 values of a bitrate / bandwidth /
 * epmode switching file.  Returns how many values the file holds (the first max_values are stored in values; values
 * may be NULL to count), -1 if it cannot be opened.  The reference takes one value per frame and wraps around:
 * frame f uses values[f % count] (per channel for bitrates, fx/codec_exe.c:342-347). */
int lc3plus_read_switching_file(const char *path, int64_t *values, int max_values);

#ifdef __cplusplus
}
#endif
#endif
