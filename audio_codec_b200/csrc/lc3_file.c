/* Bitstream containers of the reference command-line tool, as library calls.
 *
 *   .lc3plus binary format   ETSI/Readme.txt:308-333, written by fx/codec_exe.c:726-751 (header) and
 *                            :829-843 (frames: uint16 byte count, then the bytes)
 *   G.192 format             fx/codec_exe.c:799-827 (sync word 0x6B21, uint16 bit count, one int16 per bit,
 *                            LSB of each byte first, 0x0081 = one, 0x007F = zero; the 18-byte header goes to
 *                            a separate .cfg file, fx/codec_exe.c:735-740)
 *   switching files          fx/codec_exe.c:968-977 (a flat file of little-endian int64 values, read one per
 *                            frame and restarted from the top when it runs out)
 *
 * All multi-byte fields are little endian whatever the host is (the reference fwrite()s host words and
 * documents the file as little endian).  Pure host code: no device, no encoder state.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/lc3plus_b200.h"

static void put16(uint8_t *p, unsigned v) { p[0] = (uint8_t)(v & 0xff); p[1] = (uint8_t)((v >> 8) & 0xff); }

size_t lc3plus_container_header(void *dst, int samplerate, int bitrate, int channels, int frame_dms, int epmode,
                                uint32_t signal_len)
{
    uint8_t *p = (uint8_t *)dst;
    if (!p) return LC3PLUS_CONTAINER_HEADER_BYTES;
    put16(p + 0, 0xcc1c);
    put16(p + 2, LC3PLUS_CONTAINER_HEADER_BYTES);
    put16(p + 4, (unsigned)samplerate / 100);
    put16(p + 6, (unsigned)(bitrate / 100)); /* the reference stores bitrate/100 in a uint16 and lets it wrap */
    put16(p + 8, (unsigned)channels);
    put16(p + 10, (unsigned)frame_dms * 10); /* (uint16_t)(frame_ms * 100) */
    put16(p + 12, (unsigned)epmode);
    put16(p + 14, signal_len & 0xffff);
    put16(p + 16, signal_len >> 16);
    return LC3PLUS_CONTAINER_HEADER_BYTES;
}

static int frame_size(const int *num_bytes, int f, int bytes_per_frame)
{
    return num_bytes ? num_bytes[f] : bytes_per_frame;
}

size_t lc3plus_container_frames(void *dst, const void *frames, int n_frames, size_t frame_stride_bytes,
                                int bytes_per_frame, const int *num_bytes, int g192)
{
    uint8_t *p = (uint8_t *)dst;
    const uint8_t *src = (const uint8_t *)frames;
    size_t n = 0;
    if (n_frames < 0 || (!num_bytes && bytes_per_frame < 0)) return 0;
    for (int f = 0; f < n_frames; f++) {
        int sz = frame_size(num_bytes, f, bytes_per_frame);
        if (sz < 0 || sz > 0xffff / 8 * 8) return 0;
        if (!p) { n += g192 ? 4 + (size_t)sz * 16 : 2 + (size_t)sz; continue; }
        if (!src) return 0;
        const uint8_t *b = src + (size_t)f * frame_stride_bytes;
        if (g192) {
            put16(p + n, 0x6B21);
            put16(p + n + 2, (unsigned)sz * 8);
            n += 4;
            for (int i = 0; i < sz; i++)
                for (int k = 0; k < 8; k++, n += 2) put16(p + n, (b[i] >> k) & 1 ? 0x0081 : 0x007F);
        } else {
            put16(p + n, (unsigned)sz);
            memcpy(p + n + 2, b, (size_t)sz);
            n += 2 + (size_t)sz;
        }
    }
    return n;
}

static int write_all(const char *path, const void *a, size_t na, const void *b, size_t nb)
{
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    int bad = (na && fwrite(a, 1, na, f) != na) || (nb && fwrite(b, 1, nb, f) != nb);
    if (fclose(f)) bad = 1;
    return bad ? -1 : 0;
}

LC3_Error lc3plus_write_bitstream_file(const char *path, const char *cfg_path, int g192, int samplerate, int bitrate,
                                       int channels, int frame_dms, int epmode, uint32_t signal_len,
                                       const void *frames, int n_frames, size_t frame_stride_bytes,
                                       int bytes_per_frame, const int *num_bytes)
{
    uint8_t hdr[LC3PLUS_CONTAINER_HEADER_BYTES];
    if (!path || (n_frames > 0 && !frames)) return LC3_NULL_ERROR;
    lc3plus_container_header(hdr, samplerate, bitrate, channels, frame_dms, epmode, signal_len);
    size_t need = lc3plus_container_frames(NULL, NULL, n_frames, frame_stride_bytes, bytes_per_frame, num_bytes, g192);
    if (n_frames > 0 && need == 0) return LC3_NUMBYTES_ERROR;
    uint8_t *body = (uint8_t *)malloc(need ? need : 1);
    if (!body) return LC3_ERROR;
    lc3plus_container_frames(body, frames, n_frames, frame_stride_bytes, bytes_per_frame, num_bytes, g192);
    int rc;
    if (g192) {
        /* header in its own file: cfg_path, or path with ".cfg" appended (fopen_with_ext, fx/codec_exe.c:720-723) */
        char *auto_cfg = NULL;
        if (!cfg_path) {
            size_t L = strlen(path);
            auto_cfg = (char *)malloc(L + 5);
            if (!auto_cfg) { free(body); return LC3_ERROR; }
            memcpy(auto_cfg, path, L);
            memcpy(auto_cfg + L, ".cfg", 5);
            cfg_path = auto_cfg;
        }
        rc = write_all(cfg_path, hdr, sizeof hdr, NULL, 0);
        if (!rc) rc = write_all(path, body, need, NULL, 0);
        free(auto_cfg);
    } else {
        rc = write_all(path, hdr, sizeof hdr, body, need);
    }
    free(body);
    return rc ? LC3_ERROR : LC3_OK;
}

int lc3plus_read_switching_file(const char *path, int64_t *values, int max_values)
{
    FILE *f = path ? fopen(path, "rb") : NULL;
    if (!f) return -1;
    int n = 0;
    uint8_t b[8];
    while (fread(b, 1, 8, f) == 8) {
        if (values && n < max_values) {
            uint64_t v = 0;
            for (int k = 7; k >= 0; k--) v = (v << 8) | b[k];
            values[n] = (int64_t)v;
        }
        n++;
    }
    fclose(f);
    return n;
}
