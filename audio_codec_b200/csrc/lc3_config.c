/* See lc3_config.h.  Reference: fx/setup_enc_lc3.c:38-281, fx/lc3.c:94-244. */
#include <string.h>
#include "lc3_config.h"

#define ROM_TAB(type, name, n) static const type name[n] __attribute__((unused))
#include "rom_tables.inc"
#undef ROM_TAB

enum { E_OK = 0, E_NULL = 3, E_SAMPLERATE = 4, E_CHANNELS = 5, E_BITRATE = 6, E_FRAMEMS = 10, E_BITRATE_SET = 15, E_BW_WARNING = 18 };

static int norm_s(int a)
{ /* norm_s of a 16-bit value */
    if (a == 0) return 0;
    if (a == -1) return 15;
    if (a < 0) a = ~a;
    int n = 0;
    while (a < 0x4000) { a <<= 1; n++; }
    return n;
}
static int div_l(int32_t num, int den)
{ /* div_l (basop32.c:2036+) for positive operands */
    int32_t d = (int32_t)den << 16;
    if (num >= d) return 32767;
    num >>= 1; d >>= 1;
    int q = 0;
    for (int k = 0; k < 15; k++) {
        q <<= 1; num <<= 1;
        if (num >= d) { num -= d; q++; }
    }
    return q;
}

static void frame_params(LC3Cfg *c)
{ /* fx/setup_enc_lc3.c:69-127 */
    int k = c->dms_idx * 5 + c->fs_idx;
    c->N = (c->fs / 100) >> c->dms_idx;
    c->bw_bits = rom_bw_cutoff_bits[c->fs_idx];
    c->la_zeroes = rom_win_laz[k];
    c->mdct_mem_len = c->N - c->la_zeroes;
    c->nsub = c->dms == 100 ? 3 : 2;
    c->ltpf_mem_in_len = c->dms == 25 ? 232 + 32 : 232;
    c->yLen = (400 >> c->dms_idx) < c->N ? (400 >> c->dms_idx) : c->N;
    c->W_off = rom_win_off[k];
    c->W_len = rom_win_len[k];
    c->nbands = rom_bands_number[k];
    c->bands_off = rom_bands_offset_off[k];
}

int lc3cfg_init(LC3Cfg *c, int samplerate, int channels)
{ /* fx/lc3.c:124-131, fx/setup_enc_lc3.c:38-66 */
    if (!c) return E_NULL;
    if (!(samplerate == 8000 || samplerate == 16000 || samplerate == 24000 || samplerate == 32000 ||
          samplerate == 44100 || samplerate == 48000)) return E_SAMPLERATE;
    if (channels < 1 || channels > 2) return E_CHANNELS;
    memset(c, 0, sizeof(*c));
    c->fs_in = samplerate;
    c->fs = samplerate == 44100 ? 48000 : samplerate;
    c->fs_idx = c->fs / 10000;
    c->channels = channels;
    c->dms = 100;
    c->dms_idx = 0;
    c->r12k8_mem_in_len = (int)((((int64_t)c->fs * 20972) >> 15) >> 9);
    frame_params(c);
    return E_OK;
}

int lc3cfg_set_frame_dms(LC3Cfg *c, int dms)
{ /* fx/lc3.c:215-223 */
    if (!c) return E_NULL;
    if (dms != 25 && dms != 50 && dms != 100) return E_FRAMEMS;
    if (c->br_set) return E_BITRATE_SET;
    c->dms = dms;
    c->dms_idx = dms == 100 ? 0 : (dms == 50 ? 1 : 2);
    frame_params(c);
    return E_OK;
}

/* ---- error-protection sizes (fx/al_fec.c:203-358) */
static int fec_n_codewords(int slot_bytes) { return (2 * slot_bytes + 14) / 15; }
static int fec_round16(int32_t v) { return (int)((v + 0x8000) >> 16); } /* round_fx of a value far from saturation */
static int fec_n_pccw(int slot_bytes, int mode, int ccc)
{ /* :203-229 */
    int n = 0;
    if (mode == 3) n = fec_round16(2 * 2636 * slot_bytes - 117377);
    else if (mode == 4) n = fec_round16(2 * 2178 * slot_bytes - 129115);
    return (ccc == 1 || slot_bytes < 80) ? 0 : n;
}
static int fec_data_size(int mode, int ccc, int slot_bytes)
{ /* :288-327 */
    int p = slot_bytes;
    if (mode > 0) {
        p -= mode == 1 ? 1 : (mode - 1) * fec_n_codewords(slot_bytes);
        p -= (slot_bytes == 40 && mode > 1) ? 2 : 3;
        if (ccc == 0 && mode > 2 && slot_bytes >= 80) p -= 2;
    }
    return p;
}
static int fec_n_pc(int mode, int n_pccw, int slot_bytes)
{ /* :329-358 */
    if (mode == 1 || slot_bytes < 80) return 0;
    int n_cw = fec_n_codewords(slot_bytes), pc = n_pccw * -2 * (mode - 1);
    for (int i = 0; i < n_pccw; i++) pc += (2 * slot_bytes - (n_cw - i - 1) - 1) / n_cw + 1;
    return pc;
}

int lc3cfg_set_bitrate(LC3Cfg *c, int bitrate)
{ /* fx/lc3.c:186-191, fx/setup_enc_lc3.c:130-281 */
    if (!c) return E_NULL;
    if (bitrate <= 0) return E_BITRATE;
    int minBR = (int)(20 * 8.0 * (10000.0 / c->dms) * (c->fs_in == 44100 ? 441. / 480 : 1));
    int maxBR = (int)(400 * 8.0 * (10000.0 / c->dms) * (c->fs_in == 44100 ? 441. / 480 : 1));
    minBR *= c->channels;
    maxBR *= c->channels;
    if (bitrate < minBR || bitrate > maxBR) return E_BITRATE;
    c->bitrate = bitrate;
    c->br_set = 1;
    const int slot_total = (int)((int32_t)bitrate * c->N / (8 * c->fs_in));
    c->ccc = c->channels > 1 && c->epmode && bitrate * c->N / (8 * c->fs_in) <= 160;
    if (c->epmode > 0) {
        int mb = bitrate * c->N / (8 * c->fs_in * c->channels);
        if (mb < 40 || mb > 300) return E_BITRATE; /* FEC_SLOT_BYTES_MIN / _MAX, fx/defines.h:146-147 */
    }
    int total = c->ccc ? fec_data_size(c->epmode, 1, slot_total) : slot_total;
    for (int ch = 0; ch < c->channels; ch++) {
        int32_t *w = c->ch[ch];
        const int channel_bytes = total / c->channels + (ch < (total % c->channels));
        int tb = channel_bytes;
        w[CFG_EP] = c->epmode ? (c->epmode | c->epmr << 4 | c->ccc << 8) : 0;
        w[CFG_SLOT_BYTES] = 0; w[CFG_EP_DATA_BYTES] = 0; w[CFG_N_PCCW] = 0; w[CFG_N_PC] = 0;
        if (c->epmode && c->ccc) { /* one slot for the whole frame, coded from channel 0's lane; no partial concealment */
            if (ch == 0) { w[CFG_SLOT_BYTES] = slot_total; w[CFG_EP_DATA_BYTES] = total; }
        } else if (c->epmode) {
            tb = fec_data_size(c->epmode, 0, channel_bytes);
            w[CFG_SLOT_BYTES] = channel_bytes;
            w[CFG_EP_DATA_BYTES] = tb;
            w[CFG_N_PCCW] = fec_n_pccw(channel_bytes, c->epmode, 0);
            w[CFG_N_PC] = fec_n_pc(c->epmode, w[CFG_N_PCCW], channel_bytes);
        }
        if (tb < 20 || tb > 435) return E_BITRATE;
        int bits = tb * 8;
        int ti = bits - (38 + 8 + 3 + c->bw_bits) - (17 - norm_s(c->yLen - 1));
        if (bits > 1280) ti--;
        if (bits > 2560) ti--;
        int thr = 560 + 80 * c->fs_idx, ltpf, lpcw;
        if (c->dms == 25) { ltpf = (int16_t)((9830 * bits) >> 12) - thr < 0; lpcw = 0; }
        else if (c->dms == 50) { ltpf = bits * 2 - 160 - thr < 0; lpcw = bits < 240; }
        else { lpcw = bits < 480; ltpf = bits - thr < 0; }
        int q = bits / (10 * (c->fs_idx + 1));
        int attack = c->dms == 100 && ((c->fs_in >= 44100 && tb >= 100) || (c->fs_in == 32000 && tb >= 81)) && tb < 340;
        w[CFG_TARGET_BYTES] = tb;
        w[CFG_TOTAL_BITS] = bits;
        w[CFG_TBITS_INIT] = ti;
        w[CFG_LPC_WEIGHTING] = lpcw;
        w[CFG_LTPF_ENABLE] = ltpf;
        w[CFG_QGAIN_OFF] = -((q < 115 ? q : 115) + 105 + 5 * (c->fs_idx + 1));
        w[CFG_ATTACK] = attack;
        w[CFG_BW_CUTOFF_BIN] = c->bandwidth ? c->bw_cutoff_bin : 0;
        w[CFG_BW_INDEX] = c->bw_index;
        if (!attack) c->attack_reset[ch] = 1; /* the detector state is cleared whenever it is switched off */
    }
    return E_OK;
}

int lc3cfg_set_bandwidth(LC3Cfg *c, int bandwidth)
{ /* fx/lc3.c:226-244 */
    if (!c) return E_NULL;
    if (c->bandwidth != bandwidth) {
        int eff = c->fs_in > 40000 ? 40000 : c->fs_in;
        if (bandwidth * 2 > eff) return E_BW_WARNING;
        c->bandwidth = bandwidth;
        c->bw_cutoff_bin = div_l((int32_t)(int16_t)bandwidth * (int16_t)c->dms, 2500);
        int bi = div_l(bandwidth, 2000) - 1;
        c->bw_index = bi < 0 ? 0 : bi;
        for (int ch = 0; ch < c->channels; ch++) {
            c->ch[ch][CFG_BW_CUTOFF_BIN] = bandwidth ? c->bw_cutoff_bin : 0;
            c->ch[ch][CFG_BW_INDEX] = c->bw_index;
        }
    }
    return E_OK;
}

int lc3cfg_set_ep_mode(LC3Cfg *c, int epmode)
{ /* fx/lc3.c:199-205 */
    if (!c) return E_NULL;
    if ((unsigned)epmode > 4) return 9; /* LC3_EPMODE_ERROR */
    c->epmode = epmode;
    return c->br_set ? lc3cfg_set_bitrate(c, c->bitrate) : E_OK;
}

int lc3cfg_set_ep_mode_request(LC3Cfg *c, int epmr)
{ /* fx/lc3.c:207-213 */
    if (!c) return E_NULL;
    if ((unsigned)epmr > 3) return 12; /* LC3_EPMR_ERROR */
    c->epmr = epmr;
    for (int ch = 0; ch < c->channels; ch++)
        if (c->ch[ch][CFG_EP]) c->ch[ch][CFG_EP] = (c->ch[ch][CFG_EP] & ~0xf0) | epmr << 4;
    return E_OK;
}

/* fx/enc_lc3.c:279-299: a channel's slot with error protection per channel, else its payload */
int lc3cfg_out_bytes(const LC3Cfg *c, int ch)
{
    return (c->epmode && !c->ccc) ? c->ch[ch][CFG_SLOT_BYTES] : c->ch[ch][CFG_TARGET_BYTES];
}

int lc3cfg_num_bytes(const LC3Cfg *c) { return c ? (int)((int32_t)c->bitrate * c->N / (8 * c->fs_in)) : 0; }

int lc3cfg_real_bitrate(const LC3Cfg *c)
{ /* fx/lc3.c:163-184 */
    if (!c) return 0;
    if (!c->br_set) return 14;
    int tb = 0;
    for (int ch = 0; ch < c->channels; ch++) tb += c->ch[ch][CFG_TARGET_BYTES];
    int br = (int)((tb * 80000.0 + c->dms - 1) / c->dms);
    if (c->fs_in == 44100) { int rem = br % 480; br = ((br - rem) / 480) * 441 + (rem * 441) / 480; }
    return br;
}

void lc3cfg_fill_batch(const LC3Cfg *c, LC3K_Batch *b)
{
    b->fs_idx = c->fs_idx; b->dms = c->dms; b->dms_idx = c->dms_idx; b->N = c->N; b->yLen = c->yLen;
    b->nbands = c->nbands; b->fs = c->fs; b->W_off = c->W_off; b->W_len = c->W_len; b->mdct_mem_len = c->mdct_mem_len;
    b->bands_off = c->bands_off; b->nsub = c->nsub; b->ltpf_mem_in_len = c->ltpf_mem_in_len;
    b->r12k8_mem_in_len = c->r12k8_mem_in_len; b->bw_bits = c->bw_bits;
}
