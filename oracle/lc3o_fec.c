/* ORACLE (test infrastructure, not product code).
 * Error protection of the encoder: restates fx/al_fec.c:203-358 (sizes), :361-646 (encoder: data word
 * pre-processing, Reed-Solomon code over GF(16), interleaver), :2177-2314 (CRCs) and
 * fx/reorder_bitstream_fx.c:16-56 (partial-concealment block moved to the front of the payload).
 * Arithmetic is on 4-bit symbols; GF(16) uses the field polynomial x^4 + x + 1 (fx/al_fec.c:72-76), products
 * are computed here instead of looked up (fx/rom_basop_util.c:1177).
 */
#include <assert.h>
#include <string.h>
#include "lc3o_internal.h"

#define CW_MAX 15 /* RS16_CW_LEN_MAX, fx/al_fec.c:22 */

static const uint8_t hd_mode0[5] = {1, 3, 3, 5, 7};      /* fx/al_fec.c:115 : hamming distance, first codeword */
static const uint8_t hd_mode1[5] = {1, 1, 3, 5, 7};      /* :116 : remaining codewords */
static const uint8_t crc1_bytes0[5] = {0, 3, 2, 2, 2};   /* :118 : slot size 40 */
static const uint8_t crc1_bytes1[5] = {0, 3, 3, 3, 3};   /* :119 */
static const uint8_t crc2_bytes[5] = {0, 0, 2, 2, 2};    /* :120 */
static const uint8_t gp_d3[] = {8, 6}, gp_d5[] = {7, 8, 12, 13}, gp_d7[] = {12, 10, 12, 3, 9, 7}; /* :90-92 */
static const uint8_t sig_polys[4][13] = {                /* :98-101, degree 12 */
    {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0},
    {7, 15, 5, 6, 14, 9, 1, 3, 12, 10, 13, 3, 2},
    {7, 11, 14, 1, 2, 3, 12, 11, 6, 15, 7, 6, 12},
    {6, 15, 12, 2, 9, 15, 2, 8, 12, 3, 10, 5, 4}};

static uint8_t gf16_mul(uint8_t a, uint8_t b)
{
    uint8_t r = 0;
    for (int i = 0; i < 4; i++) {
        if (b & 1) r ^= a;
        b >>= 1;
        a = (uint8_t)(a << 1);
        if (a & 16) a ^= 19;
    }
    return r;
}

/* the reference's mask tables (fx/al_fec.c:2177-2182, :2246-2247): entry k is the multiple of the generator
 * (degree = shift) whose bits shift..shift+3 equal k */
static u32 crc_mask(u32 gen, int shift, int k)
{
    u32 m = 0;
    for (int bit = 3; bit >= 0; bit--)
        if (((m >> (shift + bit)) ^ ((u32)k >> bit)) & 1) m ^= gen << bit;
    return m;
}

static int n_codewords(int slot_bytes) { return (2 * slot_bytes + CW_MAX - 1) / CW_MAX; }          /* :251-266 */
static int cw_length(int n_cw, int slot_nibbles, int i) { return (slot_nibbles - i - 1) / n_cw + 1; } /* :268-284 */

/* fx/al_fec.c:203-229 */
int o_fec_n_pccw(int slot_bytes, int mode, int ccc)
{
    int n = 0;
    if (mode == 3) n = l_round(l_sub(l_mult(2636, (i16)slot_bytes), 117377));
    else if (mode == 4) n = l_round(l_sub(l_mult(2178, (i16)slot_bytes), 129115));
    if (ccc == 1 || slot_bytes < 80) n = 0;
    return n;
}

/* fx/al_fec.c:288-327 */
int o_fec_data_size(int mode, int ccc, int slot_bytes)
{
    int n_cw = n_codewords(slot_bytes), p = slot_bytes;
    if (mode > 0) {
        p -= mode == 1 ? 1 : (mode - 1) * n_cw;
        p -= slot_bytes == 40 ? crc1_bytes0[mode] : crc1_bytes1[mode];
        if (ccc == 0 && mode > 2 && slot_bytes >= 80) p -= crc2_bytes[mode];
    }
    return p;
}

/* fx/al_fec.c:329-358 */
int o_fec_n_pc(int mode, int n_pccw, int slot_bytes)
{
    int n_cw = n_codewords(slot_bytes), pc = n_pccw * -2 * (mode - 1);
    if (mode == 1 || slot_bytes < 80) return 0;
    for (int i = 0; i < n_pccw; i++) pc += cw_length(n_cw, 2 * slot_bytes, n_cw - i - 1);
    return pc;
}

/* fx/al_fec.c:2184-2244 */
static void crc1(const uint8_t *data, int n, int epmr, uint8_t *hash, int hash_size)
{
    const int shift = hash_size == 2 ? 14 : 22;
    const u32 gen = hash_size == 2 ? 17989u : 4788009u;
    u32 rem = 0;
    for (int i = n - 1; i >= 0; i--) {
        rem = (rem << 4) ^ data[i];
        rem ^= crc_mask(gen, shift, (rem >> shift) & 15);
    }
    rem = (rem << 4) ^ ((u32)epmr << 2);
    rem ^= crc_mask(gen, shift, (rem >> shift) & 15);
    for (int i = 0; i < 2 * hash_size - 1; i++) {
        rem <<= 4;
        rem ^= crc_mask(gen, shift, (rem >> shift) & 15);
    }
    rem ^= (u32)epmr << shift;
    for (int i = 0; i < 2 * hash_size; i++) hash[i] = (uint8_t)((rem >> (4 * i)) & 15);
}

/* fx/al_fec.c:2256-2314 */
static void crc2(const uint8_t *data, int n, uint8_t *hash, int hash_size)
{
    const int shift = 16;
    const u32 gen = 107243u;
    u32 rem = 0;
    assert(hash_size == 2);
    for (int i = n - 1; i >= 0; i--) {
        rem = (rem << 4) ^ data[i];
        rem ^= crc_mask(gen, shift, (rem >> shift) & 15);
    }
    for (int i = 0; i < 2 * hash_size; i++) {
        rem <<= 4;
        rem ^= crc_mask(gen, shift, (rem >> shift) & 15);
    }
    for (int i = 0; i < 2 * hash_size; i++) hash[i] = (uint8_t)((rem >> (4 * i)) & 15);
}

/* fx/al_fec.c:558-612 : systematic RS encoding; buf holds (data polynomial) * x^(hd-1) */
static void rs16_enc(uint8_t *buf, int len, int hd, int mode, int signal)
{
    const uint8_t *gp = hd == 3 ? gp_d3 : (hd == 5 ? gp_d5 : gp_d7);
    const int deg = hd - 1;
    uint8_t sb[CW_MAX + 1];
    if (deg <= 0) return;
    memset(sb, 0, sizeof(sb));
    memset(buf, 0, (size_t)deg);
    memmove(sb + 1, buf + len - deg, (size_t)deg);
    for (int i = len - deg - 1; i >= 0; i--) {
        sb[0] = buf[i];
        const uint8_t lc = sb[deg];
        for (int j = deg - 1; j >= 0; j--) sb[j + 1] = sb[j] ^ gf16_mul(gp[j], lc);
    }
    for (int i = 0; i < deg; i++) buf[i] = sb[i + 1];
    if (signal)
        for (int i = 0; i <= 12; i++) buf[i] ^= sig_polys[mode - 1][i];
}

/* fx/al_fec.c:481-556 with fec_data_preproc (:424-479), dw0_bitswap (:361-390), fec_interleave_pack (:614-646).
 * In place: iobuf holds data_bytes of payload on entry and slot_bytes of channel-coded data on return. */
void o_fec_encode(int mode, int epmr, uint8_t *iobuf, int data_bytes, int slot_bytes, int n_pccw)
{
    uint8_t cw[2 * 300 + 16], out[300];
    const int n_cw = n_codewords(slot_bytes), nn = 2 * slot_bytes;
    const int pc_split = o_fec_n_pc(mode, n_pccw, slot_bytes);
    memset(cw, 0, sizeof(cw));

    /* data nibbles reversed into the top of the buffer */
    int off = 2 * (slot_bytes - data_bytes), j = nn - 1;
    for (int i = 0; i < data_bytes; i++) { cw[j--] = iobuf[i] & 15; cw[j--] = iobuf[i] >> 4; }
    const int n1 = slot_bytes == 40 ? crc1_bytes0[mode] : crc1_bytes1[mode];
    const int n2 = (pc_split > 0 && mode > 1) ? crc2_bytes[mode] : 0;
    if (n2) crc2(cw + off + 2 * data_bytes - pc_split, pc_split, cw + off - 2 * n2, n2);
    if (n1) crc1(cw + off, 2 * data_bytes - pc_split, epmr, cw + off - 2 * (n1 + n2), n1);
    off -= 2 * (n1 + n2);
    { /* swap the request bits with the bits that end up next to the first codeword's redundancy */
        uint8_t *dw0 = cw + off;
        const int pos = cw_length(n_cw, nn, 0) - 1, i0 = 2 * n1 - 1, i1 = pos - (hd_mode0[mode] - 1);
        const uint8_t t = (dw0[i0] >> 2) & 3;
        dw0[i0] = (uint8_t)((dw0[i0] & 3) | ((dw0[i1] & 3) << 2));
        dw0[i1] = (uint8_t)((dw0[i1] & 12) | t);
    }

    int cwo = 0, dwo = off;
    for (int i = 0; i < n_cw; i++) {
        const int hd = i == 0 ? hd_mode0[mode] : hd_mode1[mode], len = cw_length(n_cw, nn, i);
        for (int k = hd - 1; k < len; k++, dwo++) cw[cwo + k] = cw[dwo];
        rs16_enc(cw + cwo, len, hd, mode, i < 6);
        cwo += len;
    }
    assert(cwo == nn && dwo == nn);

    memset(out, 0, (size_t)slot_bytes);
    cwo = 0;
    for (int i = 0; i < n_cw; i++) {
        const int len = cw_length(n_cw, nn, i);
        for (int k = 0; k < len; k++, cwo++) {
            const int o = nn - (k * n_cw + i + 1);
            out[o >> 1] |= (uint8_t)(cw[cwo] << ((o & 1) << 2));
        }
    }
    memcpy(iobuf, out, (size_t)slot_bytes);
}

/* fx/reorder_bitstream_fx.c:16-56 */
void o_reorder(uint8_t *bytes, int n_pccw, int n_pc, int b_left)
{
    uint8_t tmp[870];
    if (n_pccw == 0) return;
    const int block = (n_pc + 1) >> 1;
    memcpy(tmp, bytes + b_left, (size_t)block);
    memcpy(tmp + block, bytes, (size_t)b_left);
    memcpy(bytes, tmp, (size_t)(block + b_left));
}
