// Error protection of one coded slot, one slot per lane (fx/al_fec.c:424-646 fec_data_preproc / fec_encoder / rs16_enc /
// fec_interleave_pack, :2177-2314 crc1 / crc2): the payload nibbles are laid out in reverse, extended by one or two CRC
// hashes and the mode request, cut into n codewords of 13..15 nibbles, each protected by a shortened Reed-Solomon code
// over GF(16) with Hamming distance 1, 3, 5 or 7 (the first six also carry the mode's signalling polynomial), and
// interleaved nibble by nibble into the slot.  Everything is 4-bit work on a few hundred nibbles; the buffers live in
// the lane's local memory (the protected path is optional and not the throughput case).
#pragma once
#include "lc3_basop.cuh"

namespace lc3k {

constexpr int FEC_CW_MAX = 15;          // RS16_CW_LEN_MAX, fx/al_fec.c:22
constexpr int FEC_SLOT_MAX = 300;       // FEC_SLOT_BYTES_MAX, fx/defines.h:147

LC3_HD int fec_hd(int mode, int first) { return mode == 0 ? 1 : (mode == 1 ? (first ? 3 : 1) : 2 * mode - 1); } // fx/al_fec.c:115-116
LC3_HD int fec_crc1_bytes(int mode, int slot_bytes) { return mode == 0 ? 0 : ((slot_bytes == 40 && mode > 1) ? 2 : 3); } // :118-119
LC3_HD int fec_crc2_bytes(int mode) { return mode > 1 ? 2 : 0; }                                                         // :120
LC3_HD int fec_n_codewords(int slot_bytes) { return (2 * slot_bytes + FEC_CW_MAX - 1) / FEC_CW_MAX; }                    // :251-266
LC3_HD int fec_cw_length(int n_cw, int slot_nibbles, int i) { return (slot_nibbles - i - 1) / n_cw + 1; }                // :268-284

LC3_HD int fec_n_pc(int mode, int n_pccw, int slot_bytes)
{ // fx/al_fec.c:329-358
    if (mode == 1 || slot_bytes < 80) return 0;
    const int n_cw = fec_n_codewords(slot_bytes);
    int pc = n_pccw * -2 * (mode - 1);
    for (int i = 0; i < n_pccw; i++) pc += fec_cw_length(n_cw, 2 * slot_bytes, n_cw - i - 1);
    return pc;
}

// Look-up tables the kernel builds once per block in shared memory (the reference looks these up too, fx/rom_basop_util.c:1177,
// fx/al_fec.c:2177-2182): all 256 GF(16) products, and for each of the three CRC generators the multiple that cancels a given
// top nibble.  Lane-varying indices, so shared memory, not constant memory.
struct FecTabs { u8 mul[256]; u32 crc[3][16]; };

LC3_HD u32 gf16_mul_bits(u32 a, u32 b)
{ // GF(16), field polynomial x^4 + x + 1 (fx/al_fec.c:72-76)
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        r ^= (b >> i) & 1 ? a : 0;
        a <<= 1;
        a ^= a & 16 ? 19 : 0;
    }
    return r;
}

// one step of the reference's table-driven CRC (mask[k], fx/al_fec.c:2177-2182, :2246-2247, is the multiple of the generator
// whose bits shift..shift+3 equal k): cancel the four bits above `shift`, top one first
LC3_HD u32 fec_crc_reduce_bits(u32 rem, u32 gen, int shift)
{
#pragma unroll
    for (int bit = 3; bit >= 0; bit--) rem ^= (rem >> (shift + bit)) & 1 ? gen << bit : 0;
    return rem;
}
// generator and degree of: crc1 with a two-byte hash, crc1 with three bytes, crc2
LC3_HD u32 fec_gen(int g) { return g == 0 ? 17989u : (g == 1 ? 4788009u : 107243u); }
LC3_HD int fec_shift(int g) { return g == 0 ? 14 : (g == 1 ? 22 : 16); }
// entry i of the tables (i < 256 + 48), for a block-wide fill
LC3_HD void fec_tabs_fill(FecTabs &T, int i)
{
    if (i < 256) T.mul[i] = (u8)gf16_mul_bits((u32)i >> 4, (u32)i & 15);
    else if (i < 256 + 48) {
        const int g = (i - 256) >> 4, k = (i - 256) & 15;
        T.crc[g][k] = fec_crc_reduce_bits((u32)k << fec_shift(g), fec_gen(g), fec_shift(g)) ^ ((u32)k << fec_shift(g));
    }
}
LC3_HD u32 gf16_mul(const FecTabs &T, u32 a, u32 b) { return T.mul[(a << 4) | b]; }
LC3_HD u32 fec_crc_reduce(const FecTabs &T, u32 rem, int g) { return rem ^ T.crc[g][(rem >> fec_shift(g)) & 15]; }

LC3_HD void fec_crc1(const FecTabs &T, const u8 *data, int n, int epmr, u8 *hash, int hash_size)
{ // fx/al_fec.c:2184-2244
    const int g = hash_size == 2 ? 0 : 1, shift = fec_shift(g);
    u32 rem = 0;
    for (int i = n - 1; i >= 0; i--) rem = fec_crc_reduce(T, (rem << 4) ^ data[i], g);
    rem = fec_crc_reduce(T, (rem << 4) ^ ((u32)epmr << 2), g);
    for (int i = 0; i < 2 * hash_size - 1; i++) rem = fec_crc_reduce(T, rem << 4, g);
    rem ^= (u32)epmr << shift;
    for (int i = 0; i < 2 * hash_size; i++) hash[i] = (u8)((rem >> (4 * i)) & 15);
}

LC3_HD void fec_crc2(const FecTabs &T, const u8 *data, int n, u8 *hash)
{ // fx/al_fec.c:2256-2314 (two-byte hash)
    u32 rem = 0;
    for (int i = n - 1; i >= 0; i--) rem = fec_crc_reduce(T, (rem << 4) ^ data[i], 2);
    for (int i = 0; i < 4; i++) rem = fec_crc_reduce(T, rem << 4, 2);
    for (int i = 0; i < 4; i++) hash[i] = (u8)((rem >> (4 * i)) & 15);
}

LC3_HD void fec_rs16(const FecTabs &T, u8 *buf, int len, int hd, int mode, int signal)
{ // fx/al_fec.c:558-612: remainder of (data polynomial * x^(hd-1)) by the generator, lowest coefficient first
    const int deg = hd - 1;
    if (deg <= 0) return;
    const u32 g3[2] = {8, 6}, g5[4] = {7, 8, 12, 13}, g7[6] = {12, 10, 12, 3, 9, 7}; // fx/al_fec.c:90-92
    u32 gp[6], sb[7];
    for (int j = 0; j < 6; j++) { gp[j] = hd == 3 ? g3[j & 1] : (hd == 5 ? g5[j & 3] : g7[j]); sb[j + 1] = 0; }
    for (int j = 0; j < deg; j++) { sb[j + 1] = buf[len - deg + j]; buf[j] = 0; }
    for (int i = len - deg - 1; i >= 0; i--) {
        sb[0] = buf[i];
        const u32 lc = sb[deg];
        for (int j = deg - 1; j >= 0; j--) sb[j + 1] = sb[j] ^ gf16_mul(T, gp[j], lc);
    }
    for (int i = 0; i < deg; i++) buf[i] = (u8)sb[i + 1];
    if (signal && mode > 1) { // signalling polynomials of modes 2..4, degree 12 (fx/al_fec.c:98-101; mode 1's is zero)
        const u8 sig[3][13] = {{7, 15, 5, 6, 14, 9, 1, 3, 12, 10, 13, 3, 2},
                               {7, 11, 14, 1, 2, 3, 12, 11, 6, 15, 7, 6, 12},
                               {6, 15, 12, 2, 9, 15, 2, 8, 12, 3, 10, 5, 4}};
        for (int i = 0; i <= 12; i++) buf[i] ^= sig[mode - 2][i];
    }
}

// fx/al_fec.c:481-556.  io: data_bytes of payload on entry, slot_bytes of protected data on return (global memory).
LC3_HD void fec_encode_slot(const FecTabs &T, int mode, int epmr, u8 *io, int data_bytes, int slot_bytes, int n_pccw)
{
    u8 cw[2 * FEC_SLOT_MAX + 16], out[FEC_SLOT_MAX];
    const int n_cw = fec_n_codewords(slot_bytes), nn = 2 * slot_bytes;
    const int pc_split = fec_n_pc(mode, n_pccw, slot_bytes);
    int off = 2 * (slot_bytes - data_bytes), j = nn - 1;
    for (int i = 0; i < off; i++) cw[i] = 0;
    for (int i = 0; i < data_bytes; i++) { const u32 v = io[i]; cw[j--] = (u8)(v & 15); cw[j--] = (u8)(v >> 4); }
    const int n1 = fec_crc1_bytes(mode, slot_bytes), n2 = (pc_split > 0 && mode > 1) ? fec_crc2_bytes(mode) : 0;
    if (n2) fec_crc2(T, cw + off + 2 * data_bytes - pc_split, pc_split, cw + off - 2 * n2);
    if (n1) fec_crc1(T, cw + off, 2 * data_bytes - pc_split, epmr, cw + off - 2 * (n1 + n2), n1);
    off -= 2 * (n1 + n2);
    { // dw0_bitswap, fx/al_fec.c:361-390: the mode request moves next to the first codeword's redundancy
        u8 *dw0 = cw + off;
        const int i0 = 2 * n1 - 1, i1 = fec_cw_length(n_cw, nn, 0) - 1 - (fec_hd(mode, 1) - 1);
        const u32 a = dw0[i0], b = dw0[i1];
        dw0[i0] = (u8)((a & 3) | ((b & 3) << 2));
        dw0[i1] = (u8)((b & 12) | ((a >> 2) & 3));
    }
    int cwo = 0, dwo = off;
    for (int i = 0; i < n_cw; i++) {
        const int hd = fec_hd(mode, i == 0), len = fec_cw_length(n_cw, nn, i);
        for (int k = hd - 1; k < len; k++, dwo++) cw[cwo + k] = cw[dwo];
        fec_rs16(T, cw + cwo, len, hd, mode, i < 6);
        cwo += len;
    }
    for (int i = 0; i < slot_bytes; i++) out[i] = 0;
    cwo = 0;
    for (int i = 0; i < n_cw; i++) { // fec_interleave_pack, fx/al_fec.c:614-646
        const int len = fec_cw_length(n_cw, nn, i);
        for (int k = 0; k < len; k++, cwo++) {
            const int o = nn - (k * n_cw + i + 1);
            out[o >> 1] |= (u8)(cw[cwo] << ((o & 1) << 2));
        }
    }
    for (int i = 0; i < slot_bytes; i++) io[i] = out[i];
}

} // namespace lc3k
