/* ORACLE (test infrastructure, not product code).
 * Bitstream writer: side information packed backwards from the end of the frame
 * (fx/enc_entropy.c:19-175, fx/ari_codec.c:1311-1345) and the 24-bit range coder for TNS data and
 * the quantised spectrum written forwards from byte 0 (fx/ari_codec.c:94-406, :1348-1527). */
#include "lc3o_internal.h"

/* Partial concealment (error protection modes 3 and 4): where the block of `bytes` least-sensitive payload bytes starts.
 * The reference finds it by running its DECODER over the frame it has just written (fx/enc_lc3.c:247-259,
 * processAriDecoder_fx in mode 1) and stopping when the bytes not yet consumed between the decoder's forward (range
 * coder) and backward (side bits) read positions number exactly `bytes` (check_pc_bytes, fx/ari_codec.c:1823-1872).
 * The decoder consumes bytes in the order the encoder produces symbols and bits -- three bytes up front
 * (ac_dec_init_fx, :1559-1580), one per renormalisation shift of a decoded symbol (ac_dec_update_fx, :1776-1795), one
 * whenever a side bit is read from a fresh byte -- so the encoder can count along: cl = bytes the decoder has read from
 * the left so far; the writer's own position is the decoder's backward position. */
typedef struct { int bytes, cl, b_left; } pc_t;

typedef struct { uint8_t *p; i16 bp, mask; pc_t *pc; } bwr; /* backward bit writer */

static void pc_from_left(pc_t *pc, const bwr *w)
{ /* check_pc_bytes(from_left = 1) before the decoder reads byte cl */
    if (!pc || pc->bytes <= 0) return;
    const int side = w->bp + (w->mask == 1);
    if (pc->b_left < 0 && side - 1 - pc->cl == pc->bytes) pc->b_left = pc->cl + 1;
    pc->cl++;
}

static void put_bit(bwr *w, int bit)
{
    if (w->pc && w->pc->bytes > 0 && w->mask == 1 && w->pc->b_left < 0 && w->bp - w->pc->cl == w->pc->bytes)
        w->pc->b_left = w->pc->cl; /* check_pc_bytes(from_left = 0): the decoder is about to open byte bp */
    if (bit > 0) w->p[w->bp] |= (uint8_t)w->mask;
    w->mask = (i16)(w->mask << 1);
    if (w->mask == 0x100) w->mask = 1;
    if (w->mask == 1) w->bp--;
}
static void put_bits(bwr *w, i16 v, int n)
{
    for (int k = 0; k < n; k++) { put_bit(w, v & 1); v = s_lshr(v, 1); }
}

/* fx/enc_entropy.c:19-152 */
void o_side_info(uint8_t *bytes, i16 *bp_side, i16 *mask_side, i16 nbbits, i16 targetBytes, int L, i16 bw_bits,
                 i16 nfilt, i16 lsbMode, i16 lastnz, const i16 *tns_order, i16 fac_ns, i16 gg, i16 bw,
                 const i16 *ltpf_idx, const i32 *scf, int bfi_ext, int fs_idx)
{
    static const i16 lastnz_trigger[5] = {63, 127, 127, 255, 255};
    bwr w;
    w.p = bytes;
    w.pc = NULL;
    w.bp = s_shr(s_sub(nbbits, 1), 3);
    w.mask = s_shl(1, s_sub(8, s_sub(nbbits, s_shl(w.bp, 3))));
    memset(bytes, 0, targetBytes);

    if (bw_bits > 0) put_bits(&w, bw, bw_bits);
    int nb_lastnz = s_sub(14, s_norm(s_neg((i16)L)));
    put_bits(&w, bfi_ext == 1 ? lastnz_trigger[fs_idx] : s_sub(s_shr(lastnz, 1), 1), nb_lastnz);
    put_bit(&w, lsbMode);
    put_bits(&w, gg, 8);
    for (int n = 0; n < nfilt; n++) put_bit(&w, s_min(tns_order[n], 1));
    put_bits(&w, ltpf_idx[0], 1);

    /* SNS: stage 1 (5+5), then submode MSB, gain MSBs, leading sign, joint shape index */
    put_bits(&w, lo16(scf[0]), 5);
    put_bits(&w, lo16(scf[1]), 5);
    i16 sub = lo16(scf[2]), sub_msb = s_shr(sub, 1), sub_lsb = (i16)(sub & 1);
    put_bit(&w, sub_msb);
    i32 gain_lsb = scf[3] & 1;
    put_bits(&w, s_shr(lo16(scf[3]), rom_sns_gain_lsb_bits[sub]), rom_sns_gain_msb_bits[sub]);
    put_bit(&w, lo16(scf[4]));
    if (sub_msb == 0) {
        i32 cwB = gain_lsb;
        if (sub_lsb == 0) cwB = l_add(scf[6], (i32)rom_sns_mpvq_sz[1 * 2 + 1]);
        i32 cw = l_add((i32)((u32)cwB * rom_sns_mpvq_sz[0]), scf[5]);
        put_bits(&w, lo16(cw), 13);
        put_bits(&w, lo16(l_shr(cw, 13)), 12);
    } else {
        i32 cw = scf[5];
        if (sub_lsb != 0) cw = l_add(l_add(l_shl(cw, 1), gain_lsb), (i32)rom_sns_mpvq_sz[2 * 2]);
        put_bits(&w, lo16(cw), 12);
        put_bits(&w, lo16(l_shr(cw, 12)), 12);
    }
    if (ltpf_idx[0] != 0) {
        put_bits(&w, ltpf_idx[1], 1);
        put_bits(&w, ltpf_idx[2], 9);
    }
    put_bits(&w, fac_ns, 3);
    *bp_side = w.bp;
    *mask_side = w.mask;
}

/* ------------------------------------------------------------------ range coder */
typedef struct { u32 low, range; i16 cache, carry, carry_count; uint8_t *p; i16 bp; pc_t *pc; const bwr *w; } rc_t;

/* fx/ari_codec.c:1374-1406 : emit one byte (with deferred carry resolution) */
static void rc_shift(rc_t *s)
{
    if (s->low < 0x00ff0000u || s->carry == 1) {
        if (s->cache >= 0) s->p[s->bp++] = (uint8_t)s_add(s->cache, s->carry);
        while (s->carry_count > 0) {
            s->p[s->bp++] = (uint8_t)(s_add(s->carry, 0xff) & 255);
            s->carry_count = s_sub(s->carry_count, 1);
        }
        s->cache = (i16)(s->low >> 16);
        s->carry = 0;
    } else {
        s->carry_count = s_add(s->carry_count, 1);
    }
    s->low = (s->low << 8) & 0x00ffffffu;
}

/* 24-bit add with wrap flag (fx/ari_codec.c:1264-1267) */
static inline u32 add24(u32 a, u32 b, int *wrap)
{
    u32 x = a << 8, y = b << 8, r = x + y;
    *wrap = r < x;
    return r >> 8;
}

/* fx/ari_codec.c:1408-1446 */
static void rc_encode(rc_t *s, u32 cum, u32 freq)
{
    u32 r = s->range >> 10;
    int carry;
    s->low = add24(s->low, r * cum, &carry);
    if (carry) s->carry = 1;
    s->range = r * freq;
    while (s->range < (1u << 16)) {
        s->range <<= 8;
        rc_shift(s);
        pc_from_left(s->pc, s->w); /* the decoder shifts one byte in for every byte shifted out here */
    }
}

static void put_fwd(uint8_t *p, i16 bp, i16 v, int n)
{ /* fx/ari_codec.c:1348-1368 : overwrite the top n bits of p[bp] with the top n bits of v's low byte */
    i16 t = p[bp], mask = 0x80;
    for (int k = 0; k < n; k++) {
        t |= mask;
        if ((v & mask) == 0) t = (i16)(t - mask);
        mask = s_lshr(mask, 1);
    }
    p[bp] = (uint8_t)t;
}

/* fx/ari_codec.c:1448-1527 */
static i16 rc_finish(rc_t *s)
{
    i16 bits = s_sub(ul_norm(s->range), 7);
    u32 mask = 0x00ffffffu >> bits;
    int o1, o2;
    u32 val = add24(s->low, mask, &o1);
    u32 high = add24(s->low, s->range, &o2);
    val &= (~mask) & 0x00ffffffu;
    if ((o1 ^ o2) == 0) {
        if (val + mask >= high) {
            bits = s_add(bits, 1);
            mask >>= 1;
            val = (s->low + mask) & ((~mask) & 0x00ffffffu);
        }
        if (val < s->low) s->carry = 1;
    }
    s->low = val;
    for (; bits > 0; bits -= 8) rc_shift(s);
    bits = s_add(bits, 8);
    if (s->carry_count > 0) {
        s->p[s->bp++] = (uint8_t)s->cache;
        for (; s->carry_count > 1; s->carry_count--) s->p[s->bp++] = 0xff;
        put_fwd(s->p, s->bp, s_lshr(0xff, s_sub(8, bits)), bits);
    } else {
        put_fwd(s->p, s->bp, s->cache, bits);
    }
    return bits;
}

static inline void enc_sym(rc_t *s, int ctx, int sym)
{
    int row = rom_ari_lookup[ctx] * 17 + sym;
    rc_encode(s, rom_ari_cumfreq[row], rom_ari_freq[row]);
}

/* fx/ari_codec.c:94-406.  n_pc > 0: also returns where the partial-concealment block starts (see pc_t), else -1. */
int o_ari_encode(uint8_t *bytes, i16 bp_side, i16 mask_side, i16 nbbits, const i16 *xq, const i16 *tns_order,
                 i16 nfilt, const i16 *tns_idx, i16 lastnz, const i16 *cd, const uint8_t *res, i16 nres,
                 i16 lsbMode, int lpc_weighting, int n_pc)
{
    i16 lsb[2 * LC3O_MAX_N];
    int nlsbs = 0;
    pc_t pc = {(n_pc + 1) >> 1, 0, -1};
    bwr w = {bytes, bp_side, mask_side, &pc};
    rc_t s = {0, 0x00ffffffu, -1, 0, 0, bytes, 0, &pc, &w};
    for (int i = 0; i < 3; i++) pc_from_left(&pc, &w); /* ac_dec_init_fx */

    for (int n = 0; n < nfilt; n++) {
        if (tns_order[n] <= 0) continue;
        int o = tns_order[n] - 1;
        rc_encode(&s, (u32)rom_tns_order_cumfreq[lpc_weighting * 8 + o], (u32)rom_tns_order_freq[lpc_weighting * 8 + o]);
        for (int k = 0; k < tns_order[n]; k++) {
            int ix = k * 17 + tns_idx[8 * n + k];
            rc_encode(&s, (u32)rom_tns_coef_cumfreq[ix], (u32)rom_tns_coef_freq[ix]);
        }
    }

    for (int k = 0; k < lastnz; k += 2, cd += 3) {
        i16 a = xq[k], b = xq[k + 1], lev = cd[1];
        int ctx = cd[0];
        if (lev < 0) { enc_sym(&s, ctx, 0); continue; }
        if (lev == 0) {
            enc_sym(&s, ctx, cd[2]);
            if (a != 0) put_bit(&w, s_lshr(a, 15));
            if (b != 0) put_bit(&w, s_lshr(b, 15));
            continue;
        }
        i16 a1 = s_abs(a), b1 = s_abs(b);
        if (lsbMode == 0) {
            for (int l = 0; l < lev; l++) { /* escape symbol + one LSB plane per level */
                enc_sym(&s, ctx + rom_ari_esc_nb[l < 3 ? l : 3], 16);
                if (lev == 1) { put_bit(&w, a & 1); put_bit(&w, b & 1); }
                else { put_bit(&w, s_shr(a1, (i16)l) & 1); put_bit(&w, s_shr(b1, (i16)l) & 1); }
            }
            enc_sym(&s, ctx + rom_ari_esc_nb[lev < 3 ? lev : 3], cd[2]);
            if (a != 0) put_bit(&w, s_lshr(a, 15));
            if (b != 0) put_bit(&w, s_lshr(b, 15));
        } else { /* lowest bit-plane is diverted to the lsb[] side buffer */
            i16 am = lev == 1 ? (i16)(cd[2] & 3) : s_shr(a1, 1), bm = lev == 1 ? s_shr(cd[2], 2) : s_shr(b1, 1);
            i16 ta = lev == 1 ? (i16)(a & 1) : (i16)(a1 & 1), tb = lev == 1 ? (i16)(b & 1) : (i16)(b1 & 1);
            if (lev == 1) {
                enc_sym(&s, ctx, 16);
                enc_sym(&s, ctx + rom_ari_esc_nb[1], cd[2]);
                lsb[nlsbs++] = ta;
                if (am == 0 && ta > 0) { if (a > 0) lsb[nlsbs++] = 0; if (a < 0) lsb[nlsbs++] = 1; }
                if (am != 0) put_bit(&w, s_lshr(a, 15));
                lsb[nlsbs++] = tb;
                if (bm == 0 && tb > 0) { if (b > 0) lsb[nlsbs++] = 0; if (b < 0) lsb[nlsbs++] = 1; }
                if (bm != 0) put_bit(&w, s_lshr(b, 15));
            } else {
                lsb[nlsbs++] = ta;
                if (am == 0 && ta > 0) { if (a > 0) lsb[nlsbs++] = 0; if (a < 0) lsb[nlsbs++] = 1; }
                lsb[nlsbs++] = tb;
                if (bm == 0 && tb > 0) { if (b > 0) lsb[nlsbs++] = 0; if (b < 0) lsb[nlsbs++] = 1; }
                enc_sym(&s, ctx + rom_ari_esc_nb[0], 16);
                for (int l = 1; l < lev; l++) {
                    enc_sym(&s, ctx + rom_ari_esc_nb[l < 3 ? l : 3], 16);
                    put_bit(&w, s_shr(a1, (i16)l) & 1);
                    put_bit(&w, s_shr(b1, (i16)l) & 1);
                }
                enc_sym(&s, ctx + rom_ari_esc_nb[lev < 3 ? lev : 3], cd[2]);
                if (am != 0) put_bit(&w, s_lshr(a, 15));
                if (bm != 0) put_bit(&w, s_lshr(b, 15));
            }
        }
    }

    /* bits still free between the two cursors (in sync with the decoder) */
    i16 nbits_side = s_sub(nbbits, s_add(s_shl(w.bp, 3), s_sub(s_norm(w.mask), 6)));
    i16 extra = s_sub(ul_norm(s.range), 6);
    if (s.cache >= 0) extra = s_add(extra, 8);
    if (s.carry_count > 0) extra = s_add(extra, s_shl(s.carry_count, 3));
    i16 n = s_sub(nbbits, s_add(s_shl(s.bp, 3), s_add(extra, nbits_side)));

    if (lsbMode == 0) {
        int m = s_min(nres, n);
        for (int i = 0; i < m; i++) put_bit(&w, res[i]);
    } else {
        int m = s_min((i16)nlsbs, n);
        for (int i = 0; i < m; i++) put_bit(&w, lsb[i]);
    }
    s.pc = NULL; /* the termination bytes are not decoder reads */
    rc_finish(&s);
    if (pc.bytes <= 0) return -1;
    return pc.b_left >= 0 ? pc.b_left : w.bp - pc.bytes; /* fx/ari_codec.c:1133-1141 */
}
