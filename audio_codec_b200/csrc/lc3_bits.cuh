// Bitstream writer, one frame per lane: side information packed backwards from the end of the frame
// (fx/enc_entropy.c:19-175, fx/ari_codec.c:1311-1345) and the 24-bit range coder for TNS data and the
// quantised spectrum written forwards from byte 0 (fx/ari_codec.c:94-406, :1348-1527).
#pragma once
#include "lc3_math.cuh"
#include "lc3_tns.cuh"

namespace lc3k {

struct BackWriter { u8 *p; i32 bp, mask; }; // backward bit writer: LSB first, byte pointer descending
LC3_HD void put_bit(BackWriter &w, i32 bit)
{
    if (bit > 0) w.p[w.bp] |= (u8)w.mask;
    w.mask <<= 1;
    if (w.mask == 0x100) { w.mask = 1; w.bp--; }
}
LC3_HD void put_bits(BackWriter &w, i32 v, int n)
{
    v &= 0xffff;
    for (int k = 0; k < n; k++) { put_bit(w, v & 1); v >>= 1; }
}

struct SideInfo {
    i32 bw, bw_bits, lastnz, lsbMode, gg, nfilt, tns_order[2], ltpf_idx[3], fac_ns, bfi_ext;
    i32 scf[7];
};

// fx/enc_entropy.c:19-152.  bytes[] must be zeroed over targetBytes by the caller.
LC3_HD void side_info(BackWriter &w, u8 *bytes, i32 nbbits, int L, const SideInfo &si, int fs_idx)
{
    w.p = bytes;
    w.bp = s_sub(nbbits, 1) >> 3;
    w.mask = s_shl(1, s_sub(8, s_sub(nbbits, w.bp << 3)));

    if (si.bw_bits > 0) put_bits(w, si.bw, si.bw_bits);
    const int nb_lastnz = s_sub(14, s_norm(s_neg(L)));
    const i32 trigger = fs_idx == 0 ? 63 : (fs_idx <= 2 ? 127 : 255);
    put_bits(w, si.bfi_ext == 1 ? trigger : s_sub(si.lastnz >> 1, 1), nb_lastnz);
    put_bit(w, si.lsbMode);
    put_bits(w, si.gg, 8);
    for (int n = 0; n < si.nfilt; n++) put_bit(w, imin(si.tns_order[n], 1));
    put_bits(w, si.ltpf_idx[0], 1);

    // SNS: stage 1 (5+5), submode MSB, gain MSBs, leading sign, joint shape index
    put_bits(w, lo16(si.scf[0]), 5);
    put_bits(w, lo16(si.scf[1]), 5);
    const i32 sub = lo16(si.scf[2]), sub_msb = sub >> 1, sub_lsb = sub & 1;
    put_bit(w, sub_msb);
    const i32 gain_lsb = si.scf[3] & 1;
    put_bits(w, lo16(si.scf[3]) >> rom_sns_gain_lsb_bits[sub], rom_sns_gain_msb_bits[sub]);
    put_bit(w, lo16(si.scf[4]));
    if (sub_msb == 0) {
        i32 cwB = gain_lsb;
        if (sub_lsb == 0) cwB = l_add(si.scf[6], (i32)rom_sns_mpvq_sz[1 * 2 + 1]);
        i32 cw = l_add((i32)((u32)cwB * rom_sns_mpvq_sz[0]), si.scf[5]);
        put_bits(w, lo16(cw), 13);
        put_bits(w, lo16(cw >> 13), 12);
    } else {
        i32 cw = si.scf[5];
        if (sub_lsb != 0) cw = l_add(l_add(l_shl_p(cw, 1), gain_lsb), (i32)rom_sns_mpvq_sz[2 * 2]);
        put_bits(w, lo16(cw), 12);
        put_bits(w, lo16(cw >> 12), 12);
    }
    if (si.ltpf_idx[0] != 0) {
        put_bits(w, si.ltpf_idx[1], 1);
        put_bits(w, si.ltpf_idx[2], 9);
    }
    put_bits(w, si.fac_ns, 3);
}

// ---- range coder (fx/ari_codec.c:1374-1527)
struct RangeCoder { u32 low, range; i32 cache, carry, carry_count; u8 *p; i32 bp; };

LC3_HD void rc_shift(RangeCoder &s)
{ // emit one byte with deferred carry resolution (fx/ari_codec.c:1374-1406)
    if (s.low < 0x00ff0000u || s.carry == 1) {
        if (s.cache >= 0) s.p[s.bp++] = (u8)(s.cache + s.carry);
        while (s.carry_count > 0) {
            s.p[s.bp++] = (u8)((s.carry + 0xff) & 255);
            s.carry_count--;
        }
        s.cache = (i32)(s.low >> 16);
        s.carry = 0;
    } else {
        s.carry_count++;
    }
    s.low = (s.low << 8) & 0x00ffffffu;
}
LC3_HD u32 add24(u32 a, u32 b, int &wrap)
{ // 24-bit add with wrap flag (fx/ari_codec.c:1264-1267)
    u32 x = a << 8, y = b << 8, r = x + y;
    wrap = r < x;
    return r >> 8;
}
LC3_HD void rc_encode(RangeCoder &s, u32 cum, u32 freq)
{ // fx/ari_codec.c:1408-1446
    u32 r = s.range >> 10;
    int carry;
    s.low = add24(s.low, r * cum, carry);
    if (carry) s.carry = 1;
    s.range = r * freq;
    while (s.range < (1u << 16)) {
        s.range <<= 8;
        rc_shift(s);
    }
}
LC3_HD void put_fwd(u8 *p, i32 bp, i32 v, int n)
{ // overwrite the top n bits of p[bp] with the top n bits of v's low byte (fx/ari_codec.c:1348-1368)
    i32 t = p[bp], mask = 0x80;
    for (int k = 0; k < n; k++) {
        t |= mask;
        if ((v & mask) == 0) t -= mask;
        mask >>= 1;
    }
    p[bp] = (u8)t;
}
LC3_HD void rc_finish(RangeCoder &s)
{ // fx/ari_codec.c:1448-1527
    i32 bits = s_sub(ul_norm(s.range), 7);
    u32 mask = 0x00ffffffu >> bits;
    int o1, o2;
    u32 val = add24(s.low, mask, o1);
    u32 high = add24(s.low, s.range, o2);
    val &= (~mask) & 0x00ffffffu;
    if ((o1 ^ o2) == 0) {
        if (val + mask >= high) {
            bits = s_add(bits, 1);
            mask >>= 1;
            val = (s.low + mask) & ((~mask) & 0x00ffffffu);
        }
        if (val < s.low) s.carry = 1;
    }
    s.low = val;
    for (; bits > 0; bits -= 8) rc_shift(s);
    bits = s_add(bits, 8);
    if (s.carry_count > 0) {
        s.p[s.bp++] = (u8)s.cache;
        for (; s.carry_count > 1; s.carry_count--) s.p[s.bp++] = 0xff;
        put_fwd(s.p, s.bp, s_lshr(0xff, s_sub(8, bits)), bits);
    } else {
        put_fwd(s.p, s.bp, s.cache, bits);
    }
}

struct AriTabs { const u8 *lut; const uint16_t *cum, *freq; };
LC3_HD void enc_sym(RangeCoder &s, const AriTabs &T, int ctx, int sym)
{
    const int row = T.lut[ctx] * 17 + sym;
    rc_encode(s, T.cum[row], T.freq[row]);
}
LC3_HD i32 res_bit(const u32 *res, int i) { return (i32)((res[i >> 5] >> (i & 31)) & 1u); }

// fx/ari_codec.c:94-406.  lsbs: scratch for the diverted LSB plane (2*L bits, packed).
// The context / escape level / MSB symbol of each tuple (the reference's codingdata[], written by its quantiser,
// fx/quantize_spec_fx.c:100-190) are recomputed here from the quantised bins as the loop goes.
template <class XQ>
LC3_HD void ari_encode(u8 *bytes, BackWriter &w, i32 nbbits, XQ xq, int nt, i32 rateFlag, const TnsOut &tns, i32 lastnz,
                       const u32 *res, i32 nres, i32 lsbMode, int lpc_weighting, const AriTabs &T, u32 *lsbs)
{
    int nlsbs = 0;
    u32 lword = 0;
    auto push_lsb = [&](i32 b) {
        lword |= (u32)(b & 1) << (nlsbs & 31);
        nlsbs++;
        if ((nlsbs & 31) == 0) { lsbs[(nlsbs >> 5) - 1] = lword; lword = 0; }
    };
    RangeCoder s = {0, 0x00ffffffu, -1, 0, 0, bytes, 0};

    for (int n = 0; n < tns.nfilt; n++) {
        if (tns.order[n] <= 0) continue;
        const int o = tns.order[n] - 1;
        rc_encode(s, (u32)rom_tns_order_cumfreq[lpc_weighting * 8 + o], (u32)rom_tns_order_freq[lpc_weighting * 8 + o]);
        for (int k = 0; k < tns.order[n]; k++) {
            const int ix = k * 17 + tns.idx[8 * n + k];
            rc_encode(s, (u32)rom_tns_coef_cumfreq[ix], (u32)rom_tns_coef_freq[ix]);
        }
    }

    const int nt_half = nt >> 1;
    i32 c = 0;
    for (int k = 0; k < lastnz; k += 2) {
        const i32 a = xq[k], b = xq[k + 1];
        const i32 a1 = s_abs(a), b1 = s_abs(b), mx = imax(a1, b1);
        int ctx = c + rateFlag;
        if (k > nt_half) ctx += 1 << 8;
        if (mx == 0) { enc_sym(s, T, ctx, 0); c = ((c & 0xf) << 4) + 1; continue; }
        i32 lev, msb;
        if (mx < 4) {
            lev = 0; msb = a1 + (b1 << 2);
            c = ((c & 0xf) << 4) + a1 + b1 + 1;
        } else if (mx < 8) {
            lev = 1; msb = (a1 >> 1) + ((b1 >> 1) << 2);
            c = ((c & 0xf) << 4) + 2 * ((a1 >> 1) + (b1 >> 1)) + 1;
        } else {
            lev = 13 - s_norm(mx); msb = (a1 >> lev) + ((b1 >> lev) << 2);
            c = ((c & 0xf) << 4) + 12 + imin(lev, 3);
        }
        if (lev == 0) {
            enc_sym(s, T, ctx, msb);
            if (a != 0) put_bit(w, (a >> 15) & 1);
            if (b != 0) put_bit(w, (b >> 15) & 1);
            continue;
        }
        if (lsbMode == 0) {
            for (int l = 0; l < lev; l++) { // escape symbol + one LSB plane per level
                enc_sym(s, T, ctx + rom_ari_esc_nb[l < 3 ? l : 3], 16);
                if (lev == 1) { put_bit(w, a & 1); put_bit(w, b & 1); }
                else { put_bit(w, (a1 >> l) & 1); put_bit(w, (b1 >> l) & 1); }
            }
            enc_sym(s, T, ctx + rom_ari_esc_nb[lev < 3 ? lev : 3], msb);
            if (a != 0) put_bit(w, (a >> 15) & 1);
            if (b != 0) put_bit(w, (b >> 15) & 1);
        } else { // lowest bit-plane is diverted to the lsbs[] side buffer
            const i32 am = lev == 1 ? (msb & 3) : (a1 >> 1), bm = lev == 1 ? (msb >> 2) : (b1 >> 1);
            const i32 ta = lev == 1 ? (a & 1) : (a1 & 1), tb = lev == 1 ? (b & 1) : (b1 & 1);
            if (lev == 1) {
                enc_sym(s, T, ctx, 16);
                enc_sym(s, T, ctx + rom_ari_esc_nb[1], msb);
                push_lsb(ta);
                if (am == 0 && ta > 0) { if (a > 0) push_lsb(0); if (a < 0) push_lsb(1); }
                if (am != 0) put_bit(w, (a >> 15) & 1);
                push_lsb(tb);
                if (bm == 0 && tb > 0) { if (b > 0) push_lsb(0); if (b < 0) push_lsb(1); }
                if (bm != 0) put_bit(w, (b >> 15) & 1);
            } else {
                push_lsb(ta);
                if (am == 0 && ta > 0) { if (a > 0) push_lsb(0); if (a < 0) push_lsb(1); }
                push_lsb(tb);
                if (bm == 0 && tb > 0) { if (b > 0) push_lsb(0); if (b < 0) push_lsb(1); }
                enc_sym(s, T, ctx + rom_ari_esc_nb[0], 16);
                for (int l = 1; l < lev; l++) {
                    enc_sym(s, T, ctx + rom_ari_esc_nb[l < 3 ? l : 3], 16);
                    put_bit(w, (a1 >> l) & 1);
                    put_bit(w, (b1 >> l) & 1);
                }
                enc_sym(s, T, ctx + rom_ari_esc_nb[lev < 3 ? lev : 3], msb);
                if (am != 0) put_bit(w, (a >> 15) & 1);
                if (bm != 0) put_bit(w, (b >> 15) & 1);
            }
        }
    }
    if (nlsbs & 31) lsbs[nlsbs >> 5] = lword;

    // bits still free between the two cursors (in sync with the decoder)
    const i32 nbits_side = s_sub(nbbits, s_add(w.bp << 3, s_sub(s_norm(w.mask), 6)));
    i32 extra = s_sub(ul_norm(s.range), 6);
    if (s.cache >= 0) extra = s_add(extra, 8);
    if (s.carry_count > 0) extra = s_add(extra, s.carry_count << 3);
    const i32 n = s_sub(nbbits, s_add(s.bp << 3, s_add(extra, nbits_side)));

    const u32 *src = lsbMode == 0 ? res : lsbs;
    const int m = imin(lsbMode == 0 ? nres : nlsbs, n);
    for (int i = 0; i < m; i++) put_bit(w, res_bit(src, i));
    rc_finish(s);
}

} // namespace lc3k
