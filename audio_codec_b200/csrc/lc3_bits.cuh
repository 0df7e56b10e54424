// Bitstream writer, one frame per lane: side information packed backwards from the end of the frame
// (fx/enc_entropy.c:19-175, fx/ari_codec.c:1311-1345) and the 24-bit range coder for TNS data and the
// quantised spectrum written forwards from byte 0 (fx/ari_codec.c:94-406, :1348-1527).
#pragma once
#include "lc3_math.cuh"
#include "lc3_tns.cuh"

namespace lc3k {

// Frame bytes of one lane.  LocalBytes: a plain per-lane array (host build, fallback).  The kernel uses a shared-memory
// view (lc3_k_back.cu) in which every lane owns one bank, so the lane-varying byte positions of the two writers never
// conflict and the finished frames can be copied out by the whole warp.
struct LocalBytes {
    u8 *p;
    LC3_HDM i32 get(int i) const { return p[i]; }
    LC3_HDM void set(int i, i32 v) const { p[i] = (u8)v; }
    LC3_HDM void zero(int n) const { for (int i = 0; i < n; i++) p[i] = 0; }
};

// backward bit writer: LSB first, byte pointer descending (fx/ari_codec.c:1311-1345).  Bits collect in `acc` (n of them,
// starting at the current byte's bit 0) and leave as whole bytes; bw_close ORs the last partial byte into place.
template <class B> struct BackWriter { B buf; i32 bp; u32 acc; i32 n; };
template <class B> LC3_HD void bw_flush(BackWriter<B> &w)
{
    while (w.n >= 8) { w.buf.set(w.bp, (i32)(w.acc & 0xffu)); w.bp--; w.acc >>= 8; w.n -= 8; }
}
template <class B> LC3_HD void put_bits(BackWriter<B> &w, i32 v, int n)
{ // n <= 16
    w.acc |= ((u32)v & ((1u << n) - 1u)) << w.n;
    w.n += n;
    bw_flush(w);
}
template <class B> LC3_HD void put_bit(BackWriter<B> &w, i32 bit) { put_bits(w, bit > 0 ? 1 : 0, 1); }
template <class B> LC3_HD void put_bits8(BackWriter<B> &w, i32 v, int n)
{ // n <= 8 and fewer than 8 bits pending: at most one byte leaves, no loop
    w.acc |= ((u32)v & ((1u << n) - 1u)) << w.n;
    w.n += n;
    if (w.n >= 8) { w.buf.set(w.bp, (i32)(w.acc & 0xffu)); w.bp--; w.acc >>= 8; w.n -= 8; }
}
template <class B> LC3_HD void bw_close(BackWriter<B> &w)
{
    if (w.n > 0) w.buf.set(w.bp, w.buf.get(w.bp) | (i32)(w.acc & 0xffu));
}

struct SideInfo {
    i32 bw, bw_bits, lastnz, lsbMode, gg, nfilt, tns_order[2], ltpf_idx[3], fac_ns, bfi_ext;
    i32 scf[7];
};

// fx/enc_entropy.c:19-152.  The frame's bytes must be zeroed over targetBytes by the caller.
template <class B>
LC3_HD void side_info(BackWriter<B> &w, B bytes, i32 nbbits, int L, const SideInfo &si, int fs_idx)
{
    w.buf = bytes;
    w.bp = s_sub(nbbits, 1) >> 3;
    w.acc = 0;
    w.n = s_sub(8, s_sub(nbbits, w.bp << 3)); // first bit position inside the last byte (mask = 1 << n in the reference)

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

// ---- partial concealment (error protection modes 3 and 4).  The reference runs its decoder over the frame it has just
// written (fx/enc_lc3.c:247-259) to find where the block of `bytes` least-sensitive payload bytes starts: the point at which
// the bytes not yet consumed between the decoder's forward (range coder) and backward (side bits) read positions number
// exactly `bytes` (check_pc_bytes, fx/ari_codec.c:1823-1872).  The decoder consumes bytes in the order the encoder makes
// symbols and bits -- three up front (ac_dec_init_fx), one per renormalisation shift of a symbol (ac_dec_update_fx), one
// whenever a side bit comes from a fresh byte -- so the encoder counts along instead: cl = bytes read from the left so far.
struct PcTrack { i32 bytes, cl, b_left; };
template <class B> LC3_HD void pc_from_left(PcTrack &pc, const BackWriter<B> &w)
{ // before the decoder reads byte cl
    const i32 side = w.bp + (w.n == 0 ? 1 : 0);
    if (pc.b_left < 0 && side - 1 - pc.cl == pc.bytes) pc.b_left = pc.cl + 1;
    pc.cl++;
}
template <class B> LC3_HD void pc_from_right(PcTrack &pc, const BackWriter<B> &w, int cnt)
{ // before cnt (<= 8) bits are appended: at most one of them opens a fresh byte
    const i32 first = (8 - w.n) & 7; // index of the first bit that lands on a byte start
    if (first < cnt && pc.b_left < 0 && (w.bp - (w.n == 0 ? 0 : 1)) - pc.cl == pc.bytes) pc.b_left = pc.cl;
}

// ---- range coder (fx/ari_codec.c:1374-1527)
template <class B> struct RangeCoder { u32 low, range; i32 cache, carry, carry_count; B p; i32 bp; };

template <class B> LC3_HD void rc_shift(RangeCoder<B> &s)
{ // emit one byte with deferred carry resolution (fx/ari_codec.c:1374-1406)
    if (s.low < 0x00ff0000u || s.carry == 1) {
        if (s.cache >= 0) s.p.set(s.bp++, s.cache + s.carry);
        while (s.carry_count > 0) {
            s.p.set(s.bp++, (s.carry + 0xff) & 255);
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
template <bool PC, class B> LC3_HD void rc_encode(RangeCoder<B> &s, u32 cum, u32 freq, PcTrack &pc, const BackWriter<B> &w)
{ // fx/ari_codec.c:1408-1446
    u32 r = s.range >> 10;
    int carry;
    s.low = add24(s.low, r * cum, carry);
    if (carry) s.carry = 1;
    s.range = r * freq;
    while (s.range < (1u << 16)) {
        s.range <<= 8;
        rc_shift(s);
        if (PC && pc.bytes > 0) pc_from_left(pc, w); // the decoder shifts a byte in for every byte shifted out here
    }
}
template <class B> LC3_HD void put_fwd(B p, i32 bp, i32 v, int n)
{ // overwrite the top n bits of p[bp] with the top n bits of v's low byte (fx/ari_codec.c:1348-1368)
    i32 t = p.get(bp), mask = 0x80;
    for (int k = 0; k < n; k++) {
        t |= mask;
        if ((v & mask) == 0) t -= mask;
        mask >>= 1;
    }
    p.set(bp, t);
}
template <class B> LC3_HD void rc_finish(RangeCoder<B> &s)
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
        s.p.set(s.bp++, s.cache);
        for (; s.carry_count > 1; s.carry_count--) s.p.set(s.bp++, 0xff);
        put_fwd(s.p, s.bp, s_lshr(0xff, s_sub(8, bits)), bits);
    } else {
        put_fwd(s.p, s.bp, s.cache, bits);
    }
}

// lut: context -> probability model (fx ari_spec_lookup); cf[model*17 + symbol] = freq << 16 | cumfreq (ari_spec_freq / _cumfreq)
struct AriTabs { const u8 *lut; const u32 *cf; };
LC3_HD i32 res_bit(const u32 *res, int i) { return (i32)((res[i >> 5] >> (i & 31)) & 1u); }

// fx/ari_codec.c:94-406.  lsbs: scratch for the diverted LSB plane (2*L bits, packed).
// The context / escape level / MSB symbol of each tuple (the reference's codingdata[], written by its quantiser,
// fx/quantize_spec_fx.c:100-190) are recomputed here from the quantised bins as the loop goes.
//
// The reference walks the tuples and, per tuple, codes `lev` escape symbols and one MSB symbol, with sign / bit-plane
// bits going to the backward writer in between.  Here ONE loop iteration codes ONE range-coder symbol (a single call
// site of the coder), so the 32 lanes of a warp -- 32 different frames -- stay converged through the coder whatever
// their tuples look like; only the short "fetch the next tuple" block is predicated.  The forward (range coder) and
// backward (bits) streams are independent, so only the order inside each stream has to match the reference, and it does:
// per tuple, bit planes in rising level, then the signs.
// n_pc > 0 (partial concealment, see PcTrack): returns the first byte of the concealment block, else -1.  PC = false
// compiles the tracking out (the caller picks it when no lane of the warp needs it).
template <bool PC, class B, class XQ>
LC3_HD i32 ari_encode(B bytes, BackWriter<B> &w, i32 nbbits, XQ xq, int nt, i32 rateFlag, const TnsOut &tns, i32 lastnz,
                      const u32 *res, i32 nres, i32 lsbMode, int lpc_weighting, const AriTabs &T, u32 *lsbs, i32 n_pc)
{
    PcTrack pc = {PC ? (n_pc + 1) >> 1 : 0, 0, -1};
    if (PC && pc.bytes > 0) for (int i = 0; i < 3; i++) pc_from_left(pc, w); // ac_dec_init_fx, fx/ari_codec.c:1559-1580
    int nlsbs = 0;
    u32 lword = 0;
    auto push_lsb = [&](i32 b) {
        lword |= (u32)(b & 1) << (nlsbs & 31);
        nlsbs++;
        if ((nlsbs & 31) == 0) { lsbs[(nlsbs >> 5) - 1] = lword; lword = 0; }
    };
    RangeCoder<B> s = {0, 0x00ffffffu, -1, 0, 0, bytes, 0};

    for (int n = 0; n < tns.nfilt; n++) {
        if (tns.order[n] <= 0) continue;
        const int o = tns.order[n] - 1;
        rc_encode<PC>(s, (u32)rom_tns_order_cumfreq[lpc_weighting * 8 + o], (u32)rom_tns_order_freq[lpc_weighting * 8 + o], pc, w);
        for (int k = 0; k < tns.order[n]; k++) {
            const int ix = k * 17 + tns.idx[8 * n + k];
            rc_encode<PC>(s, (u32)rom_tns_coef_cumfreq[ix], (u32)rom_tns_coef_freq[ix], pc, w);
        }
    }

    const int nt_half = nt >> 1;
    i32 c = 0;
    int k = 0, l = 0, lev = 0, ctx = 0;
    i32 a = 0, b = 0, a1 = 0, b1 = 0, msb = 0;
    i32 na = 0, nb = 0, nna = 0, nnb = 0; // the next two tuples, fetched two tuples ahead of their use
    if (lastnz > 0) {
        na = xq[0]; nb = xq[1];
        const int k2 = 2 < lastnz ? 2 : 0;
        nna = xq[k2]; nnb = xq[k2 + 1];
    }
    while (k < lastnz) {
        if (l == 0) { // next tuple: level, MSB symbol, context (fx/quantize_spec_fx.c:100-190)
            a = na; b = nb; na = nna; nb = nnb;
            const int kn = k + 4 < lastnz ? k + 4 : k;
            nna = xq[kn]; nnb = xq[kn + 1];
            a1 = s_abs(a); b1 = s_abs(b);
            const i32 mx = imax(a1, b1);
            ctx = c + rateFlag;
            if (k > nt_half) ctx += 1 << 8;
            lev = imax(0, 13 - s_norm(mx | 1)); // 0 below 4, 1 below 8, 13 - norm_s(mx) from there
            const i32 am = a1 >> lev, bm = b1 >> lev;
            msb = am + (bm << 2);
            const i32 t = lev < 2 ? ((am + bm) << lev) + 1 : 12 + imin(lev, 3); // 1 | a1 + b1 + 1 | 2 (a1 / 2 + b1 / 2) + 1 | 12 + ..
            c = ((c & 0xf) << 4) + t;
        }
        const bool last = l == lev; // this iteration codes the MSB symbol, the earlier ones an escape each
        {
            const int row = T.lut[ctx + (imin(l, 3) << 10)] * 17 + (last ? msb : 16); // rom_ari_esc_nb[q] = 1024 q
            const u32 cf = T.cf[row];
            rc_encode<PC>(s, cf & 0xffffu, cf >> 16, pc, w);
        }
        // side bits of this iteration, without branching on `last`: bit plane l of both magnitudes after an escape, the
        // signs after the MSB symbol (of the values whose sign has not gone with a diverted LSB)
        const i32 ta = (a1 >> l) & 1, tb = (b1 >> l) & 1;
        const bool div0 = lsbMode != 0 && l == 0 && !last;
        if (div0) { // lowest plane diverted to lsbs[]; a value that is +-1 takes its sign along
            const i32 am = a1 >> 1, bm = b1 >> 1;
            push_lsb(ta);
            if (am == 0 && ta > 0) push_lsb(a < 0 ? 1 : 0);
            push_lsb(tb);
            if (bm == 0 && tb > 0) push_lsb(b < 0 ? 1 : 0);
        }
        const bool dv = lsbMode != 0 && lev > 0;
        const i32 sa = (dv ? (a1 >> 1) : a1) != 0, sb = (dv ? (b1 >> 1) : b1) != 0;
        const i32 na_ = (a >> 15) & 1, nb_ = (b >> 15) & 1;
        const i32 v = last ? (sa ? (na_ | (nb_ << 1)) : nb_) : (ta | (tb << 1));
        const int cnt = last ? sa + sb : (div0 ? 0 : 2);
        if (PC && pc.bytes > 0) pc_from_right(pc, w, cnt);
        put_bits8(w, v, cnt);
        k += last ? 2 : 0;
        l = last ? 0 : l + 1;
    }
    if (nlsbs & 31) lsbs[nlsbs >> 5] = lword;

    // bits still free between the two cursors (in sync with the decoder)
    const i32 nbits_side = s_sub(nbbits, s_add(w.bp << 3, s_sub(8, w.n)));
    i32 extra = s_sub(ul_norm(s.range), 6);
    if (s.cache >= 0) extra = s_add(extra, 8);
    if (s.carry_count > 0) extra = s_add(extra, s.carry_count << 3);
    const i32 n = s_sub(nbbits, s_add(s.bp << 3, s_add(extra, nbits_side)));

    const u32 *src = lsbMode == 0 ? res : lsbs;
    const int m = imin(lsbMode == 0 ? nres : nlsbs, n);
    for (int i = 0; i < m; i += 8) { // eight at a time
        const int cnt = imin(8, m - i);
        const u32 v = src[i >> 5] >> (i & 31); // i is a multiple of 8: the group never straddles two words
        if (PC && pc.bytes > 0) pc_from_right(pc, w, cnt);
        put_bits8(w, (i32)(v & 0xffu), cnt);
    }
    const i32 b_left = pc.bytes <= 0 ? -1 : (pc.b_left >= 0 ? pc.b_left : w.bp - pc.bytes); // fx/ari_codec.c:1133-1141
    bw_close(w);
    rc_finish(s);
    return b_left;
}

// fx/reorder_bitstream_fx.c:16-56: the concealment block [b_left, b_left + block) moves to the front, the bytes before it
// follow.  In place: a rotation of the first b_left + block bytes, done as three reversals.
template <class B> LC3_HD void reorder_payload(B bytes, i32 n_pc, i32 b_left)
{
    const int block = (n_pc + 1) >> 1, L = b_left + block;
    auto rev = [&](int lo, int hi) { // [lo, hi)
        for (hi--; lo < hi; lo++, hi--) { const i32 t = bytes.get(lo); bytes.set(lo, bytes.get(hi)); bytes.set(hi, t); }
    };
    rev(0, L); rev(0, block); rev(block, L);
}

} // namespace lc3k
