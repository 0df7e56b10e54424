// Frame-parallel half of the batched LC3plus decoder, ONE LANE PER CHANNEL-FRAME (fx/dec_lc3.c:115-232 for good frames):
// side information (fx/dec_entropy.c:22-194), arithmetic decoder and residual bits (fx/ari_codec.c:409-1214, without
// partial concealment: n_pc = 0), spectrum scaling (:1217-1256), residual decoding (fx/residual_decoding_fx.c:12-69),
// noise filling (fx/noise_filling_fx.c:14-129), global gain (fx/apply_global_gain_fx.c:14-44), TNS synthesis
// (fx/tns_decoder_fx.c:20-134), SNS decoding (fx/sns_quantize_scf_fx.c:442-570, fx/pvq_dec_fx.c, fx/pvq_index_fx.c:202-345)
// and spectral shaping (fx/sns_interpolate_scf_fx.c, fx/mdct_shaping_fx.c).
//
// Nothing here depends on another frame: the decoder's inter-frame state (overlap-add memory, LTPF memories) belongs to the
// per-stream pass in lc3_dec_synth.cuh.  Concealment (fx/plc_*.c, fx/pc_*.c) is NOT part of this library: a payload the
// reference would flag as a bad frame is reported through DO_BFI and synthesised as a zero spectrum.
#pragma once
#include "lc3_dec_common.cuh"

namespace lc3k {

// ---- bit reader of the side-information end of the frame (read_bit / read_indice, fx/ari_codec.c:1530-1556,
// fx/dec_entropy.c:284-301): bits run from the last byte towards the first, least significant bit first.  Instead of the
// reference's byte pointer and mask the reader keeps the count of bits taken and two 32-bit words of the bits ahead (a take
// is one funnel shift), refilled one aligned word at a time; pointer and mask follow from the count: bp_side = last - (nread >> 3), mask_side =
// 1 << (nread & 7).  Bytes before the start of the frame read as zero; rd_oob() tells whether a bit was TAKEN from there
// (only a corrupt frame gets that far).  B::get(i) = byte i of the frame, B::word(k) = bytes 4k..4k+3, little-endian.
template <class B> struct BackReader { B buf; i32 last /* index of the frame's last byte */, nread, pos /* bit position in cur */, next /* bytes [0, next) not loaded yet */; u32 cur, nxt; };
LC3_HD u32 bswap32(u32 v)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(v, 0, 0x0123);
#else
    return __builtin_bswap32(v);
#endif
}
LC3_HD u32 funnel_r(u32 lo, u32 hi, int sh) // bits [sh, sh + 32) of hi:lo, 0 <= sh < 32
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return (u32)((((uint64_t)hi << 32) | lo) >> sh);
#endif
}
template <class B> LC3_HD u32 rd_word(BackReader<B> &r) // the next 32 bits of the stream: four bytes going down, the highest in bits 0..7
{
    if (r.next <= 0) return 0;
    r.next -= 4;
    return bswap32(r.buf.word(r.next >> 2));
}
template <class B> LC3_HD void rd_init(BackReader<B> &r, B buf, i32 nbytes)
{
    r.buf = buf; r.last = nbytes - 1; r.nread = 0; r.next = nbytes > 0 ? nbytes : 0;
    // the frame's top (nbytes & 3) bytes form a partial first word, placed at the TOP of cur with the position past the missing
    // low part; from there on the words are aligned
    u32 part = 0;
    int k = 0;
    while (r.next & 3) { part |= (u32)buf.get(--r.next) << (8 * k); k++; }
    if (k) { r.cur = part << (32 - 8 * k); r.pos = 32 - 8 * k; }
    else { r.cur = rd_word(r); r.pos = 0; }
    r.nxt = rd_word(r);
}
template <class B> LC3_HD i32 rd_bp(const BackReader<B> &r) { return r.last - (r.nread >> 3); }
template <class B> LC3_HD i32 rd_oob(const BackReader<B> &r) { return r.nread > ((r.last + 1) << 3); }
// the next n bits (0 <= n <= 32), first bit read in bit 0
template <class B> LC3_HD u32 rd_take(BackReader<B> &r, int n)
{
    const u32 f = funnel_r(r.cur, r.nxt, r.pos);
    const u32 v = n ? f & (0xffffffffu >> (32 - n)) : 0u;
    r.pos += n; r.nread += n;
    if (r.pos >= 32) { r.cur = r.nxt; r.nxt = rd_word(r); r.pos -= 32; }
    return v;
}
template <class B> LC3_HD i32 rd_bit(BackReader<B> &r) { return (i32)rd_take(r, 1); }
template <class B> LC3_HD i32 rd_bits(BackReader<B> &r, int n) { return (i32)rd_take(r, n); }

struct DecSide {
    i32 bw_idx, nfilt, lsbMode, lastnz, gg_idx, fac_ns_idx;
    i32 tns_order[2], ltpf[3], scf[7];
};

// ac_dec_split_st2VQ_CW (fx/dec_entropy.c:303-369): de-multiplex the joint shape-A / shape-B-or-gain-LSB codeword
LC3_HD i32 split_st2_cw(u32 cw, u32 szA, u32 szB, i32 &cwA, i32 &cwB, i32 &lsb)
{
    const u32 max_size = szB * szA;
    if (cw >= max_size) { cwA = 0; cwB = 0; lsb = 0; return 1; }
    i32 start = 0;
    if (cw >= (max_size >> 1)) start = (i32)(szB >> 1);
    i32 ind = start;
    i32 t = -(i32)cw + (i32)((u32)start * szA);
    const i32 fin = start + (i32)(szB >> 1);
    for (i32 i = start; i < fin; i++) {
        ind++;
        t += (i32)szA;
        if (t > 0) ind--;
    }
    cwB = ind;
    cwA = (i32)(cw - (u32)ind * szA);
    lsb = 0;
    cwB -= 2;
    if (cwB < 0) lsb = 1;
    cwB += 2 * lsb;
    return 0;
}

// processDecoderEntropy_fx (fx/dec_entropy.c:22-194).  Returns the bad-frame flag.
template <class B> LC3_HD i32 dec_side_info(BackReader<B> &r, int yLen, int fs_idx, int bw_bits, int dms, DecSide &s)
{
    if (bw_bits > 0) {
        s.bw_idx = rd_bits(r, bw_bits);
        if (fs_idx < s.bw_idx) { s.bw_idx = fs_idx; return 1; }
    } else s.bw_idx = fs_idx;
    s.nfilt = (s.bw_idx >= 3 && dms >= 50) ? 2 : 1;
    const int L = 14 - s_norm(-yLen);
    const i32 n = rd_bits(r, L) + 1;
    s.lastnz = n << 1;
    if (s.lastnz > yLen) return 1;
    s.lsbMode = rd_bit(r);
    s.gg_idx = rd_bits(r, 8);
    s.tns_order[1] = 0;
    for (int k = 0; k < s.nfilt; k++) s.tns_order[k] = rd_bit(r);
    s.ltpf[0] = rd_bit(r);
    s.scf[0] = rd_bits(r, 5);
    s.scf[1] = rd_bits(r, 5);
    const i32 msb = rd_bit(r);
    s.scf[2] = msb << 1;
    s.scf[3] = rd_bits(r, rom_sns_gain_msb_bits[s.scf[2]]);
    s.scf[4] = rd_bit(r);
    i32 lsb, ber;
    if (msb == 0) {
        u32 cw = (u32)rd_bits(r, 13);
        cw |= (u32)rd_bits(r, 12) << 13;
        ber = split_st2_cw(cw, rom_sns_mpvq_sz[0], rom_sns_mpvq_sz[1] + rom_sns_mpvq_sz[3], s.scf[5], s.scf[6], lsb);
        if (lsb != 0) { s.scf[3] = (s.scf[3] << 1) + s.scf[6]; s.scf[6] = -2; }
    } else {
        i32 cw = rd_bits(r, 12);
        cw |= rd_bits(r, 12) << 12;
        s.scf[5] = cw;
        s.scf[6] = -1;
        lsb = 0; ber = 0;
        const i32 lim = (i32)rom_sns_mpvq_sz[4] + (i32)(rom_sns_mpvq_sz[6] << 1);
        if (cw >= lim) ber = 1;
        else {
            cw -= (i32)rom_sns_mpvq_sz[4];
            if (cw >= 0) { lsb = 1; s.scf[3] = (s.scf[3] << 1) + (cw & 1); s.scf[5] = cw >> 1; s.scf[6] = -2; }
        }
    }
    s.scf[2] += lsb;
    if (ber) return 1;
    if (s.ltpf[0]) { s.ltpf[1] = rd_bit(r); s.ltpf[2] = rd_bits(r, 9); }
    else { s.ltpf[1] = 0; s.ltpf[2] = 0; }
    s.fac_ns_idx = rd_bits(r, 3);
    return 0;
}

// ---- range decoder (ac_dec_init_fx / ac_decode_* / ac_dec_update_fx, fx/ari_codec.c:1559-1803)
template <class B> struct RangeDec { u32 low, range, help, wbuf /* bytes of the current aligned word still to come */; i32 ber, bp, nbytes; B p; };
template <class B> LC3_HD u32 rdc_byte(RangeDec<B> &s)
{
    if ((s.bp & 3) == 0 && s.bp < s.nbytes) s.wbuf = s.p.word(s.bp >> 2); // the frame's storage is whole words
    const u32 v = s.bp < s.nbytes ? s.wbuf & 0xffu : 0u;
    if (s.bp >= s.nbytes) s.ber = 1;
    s.wbuf >>= 8;
    s.bp++;
    return v;
}
template <class B> LC3_HD void rdc_init(RangeDec<B> &s)
{
    s.low = 0; s.range = 0x00ffffff; s.ber = 0;
    for (int i = 0; i < 3; i++) s.low = (s.low << 8) + rdc_byte(s);
}
// Symbol search in a cumulative-frequency row of 17 entries (spectrum models, TNS coefficients); CF(r) = row entry r.
// The reference bisects in four dependent steps (ac_decode_fx, fx/ari_codec.c:1589-1648): the result is the LARGEST r <= 15
// with low >= help * CF(r), because every row rises monotonically from 0 (tests/test_rom_bounds.py).  The same r is found here
// in two rounds of three independent comparisons (quarter, then position inside it): half the chain of dependent table
// loads, which is what a lane waits for in this loop.  r = 15 then goes on to the escape entry and the range check as there.
template <class B, class CF> LC3_HD i32 rdc_decode17(RangeDec<B> &s, CF cf)
{
    s.help = s.range >> 10;
    const u32 low = s.low, h = s.help;
    const u32 c4 = cf(4), c8 = cf(8), c12 = cf(12);
    const i32 base = 4 * ((low >= h * c4) + (low >= h * c8) + (low >= h * c12));
    const u32 d1 = cf(base + 1), d2 = cf(base + 2), d3 = cf(base + 3);
    i32 val = base + (low >= h * d1) + (low >= h * d2) + (low >= h * d3);
    if (val == 15) {
        if (low >= h * cf(16)) val = 16;
        if (low >= (h << 10)) s.ber = 1;
    }
    return val;
}
template <class B> LC3_HD i32 rdc_decode_tns_order(RangeDec<B> &s, const i16 *cum /* 8 */)
{
    s.help = s.range >> 10;
    i32 val = 0;
    if (s.low >= s.help * (u32)cum[4]) val = 4;
    if (s.low >= s.help * (u32)cum[val + 2]) val += 2;
    if (s.low >= s.help * (u32)cum[val + 1]) val += 1;
    if (s.low >= (s.help << 10)) s.ber = 1;
    return val;
}
// range >= help >= 2^6 after the update (range was >= 2^16, freq >= 1): at most two bytes are shifted in
template <class B> LC3_HD void rdc_update(RangeDec<B> &s, u32 cum, u32 freq)
{
    s.low -= cum * s.help;
    s.range = s.help * freq;
#pragma unroll
    for (int k = 0; k < 2; k++)
        if (s.range < (1u << 16)) {
            s.low = ((s.low & 0xffffu) << 8) + rdc_byte(s);
            s.range <<= 8;
        }
    if (s.ber) s.range = 0x00ffffff; // past the end of a corrupt frame: the caller stops at the flag
}

struct AriOut { i32 bfi, resBits, nf_seed, zero_frame, x_absmax; i16 tns_idx[16]; };

// processAriDecoder_fx for n_pc == 0, mode 0 (fx/ari_codec.c:409-1214).  x: SV<i16>-like, receives yLen bins.
// lsb_ind: scratch of yLen / 2 words.  res: residual bits, packed.
//
// The reference walks the tuples and, per tuple, decodes `lev` escape symbols and one MSB symbol with sign / bit-plane bits
// from the backward reader in between.  Here ONE loop iteration decodes ONE range-coder symbol (a single call site of the
// decoder), so the 32 lanes of a warp -- 32 different frames -- stay converged through the decoder whatever their tuples look
// like; what the symbol means (escape of level lev, or the MSB symbol that ends the tuple) is predicated.  The order inside
// each of the two streams is the reference's: per tuple, bit planes in rising level, then the signs.
template <class B, class XQ, class LI>
LC3_HD void ari_decode(B bytes, BackReader<B> &r, i32 nbbits, int yLen, int fs_idx, int lpc_weighting, int dms, DecSide &si,
                       const AriTabs &T, XQ x, LI lsb_ind, u32 *res, AriOut &o)
{
    RangeDec<B> st;
    st.p = bytes; st.bp = 0; st.nbytes = nbbits >> 3; st.wbuf = 0;
    const i32 rateFlag = nbbits > 160 + fs_idx * 160 ? 2 << 8 : 0;
    const i32 nt_half = yLen >> 1;
    i32 c = 0, lsb_cnt = 0;
    o.bfi = 0; o.resBits = 0; o.nf_seed = 0; o.zero_frame = 0; o.x_absmax = 0;
    for (int i = 0; i < 16; i++) o.tns_idx[i] = 0;
    rdc_init(st);
    // TNS data
    int maxord = 8;
    if (dms == 25) maxord >>= 1;
    if (dms == 50) maxord >>= 1;
    for (int n = 0; n < si.nfilt; n++) {
        if (si.tns_order[n] > 0) {
            const i32 ord = rdc_decode_tns_order(st, &rom_tns_order_cumfreq[lpc_weighting * 8]) + 1;
            si.tns_order[n] = ord;
            if (ord > maxord) { o.bfi = 1; return; }
            rdc_update(st, (u32)rom_tns_order_cumfreq[lpc_weighting * 8 + ord - 1], (u32)rom_tns_order_freq[lpc_weighting * 8 + ord - 1]);
            if (st.ber) { o.bfi = 1; return; }
            for (int k = 0; k < ord; k++) {
                if (rd_bp(r) < st.bp) { o.bfi = 1; return; }
                const i16 *cum = &rom_tns_coef_cumfreq[k * 17];
                const i32 v = rdc_decode17(st, [&](int q) { return (u32)cum[q]; });
                o.tns_idx[8 * n + k] = (i16)v;
                rdc_update(st, (u32)cum[v], (u32)rom_tns_coef_freq[k * 17 + v]);
                if (st.ber) { o.bfi = 1; return; }
            }
        }
    }
    if (st.ber) { o.bfi = 1; return; }

    const int lsbMode = si.lsbMode;
    i32 k = 0, lev = 0, a = 0, bb = 0, t = c + rateFlag, seed = 0, mx = 0, nnz = 0;
    while (k < si.lastnz) {
        const i32 tt = t + (imin(lev, 3) << 10); // rom_ari_esc_nb[q] = q << 10
        const u32 *cf = T.cf + 17 * T.lut[tt];
        const i32 sym = rdc_decode17(st, [&](int q) { return cf[q] & 0xffffu; });
        rdc_update(st, cf[sym] & 0xffffu, cf[sym] >> 16);
        // the iteration's bits from the side stream in ONE take: an escape brings one more bit plane of both coefficients (the
        // lowest one is deferred in lsbMode), the tuple's MSB symbol the signs of its non-zero coefficients
        const bool esc = sym >= 16;
        const i32 a1 = esc ? 0 : sym & 3, b1 = esc ? 0 : sym >> 2;
        a += a1 << lev; bb += b1 << lev;
        const int nb_a = esc ? (lev > 0 || !lsbMode) : a > 0, nb_b = esc ? nb_a : bb > 0;
        const u32 tk = rd_take(r, nb_a + nb_b);
        const i32 bit_a = nb_a ? (i32)(tk & 1u) : 0, bit_b = nb_b ? (i32)((tk >> nb_a) & 1u) : 0;
        if (esc) {
            a += bit_a << lev; bb += bit_b << lev;
            lev++;
            if (lev == 14 || st.ber) { o.bfi = 1; return; }
        } else {
            if (bit_a) a = -a;
            if (bit_b) bb = -bb;
            nnz += nb_a + nb_b;
            c = ((c & 0xf) << 4) + (lev == 0 ? a1 + b1 + 1 : (lev == 1 ? ((a1 + b1) << 1) + 1 : imin(lev, 3) + 12));
            if (lsbMode && lev > 0) lsb_ind[lsb_cnt++] = k;
            x[k] = (i16)a; x[k + 1] = (i16)bb;
            const i32 aa = a < 0 ? -a : a, ab = bb < 0 ? -bb : bb;
            seed = l_mac0(l_mac0(seed, aa, k), ab, k + 1);
            mx = imax(mx, imax(aa, ab));
            if (st.bp - rd_bp(r) > 3 || st.ber || rd_oob(r)) { o.bfi = 1; return; }
            k += 2; lev = 0; a = 0; bb = 0;
            t = c + rateFlag;
            if (k > nt_half) t += 1 << 8;
        }
    }
    for (; k < yLen; k++) x[k] = 0;

    const i32 nbits_side = r.nread; // = nbbits - (8 * bp_side + norm_s(mask_side) - 6) of the reference
    const i32 extra_bits = clz32(st.range) - 6; // norm_ul
    const i32 nbits_ari = (st.bp - 3) << 3;
    i32 n = nbbits - (nbits_ari + extra_bits + nbits_side);
    if (n < 0) { o.bfi = 1; return; }

    for (int w = 0; w < 13; w++) res[w] = 0;
    if (!lsbMode) { // one residual bit per non-zero coefficient, in bin order, while bits remain (:991-1014)
        const i32 cnt = imin(n, nnz);
        for (i32 q = 0; q < cnt; q += 32) res[q >> 5] = rd_take(r, imin(32, cnt - q));
        o.resBits = cnt;
    } else {
        for (int q = 0; q < lsb_cnt; q++) {
            const int p = lsb_ind[q];
            a = x[p];
            if (n == 0) break;
            i32 bit = rd_bit(r);
            n--;
            if (bit) {
                if (a > 0) a++;
                else if (a < 0) a--;
                else {
                    if (n == 0) break;
                    a = 1;
                    if (rd_bit(r)) a = -a;
                    n--;
                }
            }
            x[p] = (i16)a;
            bb = x[p + 1];
            if (n == 0) break;
            bit = rd_bit(r);
            n--;
            if (bit) {
                if (bb > 0) bb++;
                else if (bb < 0) bb--;
                else {
                    if (n == 0) break;
                    bb = 1;
                    if (rd_bit(r)) bb = -bb;
                    n--;
                }
            }
            x[p + 1] = (i16)bb;
        }
        // the low bit plane changed the values: noise-filling seed and block maximum from the finished bins
        seed = 0; mx = 0;
        for (int i0 = 0; i0 < yLen; i0 += 8) {
            i32 v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = i0 + j < yLen ? (i32)x[i0 + j] : 0;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const i32 av = s_abs(v[j]);
                seed = l_mac0(seed, av, i0 + j);
                mx = imax(mx, av);
            }
        }
    }
    if (rd_oob(r)) { o.bfi = 1; return; }
    o.nf_seed = sx16(seed);
    o.x_absmax = mx;
    o.zero_frame = (si.lastnz == 2 && x[0] == 0 && x[1] == 0 && si.gg_idx == 0 && si.fac_ns_idx == 7) ? 1 : 0;
}

// ---- SNS vector decoding
// a_bwd_fx (fx/pvq_index_fx.c:90-120): one step back of the offset recurrence, in place
LC3_HD void mpvq_a_bwd(u32 *a, int n_items)
{
    u32 a0 = 0;
    int i;
    for (i = 1; i <= n_items; i++) {
        const u32 a1 = a[i] - a0 - a[i - 1];
        a[i - 1] = a0;
        a0 = a1;
    }
    a[i - 1] = a0;
}
// pvq_dec_deidx_fx + mpvq_deindex_fx + mind2vec_fx (fx/pvq_dec_fx.c:15-52, fx/pvq_index_fx.c:202-345)
LC3_HD void pvq_deindex(i16 *y, int k_val, int dim, i32 ls_ind, u32 index)
{
    u32 h[24];
    const u32 *offs = dim == 16 ? rom_mpvq_offs16 : (dim == 10 ? rom_mpvq_offs10 : rom_mpvq_offs6);
    const int kmax_tab = dim == 16 ? 12 : (dim == 10 ? 22 : 2);
    const int kp1 = k_val + 1;
    for (int i = 0; i <= kp1; i++) h[i] = offs[i];
    if (k_val != kmax_tab) h[kp1] >>= 1;
    const u32 size = 1u + h[kp1] + (h[k_val] >> 1);
    if (index >= size) index = 0;
    for (int i = 0; i < dim; i++) y[i] = 0;
    i32 sign = ls_ind != 0 ? -1 : 1;
    if (k_val == 0) return;
    int kmax = k_val, kacc = k_val, kdelta = 0;
    for (int pos = 0; pos < dim; pos++) {
        if (index != 0) {
            kacc = kmax;
            u32 d = index - h[kacc];
            bool borrow = index < h[kacc];
            while (borrow) { --kacc; d = index - h[kacc]; borrow = index < h[kacc]; }
            index = d;
            kdelta = kmax - kacc;
        } else {
            y[pos] = (i16)(sign < 0 ? -kmax : kmax);
            break;
        }
        if (kdelta != 0) {
            y[pos] = (i16)(sign < 0 ? -kdelta : kdelta);
            sign = (index & 1u) ? -1 : 1;
            index >>= 1;
            kmax -= kdelta;
        }
        mpvq_a_bwd(h, kmax + 1);
    }
}
// processSnsQuantizeScfDecoder_fx (fx/sns_quantize_scf_fx.c:442-570); idx = L_scf_idx[0..6]
LC3_HD void sns_decode(const i32 *idx, i16 *scf_q)
{
    for (int i = 0; i < 8; i++) { scf_q[i] = rom_sns_cb_lf[idx[0] * 8 + i]; scf_q[8 + i] = rom_sns_cb_hf[idx[1] * 8 + i]; }
    const int submode = idx[2];
    const i32 gain = rom_sns_gain[submode * 8 + idx[3]];
    i16 pulses[16], en1[16], adj[16];
    const i32 maxK = rom_sns_kval[submode * 2];
    if (submode >> 1) {
        pvq_deindex(pulses, maxK, 16, idx[4], (u32)idx[5]);
    } else {
        pvq_deindex(pulses, maxK, 10, idx[4], (u32)idx[5]);
        if ((submode & 1) == 0) {
            const i32 idxB = idx[6];
            pvq_deindex(pulses + 10, rom_sns_kval[submode * 2 + 1], 6, idxB & 1, (u32)(idxB >> 1));
        } else {
            for (int i = 10; i < 16; i++) pulses[i] = 0;
        }
    }
    pvq_unit_norm(en1, pulses, maxK);
    idct16(en1, adj);
    for (int i = 0; i < 16; i++) scf_q[i] = (i16)s_add(scf_q[i], s_mult_r(gain, adj[i]));
}

// IIRLattice (fx/tns_decoder_fx.c:117-134)
LC3_HD i32 tns_iir(int order, const i16 *rc, i32 *state, i32 x)
{
    // rc: rom_tns_quant_pts entries, none of them -32768 (tests/test_rom_bounds.py): Mpy_32_16 cannot saturate
    x = l_sub(x, mpy32x16k(state[order - 1], rc[order - 1]));
    for (int i = order - 2; i >= 0; i--) {
        x = l_sub(x, mpy32x16k(state[i], rc[i]));
        state[i + 1] = l_add(state[i], mpy32x16k(x, rc[i]));
    }
    state[0] = x;
    return x;
}

// One channel-frame from payload to shaped spectrum.  bytes: accessor of the frame's payload (get(i)).
template <class B> LC3_HD void dec_frame_body(const LC3D_Batch &b, int f, int tile, int lane, const AriTabs &T, B bytes)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const size_t ft = dframe_tile(b, f, tile);
    SV<const i32> cfg = tile_view(b.cfg, (size_t)tile, DCFG_N, lane);
    SV<i32> dout = tile_view(b.dout, ft, DO_N, lane);
    SV<i16> xq = tile_view(b.xq, ft, b.N, lane);
    SV<i32> d = tile_view(b.spec, ft, b.N, lane);
    const int N = b.N, yLen = b.yLen, dms = b.dms;
    const i32 nbytes = cfg[DCFG_BYTES], nbbits = nbytes << 3;

    DecSide si;
    si.bw_idx = b.fs_idx; si.nfilt = 1; si.lsbMode = 0; si.lastnz = 2; si.gg_idx = 0; si.fac_ns_idx = 0;
    si.tns_order[0] = si.tns_order[1] = 0; si.ltpf[0] = si.ltpf[1] = si.ltpf[2] = 0;
    for (int i = 0; i < 7; i++) si.scf[i] = 0;
    BackReader<B> r;
    rd_init(r, bytes, nbytes);
    i32 bfi = nbytes <= 0 ? 1 : dec_side_info(r, yLen, b.fs_idx, b.bw_bits, dms, si);
    AriOut ao;
    u32 res[13];
    ao.bfi = 0;
    if (!bfi) {
        ari_decode(bytes, r, nbbits, yLen, b.fs_idx, cfg[DCFG_LPC_WEIGHTING], dms, si, T, xq, d, res, ao);
        bfi = ao.bfi;
    }
    if (b.bfi) b.bfi[(size_t)(b.frame0 + f) * b.n_lanes + li] = (u8)bfi;
    dout[DO_BFI] = bfi;
    if (bfi) { // no concealment here: silence goes through the synthesis so that the stream's memories stay consistent
        for (int i = 0; i < N; i++) d[i] = 0;
        dout[DO_Q_EXP] = 0;
        dout[DO_LTPF0] = 0; dout[DO_LTPF1] = 0; dout[DO_LTPF2] = 0;
        return;
    }
    dout[DO_LTPF0] = si.ltpf[0]; dout[DO_LTPF1] = si.ltpf[1]; dout[DO_LTPF2] = si.ltpf[2];

    // ---- scaling + residual decoding + noise filling + global gain in one walk over the bins
    const i32 shift = ao.x_absmax == 0 ? 15 : s_norm(ao.x_absmax);
    i32 q_exp = 15 - shift;
    const i32 rs = q_exp - 1;
    const i32 fac_m = l_shr(0xC000000, rs), fac_p = l_shr(0x14000000, rs);
    int nf_w, nf_start, nbw = rom_bw_cutoff_bin[si.bw_idx];
    if (dms == 25) { nbw >>= 2; nf_w = 1; nf_start = 6; }
    else if (dms == 50) { nbw >>= 1; nf_w = 1; nf_start = 12; }
    else { nf_w = 3; nf_start = 24; }
    const i32 nf_val = l_shr((8 - si.fac_ns_idx) << 11, q_exp - 16);
    i32 nfseed = ao.nf_seed, nzeros = -2 * nf_w - 1;
    const bool nf_on = !ao.zero_frame;
    // global gain (fx/apply_global_gain_fx.c:29-31)
    const i32 gt = l_shl_p((si.gg_idx + cfg[DCFG_QGAIN_OFF]) * 0x797D, 7);
    const i32 gg_e = s_add(sx16(gt >> 25), 1);
    const i32 gg = l_round(m_invlog2(gt | (i32)0xFE000000));
    // TNS split point for the two headroom scans (fx/tns_decoder_fx.c:33-70)
    int tns_start0 = dms == 25 ? 3 : (dms == 50 ? 6 : 12), bw_stop = rom_bw_cutoff_bin[si.bw_idx];
    if (dms == 25) bw_stop >>= 2;
    if (dms == 50) bw_stop >>= 1;
    const int tns_nf = (si.bw_idx >= 3 && dms >= 50) ? 2 : 1;
    const int tns_start1 = bw_stop >> 1;
    const bool tns_on = si.tns_order[0] > 0 || (tns_nf == 2 && si.tns_order[1] > 0);
    const int tf = (tns_nf == 2 && si.tns_order[0] == 0) ? tns_start1 : tns_start0;
    Range r1, r2; range_init(r1); range_init(r2);
    i32 bits = 0;
    if (nf_on)
        for (int k = nf_start - nf_w; k < nf_start + nf_w; k++) { if (xq[k] != 0) nzeros = -2 * nf_w - 1; else nzeros++; }
    // eight bins at a time, their loads (and the noise-filling look-ahead) issued before the first is used: the walk waits
    // on the latency of its loads, not on arithmetic
    i32 qn[8]; // the group after the current one, loaded while the current one is processed
#pragma unroll
    for (int j = 0; j < 8; j++) qn[j] = j < yLen ? (i32)xq[j] : 0;
    for (int i0 = 0; i0 < yLen; i0 += 8) {
        i32 qv[8 + 3];
#pragma unroll
        for (int j = 0; j < 8; j++) qv[j] = qn[j];
#pragma unroll
        for (int j = 0; j < 8; j++) qn[j] = i0 + 8 + j < yLen ? (i32)xq[i0 + 8 + j] : 0;
#pragma unroll
        for (int j = 0; j < 3; j++) qv[8 + j] = qn[j];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int i = i0 + j;
            if (i >= yLen) break;
            const i32 q = qv[j];
            i32 v = (i32)((u32)q << (16 + shift));
            if (q != 0 && bits < ao.resBits) {
                if (res_bit(res, bits) == 0) v = l_sub(v, q > 0 ? fac_m : fac_p);
                else v = l_add(v, q > 0 ? fac_p : fac_m);
                bits++;
            }
            if (nf_on && i >= nf_start && i < nbw) {
                if (i < nbw - nf_w) { if ((nf_w == 3 ? qv[j + 3] : qv[j + 1]) != 0) nzeros = -2 * nf_w - 1; else nzeros++; }
                else nzeros++;
                if (nzeros >= 0) {
                    nfseed = sx16(13849 + nfseed * 31821);
                    v = nfseed >= 0 ? nf_val : l_neg(nf_val);
                }
            }
            v = mpy32x16k(v, gg); // gg >= 0 (mantissa of 2^x)
            if (i < tf) range_add(r1, v); else range_add(r2, v);
            d[i] = v;
        }
    }
    q_exp = s_add(q_exp, gg_e);

    // ---- TNS synthesis + SNS shaping in a second walk
    i32 ts = 0;
    if (tns_on) {
        ts = imin(range_headroom32(r1, 0), s_sub(range_headroom32(r2, 0), 7));
        q_exp = s_sub(q_exp, ts);
    }
    i16 scf_q[16], g[64], g_exp[64];
    sns_decode(si.scf, scf_q);
    sns_interp<false>(scf_q, g, g_exp, b.nbands);
    i32 emax = g_exp[0];
    for (int i = 1; i < b.nbands; i++) emax = imax(emax, g_exp[i]);
    for (int i = 0; i < b.nbands; i++) g_exp[i] = (i16)s_sub(g_exp[i], emax);
    q_exp = s_add(q_exp, emax);
    const i16 *bo = &rom_bands_offset[b.bands_off];
    const int shaped = bo[b.nbands];
    i16 rc[2][8];
    i32 state[8];
    for (int i = 0; i < 8; i++) { state[i] = 0; rc[0][i] = rom_tns_quant_pts[ao.tns_idx[i]]; rc[1][i] = rom_tns_quant_pts[ao.tns_idx[8 + i]]; }
    // bins of filter j: [lo_j, hi_j)
    const int lo0 = tns_start0, hi0 = tns_nf == 2 ? tns_start1 : bw_stop, lo1 = tns_start1, hi1 = bw_stop;
    int band = 0, stop = bo[1];
    i32 gsh = imin(-(i32)g_exp[0], 31), gcur = g[0]; // the band exponents are <= 0 after the common one went into q_exp: plain right shifts
    const ShlPlan pts = shl_plan(ts);
    Range ry; range_init(ry);
    i32 dn[8];
#pragma unroll
    for (int j = 0; j < 8; j++) dn[j] = j < yLen ? d[j] : 0;
    for (int i0 = 0; i0 < yLen; i0 += 8) {
        i32 dv[8];
#pragma unroll
        for (int j = 0; j < 8; j++) dv[j] = dn[j];
#pragma unroll
        for (int j = 0; j < 8; j++) dn[j] = i0 + 8 + j < yLen ? d[i0 + 8 + j] : 0; // not yet written by this walk: it stores at i < i0 + 8
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int i = i0 + j;
            if (i >= yLen) break;
            i32 v = dv[j];
            if (tns_on) {
                v = l_shl_plan(v, pts);
                if (si.tns_order[0] > 0 && i >= lo0 && i < hi0) v = tns_iir(si.tns_order[0], rc[0], state, v);
                else if (tns_nf == 2 && si.tns_order[1] > 0 && i >= lo1 && i < hi1) v = tns_iir(si.tns_order[1], rc[1], state, v);
            }
            if (i < shaped) {
                while (i >= stop) { band++; stop = bo[band + 1]; gsh = imin(-(i32)g_exp[band], 31); gcur = g[band]; }
                v = mpy32x16k(v, gcur) >> gsh; // g > 0 (mantissa of 2^x)
            }
            range_add(ry, v);
            d[i] = v;
        }
    }
    for (int i = yLen; i < N; i++) d[i] = 0; // fx/imdct_fx.c:72-73
    dout[DO_Q_EXP] = q_exp;
    dout[DO_Y_S] = range_headroom32(ry, 1); // getScaleFactor32_0 of the frame (fx/imdct_fx.c:76), consumed by the DCT pass
}

} // namespace lc3k
