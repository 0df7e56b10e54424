// TEST TOOLING: equivalence of the arithmetic shortcuts in audio_codec_b200/csrc with the operator-by-operator
// BASOP forms they replace (the forms the oracle and the reference use).  Exhaustive where the input domain is
// small, edge values + a large random sample elsewhere.  Prints "ok <name>" per check; exit code 1 on a mismatch.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#define LC3_FFT_TRACE 1
#include <vector>
#include <cmath>
#include "lc3_math.cuh"
#include "lc3_fft.cuh"
#include "lc3_rate.cuh"
#include "lc3_pitch.cuh"

using namespace lc3k;

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint32_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return (uint32_t)(rng_state >> 16); }

// ---- operator-by-operator forms
static i32 quant_bin_ref(i32 x, i32 inv_gain, i32 s)
{ // quantize_spec_fx.c:53-74
    i32 v = l_shl(mpy32x16(l_abs(x), inv_gain), s);
    i32 q = s_mac_r(v, -4096, 1);
    return x < 0 ? s_neg(q) : q;
}
static i32 log2_poly_ref(i32 f)
{ // basop_util.c:25-86
    i32 acc = l_mult(f, rom_ld_coeff[0]);
    i32 p = s_mult(f, f);
    for (int k = 1; k < 7; k++) {
        acc = l_mac(acc, p, rom_ld_coeff[k]);
        if (k < 6) p = s_mult(p, f);
    }
    acc = l_mac0(acc >> 1, hi16(acc), 14506);
    return acc >> 5;
}

// operator-by-operator form of normcorr_q15 (fx/olpa_fx.c:126-155)
static i32 normcorr_ref(i32 sum0, i32 sum1, i32 sum2, int sum1_prenorm, i32 sc1_in)
{
    i32 sc1 = sc1_in, sc2;
    if (!sum1_prenorm) { sc1 = l_norm(sum1); sum1 = l_shl(sum1, sc1); }
    sc2 = l_norm(sum2);
    sum2 = l_shl(sum2, sc2);
    i32 prod = mpy32x32(sum1, sum2);
    i32 sh = l_norm(prod);
    prod = l_shl(prod, sh);
    i32 pe = s_sub(62, s_add(s_add(sc1, sc2), sh));
    i32 inv = m_isqrt32(prod, &pe);
    i32 sc0 = l_norm(sum0);
    sum0 = l_shl(sum0, sc0);
    prod = mpy32x32(sum0, inv);
    pe = s_add(s_sub(31, sc0), pe);
    if (prod == 0 || s_sub(l_norm(prod), pe) >= 0) return s_max(0, l_round(l_shl(prod, pe)));
    return 32767;
}

static int fails = 0;
#define CHECK(name, cond, ...) do { if (!(cond)) { if (fails < 10) { printf("MISMATCH %s: ", name); printf(__VA_ARGS__); printf("\n"); } fails++; } } while (0)

// ---- FFT network: no addition can leave the 32-bit range when every input component is bounded by m_in.
// The network is linear and data independent (no branches on data), so operation k of every run computes the same
// linear form of the 2*LEN inputs.  Basis runs (one input = amp, the rest 0) give the magnitude of each coefficient;
// the largest value operation k can take over ALL admissible inputs is sum_i |coef_ki| * m_in, plus rounding.
// Rounding: every shift / fractional multiply truncates by < 1 unit; the measured deviation from linearity
// (run at 2*amp against twice the run at amp) bounds what accumulates through the network.
template <int LEN> static void fft_bound(double m_in)
{
    static i32 x[2 * LEN], w[2 * LEN];
    const i64 amp = 1 << 27;
    FftTrace::acc = nullptr; FftTrace::n = 0;
    for (int j = 0; j < 2 * LEN; j++) x[j] = 0;
    cfft<LEN>(x, w);
    const int n_ops = FftTrace::n;
    std::vector<i64> acc(n_ops, 0), one(n_ops), two(n_ops);
    i64 nonlin = 0;
    for (int i = 0; i < 2 * LEN; i++) {
        for (int rep = 0; rep < 2; rep++) {
            for (int j = 0; j < 2 * LEN; j++) x[j] = 0;
            x[i] = (i32)(amp << rep);
            std::vector<i64> &dst = rep ? two : one;
            std::fill(dst.begin(), dst.end(), 0);
            FftTrace::acc = dst.data(); FftTrace::cap = n_ops; FftTrace::n = 0;
            cfft<LEN>(x, w);
            CHECK("fft ops data independent", FftTrace::n == n_ops, "LEN=%d", LEN);
        }
        for (int k = 0; k < n_ops; k++) {
            acc[k] += one[k];
            const i64 d = two[k] - 2 * one[k];
            nonlin = std::max<i64>(nonlin, d < 0 ? -d : d);
        }
    }
    FftTrace::acc = nullptr;
    // rounding slack: per basis run the recorded |value| is within `nonlin` of |coef| * amp, so the sum over 2*LEN runs is
    // within 2*LEN*nonlin; a real run deviates from its linear form by at most about the same amount again.
    double worst = 0;
    for (int k = 0; k < n_ops; k++) worst = std::max(worst, (double)acc[k] / (double)amp);
    const double slack = (2.0 * LEN * (double)(nonlin + 1)) * m_in / (double)amp + 4096.0 * (double)(nonlin + 1);
    const double bound = worst * m_in + slack;
    printf("fft%-3d ops %5d  max sum|coef| %.4f  nonlinearity %lld  bound %.4f * 2^31\n", LEN, n_ops, worst, (long long)nonlin,
           bound / 2147483648.0);
    if (FftNeedsSat<LEN>::v) printf("       -> this length keeps the saturating operators (FftNeedsSat)\n");
    CHECK("fft bound", FftNeedsSat<LEN>::v || bound < 2147483647.0, "LEN=%d bound=%.0f", LEN, bound);
}


int main(int argc, char **argv)
{
    const long n_rand = argc > 1 ? atol(argv[1]) : 20000000;

    // quant_bin: inv_gain is an Inv16 output (positive 16-bit), s any plausible shift and beyond
    {
        const i32 edge_x[] = {0, 1, -1, 2, -2, 0x7fff, 0x8000, -0x8000, 0x7fffffff, LC3_I32_MIN, LC3_I32_MIN + 1, 0x40000000, -0x40000000,
                              0x3fffffff, 12345678, -12345678, 0x7fffa000, 0x7fff9fff, 0x00010000, 0x0000ffff};
        const i32 edge_g[] = {1, 2, 16384, 16385, 23170, 32766, 32767, 8192, 100};
        for (i32 x : edge_x) for (i32 g : edge_g) for (i32 s = -40; s <= 40; s++)
            CHECK("quant_bin", quant_bin(x, g, s) == quant_bin_ref(x, g, s), "x=%d g=%d s=%d", x, g, s);
        for (long i = 0; i < n_rand; i++) {
            i32 x = (i32)rnd();
            if (i & 1) x >>= (rnd() % 31);
            const i32 g = 1 + (i32)(rnd() % 32767), s = (i32)(rnd() % 56) - 36;
            CHECK("quant_bin", quant_bin(x, g, s) == quant_bin_ref(x, g, s), "x=%d g=%d s=%d", x, g, s);
        }
        printf("ok quant_bin\n");
    }
    // m_log2 polynomial: exhaustive over its whole domain
    {
        for (i32 f = 0; f < 16384; f++) CHECK("m_log2_poly", m_log2_poly(f) == log2_poly_ref(f), "f=%d", f);
        printf("ok m_log2_poly\n");
    }
    // round_fx
    {
        auto ref = [](i32 a) { return l_add(a, 0x8000) >> 16; };
        const bool full = argc > 2;
        if (full) {
            for (int64_t a = -2147483648LL; a <= 2147483647LL; a++) CHECK("l_round", l_round((i32)a) == ref((i32)a), "a=%d", (i32)a);
        } else {
            for (int64_t c : {-2147483648LL, -65536LL, 0LL, 65536LL, 2147418112LL, 2147483647LL})
                for (int64_t a = std::max<int64_t>(c - 200000, -2147483648LL); a <= std::min<int64_t>(c + 200000, 2147483647LL); a++)
                    CHECK("l_round", l_round((i32)a) == ref((i32)a), "a=%d", (i32)a);
            for (long i = 0; i < n_rand; i++) { const i32 a = (i32)rnd(); CHECK("l_round", l_round(a) == ref(a), "a=%d", a); }
        }
        printf("ok l_round%s\n", full ? " (all 2^32 inputs)" : "");
    }
    // L_shl with a planned count
    {
        const i32 edge[] = {0, 1, -1, 2, -2, 3, -3, 0x7fffffff, LC3_I32_MIN, LC3_I32_MIN + 1, 0x40000000, -0x40000000, 0x3fffffff, -0x40000001,
                            0x7fff, -0x8000, 0x10000, -0x10000, 123456789, -123456789};
        for (i32 a : edge) for (i32 n = -70; n <= 70; n++)
            CHECK("l_shl_plan", l_shl_plan(a, shl_plan(n)) == l_shl(a, n), "a=%d n=%d", a, n);
        for (long i = 0; i < n_rand; i++) {
            i32 a = (i32)rnd();
            if (i & 1) a >>= (rnd() % 32);
            const i32 n = (i32)(rnd() % 80) - 40;
            CHECK("l_shl_plan", l_shl_plan(a, shl_plan(n)) == l_shl(a, n), "a=%d n=%d", a, n);
        }
        printf("ok l_shl_plan\n");
    }
    // FFT inputs are cmul_half() outputs: |re|, |im| <= 2^31 * max(|c| + |d|) / 2^16 + 2 over the pre-twiddle table
    {
        i32 mx = 0;
        for (int i = 0; i + 1 < 1520; i += 2) mx = imax(mx, (i32)abs(rom_dct4_twiddle[i]) + (i32)abs(rom_dct4_twiddle[i + 1]));
        i32 mx2 = 0; // post-twiddle tables (their outputs are only negated / stored)
        for (int i = 0; i + 1 < 482; i += 2) mx2 = imax(mx2, (i32)abs(rom_sin480[i]) + (i32)abs(rom_sin480[i + 1]));
        for (int i = 0; i + 1 < 322; i += 2) mx2 = imax(mx2, (i32)abs(rom_sin320[i]) + (i32)abs(rom_sin320[i + 1]));
        CHECK("twiddle sums", mx <= 46341 && mx2 <= 46341, "mx=%d mx2=%d", mx, mx2);
        const double m_in = 2147483648.0 * mx / 65536.0 + 2.0;
        printf("pre-twiddle max |c|+|d| = %d -> FFT input bound %.5f * 2^31\n", mx, m_in / 2147483648.0);
        fft_bound<10>(m_in); fft_bound<20>(m_in); fft_bound<30>(m_in); fft_bound<40>(m_in); fft_bound<60>(m_in);
        fft_bound<80>(m_in); fft_bound<120>(m_in); fft_bound<160>(m_in); fft_bound<240>(m_in);
        printf("ok fft_bound\n");
    }
    // rate-loop bisection (lc3_rate_warp.cuh): the quadruple energy written by gain_energies() is
    // (mpy32x16k(t, 0x436E) >> 6) + 0x9CCCD, monotone in t, so its range is set by the two extremes of t
    {
        const i32 lo = l_add(mpy32x16k(LC3_I32_MIN, 0x436E) >> 6, 0x9CCCD), hi = l_add(mpy32x16k(LC3_I32_MAX, 0x436E) >> 6, 0x9CCCD);
        CHECK("rate_bounds", lo > -(1 << 25) && hi < (1 << 25), "lo=%d hi=%d", lo, hi);
        // offset = (255 + qgain_off) << 16 with 255 + qgain_off in [0, 255], moved by at most 255 << 16 either way
        const int64_t off_max = (int64_t)(255 + 255) << 16, t_max = (int64_t)hi + off_max, t_min = (int64_t)lo - off_max;
        CHECK("rate_bounds", t_max < (1 << 26) && t_min > -(1 << 26), "t range");
        CHECK("rate_bounds", 4 * (2 * t_max) < 2147483647LL, "lane sum"); // four quadruples per lane, each t + (t - 0x460000)
        printf("ok rate_bounds en in [%d, %d]\n", lo, hi);
    }
    { // normcorr_q15: energies >= 1 of every magnitude, correlations of either sign, pre-normalised first energy too
        const i32 edge[] = {1, 2, 3, 0x7fffffff, 0x40000000, 0x3fffffff, 0x10000, 0xffff, 12345};
        const i32 edge0[] = {0, 1, -1, LC3_I32_MIN, LC3_I32_MAX, 0x40000000, -0x40000000, 77, -77};
        for (i32 a : edge0) for (i32 b : edge) for (i32 c : edge) {
            CHECK("normcorr_q15", normcorr_q15(a, b, c, 0, 0) == normcorr_ref(a, b, c, 0, 0), "%d %d %d", a, b, c);
            const i32 n1 = l_norm(b);
            CHECK("normcorr_q15", normcorr_q15(a, l_shl(b, n1), c, 1, n1) == normcorr_ref(a, l_shl(b, n1), c, 1, n1), "pre %d %d %d", a, b, c);
        }
        for (long it = 0; it < 20000000; it++) {
            const int s1 = rnd() % 31, s2 = rnd() % 31, s0 = rnd() % 32;
            const i32 b = (i32)((rnd() << 16 ^ rnd()) & 0x7fffffff) >> s1 | 1, c = (i32)((rnd() << 16 ^ rnd()) & 0x7fffffff) >> s2 | 1;
            const i32 a = (i32)(rnd() << 16 ^ rnd()) >> s0;
            const int pre = it & 1;
            const i32 n1 = l_norm(b);
            const i32 got = pre ? normcorr_q15(a, l_shl(b, n1), c, 1, n1) : normcorr_q15(a, b, c, 0, 0);
            const i32 want = pre ? normcorr_ref(a, l_shl(b, n1), c, 1, n1) : normcorr_ref(a, b, c, 0, 0);
            CHECK("normcorr_q15", got == want, "%d %d %d pre %d: %d vs %d", a, b, c, pre, got, want);
        }
        printf("ok normcorr_q15\n");
    }
    if (fails) { printf("%d mismatches\n", fails); return 1; }
    return 0;
}
