/* ORACLE (test infrastructure, not product code).
 * Low-delay MDCT: windowing with time-domain aliasing fold (fx/mdct_fx.c:17-96), then a DCT-IV
 * done as pre-twiddle -> complex FFT of N/2 -> post-twiddle (fx/dct4_fx.c:16-138,
 * twiddle selection fx/rom_basop_util.c:590-670). */
#include "lc3o_internal.h"

static int tw_slot(int N)
{
    static const int sizes[9] = {20, 40, 60, 80, 120, 160, 240, 320, 480};
    for (int i = 0; i < 9; i++) if (sizes[i] == N) return i;
    return -1;
}

/* (a + jb)*(c + jd) with each product halved before the add (fx/basop_mpy.h:45-50) */
static inline void cmul_half(i32 *re, i32 *im, i32 a, i32 b, i16 c, i16 d)
{
    *re = l_sub(l_shr(mpy32x16(a, c), 1), l_shr(mpy32x16(b, d), 1));
    *im = l_add(l_shr(mpy32x16(a, d), 1), l_shr(mpy32x16(b, c), 1));
}

/* fx/dct4_fx.c:16-138 */
void o_dct4(i32 *p, i16 *p_e, int L)
{
    int M = L >> 1;
    const i16 *tw = &rom_dct4_twiddle[rom_dct4_twiddle_off[tw_slot(L)]];
    /* SineTable480 for lengths divisible by 3, SineTable320 otherwise; step = table_len*2/L */
    const i16 *sn = (L % 3 == 0) ? rom_sin480 : rom_sin320;
    int step = (L % 3 == 0) ? 960 / L : 640 / L;
    i32 a1, a2, a3, a4;

    for (int i = 0, lo = 0, hi = L - 2; i < M - 1; i += 2, lo += 2, hi -= 2) {
        cmul_half(&a1, &a2, p[hi + 1], p[lo], tw[2 * i], tw[2 * i + 1]);
        cmul_half(&a3, &a4, p[hi], p[lo + 1], tw[2 * i + 2], tw[2 * i + 3]);
        p[lo] = a2; p[lo + 1] = a1;
        p[hi] = a4; p[hi + 1] = l_neg(a3);
    }

    o_cfft(p, M, p_e);

    int lo = 0, hi = L - 2, idx = step;
    int M2 = ((M + 1) >> 1) - 1;
    cmul_half(&a3, &a4, p[hi], p[hi + 1], sn[2 * idx], sn[2 * idx + 1]);
    p[hi + 1] = l_neg(l_shr(p[lo + 1], 1));
    p[lo] = l_shr(p[lo], 1);
    for (int i = 1; i < M2; i++) {
        p[lo + 1] = a3; p[hi] = a4;
        lo += 2; hi -= 2;
        cmul_half(&a1, &a2, p[lo + 1], p[lo], sn[2 * idx], sn[2 * idx + 1]);
        idx += step;
        cmul_half(&a3, &a4, p[hi], p[hi + 1], sn[2 * idx], sn[2 * idx + 1]);
        p[hi + 1] = l_neg(a1);
        p[lo] = a2;
    }
    p[lo + 1] = a3; p[hi] = a4;
    lo += 2; hi -= 2;
    cmul_half(&a3, &a4, p[lo + 1], p[lo], sn[2 * idx], sn[2 * idx + 1]);
    a1 = l_shr(mpy32x16(p[hi], 23170), 1);
    a2 = l_shr(mpy32x16(p[hi + 1], 23170), 1);
    p[hi] = l_add(a1, a2);
    p[lo + 1] = l_sub(a1, a2);
    p[hi + 1] = l_neg(a3);
    p[lo] = a4;
    *p_e = s_add(*p_e, 2);
}

/* fx/mdct_fx.c:17-96.  mem[] holds the last memLen input samples and is updated. */
void o_mdct(const i16 *x, i16 x_exp, int N, const i16 *w, int wLen, i16 *mem, int memLen, i32 *y, i16 *y_e)
{
    i16 buf[LC3O_MAX_N];
    int z = 2 * N - wLen, m = N >> 1, i;

    memcpy(buf, mem, memLen * sizeof(i16));
    memcpy(buf + memLen, x, (N - memLen) * sizeof(i16));
    memcpy(mem, x + N - memLen, memLen * sizeof(i16));

    for (i = 0; i < m; i++)
        y[m + i] = l_msu0(l_mult0(buf[i], w[i]), buf[2 * m - 1 - i], w[2 * m - 1 - i]);
    for (i = 0; i < z; i++)
        y[m - 1 - i] = l_mult0(x[2 * m - memLen + i], w[2 * m + i]);
    for (; i < m; i++)
        y[m - 1 - i] = l_mac0(l_mult0(x[2 * m - memLen + i], w[2 * m + i]), x[4 * m - memLen - 1 - i], w[4 * m - 1 - i]);

    i16 s = s_max(0, o_headroom32(y, N, 0));
    for (i = 0; i < N; i++) y[i] = l_shl(y[i], s);
    *y_e = s_sub(s_sub(x_exp, 2), s);
    if (N <= 20) *y_e = s_add(*y_e, 2);
    else if (N <= 120) *y_e = s_add(*y_e, 1);
    o_dct4(y, y_e, N);
}
