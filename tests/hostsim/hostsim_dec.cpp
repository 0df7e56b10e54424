// TEST TOOLING (not product code): the decoder's kernel bodies (audio_codec_b200/csrc/lc3_dec_*.cuh) compiled for the host
// and run lane by lane, so that they can be checked against the compiled reference decoder in a container without a GPU.
// The product library never contains or calls this.
#include <cstdlib>
#include <cstring>
#include <vector>
extern "C" {
#include "lc3_dec_config.h"
}
#include "lc3_dec_frame.cuh"
#include "lc3_dec_idct.cuh"
#include "lc3_dec_synth.cuh"

using namespace lc3k;

struct HostBytes {
    const u8 *p;
    int n;
    i32 get(int i) const { return p[i]; }
    u32 word(int k) const { u32 v = 0; const int m = n - 4 * k; if (m > 0) memcpy(&v, p + 4 * k, m < 4 ? m : 4); return v; } // never past the frame
};

struct DSim {
    LC3DCfg cfg;
    int n_streams, n_slots;
    std::vector<int32_t> slot, st32;
    std::vector<int16_t> st16;
};

extern "C" void *hostsim_dec_new(int fs, int dms, int n_streams)
{
    DSim *s = new DSim;
    if (lc3dcfg_init(&s->cfg, fs, 1) || lc3dcfg_set_frame_dms(&s->cfg, dms)) { delete s; return 0; }
    s->n_streams = n_streams;
    s->n_slots = (n_streams + 31) / 32 * 32;
    s->slot.resize(n_streams);
    for (int i = 0; i < n_streams; i++) s->slot[i] = i;
    s->st32.assign((size_t)s->n_slots * DST32_N, 0);
    s->st16.assign((size_t)s->n_slots * DST16_N, 0);
    for (int i = 0; i < s->n_slots; i++) slot_view(s->st32.data(), i, DST32_N)[DST_LTPF_SCALE_IDX] = -1;
    return s;
}
extern "C" void hostsim_dec_free(void *p) { delete (DSim *)p; }
extern "C" int hostsim_dec_frame_len(void *p) { return ((DSim *)p)->cfg.N; }

template <int N> static void run_idct_tile(const LC3D_Batch &b, int f, int tile)
{
    typedef MdctTile<N> G;
    static std::vector<i32> sm(G::smem_bytes / sizeof(i32));
    i32 *X = sm.data(), *mx = X + 32 * G::S, *mn = mx + G::PARTS * 32;
    const MtTabs T = mt_rom_tabs<N>();
    for (int t = 0; t < G::THREADS; t++) mt_dec_load<N>(b, f, tile, t, X, mx, mn);
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; for (int j = 0; j < 4; j++) mt_pre<N>(X, mx, mn, w + 8 * j, l, T); }
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; mt_leaf<N>(X, w + 8 * (l >> 3), l & 7); }
    for (int w = 0; w < 8; w++)
        for (int j = 0; j < 4; j++) {
            cpx o[32][8];
            for (int l = 0; l < G::LEAF; l++) mt_col_compute<N>(X, w + 8 * j, l, o[l], T);
            for (int l = 0; l < G::LEAF; l++) mt_col_store<N>(X, w + 8 * j, l, o[l]);
        }
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; for (int j = 0; j < 4; j++) mt_post<N>(X, w + 8 * j, l, T); }
    i32 red[64];
    for (int fr = 0; fr < 32; fr++) mt_dec_store<N>(b, f, tile, fr, X, red);
}
template <int N> static void run_idct_lane(const LC3D_Batch &b, int f, int tile)
{
    for (int l = 0; l < 32; l++) idct_lane_body<N>(b, f, tile, l);
    for (int l = 0; l < 32; l++) {
        const int li = tile * 32 + l;
        if (li >= b.n_lanes) continue;
        SV<const i32> d = tile_view((const i32 *)b.spec, dframe_tile(b, f, tile), N, l);
        i32 *row = b.yrow + dframe_row(b, f, li) * N;
        for (int e = 0; e < N; e++) row[e] = d[e];
    }
}

// in: [n_streams][n_frames][nbytes]; pcm: [n_streams][n_frames*N] (int16 or int32); bfi: [n_frames][n_streams] or NULL
extern "C" int hostsim_dec_decode(void *p, const uint8_t *in, int nbytes, int bits, int n_frames, void *pcm, uint8_t *bfi)
{
    DSim *s = (DSim *)p;
    LC3D_Batch b;
    memset(&b, 0, sizeof(b));
    lc3dcfg_fill_batch(&s->cfg, &b);
    const int N = b.N;
    b.bits = bits; b.n_lanes = s->n_streams; b.n_tiles = (s->n_streams + 31) / 32; b.n_frames = n_frames; b.max_bytes = nbytes;
    b.slot = s->slot.data(); b.st32 = s->st32.data(); b.st16 = s->st16.data();
    int32_t w[DCFG_N];
    int r = lc3dcfg_lane_words(&s->cfg, nbytes, w);
    if (r) return r;
    std::vector<int32_t> cfg((size_t)b.n_tiles * DCFG_N * 32, 0);
    for (int i = 0; i < s->n_slots; i++) for (int k = 0; k < DCFG_N; k++) tile_view(cfg.data(), i >> 5, DCFG_N, i & 31)[k] = w[k];
    b.cfg = cfg.data();
    std::vector<int64_t> in_off(s->n_streams), pcm_off(s->n_streams);
    std::vector<int32_t> in_stride(s->n_streams, nbytes);
    for (int i = 0; i < s->n_streams; i++) { in_off[i] = (int64_t)i * n_frames * nbytes; pcm_off[i] = (int64_t)i * n_frames * N; }
    b.in = in; b.in_off = in_off.data(); b.in_stride = in_stride.data(); b.pcm = pcm; b.pcm_off = pcm_off.data(); b.bfi = bfi;
    const size_t ftl = (size_t)n_frames * b.n_tiles * 32;
    std::vector<int16_t> xq(ftl * N);
    std::vector<int32_t> spec(ftl * N), dout(ftl * DO_N), yrow(ftl * N), work(ftl * N);
    b.xq = xq.data(); b.spec = spec.data(); b.dout = dout.data(); b.yrow = yrow.data(); b.work = work.data();

    static u32 ari_cf[1088];
    for (int i = 0; i < 1088; i++) ari_cf[i] = (u32)rom_ari_freq[i] << 16 | rom_ari_cumfreq[i];
    AriTabs T = {(const u8 *)rom_ari_lookup, ari_cf};
    for (int f = 0; f < n_frames; f++)
        for (int tile = 0; tile < b.n_tiles; tile++) {
            for (int l = 0; l < 32; l++) {
                const int li = tile * 32 + l;
                if (li >= b.n_lanes) continue;
                dec_frame_body(b, f, tile, l, T, HostBytes{in + in_off[li] + (size_t)f * nbytes, nbytes});
            }
            switch (N) {
            case 480: run_idct_tile<480>(b, f, tile); break;
            case 320: run_idct_tile<320>(b, f, tile); break;
            case 240: run_idct_tile<240>(b, f, tile); break;
            case 160: run_idct_tile<160>(b, f, tile); break;
            case 120: run_idct_lane<120>(b, f, tile); break;
            case 80: run_idct_lane<80>(b, f, tile); break;
            case 60: run_idct_lane<60>(b, f, tile); break;
            case 40: run_idct_lane<40>(b, f, tile); break;
            case 20: run_idct_lane<20>(b, f, tile); break;
            }
        }
    static SynthShared sh;
    for (int li = 0; li < b.n_lanes; li++) {
        if (bits == 16) synth_warp_body<16>(b, li, sh, b.yrow, &rom_win[b.W_off]);
        else synth_warp_body<24>(b, li, sh, b.yrow, &rom_win[b.W_off]);
    }
    return 0;
}
