// TEST TOOLING (not product code): compiles the kernel bodies of audio_codec_b200/csrc for the host and
// runs them lane by lane, so that the stage code can be debugged against the oracle in a container without
// a GPU.  The product library (liblc3plus_b200.so) never contains or calls this; it has the CUDA build only.
#include <cstdlib>
#include <cstring>
#include <vector>
extern "C" {
#include "lc3_config.h"
}
#include "lc3_body_frame.cuh"
#include "lc3_pitch_warp.cuh"
#include "lc3_rate_warp.cuh"
#include "lc3_quant_tile.cuh"

using namespace lc3k;

template <int BITS> static void run_front(const LC3K_Batch &b, int f, int tile, int lane)
{
    switch (b.N) {
    case 20: front_body<20, BITS>(b, f, tile, lane); break;
    case 40: front_body<40, BITS>(b, f, tile, lane); break;
    case 60: front_body<60, BITS>(b, f, tile, lane); break;
    case 80: front_body<80, BITS>(b, f, tile, lane); break;
    case 120: front_body<120, BITS>(b, f, tile, lane); break;
    case 160: front_body<160, BITS>(b, f, tile, lane); break;
    case 240: front_body<240, BITS>(b, f, tile, lane); break;
    case 320: front_body<320, BITS>(b, f, tile, lane); break;
    case 480: front_body<480, BITS>(b, f, tile, lane); break;
    }
}

// the tile MDCT kernel (k_mdct in lc3_k_front.cuh), phase by phase, each phase a plain loop over the block's threads
template <int N, int BITS> static void run_mdct_tile(const LC3K_Batch &b, int f, int tile)
{
    typedef MdctTile<N> G;
    static std::vector<i32> sm(G::smem_bytes / sizeof(i32));
    i32 *X = sm.data(), *mx = X + 32 * G::S, *mn = mx + G::PARTS * 32;
    const MtTabs T = mt_rom_tabs<N>();
    if (BITS == 16 && !b.stage_full) { // rows from the caller's PCM (the kernel's bulk copies are a memcpy here)
        for (int r = 0; r < 32; r++) {
            const int li = tile * 32 + r;
            if (li < b.n_lanes && f > 0) {
                const i16 *cur = (const i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
                memcpy(MtRowIn<N>::at(X, r), cur - N, 4 * N);
            }
            mt_fill_row<N>(b, f, tile, r, f > 0, X);
            mt_fold_row<N>(b, r, li < b.n_lanes, X, mx, mn, &rom_win[b.W_off]);
        }
    } else
    for (int t = 0; t < G::THREADS; t++) mt_fold<N, BITS>(b, f, tile, t, X, mx, mn);
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; for (int j = 0; j < 4; j++) mt_pre<N>(X, mx, mn, w + 8 * j, l, T); }
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; mt_leaf<N>(X, w + 8 * (l >> 3), l & 7); }
    for (int w = 0; w < 8; w++)
        for (int j = 0; j < 4; j++) {
            cpx o[32][8];
            for (int l = 0; l < G::LEAF; l++) mt_col_compute<N>(X, w + 8 * j, l, o[l], T);
            for (int l = 0; l < G::LEAF; l++) mt_col_store<N>(X, w + 8 * j, l, o[l]);
        }
    for (int t = 0; t < G::THREADS; t++) { const int w = t >> 5, l = t & 31; for (int j = 0; j < 4; j++) mt_post<N>(X, w + 8 * j, l, T); }
    i32 *smx = mn + G::PARTS * 32, *ben = smx + G::PARTS * 32, *bexp = ben + 32 * G::BROW;
    for (int t = 0; t < G::THREADS; t++) mt_bands<N>(b, t, X, ben, bexp, smx);
    for (int t = 0; t < G::THREADS; t++) mt_store<N, BITS>(b, f, tile, t, X, mx, mn, ben, bexp, smx);
}
template <int BITS> static void run_front_tile(const LC3K_Batch &b, int f, int tile)
{ // same split as front_launch_tiled / front_launch
    switch (b.N) {
    case 160: run_mdct_tile<160, BITS>(b, f, tile); break;
    case 240: run_mdct_tile<240, BITS>(b, f, tile); break;
    case 320: run_mdct_tile<320, BITS>(b, f, tile); break;
    case 480: run_mdct_tile<480, BITS>(b, f, tile); break;
    default: for (int l = 0; l < 32; l++) run_front<BITS>(b, f, tile, l); return;
    }
    for (int l = 0; l < 32; l++) sns_body(b, f, tile, l);
    for (int l = 0; l < 32; l++) {
        switch (b.N) {
        case 160: tns_body<160>(b, f, tile, l); break;
        case 240: tns_body<240>(b, f, tile, l); break;
        case 320: tns_body<320>(b, f, tile, l); break;
        case 480: tns_body<480>(b, f, tile, l); break;
        }
    }
}

struct Sim {
    LC3Cfg cfg;
    int n_streams, n_slots;
    std::vector<int32_t> slot, cfgw, st32;
    std::vector<int16_t> st16, hist16;
    std::vector<int32_t> hist32;
};

extern "C" void *hostsim_new(int fs, int dms, int n_streams)
{
    Sim *s = new Sim;
    if (lc3cfg_init(&s->cfg, fs, 1) || lc3cfg_set_frame_dms(&s->cfg, dms)) { delete s; return 0; }
    s->n_streams = n_streams;
    int nt = (n_streams + 31) / 32;
    s->n_slots = nt * 32;
    s->slot.resize(n_streams);
    for (int i = 0; i < n_streams; i++) s->slot[i] = i;
    s->cfgw.assign((size_t)s->n_slots * CFG_N, 0);
    s->st32.assign((size_t)s->n_slots * ST32_N, 0);
    s->st16.assign((size_t)s->n_slots * ST16_N, 0);
    s->hist16.assign((size_t)s->n_slots * LC3K_MAX_N, 0);
    s->hist32.assign((size_t)s->n_slots * LC3K_MAX_N, 0);
    for (int i = 0; i < s->n_slots; i++) {
        slot_view(s->st32.data(), i, ST32_N)[ST_X_EXP] = 15;
        slot_view(s->st32.data(), i, ST32_N)[ST_RESAMP_EXP] = 17;
    }
    return s;
}
extern "C" void hostsim_free(void *p) { delete (Sim *)p; }

extern "C" int hostsim_set_bitrate(void *p, int bitrate)
{
    Sim *s = (Sim *)p;
    int r = lc3cfg_set_bitrate(&s->cfg, bitrate);
    if (r) return r;
    for (int i = 0; i < s->n_slots; i++) {
        for (int k = 0; k < CFG_N; k++) slot_view(s->cfgw.data(), i, CFG_N)[k] = s->cfg.ch[0][k];
        if (s->cfg.attack_reset[0])
            for (int k = ST_ATT_F0; k <= ST_ATT_ACC; k++) slot_view(s->st32.data(), i, ST32_N)[k] = 0;
    }
    s->cfg.attack_reset[0] = 0;
    return 0;
}
extern "C" int hostsim_set_ep(void *p, int epmode, int epmr)
{ // error protection mode and mode request (lc3_enc_set_ep_mode / _request); call before hostsim_set_bitrate
    Sim *s = (Sim *)p;
    int r = lc3cfg_set_ep_mode(&s->cfg, epmode);
    if (!r) r = lc3cfg_set_ep_mode_request(&s->cfg, epmr);
    if (!r && s->cfg.br_set)
        for (int i = 0; i < s->n_slots; i++)
            for (int k = 0; k < CFG_N; k++) slot_view(s->cfgw.data(), i, CFG_N)[k] = s->cfg.ch[0][k];
    return r;
}
extern "C" int hostsim_frame_len(void *p) { return ((Sim *)p)->cfg.N; }
extern "C" int hostsim_num_bytes(void *p) { return lc3cfg_num_bytes(&((Sim *)p)->cfg); }

// pcm: [n_streams][n_frames*N] (int16 or int32), out: [n_streams][n_frames][nbytes]
extern "C" int hostsim_encode(void *p, const void *pcm, int bits, int n_frames, uint8_t *out, const uint8_t *bfi)
{
    Sim *s = (Sim *)p;
    LC3K_Batch b;
    memset(&b, 0, sizeof(b));
    lc3cfg_fill_batch(&s->cfg, &b);
    const int N = b.N, nb = lc3cfg_num_bytes(&s->cfg);
    b.bits = bits;
    b.stage_full = !(bits == 16 && N >= 160);
    b.n_lanes = s->n_streams; b.n_tiles = (s->n_streams + 31) / 32; b.n_frames = n_frames;
    b.slot = s->slot.data(); b.cfg = s->cfgw.data(); b.st32 = s->st32.data(); b.st16 = s->st16.data();
    b.hist16 = s->hist16.data(); b.hist32 = s->hist32.data();
    std::vector<int64_t> out_off(s->n_streams);
    std::vector<int32_t> out_stride(s->n_streams, nb);
    for (int i = 0; i < s->n_streams; i++) out_off[i] = (int64_t)i * n_frames * nb;
    b.out = out; b.out_off = out_off.data(); b.out_stride = out_stride.data(); b.bfi = bfi;
    const size_t ftl = (size_t)n_frames * b.n_tiles * 32;
    std::vector<int32_t> pcm_t((size_t)b.n_tiles * (n_frames + 1) * N * 32, 0); // int32 storage covers both widths
    std::vector<int32_t> pout(ftl * PO_N), spec(ftl * N), fout(ftl * FO_N), en4(ftl * (N / 4)), rout(ftl * RO_N);
    b.pcm_t = pcm_t.data(); b.pout = pout.data(); b.spec = spec.data(); b.fout = fout.data(); b.en4 = en4.data();
    b.rout = rout.data();
    std::vector<int16_t> xqbuf(ftl * N);
    b.xq = xqbuf.data();
    std::vector<int32_t> benbuf(ftl * BEN_N);
    b.ben = benbuf.data();

    // staging (what the stage kernel does): history prefix from the state, frames from the caller, new history
    for (int li = 0; li < s->n_streams; li++) {
        int tile = li >> 5, lane = li & 31, slot = s->slot[li];
        size_t per_tile = (size_t)(n_frames + 1) * N;
        if (bits == 16) {
            SV<i16> t = tile_view((i16 *)b.pcm_t, tile, per_tile, lane);
            SV<i16> h = slot_view(b.hist16, slot, LC3K_MAX_N);
            const int16_t *src = (const int16_t *)pcm + (size_t)li * n_frames * N;
            for (int i = 0; i < N; i++) t[i] = h[i];
            if (b.stage_full) for (int i = 0; i < n_frames * N; i++) t[N + i] = src[i];
            for (int i = 0; i < N; i++) h[i] = src[(n_frames - 1) * N + i];
        } else {
            SV<i32> t = tile_view((i32 *)b.pcm_t, tile, per_tile, lane);
            SV<i32> h = slot_view(b.hist32, slot, LC3K_MAX_N);
            const int32_t *src = (const int32_t *)pcm + (size_t)li * n_frames * N;
            for (int i = 0; i < N; i++) t[i] = h[i];
            for (int i = 0; i < n_frames * N; i++) t[N + i] = src[i];
            for (int i = 0; i < N; i++) h[i] = t[n_frames * N + i];
        }
    }
    static u32 ari_cf[1088];
    for (int i = 0; i < 1088; i++) ari_cf[i] = (u32)rom_ari_freq[i] << 16 | rom_ari_cumfreq[i];
    AriTabs T = {(const u8 *)rom_ari_lookup, ari_cf};
    static PitchShared psh;
    static RateShared rsh;
    b.pcm = pcm;
    std::vector<int64_t> pcm_off(s->n_streams);
    for (int i = 0; i < s->n_streams; i++) pcm_off[i] = (int64_t)i * n_frames * N;
    b.pcm_off = pcm_off.data();
    {
        std::vector<int16_t> r12(ftl * 128);
        b.r12 = r12.data();
        if (bits == 16) { // three passes as the kernels run them
            static FirShared fsh;
            uint32_t filtp[120];
            for (int i = 0; i < 120; i++) filtp[i] = fir_pack_taps(rom_resamp_filt[b.fs_idx * 240 + 2 * i], rom_resamp_filt[b.fs_idx * 240 + 2 * i + 1]);
            for (int f = 0; f < n_frames; f++)
                for (int li = 0; li < b.n_lanes; li++) fir_warp_body(b, li, f, fsh, &rom_resamp_filt[b.fs_idx * 240], filtp);
            for (int tile = 0; tile < b.n_tiles; tile++)
                for (int lane = 0; lane < 32; lane++) scalar_body(b, tile, lane);
            for (int li = 0; li < b.n_lanes; li++) pitch_warp_body<16, true>(b, li, psh);
        } else {
            for (int li = 0; li < b.n_lanes; li++) pitch_warp_body<24, false>(b, li, psh);
        }
    }
    for (int f = 0; f < n_frames; f++)
        for (int tile = 0; tile < b.n_tiles; tile++) {
            if (bits == 16) run_front_tile<16>(b, f, tile);
            else run_front_tile<24>(b, f, tile);
        }
    for (int li = 0; li < b.n_lanes; li++) rate_warp_body(b, li, rsh, (const u8 *)rom_ari_lookup, rom_ari_bits);
    for (int f = 0; f < n_frames; f++)
        for (int tile = 0; tile < b.n_tiles; tile++) {
            // the tile quantiser (k_quant in lc3_k_back.cu), phase by phase
            static QuantShared qs;
            for (int t = 0; t < 32; t++) qt_setup(b, f, tile, t, qs.fr[t]);
            for (int t = 0; t < QT_THREADS; t++) qt_quantise(b, f, tile, t, qs);
            for (int w = 0; w < QT_PARTS; w++)
                for (int j = 0; j < 4; j++)
                    qt_analyse(qs.fr[w + 8 * j], qs.xq + (w + 8 * j) * QT_ROW, b.yLen, qs.w[w], (const u8 *)rom_ari_lookup, rom_ari_bits);
            for (int t = 0; t < QT_THREADS; t++) qt_store(b, f, tile, t, qs);
            for (int lane = 0; lane < 32; lane++) {
                u8 frame[440], *dst;
                const i32 nb = back_body(b, f, tile, lane, T, LocalBytes{frame}, dst);
                for (int i = 0; i < nb; i++) dst[i] = frame[i];
            }
            static FecTabs fec_tabs;
            static bool fec_tabs_ready = false;
            if (!fec_tabs_ready) { for (int i = 0; i < 256 + 48; i++) fec_tabs_fill(fec_tabs, i); fec_tabs_ready = true; }
            for (int lane = 0; lane < 32; lane++) fec_body(b, f, tile, lane, fec_tabs); // k_fec

        }
    return 0;
}
