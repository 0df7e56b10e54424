"""TEST TOOLING: run the host build of the kernel bodies against the oracle (no GPU needed)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import lc3ref
from audio_codec_b200.synth import make_streams


def lib():
    subprocess.run(["make", "-s", "-C", HERE], check=True)
    L = C.CDLL(os.path.join(HERE, "libhostsim.so"))
    L.hostsim_new.restype = C.c_void_p
    L.hostsim_free.argtypes = [C.c_void_p]
    L.hostsim_set_bitrate.argtypes = [C.c_void_p, C.c_int]
    L.hostsim_set_ep.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.hostsim_frame_len.argtypes = [C.c_void_p]
    L.hostsim_num_bytes.argtypes = [C.c_void_p]
    L.hostsim_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return L


def sim_encode(L, fs, dms, schedule, pcm, bits, bfi=None, ep=(0, 0)):
    """schedule: list of (bitrate, n_frames) segments; pcm [S, total*N]; ep = (error protection mode, mode request)"""
    S = pcm.shape[0]
    h = C.c_void_p(L.hostsim_new(fs, dms, S))
    assert L.hostsim_set_ep(h, ep[0], ep[1]) == 0
    outs = []
    pos = 0
    for br, nf in schedule:
        assert L.hostsim_set_bitrate(h, br) == 0
        N, nb = L.hostsim_frame_len(h), L.hostsim_num_bytes(h)
        seg = np.ascontiguousarray(pcm[:, pos * N:(pos + nf) * N])
        out = np.zeros((S, nf, nb), dtype=np.uint8)
        b = None
        if bfi is not None:
            b = np.ascontiguousarray(np.repeat(np.asarray(bfi[pos:pos + nf], dtype=np.uint8)[:, None], S, 1))
        L.hostsim_encode(h, seg.ctypes.data, bits, nf, out.ctypes.data, b.ctypes.data if b is not None else None)
        outs.append(out)
        pos += nf
    L.hostsim_free(h)
    return outs


def oracle_encode(fs, dms, schedule, pcm, bits, bfi=None, ep=(0, 0)):
    outs = []
    encs = [lc3ref.OracleEnc(fs, 1, dms) for _ in range(pcm.shape[0])]
    for e in encs:
        assert e.set_ep_mode(ep[0]) == 0 and e.set_ep_mode_request(ep[1]) == 0
    pos = 0
    for br, nf in schedule:
        seg_out = []
        for s, e in enumerate(encs):
            assert e.set_bitrate(br) == 0
            N = e.frame_len
            seg_out.append(e.encode(pcm[s:s + 1, pos * N:(pos + nf) * N], bits, bfi[pos:pos + nf] if bfi is not None else None))
        outs.append(np.stack(seg_out))
        pos += nf
    return outs


if __name__ == "__main__":
    L = lib()
    cases = [(48000, 100, 128000, 16)] if len(sys.argv) < 2 else [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
    for fs, dms, br, bits in cases:
        N = (48000 if fs == 44100 else fs) // 100 * dms // 100
        nf, S = 30, 9
        pcm = make_streams(S, nf * N, fs, N, bits, silence_frames=10)
        a = sim_encode(L, fs, dms, [(br, nf)], pcm, bits)[0]
        b = oracle_encode(fs, dms, [(br, nf)], pcm, bits)[0]
        bad = np.argwhere((a != b).any(2))
        print(fs, dms, br, bits, "OK" if len(bad) == 0 else f"MISMATCH first (stream,frame)={bad[0]} count={len(bad)}")
