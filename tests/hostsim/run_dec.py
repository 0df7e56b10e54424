"""TEST TOOLING: run the host build of the DECODER's kernel bodies (no GPU needed)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def lib():
    subprocess.run(["make", "-s", "-C", HERE, "libhostsim_dec.so"], check=True)
    L = C.CDLL(os.path.join(HERE, "libhostsim_dec.so"))
    L.hostsim_dec_new.restype = C.c_void_p
    L.hostsim_dec_free.argtypes = [C.c_void_p]
    L.hostsim_dec_frame_len.argtypes = [C.c_void_p]
    L.hostsim_dec_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return L


def sim_decode(L, fs, dms, frames, bits=16, cuts=()):
    """frames: uint8 [S, n_frames, nbytes] -> (pcm [S, n_frames*N], bfi [n_frames, S]); cuts: frame indices where the batch is
    split into consecutive calls (state carried by the simulator)"""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    S, nf, nb = frames.shape
    h = C.c_void_p(L.hostsim_dec_new(fs, dms, S))
    N = L.hostsim_dec_frame_len(h)
    pcm = np.zeros((S, nf * N), dtype=np.int16 if bits == 16 else np.int32)
    bfi = np.zeros((nf, S), dtype=np.uint8)
    edges = [0] + list(cuts) + [nf]
    for a, b in zip(edges[:-1], edges[1:]):
        seg = np.ascontiguousarray(frames[:, a:b])
        o = np.zeros((S, (b - a) * N), dtype=pcm.dtype)
        fl = np.zeros((b - a, S), dtype=np.uint8)
        assert L.hostsim_dec_decode(h, seg.ctypes.data, nb, bits, b - a, o.ctypes.data, fl.ctypes.data) == 0
        pcm[:, a * N:b * N] = o
        bfi[a:b] = fl
    L.hostsim_dec_free(h)
    return pcm, bfi
