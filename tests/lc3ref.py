"""ctypes access to the two CPU checkers (TEST INFRASTRUCTURE ONLY):

  * RefEnc    -- the unmodified ETSI fixed-point encoder compiled from /root/reference into
                 oracle/_ref/liblc3ref.so (API: fx/lc3.h:170-309)
  * OracleEnc -- our plain-C restatement, oracle/liblc3oracle.so (API: oracle/lc3o.h)

Both expose encode(pcm[ch][n_frames*N]) -> bytes [n_frames, nbytes].
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "liblc3ref.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "liblc3oracle.so")

_ref = None
_ora = None


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        _ref = C.CDLL(REF_SO)
        _ref.lc3_enc_get_size.restype = C.c_int
    return _ref


def oracle_lib():
    global _ora
    if _ora is None:
        _ora = C.CDLL(ORACLE_SO)
        _ora.lc3o_enc_new.restype = C.c_void_p
        _ora.lc3o_enc_free.argtypes = [C.c_void_p]
        for n in ("lc3o_enc_set_frame_dms", "lc3o_enc_set_bitrate", "lc3o_enc_set_bandwidth", "lc3o_enc_set_ep_mode",
                  "lc3o_enc_set_ep_mode_request"):
            getattr(_ora, n).argtypes = [C.c_void_p, C.c_int]
        for n in ("lc3o_enc_get_input_samples", "lc3o_enc_get_num_bytes", "lc3o_enc_get_delay",
                  "lc3o_enc_get_real_bitrate"):
            getattr(_ora, n).argtypes = [C.c_void_p]
        for n in ("lc3o_enc16", "lc3o_enc24"):
            getattr(_ora, n).argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _ora.lc3o_enc16_trace.argtypes = [C.c_void_p] * 5
        _ora.lc3o_dec_new.restype = C.c_void_p
        _ora.lc3o_dec_new.argtypes = [C.c_int, C.c_int]
        _ora.lc3o_dec_free.argtypes = [C.c_void_p]
        _ora.lc3o_dec_set_frame_dms.argtypes = [C.c_void_p, C.c_int]
        _ora.lc3o_dec_get_output_samples.argtypes = [C.c_void_p]
        _ora.lc3o_dec_get_delay.argtypes = [C.c_void_p]
        for n in ("lc3o_dec16", "lc3o_dec24"):
            getattr(_ora, n).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return _ora


class _Base:
    def encode(self, pcm, bits=16, bfi=None):
        """pcm: [channels, n_frames*N] int16/int32 -> uint8 [n_frames, nbytes]"""
        pcm = np.ascontiguousarray(pcm)
        if pcm.ndim == 1:
            pcm = pcm[None, :]
        assert pcm.shape[0] == self.channels
        N = self.frame_len
        nf = pcm.shape[1] // N
        nb = self.num_bytes
        out = np.zeros((nf, nb), dtype=np.uint8)
        ct = C.c_int16 if bits == 16 else C.c_int32
        assert pcm.dtype == (np.int16 if bits == 16 else np.int32)
        ptrs = (C.c_void_p * 2)()
        for f in range(nf):
            for c in range(self.channels):
                ptrs[c] = pcm[c, f * N:].ctypes.data
            n = C.c_int(-1 if (bfi is not None and bfi[f]) else 0)
            r = self._enc(bits, ptrs, out[f].ctypes.data, C.byref(n))
            assert r == 0, r
            assert n.value == nb
        return out


class RefEnc(_Base):
    def __init__(self, fs, channels=1, dms=100, bitrate=None):
        L = ref_lib()
        self.channels = channels
        self._mem = C.create_string_buffer(L.lc3_enc_get_size(fs, channels) + 16)
        self.h = C.c_void_p((C.addressof(self._mem) + 15) & ~15)
        r = L.lc3_enc_init(self.h, fs, channels)
        assert r == 0, r
        r = L.lc3_enc_set_frame_dms(self.h, dms)
        assert r == 0, r
        self._scratch = C.create_string_buffer(L.lc3_enc_get_scratch_size(self.h) + 64)
        if bitrate is not None:
            assert self.set_bitrate(bitrate) == 0

    def set_bitrate(self, br):
        return ref_lib().lc3_enc_set_bitrate(self.h, br)

    def set_bandwidth(self, bw):
        return ref_lib().lc3_enc_set_bandwidth(self.h, bw)

    def set_ep_mode(self, epmode):
        return ref_lib().lc3_enc_set_ep_mode(self.h, epmode)

    def set_ep_mode_request(self, epmr):
        return ref_lib().lc3_enc_set_ep_mode_request(self.h, epmr)

    @property
    def frame_len(self):
        return ref_lib().lc3_enc_get_input_samples(self.h)

    @property
    def num_bytes(self):
        return ref_lib().lc3_enc_get_num_bytes(self.h)

    def _enc(self, bits, ptrs, out, n):
        L = ref_lib()
        return (L.lc3_enc16 if bits == 16 else L.lc3_enc24)(self.h, ptrs, C.c_void_p(out), n, self._scratch)


class OracleEnc(_Base):
    def __init__(self, fs, channels=1, dms=100, bitrate=None):
        L = oracle_lib()
        self.channels = channels
        self.h = C.c_void_p(L.lc3o_enc_new(fs, channels))
        assert self.h.value
        r = L.lc3o_enc_set_frame_dms(self.h, dms)
        assert r == 0, r
        if bitrate is not None:
            assert self.set_bitrate(bitrate) == 0

    def __del__(self):
        try:
            oracle_lib().lc3o_enc_free(self.h)
        except Exception:
            pass

    def set_bitrate(self, br):
        return oracle_lib().lc3o_enc_set_bitrate(self.h, br)

    def set_bandwidth(self, bw):
        return oracle_lib().lc3o_enc_set_bandwidth(self.h, bw)

    def set_ep_mode(self, epmode):
        return oracle_lib().lc3o_enc_set_ep_mode(self.h, epmode)

    def set_ep_mode_request(self, epmr):
        return oracle_lib().lc3o_enc_set_ep_mode_request(self.h, epmr)

    @property
    def frame_len(self):
        return oracle_lib().lc3o_enc_get_input_samples(self.h)

    @property
    def num_bytes(self):
        return oracle_lib().lc3o_enc_get_num_bytes(self.h)

    def _enc(self, bits, ptrs, out, n):
        L = oracle_lib()
        return (L.lc3o_enc16 if bits == 16 else L.lc3o_enc24)(self.h, ptrs, C.c_void_p(out), n)


class RefDec:
    """The unmodified reference DECODER (oracle/_ref/liblc3ref.so, fx/lc3.h:334-457): test infrastructure for round trips."""

    def __init__(self, fs, channels=1, dms=100, ep=False):
        L = ref_lib()
        L.lc3_dec_get_size.restype = C.c_int
        self.channels = channels
        self._mem = C.create_string_buffer(L.lc3_dec_get_size(fs, channels, 1) + 16)
        self.h = C.c_void_p((C.addressof(self._mem) + 15) & ~15)
        assert L.lc3_dec_init(self.h, fs, channels, 1) == 0   # LC3_PLC_ADVANCED, the only mode this build supports
        assert L.lc3_dec_set_frame_dms(self.h, dms) == 0
        if ep:
            assert L.lc3_dec_set_ep_enabled(self.h, 1) == 0
        self._scratch = C.create_string_buffer(L.lc3_dec_get_scratch_size(self.h) + 64)
        self.frame_len = L.lc3_dec_get_output_samples(self.h)
        self.delay = L.lc3_dec_get_delay(self.h)

    def decode(self, frames, bits=16, codes=None):
        """frames: uint8 [n_frames, nbytes] -> int16 / int32 [channels, n_frames * N]; returns (pcm, number of frames flagged
        bad); codes: optional list that receives the per-frame return value of lc3_dec16 / lc3_dec24"""
        L = ref_lib()
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        nf, nb = frames.shape
        N = self.frame_len
        out = np.zeros((self.channels, nf * N), dtype=np.int16 if bits == 16 else np.int32)
        ptrs = (C.c_void_p * 2)()
        bad = 0
        fn = L.lc3_dec16 if bits == 16 else L.lc3_dec24
        for f in range(nf):
            for c in range(self.channels):
                ptrs[c] = out[c, f * N:].ctypes.data
            r = fn(self.h, C.c_void_p(frames[f].ctypes.data), nb, ptrs, self._scratch, 0)
            bad += r != 0
            if codes is not None:
                codes.append(r)
        return out, bad


class OracleDec:
    """Our plain-C restatement of the decoder for good frames (oracle/lc3o_dec.c): test infrastructure.  Same interface as RefDec."""

    def __init__(self, fs, channels=1, dms=100):
        L = oracle_lib()
        self.channels = channels
        self.h = C.c_void_p(L.lc3o_dec_new(fs, channels))
        assert self.h.value
        assert L.lc3o_dec_set_frame_dms(self.h, dms) == 0
        self.frame_len = L.lc3o_dec_get_output_samples(self.h)
        self.delay = L.lc3o_dec_get_delay(self.h)

    def __del__(self):
        try:
            oracle_lib().lc3o_dec_free(self.h)
        except Exception:
            pass

    def decode(self, frames, bits=16, codes=None):
        L = oracle_lib()
        frames = np.ascontiguousarray(frames, dtype=np.uint8)
        nf, nb = frames.shape
        N = self.frame_len
        out = np.zeros((self.channels, nf * N), dtype=np.int16 if bits == 16 else np.int32)
        ptrs = (C.c_void_p * 2)()
        bad = 0
        fn = L.lc3o_dec16 if bits == 16 else L.lc3o_dec24
        for f in range(nf):
            for c in range(self.channels):
                ptrs[c] = out[c, f * N:].ctypes.data
            r = fn(self.h, C.c_void_p(frames[f].ctypes.data), nb, ptrs)
            bad += r != 0
            if codes is not None:
                codes.append(r)
        return out, bad
