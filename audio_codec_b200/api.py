"""ctypes mirror of include/lc3plus_b200.h (same names, argument meaning and error codes as fx/lc3.h).

`Encoder` wraps one LC3_Enc block; `enc_batch` wraps lc3plus_enc_batch.  Buffers may be numpy arrays (host)
or integer device addresses (e.g. torch_tensor.data_ptr()).  Loading fails loudly if the CUDA library has
not been built; encoding fails (LC3Error) if there is no CUDA device -- there is no fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ERRORS = {0: "LC3_OK", 1: "LC3_ERROR", 2: "LC3_DECODE_ERROR", 7: "LC3_NUMBYTES_ERROR", 8: "LC3_PLCMODE_ERROR", 3: "LC3_NULL_ERROR", 4: "LC3_SAMPLERATE_ERROR", 5: "LC3_CHANNELS_ERROR",
          6: "LC3_BITRATE_ERROR", 9: "LC3_EPMODE_ERROR", 10: "LC3_FRAMEMS_ERROR", 11: "LC3_ALIGN_ERROR",
          12: "LC3_EPMR_ERROR", 14: "LC3_BITRATE_UNSET_ERROR", 15: "LC3_BITRATE_SET_ERROR", 18: "LC3_BW_WARNING"}


class LC3Error(RuntimeError):
    def __init__(self, code, what=""):
        self.code = code
        super().__init__(f"{ERRORS.get(code, code)} {what} {last_error()}".strip())


def library_path():
    """The in-tree library; LC3PLUS_B200_LIB selects another build of it (the profiling build, csrc/Makefile TIMING=1)."""
    return os.environ.get("LC3PLUS_B200_LIB") or os.path.join(_HERE, "liblc3plus_b200.so")


def load_library():
    global _LIB
    if _LIB is None:
        p = library_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C audio_codec_b200/csrc); there is no CPU fallback")
        L = C.CDLL(p)
        L.lc3_enc_get_size.argtypes = [C.c_int, C.c_int]
        for n in ("lc3_enc_init",):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int]
        for n in ("lc3_enc_set_bitrate", "lc3_enc_set_frame_dms", "lc3_enc_set_bandwidth", "lc3_enc_set_ep_mode",
                  "lc3_enc_set_ep_mode_request"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int]
        for n in ("lc3_enc_get_scratch_size", "lc3_enc_get_input_samples", "lc3_enc_get_num_bytes",
                  "lc3_enc_get_real_bitrate", "lc3_enc_get_delay", "lc3plus_enc_release"):
            getattr(L, n).argtypes = [C.c_void_p]
        for n in ("lc3_enc16", "lc3_enc24"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        for n in ("lc3plus_enc_batch", "lc3plus_enc_batch_async"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p,
                                      C.c_size_t, C.c_void_p]
        L.lc3plus_set_device.argtypes = [C.c_int]
        L.lc3plus_profile_read_device.argtypes = [C.c_int, C.c_void_p]
        L.lc3plus_enc_set_bitrate_many.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.lc3plus_enc_get_device.argtypes = [C.c_void_p]
        L.lc3plus_enc_fresh_noalloc.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t,
                                                C.c_void_p, C.c_size_t, C.c_void_p]
        L.lc3plus_kernel_launches.restype = C.c_longlong
        L.lc3plus_last_error.restype = C.c_char_p
        L.lc3plus_set_chunk_frames.argtypes = [C.c_int]
        L.lc3plus_enc_release.restype = None
        L.lc3plus_container_header.restype = C.c_size_t
        L.lc3plus_container_header.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32]
        L.lc3plus_container_frames.restype = C.c_size_t
        L.lc3plus_container_frames.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_int, C.c_void_p, C.c_int]
        L.lc3plus_write_bitstream_file.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                   C.c_int, C.c_uint32, C.c_void_p, C.c_int, C.c_size_t, C.c_int,
                                                   C.c_void_p]
        L.lc3plus_read_switching_file.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
        # decoder
        L.lc3_dec_get_size.argtypes = [C.c_int, C.c_int, C.c_int]
        L.lc3_dec_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        for n in ("lc3_dec_set_frame_dms", "lc3_dec_set_ep_enabled"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int]
        for n in ("lc3_dec_get_scratch_size", "lc3_dec_get_output_samples", "lc3_dec_get_delay", "lc3_dec_free",
                  "lc3plus_dec_get_device"):
            getattr(L, n).argtypes = [C.c_void_p]
        for n in ("lc3_dec16", "lc3_dec24"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.lc3plus_dec_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                        C.c_size_t, C.c_int, C.c_void_p]
        L.lc3plus_dec_kernel_launches.restype = C.c_longlong
        L.lc3plus_dec_set_chunk_frames.argtypes = [C.c_int]
        L.lc3plus_dec_profile_enable.argtypes = [C.c_int]
        L.lc3plus_dec_profile_read.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def last_error():
    return load_library().lc3plus_last_error().decode() if _LIB is not None else ""


def kernel_launches():
    return int(load_library().lc3plus_kernel_launches())


def profile_enable(level=2):
    """0/False off, 1 device span only (asynchronous calls stay asynchronous), 2/True per-kernel events as well"""
    load_library().lc3plus_profile_enable(2 if level is True else int(level))


def set_device(device):
    """lc3plus_set_device: encoders created by this thread from now on live on `device` (-1: the current CUDA device)."""
    r = load_library().lc3plus_set_device(int(device))
    if r:
        raise LC3Error(r, "lc3plus_set_device")


def sync():
    """lc3plus_sync: wait for every batch queued with enc_batch(..., wait=False)."""
    r = load_library().lc3plus_sync()
    if r:
        raise LC3Error(r, "lc3plus_sync")


def profile_read():
    """dict of accumulated kernel milliseconds since profile_enable()"""
    ms = (C.c_double * 13)()
    load_library().lc3plus_profile_read(ms)
    return dict(zip(("stage", "fir", "scalar", "pitch", "mdct", "sns", "tns", "rate", "quant", "back", "span", "calls", "chunks"), list(ms)))


def profile_read_device(device):
    ms = (C.c_double * 13)()
    load_library().lc3plus_profile_read_device(int(device), ms)
    return dict(zip(("stage", "fir", "scalar", "pitch", "mdct", "sns", "tns", "rate", "quant", "back", "span", "calls", "chunks"), list(ms)))


def set_bitrate_many(encoders, bitrates):
    """lc3plus_enc_set_bitrate_many: encoders is an EncoderArray, bitrates an int32 array of the same length."""
    b = np.ascontiguousarray(bitrates, dtype=np.int32)
    r = load_library().lc3plus_enc_set_bitrate_many(encoders.arr, len(encoders), b.ctypes.data)
    if r:
        raise LC3Error(r, "lc3plus_enc_set_bitrate_many")


def set_chunk_frames(n):
    load_library().lc3plus_set_chunk_frames(int(n))


class Encoder:
    """One LC3_Enc: lc3_enc_init -> lc3_enc_set_frame_dms -> lc3_enc_set_bitrate -> lc3_enc16/24 (fx/codec_exe.c:192-205)."""

    def __init__(self, samplerate, channels=1, frame_dms=100, bitrate=None):
        L = load_library()
        size = L.lc3_enc_get_size(samplerate, channels)
        self._mem = C.create_string_buffer(max(size, 64) + 16)
        self.h = C.c_void_p((C.addressof(self._mem) + 15) & ~15)
        self.channels = channels
        self._live = False
        r = L.lc3_enc_init(self.h, samplerate, channels)
        if r:
            raise LC3Error(r, "lc3_enc_init")
        self._live = True
        self._scratch = C.create_string_buffer(16)   # must not be NULL (fx/lc3.c:250); never used by this library
        r = L.lc3_enc_set_frame_dms(self.h, frame_dms)
        if r:
            raise LC3Error(r, "lc3_enc_set_frame_dms")
        if bitrate is not None:
            self.set_bitrate(bitrate)

    def close(self):
        if self._live:
            load_library().lc3plus_enc_release(self.h)
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_bitrate(self, bitrate, check=True):
        r = load_library().lc3_enc_set_bitrate(self.h, bitrate)
        if r and check:
            raise LC3Error(r, "lc3_enc_set_bitrate")
        return r

    def set_frame_dms(self, dms):
        return load_library().lc3_enc_set_frame_dms(self.h, dms)

    def set_bandwidth(self, bw):
        return load_library().lc3_enc_set_bandwidth(self.h, bw)

    def set_ep_mode(self, epmode):
        """lc3_enc_set_ep_mode: 0 off, 1..4 (fx/lc3.h:74-82); changes num_bytes' payload share, not num_bytes."""
        return load_library().lc3_enc_set_ep_mode(self.h, epmode)

    def set_ep_mode_request(self, epmr):
        return load_library().lc3_enc_set_ep_mode_request(self.h, epmr)

    @property
    def frame_len(self):
        return load_library().lc3_enc_get_input_samples(self.h)

    @property
    def num_bytes(self):
        return load_library().lc3_enc_get_num_bytes(self.h)

    @property
    def real_bitrate(self):
        return load_library().lc3_enc_get_real_bitrate(self.h)

    @property
    def device(self):
        return load_library().lc3plus_enc_get_device(self.h)

    @property
    def delay(self):
        return load_library().lc3_enc_get_delay(self.h)

    def enc(self, frame, bits=16, lost=False):
        """frame: [channels, N] -> bytes of one frame (lc3_enc16 / lc3_enc24)."""
        L = load_library()
        frame = np.ascontiguousarray(frame, dtype=np.int16 if bits == 16 else np.int32)
        if frame.ndim == 1:
            frame = frame[None, :]
        ptrs = (C.c_void_p * 2)()
        for c in range(self.channels):
            ptrs[c] = frame[c].ctypes.data
        out = np.zeros(870, dtype=np.uint8)
        n = C.c_int(-1 if lost else 0)
        r = (L.lc3_enc16 if bits == 16 else L.lc3_enc24)(self.h, ptrs, out.ctypes.data, C.byref(n), self._scratch)
        if r:
            raise LC3Error(r, "lc3_enc")
        return out[:n.value].copy()


def _addr(x):
    return x.ctypes.data if isinstance(x, np.ndarray) else int(x)


class EncoderArray:
    """The `LC3_Enc *const *encoders` argument of lc3plus_enc_batch, built once for a fixed set of encoders (a C caller
    keeps such an array anyway; building it from a Python list costs about 0.1 us per encoder per call)."""

    def __init__(self, encoders):
        self.encoders = list(encoders)  # keeps the Encoder objects (and their handle blocks) alive
        self.arr = (C.c_void_p * len(self.encoders))(*[e.h.value for e in self.encoders])

    def __len__(self):
        return len(self.encoders)


def enc_batch(encoders, n_frames, pcm, stream_stride_bytes, bits, out, out_stream_stride_bytes, num_bytes=None, wait=True):
    """lc3plus_enc_batch (wait=False: lc3plus_enc_batch_async, results valid after sync()).  encoders: list of Encoder or an
    EncoderArray.  pcm/out: numpy arrays or device addresses.  num_bytes: optional int32 array [S*n_frames]."""
    L = load_library()
    arr = encoders.arr if isinstance(encoders, EncoderArray) else (C.c_void_p * len(encoders))(*[e.h.value for e in encoders])
    nb = num_bytes.ctypes.data if num_bytes is not None else None
    r = (L.lc3plus_enc_batch if wait else L.lc3plus_enc_batch_async)(arr, len(encoders), n_frames, _addr(pcm), stream_stride_bytes, bits,
                                                                     _addr(out), out_stream_stride_bytes, nb)
    if r:
        raise LC3Error(r, "lc3plus_enc_batch")


def enc_fresh_noalloc(samplerate, frame_dms, bitrate, bits, n_streams, n_frames, d_pcm, stream_stride_bytes, d_out,
                      out_stream_stride_bytes, cuda_stream=None):
    """lc3plus_enc_fresh_noalloc: fresh mono streams, device pointers only, nothing allocated, nothing synchronised."""
    r = load_library().lc3plus_enc_fresh_noalloc(samplerate, frame_dms, bitrate, bits, n_streams, n_frames, int(d_pcm),
                                                 stream_stride_bytes, int(d_out), out_stream_stride_bytes, cuda_stream)
    if r:
        raise LC3Error(r, "lc3plus_enc_fresh_noalloc")


# ---- bitstream containers of the reference tool (include/lc3plus_b200.h); host only

def container_header(samplerate, bitrate, channels, frame_dms, epmode=0, signal_len=0):
    """The 18-byte .lc3plus header (fx/codec_exe.c:742-746)."""
    buf = np.zeros(18, dtype=np.uint8)
    n = load_library().lc3plus_container_header(buf.ctypes.data, samplerate, bitrate, channels, frame_dms, epmode, signal_len)
    return buf[:n].tobytes()


def container_frames(frames, num_bytes=None, g192=False):
    """frames: uint8 [n_frames, stride]; num_bytes: optional per-frame sizes (default: the row length)."""
    L = load_library()
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    nb = np.ascontiguousarray(num_bytes, dtype=np.int32) if num_bytes is not None else None
    args = (frames.ctypes.data, frames.shape[0], frames.shape[1], frames.shape[1], nb.ctypes.data if nb is not None else None,
            1 if g192 else 0)
    need = L.lc3plus_container_frames(None, *args)
    buf = np.zeros(max(need, 1), dtype=np.uint8)
    got = L.lc3plus_container_frames(buf.ctypes.data, *args)
    if got != need or (need == 0 and frames.shape[0] > 0):
        raise LC3Error(7, "lc3plus_container_frames")
    return buf[:need].tobytes()


def write_bitstream_file(path, samplerate, bitrate, channels, frame_dms, signal_len, frames, num_bytes=None, epmode=0,
                         g192=False, cfg_path=None):
    """lc3plus_write_bitstream_file: the file the reference tool writes for these frames (fx/codec_exe.c:255-260, :490)."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    nb = np.ascontiguousarray(num_bytes, dtype=np.int32) if num_bytes is not None else None
    r = load_library().lc3plus_write_bitstream_file(os.fsencode(path), os.fsencode(cfg_path) if cfg_path else None,
                                                    1 if g192 else 0, samplerate, bitrate, channels, frame_dms, epmode,
                                                    signal_len, frames.ctypes.data, frames.shape[0], frames.shape[1],
                                                    frames.shape[1], nb.ctypes.data if nb is not None else None)
    if r:
        raise LC3Error(r, "lc3plus_write_bitstream_file")


def read_switching_file(path):
    """int64 values of a switching file (fx/codec_exe.c:968-977); frame f uses values[f % len(values)]."""
    L = load_library()
    n = L.lc3plus_read_switching_file(os.fsencode(path), None, 0)
    if n < 0:
        raise OSError(f"cannot open {path}")
    v = np.zeros(max(n, 1), dtype=np.int64)
    L.lc3plus_read_switching_file(os.fsencode(path), v.ctypes.data, n)
    return v[:n]


# ---- decoder (include/lc3plus_b200.h): error-free streams, no concealment

class Decoder:
    """One LC3_Dec: lc3_dec_init -> lc3_dec_set_frame_dms -> lc3_dec16/24 (fx/codec_exe.c:206-215)."""

    def __init__(self, samplerate, channels=1, frame_dms=100, plc_mode=1):
        L = load_library()
        size = L.lc3_dec_get_size(samplerate, channels, plc_mode)
        self._mem = C.create_string_buffer(max(size, 64) + 16)
        self.h = C.c_void_p((C.addressof(self._mem) + 15) & ~15)
        self.channels = channels
        self._live = False
        r = L.lc3_dec_init(self.h, samplerate, channels, plc_mode)
        if r:
            raise LC3Error(r, "lc3_dec_init")
        self._live = True
        self._scratch = C.create_string_buffer(16)
        r = L.lc3_dec_set_frame_dms(self.h, frame_dms)
        if r:
            raise LC3Error(r, "lc3_dec_set_frame_dms")

    def close(self):
        if self._live:
            load_library().lc3_dec_free(self.h)
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def frame_len(self):
        return load_library().lc3_dec_get_output_samples(self.h)

    @property
    def delay(self):
        return load_library().lc3_dec_get_delay(self.h)

    @property
    def device(self):
        return load_library().lc3plus_dec_get_device(self.h)

    def dec(self, payload, bits=16, bfi_ext=0):
        """payload: bytes of one frame -> ([channels, N] samples, return code 0 or 2 = LC3_DECODE_ERROR)"""
        L = load_library()
        payload = np.ascontiguousarray(payload, dtype=np.uint8)
        out = np.zeros((self.channels, self.frame_len), dtype=np.int16 if bits == 16 else np.int32)
        ptrs = (C.c_void_p * 2)()
        for c in range(self.channels):
            ptrs[c] = out[c].ctypes.data
        r = (L.lc3_dec16 if bits == 16 else L.lc3_dec24)(self.h, payload.ctypes.data, len(payload), ptrs, self._scratch, bfi_ext)
        if r not in (0, 2):
            raise LC3Error(r, "lc3_dec")
        return out, r


class DecoderArray:
    """The `LC3_Dec *const *decoders` argument of lc3plus_dec_batch, built once for a fixed set of decoders."""

    def __init__(self, decoders):
        self.decoders = list(decoders)
        self.arr = (C.c_void_p * len(self.decoders))(*[d.h.value for d in self.decoders])

    def __len__(self):
        return len(self.decoders)


def dec_batch(decoders, n_frames, in_bytes, in_stream_stride_bytes, num_bytes, pcm, pcm_stream_stride_bytes, bits=16, bfi=None):
    """lc3plus_dec_batch.  in_bytes / pcm: numpy arrays or device addresses; num_bytes: int32 [n_streams]; bfi: optional uint8
    [n_streams*n_frames].  Returns 0 or 2 (LC3_DECODE_ERROR: at least one frame did not decode); raises on anything else."""
    L = load_library()
    arr = decoders.arr if isinstance(decoders, DecoderArray) else (C.c_void_p * len(decoders))(*[d.h.value for d in decoders])
    nb = np.ascontiguousarray(num_bytes, dtype=np.int32)
    r = L.lc3plus_dec_batch(arr, len(decoders), n_frames, _addr(in_bytes), in_stream_stride_bytes, nb.ctypes.data, _addr(pcm),
                            pcm_stream_stride_bytes, bits, bfi.ctypes.data if bfi is not None else None)
    if r not in (0, 2):
        raise LC3Error(r, "lc3plus_dec_batch")
    return r


def dec_kernel_launches():
    return int(load_library().lc3plus_dec_kernel_launches())


def dec_profile_enable(on=True):
    load_library().lc3plus_dec_profile_enable(1 if on else 0)


def dec_profile_read():
    ms = (C.c_double * 4)()
    load_library().lc3plus_dec_profile_read(ms)
    return dict(zip(("entropy_chunk0", "idct_chunk0", "synth_chunk0", "span"), list(ms)))
