"""Synthetic multichannel PCM used by the parity tests and bench.py.

Follows the reference repo's own stimulus definition
(/root/reference/sim/scripts/generate_test_vectors.py:94-130): a fundamental plus three harmonics
(0.5/0.3/0.2/0.1), Gaussian noise sigma 0.01, 10 ms linear fade-in, scaled to full range and
clipped.  Streams are de-correlated as SURVEY.md section 8d states: f_s = 200 + (37 s mod 3000) Hz,
amplitude 0.1 + 0.9 ((7919 s) mod 1000)/1000, noise seed 1234 + s; every 8th stream is digital
silence for its first 50 frames (exercises the all-zero spectrum paths).
"""
import numpy as np


def make_streams(n_streams, n_samples, fs, frame_len, bits=16, first_stream=0, silence_frames=50):
    """Return int16 (bits=16) or int32 (bits=24) array [n_streams, n_samples]."""
    t = np.arange(n_samples, dtype=np.float64) / fs
    out = np.empty((n_streams, n_samples), dtype=np.int16 if bits == 16 else np.int32)
    full = 32767.0 if bits == 16 else 8388607.0
    fade = int(0.01 * fs)
    env = np.ones(n_samples)
    env[:min(fade, n_samples)] = np.linspace(0, 1, fade)[:min(fade, n_samples)]
    for i in range(n_streams):
        s = first_stream + i
        f = 200.0 + (37 * s) % 3000
        amp = 0.1 + 0.9 * ((s * 7919) % 1000) / 1000.0
        rng = np.random.RandomState(1234 + s)
        x = (0.5 * np.sin(2 * np.pi * f * t) + 0.3 * np.sin(2 * np.pi * 2 * f * t)
             + 0.2 * np.sin(2 * np.pi * 3 * f * t) + 0.1 * np.sin(2 * np.pi * 4 * f * t))
        x = (x + rng.normal(0, 0.01, n_samples)) * env * amp
        if s % 8 == 7:
            x[:min(silence_frames * frame_len, n_samples)] = 0.0
        out[i] = np.clip(x * full, -full - 1, full).astype(out.dtype)
    return out


def make_streams_fast(n_streams, n_samples, fs, frame_len, bits=16, n_unique=64):
    """Large batches for timing: n_unique distinct streams from make_streams(), tiled with a
    per-copy circular shift so that no two streams are sample-identical."""
    base = make_streams(min(n_unique, n_streams), n_samples, fs, frame_len, bits)
    if n_streams <= base.shape[0]:
        return base
    out = np.empty((n_streams, n_samples), dtype=base.dtype)
    for i in range(n_streams):
        out[i] = np.roll(base[i % base.shape[0]], (i // base.shape[0]) * 97)
    return out


def make_bursts(n_streams, n_samples, fs, bits=16, seed=99, n_unique=64):
    """Transient-rich streams (timing robustness, not in the reference's stimulus set): a harmonic tone gated by a
    random on/off envelope with abrupt onsets every 20-120 ms plus band-limited noise bursts and clicks, so that
    the temporal-noise-shaping, attack and pitch-change paths run in most frames instead of a few."""
    full = 32767.0 if bits == 16 else 8388607.0
    t = np.arange(n_samples, dtype=np.float64) / fs
    base = np.empty((min(n_unique, n_streams), n_samples), dtype=np.int16 if bits == 16 else np.int32)
    for i in range(base.shape[0]):
        rng = np.random.RandomState(seed + i)
        env = np.zeros(n_samples)
        pos, on = 0, False
        while pos < n_samples:
            seg = int(fs * rng.uniform(0.02, 0.12))
            if on:
                env[pos:pos + seg] = rng.uniform(0.2, 1.0) * np.exp(-np.arange(min(seg, n_samples - pos)) / (fs * rng.uniform(0.01, 0.08)))
            on = not on
            pos += seg
        f0 = rng.uniform(80, 400) * (1 + 0.2 * np.sin(2 * np.pi * rng.uniform(0.5, 3) * t))
        ph = 2 * np.pi * np.cumsum(f0) / fs
        x = sum(np.sin(k * ph) / k for k in range(1, 8)) * 0.3 + rng.normal(0, 0.25, n_samples)
        x = x * env
        clicks = rng.randint(0, n_samples, max(1, n_samples // (fs // 5)))
        x[clicks] += rng.uniform(-1, 1, len(clicks))
        base[i] = np.clip(x * full * 0.7, -full - 1, full).astype(base.dtype)
    if n_streams <= base.shape[0]:
        return base
    out = np.empty((n_streams, n_samples), dtype=base.dtype)
    for i in range(n_streams):
        out[i] = np.roll(base[i % base.shape[0]], (i // base.shape[0]) * 97)
    return out


def make_torture(n_samples, fs, bits=16, seed=7):
    """Adversarial streams for the saturation / headroom paths: [10, n_samples].
    full-scale noise, full-scale square, Nyquist alternation, negative-rail DC, sparse impulses, silence then
    a full-scale step, +-1 LSB noise, full-scale chirp, full-scale low sine, clipped sum of tones."""
    full = 32767 if bits == 16 else 8388607
    lo = -full - 1
    rng = np.random.RandomState(seed)
    t = np.arange(n_samples, dtype=np.float64) / fs
    x = np.zeros((10, n_samples), dtype=np.float64)
    x[0] = rng.randint(lo, full + 1, n_samples)
    x[1] = np.where((np.arange(n_samples) // 37) % 2 == 0, full, lo)
    x[2] = np.where(np.arange(n_samples) % 2 == 0, full, lo)
    x[3] = lo
    x[4, ::211] = full
    x[4, 100::503] = lo
    x[5, n_samples // 3:] = full
    x[6] = rng.randint(-1, 2, n_samples)
    x[7] = full * np.sin(2 * np.pi * (50 + 0.45 * fs * t / max(t[-1], 1e-9) / 2) * t)
    x[8] = full * np.sin(2 * np.pi * 60 * t)
    x[9] = np.clip(3.0 * full * (np.sin(2 * np.pi * 440 * t) + np.sin(2 * np.pi * 3000 * t)), lo, full)
    return np.clip(np.round(x), lo, full).astype(np.int16 if bits == 16 else np.int32)
