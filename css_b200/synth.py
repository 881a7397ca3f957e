"""Seeded synthetic LibriCSS-shaped recordings.

Host-side only (numpy): generates the D x T fp32 mixtures the bench and the parity tests
feed to both the CUDA path and the CPU oracle.  Geometry follows the array the reference
assumes in css/executor/separator.py:129-140 (mic 0 in the centre, mics 1-6 on a
4.25 cm circle, c = 340 m/s).
"""
import math

import numpy as np
import torch

_MIC_OFFSETS = (None, math.pi / 6, -math.pi / 6, -math.pi / 2, -5 * math.pi / 6, 5 * math.pi / 6, math.pi / 2)


def mic_delays(theta, sr=16000, radius=0.0425, c=340.0):
    """Far-field delay (in samples) of each of the 7 mics for a source at angle theta (rad)."""
    return np.array([0.0 if o is None else radius * math.cos(theta + o) / c * sr for o in _MIC_OFFSETS])


def _speech_like(rng, n, sr):
    """Speech-shaped noise: -6 dB/oct 100 Hz-4 kHz envelope, 3-5 Hz syllabic modulation."""
    spec = np.fft.rfft(rng.standard_normal(n))
    f = np.fft.rfftfreq(n, 1.0 / sr)
    shape = np.where((f >= 100) & (f <= 4000), 100.0 / np.maximum(f, 100.0), 0.0)
    x = np.fft.irfft(spec * shape, n)
    t = np.arange(n) / sr
    rate = rng.uniform(3.0, 5.0)
    env = 0.55 + 0.45 * np.sin(2 * math.pi * rate * t + rng.uniform(0, 2 * math.pi))
    return x * env / (np.abs(x).max() + 1e-9)


def _activity(rng, n, sr, lo=2.0, hi=10.0, pause=(0.3, 1.5), start_talking=True):
    """0/1 gate: utterances of lo..hi seconds separated by pauses, 10 ms ramps."""
    gate = np.zeros(n)
    pos = 0 if start_talking else int(rng.uniform(*pause) * sr)
    while pos < n:
        dur = int(rng.uniform(lo, hi) * sr)
        gate[pos : pos + dur] = 1.0
        pos += dur + int(rng.uniform(*pause) * sr)
    ramp = int(0.01 * sr)
    k = np.ones(ramp) / ramp
    return np.convolve(gate, k, mode="same")


def make_recording(seconds, channels=7, sr=16000, seed=1234, speakers=2, snr_db=20.0, return_sources=False):
    """Return a (channels, seconds*sr) float32 torch tensor (peak 0.5) of overlapping talkers."""
    rng = np.random.default_rng(seed)
    n = int(seconds * sr)
    base = rng.uniform(0, 2 * math.pi)
    thetas = [base + k * (2 * math.pi / speakers) + rng.uniform(-0.3, 0.3) for k in range(speakers)]
    mix = np.zeros((channels, n))
    sources = []
    freqs = np.fft.rfftfreq(n, 1.0 / sr)
    for k in range(speakers):
        src = _speech_like(rng, n, sr) * _activity(rng, n, sr, start_talking=(k == 0)) * rng.uniform(0.6, 1.0)
        sources.append(src)
        if channels == 1:
            mix[0] += src
            continue
        spec = np.fft.rfft(src)
        delays = mic_delays(thetas[k], sr)
        for m in range(channels):
            mix[m] += np.fft.irfft(spec * np.exp(-2j * math.pi * freqs * delays[m] / sr), n)
    sig_pow = np.mean(mix ** 2)
    mix += rng.standard_normal(mix.shape) * math.sqrt(sig_pow / (10 ** (snr_db / 10)))
    scale = 0.5 / np.abs(mix).max()
    out = torch.from_numpy((mix * scale).astype(np.float32))
    if return_sources:
        return out, torch.from_numpy((np.stack(sources) * scale).astype(np.float32))
    return out


def n_chunks(T, sr=16000, eval_win=2.4, eval_hop=0.8):
    """Number of full windows ``unfold`` yields (css/executor/separator.py:67)."""
    win = int(eval_win * sr) + 256
    return 0 if T < win else (T - win) // int(eval_hop * sr) + 1
