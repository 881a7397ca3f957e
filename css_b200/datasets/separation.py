"""On-the-fly 2-speaker training mixtures, mixed on the GPU (reference: css/datasets/separation.py:26-267).

The reference builds every batch on the host with lhotse cut arithmetic and a DataLoader: sample two speakers and one
utterance each, start the shorter one at a random offset inside the first half of the longer one, pad both to the mixture
length, cut the mixture and the two padded sources into windows of a per-batch random size, reverberate (p = 0.5) and add
noise at a random SNR (p = 0.5) per window, extract STFT features (separation.py:163-267).

Here the utterances (and optional noise / impulse responses) are resident in HBM as one flat pool; the host keeps ONLY the
sampling decisions (same order of ``random`` calls as the reference's ``__next__``, so a seed yields the same choices for the
same utterance list), and three kernels do the arithmetic for a whole batch (``css_mix_windows``, ``css_mix_reverb``,
``css_mix_add_noise``), followed by the STFT / feature kernels the inference path uses.  Batches never touch the host.

Parity status: lhotse is not installed and the reference pins no version of it, so this path is "parity unpinned"; the
arithmetic is checked against a numpy restatement of the steps above (tests/test_gpu_mixer.py).
"""
import math
import random

import torch

from .. import _native as nv
from ..executor.feature import FeatureExtractor


class UtterancePool:
    """Utterances of equal channel count packed into one (C, total) fp32 CUDA buffer."""

    def __init__(self, utterances, speakers=None, device="cuda", sampling_rate=16000):
        """utterances: list of (C, n_i) or (n_i,) tensors; speakers: list of speaker ids (one per utterance)."""
        us = [u if u.dim() == 2 else u.unsqueeze(0) for u in utterances]
        self.channels = us[0].shape[0]
        if any(u.shape[0] != self.channels for u in us):
            raise ValueError("all utterances of a pool must have the same number of channels")
        self.lengths = [int(u.shape[1]) for u in us]
        self.offsets = [0]
        for n in self.lengths[:-1]:
            self.offsets.append(self.offsets[-1] + n)
        self.total = sum(self.lengths)
        self.data = torch.cat([u.float() for u in us], dim=1).to(device).contiguous()
        self.speakers = list(speakers) if speakers is not None else list(range(len(us)))
        self.sampling_rate = sampling_rate
        self.by_speaker = {}
        for i, s in enumerate(self.speakers):
            self.by_speaker.setdefault(s, []).append(i)
        self.speaker_ids = sorted(self.by_speaker)

    def duration(self, i):
        return self.lengths[i] / self.sampling_rate


class ContinuousSpeechSeparationDataset(torch.utils.data.IterableDataset):
    """Reference surface (separation.py:84-161): an iterable yielding dicts with mix / lens / source1 / source2 / ovl."""

    @staticmethod
    def add_args(parser):
        parser.add_argument("--batch-size", type=int, default=32)
        parser.add_argument("--min-window-size", type=float, default=2)
        parser.add_argument("--max-window-size", type=float, default=5)
        parser.add_argument("--min-snr", type=float, default=5, help="Minimum SNR for isotropic noise")
        parser.add_argument("--max-snr", type=float, default=20, help="Maximum SNR for isotropic noise")

    def __init__(self, pool, rir=None, noise=None, batch_size=32, use_stft=True, stft_frame_length=512, stft_frame_shift=256,
                 stft_window_type="hann", min_window_size=2, max_window_size=5, min_snr=0, max_snr=10, num_workers=0,
                 ipd_index=None, seed=None, window_frames=None):
        """pool: UtterancePool (the reference takes a lhotse manifest path); rir: (n_rirs, taps) tensor or None; noise:
        (C, n) tensor or None.  ipd_index: 7-ch training features (mag + IPD, as FeatureExtractor); None: magnitudes.
        window_frames: fix the window to this many STFT frames (BASELINE config 3 uses 250) instead of sampling a size."""
        super().__init__()
        if stft_frame_length != 512 or stft_frame_shift != 256 or stft_window_type != "hann":
            raise ValueError("the sm_100a FFT kernel is built for a 512/256 Hann STFT (conf/*.yaml)")
        self.pool = pool
        self.device = pool.data.device
        self.batch_size = batch_size
        self.use_stft = use_stft
        self.min_window_size, self.max_window_size = min_window_size, max_window_size
        self.min_snr, self.max_snr = min_snr, max_snr
        self.window_frames = window_frames
        self.rng = random.Random(seed)
        self.rir = rir.float().to(self.device).contiguous() if rir is not None else None
        self.noise = None
        if noise is not None:
            nz = noise if noise.dim() == 2 else noise.unsqueeze(0)
            if nz.shape[0] not in (1, pool.channels):
                raise ValueError("noise must have 1 channel or as many as the utterances")
            self.noise = nz.float().expand(pool.channels, -1).to(self.device).contiguous()
        self.extractor = FeatureExtractor(stft_frame_length, stft_frame_shift, ipd_index=ipd_index) if use_stft else None
        self.last_plan = None

    def __iter__(self):
        return self

    # ------------------------------------------------------------------ host: the sampling decisions of separation.py:175-236
    def sample_plan(self):
        sr = self.pool.sampling_rate
        rng = self.rng
        window_size = rng.uniform(self.min_window_size, self.max_window_size)
        if self.window_frames is not None:
            window_size = (512 + 256 * (self.window_frames - 1)) / sr
        W = int(window_size * sr)
        items, total_length, total_overlap = [], 0.0, 0.0
        while len(items) < self.batch_size:
            spk1, spk2 = rng.sample(self.pool.speaker_ids, 2)
            c1 = rng.choice(self.pool.by_speaker[spk1])
            c2 = rng.choice(self.pool.by_speaker[spk2])
            if self.pool.duration(c1) < self.pool.duration(c2):
                c1, c2 = c2, c1
            d1, d2 = self.pool.duration(c1), self.pool.duration(c2)
            start2 = rng.uniform(0, d1 / 2)
            end2 = start2 + d2
            mix_end = max(d1, end2)
            total_length += mix_end
            total_overlap += min(d2, d1 - start2)
            if mix_end < window_size:
                continue
            for k in range(int(mix_end / window_size)):
                items.append((self.pool.offsets[c1], self.pool.lengths[c1], self.pool.offsets[c2], self.pool.lengths[c2],
                              int(round(start2 * sr)), k * W))
        items = items[: self.batch_size]
        B = len(items)
        rir_idx = [-1] * B
        if self.rir is not None:
            rir_idx = [rng.randrange(self.rir.shape[0]) if rng.random() < 0.5 else -1 for _ in range(B)]
        snr, noise_off = [float("nan")] * B, [0] * B
        if self.noise is not None:
            for b in range(B):
                if rng.random() < 0.5:
                    snr[b] = rng.uniform(self.min_snr, self.max_snr)
                    noise_off[b] = rng.randrange(self.noise.shape[1])
        return {"W": W, "items": items, "rir_idx": rir_idx, "snr": snr, "noise_off": noise_off,
                "ovl": total_overlap / total_length, "window_size": window_size}

    # ------------------------------------------------------------------ device
    def mix(self, plan):
        """-> mix (B, C, W), s1 (B, W), s2 (B, W) on the GPU for a plan from sample_plan()."""
        dev, C, W = self.device, self.pool.channels, plan["W"]
        B = len(plan["items"])
        items = torch.tensor(plan["items"], dtype=torch.int64).to(dev)
        mix = torch.empty(B, C, W, dtype=torch.float32, device=dev)
        s1 = torch.empty(B, W, dtype=torch.float32, device=dev)
        s2 = torch.empty(B, W, dtype=torch.float32, device=dev)
        nv.call("css_mix_windows", nv.ptr(self.pool.data), nv.c_long(self.pool.total), nv.c_int(C), nv.ptr(items), nv.c_int(B),
                nv.c_int(W), nv.ptr(mix), nv.ptr(s1), nv.ptr(s2))
        if self.rir is not None and any(r >= 0 for r in plan["rir_idx"]):
            idx = torch.tensor(plan["rir_idx"], dtype=torch.int32).to(dev)
            wet = mix.clone()
            energy = torch.empty(2 * B, dtype=torch.float64, device=dev)
            nv.call("css_mix_reverb", nv.ptr(mix), nv.ptr(wet), nv.c_int(B), nv.c_int(C), nv.c_int(W), nv.ptr(self.rir),
                    nv.c_int(self.rir.shape[1]), nv.ptr(idx), nv.ptr(energy))
            mix = wet
        if self.noise is not None and any(not math.isnan(s) for s in plan["snr"]):
            off = torch.tensor(plan["noise_off"], dtype=torch.int64).to(dev)
            snr = torch.tensor(plan["snr"], dtype=torch.float32).to(dev)
            nv.call("css_mix_add_noise", nv.ptr(mix), nv.c_int(B), nv.c_int(C), nv.c_int(W), nv.ptr(self.noise),
                    nv.c_long(self.noise.shape[1]), nv.c_long(self.noise.shape[1]), nv.ptr(off), nv.ptr(snr))
        return mix, s1, s2

    def __next__(self):
        plan = self.sample_plan()
        self.last_plan = plan
        mix, s1, s2 = self.mix(plan)
        B = mix.shape[0]
        lens = torch.full((B,), plan["window_size"])
        if self.extractor is None:
            return {"mix": mix.squeeze(1) if mix.shape[1] == 1 else mix, "lens": lens, "source1": s1, "source2": s2, "ovl": plan["ovl"]}
        # features as the model consumes them: (B, T, F) mixture features, (B, T, 257) source magnitudes
        _, f, _, _ = self.extractor(mix if mix.shape[1] > 1 else mix[:, 0])
        m1 = self.extractor.forward_stft.forward(s1)[0]
        m2 = self.extractor.forward_stft.forward(s2)[0]
        if self.extractor.ipd_index is None:
            f = self.extractor.forward_stft.forward(mix[:, 0])[0]          # magnitude spectrogram of the mixture
        return {"mix": f.transpose(1, 2).contiguous(), "lens": lens, "source1": m1.transpose(1, 2).contiguous(),
                "source2": m2.transpose(1, 2).contiguous(), "ovl": plan["ovl"]}
