"""Mask-based Souden MVDR + iSTFT + slice placement (reference: css/executor/beamformer.py).

The per-chunk centred STFT of the reference is served from ONE global uncentred STFT (centred frame
t>=1 of chunk n == global frame 50n+t-1; frame 0 is a separate reflect-padded FFT), the three masked
covariances are accumulated once per (chunk, bin), the 7x7 complex128 solves run in registers/local
memory, and only the frames consumed by the placement are inverse-transformed.  The reference's
placement quirks (seam double-add, ``en == T`` branch, zero tail -- SURVEY.md A.6) are reproduced by
``placement_plan``.
"""
import math

import torch

from .. import _native as nv
from .feature import stft, stft_edge

_SUPPORTED = ("SoudenMVDRBeamformer",)


def placement_plan(n_chunks, T, hop, margin, wav_len, batch_size):
    """Slice additions of beamformer.py:54-57,104-118 as (chunk, src_start, src_end, dst_start), in order."""
    plan = []
    for i in range(math.ceil(n_chunks / batch_size)):
        first = i * batch_size
        last = min((i + 1) * batch_size, n_chunks) + 1      # the reference's overlapping batch end
        st, en = hop * first, hop * last
        for c in range(first, min(last, n_chunks)):
            if st == 0:
                plan.append((c, 0, margin, 0))
            elif en == T:
                plan.append((c, margin - hop, wav_len, st + margin - hop))
            else:
                plan.append((c, margin - hop, margin, st + margin - hop))
            st += hop
            en += hop
    return plan


def layered_items(plan, T, max_blocks=8):
    """Split plan entries into <= 8x256-sample items and group them into layers whose destinations are disjoint."""
    layers, ends = [], []
    for (c, s0, s1, dst) in plan:
        s1 = min(s1, s0 + max(0, T - dst))                  # clip at the end of the output
        if s1 <= s0:
            continue
        for li, e in enumerate(ends):
            if e <= dst:
                break
        else:
            li = len(ends)
            ends.append(0)
            layers.append([])
        ends[li] = dst + (s1 - s0)
        a = s0
        while a < s1:
            b = min(s1, ((a >> 8) + max_blocks) << 8)
            layers[li].append((c, a, b, dst + (a - s0)))
            a = b
    return layers


class Beamformer:
    def __init__(self, beamforming_config, sr=16000, rescale=True):
        self.type = beamforming_config["type"]
        if self.type not in _SUPPORTED:
            raise ValueError(f"Beamformer type '{self.type}' not found")
        self.n_fft = beamforming_config["n_fft"]
        self.hop_length = beamforming_config["hop_size"]
        if self.n_fft != 512 or self.hop_length != 256:
            raise ValueError("the sm_100a FFT kernels are built for n_fft 512 / hop_size 256 (conf/*.yaml)")
        self.eval_win = int(beamforming_config["eval_win"] * sr)
        self.eval_hop = int(beamforming_config["eval_hop"] * sr)
        self.proceed_margin = beamforming_config["proceed_margin"] * sr
        self.batch_size = beamforming_config["batch_size"]
        self.sr = sr
        self.mask_hop = int(self.eval_hop / self.hop_length)
        self.mask_win = int(self.eval_win / self.hop_length)
        self.rescale = rescale
        self._plans = {}

    # ------------------------------------------------------------------ native pieces
    def _mvdr(self, spec, edge, stitched, B, frames):
        """-> Y (B, 2, mask_win, KP, 2) raw beamformed spectra, energy partials."""
        dev = spec.device
        T = self.mask_win
        nblk = nv.lib().css_mvdr_energy_blocks()
        Y = torch.empty(B, 2, T, nv.KP, 2, dtype=torch.float32, device=dev)
        energy = torch.empty(B, 2, nblk, 2, dtype=torch.float32, device=dev)
        nv.call("css_mvdr", nv.ptr(spec), nv.ptr(edge), nv.c_long(frames), nv.c_long(B), nv.ptr(stitched),
                nv.c_long(stitched.shape[1]), nv.c_int(self.mask_hop), nv.c_int(T), nv.ptr(Y), nv.ptr(energy))
        return Y, energy

    def _finalize(self, Y, energy, dedup):
        B = Y.shape[0]
        scal = torch.empty(B, 5, dtype=torch.float32, device=Y.device)
        nv.call("css_mvdr_finalize", nv.ptr(Y), nv.ptr(energy), nv.c_long(B), nv.c_int(self.mask_win),
                nv.c_int(1 if self.rescale else 0), nv.c_int(1 if dedup else 0), nv.ptr(scal))
        return scal

    def _plan(self, B, T, device):
        key = (B, T, str(device))
        if key not in self._plans:
            wav_len = self.hop_length * (self.mask_win - 1)
            plan = placement_plan(B, T, self.eval_hop, self.proceed_margin, wav_len, self.batch_size)
            layers = layered_items(plan, T)
            self._plans[key] = [torch.tensor(l, dtype=torch.int32, device=device).reshape(-1, 4) for l in layers]
        return self._plans[key]

    @staticmethod
    def stitched_layout(masks, device):
        """3 x (257, L) tensors -> [3][L][KP] fp32 (reuses the stitcher's buffer when the views come from it)."""
        base = getattr(masks[0], "_css_stitched", None)
        if base is not None and all(getattr(m, "_css_stitched", None) is base for m in masks):
            return base
        L = masks[0].shape[1]
        out = torch.zeros(3, L, nv.KP, dtype=torch.float32, device=device)
        for c in range(3):
            out[c, :, : nv.NBINS] = masks[c].to(device, torch.float32).t()
        return out

    # ------------------------------------------------------------------ reference surface
    def continuous_process(self, X, masks, spec=None):
        """X: (D, T) waveform; masks: (speech1, speech2, noise) each (257, L) -> two (1, T) waveforms (CUDA).

        reference: beamformer.py:32-124.  ``spec`` may pass the Separator's cached global STFT.
        """
        dev = masks[0].device if masks[0].is_cuda else torch.device("cuda")
        X = X.to(dev, torch.float32, non_blocking=True)
        D, T = X.shape
        if D != 7:
            raise RuntimeError(f"the MVDR kernel is built for the 7-microphone array, got {D} channels")
        chunk_len = self.eval_win + 256
        B = 0 if T < chunk_len else (T - chunk_len) // self.eval_hop + 1
        if B < 1:
            raise RuntimeError(f"waveform of {T} samples is shorter than one window of {chunk_len}")
        stitched = self.stitched_layout(masks, dev)
        Y = self.beamform_chunks(X, stitched, B, spec)
        out = torch.zeros(2, T, dtype=torch.float32, device=dev)
        self.place(Y, self._plan(B, T, dev), out)
        self.peak_normalize(out)
        return out[0:1], out[1:2]

    def beamform_chunks(self, X, stitched, B, spec=None):
        """MVDR + rescale + dedup of chunks 0..B-1 of X (beamformer.py:60-80) -> Y (B, 2, mask_win, KP, 2)."""
        frames = self.mask_hop * (B - 1) + self.mask_win
        if stitched.shape[1] < frames:
            raise RuntimeError(f"masks cover {stitched.shape[1]} frames, need {frames}")
        if spec is None:
            spec = stft(X, n_frames=frames)
        edge = stft_edge(X, B, self.eval_hop)
        Y, energy = self._mvdr(spec, edge, stitched, B, spec.shape[1])
        self.last_scal = self._finalize(Y, energy, dedup=True)
        return Y

    def place(self, Y, layers, out):
        """iSTFT + slice placement (beamformer.py:83-118) of the plan layers into out (2, T), accumulated."""
        for items in layers:
            if items.shape[0]:
                nv.call("css_istft_place", nv.ptr(Y), nv.c_int(self.mask_win), nv.ptr(items), nv.c_int(items.shape[0]),
                        nv.ptr(out), nv.c_long(out.shape[1]))

    @staticmethod
    def peak_normalize(out):
        """beamformer.py:122-123: every output scaled to a peak of 0.9, in place."""
        scratch = torch.empty(out.shape[0], dtype=torch.int32, device=out.device)
        nv.call("css_peak_normalize", nv.ptr(out), nv.c_int(out.shape[0]), nv.c_long(out.shape[1]), nv.c_float(0.9), nv.ptr(scratch))

    def normalize_scale(self, result, ref_spec, mask):
        """result (B,F,T) complex, ref_spec (B,F,T) complex (channel 0), mask (B,F,T) real -> result scaled so that its rms
        over (F,T) equals the rms of mask * ref_spec, per chunk (beamformer.py:165-183).  Inside continuous_process the same
        scale is applied by css_mvdr_finalize from energies reduced in the MVDR kernel; this method serves callers of the
        reference surface that hold the spectra themselves."""
        nv.require_cuda(result, "result")
        dev = result.device
        B = result.shape[0]
        r = torch.view_as_real(result.to(torch.complex64).contiguous())
        x = torch.view_as_real(ref_spec.to(dev, torch.complex64).contiguous())
        m = mask.to(dev, torch.float32).contiguous()
        if x.shape != r.shape or m.shape != result.shape:
            raise RuntimeError(f"normalize_scale: shapes differ {tuple(result.shape)} {tuple(ref_spec.shape)} {tuple(mask.shape)}")
        out = torch.empty_like(r)
        nv.call("css_normalize_scale", nv.ptr(r), nv.ptr(x), nv.ptr(m), nv.c_long(B), nv.c_long(result[0].numel()), nv.ptr(out))
        return torch.view_as_complex(out)

    def process_one_batch(self, x, speech_mask, noise_mask):
        """x: (B, D, N) windows, masks (B, F, T) -> beamformed (B, F, T) complex64 (beamformer.py:126-163)."""
        nv.require_cuda(x, "x")
        B, D, N = x.shape
        if D != 7 or N != self.eval_win + 256:
            raise RuntimeError(f"process_one_batch expects (B, 7, {self.eval_win + 256}) windows, got {tuple(x.shape)}")
        T = self.mask_win
        dev = x.device
        # treat every window as its own one-chunk "recording": chunk b of channel d is signal d*B+b
        sig = x.float().permute(1, 0, 2).reshape(D * B, N)
        nfr = (N - 512) // 256 + 1
        spec = stft(sig)                                  # (D*B, nfr, KP, 2) ; frames of window b contiguous
        edge = stft_edge(sig, 1, self.eval_hop)           # (D*B, 1, KP, 2) == [d][b][KP]
        stitched = torch.zeros(3, B * nfr, nv.KP, dtype=torch.float32, device=dev)
        # speech mask for both "speakers"; output 0 is the requested beam
        sm = speech_mask.to(dev, torch.float32).permute(0, 2, 1)[:, :T]
        nm = noise_mask.to(dev, torch.float32).permute(0, 2, 1)[:, :T]
        view = stitched.view(3, B, nfr, nv.KP)
        view[0, :, :T, : nv.NBINS] = sm
        view[1, :, :T, : nv.NBINS] = sm
        view[2, :, :T, : nv.NBINS] = nm
        # spec as [7][B*nfr][KP] with hop_frames = nfr: chunk b frame t-1 -> global frame b*nfr + t-1
        nblk = nv.lib().css_mvdr_energy_blocks()
        Y = torch.empty(B, 2, T, nv.KP, 2, dtype=torch.float32, device=dev)
        energy = torch.empty(B, 2, nblk, 2, dtype=torch.float32, device=dev)
        nv.call("css_mvdr", nv.ptr(spec), nv.ptr(edge), nv.c_long(B * nfr), nv.c_long(B), nv.ptr(stitched),
                nv.c_long(B * nfr), nv.c_int(nfr), nv.c_int(T), nv.ptr(Y), nv.ptr(energy))
        if self.rescale:
            scal = torch.empty(B, 5, dtype=torch.float32, device=dev)
            nv.call("css_mvdr_finalize", nv.ptr(Y), nv.ptr(energy), nv.c_long(B), nv.c_int(T), nv.c_int(1), nv.c_int(0), nv.ptr(scal))
        return torch.view_as_complex(Y[:, 0, :, : nv.NBINS]).transpose(1, 2)
