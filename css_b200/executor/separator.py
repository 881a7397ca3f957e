"""Chunked mask estimation (reference: css/executor/separator.py).

``Separator`` keeps the reference constructor/``separate``/``separate_one_batch`` surface, but the
whole recording is one device-side batch: a single global STFT (chunk frame t == global frame
50*n + t), fused features, ONE mask-network launch over all chunks, clamp and the
DOA merge -- no per-chunk Python loop and no numpy round trip.

backend "b200": ``model`` is a ``css_b200.models`` Conformer/BLSTM on a CUDA device (or
``model_path`` names a checkpoint saved by ``css_b200.utils.model_util.save_model``).
backend "onnx" (the reference's only backend, separator.py:26-34): ``model_path`` names an ONNX file; its weights are
imported into the native model (css_b200/utils/onnx_import.py), no onnxruntime involved.  As in the reference the ONNX
backend asserts ``batch_size == 1`` (separator.py:21-24).
"""
import math
import os

import torch

from .. import _native as nv
from ..utils.model_util import B200Session, get_b200_model
from .feature import FeatureExtractor, features, stft, stft_range

CHUNK_EXTRA = 256   # separator.py:67


def steervec_7ch(nfreqs=257, nvecs=30, radius=0.0425, sndvelocity=340, samplerate=16000):
    """Far-field steering vectors of the 7-mic circular array (separator.py:113-163).  Host, one-off."""
    ang = torch.tensor([2 * math.pi * k / nvecs for k in range(nvecs)])
    shifts = (math.pi / 6, -math.pi / 6, -math.pi / 2, -5 * math.pi / 6, 5 * math.pi / 6, math.pi / 2)
    dist = radius * torch.stack([torch.zeros(nvecs)] + [torch.cos(ang + s) for s in shifts], dim=1)
    deltas = dist / sndvelocity * samplerate
    f = torch.arange(nfreqs, dtype=torch.float32).view(-1, 1, 1)
    sv = torch.exp(1j * deltas.unsqueeze(0) * math.pi * f / (nfreqs - 1)) / math.sqrt(7)
    return sv.to(torch.complex64), ang * 360 / (2 * math.pi)


class Separator:
    def __init__(self, separation_config, sr=16000, sess_options=None, device="cuda", backend="b200",
                 lowcut=80, highcut=2000, model=None, max_chunks_per_launch=2048, stream_first_fraction=0.125):
        if backend not in ("b200", "native", "onnx"):
            raise ValueError(f"Unknown backend: {backend}")
        if backend == "onnx":
            assert separation_config["batch_size"] == 1, "ONNX backend only supports batch size 1"      # separator.py:21-24
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise nv.NativeError("css_b200 runs only on CUDA devices (no CPU fallback)")
        if model is None:
            path = separation_config["model_path"]
            if backend == "onnx" and not str(path).endswith(".onnx"):
                raise ValueError(f"backend 'onnx' expects an .onnx model_path, got {path}")
            model = get_b200_model(path, sess_options, self.device)[0]
        self.separator = model if isinstance(model, B200Session) else B200Session(model)
        self.model = self.separator.model
        self.batch_size = separation_config["batch_size"]
        self.eval_win = int(separation_config["eval_win"] * sr)
        self.eval_hop = int(separation_config["eval_hop"] * sr)
        self.sr = sr
        self.frame_length = separation_config["frame_length"]
        self.frame_shift = separation_config["frame_shift"]
        self.ipd = separation_config.get("ipd", None)
        self.feature = FeatureExtractor(frame_len=self.frame_length, frame_hop=self.frame_shift, ipd_index=self.ipd)
        if self.feature.num_features != self.model.in_features:
            raise ValueError(f"model expects {self.model.in_features} features, config yields {self.feature.num_features}")
        if (self.eval_win + CHUNK_EXTRA - self.frame_length) % self.frame_shift or self.eval_hop % self.frame_shift:
            raise ValueError("eval_win/eval_hop must be whole STFT hops")
        self.chunk_len = self.eval_win + CHUNK_EXTRA
        self.chunk_frames = (self.chunk_len - self.frame_length) // self.frame_shift + 1
        self.hop_frames = self.eval_hop // self.frame_shift
        self.max_chunks = max_chunks_per_launch
        # spectrum -> features -> MVN -> planes fused in front of the network (css_conformer_forward_spec); off = the two-call
        # form css_features + css_conformer_forward (A/B switch, and what models without the fused entry point use)
        self.fused_front_end = os.environ.get("CSS_B200_FUSED_FRONT", "1") != "0"
        # a pinned host recording is copied in pieces: a small first one, whose chunks keep the GPU busy while the rest
        # arrives (the mask network of 1/8 of the chunks takes about as long as the copy of the other 7/8)
        self.stream_first_fraction = float(stream_first_fraction)
        self._copy_stream = None
        # DOA merge; the 1-ch config has no `merge` keys (defect D1) -> default off
        self.merge = separation_config.get("merge", False)
        self.merge_threshold = separation_config.get("merge_threshold", 16)
        if self.merge:
            sv, angles = steervec_7ch(nfreqs=257, nvecs=30, samplerate=sr)
            freq_step = (sr // 2) / (257 - 1)
            self._lo = int(math.floor(lowcut / freq_step))
            self._hi = int(math.ceil(highcut / freq_step))
            self.steervec, self.angles = sv, angles
            svT = torch.view_as_real(sv[self._lo:self._hi].permute(1, 2, 0).contiguous())   # (30,7,nf,2)
            self._svT = svT.to(self.device)
            self._angles = angles.float().to(self.device)
        self.last = {}

    # ------------------------------------------------------------------
    def n_chunks(self, T):
        return 0 if T < self.chunk_len else (T - self.chunk_len) // self.eval_hop + 1

    def _batches(self, B, first=0):
        """[(b0, nb)]: an optional short first batch, then launches of at most max_chunks_per_launch chunks."""
        out, b0 = [], 0
        if 0 < first < B:
            out.append((0, first))
            b0 = first
        while b0 < B:
            nb = min(self.max_chunks, B - b0)
            out.append((b0, nb))
            b0 += nb
        return out

    def _masks(self, spec, stride_d, stride_chunk, B, D, batches=None, before_batch=None):
        il, ir = self.feature.ipd_tensors(spec.device)
        ldin = self.model.engine()["ldin"]
        T = self.chunk_frames
        mags, masks = [], []
        for i, (b0, nb) in enumerate(batches or self._batches(B)):
            if before_batch is not None:
                before_batch(i, b0, nb)
            sub = spec.reshape(-1)[b0 * stride_chunk * nv.KP * 2:]
            fused = self.model.masks_from_spec(sub, stride_d, stride_chunk, nb, T, il, ir, clamp_max=1.0,
                                               shared_right=len(set(self.feature.ipd_r)) <= 1) \
                if self.fused_front_end and hasattr(self.model, "masks_from_spec") else None
            if fused is not None:
                m, mag = fused
            else:
                mag, feat = features(sub, stride_d, stride_chunk, nb, T, il, ir, ldin)
                m = self.model.masks_layout(feat, nb, T, clamp_max=1.0)
            if self.merge:
                like = torch.empty(nb, 2, 30, dtype=torch.float32, device=spec.device)
                idx = torch.empty(nb, 3, dtype=torch.int32, device=spec.device)
                nv.call("css_doa_merge", nv.ptr(sub), nv.c_long(stride_d), nv.c_long(stride_chunk), nv.c_long(nb),
                        nv.c_int(T), nv.ptr(self._svT), nv.ptr(self._angles), nv.c_int(self._lo), nv.c_int(self._hi),
                        nv.c_float(float(self.merge_threshold)), nv.ptr(m), nv.c_int(m.shape[-1]), nv.ptr(like), nv.ptr(idx))
                self.last.setdefault("doa", []).append((like, idx))
            mags.append(mag)
            masks.append(m)
        if len(masks) == 1:
            return masks[0], mags[0]
        return torch.cat(masks, 0), torch.cat(mags, 0)

    def separate(self, s):
        """s: (D, T) waveform (CPU or CUDA) -> masks (B,257,150,3), mags (B,257,150) on the GPU."""
        if s.dim() != 2:
            raise RuntimeError(f"separate expects a (D, T) waveform, got {tuple(s.shape)}")
        D, T = s.shape
        B = self.n_chunks(T)
        if B < 1:
            raise RuntimeError(f"waveform of {T} samples is shorter than one window of {self.chunk_len}")
        L = self.hop_frames * (B - 1) + self.chunk_frames
        if not s.is_cuda and s.is_pinned() and s.dtype == torch.float32 and self.stream_first_fraction > 0 and B >= 64:
            return self._separate_streamed(s, B, L)
        s = s.to(self.device, torch.float32, non_blocking=True)
        spec = stft(s, n_frames=L)
        self.last = {"spec": spec, "wav": s, "n_chunks": B, "frames": L}
        return self._masks(spec, L, self.hop_frames, B, D)

    def _separate_streamed(self, s, B, L):
        """Pinned host recording: the H2D copy runs in pieces (one per batch of chunks) on a copy stream while the STFT, the
        features and the mask network of the chunks already on the device run on the compute stream (each batch of
        chunks waits only for the piece that holds its last sample).  Same arithmetic as the resident path: the
        STFT frames of a batch are written into the one global spec buffer (frames shared with the previous batch
        are recomputed to the same values)."""
        D, T = s.shape
        dev = self.device
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)
        wav = torch.empty(D, T, dtype=torch.float32, device=dev)
        spec = torch.empty(D, L, nv.KP, 2, dtype=torch.float32, device=dev)
        batches = self._batches(B, first=max(1, int(B * self.stream_first_fraction)))
        events = []
        self._copy_stream.wait_stream(main)            # wav/spec allocations are ordered before the copies
        with torch.cuda.stream(self._copy_stream):
            lo = 0
            for b0, nb in batches:
                last = b0 + nb - 1
                hi = T if last == B - 1 else last * self.eval_hop + self.chunk_len
                for d in range(D):
                    wav[d, lo:hi].copy_(s[d, lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
                events.append(ev)
                lo = hi
        wav.record_stream(self._copy_stream)

        def before_batch(i, b0, nb):
            main.wait_event(events[i])
            f0 = b0 * self.hop_frames
            f1 = (b0 + nb - 1) * self.hop_frames + self.chunk_frames
            stft_range(wav, spec, f0, f1)

        self.last = {"spec": spec, "wav": wav, "n_chunks": B, "frames": L}
        return self._masks(spec, L, self.hop_frames, B, D, batches=batches, before_batch=before_batch)

    def separate_one_batch(self, s):
        """s: (B, D, T) batch of windows -> masks (B,257,T',3), mags (B,257,T')."""
        if s.dim() != 3:
            raise RuntimeError(f"separate_one_batch expects (B, D, T), got {tuple(s.shape)}")
        s = s.to(self.device, torch.float32)
        B, D, N = s.shape
        if N != self.chunk_len:
            raise RuntimeError(f"windows must be {self.chunk_len} samples long, got {N}")
        spec = stft(s.reshape(B * D, N))
        self.last = {}
        return self._masks(spec, self.chunk_frames, D * self.chunk_frames, B, D)

    def angle_merge_debug(self):
        """(likelihood (B,2,30), idx (B,3) = argmax0, argmax1, merged) of the last call, or None."""
        if "doa" not in self.last:
            return None
        like = torch.cat([d[0] for d in self.last["doa"]], 0)
        idx = torch.cat([d[1] for d in self.last["doa"]], 0)
        return like, idx
