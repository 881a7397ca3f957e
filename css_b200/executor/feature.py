"""Spectral + spatial features (reference: css/executor/feature.py).

``FeatureExtractor.forward`` keeps the reference's return layout (mag (B,F,T), f (B,*,T), re/im
(B,D,F,T)); the numbers come from the framing + shared-memory FFT kernel (``css_stft``) and the fused
MVN/IPD kernel (``css_features``).  The executor's fast path uses ``stft``/``features`` directly
and stays in the bin-fastest device layout.
"""
import ctypes

import torch

from .. import _native as nv

EPSILON = torch.finfo(torch.float32).eps


def parse_ipd(ipd_index):
    pairs = [tuple(map(int, p.split(","))) for p in ipd_index.split(";")]
    return [t[0] for t in pairs], [t[1] for t in pairs]


def stft(x, n_frames=None):
    """x: (n_sig, N) fp32 CUDA (row stride may exceed N) -> spec (n_sig, n_frames, KP) complex64 (as float2)."""
    nv.require_cuda(x, "x")
    n_sig, N = x.shape
    if n_frames is None:
        n_frames = (N - 512) // 256 + 1
    if N < 512 or n_frames < 1:
        raise RuntimeError(f"signal of {N} samples is shorter than one 512-sample frame")
    if x.stride(1) != 1:
        x = x.contiguous()
    spec = torch.empty(n_sig, n_frames, nv.KP, 2, dtype=torch.float32, device=x.device)
    nv.call("css_stft", nv.ptr(x), nv.c_long(n_sig), nv.c_long(x.stride(0)), nv.c_long(n_frames), nv.ptr(spec))
    return spec


def stft_range(x, spec, f0, f1):
    """Frames [f0, f1) of x (n_sig, N) into the matching rows of spec (n_sig, L, KP, 2), in place."""
    n_sig, L = spec.shape[0], spec.shape[1]
    nv.call("css_stft_range", ctypes.c_void_p(x.data_ptr() + 4 * 256 * f0), nv.c_long(n_sig), nv.c_long(x.stride(0)), nv.c_long(f1 - f0),
            ctypes.c_void_p(spec.data_ptr() + 8 * nv.KP * f0), nv.c_long(L))


def stft_edge(x, n_chunks, chunk_hop):
    """Centred-STFT frame 0 (reflect pad) at every chunk start: (n_sig, n_chunks, KP, 2)."""
    nv.require_cuda(x, "x")
    if x.stride(1) != 1:
        x = x.contiguous()
    edge = torch.empty(x.shape[0], n_chunks, nv.KP, 2, dtype=torch.float32, device=x.device)
    nv.call("css_stft_edge", nv.ptr(x), nv.c_long(x.shape[0]), nv.c_long(x.stride(0)), nv.c_long(n_chunks),
            nv.c_long(chunk_hop), nv.ptr(edge))
    return edge


def features(spec, stride_d, stride_chunk, n_chunks, T, ipd_l, ipd_r, ldf):
    """-> mag (B,257,T), feat (B*T, ldf) zero padded.  ipd_l/ipd_r: int32 CUDA tensors (may be empty)."""
    dev = spec.device
    mag = torch.empty(n_chunks, nv.NBINS, T, dtype=torch.float32, device=dev)
    feat = torch.empty(n_chunks * T, ldf, dtype=torch.float32, device=dev)
    nv.call("css_features", nv.ptr(spec), nv.c_long(stride_d), nv.c_long(stride_chunk), nv.c_long(n_chunks), nv.c_int(T),
            nv.ptr(ipd_l), nv.ptr(ipd_r), nv.c_int(int(ipd_l.numel())), nv.ptr(mag), nv.ptr(feat), nv.c_int(ldf))
    return mag, feat


class STFT:
    """Uncentred Hann STFT, frame 512 / hop 256 (reference: css/executor/feature.py:33-82)."""

    def __init__(self, frame_len=512, frame_hop=256, round_pow_of_two=True):
        if frame_len != 512 or frame_hop != 256:
            raise ValueError("the sm_100a FFT kernel is built for frame_len 512 / frame_hop 256 (conf/*.yaml)")
        self.stride = frame_hop
        self.num_bins = frame_len // 2 + 1

    def forward(self, x):
        if x.dim() not in [2, 3]:
            raise RuntimeError("STFT expect 2D/3D tensor, but got {:d}D signal".format(x.dim()))
        nv.require_cuda(x, "x")
        shape = x.shape
        spec = stft(x.reshape(-1, shape[-1]).float())
        c = torch.view_as_complex(spec[:, :, : nv.NBINS]).transpose(1, 2)   # (n_sig, F, T)
        c = c.reshape(*shape[:-1], nv.NBINS, c.shape[-1])
        r, i = c.real, c.imag
        return torch.abs(c), torch.atan2(i, r), r, i


class FeatureExtractor(torch.nn.Module):
    """reference: css/executor/feature.py:137-175"""

    def __init__(self, frame_len=512, frame_hop=256, round_pow_of_two=True, ipd_index=None):
        super().__init__()
        self.forward_stft = STFT(frame_len, frame_hop, round_pow_of_two)
        self.num_bins = self.forward_stft.num_bins
        self.feature_dim = self.num_bins
        self.ipd_index = ipd_index
        self.ipd_l, self.ipd_r = parse_ipd(ipd_index) if ipd_index else ([], [])
        self._dev = {}

    def ipd_tensors(self, device):
        if device not in self._dev:
            self._dev[device] = (torch.tensor(self.ipd_l, dtype=torch.int32, device=device),
                                 torch.tensor(self.ipd_r, dtype=torch.int32, device=device))
        return self._dev[device]

    @property
    def num_features(self):
        return self.num_bins * (1 + len(self.ipd_l))

    def forward(self, x):
        """x: (B, N) or (B, D, N) CUDA -> mag (B,F,T), f (B,*,T), re, im ((B,[D,]F,T))."""
        nv.require_cuda(x, "x")
        if x.dim() not in [2, 3]:
            raise RuntimeError("FeatureExtractor expect 2D/3D tensor, but got {:d}D signal".format(x.dim()))
        multi = x.dim() == 3
        xs = x.float() if multi else x.float().unsqueeze(1)
        B, D, N = xs.shape
        if self.ipd_l and max(self.ipd_l + self.ipd_r) >= D:
            raise RuntimeError(f"ipd index out of range for {D} channels")
        spec = stft(xs.reshape(B * D, N))
        T = spec.shape[1]
        il, ir = self.ipd_tensors(x.device)
        F = self.num_features
        mag, feat = features(spec, T, D * T, B, T, il, ir, F)
        f = feat.reshape(B, T, F).transpose(1, 2)
        c = torch.view_as_complex(spec[:, :, : nv.NBINS]).transpose(1, 2).reshape(B, D, nv.NBINS, T)
        r, i = c.real, c.imag
        if not multi:
            r, i = r[:, 0], i[:, 0]
        return mag, f, r, i
