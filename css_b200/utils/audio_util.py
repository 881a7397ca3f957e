"""Audio IO either side of the hot path (reference: css/utils/audio_util.py:7-54 and the
``torchaudio.save`` calls of scripts/python/separate_libricss.py:81-82).

The reference reads recordings through lhotse (``recordings[key].load_audio()``) and writes with torchaudio; neither is
needed here.  ``read_wav`` decodes RIFF/WAVE (PCM 16/24/32-bit, IEEE float32, plain or WAVE_FORMAT_EXTENSIBLE) straight
into a *pinned* ``(C, N)`` float32 tensor scaled to [-1, 1) -- the layout and scaling lhotse/torchaudio produce -- so the
executor can stream it to the GPU asynchronously (``CssPipeline.run_many``).  ``write_wav`` writes 16-bit PCM with
torchaudio's rounding/clipping convention.  ``WaveReader`` / ``EgsReader`` keep the reference's iteration protocol over
a ``{key: path}`` mapping instead of a lhotse RecordingSet.
"""
import struct

import numpy as np
import torch

_PCM, _FLOAT, _EXTENSIBLE = 1, 3, 0xFFFE


def _chunks(buf):
    if len(buf) < 12 or buf[:4] != b"RIFF" or buf[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file")
    pos = 12
    while pos + 8 <= len(buf):
        cid, size = buf[pos : pos + 4], struct.unpack_from("<I", buf, pos + 4)[0]
        yield cid, pos + 8, size
        pos += 8 + size + (size & 1)


def read_wav(path, channels=None, pin=True):
    """-> (samples (C, N) float32 in [-1, 1), sampling rate).  channels: optional list of channel indices to keep."""
    with open(path, "rb") as f:
        buf = f.read()
    fmt = data = None
    for cid, off, size in _chunks(buf):
        if cid == b"fmt ":
            fmt = struct.unpack_from("<HHIIHH", buf, off)
            if fmt[0] == _EXTENSIBLE and size >= 26:
                fmt = (struct.unpack_from("<H", buf, off + 24)[0],) + fmt[1:]
        elif cid == b"data":
            data = (off, min(size, len(buf) - off))
    if fmt is None or data is None:
        raise ValueError(f"{path}: missing fmt or data chunk")
    code, n_ch, sr, _, block, bits = fmt
    off, size = data
    n = size // block
    raw = np.frombuffer(buf, dtype=np.uint8, count=n * block, offset=off)
    if code == _PCM and bits == 16:
        x = raw.view("<i2").astype(np.float32) / 32768.0
    elif code == _PCM and bits == 32:
        x = raw.view("<i4").astype(np.float32) / 2147483648.0
    elif code == _PCM and bits == 24:
        b = raw.reshape(-1, 3).astype(np.int32)
        v = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
        x = ((v ^ 0x800000) - 0x800000).astype(np.float32) / 8388608.0
    elif code == _FLOAT and bits == 32:
        x = raw.view("<f4").astype(np.float32)
    else:
        raise ValueError(f"{path}: unsupported wav encoding (format {code}, {bits} bits)")
    x = x.reshape(n, n_ch)
    if channels is not None:
        x = x[:, list(channels)]
    out = torch.empty(x.shape[1], n, dtype=torch.float32)
    if pin and torch.cuda.is_available():
        out = out.pin_memory()
    out.numpy()[...] = x.T          # de-interleave into the (C, N) layout the executor reads
    return out, sr


def write_wav(path, x, sr):
    """x: (C, N) or (N,) float tensor/array in [-1, 1] -> 16-bit PCM wav (values clipped, scaled by 32768 as torchaudio.save does)."""
    a = x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)
    if a.ndim == 1:
        a = a[None]
    pcm = np.clip(np.round(a.astype(np.float64) * 32768.0), -32768, 32767).astype("<i2").T.copy()   # (N, C) interleaved
    n_ch, nbytes = pcm.shape[1], pcm.size * 2
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + nbytes) + b"WAVE")
        f.write(b"fmt " + struct.pack("<IHHIIHH", 16, _PCM, n_ch, sr, sr * n_ch * 2, n_ch * 2, 16))
        f.write(b"data" + struct.pack("<I", nbytes))
        f.write(pcm.tobytes())


class WaveReader(object):
    """Sequential reader over ``{key: wav path}`` (reference: audio_util.py:25-54, there over a lhotse RecordingSet)."""

    def __init__(self, recordings, multi_channel=False):
        self.recordings = dict(recordings)
        self.multi_channel = multi_channel

    def _load(self, key):
        x, _ = read_wav(self.recordings[key], channels=None if self.multi_channel else [0])
        return x if self.multi_channel else x[0]        # C x N or N, as the reference

    def __len__(self):
        return len(self.recordings)

    def __contains__(self, key):
        return key in self.recordings

    def __iter__(self):
        for key in self.recordings:
            yield key, self._load(key)


class EgsReader(object):
    """reference: audio_util.py:7-22 -- yields (key, {"mix": samples})."""

    def __init__(self, recordings, multi_channel=False):
        self.mix_reader = WaveReader(recordings, multi_channel=multi_channel)

    def __len__(self):
        return len(self.mix_reader)

    def __iter__(self):
        for key, mix in self.mix_reader:
            yield key, {"mix": mix}
