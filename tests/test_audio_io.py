"""Audio IO against independent implementations: scipy.io.wavfile and torchaudio's scaling rule."""
import numpy as np
import pytest
import torch
from scipy.io import wavfile

from css_b200.utils.audio_util import EgsReader, read_wav, write_wav


def test_pcm16_roundtrip_matches_scipy(tmp_path):
    g = np.random.default_rng(0)
    pcm = g.integers(-32768, 32767, size=(4000, 7), dtype=np.int16)
    p = tmp_path / "a.wav"
    wavfile.write(p, 16000, pcm)
    x, sr = read_wav(p, pin=False)
    assert sr == 16000 and x.shape == (7, 4000) and x.dtype == torch.float32
    assert np.array_equal(x.numpy(), pcm.T.astype(np.float32) / 32768.0)      # torchaudio/lhotse scaling
    q = tmp_path / "b.wav"
    write_wav(q, x, sr)
    sr2, back = wavfile.read(q)
    assert sr2 == 16000 and np.array_equal(back, pcm)                          # bit-exact PCM round trip


@pytest.mark.parametrize("dtype", [np.int32, np.float32])
def test_other_encodings(tmp_path, dtype):
    g = np.random.default_rng(1)
    a = (g.standard_normal((1000, 2)) * 0.1).astype(np.float32)
    data = a if dtype == np.float32 else (a * 2147483648.0).astype(np.int32)
    p = tmp_path / "c.wav"
    wavfile.write(p, 8000, data)
    x, sr = read_wav(p, pin=False)
    ref = a if dtype == np.float32 else data.astype(np.float32) / 2147483648.0
    assert sr == 8000 and np.array_equal(x.numpy(), ref.T)


def test_write_clips_and_rounds_like_torchaudio(tmp_path):
    x = torch.tensor([[0.0, 1.0, -1.0, 1.5, -1.5, 0.5 / 32768.0 * 3]])
    p = tmp_path / "d.wav"
    write_wav(p, x, 16000)
    _, back = wavfile.read(p)
    assert back.tolist() == [0, 32767, -32768, 32767, -32768, 2]


def test_egs_reader_protocol(tmp_path):
    pcm = (np.arange(7 * 50, dtype=np.int16).reshape(50, 7) - 100)
    wavfile.write(tmp_path / "s0.wav", 16000, pcm)
    multi = list(EgsReader({"session0": tmp_path / "s0.wav"}, multi_channel=True))
    single = list(EgsReader({"session0": tmp_path / "s0.wav"}, multi_channel=False))
    assert multi[0][0] == "session0" and multi[0][1]["mix"].shape == (7, 50)
    assert single[0][1]["mix"].shape == (50,) and torch.equal(single[0][1]["mix"], multi[0][1]["mix"][0])


def test_rejects_non_wav(tmp_path):
    p = tmp_path / "x.wav"
    p.write_bytes(b"not a wav file at all")
    with pytest.raises(ValueError):
        read_wav(p)
