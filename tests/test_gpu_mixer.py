"""the on-the-fly training mixer on the GPU (css/datasets/separation.py:163-267) against a numpy
restatement of the same steps (lhotse is absent and unpinned: parity unpinned, arithmetic checked here)."""
import math

import numpy as np
import pytest
import torch

from tests.helpers import rel_inf

pytestmark = pytest.mark.gpu


def _pool(C=1, n_utt=12, seed=0):
    from css_b200.datasets import UtterancePool

    g = torch.Generator().manual_seed(seed)
    utts = [torch.randn(C, int(16000 * float(torch.empty(1).uniform_(2.0, 6.0, generator=g))), generator=g) * 0.1 for _ in range(n_utt)]
    return UtterancePool(utts, speakers=[i % 4 for i in range(n_utt)]), utts


def _ref_mix(utts, pool, plan, rir, noise):
    """numpy: pad + mix + window (separation.py:205-241), FIR reverb with energy normalisation, noise at SNR."""
    out, s1s, s2s = [], [], []
    W = plan["W"]
    C = utts[0].shape[0]
    offs = {o: i for i, o in enumerate(pool.offsets)}
    for b, (o1, l1, o2, l2, start2, ws) in enumerate(plan["items"]):
        u1, u2 = utts[offs[o1]].numpy(), utts[offs[o2]].numpy()
        L = max(l1, start2 + l2, ws + W)
        a = np.zeros((C, L), np.float32); a[:, :l1] = u1
        v = np.zeros((C, L), np.float32); v[:, start2:start2 + l2] = u2
        m = (a + v)[:, ws:ws + W]
        s1s.append(a[0, ws:ws + W]); s2s.append(v[0, ws:ws + W])
        r = plan["rir_idx"][b]
        if r >= 0:
            h = rir[r].numpy().astype(np.float64)
            wet = np.stack([np.convolve(m[c].astype(np.float64), h)[:W] for c in range(C)])
            ex, ey = (m.astype(np.float64) ** 2).sum(), (wet ** 2).sum()
            m = (wet * math.sqrt(ex / ey)).astype(np.float32)
        if not math.isnan(plan["snr"][b]):
            nz = noise.numpy()
            idx = (plan["noise_off"][b] + np.arange(W)) % nz.shape[1]
            seg = nz[:, idx]
            g = math.sqrt((m[0].astype(np.float64) ** 2).sum() / ((seg[0].astype(np.float64) ** 2).sum() * 10 ** (plan["snr"][b] / 10)))
            m = m + np.float32(g) * seg
        out.append(m)
    return np.stack(out), np.stack(s1s), np.stack(s2s)


@pytest.mark.parametrize("C", [1, 7])
def test_mixer_matches_numpy(cuda, C):
    from css_b200.datasets import ContinuousSpeechSeparationDataset

    pool, utts = _pool(C=C)
    g = torch.Generator().manual_seed(5)
    rir = torch.randn(3, 600, generator=g) * torch.exp(-torch.arange(600) / 80.0)
    noise = torch.randn(C, 50000, generator=g) * 0.05
    ds = ContinuousSpeechSeparationDataset(pool, rir=rir, noise=noise, batch_size=16, min_snr=5, max_snr=20, seed=3, use_stft=False)
    batch = next(ds)
    plan = ds.last_plan
    assert len(plan["items"]) == 16 and 0.0 < batch["ovl"] <= 1.0
    assert any(r >= 0 for r in plan["rir_idx"]) and any(r < 0 for r in plan["rir_idx"])
    assert any(not math.isnan(s) for s in plan["snr"]) and any(math.isnan(s) for s in plan["snr"])
    mix, s1, s2 = _ref_mix(utts, pool, plan, rir, noise)
    got = batch["mix"].cpu().numpy().reshape(mix.shape)
    assert np.array_equal(batch["source1"].cpu().numpy(), s1) and np.array_equal(batch["source2"].cpu().numpy(), s2)
    dry = [b for b in range(16) if plan["rir_idx"][b] < 0 and math.isnan(plan["snr"][b])]
    assert dry and all(np.array_equal(got[b], mix[b]) for b in dry)          # plain mixing is exact
    assert rel_inf(torch.from_numpy(got), torch.from_numpy(mix)) < 1e-5
    # same seed -> same decisions
    ds2 = ContinuousSpeechSeparationDataset(pool, rir=rir, noise=noise, batch_size=16, min_snr=5, max_snr=20, seed=3, use_stft=False)
    p2 = ds2.sample_plan()
    assert p2["items"] == plan["items"] and p2["rir_idx"] == plan["rir_idx"] and p2["noise_off"] == plan["noise_off"]
    assert all(a == b or (math.isnan(a) and math.isnan(b)) for a, b in zip(p2["snr"], plan["snr"]))


def test_mixer_feeds_the_training_step(cuda):
    """7-ch pool, 250-frame windows (BASELINE configs[3]): batches with the model's feature layout go straight into the
    native PIT training step."""
    from css_b200 import _native as nv
    from css_b200.datasets import ContinuousSpeechSeparationDataset
    from css_b200.objectives import build_objective
    from css_b200.trainer import FusedAdam
    from tests.helpers import cfg7, train_conformer

    pool, _ = _pool(C=7, n_utt=8, seed=2)
    ds = ContinuousSpeechSeparationDataset(pool, batch_size=4, seed=1, ipd_index=cfg7()["separation"]["ipd"], window_frames=250)
    batch = next(ds)
    assert batch["mix"].shape == (4, 250, 1799) and batch["source1"].shape == (4, 250, 257) and batch["mix"].is_cuda
    nv.set_gemm_mode(1)
    model = train_conformer(num_blocks=1).to(cuda)
    opt = FusedAdam(model.train_engine(), lr=1e-3)
    obj = build_objective("MSE", {})
    losses = []
    for _ in range(3):
        loss = obj(model, batch, device=cuda)
        loss.backward()
        opt.clip_grad_norm_(5.0)
        opt.step()
        opt.zero_grad()
        losses.append(float(loss.detach()))
    assert all(math.isfinite(v) for v in losses) and losses[-1] < losses[0]
