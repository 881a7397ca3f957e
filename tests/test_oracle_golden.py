"""CPU tests: the oracle (oracle/css_oracle.py) replays the reference-generated golden fixtures."""
import numpy as np
import torch

from oracle import css_oracle as O
from tests.helpers import cfg1, cfg7, cpu_state_dict, golden_blstm, golden_conformer, rel_inf


def test_case_a_features_and_masks(golden):
    wav = torch.from_numpy(golden["A_wav_i16"]).float() / 32767
    chunks = wav.unfold(-1, 38656, 12800).permute(1, 0, 2).contiguous()[:2]
    mag, f, re, im = O.feature_extract(chunks, ipd_index=cfg7()["separation"]["ipd"])
    assert rel_inf(f[:, ::7, ::3], golden["A_feat"]) < 1e-6
    assert rel_inf(re[:, :, ::5, ::7], golden["A_re"]) < 1e-6 and rel_inf(im[:, :, ::5, ::7], golden["A_im"]) < 1e-6
    sd = cpu_state_dict(golden_conformer(num_blocks=2))
    with torch.no_grad():
        masks = O.conformer_masks(f.transpose(1, 2), sd)
    for c in range(3):
        assert rel_inf(masks[c][:, ::3, ::3], golden["A_masks2"][c]) < 1e-5


def test_case_a_pipeline(golden):
    wav = torch.from_numpy(golden["A_wav_i16"]).float() / 32767
    sd = cpu_state_dict(golden_conformer(num_blocks=2))
    with torch.no_grad():
        o = O.css_pipeline(wav, cfg7(), lambda f: O.conformer_masks(f, sd))
        dbg = O.separate(wav, cfg7()["separation"], lambda f: O.conformer_masks(f, sd), return_debug=True)[2]
    assert rel_inf(o["masks"][:, ::3, ::3], golden["A_masks"]) < 1e-5
    assert o["perms"] == golden["A_perms"].tolist()
    assert [d[0] for d in dbg] == golden["A_doa_idx"].tolist() and [int(d[1]) for d in dbg] == golden["A_doa_hit"].tolist()
    st = torch.stack(o["stitched"])
    assert rel_inf(st[:, ::3, ::5], golden["A_stitched"]) < 1e-5
    # the reference's fp32 MVDR chain amplifies 1e-6 mask differences ~500x: 2e-3 is the reproducibility floor
    assert rel_inf(o["wavs"][0][0, ::16], golden["A_wav0"]) < 2e-3
    assert rel_inf(o["wavs"][1][0, ::16], golden["A_wav1"]) < 2e-3


def test_case_b_blstm(golden):
    wav = torch.from_numpy(golden["B_wav_i16"]).float() / 32767
    sd = cpu_state_dict(golden_blstm())
    with torch.no_grad():
        o = O.css_pipeline(wav, cfg1(), lambda f: O.blstm_masks(f, sd))
    assert rel_inf(o["masks"][:, ::3, ::3], golden["B_masks"]) < 1e-5
    assert o["perms"] == golden["B_perms"].tolist()
    assert rel_inf(torch.stack(o["stitched"])[:, ::3, ::5], golden["B_stitched"]) < 1e-5


def test_blstm_restatement_matches_torch_lstm():
    """The explicit recurrence in the oracle equals torch.nn.LSTM (what the reference calls)."""
    m = golden_blstm()
    sd = cpu_state_dict(m)
    x = torch.rand(2, 30, 256)
    with torch.no_grad():
        ref = m.blstm.encoders[0].lstm(x)[0]
        got = O._bilstm(x, sd, "blstm.encoders.0.lstm")
    assert rel_inf(got, ref) < 1e-5


def test_case_d_pit(golden):
    g = torch.Generator().manual_seed(int(golden["D_seed"]))
    est = torch.rand(6, 2, 50, 257, generator=g)
    tgt = torch.rand(6, 2, 50, 257, generator=g)
    tgt[1] = est[1].flip(0) + 0.01 * torch.rand(2, 50, 257, generator=g)
    tgt[2] = est[2]
    tgt[3, 1] = tgt[3, 0]; est[3, 1] = est[3, 0]
    loss, perms = O.pit_mse(est, tgt, return_permutation=True)
    assert [list(p) for p in perms] == golden["D_perm"].tolist()
    assert abs(float(loss) - float(np.mean(golden["D_loss"]))) < 1e-6


def test_placement_plan_quirks():
    """seam double-add (Q3) and the spurious en == T branch (Q4) for a 60 s file."""
    T, hop = 960000, 12800
    plan = O.placement_plan(72, T, hop, 32000, 38144, 32)
    chunks = [p[0] for p in plan]
    assert chunks.count(32) == 2 and chunks.count(64) == 2 and len(plan) == 74
    long_tail = [p[0] for p in plan if p[2] == 38144]
    assert long_tail == [42, 66]
    covered = max(p[3] + p[2] - p[1] for p in plan)
    assert T - covered >= 19200            # Q6: the tail stays zero


# ---------------------------------------------------------------------------------------------------------------
# round-2 fixtures (tests/golden/golden_r2.npz, make_golden_r2.py): the benchmarked 16-block configuration on 72 chunks
# ---------------------------------------------------------------------------------------------------------------
def _g2():
    import os

    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_r2.npz"))


def test_case_e_16_blocks_72_chunks_oracle_replay():
    """Oracle == reference on the bench.py model (16 blocks, seed 0), 60 s: masks, DOA, perms, stitched, the Q3/Q4
    placement seams sample by sample, zero tail (Q6)."""
    import hashlib

    import css_b200.models as M
    from css_b200.synth import make_recording
    from tests.test_gpu_parity_r2 import CONF16

    g = _g2()
    wav = make_recording(int(g["E_seconds"]), channels=7, seed=int(g["E_seed"]))
    assert np.array_equal(np.frombuffer(hashlib.sha256(wav.numpy().tobytes()).digest(), dtype=np.uint8), g["E_wav_sha256"])
    torch.manual_seed(0)
    sd = cpu_state_dict(M.build_model("Conformer", CONF16).eval())
    with torch.no_grad():
        o = O.css_pipeline(wav, cfg7(), lambda f: O.conformer_masks(f, sd))
    assert rel_inf(o["masks"][:, ::5, ::7], g["E_masks"]) < 1e-5
    assert o["perms"] == g["E_perms"].tolist()
    assert rel_inf(torch.stack(o["stitched"])[:, ::5, ::5], g["E_stitched"]) < 2e-4     # WTA tie flips: one floor value
    w = torch.cat(o["wavs"])
    assert int(w.abs().amax(0).nonzero().max()) == int(g["E_last_nonzero"])
    peak = float(np.abs(g["E_wavs"]).max())
    for name, c, n in (("E_seam32", 32, 12800), ("E_seam64", 64, 12800), ("E_tail42", 42, 18944), ("E_tail66", 66, 18944)):
        a = 12800 * c + 19200
        assert float((w[:, a: a + n] - torch.from_numpy(g[name])).abs().max()) / peak < 5e-3
    for k in range(2):
        assert abs(O.si_sdr(w[k], wav[0]) - float(g["E_sisdr"][k])) < 0.01
    # fp64 bracket: the fp32 chain of the reference is itself > 1e-3 away from the exact answer on this recording
    w64 = torch.cat(O.continuous_process(wav, o["stitched"], cfg7()["beamforming"], dtype=torch.float64))
    assert max(rel_inf(w[k], w64[k]) for k in range(2)) > 1e-3


def test_case_g_doa_oracle():
    from tests.helpers import doa_case

    g = _g2()
    X, mask, _ = doa_case(int(g["G_seed"]))
    sv, angles = O.steervec_7ch(257, 30)
    lo, hi = int(g["G_lo"]), int(g["G_hi"])
    like = O.doa_likelihood(X, sv, (mask > 0.5).float(), lo, hi)
    ref = torch.from_numpy(g["G_like"])
    assert torch.equal(torch.isnan(like), torch.isnan(ref)) and bool(torch.isnan(ref).any())
    assert rel_inf(torch.nan_to_num(like), torch.nan_to_num(ref)) < 1e-6
    _, idx, hit = O.angle_merge(X, mask, sv, angles, lo, hi, 16)
    assert idx.tolist() == g["G_idx"].tolist() and [int(h) for h in hit] == g["G_hit"].tolist() == [0, 1, 0, 1]
