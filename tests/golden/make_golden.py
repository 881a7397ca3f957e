"""Generate the golden fixtures by running the REFERENCE (desh2608/css at /root/reference) on CPU.

Run once in the build container (the reference is not available on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The reference raises as written on several paths; the stand-ins below are the
minimal documented patches: an ``onnxruntime`` shim whose session runs the reference's own PyTorch
model (forward up to the masks, D3), an ``asteroid.dsp.beamforming`` shim backed by the restatement in
oracle/css_oracle.py (asteroid is not installed: that seam stays "parity unpinned" and is
cross-checked against torchaudio here), ``merge: False`` for the 1-ch config (D1).
Every other line executed is the reference's.  The script also asserts that oracle/css_oracle.py
reproduces each reference output, and that css_b200's model classes initialise to the reference's
weights under the same seed.
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
REF = os.environ.get("CSS_REFERENCE", "/root/reference")

from oracle import css_oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
_MODELS = {}


def install_stubs():
    ort = types.ModuleType("onnxruntime")

    class InferenceSession:
        def __init__(self, path, sess_options=None, providers=None):
            self.fn = _MODELS[path]

        def get_inputs(self):
            return [types.SimpleNamespace(name="input")]

        def run(self, _names, feed):
            f = torch.from_numpy(feed["input"])            # (1, F, T)
            with torch.no_grad():
                masks = self.fn(f.transpose(1, 2))         # list of (1, T, 257)
            return [m[0].transpose(0, 1).numpy() for m in masks]

    ort.InferenceSession = InferenceSession
    ort.SessionOptions = lambda: types.SimpleNamespace()
    ort.get_device = lambda: "CPU"
    sys.modules["onnxruntime"] = ort
    ast = types.ModuleType("asteroid")
    dsp = types.ModuleType("asteroid.dsp")
    bf = types.ModuleType("asteroid.dsp.beamforming")
    bf.compute_scm = O.compute_scm

    class SoudenMVDRBeamformer:
        def forward(self, mix, target_scm, noise_scm, ref_mic=0, eps=1e-8):
            return O.souden_mvdr(mix, target_scm, noise_scm, ref_mic, eps)

    bf.SoudenMVDRBeamformer = SoudenMVDRBeamformer
    ast.dsp, dsp.beamforming = dsp, bf
    sys.modules.update({"asteroid": ast, "asteroid.dsp": dsp, "asteroid.dsp.beamforming": bf})
    import importlib.machinery

    for mod in (ort, ast, dsp, bf):           # importlib.util.find_spec (used by torch._dynamo) rejects a module without a spec
        mod.__spec__ = importlib.machinery.ModuleSpec(mod.__name__, None)
    sys.path.insert(0, REF)


def ref_mask_fn(model):
    """Reference forward up to the masks (css/models/conformer.py:92-101 / blstm.py:77-86), D3 patch."""
    enc = model.conformer if hasattr(model, "conformer") else model.blstm

    def fn(f):
        if f.ndim == 4:
            f = f.squeeze(0)
        x = (f - f.mean(-2, keepdim=True)) / (f.std(-2, keepdim=True) + O.EPS)
        x = enc(x, masks=None)[0] if hasattr(model, "conformer") else enc(x)
        m = torch.nn.functional.relu(model.linear(x))
        return list(torch.chunk(m, model.num_spk + model.num_noise, -1))

    return fn


def synth_int16(seconds, channels, seed):
    """Deterministic LibriCSS-shaped mixture quantised to 16-bit PCM (what a wav file would hold)."""
    from css_b200.synth import make_recording

    x = make_recording(seconds, channels=channels, seed=seed)
    return torch.round(x * 32767).to(torch.int16)


def close(a, b, tol, what):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    err = float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30)
    print(f"  oracle vs reference  {what:34s} rel-inf err {err:.3e}")
    assert err <= tol, f"{what}: {err} > {tol}"


def sub(t, *steps):
    """Strided subsample along the leading len(steps) dims -> numpy."""
    idx = tuple(slice(None, None, s) for s in steps)
    return t[idx].contiguous().numpy()


def main():
    torch.set_num_threads(8)
    install_stubs()
    import yaml
    from css.executor.beamformer import Beamformer
    from css.executor.feature import FeatureExtractor
    from css.executor.separator import Separator
    from css.executor.stitcher import Stitcher
    from css.models import build_model as ref_build
    from css.objectives.pit import pit_loss as ref_pit
    import css_b200.models as ours

    cfg7 = yaml.safe_load(open(os.path.join(REF, "conf/config_7ch.yaml")))
    cfg1 = yaml.safe_load(open(os.path.join(REF, "conf/config_1ch.yaml")))
    cfg1["separation"].update(merge=False, merge_threshold=16)   # D1
    G = {}

    # ---------------- model init parity: our parameter containers == reference under the same seed
    conf_c7 = dict(idim=1799, num_bins=257, num_spk=2, num_noise=1, conformer_attention_dim=256, conformer_attention_heads=4,
                   conformer_linear_units=1024, conformer_num_blocks=2, conformer_kernel_size=33, conformer_dropout_rate=0.1,
                   conformer_relative_pos_emb=True)
    conf_b1 = dict(idim=257, num_bins=257, num_spk=2, num_noise=1, blstm_hdim=256, blstm_num_layers=2, blstm_dropout_rate=0.1)
    for name, conf, seed in (("Conformer", conf_c7, 0), ("BLSTM", conf_b1, 1)):
        torch.manual_seed(seed)
        ref_m = ref_build(name, conf).eval()
        torch.manual_seed(seed)
        our_m = ours.build_model(name, conf).eval()
        sd_r, sd_o = ref_m.state_dict(), our_m.state_dict()
        assert list(sd_r.keys()) == list(sd_o.keys()), f"{name}: state_dict keys differ"
        assert all(torch.equal(sd_r[k], sd_o[k]) for k in sd_r), f"{name}: seeded init differs"
        print(f"  {name}: {len(sd_r)} state_dict entries identical to the reference under seed {seed}")
        _MODELS[name] = (ref_m, conf, seed)

    # ---------------- case A: 7-ch Conformer (2 blocks), 10 s, full pipeline incl. DOA merge + MVDR
    wav7_i16 = synth_int16(10, 7, seed=1234)
    wav7 = wav7_i16.float() / 32767
    ref_c, _, _ = _MODELS["Conformer"]
    # make BatchNorm running stats non-trivial so the folded-BN path is exercised
    g = torch.Generator().manual_seed(7)
    for blk in ref_c.conformer.encoders:
        blk.conv.BN.running_mean.copy_(torch.randn(256, generator=g) * 0.1)
        blk.conv.BN.running_var.copy_(torch.rand(256, generator=g) + 0.5)
    sd_c = {k: v.clone() for k, v in ref_c.state_dict().items()}
    _MODELS["c7.onnx"] = ref_mask_fn(ref_c)
    cfg7["separation"]["model_path"] = "c7.onnx"
    sep = Separator(cfg7["separation"], sess_options=None, device="cpu")
    with torch.no_grad():
        masks, mags = sep.separate(wav7)
    st = Stitcher(cfg7["stitching"])
    perms = st.get_stitch(mags, masks)
    stitched = st.get_connect(perms, masks)
    bfm = Beamformer(cfg7["beamforming"])
    out0, out1 = bfm.continuous_process(wav7, stitched)
    # oracle replay
    stages = {}
    o = O.css_pipeline(wav7, cfg7, lambda f: O.conformer_masks(f, sd_c), stages=stages)
    close(o["mags"], mags, 1e-6, "A mags")
    close(o["masks"], masks, 1e-5, "A masks")
    assert o["perms"] == perms, "A perms"
    for k in range(3):
        close(o["stitched"][k], stitched[k], 1e-6, f"A stitched[{k}]")
    close(o["wavs"][0], out0, 2e-3, "A wav0")   # MVDR amplifies 1e-6 mask differences ~500x (cond(Phi_n))
    close(o["wavs"][1], out1, 2e-3, "A wav1")
    feat_ref = FeatureExtractor(512, 256, ipd_index=cfg7["separation"]["ipd"])
    chunks = wav7.unfold(-1, 38656, 12800).permute(1, 0, 2).contiguous()
    mag_r, f_r, re_r, im_r = feat_ref.forward(chunks[:2])
    mag_o, f_o, re_o, im_o = O.feature_extract(chunks[:2], ipd_index=cfg7["separation"]["ipd"])
    close(f_o, f_r, 1e-6, "A features")
    with torch.no_grad():
        m2_ref = ref_mask_fn(ref_c)(f_r.transpose(1, 2))
        m2_or = O.conformer_masks(f_r.transpose(1, 2), sd_c)
    for a, b in zip(m2_or, m2_ref):
        close(a, b, 1e-5, "A conformer masks (batch of 2)")
    G.update(A_wav_i16=wav7_i16.numpy(), A_bn_seed=np.int64(7), A_feat=sub(f_r, 1, 7, 3), A_re=sub(re_r, 1, 1, 5, 7),
             A_im=sub(im_r, 1, 1, 5, 7), A_mag=sub(mags, 1, 3, 3), A_masks=sub(masks, 1, 3, 3, 1),
             A_perms=np.array(perms, dtype=np.int32), A_stitched=np.stack([sub(s, 3, 5) for s in stitched]),
             A_wav0=sub(out0[0], 16), A_wav1=sub(out1[0], 16), A_wav0_full_absmax=np.float32(out0.abs().max()),
             A_masks2=np.stack([sub(m, 1, 3, 3) for m in m2_ref]))
    dbg = O.separate(wav7, cfg7["separation"], lambda f: O.conformer_masks(f, sd_c), return_debug=True)[2]
    G["A_doa_idx"] = np.array([d[0] for d in dbg], dtype=np.int32)
    G["A_doa_hit"] = np.array([d[1] for d in dbg], dtype=np.int32)

    # ---------------- case B: 1-ch BLSTM (config 0 of BASELINE.json, reduced width), 30 s
    wav1_i16 = synth_int16(30, 1, seed=4321)
    wav1 = wav1_i16.float() / 32767
    ref_b, _, _ = _MODELS["BLSTM"]
    sd_b = {k: v.clone() for k, v in ref_b.state_dict().items()}
    _MODELS["b1.onnx"] = lambda f: list(ref_b(f)[1])          # unmodified reference forward (idim == num_bins)
    cfg1["separation"]["model_path"] = "b1.onnx"
    sep1 = Separator(cfg1["separation"], sess_options=None, device="cpu")
    with torch.no_grad():
        masks1, mags1 = sep1.separate(wav1)
    st1 = Stitcher(cfg1["stitching"])
    perms1 = st1.get_stitch(mags1, masks1)
    stitched1 = st1.get_connect(perms1, masks1)
    o1 = O.css_pipeline(wav1, cfg1, lambda f: O.blstm_masks(f, sd_b))
    close(o1["masks"], masks1, 1e-5, "B masks")
    assert o1["perms"] == perms1, "B perms"
    for k in range(3):
        close(o1["stitched"][k], stitched1[k], 1e-5, f"B stitched[{k}]")
    G.update(B_wav_i16=wav1_i16.numpy(), B_masks=sub(masks1, 1, 3, 3, 1), B_mags=sub(mags1, 1, 3, 3),
             B_perms=np.array(perms1, dtype=np.int32), B_stitched=np.stack([sub(s, 3, 5) for s in stitched1]))

    # ---------------- case C: stitcher stress -- constructed masks with known random swaps + noise
    g = torch.Generator().manual_seed(99)
    nC = 40
    L = 50 * (nC - 1) + 150
    base = torch.rand(2, 257, L, generator=g)
    magC = torch.rand(257, L, generator=g) + 0.1
    flips = torch.randint(0, 2, (nC,), generator=g)
    mC = torch.empty(nC, 257, 150, 3)
    for n in range(nC):
        seg = base[:, :, 50 * n : 50 * n + 150] + 0.05 * torch.rand(2, 257, 150, generator=g)
        if flips[n]:
            seg = seg.flip(0)
        mC[n, :, :, 0], mC[n, :, :, 1] = seg[0], seg[1]
        mC[n, :, :, 2] = 0.3 * torch.rand(257, 150, generator=g)
    magsC = magC.unfold(1, 150, 50).permute(1, 0, 2).contiguous()
    permsC = st.get_stitch(magsC, mC)
    assert O.get_stitch(magsC, mC) == permsC
    expect = [[0, 1] if flips[n] == flips[n + 1] else [1, 0] for n in range(nC - 1)]
    assert permsC == expect, "constructed swaps not recovered by the reference"
    conC = st.get_connect(permsC, mC)
    for k in range(3):
        close(O.get_connect(permsC, mC)[k], conC[k], 1e-6, f"C connect[{k}]")
    G.update(C_seed=np.int64(99), C_perms=np.array(permsC, dtype=np.int32), C_connect=np.stack([sub(s, 3, 5) for s in conC]))

    # ---------------- case D: PIT loss
    g = torch.Generator().manual_seed(5)
    est = torch.rand(6, 2, 50, 257, generator=g)
    tgt = torch.rand(6, 2, 50, 257, generator=g)
    tgt[1] = est[1].flip(0) + 0.01 * torch.rand(2, 50, 257, generator=g)    # swapped is best
    tgt[2] = est[2]                                                         # exact match
    tgt[3, 1] = tgt[3, 0]; est[3, 1] = est[3, 0]                            # tie -> first permutation
    losses, pp = [], []
    for b in range(6):
        l, p = ref_pit(est[b], tgt[b], axis=0, return_permutation=True)
        lo, po = O.pit_loss(est[b], tgt[b], axis=0, return_permutation=True)
        assert p == po and abs(float(l) - float(lo)) <= 1e-7 * max(1.0, abs(float(l)))
        losses.append(float(l)); pp.append(p)
    est3 = torch.rand(2, 3, 20, 257, generator=g); tgt3 = est3[:, [2, 0, 1]] + 0.01 * torch.rand(2, 3, 20, 257, generator=g)
    l3 = [ref_pit(est3[b], tgt3[b], axis=0, return_permutation=True) for b in range(2)]
    G.update(D_seed=np.int64(5), D_loss=np.array(losses, dtype=np.float32), D_perm=np.array(pp, dtype=np.int32),
             D_loss3=np.array([float(l[0]) for l in l3], dtype=np.float32), D_perm3=np.array([l[1] for l in l3], dtype=np.int32))

    # ---------------- asteroid seam: restatement vs torchaudio (independent implementation)
    import torchaudio.functional as AF
    g = torch.Generator().manual_seed(3)
    X = torch.randn(2, 7, 33, 60, generator=g) + 1j * torch.randn(2, 7, 33, 60, generator=g)
    ms, mn = torch.rand(2, 33, 60, generator=g), torch.rand(2, 33, 60, generator=g)
    ours_bf = O.souden_mvdr(X, O.compute_scm(X, ms.unsqueeze(1)), O.compute_scm(X, mn.unsqueeze(1)))
    w = AF.mvdr_weights_souden(AF.psd(X, ms, normalize=True, eps=0.0), AF.psd(X, mn, normalize=True, eps=0.0), 0,
                               diagonal_loading=False, eps=1e-8)
    close(ours_bf, AF.apply_beamforming(w, X), 1e-4, "asteroid restatement vs torchaudio")

    np.savez_compressed(os.path.join(OUT, "golden.npz"), **G)
    print("wrote", os.path.join(OUT, "golden.npz"), os.path.getsize(os.path.join(OUT, "golden.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
