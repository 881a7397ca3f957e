"""checkpoint averaging tool (scripts/python/combine_models.py:17-154).  Host-side file work: CPU test."""
import json
import os

import torch

from tests.helpers import CONF_C7


def _conf(path):
    conf = dict(CONF_C7, conformer_num_blocks=1, idim=257, conformer_linear_units=64, model="Conformer", objective="MSE", optim="adam",
                lr=1e-4, weight_decay=1e-2)
    with open(path, "w") as f:
        json.dump(conf, f)
    return conf


def _checkpoints(tmp, conf, n, fused):
    import css_b200.models as M

    paths, sds = [], []
    for i in range(n):
        torch.manual_seed(100 + i)
        m = M.build_model("Conformer", conf)
        for blk in m.conformer.encoders:
            blk.conv.BN.num_batches_tracked.fill_(10 * (i + 1))
        params = list(m.parameters())
        if fused:
            nflat = sum(p.numel() for p in params)
            opt = {"steps": 7 + i, "m": torch.full((nflat,), float(i + 1)), "v": torch.full((nflat,), 2.0 * (i + 1)),
                   "param_groups": [{"lr": 1e-4, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 1e-2}]}
        else:
            o = torch.optim.Adam(params, lr=1e-4, weight_decay=1e-2)
            for p in params:
                p.grad = torch.ones_like(p) * (i + 1)
            o.step()
            opt = o.state_dict()
        p = os.path.join(tmp, f"m{i}.mdl")
        torch.save({"model": m.state_dict(), "objective": {}, "optimizer": opt, "lr_sched": {"curr_lr": 1e-4 * (i + 1)}, "epoch": i}, p)
        paths.append(p)
        sds.append({k: v.clone() for k, v in m.state_dict().items()})
    return paths, sds


def test_average_model_optimizer_and_cli(tmp_path):
    from css_b200.tools import combine_models as C

    conf = _conf(tmp_path / "conf.json")
    for fused in (False, True):
        paths, sds = _checkpoints(str(tmp_path), conf, 3, fused)
        out_path = str(tmp_path / f"avg{int(fused)}.mdl")
        C.main([out_path, str(tmp_path / "conf.json"), "--save-models", "--models", *paths, str(tmp_path / "missing.mdl")])
        out = torch.load(out_path, weights_only=False)
        for k, v in out["model"].items():
            if v.dim() == 0:        # reference rule: last model's value * fraction, cast back (combine_models.py:77-78)
                assert int(v) == int(sds[-1][k] * (1.0 / 3))
            else:
                want = sum(sd[k].double() for sd in sds) / 3
                assert torch.allclose(v.double(), want, atol=1e-6), k
        assert out["epoch"] == 2 and abs(out["lr_sched"]["curr_lr"] - 3e-4) < 1e-12
        assert all(os.path.exists(p) for p in paths)                           # --save-models keeps the inputs
    # FusedAdam moments accumulate with weight 1/n
    paths, _ = _checkpoints(str(tmp_path), conf, 2, True)
    state = {"m": None, "v": None}
    for p in paths:
        C.update_opt_state_dict(state, torch.load(p, weights_only=False)["optimizer"], 0.5)
    assert float(state["m"][0]) == 1.5 and float(state["v"][0]) == 3.0 and state["steps"] == 8
    # without --save-models the inputs are removed, as in the reference
    C.main([str(tmp_path / "o.mdl"), str(tmp_path / "conf.json"), "--models", *paths])
    assert not any(os.path.exists(p) for p in paths)
