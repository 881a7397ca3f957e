#!/usr/bin/env python
"""Checkpoint averaging (reference: scripts/python/combine_models.py:17-154).

    python -m css_b200.tools.combine_models OMODEL CONF.json --models m1.mdl m2.mdl ... [--save-models]

Averages the ``model`` tensors of the listed ``.mdl`` checkpoints (scripts/python/train.py:304-317: keys model, objective,
optimizer, lr_sched, epoch), accumulates the optimizer moments and the objective state with weight 1/len(models), and keeps
the last checkpoint's lr_sched / epoch.  Missing files are skipped, as in the reference (only jobs that finished write a
model).  0-d entries (BatchNorm ``num_batches_tracked``) follow the reference's rule: ``value * fraction`` of the LAST model,
cast back to the entry's dtype (combine_models.py:74-78).

In css_b200 the parallel-training use of this tool (train N jobs, average, repeat) is replaced by the NCCL gradient
all-reduce of ``GradientExchange``; the tool remains for interchange with reference-trained checkpoints.  Host-side file
work: plain torch CPU tensors, no kernels.
"""
import argparse
import json
import os
import sys
from itertools import chain

import torch


def update_opt_state_dict(state_dict1, state_dict2, fraction):
    """state_dict1 += fraction * state_dict2 for a torch.optim state_dict (combine_models.py:113-154) or a FusedAdam one."""
    if "m" in state_dict2 and "v" in state_dict2:                   # css_b200.trainer.optim.FusedAdam: flat moments
        for k in ("m", "v"):
            inc = fraction * state_dict2[k].detach().cpu()
            state_dict1[k] = inc if state_dict1.get(k) is None else state_dict1[k] + inc
        state_dict1["steps"] = state_dict2["steps"]
        state_dict1["param_groups"] = state_dict2["param_groups"]
        return
    g1, g2 = state_dict1["param_groups"], state_dict2["param_groups"]
    if len(g1) != len(g2):
        raise ValueError("state dict as a different number of parameter groups")
    if any(len(a["params"]) != len(b["params"]) for a, b in zip(g1, g2)):
        raise ValueError("loaded state dict contains a parameter group that doesn't match the size of the optimizer's group")
    id_map = dict(zip(chain(*(g["params"] for g in g2)), chain(*(g["params"] for g in g1))))
    for k, v in state_dict2["state"].items():
        dst = id_map.get(k, k)
        if dst in state_dict1["state"]:
            for name, p in v.items():
                if isinstance(p, torch.Tensor):
                    state_dict1["state"][dst][name] += fraction * p
        else:
            state_dict1["state"][dst] = {name: fraction * p for name, p in v.items()}


def average_checkpoints(paths, new_model, objective, optimizer_state):
    """-> averaged checkpoint dict.  ``new_model`` supplies names / dtypes / shapes (its values are discarded)."""
    model_list = [m for m in paths if os.path.exists(m)]
    if not model_list:
        raise FileNotFoundError("No remaining models")
    fraction = 1.0 / len(model_list)
    new_mdl = {k: (v.detach().cpu().clone().mul_(0.0) if v.dim() > 0 else v.detach().cpu().clone())
               for k, v in new_model.state_dict().items()}
    new_obj = objective.state_dict() if objective is not None else {}
    new_opt = optimizer_state
    state = None
    for i, path in enumerate(model_list):
        print("Combining Model ", i, " : ", path, " ...")
        state = torch.load(path, map_location=torch.device("cpu"), weights_only=False)
        for name, p in state["model"].items():
            if name in new_mdl:
                if p.dim() != 0:
                    new_mdl[name].add_(p.to(new_mdl[name].dtype), alpha=fraction)
                else:
                    new_mdl[name] = (p * fraction).type(new_mdl[name].dtype)
        if "optimizer" in state and new_opt is not None:
            update_opt_state_dict(new_opt, state["optimizer"], fraction)
        if objective is not None and "objective" in state:
            new_obj = objective.add_state_dict(new_obj, state["objective"], fraction, iteration=i)
    out = {"model": new_mdl, "objective": new_obj, "optimizer": state.get("optimizer"), "lr_sched": state.get("lr_sched"),
           "epoch": state.get("epoch")}
    return out, model_list


def main(argv=None):
    import css_b200.models as models
    import css_b200.objectives as objectives

    parser = argparse.ArgumentParser()
    parser.add_argument("omodel", help="path to output model", type=str)
    parser.add_argument("conf", type=str)
    parser.add_argument("--save-models", action="store_true")
    parser.add_argument("--models", nargs="+", type=str, help="paths to models")
    args = parser.parse_args(argv)
    with open(args.conf) as f:
        conf = json.load(f)
    new_model = models.MODELS[conf["model"]].build_model(conf)
    objective = objectives.OBJECTIVES[conf["objective"]].build_objective(conf)
    params = [p for p in chain(new_model.parameters(), objective.parameters()) if p.requires_grad]
    optimizers = {
        "sgd": lambda: torch.optim.SGD(params, lr=conf["lr"], momentum=0.0),
        "adadelta": lambda: torch.optim.Adadelta(params, lr=conf["lr"]),
        "adam": lambda: torch.optim.Adam(params, lr=conf["lr"], weight_decay=conf["weight_decay"]),
    }
    new_opt = optimizers[conf["optim"]]().state_dict()
    try:
        out, used = average_checkpoints(args.models or [], new_model, objective, new_opt)
    except FileNotFoundError as e:
        print(e, file=sys.stderr)
        sys.exit(1)
    torch.save(out, args.omodel)
    if not args.save_models:
        for m in used:
            os.remove(m)


if __name__ == "__main__":
    main()
