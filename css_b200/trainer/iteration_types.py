"""One epoch of PIT training (reference: css/trainer/iteration_types.py:17-77), on the native step.

Same signature and control flow as the reference; the differences are under the hood:
forward/backward are the native Conformer train kernels (bf16 tensor-core GEMMs when ``args.fp16`` -- the reference's
AMP switch -- selects GEMM mode 2, split-bf16 fp32-class GEMMs otherwise), the gradient norm / clip / Adam update are the
fused flat-buffer kernels of ``FusedAdam``, and with ``torch.distributed`` initialised ``engine.enable_data_parallel``
is switched on (parameters broadcast from rank 0, gradients all-reduced over NCCL while the backward is still running,
the mean folded into the optimiser).  Loss and gradient norm are logged without stalling the device every
iteration: they are read back once per epoch.
"""
import logging

import torch

from .. import _native as nv


def train_one_epoch(args, generator, model, objective, optim, lr_sched, device="cuda"):
    engine = model.train_engine()
    if getattr(engine, "exchange", None) is None and torch.distributed.is_available() and torch.distributed.is_initialized():
        engine.enable_data_parallel(optim)        # broadcast from rank 0, bucketed all-reduce, grad_scale = 1/world
    exch = getattr(engine, "exchange", None)
    # args.fp16 (the reference's AMP switch) selects single-plane bf16 GEMMs for the TRAINING forward/backward only: the
    # reference scopes autocast to the training forward and validates in fp32, so the mode is restored on the way out
    saved_mode = nv.get_gemm_mode()
    losses, norms = [], []
    try:
        if getattr(args, "gpu", True) and getattr(args, "fp16", False):
            nv.set_gemm_mode(2)
        for i in range(1, args.batches_per_epoch + 1):
            b = next(generator)
            loss = objective(model, b, device=device)
            loss.backward()
            if exch is not None:
                exch.finish()
            norms.append(optim.clip_grad_norm_(args.grad_thresh))
            optim.step()
            optim.zero_grad()
            lr_sched.step(1.0)
            losses.append(loss.detach())
            del b
    finally:
        nv.set_gemm_mode(saved_mode)
    losses = torch.stack(losses).cpu()
    norms = torch.stack(norms).cpu()
    for i, (l, n) in enumerate(zip(losses.tolist(), norms.tolist()), 1):
        logging.info(f"Iter: {i} of {args.batches_per_epoch} Loss: {l:0.5f} Grad_norm: {n:0.5f}")
    return float(losses.sum()) / args.batches_per_epoch


def validate(generator, model, objective, device="cuda", batches=100):
    model.eval()
    with torch.no_grad():
        avg_loss = 0.0
        for _ in range(batches):
            b = next(generator)
            avg_loss += objective(model, b, device=device)
        avg_loss /= batches
    model.train()
    return avg_loss
