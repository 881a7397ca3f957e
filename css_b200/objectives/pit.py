"""Permutation-invariant loss (reference: css/objectives/pit.py:5-62), B200-native for MSE.

``pit_loss`` keeps the reference signature.  For ``loss_fn=torch.nn.functional.mse_loss`` (the only
loss the reference trains with, css/objectives/mse.py:33-35) the pairwise error matrix, the
warp-parallel permutation search and the gradient run in ``css_pit_mse``; any other loss_fn is
rejected (no PyTorch fallback on the hot path).
"""
import torch

from .. import _native as nv


class _PitMse(torch.autograd.Function):
    """(B,K,n) estimates/targets -> batch-mean PIT-MSE; gradient only through the chosen permutation."""

    @staticmethod
    def forward(ctx, est, tgt):
        B, K, n = est.shape
        dev = est.device
        pair = torch.empty(B, K, K, dtype=torch.float64, device=dev)
        loss_b = torch.empty(B, dtype=torch.float32, device=dev)
        perm = torch.empty(B, K, dtype=torch.int32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        grad = torch.empty_like(est) if ctx.needs_input_grad[0] else None
        nv.call("css_pit_mse", nv.ptr(est), nv.ptr(tgt), nv.c_long(B), nv.c_int(K), nv.c_long(n), nv.ptr(pair),
                nv.ptr(loss_b), nv.ptr(perm), nv.ptr(loss), nv.ptr(grad), nv.c_float(1.0))
        ctx.grad = grad
        ctx.mark_non_differentiable(perm, loss_b)
        return loss.reshape(()), perm, loss_b

    @staticmethod
    def backward(ctx, g, _gp, _gl):
        return (ctx.grad * g if ctx.grad is not None else None), None


def pit_mse_batch(y_pred, y_true):
    """(B,K,T,F) x2 -> (mean loss, perms (B,K) int32, per-utterance losses)."""
    nv.require_cuda(y_pred, "y_pred")
    assert y_pred.shape == y_true.shape, f"{y_pred.shape} != {y_true.shape}"
    B, K = y_pred.shape[:2]
    est = y_pred.reshape(B, K, -1).float().contiguous()
    tgt = y_true.reshape(B, K, -1).float().contiguous()
    return _PitMse.apply(est, tgt)


def pit_loss(estimate, target, axis, loss_fn=torch.nn.functional.mse_loss, return_permutation=False, **kwargs):
    """Reference signature (css/objectives/pit.py:5-12).  Speaker axis ``axis``; no batch dimension."""
    sources = estimate.size()[axis]
    assert sources < 30, f"Are you sure? sources={sources}, estimate.shape={estimate.shape}, target.shape={target.shape}"
    assert estimate.size() == target.size(), f"{estimate.size()} != {target.size()}"
    if loss_fn is not torch.nn.functional.mse_loss or kwargs:
        raise NotImplementedError("css_b200.pit_loss implements the MSE criterion the reference trains with")
    est = estimate.movedim(axis, 0).unsqueeze(0)
    tgt = target.movedim(axis, 0).unsqueeze(0)
    loss, perm, _ = pit_mse_batch(est, tgt)
    if return_permutation:
        return loss, tuple(int(v) for v in perm[0].tolist())
    return loss
