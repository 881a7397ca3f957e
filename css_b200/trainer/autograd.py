"""torch.autograd bridge of the native training path: ``model(f)`` in train mode returns a ``y_pred`` whose
``backward`` runs ``css_trainer_backward_part`` and leaves the parameter gradients in the engine's flat buffer
(``p.grad`` of every Parameter is a view into it), exactly where ``loss.backward()`` would have put them
(css/trainer/iteration_types.py:49-52)."""
import torch


class ConformerTrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, anchor, engine, want_masks):
        y_pred, head = engine.forward(feat, want_masks)
        ctx.engine = engine
        ctx.feat = feat
        if head is None:
            head = feat.new_zeros(())
        ctx.mark_non_differentiable(head)
        return y_pred, head

    @staticmethod
    def backward(ctx, grad_y, _grad_head):
        engine = ctx.engine
        grad_y = grad_y.contiguous().float()
        exch = getattr(engine, "exchange", None)
        if exch is None:
            engine.backward_part(ctx.feat, grad_y, engine.num_blocks, -1)
        else:
            for hi, lo in exch.ranges:
                engine.backward_part(ctx.feat, grad_y, hi, lo)
                exch.after_parts(hi, lo)
        return None, None, None, None


def train_forward(model, f):
    """(N,T,F) -> (y_pred, masks) through the native train-mode forward."""
    engine = model.train_engine()
    feat = f.contiguous().float()
    y_pred, head = ConformerTrainFn.apply(feat, engine.anchor, engine, True)
    N, T, _ = feat.shape
    nb = model.num_bins
    masks = tuple(head[:, c * nb : (c + 1) * nb].reshape(N, T, nb) for c in range(model.num_spk + model.num_noise))
    return y_pred, masks
