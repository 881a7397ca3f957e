"""Fused optimiser step and data-parallel gradient exchange over the flat buffers of ``ConformerTrainEngine``.

Replaces, for the training hot path, ``torch.nn.utils.clip_grad_norm_`` + ``torch.optim.Adam(lr, weight_decay)``
(reference: css/trainer/iteration_types.py:54-61, scripts/python/train.py:209-211) by ``css_grad_sqnorm`` +
``css_adam_step``, and the reference's file-based model averaging (scripts/python/combine_models.py) by an NCCL
all-reduce of the gradients, issued per bucket of layers while the backward of earlier layers is still running.
"""
import torch
import torch.distributed as dist

from .. import _native as nv


class FusedAdam:
    """torch.optim.Adam-compatible surface (param_groups / state_dict / step / zero_grad) over one flat buffer."""

    def __init__(self, engine, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.engine = engine
        self.param_groups = [{"lr": lr, "betas": tuple(betas), "eps": eps, "weight_decay": weight_decay,
                              "params": [p for p in engine.model.parameters()]}]
        self.m = torch.zeros_like(engine.flat_p)
        self.v = torch.zeros_like(engine.flat_p)
        self.steps = 0
        self.sqnorm = torch.zeros(1, dtype=torch.float64, device=engine.device)
        self.norm = torch.zeros(1, dtype=torch.float32, device=engine.device)
        self.max_norm = 0.0      # set by clip_grad_norm_()
        self.grad_scale = 1.0    # 1/world after a sum all-reduce
        self._have_norm = False

    def clip_grad_norm_(self, max_norm):
        """Records the clip threshold and computes the global gradient norm on the device (no host sync).  The clip
        itself is applied inside step().  Returns a 0-d CUDA tensor like torch.nn.utils.clip_grad_norm_."""
        e = self.engine
        nv.call("css_grad_sqnorm", nv.ptr(e.flat_g), nv.c_long(e.size), nv.ptr(self.sqnorm))
        self.max_norm = float(max_norm)
        self._have_norm = True
        return (self.sqnorm.sqrt() * self.grad_scale).float().reshape(())

    def step(self):
        e = self.engine
        g = self.param_groups[0]
        self.steps += 1
        nv.call("css_adam_step", nv.ptr(e.flat_p), nv.ptr(e.flat_g), nv.ptr(self.m), nv.ptr(self.v), nv.c_long(e.size),
                nv.c_float(g["lr"]), nv.c_float(g["betas"][0]), nv.c_float(g["betas"][1]), nv.c_float(g["eps"]),
                nv.c_float(g["weight_decay"]), nv.c_long(self.steps), nv.c_float(self.grad_scale),
                nv.ptr(self.sqnorm if self._have_norm else None), nv.c_float(self.max_norm if self._have_norm else 0.0),
                nv.ptr(self.norm))
        self._have_norm = False
        e.model.invalidate()     # packed inference weights are stale now

    def zero_grad(self, set_to_none=False):
        self.engine.zero_grad()

    @property
    def state(self):
        """torch.optim-style ``optimizer.state`` (scripts/python/train.py:219-222 walks it after a resume to move the
        moment tensors to the GPU): one entry for the flat buffer."""
        return {"flat": {"step": self.steps, "exp_avg": self.m, "exp_avg_sq": self.v}}

    def state_dict(self):
        return {"steps": self.steps, "m": self.m, "v": self.v,
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}

    def load_state_dict(self, sd):
        self.steps = int(sd["steps"])
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.param_groups[0].update(sd["param_groups"][0])


def bucket_ranges(num_blocks, buckets):
    """Backward parts run num_blocks (mask head), num_blocks-1, ..., 0, -1 (embedding); split that sequence into at most
    `buckets` contiguous (part_hi, part_lo) ranges of near-equal length."""
    parts = list(range(num_blocks, -2, -1))
    n = max(1, min(buckets, len(parts)))
    per = (len(parts) + n - 1) // n
    return [(parts[i], parts[min(i + per, len(parts)) - 1]) for i in range(0, len(parts), per)]


def slice_bounds(part_bounds, size, part_hi, part_lo):
    """[start, end) of the flat gradient that is final once parts part_hi..part_lo have run.  part_bounds[l] is the
    offset of block l, part_bounds[L] that of the head; the embedding and the relative-position table sit in front of
    block 0 and are only final after part -1."""
    L = len(part_bounds) - 1
    end = size if part_hi >= L else part_bounds[part_hi + 1]
    start = 0 if part_lo < 0 else part_bounds[part_lo]
    return start, end


class GradientExchange:
    """Bucketed sum all-reduce of the flat gradient, overlapped with the backward (one comm stream).

    ``buckets`` layer groups are reduced as soon as their backward kernels have been enqueued: the comm stream waits on
    an event recorded after the group's last kernel, so NCCL runs over NVLink while the compute stream continues with
    the earlier layers.  The mean is taken inside the optimiser kernel (``grad_scale = 1/world``)."""

    def __init__(self, engine, buckets=4, group=None):
        self.engine = engine
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.ranges = bucket_ranges(engine.num_blocks, buckets)
        self.on_gpu = engine.flat_g.is_cuda
        self.comm_stream = torch.cuda.Stream(device=engine.device) if self.world > 1 and self.on_gpu else None
        self.pending = []

    def after_parts(self, hi, lo):
        if self.world == 1:
            return
        if not self.on_gpu:     # host-logic tests (gloo): same slices, no streams
            dist.all_reduce(self.engine.grad_slice(hi, lo), op=dist.ReduceOp.SUM, group=self.group)
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        sl = self.engine.grad_slice(hi, lo)
        with torch.cuda.stream(self.comm_stream):
            self.comm_stream.wait_event(ev)
            dist.all_reduce(sl, op=dist.ReduceOp.SUM, group=self.group)
        done = torch.cuda.Event()
        done.record(self.comm_stream)
        self.pending.append(done)

    def finish(self):
        for ev in self.pending:
            torch.cuda.current_stream().wait_event(ev)
        self.pending = []
