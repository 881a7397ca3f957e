"""Native training engine of the Conformer mask estimator.

The reference trains with plain PyTorch autograd (css/trainer/iteration_types.py:34-63).  Here all parameters of a
``css_b200.models.Conformer`` are re-homed into ONE flat fp32 buffer (and their ``.grad`` into a second one of the same
layout), so that

* the native forward/backward (``css_trainer_forward`` / ``css_trainer_backward_part`` in libcss_b200.so) address every
  weight by offset,
* the data-parallel gradient exchange is a handful of large NCCL all-reduces over contiguous slices that overlap
  the backward of the earlier layers,
* gradient clipping and Adam are one fused kernel each over the flat buffers.

The torch ``Parameter`` objects stay valid views, so ``state_dict()`` / ``.mdl`` interchange is unchanged.
"""
import ctypes

import torch

from .. import _native as nv

ALIGN = 64  # floats: every tensor starts on a 256-byte boundary


class FfnOffsets(ctypes.Structure):
    _fields_ = [(n, ctypes.c_long) for n in ("ln_g", "ln_b", "w1", "b1", "w2", "b2")]


class BlockOffsets(ctypes.Structure):
    _fields_ = (
        [("ff_in", FfnOffsets)]
        + [(n, ctypes.c_long) for n in ("att_ln_g", "att_ln_b", "wqkv", "bqkv", "wo", "bo")]
        + [(n, ctypes.c_long) for n in ("conv_ln_g", "conv_ln_b", "pw1_w", "pw1_b", "dw_w", "dw_b", "bn_g", "bn_b", "pw2_w", "pw2_b")]
        + [("ff_out", FfnOffsets)]
        + [(n, ctypes.c_long) for n in ("out_ln_g", "out_ln_b")]
    )


class TrainDesc(ctypes.Structure):
    _fields_ = (
        [(n, ctypes.c_int) for n in ("idim", "d_model", "heads", "d_ffn", "num_blocks", "num_bins", "num_out", "num_spk", "pe_rows", "ksize")]
        + [(n, ctypes.c_long) for n in ("embed_w", "embed_b", "embed_ln_g", "embed_ln_b", "pe", "head_w", "head_b")]
    )


def _bind():
    lib = nv.lib()
    lib.css_trainer_create.restype = ctypes.c_void_p
    lib.css_trainer_workspace_bytes.restype = ctypes.c_size_t
    return lib


class _Layout:
    """Assigns flat offsets in forward order: [embed | pos_emb | block 0 .. L-1 | head]."""

    def __init__(self):
        self.items = []   # (parameter, offset)
        self.size = 0

    def add(self, p):
        off = self.size
        self.items.append((p, off))
        self.size = (off + p.numel() + ALIGN - 1) // ALIGN * ALIGN
        return off


class ConformerTrainEngine:
    def __init__(self, model):
        dev = model.linear.weight.device
        if dev.type != "cuda":
            raise nv.NativeError("training runs only on a CUDA device (model.cuda()): no CPU fallback")
        self.lib = _bind()
        self.model = model
        self.device = dev
        enc = model.conformer
        conf = model.conf
        lay = _Layout()
        desc = TrainDesc()
        desc.idim, desc.d_model, desc.heads = model.in_features, conf["attention_dim"], conf["attention_heads"]
        desc.d_ffn, desc.num_blocks = conf["linear_units"], conf["num_blocks"]
        desc.num_bins, desc.num_out, desc.num_spk = model.num_bins, model.num_spk + model.num_noise, model.num_spk
        desc.ksize = conf["kernel_size"]
        desc.embed_w, desc.embed_b = lay.add(enc.embed[0].weight), lay.add(enc.embed[0].bias)
        desc.embed_ln_g, desc.embed_ln_b = lay.add(enc.embed[1].weight), lay.add(enc.embed[1].bias)
        if enc.pos_emb is not None:
            desc.pe, desc.pe_rows = lay.add(enc.pos_emb.pe_k.weight), enc.pos_emb.pe_k.weight.shape[0]
        else:
            desc.pe, desc.pe_rows = -1, 0
        self.part_bounds = [lay.size]     # flat offset where each block starts (then the head)
        blocks = (BlockOffsets * len(enc.encoders))()
        for i, blk in enumerate(enc.encoders):
            b = blocks[i]
            for dst, src in ((b.ff_in, blk.feed_forward_in),):
                self._ffn(lay, dst, src)
            at = blk.self_attn
            b.att_ln_g, b.att_ln_b = lay.add(at.layer_norm.weight), lay.add(at.layer_norm.bias)
            b.wqkv = self._adjacent(lay, [at.linear_q.weight, at.linear_k.weight, at.linear_v.weight])
            b.bqkv = self._adjacent(lay, [at.linear_q.bias, at.linear_k.bias, at.linear_v.bias])
            b.wo, b.bo = lay.add(at.linear_out.weight), lay.add(at.linear_out.bias)
            cv = blk.conv
            b.conv_ln_g, b.conv_ln_b = lay.add(cv.layer_norm.weight), lay.add(cv.layer_norm.bias)
            b.pw1_w, b.pw1_b = lay.add(cv.pw_conv_1.weight), lay.add(cv.pw_conv_1.bias)
            b.dw_w, b.dw_b = lay.add(cv.dw_conv_1d.weight), lay.add(cv.dw_conv_1d.bias)
            b.bn_g, b.bn_b = lay.add(cv.BN.weight), lay.add(cv.BN.bias)
            b.pw2_w, b.pw2_b = lay.add(cv.pw_conv_2.weight), lay.add(cv.pw_conv_2.bias)
            self._ffn(lay, b.ff_out, blk.feed_forward_out)
            b.out_ln_g, b.out_ln_b = lay.add(blk.layer_norm.weight), lay.add(blk.layer_norm.bias)
            self.part_bounds.append(lay.size)
        desc.head_w, desc.head_b = lay.add(model.linear.weight), lay.add(model.linear.bias)
        covered = {id(p) for p, _ in lay.items}
        missing = [n for n, p in model.named_parameters() if id(p) not in covered]
        if missing:
            raise nv.NativeError(f"parameters without a slot in the flat layout: {missing}")
        self.size = lay.size
        self.flat_p = torch.zeros(lay.size, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(lay.size, dtype=torch.float32, device=dev)
        for p, off in lay.items:
            n = p.numel()
            self.flat_p[off : off + n].copy_(p.detach().reshape(-1).float())
            p.data = self.flat_p[off : off + n].view(p.shape)
            p.grad = self.flat_g[off : off + n].view(p.shape)
        d = desc.d_model
        self.bn_running = torch.zeros(max(1, len(enc.encoders)), 2, d, dtype=torch.float32, device=dev)
        for i, blk in enumerate(enc.encoders):
            self.bn_running[i, 0].copy_(blk.conv.BN.running_mean)
            self.bn_running[i, 1].copy_(blk.conv.BN.running_var)
            blk.conv.BN.running_mean.data = self.bn_running[i, 0]
            blk.conv.BN.running_var.data = self.bn_running[i, 1]
        # BatchNorm1d.num_batches_tracked (a state_dict entry of the reference model): one vector, advanced once per
        # train-mode forward
        self.bn_tracked = torch.zeros(max(1, len(enc.encoders)), dtype=torch.long, device=dev)
        for i, blk in enumerate(enc.encoders):
            self.bn_tracked[i] = blk.conv.BN.num_batches_tracked.to(dev)
            blk.conv.BN.num_batches_tracked.data = self.bn_tracked[i]
        self.desc, self.blocks = desc, blocks
        self.handle = self.lib.css_trainer_create(ctypes.byref(desc), blocks, nv.ptr(self.flat_p), nv.ptr(self.flat_g),
                                                  nv.ptr(self.bn_running))
        if not self.handle:
            raise nv.NativeError(self.lib.css_last_error().decode())
        self.ws = None
        self.ws_shape = None
        self.num_blocks = len(enc.encoders)
        # nn.Dropout at the reference's six sites (css/models/conformer.py:127,249,254,302,317,319): counter-based masks,
        # a fresh seed per forward (dropout_seed + number of forwards so far); the backward re-derives the same masks
        self.dropout_p = float(conf.get("dropout_rate", 0.0))
        self.dropout_seed = int(torch.initial_seed()) & 0x7FFFFFFF
        self.forward_calls = 0
        self.last_dropout_seed = None
        self.anchor = torch.zeros(1, device=dev, requires_grad=True)   # makes the native forward's output require grad
        self.exchange = None   # GradientExchange for data-parallel runs
        model.invalidate()     # packed inference weights referenced the old parameter storage

    @staticmethod
    def _ffn(lay, dst, src):
        dst.ln_g, dst.ln_b = lay.add(src.layer_norm.weight), lay.add(src.layer_norm.bias)
        dst.w1, dst.b1 = lay.add(src.net[0].weight), lay.add(src.net[0].bias)
        dst.w2, dst.b2 = lay.add(src.net[3].weight), lay.add(src.net[3].bias)

    @staticmethod
    def _adjacent(lay, params):
        """Place tensors back to back (their sizes are multiples of ALIGN) and return the first offset."""
        first = None
        for p in params:
            if p.numel() % ALIGN:
                raise nv.NativeError("q/k/v tensors must be multiples of 64 elements to be stored adjacently")
            off = lay.add(p)
            if first is None:
                first = off
            elif off != prev_end:
                raise nv.NativeError("flat layout lost q/k/v adjacency")
            prev_end = off + p.numel()
        return first

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.css_trainer_destroy(ctypes.c_void_p(self.handle))
                self.handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------ native calls
    def workspace(self, N, T):
        need = self.lib.css_trainer_workspace_bytes(ctypes.c_void_p(self.handle), nv.c_long(N), nv.c_int(T))
        if self.ws is None or self.ws.numel() < need:
            self.ws = None
            self.ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        self.ws_shape = (N, T)
        return self.ws

    def forward(self, feat, want_masks=False):
        """feat (N,T,idim) fp32 CUDA contiguous -> y_pred (N,num_spk,T,num_bins) [, head activations (N*T, 3*num_bins)]."""
        N, T, F = feat.shape
        ws = self.workspace(N, T)
        d = self.desc
        y_pred = torch.empty(N, d.num_spk, T, d.num_bins, dtype=torch.float32, device=feat.device)
        head = torch.empty(N * T, d.num_bins * d.num_out, dtype=torch.float32, device=feat.device) if want_masks else None
        p = self.dropout_p if self.model.training else 0.0
        self.last_dropout_seed = (self.dropout_seed + self.forward_calls) & 0xFFFFFFFF
        self.forward_calls += 1
        nv.check(self.lib.css_trainer_set_dropout(ctypes.c_void_p(self.handle), nv.c_float(p), ctypes.c_uint(self.last_dropout_seed)),
                 "css_trainer_set_dropout")
        nv.call("css_trainer_forward", ctypes.c_void_p(self.handle), nv.ptr(feat), nv.c_long(N), nv.c_int(T), nv.ptr(y_pred),
                nv.ptr(head), nv.c_int(0 if head is None else head.stride(0)), nv.ptr(ws), nv.c_size_t(ws.numel()))
        self.bn_tracked += 1
        return y_pred, head

    def enable_data_parallel(self, optim=None, group=None, buckets=8):
        """Data-parallel training over ``torch.distributed``: rank 0's parameters and BatchNorm running
        statistics are broadcast, the flat gradient is sum-all-reduced in ``buckets`` layer groups while the backward is
        still running (``GradientExchange``), and the mean is folded into the optimiser (``grad_scale = 1/world``)."""
        import torch.distributed as dist

        from .optim import GradientExchange

        if not dist.is_initialized() or dist.get_world_size(group) == 1:
            return None
        dist.broadcast(self.flat_p, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        dist.broadcast(self.bn_running, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self.exchange = GradientExchange(self, buckets=buckets, group=group)
        if optim is not None:
            optim.grad_scale = 1.0 / self.exchange.world
        self.model.invalidate()
        return self.exchange

    def backward_part(self, feat, grad_y, hi, lo):
        N, T = self.ws_shape
        nv.call("css_trainer_backward_part", ctypes.c_void_p(self.handle), nv.ptr(feat), nv.ptr(grad_y), nv.c_long(N), nv.c_int(T),
                nv.c_int(hi), nv.c_int(lo), nv.ptr(self.ws), nv.c_size_t(self.ws.numel()))

    def grad_slice(self, part_hi, part_lo):
        """Flat gradient slice that is final once parts part_hi..part_lo have run (the relative-position table and the
        embedding, which sit in front of block 0, are only final after part -1)."""
        from .optim import slice_bounds

        start, end = slice_bounds(self.part_bounds, self.size, part_hi, part_lo)
        return self.flat_g[start:end]

    def zero_grad(self):
        self.flat_g.zero_()
