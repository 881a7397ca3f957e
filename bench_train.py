#!/usr/bin/env python
"""Secondary benchmark: the utterance-level PIT training step (BASELINE.json configs[3]).

    python bench_train.py [--gpus N] [--steps K] [--warmup W] [--gemm-mode 2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench_train.py --gpus N

One step = forward + PIT-MSE + backward + (N>1: bucketed NCCL gradient all-reduce overlapped with the backward) +
clip_grad_norm(5.0) + Adam(lr 1e-4, wd 1e-2) on a batch of 32 synthetic 2-speaker 4 s (250-frame) 7-ch feature
utterances per GPU, 16-block Conformer (22.0 M parameters, dropout 0.1 as in the reference), as css/trainer/iteration_types.py:34-63.
Prints one JSON line in bench.py's format (value = utterances/s over all GPUs; weak scaling: 32 utterances per GPU).
The headline benchmark of the repo is bench.py (inference xRT); this file measures the training row of SURVEY 8.
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from bench import CONF_7CH, ClockSampler, peaks  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--frames", type=int, default=250)
    ap.add_argument("--blocks", type=int, default=16)
    ap.add_argument("--buckets", type=int, default=4)
    ap.add_argument("--dropout", type=float, default=0.1, help="the reference trains with --conformer-dropout-rate 0.1")
    ap.add_argument("--gemm-mode", type=int, default=2, help="2 bf16 tcgen05 (config 4's compute type), 1 bf16x3 (fp32-class)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    import css_b200.models as M
    from css_b200 import _native as nv
    from css_b200.dist import init_from_env
    from css_b200.objectives import build_objective
    from css_b200.trainer import FusedAdam, GradientExchange

    if not torch.cuda.is_available():
        raise SystemExit("bench_train.py: no CUDA device -- css_b200 has no CPU fallback")
    rank, world, local = init_from_env("nccl")
    if world != args.gpus:
        raise SystemExit(f"bench_train.py: --gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lib = nv.lib()
    nv.set_gemm_mode(args.gemm_mode)
    conf = dict(CONF_7CH, conformer_num_blocks=args.blocks, conformer_dropout_rate=args.dropout)
    torch.manual_seed(0)
    model = M.build_model("Conformer", conf).train().to(dev)
    engine = model.train_engine()
    opt = FusedAdam(engine, lr=1e-4, weight_decay=1e-2)
    if world > 1:
        engine.exchange = GradientExchange(engine, buckets=args.buckets)
        opt.grad_scale = 1.0 / world
    objective = build_objective("MSE", {})
    g = torch.Generator().manual_seed(100 + rank)
    B, T = args.batch, args.frames
    s1 = (torch.rand(B, T, 257, generator=g) ** 2).pin_memory()
    s2 = (torch.rand(B, T, 257, generator=g) ** 2).pin_memory()
    mix = torch.empty(B, T, 1799)
    mix[..., :257] = s1 + s2
    mix[..., 257:] = (torch.rand(B, T, 1542, generator=g) * 2 - 1) * 3.14159
    mix = mix.pin_memory()
    host = {"mix": mix, "source1": s1, "source2": s2, "ovl": 0.5}
    resident = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in host.items()}
    loss_host = torch.zeros(1).pin_memory()

    def step(batch, read_loss=False):
        if batch is host:
            batch = {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) else v) for k, v in host.items()}
        loss = objective(model, batch, device=dev)
        loss.backward()
        if engine.exchange is not None:
            engine.exchange.finish()
        opt.clip_grad_norm_(5.0)
        opt.step()
        opt.zero_grad()
        if read_loss:
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        step(resident)
    step(host, True)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.css_profile_enable(1)
    nv.launch_count(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        last = step(resident)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = nv.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    ncat = 5
    ms_c, work_c, cnt_c = (ctypes.c_double * ncat)(), (ctypes.c_double * ncat)(), (ctypes.c_long * ncat)()
    nv.check(lib.css_profile_collect(ms_c, work_c, cnt_c, ncat), "css_profile_collect")
    lib.css_profile_enable(0)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step(host, True)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    final_loss = float(last)
    if rank == 0:
        pk = peaks()
        g_ms, g_flop, g_n = ms_c[0], work_c[0], cnt_c[0]
        achieved = g_flop / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
        names = ("gemm", "attention", "conv_module", "layernorm", "lstm_step")
        h2d = (mix.numel() + s1.numel() + s2.numel()) * 4
        line = {"metric": "pit_training_utterances_per_second_7ch_conformer", "value": world * B * args.steps / (ms_total * 1e-3),
                "unit": "utt/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {1: "bf16x3->f32", 2: "bf16"}[args.gemm_mode], "data": "synthetic",
                "config": {"workload": "configs[3]: utterance-level PIT training step, 7-ch Conformer, batch 32 x 250 frames per GPU, "
                                       f"Adam(1e-4, wd 1e-2), clip 5.0, dropout {args.dropout:g}", "blocks": args.blocks, "batch_per_gpu": B,
                           "frames": T, "audio_seconds_per_step": world * B * T * 256 / 16000.0,
                           "parallelism": f"dp{world}, gradient all-reduce over NCCL in {args.buckets} buckets overlapped with backward",
                           "l2": "activations saved per step (~5 GB at 16 blocks) exceed L2", "gemm_mode": args.gemm_mode},
                "clocks": clocks,
                "roofline": {"kernel": "gemm_tc (tcgen05) fwd + dX + dW", "bound": "tensor", "achieved": achieved, "peak": pk["tensor"],
                             "unit": "TFLOP/s", "frac": achieved / pk["tensor"], "traffic": None, "launches": int(g_n),
                             "avg_launch_ms": g_ms / max(1, g_n),
                             "share_of_step": {names[i]: round(ms_c[i] / ms_total, 4) for i in range(ncat) if cnt_c[i]}},
                "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": "utt/s", "ms_per_step": ms_e2e / args.steps,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
                "gpu_launches": int(launches), "final_loss": final_loss}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
