#!/usr/bin/env python
"""Long-form stress (BASELINE.json configs[4]): ONE 1-hour 7-ch synthetic meeting, chunk stitching + permutation alignment
+ MVDR, split across the GPUs of a box by contiguous chunk ranges (css_b200.dist.ShardedCssPipeline).

    python bench_long.py [--gpus N] [--seconds 3600] [--steps K] [--warmup W]      (N>1: launch with torch.distributed.run)

Strong scaling: the work per step is fixed (one recording), value = audio seconds / step time (max over ranks, CUDA
events, barrier on both sides).  The waveform is resident on every rank when the timed region starts; `e2e` re-runs
with the rank's slice copied from pinned host memory inside the timed region and the result copied back on rank 0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from bench import CONF_7CH, ClockSampler, build_cfg, make_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--seconds", type=float, default=3600.0)
    ap.add_argument("--blocks", type=int, default=16)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from css_b200 import _native as nv
    from css_b200.dist import ShardedCssPipeline, chunk_ranges, init_from_env
    from css_b200.synth import make_recording, n_chunks

    rank, world, local = init_from_env("nccl")
    if world != args.gpus:
        raise SystemExit(f"bench_long.py: --gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cfg, conf = build_cfg(args.blocks)
    sh = ShardedCssPipeline(cfg, make_model(conf).to(dev), dev, rank, world)
    # a 1-hour session is built by tiling a seeded 10-minute one (generation cost, not content, is what matters here)
    base = make_recording(min(args.seconds, 600.0), channels=7, seed=4321)
    T = int(args.seconds * 16000)
    reps = (T + base.shape[1] - 1) // base.shape[1]
    wav_host = base.repeat(1, reps)[:, :T].contiguous().pin_memory()
    wav_dev = wav_host.to(dev)
    out_host = torch.empty(2, T).pin_memory() if rank == 0 else None
    B = n_chunks(T)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def step_e2e():
        out = sh(wav_host)
        if rank == 0:
            out_host.copy_(out, non_blocking=True)

    for _ in range(max(3, args.warmup)):
        sh(wav_dev)
    step_e2e()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    nv.launch_count(reset=True)
    ms = timed(lambda: sh(wav_dev), args.steps)
    launches = nv.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = timed(step_e2e, args.steps)
    if rank == 0:
        c0, c1 = chunk_ranges(B, world)[0]
        line = {"metric": "audio_seconds_separated_per_second_7ch_conformer_css_long_form", "value": args.seconds * args.steps / (ms * 1e-3),
                "unit": "xRT", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16x3->f32", "data": "synthetic",
                "config": {"workload": f"configs[4]: one {args.seconds:g} s 7-ch meeting, {B} chunks, {B - 1} pair decisions, split into "
                                       f"contiguous chunk ranges over {world} GPU(s) with a 2-chunk recomputed halo", "blocks": args.blocks,
                           "chunks_rank0": c1 - c0, "parallelism": f"{world} ranks: parity all-gather ({world} x 4 B) + NCCL reduce of (2, T) fp32",
                           "l2": "inputs larger than L2"},
                "clocks": clocks, "e2e": {"value": args.seconds * args.steps / (ms_e2e * 1e-3), "unit": "xRT", "ms_per_step": ms_e2e / args.steps,
                                          "h2d_bytes_per_step": int(7 * 4 * T / world), "d2h_bytes_per_step": 2 * 4 * T},
                "gpu_launches": int(launches)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
