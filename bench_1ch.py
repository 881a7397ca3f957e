#!/usr/bin/env python
"""Single-channel workloads of BASELINE.json: configs[0] (BLSTM mask estimator, conf/config_1ch.yaml, one 60 s mixture) and
configs[1] (1-ch Conformer, 10-minute LibriCSS-shaped sessions, 8 recordings per GPU = 64 on 8 GPUs).

    python bench_1ch.py --workload conformer|blstm [--gpus N] [--steps K] [--warmup W]    (N>1: torch.distributed.run)

The 1-ch flow of the reference has no beamformer input to work with (one microphone): the result of a recording is the three
stitched session-level masks (`stitcher.get_connect`, stitcher.py:54-110).  One step = every rank separates its
`--recordings` recordings; `value` = audio seconds / step time with the waveforms resident in HBM (CUDA events, barrier on
both sides, max over ranks); `e2e` runs `CssPipeline.run_many` from pinned host waveforms to stitched masks in pinned host
memory.  Weak scaling: recordings per rank are fixed, no data-path collective.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from bench import CONF_7CH, ClockSampler  # noqa: E402

SR = 16000
# blstm.py:28-29 argparse defaults (what train.sh produces): 1024 x 3
CONF_BLSTM = dict(idim=257, num_bins=257, num_spk=2, num_noise=1, blstm_hdim=1024, blstm_num_layers=3, blstm_dropout_rate=0.1)


def build(workload, blocks):
    import copy

    import torch

    import css_b200.models as M
    from css_b200.pipeline import CONFIG_1CH

    torch.manual_seed(0)
    cfg = copy.deepcopy(CONFIG_1CH)
    cfg["separation"].update(merge=False, merge_threshold=16)      # the 1-ch yaml has no merge keys
    if workload == "blstm":
        return cfg, M.build_model("BLSTM", dict(CONF_BLSTM)).eval()
    return cfg, M.build_model("Conformer", dict(CONF_7CH, idim=257, conformer_num_blocks=blocks)).eval()


def cpu_port(workload, blocks, seconds):
    """The oracle's 1-ch pipeline (separate -> stitch -> connect) on all host threads: (xRT, cores)."""
    import torch

    from css_b200.synth import make_recording
    from oracle import css_oracle as O

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg, model = build(workload, blocks)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    wav = make_recording(seconds, channels=1, seed=1234)
    fn = (lambda f: O.blstm_masks(f, sd)) if workload == "blstm" else (lambda f: O.conformer_masks(f, sd))
    with torch.no_grad():
        t0 = time.perf_counter()
        O.css_pipeline(wav, cfg, fn)
        dt = time.perf_counter() - t0
    return seconds / dt, cores


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="conformer", choices=["conformer", "blstm"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--recordings", type=int, default=0, help="recordings per GPU per step (default 8 conformer, 1 blstm)")
    ap.add_argument("--seconds", type=float, default=0.0, help="recording length (default 600 conformer, 60 blstm)")
    ap.add_argument("--blocks", type=int, default=16)
    ap.add_argument("--cpu-seconds", type=float, default=0.0, help="CPU sample length (default 24 conformer, 60 = whole workload blstm)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    wl = args.workload
    n_rec = args.recordings or (8 if wl == "conformer" else 1)
    seconds = args.seconds or (600.0 if wl == "conformer" else 60.0)
    cpu_seconds = args.cpu_seconds or (24.0 if wl == "conformer" else 60.0)

    import torch
    import torch.distributed as dist

    from css_b200 import _native as nv
    from css_b200.dist import init_from_env
    from css_b200.pipeline import CssPipeline
    from css_b200.synth import make_recording, n_chunks

    if not torch.cuda.is_available():
        raise SystemExit("bench_1ch.py: no CUDA device -- css_b200 has no CPU fallback")
    rank, world, local = init_from_env("nccl")
    if world != args.gpus:
        raise SystemExit(f"bench_1ch.py: --gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cfg, model = build(wl, args.blocks)
    pipe = CssPipeline(cfg, model.to(dev), device=dev)
    T = int(seconds * SR)
    B = n_chunks(T)
    # two distinct seeded recordings per rank, dealt alternately (generation cost, not content, limits more)
    base = [make_recording(seconds, channels=1, seed=1234 + 2 * rank + i).pin_memory() for i in range(min(2, n_rec))]
    host = [base[i % len(base)] for i in range(n_rec)]
    devs = [h.to(dev) for h in base]
    resident = [devs[i % len(devs)] for i in range(n_rec)]
    probe = pipe(resident[0])
    outs = [torch.empty(probe.shape, dtype=torch.float32).pin_memory() for _ in range(min(2, n_rec))]
    out_list = [outs[i % len(outs)] for i in range(n_rec)]

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

    def step_resident():
        for w in resident:
            pipe(w)

    def step_e2e():
        pipe.run_many(host, out_list)

    for _ in range(max(3, args.warmup)):
        step_resident()
    step_e2e()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    nv.launch_count(reset=True)
    ms = timed(step_resident, args.steps)
    launches = nv.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    ms_e2e = min(timed(step_e2e, args.steps), timed(step_e2e, args.steps))
    # per-category kernel time (CUDA events around every launch on the launching stream) in a separate, untimed pass
    import ctypes
    lib = nv.lib()
    ncat = 11
    ms_c, work_c, cnt_c = (ctypes.c_double * ncat)(), (ctypes.c_double * ncat)(), (ctypes.c_long * ncat)()
    lib.css_profile_enable(1)
    step_resident()
    torch.cuda.synchronize()
    nv.check(lib.css_profile_collect(ms_c, work_c, cnt_c, ncat), "css_profile_collect")
    lib.css_profile_enable(0)
    names = ("gemm", "attention", "conv_module", "layernorm", "lstm_step", "stft", "features_mvn", "doa_merge", "stitch", "mvdr", "istft_peak")
    kernels = {names[i]: {"ms_per_step": round(ms_c[i], 3), "launches": int(cnt_c[i]),
                          "achieved": round(work_c[i] / (ms_c[i] * 1e-3) / (1e12 if i in (0, 1, 4) else 1e9), 1),
                          "unit": "TFLOP/s" if i in (0, 1, 4) else "GB/s"} for i in range(ncat) if cnt_c[i] and ms_c[i] > 0}
    if rank == 0:
        audio = world * n_rec * seconds * args.steps
        name = {"conformer": "configs[1]: 1-ch Conformer CSS inference", "blstm": "configs[0]: 1-ch BLSTM 1024x3 mask estimator (conf/config_1ch.yaml) chunked CSS inference"}[wl]
        line = {"metric": f"audio_seconds_separated_per_second_1ch_{wl}_css", "value": audio / (ms * 1e-3), "unit": "xRT", "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16x3->f32", "data": "synthetic",
                "config": {"workload": f"{name}, {n_rec} x {seconds:g} s synthetic recording(s) per GPU per step ({B} chunks each), "
                                       "separate -> stitch -> connect (stitched session masks are the 1-ch result)",
                           "recordings_per_gpu": n_rec, "chunks_per_recording": B, "blocks": args.blocks if wl == "conformer" else None,
                           "parallelism": f"dp{world} (independent recordings, no data-path collective)",
                           "l2": "features + activations of a step larger than L2" if wl == "conformer" else
                                 "60 s recording: working set below L2 -- latency-bound workload, the reference's own CPU-runnable case"},
                "clocks": clocks,
                "e2e": {"value": audio / (ms_e2e * 1e-3), "unit": "xRT", "ms_per_step": ms_e2e / args.steps,
                        "h2d_bytes_per_step": n_rec * T * 4, "d2h_bytes_per_step": n_rec * probe.numel() * 4},
                "gpu_launches": int(launches), "kernels": kernels}
        if world == 1 and not args.no_cpu_baseline:
            xrt, cores = cpu_port(wl, args.blocks, cpu_seconds)
            line["cpu_baseline"] = {"value": xrt, "unit": "xRT", "cores": cores, "kind": "port",
                                    "sample": f"{cpu_seconds:g} s 1-ch recording through oracle/css_oracle.py (torch CPU fp32), 1 pass, {cores} threads"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
