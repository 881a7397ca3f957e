#!/usr/bin/env python
"""Headline benchmark: audio-seconds separated per second (xRT), 7-ch Conformer CSS (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W]             # this repo's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]       # the reference's CPU algorithm (oracle)

One step = one synthetic 7-mic LibriCSS-shaped session (default 600 s, 747 chunks) through the whole hot path:
global STFT -> IPD features -> 16-block Conformer masks -> DOA merge -> stitch -> MVDR -> iSTFT/placement ->
peak normalisation.  `value` times the path with the waveform already resident in HBM; `e2e` times the
public call (`CssPipeline.run_many`, one recording per step) from pinned HOST waveforms to the separated waveforms back in
pinned host memory, every step's H2D and D2H inside the timed region (double-buffered around the compute).  N > 1 (torchrun, one rank per GPU): every rank separates its own session (weak scaling,
no data-path collective -- sessions are independent, run.sh:21-36 of the reference); time = max over ranks.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "audio_seconds_separated_per_second_7ch_conformer_css"
UNIT = "xRT"
SR = 16000
TRAIN_DROPOUT = 0.1      # the reference's --conformer-dropout-rate default (css/models/conformer.py:40), used by train.sh
CONF_7CH = dict(idim=1799, num_bins=257, num_spk=2, num_noise=1, conformer_attention_dim=256,
                conformer_attention_heads=4, conformer_linear_units=1024, conformer_num_blocks=16,
                conformer_kernel_size=33, conformer_dropout_rate=0.1, conformer_relative_pos_emb=True)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--seconds", type=float, default=600.0, help="session length (LibriCSS sessions are ~10 min)")
    p.add_argument("--blocks", type=int, default=16)
    p.add_argument("--gemm-mode", type=int, default=-1, help="-1 library default, 0 fp32 SIMT, 1 bf16x3 tcgen05, 2 bf16 tcgen05")
    p.add_argument("--cpu-seconds", type=float, default=24.0, help="audio length of the bounded CPU-baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-extra", action="store_true", help="skip the long_form / train regions (configs[4] / configs[3])")
    p.add_argument("--extra-steps", type=int, default=3, help="timed steps of each extra region")
    p.add_argument("--long-seconds", type=float, default=3600.0, help="recording length of the long_form region (configs[4]: 1 hour)")
    return p.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"hbm": d["hbm_gbs"], "tensor_burst": d["bf16_tflops"], "tensor": d["bf16_tflops_sustained"], "src": "measured"}
    return {"hbm": 6650.0, "tensor_burst": 1590.0, "tensor": 1400.0, "src": "fallback"}


def build_cfg(conf_blocks):
    from css_b200.pipeline import CONFIG_7CH
    import copy

    return copy.deepcopy(CONFIG_7CH), dict(CONF_7CH, conformer_num_blocks=conf_blocks)


def make_model(conf):
    """Random-init weights of the reference architecture (torch.manual_seed(0), default init), eval mode."""
    import torch
    import css_b200.models as M

    torch.manual_seed(0)
    return M.build_model("Conformer", conf).eval()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out = self.proc.communicate(timeout=5)[0]
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out = self.proc.communicate()[0]
        sm, mx, reasons, power = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.strip().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); power.append(float(c[3]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(power)}


def cpu_reference_run(seconds, blocks, steps, warmup, threads=None):
    """The reference's CPU algorithm for the path (oracle/css_oracle.py; the reference is Python and does not
    travel to the GPU box) on a bounded `seconds`-long 7-ch sample, all host threads.  -> (xRT, ms/step, cores)."""
    import torch
    from css_b200.synth import make_recording
    from oracle import css_oracle as O

    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg, conf = build_cfg(blocks)
    model = make_model(conf)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    wav = make_recording(seconds, channels=7, seed=1234)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            O.css_pipeline(wav, cfg, lambda f: O.conformer_masks(f, sd))
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    mean = sum(times) / len(times)
    return seconds / mean, mean * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    secs = args.cpu_seconds
    xrt, ms, cores = cpu_reference_run(secs, args.blocks, max(1, args.steps), max(0, args.warmup))
    sample = f"{secs:g} s 7-ch session, {args.blocks}-block Conformer, oracle/css_oracle.py (torch CPU fp32), {cores} threads"
    line = {"impl": "reference", "metric": METRIC, "value": xrt, "unit": UNIT, "n_gpus": args.gpus, "steps": max(1, args.steps),
            "warmup": max(0, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"7-ch Conformer CSS (conf/config_7ch.yaml), bounded CPU sample of {secs:g} s per step",
                       "blocks": args.blocks, "session_seconds": secs},
            "cpu_baseline": {"value": xrt, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "reference_processes": 1, "reference_note": "ONE host process on this box's cores, whatever --gpus says: at N > 1 the "
            "driver's ratio compares N GPUs with one host",
            "e2e": {"value": xrt, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def region_long_form(args, cfg, model, dev, rank, world, barrier, max_over_ranks):
    """BASELINE configs[4] (long-form, strong scaling): ONE recording split into contiguous chunk ranges over the ranks
    (css_b200.dist.ShardedCssPipeline).  Timed: resident waveform -> every rank's normalised sample range on its own GPU;
    `e2e`: from the pinned host recording (each rank copies only its slice) to each rank's range back in ITS pinned host
    buffer (N parallel D2H copies, no rank moves the whole output); `gathered`: the ranges collected on rank 0 over NCCL."""
    import torch

    from css_b200.dist import ShardedCssPipeline
    from css_b200.synth import make_recording, n_chunks

    secs = args.long_seconds
    steps = max(1, args.extra_steps)
    # a 1-hour meeting is built by tiling a seeded 10-minute one (generation cost, not content, is what matters here)
    base = make_recording(min(secs, 600.0), channels=7, seed=4321)
    T = int(secs * SR)
    wav_host = base.repeat(1, (T + base.shape[1] - 1) // base.shape[1])[:, :T].contiguous().pin_memory()
    del base
    wav_dev = wav_host.to(dev)
    sh = ShardedCssPipeline(cfg, model, dev, rank, world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn, n):
        barrier()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / n

    piece, (s0, s1) = sh(wav_dev, gather=False)
    host_piece = torch.empty(2, max(1, s1 - s0), dtype=torch.float32).pin_memory()

    def step_e2e():
        p, _ = sh(wav_host, gather=False)
        if p.shape[1]:
            host_piece[:, : p.shape[1]].copy_(p, non_blocking=True)

    for _ in range(3):
        sh(wav_dev, gather=False)
    ms = timed(lambda: sh(wav_dev, gather=False), steps)
    coll = dict(sh.last_collective_bytes)
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, steps)
    out = {"workload": f"configs[4] long-form, strong scaling: one {secs:g} s 7-ch recording ({n_chunks(T)} chunks) split into "
                       f"contiguous chunk ranges over {world} GPU(s), 2-chunk recomputed halo",
           "value": secs / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": steps, "scaling": "strong",
           "collective_bytes_per_step": coll,
           "e2e": {"value": secs / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                   "h2d_bytes_per_step_per_rank": int(7 * 4 * T / world), "d2h_bytes_per_step_per_rank": int(2 * 4 * T / world)}}
    if world > 1:
        sh(wav_dev, gather=True)
        ms_g = timed(lambda: sh(wav_dev, gather=True), steps)
        out["gathered"] = {"value": secs / (ms_g * 1e-3), "unit": UNIT, "ms_per_step": ms_g, "nccl_gather_bytes_per_step": 8 * T}
        # one GPU on the same recording in the same run (rank 0 works, the others wait): strong-scaling efficiency
        one = ShardedCssPipeline(cfg, model, dev, 0, 1)
        if rank == 0:
            for _ in range(3):
                one(wav_dev, gather=False)
        barrier()
        e0.record()
        if rank == 0:
            for _ in range(steps):
                one(wav_dev, gather=False)
        e1.record()
        barrier()
        ms1 = max_over_ranks(e0.elapsed_time(e1)) / steps
        out["one_gpu_ms_per_step"] = ms1
        out["efficiency_vs_one_gpu"] = ms1 / (world * ms)
    return out


def region_train(args, dev, rank, world, barrier, max_over_ranks):
    """BASELINE configs[3]: utterance-level PIT training step (css/trainer/iteration_types.py:34-63), 16-block 7-ch
    Conformer, batch 32 x 250 frames per GPU, bf16 tensor-core GEMMs (gemm mode 2), clip 5.0 + Adam(1e-4, wd 1e-2);
    N > 1: bucketed NCCL gradient all-reduce overlapped with the backward.  Inputs come from pinned host memory each step."""
    import torch

    import css_b200.models as M
    from css_b200 import _native as nv
    from css_b200.objectives import build_objective
    from css_b200.trainer import FusedAdam

    steps = max(1, args.extra_steps)
    saved_mode = nv.get_gemm_mode()
    nv.set_gemm_mode(2)
    try:
        conf = dict(CONF_7CH, conformer_num_blocks=args.blocks, conformer_dropout_rate=TRAIN_DROPOUT)
        torch.manual_seed(0)
        model = M.build_model("Conformer", conf).train().to(dev)
        engine = model.train_engine()
        opt = FusedAdam(engine, lr=1e-4, weight_decay=1e-2)
        objective = build_objective("MSE", {})
        g = torch.Generator().manual_seed(100 + rank)
        B, T = 32, 250
        s1 = (torch.rand(B, T, 257, generator=g) ** 2).pin_memory()
        s2 = (torch.rand(B, T, 257, generator=g) ** 2).pin_memory()
        mix = torch.empty(B, T, 1799)
        mix[..., :257] = s1 + s2
        mix[..., 257:] = (torch.rand(B, T, 1542, generator=g) * 2 - 1) * 3.14159
        host = {"mix": mix.pin_memory(), "source1": s1, "source2": s2, "ovl": 0.5}
        loss_host = torch.zeros(1).pin_memory()

        def step():
            batch = {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) else v) for k, v in host.items()}
            loss = objective(model, batch, device=dev)
            loss.backward()
            if engine.exchange is not None:
                engine.exchange.finish()
            opt.clip_grad_norm_(5.0)
            opt.step()
            opt.zero_grad()
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)

        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def timed(n):
            barrier()
            e0.record()
            for _ in range(n):
                step()
            e1.record()
            barrier()
            return max_over_ranks(e0.elapsed_time(e1)) / n

        for _ in range(3):
            step()
        ms_local = timed(steps)                                  # every rank alone: no exchange
        out = {"workload": "configs[3]: PIT training step, 16-block 7-ch Conformer, batch 32 x 250 frames per GPU, bf16 GEMMs, "
                           f"dropout {TRAIN_DROPOUT}, clip 5.0, Adam(1e-4, wd 1e-2); inputs from pinned host memory every step",
               "unit": "utt/s", "steps": steps, "scaling": "weak", "h2d_bytes_per_step": (mix.numel() + 2 * s1.numel()) * 4,
               "d2h_bytes_per_step": 4}
        if world > 1:
            engine.enable_data_parallel(opt)
            step()
            ms = timed(steps)
            out.update(value=world * B / (ms * 1e-3), ms_per_step=ms, ms_per_step_without_exchange=ms_local,
                       efficiency_vs_n_times_one_gpu=ms_local / ms, allreduce_bytes_per_step=int(engine.size) * 4,
                       allreduce_buckets=len(engine.exchange.ranges))
        else:
            out.update(value=B / (ms_local * 1e-3), ms_per_step=ms_local, allreduce_bytes_per_step=0)
        out["final_loss"] = float(loss_host[0])
        return out
    finally:
        nv.set_gemm_mode(saved_mode)


def run_b200(args):
    import torch
    import torch.distributed as dist

    from css_b200 import _native as nv
    from css_b200.dist import init_from_env
    from css_b200.pipeline import CssPipeline
    from css_b200.synth import make_recording, n_chunks

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- css_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    rank, world, local = init_from_env("nccl")
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torch.distributed.run)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lib = nv.lib()
    if args.gemm_mode >= 0:
        nv.check(lib.css_set_gemm_mode(args.gemm_mode), "css_set_gemm_mode")
    gemm_mode = int(lib.css_get_gemm_mode())

    cfg, conf = build_cfg(args.blocks)
    model = make_model(conf).to(dev)
    pipe = CssPipeline(cfg, model, device=dev)
    T = int(args.seconds * SR)
    wav_host = make_recording(args.seconds, channels=7, seed=1234 + rank).pin_memory()
    wav_dev = wav_host.to(dev)
    out_host = torch.empty(2, T, dtype=torch.float32).pin_memory()
    B = n_chunks(T)
    audio_s = T / SR

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

    # One CUDA graph per step (CssPipeline.capture: the ~800 launches of a session replayed by one call): issuing them one by
    # one takes the host ~18 ms per 26 ms step, so a host that is descheduled for a moment (seen on the GPU boxes: 0.5 s stalls)
    # starves the GPU and halves the number.  CSS_B200_GRAPHS=0: eager launches.
    graph = None

    def step_resident():
        return graph.replay() if graph is not None else pipe(wav_dev)

    def step_e2e():
        out = pipe(wav_host)                       # H2D of the session inside the public call
        out_host.copy_(out, non_blocking=True)     # D2H of the two separated waveforms
        return out

    for _ in range(max(args.warmup, 3)):
        step_resident()
    step_e2e()
    if os.environ.get("CSS_B200_GRAPHS", "1") != "0":
        try:
            graph = pipe.capture(wav_dev)
            for _ in range(3):
                step_resident()
        except Exception as e:      # capture is an optimisation of the launch path, not of the kernels: fall back to eager launches
            print(f"[bench] CUDA graph capture failed, eager launches: {e!r}", file=sys.stderr, flush=True)
            graph = None
            torch.cuda.synchronize()
    barrier()

    # ---- timed region 1: inputs resident in HBM (value) ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if graph is None:
        lib.css_profile_enable(1)
    nv.launch_count(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_resident()
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = nv.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    import ctypes
    ncat = 11
    ms_c, work_c, cnt_c = (ctypes.c_double * ncat)(), (ctypes.c_double * ncat)(), (ctypes.c_long * ncat)()
    ms_prof_total = ms_total
    if graph is not None:
        # per-kernel CUDA events cannot bracket the kernels of a replayed graph: the same K steps once more, launched eagerly
        # with the events around every instrumented launch (roofline, hbm_kernels, share_of_step, gpu_launches come from here)
        lib.css_profile_enable(1)
        nv.launch_count(reset=True)
        barrier()
        e0.record()
        for _ in range(args.steps):
            pipe(wav_dev)
        e1.record()
        barrier()
        ms_prof_total = max_over_ranks(e0.elapsed_time(e1))
        launches = nv.launch_count()
    nv.check(lib.css_profile_collect(ms_c, work_c, cnt_c, ncat), "css_profile_collect")
    lib.css_profile_enable(0)

    # ---- timed region 2: public call from pinned host memory, results back on the host (e2e): one recording per step
    # through CssPipeline.run_many, which double-buffers the H2D / D2H copies of consecutive recordings around the compute
    outs_host = [out_host, torch.empty_like(out_host).pin_memory()]
    pipe.run_many([wav_host] * 3, [outs_host[i % 2] for i in range(3)])
    # the region is run twice and the MEAN reported (both passes are kept in the JSON): host<->device copy rates on a shared
    # box vary from one second to the next, the kernels do not
    e2e_runs = []
    for _ in range(2):
        barrier()
        e0.record()
        pipe.run_many([wav_host] * args.steps, [outs_host[i % 2] for i in range(args.steps)])
        e1.record()
        barrier()
        e2e_runs.append(max_over_ranks(e0.elapsed_time(e1)))
    ms_e2e = sum(e2e_runs) / len(e2e_runs)          # mean of the passes (both are kept in the JSON)

    # ---- N > 1: the same e2e region with the NCCL gather of every rank's separated waveforms onto rank 0 (north_star: "NCCL
    # over NVLink is used only to gather stitched outputs"); rank 0 alone copies all N results to the host
    gathered = None
    if world > 1:
        from css_b200.dist import gather_waveforms

        lengths = [T] * world
        big_host = torch.empty(world, 2, T, dtype=torch.float32).pin_memory() if rank == 0 else None

        def step_gather():
            out = pipe(wav_host)
            res = gather_waveforms({rank: out}, lengths, rank, world, dst=0)
            if rank == 0:
                for i, r in enumerate(res):
                    big_host[i].copy_(r, non_blocking=True)

        step_gather()
        barrier()
        e0.record()
        for _ in range(args.steps):
            step_gather()
        e1.record()
        barrier()
        ms_g = max_over_ranks(e0.elapsed_time(e1))
        gathered = {"value": world * audio_s * args.steps / (ms_g * 1e-3), "unit": UNIT, "ms_per_step": ms_g / args.steps,
                    "nccl_gather_bytes_per_step": (world - 1) * 2 * T * 4, "d2h_bytes_per_step_rank0": world * 2 * T * 4,
                    "note": "every rank's (2, T) result gathered onto rank 0 over NCCL, rank 0 copies all of them to the host"}
    del pipe
    extra = {}
    if not args.no_extra:
        extra["long_form"] = region_long_form(args, cfg, model, dev, rank, world, barrier, max_over_ranks)
        del model
        torch.cuda.empty_cache()
        extra["train"] = region_train(args, dev, rank, world, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    ms_step = ms_total / args.steps
    value = world * audio_s * args.steps / (ms_total * 1e-3)
    e2e = world * audio_s * args.steps / (ms_e2e * 1e-3)
    # dominant kernel: the dense Linear GEMMs of the Conformer (category 0)
    g_ms, g_flop, g_n = ms_c[0], work_c[0], cnt_c[0]
    achieved = g_flop / (g_ms * 1e-3) / 1e12 if g_ms > 0 else 0.0
    names = ("gemm", "attention", "conv_module", "layernorm", "lstm_step", "stft", "features_mvn", "doa_merge", "stitch", "mvdr",
             "istft_peak")
    shares = {names[i]: round(ms_c[i] / ms_prof_total, 4) for i in range(ncat) if cnt_c[i]}
    # HBM-bound kernels (north_star: STFT, mask, MVDR, stitch): algorithmic bytes / CUDA-event time vs measured copy peak
    hbm = {names[i]: {"achieved_GBps": round(work_c[i] / (ms_c[i] * 1e-3) / 1e9, 1), "frac": round(work_c[i] / (ms_c[i] * 1e-3) / 1e9 / pk["hbm"], 4),
                      "ms_per_step": round(ms_c[i] / args.steps, 4), "algorithmic_MB_per_step": round(work_c[i] / args.steps / 1e6, 1)}
           for i in (2, 3, 5, 6, 7, 8, 9, 10) if cnt_c[i] and ms_c[i] > 0}
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("gemm_tc", {}).get("bytes_per_launch")
    roof = {"kernel": {0: "gemm_simt_f32", 1: "gemm_tc_bf16x3 (tcgen05)", 2: "gemm_tc_bf16 (tcgen05)"}[gemm_mode],
            "bound": "tensor", "achieved": achieved, "peak": pk["tensor"], "unit": "TFLOP/s",
            "frac": achieved / pk["tensor"], "peak_source": f"{pk['src']} bf16 sustained (kernel timed inside a long step)",
            "traffic": traffic, "traffic_note": "ncu dram bytes per launch, mean over the six GEMMs of one block (profiles/r2_traffic.json)",
            "launches": int(g_n), "avg_launch_ms": g_ms / max(1, g_n),
            "algorithmic_flop_per_step": g_flop / args.steps, "share_of_step": shares,
            "measured_in": ("the same K steps launched eagerly right after the timed region, CUDA events around every instrumented "
                            f"launch ({ms_prof_total / args.steps:.3f} ms per step there); the timed region replays one CUDA graph per step"
                            if graph is not None else "the timed region (eager launches)")}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {0: "f32", 1: "bf16x3->f32", 2: "bf16"}[gemm_mode], "data": "synthetic",
            "config": {"workload": "configs[2]: 7-ch Conformer (conf/config_7ch.yaml) mask estimation + MVDR on a synthetic "
                                   f"7-mic LibriCSS-shaped {audio_s:g} s session per GPU per step",
                       "session_seconds": audio_s, "chunks_per_session": B, "blocks": args.blocks, "sessions_per_step": world,
                       "l2": f"inputs larger than L2 (waveform {wav_dev.numel() * 4 / 1e6:.0f} MB, spectrum {7 * (50 * (B - 1) + 150) * 264 * 8 / 1e6:.0f} MB, network input planes {B * 150 * 1856 * 4 / 1e6:.0f} MB vs 126 MB L2)",
                       "parallelism": f"dp{world} (independent sessions, no data-path collective)", "gemm_mode": gemm_mode,
                       "launch": (f"one CUDA graph replay per step ({int(launches) // max(1, args.steps)} kernels captured once, CssPipeline.capture); "
                                  "e2e: CssPipeline.run_many over two captured graphs" if graph is not None else "eager launches")},
            "clocks": clocks, "roofline": roof, "hbm_kernels": {"peak_GBps": pk["hbm"], "kernels": hbm},
            "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps, "ms_per_step_runs": [round(t / args.steps, 3) for t in e2e_runs],
                    "h2d_bytes_per_step": wav_host.numel() * 4, "d2h_bytes_per_step": out_host.numel() * 4},
            "gpu_launches": int(launches)}
    if gathered:
        line["e2e_gathered_to_rank0"] = gathered
    line.update(extra)
    if world == 1 and not args.no_cpu_baseline:
        torch.cuda.synchronize()
        xrt, ms, cores = cpu_reference_run(args.cpu_seconds, args.blocks, 1, 0)
        line["cpu_baseline"] = {"value": xrt, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{args.cpu_seconds:g} s 7-ch session through oracle/css_oracle.py (torch CPU fp32, "
                                          f"{args.blocks}-block Conformer), 1 pass, {cores} threads"}
        # the reference's own session setting is ONE thread (ORT intra/inter-op 1, scripts/python/separate_libricss.py:40-44)
        one_s = min(args.cpu_seconds, 8.0)
        xrt1, _, _ = cpu_reference_run(one_s, args.blocks, 1, 0, threads=1)
        line["cpu_baseline_1thread"] = {"value": xrt1, "unit": UNIT, "cores": 1, "kind": "port",
                                        "sample": f"{one_s:g} s 7-ch session, same oracle path, torch.set_num_threads(1)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
