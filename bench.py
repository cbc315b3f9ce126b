#!/usr/bin/env python
"""Benchmark of the CFM mel-infilling decode hot path (BASELINE.json metric: CFM-decoded mel frames/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C1|C2|C3|C4|C5]

Default workload = BASELINE config 2 (ssc1 decoder, random-init, batch 16, 10 s clips = 1000 frames, 50 % infill,
10 Euler steps, bf16 tensor-core math); a "step" is one full decode of one batch (C2: 16 000 mel frames).
--config selects another BASELINE config: C1 one 5 s clip; C3 64 ragged clips of 3-20 s, 32 steps;
C5 the per-GPU share of the long-form config (8 x 60 s, 50 steps); C4 the utterance-sharded decode of 1024 clips
(strong scaling: the clips are split over the ranks, a step is one pass over all of them through the pipelined
host loop of serenade_b200/sharding.py).
One process per GPU (torchrun for N > 1); utterances are sharded by rank with NO collective in the timed region
(C1/C2/C3/C5: weak scaling, every rank decodes its own batch); an NCCL all_gather of the output mels / checksums runs
after the timed region for reporting only.

Printed JSON (rank 0, one line):
  value          whole-job valid frames/s with inputs resident in HBM (CUDA events, L2 flushed between steps, max over ranks)
  e2e            the same through the host-buffer C-ABI call (H2D of z/mu/mask/spk and D2H of the mel inside the timed region)
  roofline       SURVEY 8d algorithmic tensor FLOPs of the step / time vs the measured sustained bf16 peak; per-kernel-class
                 views from srn_profile_ops (algorithmic AND issued FLOPs); `traffic` from the committed ncu launch list
  softmax_path   how many attention entries ran shift-free in the timed solve, and the solve time with that path off
  tf32_mode      this library's precision="tf32" (the mode that passes the 1e-3 per-step gate) on the same workload
  gpu_reference  the reference modules' math (oracle/cfm_oracle.py functions) in eager PyTorch on the SAME GPU, TF32 and
                 bf16 autocast: what a user gets today by moving the reference to a B200 (SURVEY 8d "GPU reference beside it")
  cpu_baseline   the CPU oracle (a port of the reference's torch CPU path) on a bounded sample, host cores stated
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "cfm_decoded_mel_frames_per_sec"
UNIT = "frames/s"


def _c3_lengths():
    g = torch.Generator().manual_seed(1234 + 2)
    return sorted(torch.randint(300, 2001, (64,), generator=g).tolist(), reverse=True)


# name -> workload.  cpu = (clips, Euler steps) of the bounded CPU sample (scaled to the full step count)
CONFIGS = {
    "C1": dict(workload="C1: ssc1 decoder random-init, batch 1, one 5 s clip (500 frames), infill = last half, 10 Euler steps",
               lengths=[500], n=10, seed=1234, infill=(0.5, 1.0), cpu=(1, 10), ref=(1, 10)),
    "C2": dict(workload="C2: ssc1 decoder random-init, batch 16 x 1000 frames (10 s clips), 50% infill, 10 Euler steps",
               lengths=[1000] * 16, n=10, seed=1235, infill=(0.25, 0.75), cpu=(16, 10), ref=(4, 10)),
    "C3": dict(workload="C3: ssc1 decoder random-init, batch 64, mixed lengths 300..2000 frames (padded + masked, sorted "
                        "descending), 32 Euler steps",
               lengths=_c3_lengths(), n=32, seed=1236, infill=(0.25, 0.75), cpu=(4, 4), ref=(2, 2)),
    "C5": dict(workload="C5 (per-GPU share): ssc1 decoder random-init, batch 8 x 6000 frames (60 s clips), 50% infill, "
                        "50 Euler steps",
               lengths=[6000] * 8, n=50, seed=1238, infill=(0.25, 0.75), cpu=(1, 4), ref=(1, 2)),
}
C4 = dict(workload="C4: utterance-sharded decode of 1024 synthetic clips (each as C2's: 1000 frames, 50% infill, 10 Euler "
                   "steps), batches of 16, clips split over the ranks", clips=1024, frames=1000, n=10, batch=16)


def _traffic(config):
    """DRAM bytes (read + write) of one timed step, summed over its kernels, from the committed ncu launch list
    (profiles/r02_traffic.json, written by tools/launch_summary.py from the same bench command under ncu)."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        try:
            d = json.load(open(p))
            if d.get("config", "C2") != config:
                continue
            return float(d["dram_bytes_per_step"]), d.get("source", p), d.get("dram_bytes_per_step_warm")
        except Exception:
            continue
    return None, None, None


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                        tf_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
        except Exception:
            pass
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        time.sleep(0.05)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm),
                "reasons": sorted(reasons)}


def _cpu_decode_sample(cfg, threads: int, nclips: int, nsteps: int):
    """Times the CPU oracle (port of the reference's torch CPU fp32 path) on the first `nclips` clips of the workload
    for `nsteps` Euler steps; returns (valid frames/s scaled to the config's full step count, seconds)."""
    import serenade_b200 as S
    from oracle import cfm_oracle as O          # CPU baseline leg: allowed to execute oracle/
    torch.set_num_threads(threads)
    sd = S.make_state_dict(seed=0, jitter=False)
    lengths = cfg["lengths"][:nclips]
    b = S.make_batch(lengths, n_steps=nsteps, seed=cfg["seed"], infill=cfg["infill"])
    t0 = time.perf_counter()
    with torch.inference_mode():
        O.solve_euler(sd, b.z, b.t_span, b.mu, b.mask, b.spks)
    dt = time.perf_counter() - t0
    return sum(lengths) / (dt * cfg["n"] / nsteps), dt


def _sample_text(cfg, nclips, nsteps, threads):
    scaled = "" if nsteps == cfg["n"] else f" (time scaled x{cfg['n'] / nsteps:g} to the config's {cfg['n']} steps)"
    return (f"first {nclips} clip(s) of the batch ({sum(cfg['lengths'][:nclips])} valid frames) x {nsteps} Euler steps{scaled}, "
            f"oracle port of the reference torch CPU fp32 path, {threads} threads")


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (its torch modules' arithmetic, restated in
    oracle/cfm_oracle.py and pinned on goldens the unmodified modules produced) on the box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = "C2" if args.config == "C4" else args.config      # (C4 decodes C2-shaped clips)
    cfg = CONFIGS[name]
    threads = os.cpu_count() or 1
    nclips, nsteps = cfg["ref"]
    for _ in range(args.warmup):
        _cpu_decode_sample(cfg, threads, nclips, nsteps)
    rates, times = [], []
    for _ in range(args.steps):
        r, dt = _cpu_decode_sample(cfg, threads, nclips, nsteps)
        rates.append(r); times.append(dt)
    frames = sum(cfg["lengths"][:nclips])
    value = args.steps * frames / (sum(times) * cfg["n"] / nsteps)
    workload = C4["workload"] if args.config == "C4" else cfg["workload"]
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / args.steps,
            "higher_is_better": True, "scaling": "strong" if args.config == "C4" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "device": "host CPU", "clips_per_step": nclips, "euler_steps_per_step": nsteps,
                       "ranks_running": 1},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": _sample_text(cfg, nclips, nsteps, threads) + " per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def _dist_setup():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: serenade_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    return dist, world, rank, local, dev, barrier, max_over_ranks


def _make_cfm(S, sd, precision="bf16"):
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512, precision=precision)
    cfm.load_state_dict(sd, strict=True)
    return cfm


def _timed_solves(cfm, dargs, flush, n, warm):
    """ms per solve over n solves (CUDA events around each, L2 flushed in between, untimed)."""
    for _ in range(warm):
        cfm.solve_euler(*dargs)
    torch.cuda.synchronize()
    evs = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        x = cfm.solve_euler(*dargs)
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / n, x


def _gpu_reference(sd, batch, dev, frames, reps=2):
    """The reference modules' arithmetic (oracle/cfm_oracle.py functions) in eager PyTorch on this GPU: fp32 with TF32
    tensor cores (torch's default for cuDNN convs + allow_tf32 matmul) and bf16 autocast.  Outside the product path."""
    from oracle import cfm_oracle as O          # reference leg: allowed to execute oracle/
    sd_dev = {k: v.to(dev) for k, v in sd.items()}
    a = [batch.z.to(dev), batch.t_span.to(dev), batch.mu.to(dev), batch.mask.to(dev), batch.spks.to(dev)]
    out = {"what": "oracle/cfm_oracle.py (restatement of the reference torch modules) in eager PyTorch %s on the same GPU"
                   % torch.__version__}
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    try:
        for mode in ("tf32", "bf16_autocast"):
            try:
                with torch.inference_mode(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode != "tf32")):
                    O.solve_euler(sd_dev, *a)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(reps):
                        O.solve_euler(sd_dev, *a)
                    e1.record()
                    torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                out[mode] = {"ms_per_step": ms, "value": frames / ms * 1e3, "unit": UNIT}
            except torch.OutOfMemoryError:
                out[mode] = {"error": "out of memory"}
            torch.cuda.empty_cache()
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    return out


def run_ours(args):
    import serenade_b200 as S
    dist, world, rank, local, dev, barrier, max_over_ranks = _dist_setup()
    cfg = CONFIGS[args.config]
    lengths, n_euler = cfg["lengths"], cfg["n"]
    B, T = len(lengths), max(lengths)
    frames_per_step = sum(lengths)                          # valid frames: padded ones earn nothing

    sd = S.make_state_dict(seed=0, jitter=False)        # the reference's own random init (decoder.py:366-382)
    cfm = _make_cfm(S, sd)
    batch = S.make_batch(lengths, n_steps=n_euler, seed=cfg["seed"] + rank, infill=cfg["infill"])
    # (t_span stays on the host: srn_solve_euler takes the grid from host memory, a device tensor would be read back)
    dargs = [batch.z.to(dev), batch.t_span, batch.mu.to(dev), batch.mask.to(dev), batch.spks.to(dev)]
    eng = cfm.estimator.engine(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    # ---------------------------------------------------------------- device-resident timing
    for _ in range(args.warmup):
        x = cfm.solve_euler(*dargs)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = eng.launches
    evs = []
    for _ in range(args.steps):
        flush.zero_()                                   # L2 flush between timed iterations (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        x = cfm.solve_euler(*dargs)
        e1.record()
        evs.append((e0, e1))
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.launches - launches0
    ms_total = max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))
    value = world * frames_per_step * args.steps / (ms_total / 1e3)
    ns_free, ns_all = eng.softmax_stats()                # of the last timed solve

    # ---------------------------------------------------------------- end to end through the host-buffer C-ABI call
    hz, hmu, hspk = batch.z.pin_memory(), batch.mu.pin_memory(), batch.spks.pin_memory()
    hout = torch.empty_like(batch.z).pin_memory()
    for _ in range(max(1, min(args.warmup, 2))):
        cfm.decode_host(hz, hmu, batch.mask, hspk, batch.t_span, out=hout, device=dev)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cfm.decode_host(hz, hmu, batch.mask, hspk, batch.t_span, out=hout, device=dev)   # syncs the stream itself
    torch.cuda.synchronize()
    e2e_value = world * frames_per_step * args.steps / max_over_ranks(time.perf_counter() - t0)
    h2d = batch.z.numel() * 4 + batch.mu.numel() * 4 + B * T + batch.spks.numel() * 4
    d2h = batch.z.numel() * 4

    # reporting-only gather of the output mels over NCCL (outside every timed region)
    if world > 1:
        gathered = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(gathered, x)
        finite = all(bool(torch.isfinite(g).all()) for g in gathered)
    else:
        finite = bool(torch.isfinite(x).all())

    if rank == 0 and args.core:
        ms_step = ms_total / args.steps
        print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": ms_step, "core_legs_only": True,
                          "e2e": {"value": e2e_value, "unit": UNIT}, "gpu_launches": launches,
                          "config": {"workload": cfg["workload"]}}), flush=True)
    elif rank == 0:
        peaks = _peaks()
        flops_step = sum(S.synth.estimator_flops(L) for L in lengths) * n_euler
        ms_step = ms_total / args.steps
        achieved = flops_step / (ms_step / 1e3) / 1e12
        few = max(1, min(args.steps, 3))
        # the same solve with the shift-free softmax path switched off (every entry takes the row-max pass): what a
        # checkpoint whose logits exceed the bound pays
        eng.set_option("noshift", 0)
        ms_shift, _ = _timed_solves(cfm, dargs, flush, few, 1)
        eng.set_option("noshift", 1)
        # per-kernel view (isolated CUDA-event timing of every op of one estimator call, hot L2)
        ops = eng.profile_ops(B, T, reps=3)
        gemm = [o for o in ops if o["kernel"].startswith("gemm")]
        rowk = [o for o in ops if not o["kernel"].startswith("gemm")]
        gemm_ms, row_ms = sum(o["ms"] for o in gemm), sum(o["ms"] for o in rowk)
        gemm_alg, gemm_iss = sum(o["flops"] for o in gemm), sum(o.get("flops_issued", o["flops"]) for o in gemm)
        row_by = sum(o["bytes"] for o in rowk)
        # this library's TF32-class mode (the mode that meets the 1e-3 per-step parity gate) on the same workload
        eng.close()
        cfm.estimator._engine = None
        del cfm
        torch.cuda.empty_cache()
        cfm32 = _make_cfm(S, sd, "tf32")
        ms_tf32, _ = _timed_solves(cfm32, dargs, flush, few, 2)
        cfm32.estimator._engine.close()
        cfm32.estimator._engine = None
        del cfm32
        torch.cuda.empty_cache()
        gpu_ref = _gpu_reference(sd, batch, dev, frames_per_step)
        best_ref = min((v["ms_per_step"] for v in gpu_ref.values() if isinstance(v, dict) and "ms_per_step" in v),
                       default=None)
        gpu_ref["speedup_over_best_eager"] = best_ref / ms_step if best_ref else None
        traffic, traffic_src, traffic_warm = _traffic(args.config)
        threads = os.cpu_count() or 1
        cpu_fps, cpu_s = _cpu_decode_sample(cfg, threads, *cfg["cpu"])
        tf = lambda fl, ms: fl / (ms / 1e3) / 1e12 if ms else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": cfg["workload"], "batch_per_gpu": B, "frames": T, "valid_frames_per_step": frames_per_step,
                       "euler_steps": n_euler,
                       "sharding": f"utterance-sharded, {world} rank(s), no collective in the timed region",
                       "l2": "256 MiB buffer rewritten between timed iterations (untimed)",
                       "rtf": (ms_step / 1e3) / (frames_per_step / 100.0), "outputs_finite": finite},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tf_sustained"], "traffic": traffic, "traffic_source": traffic_src,
                         "traffic_warm_caches": traffic_warm,
                         "launch": f"one CUDA-graph launch = one step ({n_euler} Euler steps x {B} clips); gemm_kernel is the "
                                   "dominant kernel (its share of the step is gemm_share_of_step); FLOPs are SURVEY 8d's "
                                   "F(L) summed over the clips' VALID lengths",
                         "gemm_share_of_step": gemm_ms / (gemm_ms + row_ms) if gemm_ms else None,
                         "peak_source": peaks["source"] + " (sustained bf16; whole Euler solve, SURVEY 8d F(L))",
                         "gemm_kernel": {"ms_per_estimator_call": gemm_ms,
                                         "tflops_algorithmic_isolated": tf(gemm_alg, gemm_ms),
                                         "tflops_issued_isolated": tf(gemm_iss, gemm_ms),
                                         "frac_algorithmic": tf(gemm_alg, gemm_ms) / peaks["tf_burst"] if gemm_ms else None,
                                         "frac_issued": tf(gemm_iss, gemm_ms) / peaks["tf_burst"] if gemm_ms else None,
                                         "peak": peaks["tf_burst"], "peak_kind": "burst bf16 (kernels timed alone, hot L2)",
                                         "note": "algorithmic = no 242->256 channel padding, nothing for the row-max pass; "
                                                 "issued = every K block the tensor cores ran"},
                         "row_kernels": {"gbs_isolated": row_by / (row_ms / 1e3) / 1e9 if row_ms else None,
                                         "ms_per_estimator_call": row_ms,
                                         "frac_of_hbm_peak": (row_by / (row_ms / 1e3) / 1e9) / peaks["hbm"] if row_ms else None}},
            "softmax_path": {"shift_free_entries": ns_free, "entries": ns_all,
                             "note": "entries = (clip, head) x attention block x Euler step of one timed solve; shift-free = "
                                     "score bound <= 64 proven from the Q / K row norms, no row-max pass",
                             "all_shifted": {"ms_per_step": ms_shift, "value": frames_per_step / ms_shift * 1e3, "unit": UNIT,
                                             "steps": few}},
            "tf32_mode": {"ms_per_step": ms_tf32, "value": frames_per_step / ms_tf32 * 1e3, "unit": UNIT, "steps": few,
                          "note": "precision='tf32': rounded-to-nearest TF32 activations x hi+lo split weights, the mode "
                                  "that meets the 1e-3 per-step parity gate (1 GPU, rank 0)"},
            "gpu_reference": gpu_ref,
            "cpu_baseline": {"value": cpu_fps, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": _sample_text(cfg, *cfg["cpu"], threads) + f", {cpu_s:.1f} s"},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_c4(args):
    """BASELINE config 4: 1024 clips as C2's, batches of 16, clip i on rank i mod N (strong scaling).
    A step = one pass over all clips.  `value`: the ranks' decode loops over batches already resident in HBM; `e2e`: the
    pipelined host loop (serenade_b200/sharding.py: decode_sharded_pipelined) from clips in host memory to mels in host
    memory - padding, H2D and D2H inside the timed region, overlapped with the solves.  No collective in either."""
    import serenade_b200 as S
    from serenade_b200 import sharding
    dist, world, rank, local, dev, barrier, max_over_ranks = _dist_setup()
    n_clips, T, n_euler, bs = C4["clips"], C4["frames"], C4["n"], C4["batch"]
    n_clips = int(os.environ.get("SRN_C4_CLIPS", n_clips))
    sd = S.make_state_dict(seed=0, jitter=False)
    cfm = _make_cfm(S, sd)
    eng = cfm.estimator.engine(dev)
    lengths = [T] * n_clips
    mine = sharding.assign_clips(lengths, world)[rank]          # equal lengths: i mod N
    mine_set = set(mine)

    def make_clip(i):
        b = S.make_batch([T], n_steps=n_euler, seed=100000 + i, infill=(0.25, 0.75))
        return dict(mu=b.mu[0], z=b.z[0], spk=b.spks[0])

    placeholder = dict(mu=torch.empty(162, T, device="meta"), z=None, spk=None)      # (only the length is read)
    clips = [make_clip(i) if i in mine_set else placeholder for i in range(n_clips)]
    t_span = torch.linspace(0, 1, n_euler + 1)
    batches = sharding.make_batches(mine, lengths, max_batch=bs)
    # device-resident copies of this rank's batches (value leg)
    dev_batches = []
    for idx in batches:
        z, mu, mask, spk = sharding.pad_batch(clips, idx)
        dev_batches.append((z.to(dev), t_span, mu.to(dev), mask.to(dev), spk.to(dev)))
    frames_total = n_clips * T

    def resident_pass():
        for a in dev_batches:
            x = cfm.solve_euler(*a)
        return x

    for _ in range(args.warmup):
        resident_pass()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = eng.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        resident_pass()
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.launches - launches0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    value = frames_total * args.steps / (ms_total / 1e3)
    del dev_batches
    torch.cuda.empty_cache()

    # ---- end to end: host clips -> host mels through the pipelined loop
    mels = {}
    for _ in range(max(1, min(args.warmup, 2))):
        sharding.decode_sharded_pipelined(cfm, clips, t_span, rank=rank, world_size=world, max_batch=bs, device=dev, out=mels)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sharding.decode_sharded_pipelined(cfm, clips, t_span, rank=rank, world_size=world, max_batch=bs, device=dev, out=mels)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = frames_total * args.steps / e2e_s
    per_clip_in = (80 + 162) * T * 4 + T + 256 * 4
    # reporting only: a checksum over all decoded mels (must not depend on the number of ranks)
    sums = torch.zeros(n_clips, dtype=torch.float64, device=dev)
    for i, m in mels.items():
        sums[i] = m.double().sum()
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    digest = float(sums.sum())
    finite = bool(torch.isfinite(sums).all())
    if rank == 0:
        peaks = _peaks()
        flops_step = n_clips * n_euler * S.synth.estimator_flops(T)
        ms_step = ms_total / args.steps
        achieved = flops_step / (ms_step / 1e3) / 1e12 / world           # per GPU
        threads = os.cpu_count() or 1
        cfg = CONFIGS["C2"]
        if args.core:      # (scaling sweeps: the CPU leg is the same at every N)
            cpu_fps, cpu_s = None, 0.0
        else:
            cpu_fps, cpu_s = _cpu_decode_sample(cfg, threads, *cfg["cpu"])
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": C4["workload"], "clips": n_clips, "frames": T, "euler_steps": n_euler, "batch": bs,
                       "clips_per_rank": len(mine),
                       "sharding": f"clip i -> rank i mod {world}; no collective in the timed region",
                       "l2": "every batch's working set (about 1 GB of activations) exceeds the 126 MB L2; consecutive "
                             "batches are different clips",
                       "checksum_all_mels": digest, "outputs_finite": finite},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n_clips * per_clip_in,
                    "d2h_bytes_per_step": n_clips * 80 * T * 4, "ms_per_step": 1e3 * e2e_s / args.steps,
                    "path": "sharding.decode_sharded_pipelined: pinned double-buffered staging, H2D / D2H on copy streams "
                            "overlapped with the solves"},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tf_sustained"], "traffic": None,
                         "launch": "per GPU; one CUDA-graph launch per batch of 16 clips (10 Euler steps)",
                         "peak_source": peaks["source"] + " (sustained bf16)"},
            "cpu_baseline": {"value": cpu_fps, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": _sample_text(cfg, *cfg["cpu"], threads) + f", {cpu_s:.1f} s"},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--core", action="store_true",
                    help="only the two timed legs (device-resident value, host-buffer e2e): the command profiled under ncu "
                         "for the launch list, without the eager-PyTorch / TF32-mode / CPU legs' kernels")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    args.warmup = max(args.warmup, 3)
    if args.config == "C4":
        run_c4(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
