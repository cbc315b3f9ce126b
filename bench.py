#!/usr/bin/env python
"""Benchmark of the CFM mel-infilling decode hot path (BASELINE.json metric: CFM-decoded mel frames/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one full decode of one batch of BASELINE config 2 (ssc1 decoder, random-init, batch 16,
10 s clips = 1000 frames, 50 % infill, 10 Euler steps, bf16 tensor-core math) = 16 000 mel frames.
One process per GPU (torchrun for N > 1); utterances are sharded by rank with NO collective in the timed
region (weak scaling: every rank decodes its own batch); an NCCL all_gather of the output mels runs
after the timed region for reporting only.

Printed JSON (rank 0, one line): value = whole-job frames/s with inputs resident in HBM (CUDA events, max
over ranks); e2e = the same through the host-buffer C-ABI call (H2D of z/mu/mask/spk and D2H of the mel
inside the timed region); roofline = algorithmic tensor FLOPs of the step / time vs the measured bf16
peak; cpu_baseline = the CPU oracle (a port of the reference's torch CPU path) on a bounded sample.
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
WORKLOAD = "C2: ssc1 decoder random-init, batch 16 x 1000 frames (10 s clips), 50% infill, 10 Euler steps"
B, T, N_EULER = 16, 1000, 10
CPU_SAMPLE_B = 16       # cpu_baseline leg of our arm: the whole 16-clip batch once (about 15 s on the box's cores)
REF_SAMPLE_B = 4        # --impl reference: clips per timed step (bounded so that K + W steps end within minutes)


def _dram_traffic_per_step():
    """DRAM bytes (read + write) of one timed step, summed over its kernels, from the committed ncu launch list
    (profiles/r01_traffic.json, written by tools/launch_summary.py from the same bench command under ncu)."""
    p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    try:
        d = json.load(open(p))
        return float(d["dram_bytes_per_step"]), d.get("source", p)
    except Exception:
        return None, None


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


def _cpu_decode_sample(threads: int, nclips: int = CPU_SAMPLE_B):
    """Times the CPU oracle (port of the reference's torch CPU fp32 path) on CPU_SAMPLE_B clips of the workload."""
    import serenade_b200 as S
    from oracle import cfm_oracle as O          # CPU baseline leg: allowed to execute oracle/
    torch.set_num_threads(threads)
    sd = S.make_state_dict(seed=0, jitter=False)
    b = S.make_batch([T] * nclips, n_steps=N_EULER, seed=1235)
    t0 = time.perf_counter()
    with torch.inference_mode():
        O.solve_euler(sd, b.z, b.t_span, b.mu, b.mask, b.spks)
    dt = time.perf_counter() - t0
    return nclips * T / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = (f"{REF_SAMPLE_B} clip(s) x {T} frames x {N_EULER} Euler steps per step (of the 16-clip batch), "
              f"oracle port of the reference torch CPU fp32 path, {threads} threads")
    for _ in range(args.warmup):
        _cpu_decode_sample(threads, REF_SAMPLE_B)
    times = []
    for _ in range(args.steps):
        times.append(_cpu_decode_sample(threads, REF_SAMPLE_B)[1])
    total = sum(times)
    value = args.steps * REF_SAMPLE_B * T / total
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "device": "host CPU"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist
    import serenade_b200 as S

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

    sd = S.make_state_dict(seed=0, jitter=False)        # the reference's own random init (decoder.py:366-382)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512, precision="bf16")
    cfm.load_state_dict(sd, strict=True)
    batch = S.make_batch([T] * B, n_steps=N_EULER, seed=1235 + rank, infill=(0.25, 0.75))
    dargs = [batch.z.to(dev), batch.t_span.to(dev), batch.mu.to(dev), batch.mask.to(dev), batch.spks.to(dev)]
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
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    tt = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_total = float(tt.item())
    frames_per_step = B * T
    value = world * frames_per_step * args.steps / (ms_total / 1e3)

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
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * frames_per_step * args.steps / float(te.item())
    h2d = batch.z.numel() * 4 + batch.mu.numel() * 4 + B * T + batch.spks.numel() * 4
    d2h = batch.z.numel() * 4

    # reporting-only gather of the output mels over NCCL (outside every timed region)
    if world > 1:
        gathered = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(gathered, x)
        finite = all(bool(torch.isfinite(g).all()) for g in gathered)
    else:
        finite = bool(torch.isfinite(x).all())

    if rank == 0:
        peaks = _peaks()
        flops_step = sum(S.synth.estimator_flops(T) for _ in range(B)) * N_EULER
        ms_step = ms_total / args.steps
        achieved = flops_step / (ms_step / 1e3) / 1e12
        # per-kernel view (isolated CUDA-event timing of every op of one estimator call, hot L2)
        ops = eng.profile_ops(B, T, reps=3)
        gemm_ms = sum(o["ms"] for o in ops if o["kernel"].startswith("gemm"))
        gemm_fl = sum(o["flops"] for o in ops if o["kernel"].startswith("gemm"))
        row_ms = sum(o["ms"] for o in ops if not o["kernel"].startswith("gemm"))
        row_by = sum(o["bytes"] for o in ops if not o["kernel"].startswith("gemm"))
        traffic, traffic_src = _dram_traffic_per_step()
        threads = os.cpu_count() or 1
        cpu_fps, cpu_s = _cpu_decode_sample(threads)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": B, "frames": T, "euler_steps": N_EULER,
                       "sharding": f"utterance-sharded, {world} rank(s), no collective in the timed region",
                       "l2": "256 MiB buffer rewritten between timed iterations (untimed)",
                       "rtf": (ms_step / 1e3) / (frames_per_step / 100.0), "outputs_finite": finite},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tf_sustained"], "traffic": traffic, "traffic_source": traffic_src,
                         "launch": "one CUDA-graph launch = one step (10 Euler steps x 16 clips); gemm_kernel is the "
                                   "dominant kernel (its share of the step is gemm_share_of_step)",
                         "gemm_share_of_step": gemm_ms / (gemm_ms + row_ms) if gemm_ms else None,
                         "peak_source": peaks["source"] + " (sustained bf16; whole Euler solve, SURVEY 8d F(L))",
                         "gemm_kernel": {"tflops_isolated": gemm_fl / (gemm_ms / 1e3) / 1e12 if gemm_ms else None,
                                         "ms_per_estimator_call": gemm_ms,
                                         "frac_of_burst_peak": (gemm_fl / (gemm_ms / 1e3) / 1e12) / peaks["tf_burst"]
                                         if gemm_ms else None},
                         "row_kernels": {"gbs_isolated": row_by / (row_ms / 1e3) / 1e9 if row_ms else None,
                                         "ms_per_estimator_call": row_ms,
                                         "frac_of_hbm_peak": (row_by / (row_ms / 1e3) / 1e9) / peaks["hbm"]
                                         if row_ms else None}},
            "cpu_baseline": {"value": cpu_fps, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{CPU_SAMPLE_B} clips x {T} frames x {N_EULER} Euler steps, oracle port of the "
                                       f"reference torch CPU fp32 path, {threads} threads, {cpu_s:.1f} s"},
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
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        args.warmup = max(args.warmup, 3)
        run_ours(args)


if __name__ == "__main__":
    main()
