"""BASELINE config 4: utterance-sharded decode of 1024 synthetic clips (each as config 2's clip: T = 1000, 10 Euler
steps, 50 % infill), batches of 16, one process per GPU, NO collective inside the timed region; and config 5's
per-GPU share (B = 8, T = 6000, 50 steps) on every rank at once.

    python tools/run_c4.py [out.txt]                                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/run_c4.py [out.txt]

Timed region = the slowest rank's decode loop (barrier + synchronize on both sides, CUDA events, max over ranks),
including the host-side padding of each batch, its host->device copy and the device->host copy of the decoded
mels.  After it, the mels are gathered (NCCL all_gather of a checksum per clip + gather_mels on a sample) for
reporting only: the per-clip checksums must not depend on the number of GPUs.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import serenade_b200 as S  # noqa: E402
from serenade_b200 import sharding  # noqa: E402


def make_clip(i, T=1000):
    b = S.make_batch([T], n_steps=10, seed=100000 + i, infill=(0.25, 0.75))
    return dict(mu=b.mu[0], z=b.z[0], spk=b.spks[0])


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    n_clips = int(os.environ.get("SRN_C4_CLIPS", "1024"))
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    t_span = torch.linspace(0, 1, 11)

    # ---- C4: every rank builds only its own clips (assignment is deterministic and identical on all ranks)
    lengths = [1000] * n_clips
    mine = set(sharding.assign_clips(lengths, world)[rank])
    clips = [make_clip(i) if i in mine else dict(mu=torch.empty(162, 1000, device="meta"), z=None, spk=None)
             for i in range(n_clips)]
    warm = S.make_batch([1000] * 16, n_steps=10, seed=1)
    cfm.solve_euler(warm.z.to(dev), t_span.to(dev), warm.mu.to(dev), warm.mask.to(dev), warm.spks.to(dev))   # plan + graph
    t_dev = t_span.to(dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    mels = sharding.decode_sharded(cfm.solve_euler, clips, t_dev, rank=rank, world_size=world, max_batch=16, device=dev)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    c4_ms = float(ms)
    # reporting only: per-clip checksums from every rank
    sums = torch.zeros(n_clips, dtype=torch.float64, device=dev)
    for i, m in mels.items():
        sums[i] = m.double().sum()
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    digest = float(sums.sum())
    sample = sharding.gather_mels({i: m for i, m in mels.items() if i < 8}, 8)
    assert all(m is not None and m.shape == (80, 1000) for m in sample)
    del clips, mels

    # ---- C5 per-GPU share on every rank at once
    lines = []
    if os.environ.get("SRN_SKIP_C5", "0") != "1":
        b = S.make_batch([6000] * 8, n_steps=50, seed=1238 + rank, infill=(0.25, 0.75))
        args = [b.z.to(dev), b.t_span.to(dev), b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
        x = cfm.solve_euler(*args)
        barrier()
        e0.record()
        x = cfm.solve_euler(*args)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        c5_ms = float(ms)
        fl = 8 * 50 * S.synth.estimator_flops(6000)
        lines.append(f"C5 n_gpus={world}: B=8 x T=6000 x 50 steps per GPU, {c5_ms:.1f} ms (max over ranks) -> "
                     f"{world * 48000 / c5_ms * 1e3:.0f} frames/s aggregate, {fl / c5_ms / 1e9:.0f} TFLOP/s per GPU, "
                     f"finite={bool(torch.isfinite(x).all())}")
    fl4 = n_clips * 10 * S.synth.estimator_flops(1000)
    lines.insert(0, f"C4 n_gpus={world}: {n_clips} clips x 1000 frames, batches of 16, 10 steps: {c4_ms:.1f} ms (slowest rank, "
                    f"incl. host padding, H2D and D2H per batch) -> {n_clips * 1000 / c4_ms * 1e3:.0f} frames/s aggregate, "
                    f"{fl4 / c4_ms / 1e9 / world:.0f} TFLOP/s per GPU; checksum over all mels {digest:.6f}")
    if rank == 0:
        print("\n".join(lines), flush=True)
        if out_path:
            with open(out_path, "a") as f:
                f.write("\n".join(lines) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
