"""BASELINE.json configs C1, C3, C5 (per-GPU share) on one GPU: solve time, frames/s, algorithmic TFLOP/s, and the
size-independent properties (finite, zero velocity at padded frames => x keeps its noise there).

    python tools/run_configs.py [out.txt]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else None
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    g = torch.Generator().manual_seed(1234 + 2)
    c3_lengths = sorted(torch.randint(300, 2001, (64,), generator=g).tolist(), reverse=True)
    configs = [
        ("C1 B=1 T=500 N=10", [500], 10, 1234),
        ("C2 B=16 T=1000 N=10", [1000] * 16, 10, 1235),
        ("C3 B=64 T=300..2000 N=32", c3_lengths, 32, 1236),
        ("C5 (per GPU) B=8 T=6000 N=50", [6000] * 8, 50, 1238),
    ]
    lines = []
    for name, lengths, n, seed in configs:
        b = S.make_batch(lengths, n_steps=n, seed=seed, infill=(0.25, 0.75))
        args = [b.z.to(dev), b.t_span.to(dev), b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
        x = cfm.solve_euler(*args)          # builds the plan + graph
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3 if sum(lengths) < 40000 else 1
        e0.record()
        for _ in range(reps):
            x = cfm.solve_euler(*args)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        fl = sum(S.synth.estimator_flops(L) for L in lengths) * n
        pad = ~b.mask.expand_as(b.z)
        keeps_noise = bool(torch.equal(x.cpu()[pad], b.z[pad]))
        finite = bool(torch.isfinite(x).all())
        line = (f"{name}: {ms:9.2f} ms/solve  {sum(lengths) / ms * 1e3:10.0f} valid frames/s  "
                f"{fl / ms / 1e9:7.1f} TFLOP/s algorithmic (valid frames)  finite={finite}  padded frames keep noise={keeps_noise}  "
                f"mem={torch.cuda.max_memory_allocated() / 2**30:.1f} GiB torch + plan")
        print(line, flush=True)
        lines.append(line)
        del x, args
        cfm.estimator._engine.close()
        cfm.estimator._engine = None
        torch.cuda.empty_cache()
    if out:
        open(out, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
