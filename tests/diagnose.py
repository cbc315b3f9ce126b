"""GPU-box diagnostic: stage-by-stage error table vs the oracle (never stops at the first bad stage) and a
quick CUDA-event timing of the solve.  Writes gpurun_out/diagnose.txt.   python tests/diagnose.py [--time]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # (this file lives there: it runs the oracle as the checker)

import torch  # noqa: E402

import serenade_b200 as S  # noqa: E402
from stagewise import compare, format_report  # noqa: E402


def main():
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out = open(os.path.join(ROOT, "gpurun_out", "diagnose.txt"), "w")

    def emit(*a):
        print(*a, flush=True)
        print(*a, file=out, flush=True)

    precision = "tf32" if "--tf32" in sys.argv else "bf16"
    dev = torch.device("cuda", 0)
    emit(torch.cuda.get_device_name(0), "precision", precision)
    sd = S.make_state_dict(seed=0, jitter=True)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512, precision=precision)
    cfm.load_state_dict(sd, strict=True)
    for lengths in ([200, 150], [131], [64]):
        b = S.make_batch(lengths, n_steps=2, seed=21)
        try:
            rows, v, v_ref = compare(cfm, sd, b, 0.37, fine=True)
            emit(f"--- lengths {lengths}")
            emit(format_report(rows))
        except Exception as e:  # keep going: the next shape may still be informative
            emit(f"--- lengths {lengths}: FAILED {type(e).__name__}: {e}")
            break
    if "--time" in sys.argv:
        for (B, T, n) in ((1, 500, 10), (16, 1000, 10)):
            b = S.make_batch([T] * B, n_steps=n, seed=1235)
            args = [b.z.to(dev), b.t_span.to(dev), b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
            for use_graph in (False, True):
                cfm.estimator.engine(dev).set_use_graph(use_graph)
                for _ in range(2):
                    cfm.solve_euler(*args)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                reps = 3
                for _ in range(reps):
                    x = cfm.solve_euler(*args)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                fl = sum(S.synth.estimator_flops(T) for _ in range(B)) * n
                emit(f"B={B} T={T} steps={n} graph={use_graph}: {ms:.2f} ms/solve, {B * T / ms * 1e3:.0f} frames/s, "
                     f"{fl / ms / 1e9:.1f} TFLOP/s algorithmic, finite={bool(torch.isfinite(x).all())}")
    out.close()


if __name__ == "__main__":
    main()
