"""CUDA-event time of one CFM solve at (B, T, n_steps), bf16 mode.  python tools/time_solve.py B T N [reps] [precision]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    B, T, n = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    precision = sys.argv[5] if len(sys.argv) > 5 else "bf16"
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512, precision=precision)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    b = S.make_batch([T] * B, n_steps=n, seed=1238, infill=(0.25, 0.75))
    args = [b.z.to(dev), b.t_span, b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
    cfm.solve_euler(*args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        x = cfm.solve_euler(*args)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    fl = B * n * S.synth.estimator_flops(T)
    print(f"B={B} T={T} steps={n} {precision}: {ms:.2f} ms/solve, {B * T / ms * 1e3:.0f} frames/s, {fl / ms / 1e9:.1f} TFLOP/s algorithmic, "
          f"finite={bool(torch.isfinite(x).all())}", flush=True)


if __name__ == "__main__":
    main()
