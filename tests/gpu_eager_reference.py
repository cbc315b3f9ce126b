"""The oracle's restatement executed by eager PyTorch ON THE GPU (SURVEY.md 8d "GPU reference beside it").

Two uses, both test infrastructure:
  * full-size parity: the CPU oracle needs minutes for BASELINE config 2 (16 x 1000 frames, 10 steps); the same
    functions in strict fp32 on the GPU (TF32 off for matmul and cuDNN) finish in a second, so the CUDA path can be
    compared with them at the full benchmark sizes (tests/test_gpu_parity.py);
  * "the kernels to beat on the same box": the reference ships no Blackwell kernels, its modules in eager PyTorch
    (fp32, fp32 with TF32 tensor cores, bf16 autocast) are what a user gets by moving the reference to a B200.

    python tests/gpu_eager_reference.py [out.txt]      # timing table for C1, C2, C3, C5 + parity at those sizes
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import cfm_oracle as O  # noqa: E402


def strict_fp32(on: bool):
    torch.backends.cuda.matmul.allow_tf32 = not on
    torch.backends.cudnn.allow_tf32 = not on


def to_dev(sd, dev):
    return {k: v.to(dev) for k, v in sd.items()}


def eager_solve(sd_dev, batch, dev, mode="fp32", return_all=False):
    """mode: 'fp32' (no tensor-core shortcuts), 'tf32' (torch's allow_tf32), 'bf16' (autocast)."""
    args = dict(x=batch.z.to(dev), t_span=batch.t_span.to(dev), mu=batch.mu.to(dev), mask=batch.mask.to(dev),
                spk=batch.spks.to(dev))
    strict_fp32(mode == "fp32")
    try:
        with torch.inference_mode(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16")):
            return O.solve_euler(sd_dev, args["x"], args["t_span"], args["mu"], args["mask"], args["spk"],
                                 return_all=return_all)
    finally:
        strict_fp32(False)


def eager_velocity(sd_dev, batch, dev, x, t):
    strict_fp32(True)
    try:
        with torch.inference_mode():
            return O.estimator_forward(sd_dev, x, batch.mask.to(dev), batch.mu.to(dev), torch.as_tensor(t, device=dev),
                                       batch.spks.to(dev))
    finally:
        strict_fp32(False)


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm())


def _time(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def main():
    import serenade_b200 as S
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    dev = torch.device("cuda", 0)
    sd = S.make_state_dict(0, False)
    sd_dev = to_dev(sd, dev)
    g = torch.Generator().manual_seed(1234 + 2)
    c3_lengths = sorted(torch.randint(300, 2001, (64,), generator=g).tolist(), reverse=True)
    configs = [
        ("C1 B=1 T=500 N=10", [500], 10, 1234, 3),
        ("C2 B=16 T=1000 N=10", [1000] * 16, 10, 1235, 2),
        ("C3 B=64 T=300..2000 N=32", c3_lengths, 32, 1236, 1),
        ("C5 (per GPU) B=8 T=6000 N=50", [6000] * 8, 50, 1238, 1),
    ]
    only = os.environ.get("SRN_EAGER_CONFIGS")
    lines = ["# eager PyTorch on the same B200 (oracle/cfm_oracle.py functions, torch %s) vs this library; "
             "ms per solve, valid frames/s; rel-L2 of the final mel against eager strict fp32" % torch.__version__]
    for name, lengths, n, seed, reps in configs:
        if only and name.split()[0] not in only.split(","):
            continue
        b = S.make_batch(lengths, n_steps=n, seed=seed, infill=(0.25, 0.75))
        frames = sum(lengths)
        row = [name]
        ref = None
        for mode in ("fp32", "tf32", "bf16"):
            try:
                ms, x = _time(lambda: eager_solve(sd_dev, b, dev, mode), reps)
            except torch.OutOfMemoryError:
                row.append(f"eager {mode}: out of memory")
                torch.cuda.empty_cache()
                continue
            if mode == "fp32":
                ref = x.clone()
                row.append(f"eager fp32 {ms:9.1f} ms {frames / ms * 1e3:9.0f} f/s")
            else:
                row.append(f"eager {mode} {ms:9.1f} ms {frames / ms * 1e3:9.0f} f/s err {rel(x, ref):.1e}")
            del x
            torch.cuda.empty_cache()
        for prec in ("bf16", "tf32"):
            cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                        decoder_attention_head_dim=512, precision=prec)
            cfm.load_state_dict(sd, strict=True)
            args = [b.z.to(dev), b.t_span.to(dev), b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
            ms, x = _time(lambda: cfm.solve_euler(*args), reps)
            err = rel(x, ref) if ref is not None else float("nan")
            row.append(f"this library {prec} {ms:9.1f} ms {frames / ms * 1e3:9.0f} f/s err {err:.1e}")
            cfm.estimator._engine.close()
            cfm.estimator._engine = None
            del cfm, x, args
            torch.cuda.empty_cache()
        line = " | ".join(row)
        print(line, flush=True)
        lines.append(line)
    if out_path:
        with open(out_path, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
