"""Vocoder timing on one B200 (SURVEY.md 8f-2 measurement): this library vs the oracle's functions in eager PyTorch on
the same GPU (what moving the reference vocoder to the B200 gives) vs the oracle on the host cores.

    python tests/bench_vocoder.py [out.txt]

Workloads: one 5 s clip (C1's mel: 500 frames) and 16 x 10 s (C2's mels: 16 x 1000 frames), class-default generator
(512 channels, strides 8 8 2 2 = 256 samples per frame), synthetic weights.  Times are CUDA events around repeated
calls after warm-up; the metric is mel frames per second and the real-time factor at 100 frames/s.
(Lives under tests/ because it executes oracle/ code as the baseline.)
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import vocoder_oracle as V  # noqa: E402


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


def conv_flops(sd, frames):
    """2 * MACs of one generator forward for `frames` mel frames (per clip)."""
    fl, t = 0.0, frames
    fl += 2.0 * t * sd["input_conv.weight"].numel()
    i = 0
    while f"upsamples.{i}.1.weight" in sd:
        w = sd[f"upsamples.{i}.1.weight"]
        s = w.shape[-1] // 2
        fl += 2.0 * t * w.numel()                      # every input frame meets every tap once
        t *= s
        for k, v in sd.items():
            if k.endswith(".weight") and k.startswith("blocks.") and v.shape[0] == w.shape[1]:
                fl += 2.0 * t * v.numel()
        i += 1
    fl += 2.0 * t * sd["output_conv.1.weight"].numel()
    return fl


def main():
    import serenade_b200 as S
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    dev = torch.device("cuda", 0)
    sd = S.make_vocoder_state_dict(seed=0)
    sd_dev = {k: v.to(dev) for k, v in sd.items()}
    lines = ["# HiFi-GAN generator (512 ch, strides 8 8 2 2, 13.9 M parameters), synthetic weights; ms per call, mel frames/s, "
             "RTF = time / audio seconds, rel-L2 vs eager fp32 on the GPU"]
    for name, B, T, reps in (("1 x 500 frames (5 s)", 1, 500, 5), ("16 x 1000 frames (16 x 10 s)", 16, 1000, 3)):
        mel = S.make_vocoder_input(T, batch=B, seed=40).to(dev)
        fl = conv_flops(sd, T) * B
        row = [f"{name}, {fl / 1e12:.2f} TFLOP"]
        ref = None
        for mode in ("fp32", "tf32", "bf16"):
            torch.backends.cuda.matmul.allow_tf32 = mode != "fp32"
            torch.backends.cudnn.allow_tf32 = mode != "fp32"

            def run():
                with torch.inference_mode(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16")):
                    return V.hifigan_forward(sd_dev, mel.transpose(1, 2)).squeeze(1).float()
            ms, wav = _time(run, reps)
            if mode == "fp32":
                ref = wav.clone()
                row.append(f"eager fp32 {ms:8.2f} ms")
            else:
                err = float((wav.double() - ref.double()).norm() / ref.double().norm())
                row.append(f"eager {mode} {ms:8.2f} ms err {err:.1e}")
            del wav
            torch.cuda.empty_cache()
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
        for prec in ("bf16", "tf32"):
            g = S.HiFiGANGenerator(precision=prec)
            g.load_state_dict(sd, strict=True)
            ms, wav = _time(lambda: g.decode(mel), reps)
            err = float((wav.double() - ref.double()).norm() / ref.double().norm())
            row.append(f"this library {prec} {ms:8.2f} ms = {B * T / ms * 1e3:9.0f} frames/s, RTF {ms / 1e3 / (B * T / 100):.1e}, "
                       f"{fl / ms / 1e9:6.1f} TFLOP/s, err {err:.1e}")
            g.close()
            del g, wav
            torch.cuda.empty_cache()
        if B == 1:
            torch.set_num_threads(os.cpu_count() or 1)
            c = mel.cpu().transpose(1, 2)
            with torch.inference_mode():
                V.hifigan_forward(sd, c[:, :, :50])
                t0 = time.perf_counter()
                V.hifigan_forward(sd, c)
                dt = time.perf_counter() - t0
            row.append(f"CPU oracle, {os.cpu_count()} threads {dt * 1e3:8.1f} ms")
        line = " | ".join(row)
        print(line, flush=True)
        lines.append(line)
    if out_path:
        with open(out_path, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
