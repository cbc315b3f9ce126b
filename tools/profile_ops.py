"""Per-op CUDA-event timing table of one estimator call (srn_profile_ops).  python tools/profile_ops.py B T [out] [reps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    B, T = int(sys.argv[1]), int(sys.argv[2])
    out = sys.argv[3] if len(sys.argv) > 3 else None
    reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    ops = cfm.estimator.engine(dev).profile_ops(B, T, reps=reps)
    tot = sum(o["ms"] for o in ops)
    lines = [f"# per-op timing, one estimator call, B={B} T={T}, total {tot:.3f} ms (isolated, hot L2, CUDA events, {reps} launches per op)"]
    agg = {}
    for o in ops:
        tf = o["flops"] / o["ms"] / 1e9 if o["flops"] else 0.0
        gb = o["bytes"] / o["ms"] / 1e6 if o["bytes"] else 0.0
        lines.append(f"{o['op']:22s} {o['kernel']:14s} {o['ms']*1e3:9.1f} us  {tf:8.1f} TFLOP/s {gb:8.1f} GB/s")
        short = o["op"].split(".")[-1] + ":" + o["kernel"]
        a = agg.setdefault(short, [0.0, 0.0, 0.0, 0])
        a[0] += o["ms"]; a[1] += o["flops"]; a[2] += o["bytes"]; a[3] += 1
    lines.append("# aggregated by op type")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        tf = a[1] / a[0] / 1e9 if a[1] else 0.0
        gb = a[2] / a[0] / 1e6 if a[2] else 0.0
        lines.append(f"{k:34s} x{a[3]:2d} {a[0]*1e3:9.1f} us {100*a[0]/tot:5.1f}%  {tf:8.1f} TFLOP/s {gb:8.1f} GB/s")
    text = "\n".join(lines)
    print(text)
    if out:
        os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
        open(out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
