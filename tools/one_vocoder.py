"""One vocoder decode at (B, T) (for ncu captures / launch lists).  python tools/one_vocoder.py B T [calls] [precision]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    B, T = int(sys.argv[1]), int(sys.argv[2])
    calls = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    prec = sys.argv[4] if len(sys.argv) > 4 else "bf16"
    dev = torch.device("cuda", 0)
    g = S.HiFiGANGenerator(precision=prec)
    g.load_state_dict(S.make_vocoder_state_dict(seed=0), strict=True)
    mel = S.make_vocoder_input(T, batch=B, seed=40).to(dev)
    for _ in range(calls):
        w = g.decode(mel)
    torch.cuda.synchronize()
    print("ok", tuple(w.shape), float(w.abs().mean()))


if __name__ == "__main__":
    main()
