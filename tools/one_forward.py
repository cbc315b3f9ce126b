"""One eager estimator forward at (B, T) (for ncu captures).  python tools/one_forward.py B T [calls] [precision]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    B, T = int(sys.argv[1]), int(sys.argv[2])
    calls = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    precision = sys.argv[4] if len(sys.argv) > 4 else "bf16"
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512, precision=precision)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    b = S.make_batch([T] * B, n_steps=2, seed=1235, infill=(0.25, 0.75))
    x, mu, mask, spk = b.z.to(dev), b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)
    for i in range(calls):
        v = cfm.estimator(x, mask, mu, torch.tensor(0.3), spk)
        torch.cuda.synchronize()
        print("call", i, "launches so far", cfm.estimator.engine(dev).launches, "finite", bool(torch.isfinite(v).all()))


if __name__ == "__main__":
    main()
