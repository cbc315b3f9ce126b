"""Exploratory probe (not collected by pytest): how far the library and the CPU oracle drift apart when the attention
logits grow - W_q scaled by 4 ... 64, both precision modes.  The softmax turns into an arg-max at these scales, so operand
rounding of Q and K (2^-9 relative in bf16) moves probability mass between neighbouring keys; the numbers printed here are
what the tolerances of test_very_large_logits_* in tests/test_gpu_parity.py were taken from.

python tests/probe_large_logits.py [out.txt]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import serenade_b200 as S
    from golden_util import rel_l2
    from oracle import cfm_oracle as O
    dev = torch.device("cuda", 0)
    base = S.make_state_dict(seed=0, jitter=True)
    b = S.make_batch([200, 150], n_steps=2, seed=33)
    t = torch.tensor(0.41)
    lines = []
    for scale in (4.0, 16.0, 32.0, 64.0):
        sd = {k: (v * scale if k.endswith("attn1.to_q.weight") else v.clone()) for k, v in base.items()}
        with torch.inference_mode():
            v_ref = O.estimator_forward(sd, b.z, b.mask, b.mu, t, b.spks)
        for precision in ("tf32", "bf16"):
            cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                        decoder_attention_head_dim=512, precision=precision)
            cfm.load_state_dict(sd, strict=True)
            v = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), t, b.spks.to(dev)).cpu()
            free, total = cfm.estimator.engine(dev).softmax_stats()
            lines.append(f"W_q x {scale:g} ({precision}): rel_l2={rel_l2(v, v_ref):.3e} finite={bool(torch.isfinite(v).all())} "
                         f"max|v|={float(v.abs().max()):.3f} shift-free entries {free}/{total}")
            print(lines[-1], flush=True)
            del cfm
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
