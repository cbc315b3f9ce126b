"""High-volume reproducibility check: ONE device-resident batch solved `reps` times; every result compared with the first
on the device (no host sync inside the loop).  python tools/determinism_solve.py [reps] [B] [T] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    n = int(sys.argv[4]) if len(sys.argv) > 4 else 10
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    if os.environ.get("SRN_NO_GRAPH") == "1":
        cfm.estimator.engine(dev).set_use_graph(False)
    b = S.make_batch([T] * B, n_steps=n, seed=1235, infill=(0.25, 0.75))
    args = [b.z.to(dev), b.t_span, b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
    ref = cfm.solve_euler(*args)
    flags = torch.zeros(reps, B, dtype=torch.bool, device=dev)
    for i in range(reps):
        x = cfm.solve_euler(*args)
        flags[i] = (x != ref).flatten(1).any(dim=1)
    torch.cuda.synchronize()
    f = flags.cpu()
    bad = f.any(dim=1).nonzero().flatten().tolist()
    bits = int(ref.view(torch.int32).to(torch.int64).sum().item())        # (exact: integer sum of the bit patterns)
    print(f"bit checksum of the first solve: {bits}  (sum {float(ref.double().sum()):.6f})")
    print(f"{reps} solves of B={B} T={T} steps={n}: {len(bad)} differ from the first"
          + (f"; solves {bad[:12]}, clips {[f[i].nonzero().flatten().tolist() for i in bad[:12]]}" if bad else " (deterministic)"),
          flush=True)


if __name__ == "__main__":
    main()
