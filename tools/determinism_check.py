"""Run-to-run determinism of the decode loops: the same clips decoded `passes` times (alternating the pipelined and the
serial loop of serenade_b200/sharding.py) must give bit-identical mels every time.
    python tools/determinism_check.py [clips] [passes] [serial|both]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402
from serenade_b200 import sharding  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    passes = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    mode = sys.argv[3] if len(sys.argv) > 3 else "both"
    dev = torch.device("cuda", 0)
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
    if os.environ.get("SRN_NO_GRAPH") == "1":
        cfm.estimator.engine(dev).set_use_graph(False)
    clips = []
    for i in range(n):
        b = S.make_batch([1000], n_steps=10, seed=100000 + i, infill=(0.25, 0.75))
        clips.append(dict(mu=b.mu[0], z=b.z[0], spk=b.spks[0]))
    t_span = torch.linspace(0, 1, 11)
    ref = sharding.decode_sharded(cfm.solve_euler, clips, t_span, max_batch=16, device=dev)
    bad_total = 0
    for p in range(passes):
        if mode == "both" and p % 2 == 0 and hasattr(sharding, "decode_sharded_pipelined"):
            got, kind = sharding.decode_sharded_pipelined(cfm, clips, t_span, max_batch=16, device=dev), "pipelined"
        else:
            got, kind = sharding.decode_sharded(cfm.solve_euler, clips, t_span, max_batch=16, device=dev), "serial"
        bad = []
        for i in range(n):
            if not torch.equal(got[i], ref[i]):
                d = (got[i] - ref[i]).abs()
                bad.append((i, int((d > 0).sum()), float(d.max())))
        bad_total += len(bad)
        if bad:
            print(f"pass {p} ({kind}): {len(bad)} of {n} clips differ from the first serial pass", bad[:6], flush=True)
    print(f"{passes} passes x {n // 16} solves: " + ("DETERMINISTIC" if bad_total == 0 else f"NONDETERMINISTIC, {bad_total} clip results differed"),
          flush=True)


if __name__ == "__main__":
    main()
