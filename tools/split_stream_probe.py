"""Probe: one 16-clip solve vs the same batch split over k engines on k CUDA streams (each engine has its own plan, workspace
and graph, so the solves are independent and their kernels may overlap at their heads and tails).
python tools/split_stream_probe.py [B] [T] [steps] [ways...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    ways = [int(a) for a in sys.argv[4:]] or [1, 2, 4]
    dev = torch.device("cuda", 0)
    sd = S.make_state_dict(0, False)
    b = S.make_batch([T] * B, n_steps=n, seed=1235, infill=(0.25, 0.75))
    full = [b.z.to(dev), b.t_span, b.mu.to(dev), b.mask.to(dev), b.spks.to(dev)]
    ref = None
    for k in ways:
        cfms, streams, parts = [], [], []
        for i in range(k):
            c = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512), decoder_attention_head_dim=512)
            c.load_state_dict(sd, strict=True)
            cfms.append(c)
            streams.append(torch.cuda.Stream(dev))
            sl = slice(i * B // k, (i + 1) * B // k)
            parts.append([full[0][sl].contiguous(), full[1], full[2][sl].contiguous(), full[3][sl].contiguous(), full[4][sl].contiguous()])

        def run():
            outs = []
            cur = torch.cuda.current_stream(dev)
            for c, s, p in zip(cfms, streams, parts):
                s.wait_stream(cur)
                with torch.cuda.stream(s):
                    outs.append(c.solve_euler(*p))
            for s in streams:
                cur.wait_stream(s)
            return outs
        for _ in range(3):
            outs = run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            outs = run()
        e1.record()
        torch.cuda.synchronize()
        x = torch.cat(outs, dim=0)
        if ref is None:
            ref = x
        print(f"{k} engine(s) x {B // k} clips on {k} stream(s): {e0.elapsed_time(e1) / reps:.3f} ms per {B}-clip solve, "
              f"bit-identical to the single solve: {bool(torch.equal(x, ref))}", flush=True)
        del cfms


if __name__ == "__main__":
    main()
