"""Timing of the conditioning front half (content encoder, GST style encoder) and of the whole `Serenade.inference`
on one GPU, next to the CPU oracle (port of the reference modules) on the host cores.

    python tests/bench_frontend.py [out.txt]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402
from oracle import cfm_oracle as O  # noqa: E402   (CPU baseline leg)
from oracle import frontend_oracle as FO  # noqa: E402

ENC_FLOP_PER_FRAME = 2.0 * (7 * 768 * 512 + 2 * (3 * 512 * 512 + 2 * 512 * 512) + 7 * 512 * 80)


def gpu_time(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else None
    dev = torch.device("cuda", 0)
    m = S.models.Serenade()
    sd = {("cfm_decoder." + k): v for k, v in S.make_state_dict(0, False).items()}
    fsd = S.make_frontend_state_dict(0)
    sd.update(fsd)
    m.load_state_dict(sd, strict=True)
    dsd = S.make_state_dict(0, False)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    lines = []
    for (B, L) in ((1, 500), (16, 1000)):
        inp = S.make_frontend_inputs(L, L, seed=7, batch=B)
        x, mel = inp["x"].to(dev), inp["ref_logmel"].to(dev)
        t_enc = gpu_time(lambda: m.encode_content(x))
        t_sty = gpu_time(lambda: m.encode_style(mel))
        with torch.inference_mode():
            t0 = time.perf_counter(); FO.content_encoder(fsd, inp["x"]); c_enc = (time.perf_counter() - t0) * 1e3
            t0 = time.perf_counter(); FO.style_encoder(fsd, inp["ref_logmel"]); c_sty = (time.perf_counter() - t0) * 1e3
        fl = ENC_FLOP_PER_FRAME * B * L
        lines.append(f"B={B} L={L}: content encoder {t_enc * 1e3:8.1f} us ({fl / t_enc / 1e9:6.1f} TFLOP/s algorithmic, "
                     f"{B * L / t_enc * 1e3:.0f} frames/s; CPU oracle {c_enc:.1f} ms on {threads} threads), "
                     f"style encoder {t_sty * 1e3:8.1f} us (CPU oracle {c_sty:.1f} ms)")
        print(lines[-1], flush=True)
    # whole Serenade.inference, batch 1 like the reference CLI (bin/ssc_decode.py:428): 5 s source + 5 s reference prompt
    inp = S.make_frontend_inputs(500, 500, seed=8)
    dinp = {k: v.to(dev) for k, v in inp.items()}
    t_inf = gpu_time(lambda: m.inference(**dinp), reps=10)
    with torch.inference_mode():
        t0 = time.perf_counter()
        enc = FO.content_encoder(fsd, inp["x"]); renc = FO.content_encoder(fsd, inp["ref_x"]); sty = FO.style_encoder(fsd, inp["ref_logmel"])
        mu = FO.assemble_mu(enc, inp["midi"], inp["lft"], renc, inp["ref_midi"], inp["ref_lft"], inp["ref_logmel"])
        z = torch.randn(1, 80, mu.shape[2]) * 0.667
        O.solve_euler(dsd, z, torch.linspace(0, 1, 11), mu, torch.ones(1, 1, mu.shape[2], dtype=torch.bool), sty)
        c_inf = time.perf_counter() - t0
    lines.append(f"Serenade.inference (5 s source + 5 s reference prompt, T = 1000 decoder frames, 10 Euler steps): "
                 f"{t_inf:.2f} ms on the B200 (500 source frames -> {500 / t_inf * 1e3:.0f} source frames/s, RTF {t_inf / 5e3:.2e}); "
                 f"CPU oracle {c_inf * 1e3:.0f} ms on {threads} threads (RTF {c_inf / 5:.2e})")
    print(lines[-1], flush=True)
    if out:
        open(out, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
