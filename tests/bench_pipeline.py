"""The whole decode path of `serenade-decode` on one B200: content / style encoders -> CFM sampler -> HiFi-GAN vocoder
(`bin/ssc_decode.py:346-455` minus file I/O), next to the same chain on the host cores through the oracles.

    python tests/bench_pipeline.py [out.txt]

(a) one utterance, as the reference CLI decodes it: 5 s source + 5 s reference prompt, 10 Euler steps, batch 1;
(b) 16 utterances of 10 s (+ 5 s prompts) through `Serenade.inference_batch` + `Vocoder.decode_batch`.
Synthetic weights and features; identity statistics for the vocoder.  Times are CUDA events over whole calls
(host-side Python included), after warm-up.  (Lives under tests/: it executes oracle/ code as the baseline.)
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import cfm_oracle as O  # noqa: E402
from oracle import frontend_oracle as FO  # noqa: E402
from oracle import vocoder_oracle as VO  # noqa: E402

ORDER = ["x", "lengths", "midi", "lft", "ref_x", "ref_lengths", "ref_logmel", "ref_midi", "ref_lft"]


def gpu_time(fn, reps):
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
    model = S.models.Serenade()
    dsd = S.make_state_dict(0, False)
    fsd = S.make_frontend_state_dict(0)
    sd = {("cfm_decoder." + k): v for k, v in dsd.items()}
    sd.update(fsd)
    model.load_state_dict(sd, strict=True)
    vsd = S.make_vocoder_state_dict(0)
    gen = S.HiFiGANGenerator()
    gen.load_state_dict(vsd, strict=True)
    ident = dict(mean=torch.zeros(80), scale=torch.ones(80))
    voc = S.Vocoder(gen, ident, dev, trg_stats=ident)
    lines = []

    # ---- (a) one utterance
    inp = S.make_frontend_inputs(500, 500, seed=8)
    dinp = [inp[k].to(dev) for k in ORDER]

    def one():
        torch.manual_seed(11)
        mel = model.inference(*dinp)
        wav, _ = voc.decode(mel)
        return mel, wav
    ms, (mel, wav) = gpu_time(one, 10)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    with torch.inference_mode():
        t0 = time.perf_counter()
        enc = FO.content_encoder(fsd, inp["x"])
        renc = FO.content_encoder(fsd, inp["ref_x"])
        sty = FO.style_encoder(fsd, inp["ref_logmel"])
        mu = FO.assemble_mu(enc, inp["midi"], inp["lft"], renc, inp["ref_midi"], inp["ref_lft"], inp["ref_logmel"])
        torch.manual_seed(11)
        z = torch.randn(1, 80, mu.shape[2]) * 0.667
        x = O.solve_euler(dsd, z, torch.linspace(0, 1, 11), mu, torch.ones(1, 1, mu.shape[2], dtype=torch.bool), sty)
        mel_ref = x[0, :, 500:].transpose(0, 1)
        wav_ref = VO.hifigan_forward(vsd, mel_ref.transpose(0, 1).unsqueeze(0)).view(-1)
        cpu_s = time.perf_counter() - t0
    e_mel = float((mel.cpu().double() - mel_ref.double()).norm() / mel_ref.double().norm())
    e_wav = float((wav.cpu().double() - wav_ref.double()).norm() / wav_ref.double().norm())
    lines.append(f"one utterance (5 s source + 5 s prompt, 10 Euler steps, {wav.numel()} samples): {ms:.2f} ms on the B200 "
                 f"(RTF {ms / 5e3:.2e}); the same chain through the oracles on {threads} host threads {cpu_s * 1e3:.0f} ms "
                 f"(RTF {cpu_s / 5:.2e}); mel rel-L2 {e_mel:.2e}, waveform rel-L2 {e_wav:.2e} vs the oracle chain (bf16 mode)")
    print(lines[-1], flush=True)

    # ---- (b) 16 utterances of 10 s, batched
    items = []
    for k in range(16):
        it = S.make_frontend_inputs(1000, 500, seed=100 + k)
        items.append([it[name].to(dev) for name in ORDER])

    def batch():
        mels = model.inference_batch(items, max_batch=16)
        return voc.decode_batch(torch.stack(mels))
    ms_b, wavs = gpu_time(batch, 3)
    lines.append(f"16 utterances x (10 s source + 5 s prompt) via inference_batch + decode_batch: {ms_b:.1f} ms on the B200 for "
                 f"{wavs.shape[0]} x {wavs.shape[1]} samples = {16 * 1000 / ms_b * 1e3:.0f} source frames/s, RTF {ms_b / 1e3 / 160:.2e}; "
                 f"finite={bool(torch.isfinite(wavs).all())}")
    print(lines[-1], flush=True)
    # ---- (c) a decode loop over utterances of DISTINCT lengths (what bin/ssc_decode.py:346-378 does): every call meets a
    # new (B, T), i.e. a cold engine plan (workspace cudaMalloc, TMA descriptor encoding, CUDA-graph capture +
    # instantiation) unless lengths are bucketed (Serenade(shape_bucket=128), the default; 0 = one plan per length)
    lens = [403, 517, 288, 611, 349, 455, 572, 236, 498, 330, 420, 541]
    utts = []
    for k, L in enumerate(lens):
        it = S.make_frontend_inputs(L, 300, seed=300 + k)
        utts.append([it[name].to(dev) for name in ORDER])
    for bucket in (0, 128):
        m2 = S.models.Serenade(shape_bucket=bucket)
        m2.load_state_dict(sd, strict=True)
        m2.inference(*dinp)                      # weights packed, kernels loaded
        torch.cuda.synchronize()
        per = []
        for rep in range(2):                     # pass 0: every length is new; pass 1: the same lengths again
            t0 = time.perf_counter()
            for u in utts:
                m2.inference(*u)
            torch.cuda.synchronize()
            per.append((time.perf_counter() - t0) / len(utts) * 1e3)
        lines.append(f"variable-length loop, {len(lens)} utterances of {min(lens)}..{max(lens)} source frames + 300-frame prompt, "
                     f"shape_bucket={bucket}: {per[0]:.1f} ms per utterance on first sight of each length (cold plans), "
                     f"{per[1]:.1f} ms per utterance on the second pass (wall clock, encoders + CFM solve)")
        print(lines[-1], flush=True)
        del m2
        torch.cuda.empty_cache()
    if out_path:
        with open(out_path, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
