"""The reference CLI's decode job end to end on the B200: a dumped feature directory on disk ->
serenade_b200.decode_io.decode_dump(Serenade, Vocoder) -> wav and lf0 files, and the files hold exactly what calling the
model and the vocoder by hand gives for the same noise."""
import os
import wave

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scalers(rng):
    from sklearn.preprocessing import MinMaxScaler, StandardScaler
    return {"logmel": StandardScaler().fit(rng.standard_normal((300, 80)) * 2 - 5),
            "hubert": StandardScaler().fit(rng.standard_normal((300, 768))),
            "score": MinMaxScaler().fit(np.abs(rng.standard_normal((300, 1))) * 3 + 2),
            "loud": MinMaxScaler().fit(rng.standard_normal((300, 1)) * 10 - 30)}


def _dump_utterance(D, path, T, rng, f0_lo, f0_hi):
    feats = {"wave": (rng.standard_normal(T * 240) * 0.1).astype(np.float32),
             "hubert": rng.standard_normal((T, 768)).astype(np.float32),
             "logmel": (rng.standard_normal((T, 80)) * 2 - 5).astype(np.float32),
             "est_lf0_score": (np.abs(rng.standard_normal((T, 1))) * 3 + 2).astype(np.float32),
             "midi": rng.integers(40, 80, (T, 1)).astype(np.float32),
             "loud": (rng.standard_normal((T, 1)) * 10 - 30).astype(np.float32),
             "f0": np.where(rng.random(T) > 0.3, rng.uniform(f0_lo, f0_hi, T), 0.0)}
    for k, v in feats.items():
        D.write_hdf5(path, k, v)
    return feats


def test_decode_job_from_disk_to_wav(tmp_path):
    import serenade_b200 as S
    from serenade_b200 import decode_io as D
    assert torch.cuda.is_available(), "GPU tests need a B200"
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(5)
    scaler = _scalers(rng)
    dump = tmp_path / "dump.1"
    src, ref = "EN_spk1_songA_Breathy_Group_0", "EN_spk1_songB_Falsetto_Group_2"
    _dump_utterance(D, str(dump / f"{src}.npz"), 96, rng, 120, 240)
    ref_feats = _dump_utterance(D, str(dump / f"{ref}.npz"), 71, rng, 300, 520)

    model = S.models.Serenade()
    sd = {("cfm_decoder." + k): v for k, v in S.make_state_dict(0, False).items()}
    sd.update(S.make_frontend_state_dict(0))
    model.load_state_dict(sd, strict=True)
    gen = S.HiFiGANGenerator()
    gen.load_state_dict(S.make_vocoder_state_dict(0), strict=True)
    voc = S.Vocoder(gen, dict(mean=torch.zeros(80), scale=torch.ones(80)), dev,
                    trg_stats={"mean": scaler["logmel"].mean_, "scale": scaler["logmel"].scale_})

    ds = D.FeatsDataset(str(dump), audio_query=f"{src}.npz", scaler=scaler, return_utt_id=True)
    refs = D.get_random_ref_style(str(dump), src, ext=".npz")
    assert sorted(refs) == ["Breathy", "Falsetto"]      # (its own style is found too; the loop skips it)
    out = tmp_path / "out"
    torch.manual_seed(21)
    written = D.decode_dump(model, voc, ds, scaler, str(out), 24000, dev, ref_dict=refs)
    assert written == [f"{src}_Falsetto"]
    assert sorted(os.listdir(out)) == sorted([f"{src}_gt.wav", "00_Falsetto_reference.wav", f"{src}_Falsetto.wav",
                                              f"{src}_Falsetto.npz"])

    # the same decode by hand
    item = ds[0]
    norm = D.normalise_features(scaler, hubert=ref_feats["hubert"], logmel=ref_feats["logmel"],
                                score=ref_feats["est_lf0_score"], loud=ref_feats["loud"])
    f = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float).to(dev).unsqueeze(0)        # noqa: E731
    torch.manual_seed(21)
    mel = model.inference(f(item["hubert"]), torch.tensor([96], device=dev), f(item["score"]), f(item["loud"]),
                          f(norm["hubert"]), torch.tensor([71], device=dev), f(norm["logmel"]),
                          f(norm["score"]).view(1, -1, 1), f(norm["loud"]))
    assert mel.shape == (96, 80)
    wav, sr = voc.decode(mel)
    hop = wav.numel() // 96
    want = np.clip(np.rint(wav.cpu().numpy().astype(np.float64) * 32767), -32768, 32767).astype(np.int16)
    with wave.open(str(out / f"{src}_Falsetto.wav"), "rb") as w:
        assert w.getframerate() == 24000 == sr and w.getnframes() == 96 * hop and w.getsampwidth() == 2
        got = np.frombuffer(w.readframes(w.getnframes()), dtype="<i2")
    np.testing.assert_array_equal(got, want)            # (the library is bit-reproducible: same noise, same samples)
    assert np.isfinite(wav.cpu().numpy()).all() and np.abs(got).max() > 0
    lf0 = D.read_hdf5(str(out / f"{src}_Falsetto.npz"), "lf0")
    src_f0 = D.read_hdf5(str(dump / f"{src}.npz"), "f0")
    v = src_f0 > 0
    semis = 12 * np.log2(lf0[v].astype(np.float64) / src_f0[v])
    assert lf0.dtype == np.float32 and np.allclose(semis, np.round(semis[0]), atol=1e-4) and round(semis[0]) > 0
