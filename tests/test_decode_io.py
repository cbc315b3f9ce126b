"""On-disk half of the decode front end (serenade_b200/decode_io.py) - CPU tests.

Pinned against the reference through tests/golden/decode_io.npz (tests/golden/make_golden_decode_io.py ran the reference's own
`linear_midi_shift`, cent conversions and `FeatsDataset.__getitem__`); the containers, the wav writer and the decode loop
are checked on their stated behaviour (h5py / soundfile are not installed here: the .npz twin and the standard library's
wave reader stand in)."""
import importlib.util
import os
import wave

import numpy as np
import pytest
import torch

from serenade_b200 import decode_io as D

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "decode_io.npz"))


@pytest.fixture(scope="module")
def inputs():
    spec = importlib.util.spec_from_file_location("make_golden_decode_io", os.path.join(HERE, "golden", "make_golden_decode_io.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.make_inputs()


def test_f0_helpers_match_the_reference(golden):
    for i in range(3):
        sm = golden[f"shift{i}_sm"].copy()
        out = D.linear_midi_shift(sm, golden[f"shift{i}_tm"].copy())
        assert out is sm                                           # in place, like the reference
        np.testing.assert_array_equal(out, golden[f"shift{i}_out"])
    np.testing.assert_array_equal(D.hz_to_cent_based_c4(golden["hz"].copy()), golden["cent"])
    np.testing.assert_array_equal(D.cent_to_hz_based_c4(golden["cent"].copy()), golden["hz_back"])
    # the shift is a whole number of semitones on voiced frames, unvoiced frames stay 0
    sm, out = golden["shift0_sm"], golden["shift0_out"]
    v = sm > 0
    semis = 12 * np.log2(out[v] / sm[v])
    assert np.allclose(semis, np.round(semis[0]), atol=1e-9) and np.all(out[~v] == 0)


def test_feats_dataset_item_matches_the_reference(golden, inputs, tmp_path):
    feats, scaler = inputs
    path = tmp_path / "dump" / "EN_spk1_song_Breathy_Group_0.npz"
    for k, v in feats.items():
        D.write_hdf5(str(path), k, v)
    ds = D.FeatsDataset(str(tmp_path / "dump"), audio_query="*.npz", scaler=scaler, return_utt_id=True)
    assert len(ds) == 1
    item = ds[0]
    assert item["utt_id"] == "EN_spk1_song_Breathy_Group_0"
    for k in ("audio", "logmel", "hubert", "loud", "score", "midi", "lf0"):
        got = np.asarray(item[k])
        assert got.dtype == golden["item_" + k].dtype, k
        np.testing.assert_array_equal(got, golden["item_" + k], err_msg=k)
    # scalers through joblib, as bin/ssc_decode.py:303 loads them
    from joblib import dump
    dump(scaler, tmp_path / "stats.joblib")
    sc = D.load_scalers(tmp_path / "stats.joblib")
    np.testing.assert_array_equal(sc["logmel"].mean_, scaler["logmel"].mean_)
    with pytest.raises(AssertionError):
        D.FeatsDataset(str(tmp_path / "dump"), audio_query="*.h5")     # nothing found


def test_container_semantics(tmp_path, caplog):
    f = str(tmp_path / "a" / "b" / "x.npz")
    D.write_hdf5(f, "one", np.arange(5, dtype=np.float32))
    D.write_hdf5(f, "two", [[1, 2], [3, 4]])                           # lists are converted; the first dataset is kept
    np.testing.assert_array_equal(D.read_hdf5(f, "one"), np.arange(5, dtype=np.float32))
    assert D.read_hdf5(f, "two").tolist() == [[1, 2], [3, 4]]
    D.write_hdf5(f, "one", np.zeros(2))                                # overwrite: recreated
    assert D.read_hdf5(f, "one").shape == (2,)
    with pytest.raises(SystemExit):
        D.write_hdf5(f, "one", np.ones(2), is_overwrite=False)
    assert D.read_hdf5(f, "missing") is None
    assert D.read_hdf5(str(tmp_path / "nothing.npz"), "one") is None
    assert "There is no such a hdf5 file" in caplog.text
    # HDF5 proper: h5py when importable, a loud failure (never a silent substitute) when not
    try:
        import h5py  # noqa: F401
        have = True
    except ImportError:
        have = False
    h5 = str(tmp_path / "y.h5")
    if have:
        D.write_hdf5(h5, "lf0", np.arange(3, dtype=np.float32))
        np.testing.assert_array_equal(D.read_hdf5(h5, "lf0"), np.arange(3, dtype=np.float32))
    else:
        with pytest.raises(RuntimeError, match="h5py is not installed"):
            D.write_hdf5(h5, "lf0", np.arange(3, dtype=np.float32))
    assert sorted(os.path.basename(p) for p in D.find_files(str(tmp_path), "*.npz")) == ["x.npz"]


def test_wav_writer_is_16_bit_pcm(tmp_path):
    x = np.array([0.0, 0.5, -0.5, 1.0, -1.0, 1.7, -1.7, 1e-5, 0.25 + 1 / 65534], dtype=np.float32)
    p = str(tmp_path / "out" / "w.wav")
    D.write_wav_pcm16(p, x, 24000)
    with wave.open(p, "rb") as w:
        assert (w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()) == (1, 2, 24000, len(x))
        got = np.frombuffer(w.readframes(len(x)), dtype="<i2")
    want = np.clip(np.rint(x.astype(np.float64) * 32767), -32768, 32767).astype(np.int16)
    np.testing.assert_array_equal(got, want)
    assert got[3] == 32767 and got[4] == -32767 and got[5] == 32767 and got[6] == -32768
    D.write_wav_pcm16(p, np.stack([x, -x], axis=1), 16000)             # (n, channels)
    with wave.open(p, "rb") as w:
        assert w.getnchannels() == 2 and w.getnframes() == len(x)
    with pytest.raises(TypeError):
        D.write_wav_pcm16(p, np.arange(4, dtype=np.int32), 16000)


class _FakeModel:
    """Stands in for models.Serenade on the CPU: records what the loop passes, returns a mel of the source length."""

    def __init__(self):
        self.calls = []

    def inference(self, x, lengths, midi, lft, ref_x, ref_lengths, ref_logmel, ref_midi, ref_lft, **kw):
        self.calls.append(dict(x=x, lengths=lengths, midi=midi, lft=lft, ref_x=ref_x, ref_lengths=ref_lengths,
                               ref_logmel=ref_logmel, ref_midi=ref_midi, ref_lft=ref_lft, kw=kw))
        return torch.zeros(x.shape[1], 80)


class _FakeVocoder:
    def decode(self, c):
        return torch.linspace(-0.5, 0.5, c.shape[0] * 240), 24000


def test_decode_loop_writes_what_the_reference_cli_writes(inputs, tmp_path):
    feats, scaler = inputs
    dump = tmp_path / "dump.1"
    names = ["EN_spk1_song_Breathy_Group_0", "EN_spk1_song_Falsetto_Group_1", "EN_spk1_other_Mixed_Voice_Group_0"]
    rng = np.random.default_rng(0)
    for n in names:
        for k, v in feats.items():
            D.write_hdf5(str(dump / f"{n}.npz"), k, v if k != "f0" else v * rng.uniform(0.8, 1.6))
    ds = D.FeatsDataset(str(dump), audio_query=names[0] + ".npz", scaler=scaler, return_utt_id=True)
    refs = D.get_random_ref_style(str(dump), names[0], ext=".npz", rng=np.random.default_rng(1))
    assert sorted(refs) == ["Breathy", "Falsetto", "Mixed_Voice"]
    model, out = _FakeModel(), tmp_path / "out"
    src_f0 = D.read_hdf5(str(dump / f"{names[0]}.npz"), "f0")
    written = D.decode_dump(model, _FakeVocoder(), ds, scaler, str(out), 24000, "cpu", ref_dict=refs)
    assert written == [f"{names[0]}_Falsetto", f"{names[0]}_Mixed_Voice"]       # its own style (Breathy) is skipped
    files = sorted(os.listdir(out))
    assert files == sorted([f"{names[0]}_gt.wav", "00_Falsetto_reference.wav", "00_Mixed_Voice_reference.wav",
                            f"{names[0]}_Falsetto.wav", f"{names[0]}_Falsetto.npz",
                            f"{names[0]}_Mixed_Voice.wav", f"{names[0]}_Mixed_Voice.npz"])
    T = feats["hubert"].shape[0]
    c = model.calls[0]
    assert c["x"].shape == (1, T, 768) and c["lengths"].tolist() == [T] and c["ref_lengths"].tolist() == [T]
    assert c["midi"].shape == (1, T, 1) and c["lft"].shape == (1, T, 1) and c["ref_midi"].shape == (1, T, 1)
    assert c["ref_logmel"].shape == (1, T, 80) and c["x"].dtype == torch.float32
    want = D.normalise_features(scaler, hubert=feats["hubert"])["hubert"]
    np.testing.assert_allclose(c["ref_x"][0].numpy(), want.astype(np.float32), rtol=1e-6, atol=1e-6)
    with wave.open(str(out / f"{names[0]}_Falsetto.wav"), "rb") as w:
        assert w.getframerate() == 24000 and w.getnframes() == T * 240
    # the lf0 files: float32, voiced frames moved by whole semitones (cumulatively over the styles, as the reference's
    # in-place shift does), unvoiced frames 0
    lf0 = D.read_hdf5(str(out / f"{names[0]}_Mixed_Voice.npz"), "lf0")
    assert lf0.dtype == np.float32 and lf0.shape == src_f0.shape
    v = src_f0 > 0
    semis = 12 * np.log2(lf0[v].astype(np.float64) / src_f0[v])
    assert np.allclose(semis, np.round(semis), atol=1e-4) and np.all(lf0[~v] == 0)
    with pytest.raises(ValueError):
        D.decode_dump(model, _FakeVocoder(), ds, scaler, str(out), 24000, "cpu")     # neither ref_dict nor dumpdir
