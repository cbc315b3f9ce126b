"""Golden vectors for serenade_b200/decode_io.py from the UNMODIFIED reference (run in the build container, where
/root/reference exists; the fixture travels, the reference does not):

  * linear_midi_shift / hz_to_cent_based_c4 / cent_to_hz_based_c4 of /root/reference/serenade/bin/ssc_decode.py, loaded with
    its absent dependencies (soundfile, h5py, the serenade package itself) stubbed - those functions only use numpy;
  * FeatsDataset.__getitem__ of /root/reference/serenade/datasets/audio_mel_dataset.py with read_hdf5 served from an
    in-memory dict (h5py is absent), and scalers fitted with scikit-learn like the reference's preprocessing.

python tests/golden/make_golden_decode_io.py  ->  tests/golden/decode_io.npz
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference/serenade"
HERE = os.path.dirname(os.path.abspath(__file__))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


def load_reference(store):
    """`store`: {(file, dataset): array} served by the stubbed read_hdf5."""
    def read_hdf5(name, path):
        return store[(name, path)].copy()
    _stub("soundfile")
    _stub("h5py")
    _stub("serenade")
    _stub("serenade.models")
    _stub("serenade.vocoder", Vocoder=object)
    _stub("serenade.utils", read_hdf5=read_hdf5, write_hdf5=None, find_files=None, get_basename=None)
    _stub("serenade.utils.masking", make_non_pad_mask=None)
    _stub("serenade.datasets")
    mods = {}
    for dotted, path in (("serenade.datasets.audio_mel_dataset", f"{REF}/datasets/audio_mel_dataset.py"),
                         ("serenade.bin.ssc_decode", f"{REF}/bin/ssc_decode.py")):
        spec = importlib.util.spec_from_file_location(dotted, path)
        m = importlib.util.module_from_spec(spec)
        sys.modules[dotted] = m
        spec.loader.exec_module(m)
        mods[dotted] = m
    return mods["serenade.datasets.audio_mel_dataset"], mods["serenade.bin.ssc_decode"]


def make_inputs(seed=7):
    from sklearn.preprocessing import MinMaxScaler, StandardScaler
    rng = np.random.default_rng(seed)
    T = 37
    feats = {"wave": rng.standard_normal(T * 240).astype(np.float32) * 0.1,
             "hubert": rng.standard_normal((T, 768)).astype(np.float32),
             "logmel": (rng.standard_normal((T, 80)) * 2 - 5).astype(np.float32),
             "est_lf0_score": np.abs(rng.standard_normal((T, 1))).astype(np.float32) * 3 + 2,
             "midi": rng.integers(40, 80, (T, 1)).astype(np.float32),
             "loud": (rng.standard_normal((T, 1)) * 10 - 30).astype(np.float32),
             "f0": np.where(rng.random(T) > 0.3, rng.uniform(120, 500, T), 0.0)}
    scaler = {"logmel": StandardScaler().fit(rng.standard_normal((200, 80)) * 2 - 5),
              "hubert": StandardScaler().fit(rng.standard_normal((200, 768))),
              "score": MinMaxScaler().fit(np.abs(rng.standard_normal((200, 1))) * 3 + 2),
              "loud": MinMaxScaler().fit(rng.standard_normal((200, 1)) * 10 - 30)}
    return feats, scaler


def main():
    feats, scaler = make_inputs()
    store = {("/dump/EN_spk1_song_Breathy_Group_0.h5", k): v for k, v in feats.items()}
    ds_mod, dec = load_reference(store)
    out = {}
    # f0 helpers on three source / reference pairs: shift up, shift down, many unvoiced frames
    rng = np.random.default_rng(3)
    for i, (lo_s, hi_s, lo_t, hi_t, p) in enumerate([(110, 220, 300, 600, 0.2), (300, 700, 90, 180, 0.2), (150, 400, 160, 380, 0.7)]):
        sm = np.where(rng.random(101) > p, rng.uniform(lo_s, hi_s, 101), 0.0)
        tm = np.where(rng.random(83) > p, rng.uniform(lo_t, hi_t, 83), 0.0)
        out[f"shift{i}_sm"], out[f"shift{i}_tm"] = sm.copy(), tm.copy()
        out[f"shift{i}_out"] = dec.linear_midi_shift(sm.copy(), tm.copy())
    hz = np.where(rng.random(50) > 0.3, rng.uniform(60, 900, 50), 0.0)
    out["hz"], out["cent"] = hz, dec.hz_to_cent_based_c4(hz.copy())
    out["hz_back"] = dec.cent_to_hz_based_c4(out["cent"].copy())
    # FeatsDataset item with the reference's normalisation
    ds_mod.find_files = lambda root, query: ["/dump/EN_spk1_song_Breathy_Group_0.h5"]
    ds = ds_mod.FeatsDataset(root_dir="/dump", scaler=scaler, return_utt_id=True, score_type="est_lf0_score")
    item = ds[0]
    assert item["utt_id"] == "EN_spk1_song_Breathy_Group_0"
    for k in ("audio", "logmel", "hubert", "loud", "score", "midi", "lf0"):
        out["item_" + k] = np.asarray(item[k])
    np.savez_compressed(os.path.join(HERE, "decode_io.npz"), **out)
    print("wrote", os.path.join(HERE, "decode_io.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
