"""On-disk half of the decode front end: what `bin/ssc_decode.py` does around `model.inference` and
`vocoder.decode` - feature files in, scalers, PCM-16 wav + shifted-lf0 files out - as a library, not a CLI.

Reference interfaces mirrored (same names, arguments and failure behaviour):
  utils/utils.py:28-47     find_files
  utils/utils.py:50-75     read_hdf5   (returns None and logs when the file or the dataset is missing)
  utils/utils.py:78-119    write_hdf5  (recreates an existing dataset when is_overwrite, else exits)
  datasets/audio_mel_dataset.py:20-137   FeatsDataset (items, normalisation, utt ids)
  bin/ssc_decode.py:99-154   hz_to_cent_based_c4 / cent_to_hz_based_c4 / linear_midi_shift
  bin/ssc_decode.py:157-187  get_random_ref_style
  bin/ssc_decode.py:302-455  the decode loop -> decode_dump()

Containers.  The reference stores features in HDF5 through h5py and audio through soundfile; neither is installed in
this image.  `.h5` / `.hdf5` paths go to h5py when it can be imported and fail loudly when it cannot; `.npz` paths hold
the same named arrays in a numpy zip (the twin this image can read, write and test; `h5_to_npz` converts where h5py
exists).  Wav output is 16-bit PCM written with the standard library, sample = rint(x * 32767) like libsndfile's
default (non-clipping) float -> short conversion, with out-of-range samples saturated instead of wrapped.

Nothing here touches the GPU; the model and the vocoder are passed in.
"""
import fnmatch
import glob
import logging
import os
import sys
import wave as _wave
from typing import Callable, Dict, Iterable, Optional

import numpy as np

_H5_EXT = (".h5", ".hdf5")
_C4_HZ = 440 * 2 ** (3 / 12 - 1)
_C4_CENT = 4800


# ------------------------------------------------------------------------------------------------ containers
def _h5py():
    try:
        import h5py
    except ImportError as e:
        raise RuntimeError("this file is HDF5 and h5py is not installed in this environment; install h5py, or convert the "
                           "dump to the .npz twin (serenade_b200.decode_io.h5_to_npz on a machine that has h5py)") from e
    return h5py


def _is_npz(name: str) -> bool:
    return str(name).endswith(".npz")


def find_files(root_dir, query="*.wav", include_root_dir=True):
    """All files below root_dir whose name matches `query` (symlinked directories followed)."""
    found = []
    for root, _dirs, names in os.walk(root_dir, followlinks=True):
        found += [os.path.join(root, n) for n in fnmatch.filter(names, query)]
    if not include_root_dir:
        found = [f.replace(root_dir + "/", "") for f in found]
    return found


def read_hdf5(hdf5_name, hdf5_path):
    """The array stored under `hdf5_path` in the container `hdf5_name`; None (and an error in the log) when the file or
    the dataset does not exist."""
    if not os.path.exists(hdf5_name):
        logging.error(f"There is no such a hdf5 file ({hdf5_name}).")
        return None
    if _is_npz(hdf5_name):
        with np.load(hdf5_name, allow_pickle=False) as z:
            if hdf5_path not in z.files:
                logging.error(f"There is no such a data in hdf5 file. ({hdf5_path})")
                return None
            a = z[hdf5_path]
            return a[()] if a.ndim == 0 else a
    with _h5py().File(hdf5_name, "r") as f:
        if hdf5_path not in f:
            logging.error(f"There is no such a data in hdf5 file. ({hdf5_path})")
            return None
        return f[hdf5_path][()]


def write_hdf5(hdf5_name, hdf5_path, write_data, is_overwrite=True):
    """Store `write_data` under `hdf5_path`, keeping the container's other datasets; an existing dataset of that name is
    replaced when is_overwrite, otherwise the process exits with status 1 like the reference."""
    write_data = np.array(write_data)
    folder = os.path.dirname(hdf5_name)
    if folder and not os.path.exists(folder):
        os.makedirs(folder)
    if _is_npz(hdf5_name):
        held = {}
        if os.path.exists(hdf5_name):
            with np.load(hdf5_name, allow_pickle=False) as z:
                held = {k: z[k] for k in z.files}
        if hdf5_path in held:
            if not is_overwrite:
                logging.error("Dataset in hdf5 file already exists. if you want to overwrite, please set is_overwrite = True.")
                sys.exit(1)
            logging.warning("Dataset in hdf5 file already exists. recreate dataset in hdf5.")
        held[hdf5_path] = write_data
        tmp = hdf5_name + ".tmp.npz"
        np.savez(tmp, **held)
        os.replace(tmp, hdf5_name)
        return
    h5py = _h5py()
    if os.path.exists(hdf5_name):
        f = h5py.File(hdf5_name, "r+")
        if hdf5_path in f:
            if not is_overwrite:
                logging.error("Dataset in hdf5 file already exists. if you want to overwrite, please set is_overwrite = True.")
                f.close()
                sys.exit(1)
            logging.warning("Dataset in hdf5 file already exists. recreate dataset in hdf5.")
            del f[hdf5_path]
    else:
        f = h5py.File(hdf5_name, "w")
    f.create_dataset(hdf5_path, data=write_data)
    f.flush()
    f.close()


def h5_to_npz(h5_name, npz_name=None):
    """Every dataset of the root group of an HDF5 dump as the .npz twin (needs h5py)."""
    npz_name = npz_name or os.path.splitext(h5_name)[0] + ".npz"
    with _h5py().File(h5_name, "r") as f:
        np.savez(npz_name, **{k: f[k][()] for k in f.keys()})
    return npz_name


def write_wav_pcm16(path, wave, sampling_rate):
    """`soundfile.write(path, wave, sampling_rate, "PCM_16")` for a float waveform (n,) or (n, channels)."""
    a = np.asarray(wave)
    if a.dtype.kind == "f":
        a = np.clip(np.rint(a.astype(np.float64) * 32767.0), -32768, 32767).astype("<i2")
    elif a.dtype != np.int16:
        raise TypeError(f"write_wav_pcm16 takes float or int16 samples, got {a.dtype}")
    channels = 1 if a.ndim == 1 else a.shape[1]
    folder = os.path.dirname(path)
    if folder and not os.path.exists(folder):
        os.makedirs(folder)
    with _wave.open(path, "wb") as w:
        w.setnchannels(channels)
        w.setsampwidth(2)
        w.setframerate(int(sampling_rate))
        w.writeframes(np.ascontiguousarray(a.astype("<i2")).tobytes())


def load_scalers(path):
    """`joblib.load(stats.joblib)`: {"logmel", "hubert": StandardScaler; "score", "loud": MinMaxScaler}
    (bin/ssc_decode.py:303)."""
    from joblib import load
    return load(path)


# ------------------------------------------------------------------------------------------------ dataset
class FeatsDataset:
    """The reference's FeatsDataset: one dict per dumped utterance with `audio`, `logmel`, `hubert`, `loud`, `score`,
    `midi`, `lf0` (+ `utt_id`), normalised with the scalers of stats.joblib: standardisation for `logmel` / `hubert`,
    min-max for `score` / `loud`; `midi`, `lf0` and `audio` stay raw.  `audio_query` may name the .npz twin."""

    def __init__(self, root_dir, audio_query="*.h5", scaler=None, return_utt_id=False, allow_cache=False,
                 score_type="est_lf0_score", logmel_type="logmel", load_fn: Optional[Callable] = None):
        self.audio_files = sorted(find_files(root_dir, audio_query))
        assert len(self.audio_files) != 0, f"Not found any audio files in ${root_dir}."
        self.utt_ids = [os.path.splitext(os.path.basename(f))[0] for f in self.audio_files]
        self.scaler = scaler
        self.score_type, self.logmel_type = score_type, logmel_type
        self.return_utt_id = return_utt_id
        self.allow_cache = allow_cache
        self._load = load_fn or read_hdf5
        self._cache: Dict[int, dict] = {}

    def __len__(self):
        return len(self.audio_files)

    def __getitem__(self, idx):
        if self.allow_cache and idx in self._cache:
            return self._cache[idx]
        f = self.audio_files[idx]
        names = {"audio": "wave", "hubert": "hubert", "logmel": self.logmel_type, "score": self.score_type,
                 "midi": "midi", "loud": "loud", "lf0": "f0"}
        it = {k: self._load(f, n) for k, n in names.items()}
        if self.scaler is not None:
            it.update(normalise_features(self.scaler, logmel=it["logmel"], hubert=it["hubert"], score=it["score"],
                                         loud=it["loud"]))
            if np.isnan(it["logmel"]).any():
                logging.info(f"contains nan: {self.utt_ids[idx]}")
        if self.return_utt_id:
            it["utt_id"] = self.utt_ids[idx]
        if self.allow_cache:
            self._cache[idx] = it
        return it

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def normalise_features(scaler, **feats):
    """(x - mean) / scale for `logmel`, `hubert`; (x - min) / (max - min) for `score`, `loud` - the arithmetic of
    audio_mel_dataset.py:97-110 and bin/ssc_decode.py:398-415, in the arrays' own dtype promotion."""
    out = {}
    for k, v in feats.items():
        s = scaler[k]
        if k in ("logmel", "hubert"):
            out[k] = (v - s.mean_) / s.scale_
        else:
            out[k] = (v - s.data_min_) / (s.data_max_ - s.data_min_)
    return out


# ------------------------------------------------------------------------------------------------ f0 helpers
def hz_to_cent_based_c4(hz):
    out = np.zeros_like(hz)
    nz = np.nonzero(hz)
    out[nz] = 1200 * np.log(hz[nz] / _C4_HZ) / np.log(2) + _C4_CENT
    return out


def cent_to_hz_based_c4(cent):
    out = np.zeros_like(cent)
    nz = np.nonzero(cent)
    out[nz] = np.exp((cent[nz] - _C4_CENT) * np.log(2) / 1200) * _C4_HZ
    return out


def _mean_log_f0(f0):
    return np.mean(np.log(f0[np.nonzero(f0)]))


def linear_midi_shift(sm, tm):
    """Shift the voiced frames of the source f0 `sm` (Hz, modified IN PLACE and returned, like the reference) by a
    whole number of semitones towards the reference singer's `tm`: the difference of the mean log-f0 in cents, scaled
    by 1.4 upwards or 5/7 downwards, rounded to 100 cents."""
    voiced = sm > 0
    src_cent = 1200 * np.log(np.exp(_mean_log_f0(sm)) / _C4_HZ) / np.log(2) + _C4_CENT
    tgt_cent = 1200 * np.log(np.exp(_mean_log_f0(tm)) / _C4_HZ) / np.log(2) + _C4_CENT
    diff = tgt_cent - src_cent
    shift = round(diff * (1.4 if diff >= 0 else 5 / 7) / 100) * 100
    sm[voiced] = hz_to_cent_based_c4(sm[voiced])
    sm[voiced] = np.maximum(0, sm[voiced] + shift)
    sm[voiced] = cent_to_hz_based_c4(sm[voiced])
    return sm


def get_random_ref_style(dumpdir, utt_id, ext=".h5", rng=None):
    """One randomly chosen dump file per singing style of the same language and singer as `utt_id`
    ({ln}_{spk}_*_{style}_Group_*), looking in the sibling dump.1 / dump.2 directory when this one has none."""
    rng = rng or np.random
    dirname = os.path.dirname(os.path.join(dumpdir, f"{utt_id}{ext}"))
    ln, spk = utt_id.split(".")[0].split("_")[:2]
    chosen = {}
    for style in ("Breathy", "Falsetto", "Pharyngeal", "Mixed_Voice"):
        name = f"{ln}_{spk}_*_{style}_Group_*{ext}"
        files = glob.glob(os.path.join(dirname, name))
        if not files:
            for a, b in (("dump.2", "dump.1"), ("dump.1", "dump.2")):
                if a in dirname:
                    files = glob.glob(os.path.join(dirname.replace(a, b), name))
                    break
        if files:
            chosen[style] = rng.choice(sorted(files))
    logging.info(f"Using reference styles: {chosen}")
    return chosen


# ------------------------------------------------------------------------------------------------ the decode loop
def decode_dump(model, vocoder, dataset: Iterable[dict], scaler, outdir, sampling_rate, device, ref_dict=None,
                dumpdir=None, ext=None, out_ext=None, n_timesteps=10, temperature=0.667):
    """bin/ssc_decode.py:346-455 as a function: for every utterance of `dataset` (items with `utt_id`) and every
    reference style file of `ref_dict` ({style: path}; drawn once with get_random_ref_style(dumpdir, first utt_id) when
    None, like the reference) whose style name is not part of the utterance id:

        <outdir>/<utt_id>_gt.wav               the utterance's own audio
        <outdir>/00_<style>_reference.wav      the reference clip
        <outdir>/<utt_id>_<style>.wav          vocoder.decode(model.inference(...))
        <outdir>/<utt_id>_<style><out_ext>     dataset "lf0": the source f0 shifted towards the reference singer

    `model` has the reference's `inference(x, lengths, score, loud, ref_x, ref_lengths, ref_mel, ref_score, ref_loud)`,
    `vocoder` its `decode(mel) -> (wave, sr)`.  Returns the list of written utterance/style names."""
    import torch
    os.makedirs(outdir, exist_ok=True)
    written = []

    def dev(a, shape=None):
        t = torch.as_tensor(np.asarray(a), dtype=torch.float).to(device)
        return t.view(*shape) if shape is not None else t.unsqueeze(0)

    for batch in dataset:
        utt_id = batch["utt_id"]
        logging.info(f"Decoding {utt_id}")
        write_wav_pcm16(os.path.join(outdir, f"{utt_id}_gt.wav"), batch["audio"], sampling_rate)
        x = dev(batch["hubert"])
        lengths = torch.tensor([x.shape[1]], dtype=torch.long, device=device)
        scores, lfts = dev(batch["score"]), dev(batch["loud"])
        lf0 = batch["lf0"]
        if ref_dict is None:
            if dumpdir is None:
                raise ValueError("decode_dump needs ref_dict or dumpdir")
            ref_dict = get_random_ref_style(dumpdir, utt_id, ext=ext or ".h5")
        for style, ref_path in ref_dict.items():
            if style in utt_id:         # (no reconstruction of the utterance's own style)
                continue
            ref = {n: read_hdf5(ref_path, n) for n in ("hubert", "logmel", "loud", "wave", "est_lf0_score", "f0")}
            write_wav_pcm16(os.path.join(outdir, f"00_{style}_reference.wav"), ref["wave"], sampling_rate)
            norm = normalise_features(scaler, hubert=ref["hubert"], logmel=ref["logmel"], score=ref["est_lf0_score"],
                                      loud=ref["loud"])
            ref_cvec, ref_mel, ref_lft = dev(norm["hubert"]), dev(norm["logmel"]), dev(norm["loud"])
            ref_score = dev(norm["score"], (1, -1, 1))
            ref_lns = torch.tensor([ref_cvec.size(1)], dtype=torch.long, device=device)
            shifted_lf0 = linear_midi_shift(lf0, ref["f0"])
            with torch.no_grad():
                mel = model.inference(x, lengths, scores, lfts, ref_cvec, ref_lns, ref_mel, ref_score, ref_lft,
                                      n_timesteps=n_timesteps, temperature=temperature)
                wav, _sr = vocoder.decode(mel.squeeze(0) if mel.dim() == 3 else mel)
            name = f"{utt_id}_{style}"
            write_hdf5(os.path.join(outdir, name + (out_ext or os.path.splitext(str(ref_path))[1] or ".h5")), "lf0",
                       shifted_lf0.astype(np.float32))
            write_wav_pcm16(os.path.join(outdir, f"{name}.wav"), wav.detach().cpu().numpy(), sampling_rate)
            written.append(name)
    return written
