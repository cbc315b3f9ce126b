"""`serenade_b200.vocoder`: HiFi-GAN generator and the `Vocoder` wrapper of the decode script on the B200 library.

Mirrors reference serenade/vocoder/models/hifigan.py:21-284 (`HiFiGANGenerator`: same constructor kwargs, same
`state_dict` keys with or without weight norm, `inference` / `inference_batch`) and serenade/vocoder/vocoder.py:12-75
(`Vocoder.decode` / `decode_batch`).  All arithmetic runs in libserenade_b200.so (`srn_vocoder_decode`,
csrc/engine_vocoder.inl); there is no PyTorch fallback.  Training-side classes (discriminators) are out of scope.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import torch
import torch.nn as nn

from . import _lib
from .cfm import _attach
from .synth import vocoder_param_shapes

__all__ = ["HiFiGANGenerator", "Vocoder"]


def _fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """weight_g / weight_v (torch.nn.utils.weight_norm, dim 0) -> weight, what remove_weight_norm() leaves
    (hifigan.py:203-213; `Vocoder.__init__` calls it before decoding, vocoder.py:38)."""
    out = {}
    for k, v in sd.items():
        if k.endswith(".weight_v"):
            g = sd[k[:-1] + "g"]
            n = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.dim() - 1)))
            out[k[:-9] + ".weight"] = v * (g / n)
        elif not k.endswith(".weight_g"):
            out[k] = v
    return out


class HiFiGANGenerator(nn.Module):
    """Reference hifigan.py:21-284, inference side.  Non-causal, LeakyReLU, bias=True (the reference's defaults and
    the gtsinger vocoder's configuration)."""

    def __init__(self, in_channels=80, out_channels=1, channels=512, kernel_size=7, upsample_scales=(8, 8, 2, 2),
                 upsample_kernel_sizes=(16, 16, 4, 4), resblock_kernel_sizes=(3, 7, 11),
                 resblock_dilations=((1, 3, 5), (1, 3, 5), (1, 3, 5)), use_additional_convs=True, bias=True,
                 nonlinear_activation="LeakyReLU", nonlinear_activation_params=None, use_causal_conv=False,
                 use_weight_norm=True, precision="bf16"):
        super().__init__()
        params = {"negative_slope": 0.1} if nonlinear_activation_params is None else dict(nonlinear_activation_params)
        if (out_channels != 1 or not use_additional_convs or not bias or use_causal_conv
                or nonlinear_activation != "LeakyReLU" or abs(params.get("negative_slope", 0.01) - 0.1) > 1e-12):
            raise ValueError("only the non-causal LeakyReLU(0.1) generator with additional convs, bias and one output "
                             "channel is implemented")
        dil = [tuple(int(x) for x in d) for d in resblock_dilations]
        if len(dil) != len(resblock_kernel_sizes) or any(d != dil[0] for d in dil):
            raise ValueError("all residual blocks must share one dilation list")
        if any(int(k) != 2 * int(s) for k, s in zip(upsample_kernel_sizes, upsample_scales)):
            raise ValueError("upsample_kernel_sizes must be 2 * upsample_scales (hifigan.py:95)")
        self.in_channels = in_channels
        self.upsample_scales = tuple(int(s) for s in upsample_scales)
        self.dilations = dil[0]
        self.hop = 1
        for s in self.upsample_scales:
            self.hop *= s
        self.precision = {"bf16": _lib.PRECISION_BF16, "tf32": _lib.PRECISION_TF32}[precision]
        shapes = vocoder_param_shapes(in_channels=in_channels, out_channels=out_channels, channels=channels,
                                      kernel_size=kernel_size, upsample_scales=self.upsample_scales,
                                      upsample_kernel_sizes=upsample_kernel_sizes,
                                      resblock_kernel_sizes=resblock_kernel_sizes, n_dilations=len(self.dilations))
        for key, shape in shapes.items():
            _attach(self, key, nn.Parameter(torch.zeros(shape), requires_grad=False))
        with torch.no_grad():                       # reset_parameters (hifigan.py:188-201): N(0, 0.01) weights
            for name, p in self.named_parameters():
                if name.endswith("weight"):
                    p.normal_(0.0, 0.01)
        self._handle: Optional[C.c_void_p] = None
        self._device: Optional[torch.device] = None
        self._dirty = True
        self.register_load_state_dict_post_hook(lambda module, _keys: module.mark_weights_dirty())
        self._register_load_state_dict_pre_hook(self._pre_load)

    # the reference's checkpoints hold weight_g / weight_v (weight norm is removed after loading, vocoder.py:36-38)
    @staticmethod
    def _pre_load(state_dict, prefix, *_):
        if any(k.startswith(prefix) and k.endswith(".weight_v") for k in state_dict):
            sub = {k[len(prefix):]: state_dict.pop(k) for k in [k for k in state_dict if k.startswith(prefix)]}
            for k, v in _fold_weight_norm(sub).items():
                state_dict[prefix + k] = v

    def remove_weight_norm(self):
        """No-op: weights are stored folded (see _pre_load)."""

    def apply_weight_norm(self):
        raise NotImplementedError("training is outside the B200 decode path")

    def register_stats(self, stats):
        raise NotImplementedError("pass mean / scale tensors to Vocoder(...) instead (h5py is not a dependency)")

    def mark_weights_dirty(self):
        self._dirty = True

    def _apply(self, fn, *args, **kwargs):
        self._dirty = True
        return super()._apply(fn, *args, **kwargs)

    def close(self):
        if self._handle is not None:
            _lib.load_library().srn_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _engine(self, device: torch.device):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("serenade_b200 runs on a B200 (sm_100a) only; there is no CPU path. "
                               f"Got tensors on '{device}'.")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        lib = _lib.load_library()
        if self._handle is None or self._device != device or self._dirty:
            self.close()
            cfg = _lib.SrnConfig(242, 80, 256, 512, 4, 512, 2, self.precision)      # decoder geometry unused here
            h = C.c_void_p()
            with torch.cuda.device(device):
                _lib.check(lib.srn_create(C.byref(cfg), C.byref(h)), "srn_create")
                d = (C.c_int32 * len(self.dilations))(*self.dilations)
                _lib.check(lib.srn_vocoder_set_dilations(h, d, len(self.dilations)), "srn_vocoder_set_dilations")
                for key, v in self.state_dict().items():
                    t = v.detach().to("cpu", torch.float32).contiguous()
                    dims = (C.c_int64 * t.dim())(*t.shape)
                    _lib.check(lib.srn_load_weight(h, ("vocoder." + key).encode(), C.c_void_p(t.data_ptr()), t.dim(), dims),
                               f"srn_load_weight(vocoder.{key})")
                _lib.check(lib.srn_finalize_weights(h), "srn_finalize_weights")
            self._handle, self._device, self._dirty = h, device, False
        return lib, self._handle

    @torch.inference_mode()
    def decode(self, mel: torch.Tensor, scale: Optional[torch.Tensor] = None, shift: Optional[torch.Tensor] = None):
        """mel (B, T, in_channels) on the GPU -> wave (B, T * hop).  Optional per-channel `mel * scale + shift` first."""
        if mel.dim() != 3 or mel.shape[-1] != self.in_channels:
            raise ValueError(f"mel must be (B, T, {self.in_channels}), got {tuple(mel.shape)}")
        lib, h = self._engine(mel.device)
        dev = mel.device
        m = mel.detach().to(torch.float32).contiguous()
        B, T, _ = m.shape
        wave = torch.empty(B, T * self.hop, dtype=torch.float32, device=dev)
        sc = sh = None
        if scale is not None:
            sc = scale.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
            sh = shift.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
            if sc.numel() != self.in_channels or sh.numel() != self.in_channels:
                raise ValueError("scale / shift must have one value per mel channel")
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(lib.srn_vocoder_decode(h, m.data_ptr(), sc.data_ptr() if sc is not None else None,
                                              sh.data_ptr() if sh is not None else None, B, T, wave.data_ptr(),
                                              C.c_void_p(st)), "srn_vocoder_decode")
        return wave

    def forward(self, c):
        """c (B, in_channels, T) -> (B, 1, T * hop), reference hifigan.py:167-186."""
        return self.decode(c.transpose(1, 2)).unsqueeze(1)

    def inference(self, c, normalize_before=False):
        """c (T, in_channels) -> (T * hop, 1), reference hifigan.py:247-266."""
        if normalize_before:
            raise NotImplementedError("normalize with Vocoder(...) (statistics are not stored in the generator here)")
        if not isinstance(c, torch.Tensor):
            c = torch.tensor(c, dtype=torch.float)
        if c.device.type != "cuda":
            c = c.to(self._device or torch.device("cuda", torch.cuda.current_device()))
        return self.decode(c.unsqueeze(0)).squeeze(0).unsqueeze(-1)

    def inference_batch(self, c, normalize_before=False):
        """c (B, T, in_channels) -> (B, 1, T * hop), reference hifigan.py:268-284."""
        if normalize_before:
            raise NotImplementedError("normalize with Vocoder(...)")
        return self.decode(c).unsqueeze(1)


def load_vocoder(checkpoint, config=None, precision="bf16") -> "HiFiGANGenerator":
    """Reference vocoder/utils.py:14-47 `load_vocoder`: generator built from `config["generator_params"]` (a dict, a yaml
    path, or config.yml next to the checkpoint), parameters from `torch.load(checkpoint)["model"]["generator"]`."""
    import os
    if config is None:
        config = os.path.join(os.path.dirname(checkpoint), "config.yml")
    if not isinstance(config, dict):
        import yaml
        with open(config) as f:
            config = yaml.load(f, Loader=yaml.Loader)
    if config.get("generator_type", "HiFiGANGenerator") != "HiFiGANGenerator":
        raise ValueError(f"only HiFiGANGenerator is implemented, got {config.get('generator_type')}")
    params = {k.replace("upsample_kernal_sizes", "upsample_kernel_sizes"): v for k, v in config["generator_params"].items()}
    model = HiFiGANGenerator(**params, precision=precision)
    state = torch.load(checkpoint, map_location="cpu")
    model.load_state_dict(state["model"]["generator"], strict=True)
    return model


def read_stats(path) -> Dict[str, torch.Tensor]:
    """{"mean", "scale"} from the reference's stats.h5 (datasets "mean" / "scale", utils.py read_hdf5; needs h5py), from
    ParallelWaveGAN's stats.npy ((2, n): mean, scale) or from an .npz with those two arrays."""
    import numpy as np
    path = str(path)
    if path.endswith((".h5", ".hdf5")):
        try:
            import h5py
        except ImportError as e:
            raise RuntimeError("reading stats.h5 needs h5py, which this environment does not have: convert the statistics "
                               "to .npz (mean, scale) or pass them as a dict") from e
        with h5py.File(path, "r") as f:
            mean, scale = np.asarray(f["mean"]), np.asarray(f["scale"])
    elif path.endswith(".npz"):
        z = np.load(path)
        mean, scale = z["mean"], z["scale"]
    else:
        a = np.load(path)
        mean, scale = a[0], a[1]
    return {"mean": torch.as_tensor(mean, dtype=torch.float32), "scale": torch.as_tensor(scale, dtype=torch.float32)}


class Vocoder:
    """Reference vocoder/vocoder.py:12-75.  Two call forms:
      Vocoder(checkpoint, config, stats, device, trg_stats=None, take_norm_feat=True)   the reference's own (file paths:
          checkpoint .pkl, config .yml, stats .h5 / .npy / .npz), as bin/ssc_decode.py:310-318 calls it;
      Vocoder(model, stats, device, trg_stats=None, take_norm_feat=True, sampling_rate=24000)   a loaded
          HiFiGANGenerator and {"mean", "scale"} dicts."""

    def __init__(self, *args, **kwargs):
        import os
        first = args[0] if args else kwargs.get("checkpoint", kwargs.get("model"))
        if isinstance(first, (str, os.PathLike)):
            self._init_from_files(*args, **kwargs)
        else:
            self._init_from_objects(*args, **kwargs)

    def _init_from_files(self, checkpoint, config, stats, device, trg_stats=None, take_norm_feat=True, precision="bf16"):
        import yaml
        with open(config) as f:
            self.config = yaml.load(f, Loader=yaml.Loader)
        model = load_vocoder(checkpoint, self.config, precision=precision)
        st = stats if isinstance(stats, dict) else read_stats(stats)
        self._init_from_objects(model, st, device, trg_stats=trg_stats, take_norm_feat=take_norm_feat,
                                sampling_rate=self.config.get("sampling_rate", 24000))

    def _init_from_objects(self, model: HiFiGANGenerator, stats: Dict[str, Sequence[float]], device, trg_stats=None,
                           take_norm_feat=True, sampling_rate=24000):
        if take_norm_feat and trg_stats is None:
            raise AssertionError("trg_stats must be given if take_norm_feat=True")
        self.device = torch.device(device)
        self.model = model
        self.sampling_rate = sampling_rate
        self.take_norm_feat = take_norm_feat
        mean = torch.as_tensor(stats["mean"], dtype=torch.float32).reshape(-1)
        scale = torch.as_tensor(stats["scale"], dtype=torch.float32).reshape(-1)
        if take_norm_feat:
            tm = torch.as_tensor(trg_stats["mean"], dtype=torch.float32).reshape(-1)
            ts = torch.as_tensor(trg_stats["scale"], dtype=torch.float32).reshape(-1)
            # (c * ts + tm - mean) / scale  (vocoder.py:52-56)
            self._scale, self._shift = (ts / scale).to(self.device), ((tm - mean) / scale).to(self.device)
        else:
            self._scale, self._shift = (1.0 / scale).to(self.device), (-mean / scale).to(self.device)

    def decode(self, c):
        """c (T, 80) -> (wave (T * hop,), sampling rate), vocoder.py:51-62."""
        y = self.model.decode(c.to(self.device).unsqueeze(0), self._scale, self._shift).view(-1)
        return y, self.sampling_rate

    def decode_batch(self, c):
        """c (B, T, 80) -> (B, T * hop), vocoder.py:64-75."""
        return self.model.decode(c.to(self.device), self._scale, self._shift)
