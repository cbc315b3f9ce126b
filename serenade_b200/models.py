"""`serenade_b200.models`: the reference's model registry namespace for the decode path.

`bin/ssc_decode.py:337-340` selects the model with `getattr(serenade.models, config["model_type"])(**config[
"model_params"])`, loads `checkpoint["model"]` with `load_state_dict` and calls `model.inference(...)`
(`:428-438`).  `Serenade` below has the same constructor kwargs, the same `state_dict` keys (`encoder.*`, `gst.*`,
`cfm_decoder.estimator.*`) and the same `inference` signature / return value as reference
serenade/models/serenade.py:35-221, with every stage running in libserenade_b200.so:

    content encoder (Conv1dResnet, :224-376)     -> srn_content_encode  (tcgen05 GEMM core, reflection-padded operands)
    GST style encoder (modules/gst/style_encoder) -> srn_style_encode    (im2col GEMMs + GRU / style-token kernel)
    CFM Euler sampler + UNet estimator            -> srn_solve_euler     (serenade_b200.cfm)

PyTorch only concatenates the conditioning tensors (serenade.py:186-206).  Training (`forward`) is out of scope.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from . import _lib
from .cfm import CFM, _ParamNode
from .synth import frontend_param_shapes

__all__ = ["Serenade"]


def _attach_any(root: nn.Module, dotted: str, t: torch.Tensor, buffer: bool) -> None:
    parts = dotted.split(".")
    node = root
    for name in parts[:-1]:
        if name not in node._modules:
            node.add_module(name, _ParamNode())
        node = node._modules[name]
    if buffer:
        node.register_buffer(parts[-1], t)
    else:
        node.register_parameter(parts[-1], nn.Parameter(t, requires_grad=False))


class Serenade(nn.Module):
    """Drop-in for reference `serenade.models.Serenade` (inference side)."""

    def __init__(self, input_dim=768, output_dim=80, encoder_channels=80, decoder_channels=512, gst_embed_dim=256,
                 decoder_attention_head_dim=512, mask_size=(0.1, 0.5), cfg_prob=0.1, precision="bf16",
                 shape_bucket=128):
        """The reference's constructor kwargs (serenade.py:36-47) + two of this build: `precision` ("bf16" | "tf32") and
        `shape_bucket` (frames; 0 = off): the CFM decode of one unpadded utterance / of a valid-frame batch runs on T
        rounded up to the bucket, so a decode loop over arbitrary lengths reuses a handful of engine plans instead of
        building one (workspace + TMA maps + CUDA graph) per distinct length.  Results do not change, bit for bit."""
        super().__init__()
        if output_dim != 80 or gst_embed_dim != 256:
            raise ValueError("this build is specialised for 80-bin mels and a 256-dim style embedding")
        self.input_dim, self.output_dim, self.cfg_prob, self.mask_size = input_dim, output_dim, cfg_prob, list(mask_size)
        self.encoder_channels = encoder_channels
        for key, shape in frontend_param_shapes(input_dim=input_dim, out_dim=encoder_channels,
                                                gst_dim=gst_embed_dim).items():
            is_buf = key.endswith(("running_mean", "running_var", "num_batches_tracked"))
            if key.endswith("num_batches_tracked"):
                t = torch.zeros((), dtype=torch.long)
            elif key.endswith("running_var") or (".convs." in key and key.endswith(".weight") and len(shape) == 1):
                t = torch.ones(shape)
            elif key.endswith(("bias", "running_mean")):
                t = torch.zeros(shape)
            else:
                t = torch.randn(shape) * (1.0 / max(1, int(torch.tensor(shape[1:]).prod())) ** 0.5)
            _attach_any(self, key, t, is_buf)
        conditioning_dim = output_dim + encoder_channels + 1 + 1
        self.cfm_decoder = CFM(in_channels=conditioning_dim + output_dim, out_channels=output_dim,
                               spk_embed_dim=gst_embed_dim, decoder_channels=(decoder_channels, decoder_channels),
                               decoder_attention_head_dim=decoder_attention_head_dim, precision=precision,
                               shape_bucket=shape_bucket)
        self._front_refs = None
        self.cfm_decoder.estimator._extra_state = self._frontend_state
        self.register_load_state_dict_post_hook(lambda module, _keys: module.cfm_decoder.estimator.mark_weights_dirty())

    # ------------------------------------------------------------------ engine plumbing
    def _frontend_state(self) -> Dict[str, torch.Tensor]:
        """encoder.* / gst.* tensors by reference key (0-d BatchNorm counters are not part of the C ABI).  Looked up
        through (module, attribute) pairs cached once: `state_dict()` per call costs more than the kernels at B=1."""
        if self._front_refs is None:
            refs = []
            for prefix in ("encoder", "gst"):
                root = self._modules[prefix]
                for mname, mod in root.named_modules():
                    for kind in (mod._parameters, mod._buffers):
                        for pname in kind:
                            refs.append((".".join(x for x in (prefix, mname, pname) if x), kind, pname))
            self._front_refs = refs
        return {key: store[name] for key, store, name in self._front_refs if store[name].dim() > 0}

    def _apply(self, fn, *args, **kwargs):
        self.cfm_decoder.estimator.mark_weights_dirty()
        return super()._apply(fn, *args, **kwargs)

    def _engine(self, device):
        return self.cfm_decoder.estimator.engine(device)

    def _run(self, fn_name: str, x: torch.Tensor, out_shape):
        if x.device.type != "cuda":
            raise RuntimeError("serenade_b200 runs on a B200 (sm_100a) only; there is no CPU path")
        eng = self._engine(x.device)
        xf = x.detach().to(torch.float32).contiguous()
        out = torch.empty(out_shape, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            st = torch.cuda.current_stream(x.device).cuda_stream
            _lib.check(getattr(eng.lib, fn_name)(eng.h, xf.data_ptr(), xf.shape[0], xf.shape[1], out.data_ptr(),
                                                 C.c_void_p(st)), fn_name)
        return out

    @torch.inference_mode()
    def encode_content(self, x: torch.Tensor, lengths=None) -> torch.Tensor:
        """`Serenade.encoder(x, lengths)`: (B, L, input_dim) -> (B, L, encoder_channels).  `lengths` is unused, as in
        the reference (Conv1dResnet.forward ignores it, serenade.py:322-352)."""
        if x.dim() != 3 or x.shape[2] != self.input_dim:
            raise ValueError(f"x must be (B, L, {self.input_dim}), got {tuple(x.shape)}")
        return self._run("srn_content_encode", x, (x.shape[0], x.shape[1], self.encoder_channels))

    @torch.inference_mode()
    def encode_style(self, logmel: torch.Tensor) -> torch.Tensor:
        """`Serenade.gst(logmel)`: (B, L, 80) -> (B, 256)."""
        if logmel.dim() != 3 or logmel.shape[2] != self.output_dim:
            raise ValueError(f"logmel must be (B, L, {self.output_dim}), got {tuple(logmel.shape)}")
        return self._run("srn_style_encode", logmel, (logmel.shape[0], 256))

    # ------------------------------------------------------------------ reference API
    def forward(self, *a, **k):
        raise NotImplementedError("training (Serenade.forward / CFM.compute_loss) is outside the B200 decode path; "
                                  "train with the reference and load the checkpoint here")

    @torch.inference_mode()
    def inference(self, x, lengths, midi, lft, ref_x, ref_lengths, ref_logmel, ref_midi, ref_lft,
                  n_timesteps=10, temperature=0.667):
        """Reference serenade.py:168-221.  Returns the source part of the predicted mel, squeezed like the reference."""
        dev = x.device
        enc_outs_ = self.encode_content(x, lengths)
        ref_speaker_features = self.encode_style(ref_logmel)
        cond = ref_logmel.to(torch.float32)
        ref_enc_outs = self.encode_content(ref_x, ref_lengths)
        ref_enc_outs = torch.cat([ref_enc_outs, ref_midi.float(), ref_lft.float(), cond], dim=-1)
        zero_cond = torch.zeros(enc_outs_.shape[0], enc_outs_.shape[1], cond.shape[-1], device=dev)
        enc_outs = torch.cat([enc_outs_, midi.float(), lft.float(), zero_cond], dim=-1)
        enc_outs = torch.cat([ref_enc_outs, enc_outs], dim=1)
        total_lengths = (lengths + ref_lengths).to(dev)
        T = enc_outs.shape[1]
        mask = (torch.arange(T, device=dev)[None, :] < total_lengths[:, None]).unsqueeze(1)     # make_non_pad_mask
        # One unpadded utterance (the reference CLI's call, bin/ssc_decode.py:428): valid-frame GroupNorm statistics are
        # the reference's statistics, so the decode may run on a length bucket (CFM.shape_bucket) and reuse its plan.
        est = self.cfm_decoder.estimator
        saved = est.gn_stats
        if self.cfm_decoder.shape_bucket > 0 and enc_outs.shape[0] == 1 and int(total_lengths[0]) == T:
            est.gn_stats = "valid"
        try:
            mel_pred = self.cfm_decoder.inference(enc_outs.permute(0, 2, 1), mask, n_timesteps=n_timesteps,
                                                  temperature=temperature, spks=ref_speaker_features).permute(0, 2, 1)
        finally:
            est.gn_stats = saved
        mel_pred = mel_pred[:, int(ref_lengths[0]):, :]
        return mel_pred.squeeze(0)

    @torch.inference_mode()
    def inference_batch(self, items: Sequence[Sequence[torch.Tensor]], n_timesteps=10, temperature=0.667,
                        max_batch=16, max_padded_frames: Optional[int] = None) -> List[torch.Tensor]:
        """Length-bucketed batched decode: `items[i]` holds the positional arguments of `inference`
        for ONE utterance (batch dimension 1); returns `[self.inference(*it) for it in items]` without the reference's
        batch-1 restriction (the FIXMEs at serenade.py:209-210, :218: its concat / cut assume one reference length).

        The encoders run per utterance (their reflection padding and the GRU depend on the true lengths); the flow-
        matching decode, which is all the time, runs on padded batches of similar length with GroupNorm statistics
        over each utterance's own frames (`gn_stats="valid"`), so every mel equals the one `inference` returns for that
        utterance alone.  Noise is drawn per utterance, in order, with the same generator calls as sequential
        `inference` calls make (flow_matching.py:57)."""
        from . import sharding
        if len(items) == 0:
            return []
        dev = items[0][0].device
        clips, ref_lens = [], []
        for it in items:
            x, lengths, midi, lft, ref_x, ref_lengths, ref_logmel, ref_midi, ref_lft = it
            if x.shape[0] != 1:
                raise ValueError("inference_batch takes one utterance per item (batch dimension 1)")
            enc = self.encode_content(x, lengths)
            spk = self.encode_style(ref_logmel)
            cond = ref_logmel.to(torch.float32)
            ref_enc = torch.cat([self.encode_content(ref_x, ref_lengths), ref_midi.float(), ref_lft.float(), cond], dim=-1)
            zero_cond = torch.zeros(1, enc.shape[1], cond.shape[-1], device=dev)
            mu = torch.cat([ref_enc, torch.cat([enc, midi.float(), lft.float(), zero_cond], dim=-1)], dim=1)
            total = int(lengths[0]) + int(ref_lengths[0])
            mu = mu[0, :total].transpose(0, 1).contiguous()                   # (162, T_i); frames past the lengths are padding
            z = torch.randn((1, self.cfm_decoder.out_channels, mu.shape[1])).to(dev) * temperature
            clips.append(dict(mu=mu, z=z[0], spk=spk[0]))
            ref_lens.append(int(ref_lengths[0]))
        lens = [int(c["mu"].shape[-1]) for c in clips]
        t_span = torch.linspace(0, 1, n_timesteps + 1, device=dev)
        est = self.cfm_decoder.estimator
        saved = est.gn_stats
        est.gn_stats = "valid"
        out: List[Optional[torch.Tensor]] = [None] * len(clips)
        try:
            for batch in sharding.make_batches(range(len(clips)), lens, max_batch=max_batch,
                                               max_padded_frames=max_padded_frames):
                T = max(lens[i] for i in batch)
                z = torch.zeros(len(batch), clips[batch[0]]["z"].shape[0], T, device=dev)
                mu = torch.zeros(len(batch), clips[batch[0]]["mu"].shape[0], T, device=dev)
                mask = torch.zeros(len(batch), 1, T, dtype=torch.bool, device=dev)
                for b, i in enumerate(batch):
                    z[b, :, :lens[i]] = clips[i]["z"]
                    mu[b, :, :lens[i]] = clips[i]["mu"]
                    mask[b, :, :lens[i]] = True
                spk = torch.stack([clips[i]["spk"] for i in batch])
                mel = self.cfm_decoder.solve_euler(z, t_span, mu, mask, spk)
                for b, i in enumerate(batch):
                    out[i] = mel[b, :, ref_lens[i]:lens[i]].transpose(0, 1).contiguous()
        finally:
            est.gn_stats = saved
        return out  # type: ignore[return-value]

