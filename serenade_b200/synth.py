"""Synthetic ssc1 weights and decoder inputs (no checkpoints or datasets exist offline).

Weights follow the reference's own random init (`Decoder.initialize_weights`,
reference decoder.py:366-382): Kaiming-normal Conv1d/Linear with zero bias,
GroupNorm/LayerNorm gamma=1 beta=0, and torch's default init for the
ConvTranspose1d upsampler, which `initialize_weights` does not touch.  They are
drawn from numpy's PCG64 stream so that the build container and the GPU box
produce bit-identical tensors without shipping 294 MB of weights.

`jitter=True` additionally randomises every bias and norm affine parameter so
that parity tests exercise bias / gamma / beta handling (the stock init leaves
them at 0 / 1, which would hide bugs).

Inputs follow SURVEY.md section 8d (value ranges of the real features).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch


@dataclass(frozen=True)
class DecoderGeometry:
    """Architecture of the CFM estimator (reference serenade.py:77-87, conf/serenade.yaml:27-34)."""
    in_channels: int = 242          # 80 (x) + 162 (mu)
    out_channels: int = 80
    spk_embed_dim: int = 256
    channels: Tuple[int, ...] = (512, 512)
    attention_head_dim: int = 512
    num_heads: int = 4
    num_mid_blocks: int = 2
    ff_mult: int = 4

    @property
    def mu_channels(self) -> int:
        return self.in_channels - self.out_channels

    @property
    def time_embed_dim(self) -> int:
        return self.channels[0] * 4


SSC1 = DecoderGeometry()


def _kaiming(rng, shape, fan_in):
    std = math.sqrt(2.0 / fan_in)
    return torch.from_numpy((rng.standard_normal(shape, dtype=np.float32) * np.float32(std)))


def _uniform(rng, shape, bound):
    return torch.from_numpy(rng.uniform(-bound, bound, size=shape).astype(np.float32))


def estimator_param_shapes(g: DecoderGeometry = SSC1) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape, in the reference's state_dict order (prefix `estimator.`)."""
    C = g.channels[0]
    assert all(c == C for c in g.channels), "ssc1-style decoders use one width"
    inner = g.num_heads * g.attention_head_dim
    ffi = C * g.ff_mult
    te = g.time_embed_dim
    out: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()

    def lin(name, o, i, bias=True):
        out[name + ".weight"] = (o, i)
        if bias:
            out[name + ".bias"] = (o,)

    def conv(name, o, i, k):
        out[name + ".weight"] = (o, i, k)
        out[name + ".bias"] = (o,)

    def norm(name, c):
        out[name + ".weight"] = (c,)
        out[name + ".bias"] = (c,)

    def resnet(p, cin):
        lin(p + ".mlp.1", C, te)
        conv(p + ".block1.block.0", C, cin, 3); norm(p + ".block1.block.1", C)
        conv(p + ".block2.block.0", C, C, 3); norm(p + ".block2.block.1", C)
        conv(p + ".res_conv", C, cin, 1)
        lin(p + ".speaker_projection.W_scale", C, g.spk_embed_dim)
        lin(p + ".speaker_projection.W_bias", C, g.spk_embed_dim)

    def tfm(p):
        norm(p + ".norm1", C)
        lin(p + ".attn1.to_q", inner, C, bias=False)
        lin(p + ".attn1.to_k", inner, C, bias=False)
        lin(p + ".attn1.to_v", inner, C, bias=False)
        lin(p + ".attn1.to_out.0", C, inner)
        norm(p + ".norm3", C)
        lin(p + ".ff.net.0.proj", 2 * ffi, C)
        lin(p + ".ff.net.2", C, ffi)

    lin("time_mlp.linear_1", te, g.in_channels)
    lin("time_mlp.linear_2", te, te)
    n = len(g.channels)
    for i in range(n):
        p = f"down_blocks.{i}"
        resnet(p + ".0", g.in_channels if i == 0 else C)
        tfm(p + ".1.0")
        conv(p + (".2.conv" if i < n - 1 else ".2"), C, C, 3)
    for i in range(g.num_mid_blocks):
        resnet(f"mid_blocks.{i}.0", C)
        tfm(f"mid_blocks.{i}.1.0")
    for i in range(n):
        p = f"up_blocks.{i}"
        resnet(p + ".0", 2 * C)
        tfm(p + ".1.0")
        if i < n - 1:
            out[p + ".2.conv.weight"] = (C, C, 4)   # ConvTranspose1d: (in, out, k)
            out[p + ".2.conv.bias"] = (C,)
        else:
            conv(p + ".2", C, C, 3)
    conv("final_block.block.0", C, C, 3); norm("final_block.block.1", C)
    conv("final_proj", g.out_channels, C, 1)
    return OrderedDict(("estimator." + k, v) for k, v in out.items())


def make_state_dict(seed: int = 0, jitter: bool = False, g: DecoderGeometry = SSC1) -> "OrderedDict[str, torch.Tensor]":
    """Random-init CFM state dict with the reference's key layout (`estimator.*`)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for key, shape in estimator_param_shapes(g).items():
        is_norm = (".block.1." in key) or (".norm1." in key) or (".norm3." in key)
        is_convT = key.startswith("estimator.up_blocks.") and ".2.conv." in key
        if key.endswith(".weight"):
            if is_norm:
                w = torch.ones(shape)
                if jitter:
                    w = w + 0.2 * torch.from_numpy(rng.standard_normal(shape, dtype=np.float32))
            elif is_convT:
                fan_in = shape[1] * shape[2]
                w = _uniform(rng, shape, 1.0 / math.sqrt(fan_in))
            else:
                fan_in = int(np.prod(shape[1:]))
                w = _kaiming(rng, shape, fan_in)
            sd[key] = w
        else:
            if is_convT:
                C = shape[0]
                sd[key] = _uniform(rng, shape, 1.0 / math.sqrt(C * 4))
            elif jitter:
                sd[key] = 0.1 * torch.from_numpy(rng.standard_normal(shape, dtype=np.float32))
            else:
                sd[key] = torch.zeros(shape)
    return sd


def state_dict_checksum(sd) -> float:
    """Order-dependent fingerprint used by the golden fixtures to prove identical weights."""
    acc = 0.0
    for i, (k, v) in enumerate(sd.items()):
        acc += float(v.double().sum()) * (1.0 + 0.001 * i) + float(v.double().abs().sum()) * 1e-3
    return acc


@dataclass
class DecodeBatch:
    mu: torch.Tensor        # (B, 162, T) f32
    mask: torch.Tensor      # (B, 1, T) bool, True = valid frame
    spks: torch.Tensor      # (B, 256) f32
    z: torch.Tensor         # (B, 80, T) f32, pre-drawn noise * temperature
    lengths: List[int]
    t_span: torch.Tensor    # (n_steps + 1,) f32

    @property
    def n_steps(self) -> int:
        return self.t_span.numel() - 1


def make_batch(lengths: Sequence[int], n_steps: int = 10, seed: int = 1234,
               infill: Optional[Tuple[float, float]] = (0.25, 0.75), temperature: float = 0.667,
               g: DecoderGeometry = SSC1, t_max: Optional[int] = None) -> DecodeBatch:
    """Synthetic decoder inputs.  `infill` = fraction range of each clip whose
    context-mel channels are exactly zero (the frames to generate, reference serenade.py:193-199)."""
    lengths = [int(l) for l in lengths]
    B = len(lengths)
    T = int(t_max) if t_max is not None else max(lengths)
    rng = np.random.Generator(np.random.PCG64(seed))
    nm = g.out_channels
    mu = np.zeros((B, g.mu_channels, T), dtype=np.float32)
    mu[:, :nm] = rng.standard_normal((B, nm, T), dtype=np.float32)             # content-encoder output
    score = rng.uniform(0.0, 1.0, size=(B, T)).astype(np.float32)
    score[rng.uniform(size=(B, T)) < 0.3] = 0.0                                  # rests / unvoiced
    mu[:, nm] = score
    mu[:, nm + 1] = rng.uniform(0.0, 1.0, size=(B, T)).astype(np.float32)       # loudness
    ctx = rng.standard_normal((B, g.mu_channels - nm - 2, T), dtype=np.float32)
    for b, L in enumerate(lengths):
        if infill is not None:
            a, e = int(round(infill[0] * L)), int(round(infill[1] * L))
            ctx[b, :, a:e] = 0.0
    mu[:, nm + 2:] = ctx
    spk = (0.5 * np.tanh(rng.standard_normal((B, g.spk_embed_dim), dtype=np.float32))).astype(np.float32)
    z = rng.standard_normal((B, g.out_channels, T), dtype=np.float32) * np.float32(temperature)
    mask = (np.arange(T)[None, :] < np.asarray(lengths)[:, None])[:, None, :]
    for b, L in enumerate(lengths):     # padded frames of mu are zeros, as a pad collater leaves them
        mu[b, :, L:] = 0.0
    return DecodeBatch(mu=torch.from_numpy(mu), mask=torch.from_numpy(mask), spks=torch.from_numpy(spk),
                       z=torch.from_numpy(z), lengths=lengths,
                       t_span=torch.linspace(0, 1, n_steps + 1, dtype=torch.float32))


def estimator_flops(L: int) -> float:
    """Algorithmic FLOPs of one estimator call on one utterance of L valid frames."""
    L2 = (L + 1) // 2
    return 2.0 * (20_983_808 * L + 39_321_600 * L2) + 8192.0 * (2 * L * L + 4 * L2 * L2) + 25.1e6


# ------------------------------------------------------------------------------------------------------------
# Conditioning front half: content encoder + GST style encoder of `Serenade` (serenade.py:59-73)

def frontend_param_shapes(input_dim: int = 768, hidden: int = 512, out_dim: int = 80,
                          gst_chans: Sequence[int] = (128, 128, 256, 256, 512, 512), gst_tokens: int = 50,
                          gst_dim: int = 256, gru_units: int = 128, heads: int = 4) -> "OrderedDict[str, Tuple[int, ...]]":
    """Keys / shapes of `encoder.*` and `gst.*` in the reference `Serenade` state dict (old-style weight_norm:
    weight_g (O,1,1) + weight_v (O,I,k))."""
    out: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()

    def wn(name, o, i, k):
        out[name + ".bias"] = (o,)
        out[name + ".weight_g"] = (o, 1, 1)
        out[name + ".weight_v"] = (o, i, k)

    wn("encoder.model.1", hidden, input_dim, 7)
    for idx in (2, 3):
        wn(f"encoder.model.{idx}.block.2", hidden, hidden, 3)
        wn(f"encoder.model.{idx}.block.4", hidden, hidden, 1)
        wn(f"encoder.model.{idx}.shortcut", hidden, hidden, 1)
    wn("encoder.model.6", out_dim, hidden, 7)
    cin, f = 1, 80
    for i, c in enumerate(gst_chans):
        out[f"gst.ref_enc.convs.{3 * i}.weight"] = (c, cin, 3, 3)
        for nm in ("weight", "bias", "running_mean", "running_var"):
            out[f"gst.ref_enc.convs.{3 * i + 1}.{nm}"] = (c,)
        out[f"gst.ref_enc.convs.{3 * i + 1}.num_batches_tracked"] = ()
        cin, f = c, (f - 1) // 2 + 1
    gin = f * gst_chans[-1]
    out["gst.ref_enc.gru.weight_ih_l0"] = (3 * gru_units, gin)
    out["gst.ref_enc.gru.weight_hh_l0"] = (3 * gru_units, gru_units)
    out["gst.ref_enc.gru.bias_ih_l0"] = (3 * gru_units,)
    out["gst.ref_enc.gru.bias_hh_l0"] = (3 * gru_units,)
    dk = gst_dim // heads
    out["gst.stl.gst_embs"] = (gst_tokens, dk)
    for nm, i in (("linear_q", gru_units), ("linear_k", dk), ("linear_v", dk), ("linear_out", gst_dim)):
        out[f"gst.stl.mha.{nm}.weight"] = (gst_dim, i)
        out[f"gst.stl.mha.{nm}.bias"] = (gst_dim,)
    return out


def make_frontend_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Random `encoder.*` / `gst.*` parameters in torch's default-init ranges (uniform +-1/sqrt(fan_in); weight_g
    jittered around ||v||; BatchNorm running statistics away from (0, 1) so that eval-mode folding is exercised)."""
    rng = np.random.Generator(np.random.PCG64(seed + 7919))
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    shapes = frontend_param_shapes()
    for key, shape in shapes.items():
        if key.endswith("num_batches_tracked"):
            sd[key] = torch.tensor(100, dtype=torch.long)
        elif key.endswith("weight_g"):
            continue            # filled after its weight_v
        elif key.endswith("weight_v"):
            fan_in = shape[1] * shape[2]
            v = _uniform(rng, shape, 1.0 / math.sqrt(fan_in))
            sd[key] = v
            g = v.flatten(1).norm(dim=1).view(-1, 1, 1)
            sd[key[:-1] + "g"] = g * torch.from_numpy(rng.uniform(0.8, 1.25, size=(shape[0], 1, 1)).astype(np.float32))
        elif key.endswith("running_var"):
            sd[key] = torch.from_numpy(rng.uniform(0.5, 1.5, size=shape).astype(np.float32))
        elif key.endswith("running_mean"):
            sd[key] = 0.1 * torch.from_numpy(rng.standard_normal(shape, dtype=np.float32))
        elif ".convs." in key and key.endswith(".weight") and len(shape) == 1:
            sd[key] = torch.from_numpy(rng.uniform(0.8, 1.2, size=shape).astype(np.float32))
        elif key == "gst.stl.gst_embs":
            sd[key] = torch.from_numpy(rng.standard_normal(shape, dtype=np.float32))
        elif len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            if ".convs." in key:
                sd[key] = _kaiming(rng, shape, fan_in)          # keeps ReLU activations O(1) through six layers
            else:
                sd[key] = _uniform(rng, shape, 1.0 / math.sqrt(fan_in))
        else:
            sd[key] = 0.1 * torch.from_numpy(rng.standard_normal(shape, dtype=np.float32))
    return OrderedDict((k, sd[k]) for k in shapes)


def make_frontend_inputs(src_len: int, ref_len: int, seed: int = 99, batch: int = 1, input_dim: int = 768):
    """Synthetic `Serenade.inference` inputs: cvec-like content features, MIDI score and loudness in [0, 1]
    (audio_mel_dataset.py:104-110), reference log-mel ~ N(0, 1) (normalised features)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = lambda *s: torch.from_numpy(rng.standard_normal(s, dtype=np.float32))
    u = lambda *s: torch.from_numpy(rng.uniform(0.0, 1.0, size=s).astype(np.float32))
    return dict(x=n(batch, src_len, input_dim), lengths=torch.full((batch,), src_len, dtype=torch.long),
                midi=u(batch, src_len, 1), lft=u(batch, src_len, 1),
                ref_x=n(batch, ref_len, input_dim), ref_lengths=torch.full((batch,), ref_len, dtype=torch.long),
                ref_logmel=n(batch, ref_len, 80), ref_midi=u(batch, ref_len, 1), ref_lft=u(batch, ref_len, 1))


# ----------------------------------------------------------------------------------------------------------
# HiFi-GAN vocoder: parameter shapes of reference serenade/vocoder/models/hifigan.py:21-161 after
# remove_weight_norm(), and a seeded synthetic state dict whose activations stay O(1) through all stages (the
# reference's own reset_parameters draws N(0, 0.01), under which the signal decays to 1e-3: fine for parity, useless
# as a numerical stress).
def vocoder_param_shapes(in_channels: int = 80, out_channels: int = 1, channels: int = 512, kernel_size: int = 7,
                         upsample_scales: Sequence[int] = (8, 8, 2, 2),
                         upsample_kernel_sizes: Optional[Sequence[int]] = None,
                         resblock_kernel_sizes: Sequence[int] = (3, 7, 11),
                         n_dilations: int = 3) -> "OrderedDict[str, Tuple[int, ...]]":
    if upsample_kernel_sizes is None:
        upsample_kernel_sizes = [2 * s for s in upsample_scales]
    sh: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    sh["input_conv.weight"] = (channels, in_channels, kernel_size)
    sh["input_conv.bias"] = (channels,)
    for i, (s, k) in enumerate(zip(upsample_scales, upsample_kernel_sizes)):
        ci, co = channels // (2 ** i), channels // (2 ** (i + 1))
        sh[f"upsamples.{i}.1.weight"] = (ci, co, k)          # ConvTranspose1d: (in, out, k)
        sh[f"upsamples.{i}.1.bias"] = (co,)
        for j, rk in enumerate(resblock_kernel_sizes):
            b = i * len(resblock_kernel_sizes) + j
            for d in range(n_dilations):
                for cv in ("convs1", "convs2"):
                    sh[f"blocks.{b}.{cv}.{d}.1.weight"] = (co, co, rk)
                    sh[f"blocks.{b}.{cv}.{d}.1.bias"] = (co,)
    cl = channels // (2 ** len(upsample_scales))
    sh["output_conv.1.weight"] = (out_channels, cl, kernel_size)
    sh["output_conv.1.bias"] = (out_channels,)
    return sh


def make_vocoder_state_dict(seed: int = 0, **geometry) -> "OrderedDict[str, torch.Tensor]":
    rng = np.random.Generator(np.random.PCG64(1000 + seed))
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for key, shape in vocoder_param_shapes(**geometry).items():
        if key.endswith(".bias"):
            a = rng.uniform(-0.05, 0.05, size=shape)
        elif key.startswith("upsamples."):
            stride = shape[2] // 2
            a = rng.standard_normal(shape) * np.sqrt(1.5 * stride / (shape[0] * shape[2]))    # taps per output = k / stride
        elif ".convs2." in key:
            a = rng.standard_normal(shape) * (0.5 / np.sqrt(shape[1] * shape[2]))             # residual branch: small
        else:
            a = rng.standard_normal(shape) * np.sqrt(1.5 / (shape[1] * shape[2]))
        sd[key] = torch.from_numpy(a.astype(np.float32))
    return sd


def make_vocoder_input(frames: int, batch: int = 1, seed: int = 0, in_channels: int = 80) -> torch.Tensor:
    """(B, T, 80) normalised mel, the layout `Vocoder.decode_batch` receives (vocoder.py:64)."""
    rng = np.random.Generator(np.random.PCG64(2000 + seed))
    return torch.from_numpy(rng.standard_normal((batch, frames, in_channels), dtype=np.float32))

