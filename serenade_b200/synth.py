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
