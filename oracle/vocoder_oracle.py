"""CPU oracle for the HiFi-GAN vocoder that follows the CFM decoder in `serenade-decode`.

TEST INFRASTRUCTURE ONLY (same rule as oracle/cfm_oracle.py: nothing under serenade_b200/ imports this file).

Functional restatement (plain torch fp32, state-dict driven, weight norm already removed) of

  hifigan_forward   <- serenade/vocoder/models/hifigan.py:167-186 (HiFiGANGenerator.forward) with the layers
                       built at :71-161: input Conv1d k7; per stage LeakyReLU(0.1) + ConvTranspose1d(k = 2 s,
                       stride s, padding s // 2 + s % 2, output_padding s % 2), then the mean of the
                       HiFiGANResidualBlocks; LeakyReLU() (slope 0.01, sic) + Conv1d k7 + Tanh
  _resblock         <- serenade/vocoder/layers/residual_block.py:143-258 (convs1: dilated, convs2: dilation 1)
  vocoder_decode    <- serenade/vocoder/vocoder.py:51-62 (Vocoder.decode: de-normalise with the target statistics,
                       normalise with the vocoder's, generate, flatten)

PARITY PINNING: the reference has no tests or fixtures for the vocoder; the oracle is pinned against outputs of the
reference's own HiFiGANGenerator module executed in the build container (oracle/ref_loader.py:
load_reference_hifigan, tests/golden/make_golden_vocoder.py -> tests/golden/vocoder_*.npz, checked by
tests/test_vocoder_oracle.py).
"""
from __future__ import annotations

from typing import Dict, Sequence

import torch
import torch.nn.functional as F


def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """`weight_g` / `weight_v` pairs (torch.nn.utils.weight_norm, dim 0) -> plain `weight`, as
    HiFiGANGenerator.remove_weight_norm does (hifigan.py:203-213)."""
    out = {}
    for k, v in sd.items():
        if k.endswith(".weight_v"):
            g = sd[k[:-1] + "g"]
            n = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.dim() - 1)))
            out[k[:-9] + ".weight"] = v * (g / n)
        elif not k.endswith(".weight_g"):
            out[k] = v
    return out


def _resblock(sd, p, x, dilations: Sequence[int]):
    for i, d in enumerate(dilations):
        w1, b1 = sd[f"{p}.convs1.{i}.1.weight"], sd[f"{p}.convs1.{i}.1.bias"]
        w2, b2 = sd[f"{p}.convs2.{i}.1.weight"], sd[f"{p}.convs2.{i}.1.bias"]
        k = w1.shape[-1]
        xt = F.conv1d(F.leaky_relu(x, 0.1), w1, b1, dilation=d, padding=(k - 1) // 2 * d)
        xt = F.conv1d(F.leaky_relu(xt, 0.1), w2, b2, padding=(k - 1) // 2)
        x = xt + x
    return x


def hifigan_forward(sd: Dict[str, torch.Tensor], c: torch.Tensor, dilations: Sequence[int] = (1, 3, 5)) -> torch.Tensor:
    """c (B, 80, T) normalised mel -> (B, 1, T * prod(upsample_scales)).  The structure (stages, strides = kernel / 2,
    residual blocks per stage, kernel sizes) is read off the state dict."""
    n_up = sum(1 for k in sd if k.startswith("upsamples.") and k.endswith(".1.weight"))
    n_blk_total = sum(1 for k in sd if k.startswith("blocks.") and k.endswith(".convs1.0.1.weight"))
    n_blk = n_blk_total // n_up
    k_in = sd["input_conv.weight"].shape[-1]
    x = F.conv1d(c, sd["input_conv.weight"], sd["input_conv.bias"], padding=(k_in - 1) // 2)
    for i in range(n_up):
        w = sd[f"upsamples.{i}.1.weight"]
        s = w.shape[-1] // 2
        x = F.conv_transpose1d(F.leaky_relu(x, 0.1), w, sd[f"upsamples.{i}.1.bias"], stride=s,
                               padding=s // 2 + s % 2, output_padding=s % 2)
        cs = 0.0
        for j in range(n_blk):
            cs = cs + _resblock(sd, f"blocks.{i * n_blk + j}", x, dilations)
        x = cs / n_blk
    k_out = sd["output_conv.1.weight"].shape[-1]
    x = F.conv1d(F.leaky_relu(x), sd["output_conv.1.weight"], sd["output_conv.1.bias"], padding=(k_out - 1) // 2)
    return torch.tanh(x)


def vocoder_decode(sd, mel, stats_mean, stats_scale, trg_mean=None, trg_scale=None, dilations=(1, 3, 5)):
    """mel (B, T, 80) as the acoustic model returns it -> wave (B, T * hop).  vocoder.py:51-62 / :64-75."""
    c = mel
    if trg_mean is not None:
        c = c * trg_scale + trg_mean
    c = (c - stats_mean) / stats_scale
    with torch.inference_mode():
        return hifigan_forward(sd, c.transpose(1, 2), dilations).squeeze(1)
