"""CPU oracle for the conditioning front half of `Serenade.inference`.

TEST INFRASTRUCTURE ONLY (same rules as oracle/cfm_oracle.py: only tests/, smoke() and bench.py's CPU legs may
import it, and only as the checker / timed baseline).

Functional restatement (plain torch fp32, state-dict driven) of

  content_encoder  <- serenade/models/serenade.py:224-376  Conv1dResnet(768 -> 512 -> 80, 2 ResnetBlocks):
                      ReflectionPad1d(3) + weight-normed Conv1d k7; ResnetBlock(dilation 1, 2) =
                      shortcut 1x1 + [LeakyReLU(0.2), ReflectionPad1d(d), k3 dilated, LeakyReLU(0.2), 1x1]
                      (:361-376); LeakyReLU(0.2) + ReflectionPad1d(3) + Conv1d k7 (:289-303).
                      weight_norm: w = g * v / ||v||, norm over (in, k) per output channel.
  style_encoder    <- serenade/modules/gst/style_encoder.py:16-252: ReferenceEncoder (6 x [Conv2d k3 s2 p1 no
                      bias, BatchNorm2d (eval), ReLU], GRU(1024 -> 128) last hidden state) and StyleTokenLayer
                      (multi-head attention of the reference embedding over tanh(gst_embs), 4 heads x 64,
                      serenade/modules/gst/attention.py:111-147, :316-319).
  assemble_mu      <- serenade/models/serenade.py:181-206 (reference prompt + zeroed conditioning, concatenated
                      in time) and the mask of :208-209.

PARITY PINNING: no reference tests / fixtures exist for these modules either; the oracle is pinned against the
reference's own module files run in the build container (tests/golden/make_golden_frontend.py ->
tests/golden/frontend_*.npz, tests/test_frontend_oracle.py).
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F


def _wn(sd: Dict[str, torch.Tensor], p: str) -> torch.Tensor:
    v, g = sd[p + ".weight_v"], sd[p + ".weight_g"]
    return g * v / v.flatten(1).norm(dim=1).view(-1, 1, 1)


def content_encoder(sd, x: torch.Tensor, prefix: str = "encoder.model") -> torch.Tensor:
    """x (B, L, in_dim) -> (B, L, 80).  No masking: padded frames take part (and reflection is about L_max),
    exactly like the reference module."""
    h = x.transpose(1, 2)
    h = F.conv1d(F.pad(h, (3, 3), mode="reflect"), _wn(sd, f"{prefix}.1"), sd[f"{prefix}.1.bias"])
    for n, idx in enumerate((2, 3)):
        d = 2 ** n
        p = f"{prefix}.{idx}"
        a = F.pad(F.leaky_relu(h, 0.2), (d, d), mode="reflect")
        c = F.conv1d(a, _wn(sd, p + ".block.2"), sd[p + ".block.2.bias"], dilation=d)
        f = F.conv1d(F.leaky_relu(c, 0.2), _wn(sd, p + ".block.4"), sd[p + ".block.4.bias"])
        h = F.conv1d(h, _wn(sd, p + ".shortcut"), sd[p + ".shortcut.bias"]) + f
    h = F.pad(F.leaky_relu(h, 0.2), (3, 3), mode="reflect")
    out = F.conv1d(h, _wn(sd, f"{prefix}.6"), sd[f"{prefix}.6.bias"])
    return out.transpose(1, 2)


def reference_encoder(sd, mel: torch.Tensor, prefix: str = "gst.ref_enc") -> torch.Tensor:
    """mel (B, L, 80) -> (B, 128): conv stack + GRU last hidden state (style_encoder.py:167-190)."""
    x = mel.unsqueeze(1)
    for i in range(6):
        c, b = f"{prefix}.convs.{3 * i}", f"{prefix}.convs.{3 * i + 1}"
        x = F.conv2d(x, sd[c + ".weight"], None, stride=2, padding=1)
        x = F.batch_norm(x, sd[b + ".running_mean"], sd[b + ".running_var"], sd[b + ".weight"], sd[b + ".bias"],
                         training=False, eps=1e-5)
        x = F.relu(x)
    B, C, Lp, Fp = x.shape
    hs = x.transpose(1, 2).contiguous().view(B, Lp, C * Fp)
    w_ih, w_hh = sd[prefix + ".gru.weight_ih_l0"], sd[prefix + ".gru.weight_hh_l0"]
    b_ih, b_hh = sd[prefix + ".gru.bias_ih_l0"], sd[prefix + ".gru.bias_hh_l0"]
    H = w_hh.shape[1]
    h = torch.zeros(B, H)
    for t in range(Lp):          # torch.nn.GRU gate order: r, z, n
        gi = F.linear(hs[:, t], w_ih, b_ih)
        gh = F.linear(h, w_hh, b_hh)
        r = torch.sigmoid(gi[:, :H] + gh[:, :H])
        z = torch.sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
        n = torch.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
        h = (1 - z) * n + z * h
    return h


def style_token_layer(sd, ref: torch.Tensor, prefix: str = "gst.stl", heads: int = 4) -> torch.Tensor:
    """ref (B, 128) -> (B, 256) (style_encoder.py:236-252)."""
    tok = torch.tanh(sd[prefix + ".gst_embs"])                                   # (50, 64)
    q = F.linear(ref, sd[prefix + ".mha.linear_q.weight"], sd[prefix + ".mha.linear_q.bias"])
    k = F.linear(tok, sd[prefix + ".mha.linear_k.weight"], sd[prefix + ".mha.linear_k.bias"])
    v = F.linear(tok, sd[prefix + ".mha.linear_v.weight"], sd[prefix + ".mha.linear_v.bias"])
    B, D = q.shape
    dk = D // heads
    qh = q.view(B, heads, dk)
    kh = k.view(-1, heads, dk).transpose(0, 1)                                   # (h, 50, dk)
    vh = v.view(-1, heads, dk).transpose(0, 1)
    sc = torch.einsum("bhd,hnd->bhn", qh, kh) / math.sqrt(dk)
    p = torch.softmax(sc, dim=-1)
    ctx = torch.einsum("bhn,hnd->bhd", p, vh).reshape(B, D)
    return F.linear(ctx, sd[prefix + ".mha.linear_out.weight"], sd[prefix + ".mha.linear_out.bias"])


def style_encoder(sd, mel: torch.Tensor) -> torch.Tensor:
    return style_token_layer(sd, reference_encoder(sd, mel))


def assemble_mu(enc_src, midi, lft, ref_enc, ref_midi, ref_lft, ref_logmel):
    """serenade.py:186-206: (B, L_ref + L_src, 162) conditioning, reference prompt first, source frames with a
    zero context mel; returned channels-first (B, 162, T) as CFM.inference takes it."""
    ref = torch.cat([ref_enc, ref_midi, ref_lft, ref_logmel], dim=-1)
    src = torch.cat([enc_src, midi, lft, torch.zeros(enc_src.shape[0], enc_src.shape[1], ref_logmel.shape[-1])], dim=-1)
    return torch.cat([ref, src], dim=1).permute(0, 2, 1).contiguous()
