"""CPU oracle for the CFM mel-infilling decoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under serenade_b200/ imports this file; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may call it, and only as the checker or the timed CPU baseline.

It is a from-scratch functional restatement (plain torch fp32 on CPU, driven by a
state dict, no nn.Module) of the reference algorithm:

  solve_euler        <- serenade/models/matcha_components/flow_matching.py:65-93
  cfm_inference      <- flow_matching.py:39-63
  estimator_forward  <- decoder.py:384-467 (Decoder.forward)
  _resnet            <- decoder.py:80-101 (ResnetBlock1D), :66-77 (Block1D)
  _speaker_adapter   <- decoder.py:14-45
  _time_embedding    <- decoder.py:48-63 (SinusoidalPosEmb), :113-157 (TimestepEmbedding)
  _transformer       <- transformer.py:266-354 live branch (:286, :292-304, :322, :347, :352)
  _attention/_geglu  <- diffusers Attention / GEGLU as called at transformer.py:120, :212-220.
                        diffusers is an UNPINNED, un-vendored dependency (reference
                        setup.cfg:21); its published semantics are restated here.

PARITY PINNING: the reference ships no tests, golden vectors or fixtures for this
path, so the oracle is pinned against outputs of the
reference's own three module files executed in the build container
(oracle/ref_loader.py + tests/golden/make_golden.py -> tests/golden/*.npz, checked by
tests/test_oracle_golden.py; and directly module-vs-oracle in
tests/test_oracle_vs_reference.py whenever /root/reference is present).  The
diffusers pieces are pinned only by their restated semantics: "parity unpinned"
for that third-party boundary.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Optional

import torch
import torch.nn.functional as F

P = "estimator."


def _time_embedding(sd, t: torch.Tensor, in_channels: int) -> torch.Tensor:
    # decoder.py:54-63: scale=1000, half_dim = dim // 2, emb = log(1e4)/(half-1)
    if t.ndim < 1:
        t = t.unsqueeze(0)
    half = in_channels // 2
    e = math.log(10000) / (half - 1)
    freqs = torch.exp(torch.arange(half, device=t.device).float() * -e)
    emb = 1000 * t.unsqueeze(1) * freqs.unsqueeze(0)
    emb = torch.cat((emb.sin(), emb.cos()), dim=-1)
    # decoder.py:145-157: linear_1 -> SiLU -> linear_2
    h = F.linear(emb, sd[P + "time_mlp.linear_1.weight"], sd[P + "time_mlp.linear_1.bias"])
    h = F.silu(h)
    return F.linear(h, sd[P + "time_mlp.linear_2.weight"], sd[P + "time_mlp.linear_2.bias"])


def _block1d(sd, p, x, mask):
    # decoder.py:75-77: Mish(GroupNorm8(Conv1d_k3_p1(x*mask))) * mask ; GN spans all T_max frames
    y = F.conv1d(x * mask, sd[p + ".block.0.weight"], sd[p + ".block.0.bias"], padding=1)
    y = F.group_norm(y, 8, sd[p + ".block.1.weight"], sd[p + ".block.1.bias"], eps=1e-5)
    return F.mish(y) * mask


def _speaker_adapter(sd, p, x, spk):
    # decoder.py:34-45: per-frame LN over channels (biased var, eps 1e-5, no affine) * scale + bias
    xt = x.transpose(1, -1)
    mean = xt.mean(dim=-1, keepdim=True)
    var = ((xt - mean) ** 2).mean(dim=-1, keepdim=True)
    y = (xt - mean) / (var + 1e-5).sqrt()
    scale = F.linear(spk, sd[p + ".W_scale.weight"], sd[p + ".W_scale.bias"])
    bias = F.linear(spk, sd[p + ".W_bias.weight"], sd[p + ".W_bias.bias"])
    y = y * scale.unsqueeze(1) + bias.unsqueeze(1)
    return y.transpose(1, -1)


def _resnet(sd, p, x, mask, temb, spk):
    # decoder.py:95-101
    h = _block1d(sd, p + ".block1", x, mask)
    h = h + F.linear(F.mish(temb), sd[p + ".mlp.1.weight"], sd[p + ".mlp.1.bias"]).unsqueeze(-1)
    h = _block1d(sd, p + ".block2", h, mask)
    out = h + F.conv1d(x * mask, sd[p + ".res_conv.weight"], sd[p + ".res_conv.bias"])
    return _speaker_adapter(sd, p + ".speaker_projection", out, spk)


def _attention(sd, p, x, key_mask, heads):
    # diffusers Attention (self-attention, bias-free q/k/v, boolean key-padding mask, True = attend)
    b, s, _ = x.shape
    q = F.linear(x, sd[p + ".to_q.weight"]).view(b, s, heads, -1).transpose(1, 2)
    k = F.linear(x, sd[p + ".to_k.weight"]).view(b, s, heads, -1).transpose(1, 2)
    v = F.linear(x, sd[p + ".to_v.weight"]).view(b, s, heads, -1).transpose(1, 2)
    d = q.shape[-1]
    scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d)
    scores = scores.masked_fill(~key_mask.view(b, 1, 1, s), float("-inf"))
    o = torch.matmul(torch.softmax(scores, dim=-1), v)
    o = o.transpose(1, 2).reshape(b, s, -1)
    return F.linear(o, sd[p + ".to_out.0.weight"], sd[p + ".to_out.0.bias"])


def _geglu_ff(sd, p, x):
    # transformer.py:120,138 + diffusers GEGLU: proj -> chunk -> h * gelu_erf(gate) -> Linear
    pr = F.linear(x, sd[p + ".net.0.proj.weight"], sd[p + ".net.0.proj.bias"])
    h, g = pr.chunk(2, dim=-1)
    return F.linear(h * F.gelu(g), sd[p + ".net.2.weight"], sd[p + ".net.2.bias"])


def _transformer(sd, p, x, key_mask, heads):
    # transformer.py:286-304 (self-attn + residual), :322-352 (FF + residual); dropout inert in eval
    n1 = F.layer_norm(x, x.shape[-1:], sd[p + ".norm1.weight"], sd[p + ".norm1.bias"], eps=1e-5)
    x = _attention(sd, p + ".attn1", n1, key_mask, heads) + x
    n3 = F.layer_norm(x, x.shape[-1:], sd[p + ".norm3.weight"], sd[p + ".norm3.bias"], eps=1e-5)
    return _geglu_ff(sd, p + ".ff", n3) + x


def estimator_forward(sd: Dict[str, torch.Tensor], x, mask, mu, t, spk, heads: int = 4,
                      tap: Optional[Callable[[str, torch.Tensor], None]] = None) -> torch.Tensor:
    """v = Decoder.forward(x, mask, mu, t, spk)  (decoder.py:384-467).

    x (B,80,T) f32, mask (B,1,T) bool, mu (B,C_mu,T) f32, t 0-d f32, spk (B,256) -> (B,80,T).
    `tap(name, tensor)` receives named intermediates (used by the stage-by-stage GPU tests).
    """
    tap = tap or (lambda n, v: None)
    mask = mask.bool()
    in_ch = x.shape[1] + mu.shape[1]
    temb = _time_embedding(sd, torch.as_tensor(t, dtype=torch.float32), in_ch)
    T = x.shape[-1]
    h = torch.cat([x, mu], dim=1)
    n_down = sum(1 for k in sd if k.startswith(P + "down_blocks.") and k.endswith(".0.mlp.1.weight"))
    n_mid = sum(1 for k in sd if k.startswith(P + "mid_blocks.") and k.endswith(".0.mlp.1.weight"))
    hiddens, masks = [], [mask]
    for i in range(n_down):
        p = f"{P}down_blocks.{i}"
        m = masks[-1]
        h = _resnet(sd, p + ".0", h, m, temb, spk); tap(f"down{i}.resnet", h)
        h = _transformer(sd, p + ".1.0", h.transpose(1, 2), m[:, 0], heads).transpose(1, 2); tap(f"down{i}.tfm", h)
        hiddens.append(h)
        if i < n_down - 1:   # Downsample1D: Conv1d k3 s2 p1 (decoder.py:104-110)
            h = F.conv1d(h * m, sd[p + ".2.conv.weight"], sd[p + ".2.conv.bias"], stride=2, padding=1)
        else:                # plain k3 conv (decoder.py:271)
            h = F.conv1d(h * m, sd[p + ".2.weight"], sd[p + ".2.bias"], padding=1)
        tap(f"down{i}.out", h)
        masks.append(m[:, :, ::2])
    masks = masks[:-1]
    m = masks[-1]
    for i in range(n_mid):
        p = f"{P}mid_blocks.{i}"
        h = _resnet(sd, p + ".0", h, m, temb, spk); tap(f"mid{i}.resnet", h)
        h = _transformer(sd, p + ".1.0", h.transpose(1, 2), m[:, 0], heads).transpose(1, 2); tap(f"mid{i}.tfm", h)
    for i in range(n_down):
        p = f"{P}up_blocks.{i}"
        m = masks.pop()
        h = h[:, :, :T]                                   # decoder.py:447
        h = torch.cat([h, hiddens.pop()], dim=1)          # decoder.py:449
        h = _resnet(sd, p + ".0", h, m, temb, spk); tap(f"up{i}.resnet", h)
        h = _transformer(sd, p + ".1.0", h.transpose(1, 2), m[:, 0], heads).transpose(1, 2); tap(f"up{i}.tfm", h)
        if i < n_down - 1:   # ConvTranspose1d(4, 2, 1), weight (in, out, k) (decoder.py:191)
            h = F.conv_transpose1d(h * m, sd[p + ".2.conv.weight"], sd[p + ".2.conv.bias"], stride=2, padding=1)
        else:
            h = F.conv1d(h * m, sd[p + ".2.weight"], sd[p + ".2.bias"], padding=1)
        tap(f"up{i}.out", h)
    h = _block1d(sd, P + "final_block", h, m); tap("final_block", h)
    out = F.conv1d(h * m, sd[P + "final_proj.weight"], sd[P + "final_proj.bias"])
    return out * mask


def solve_euler(sd, x, t_span, mu, mask, spk, heads: int = 4, return_all: bool = False):
    """flow_matching.py:65-93: x += dt * v; t += dt; dt = t_span[k+1] - t (fp32 running t)."""
    t, dt = t_span[0], t_span[1] - t_span[0]
    sol = []
    for step in range(1, len(t_span)):
        v = estimator_forward(sd, x, mask, mu, t, spk, heads)
        x = x + dt * v
        t = t + dt
        sol.append(x)
        if step < len(t_span) - 1:
            dt = t_span[step + 1] - t
    return sol if return_all else sol[-1]


def cfm_inference(sd, mu, mask, n_timesteps=10, temperature=0.667, spks=None, out_channels=80, z=None):
    """flow_matching.py:39-63 (noise drawn on the CPU generator unless pre-drawn `z` is given)."""
    if z is None:
        z = torch.randn((mu.shape[0], out_channels, mu.shape[2])) * temperature
    t_span = torch.linspace(0, 1, n_timesteps + 1)
    with torch.inference_mode():
        return solve_euler(sd, z, t_span, mu, mask, spks)


def compute_loss(sd, x1, mask, mu, spks, z, t, mask_l=None, sigma_min: float = 1e-4, heads: int = 4):
    """flow_matching.py:95-133 with the random draws made explicit: `t` (B,) is what `torch.rand([b, 1, 1])` gave and
    `z` what `torch.randn_like(x1)` gave.  Returns (loss, y) like the reference."""
    t = t.reshape(-1, 1, 1).to(torch.float32)
    y = (1 - (1 - sigma_min) * t) * z + t * x1
    u = x1 - (1 - sigma_min) * z
    denoised = estimator_forward(sd, y, mask, mu, t.squeeze(), spks, heads)
    if mask_l is not None:
        denoised = denoised * mask_l
        u = u * mask_l
        den = torch.sum(mask_l) * u.shape[1]
    else:
        den = torch.sum(mask) * u.shape[1]
    loss = F.mse_loss(denoised, u, reduction="sum") / den
    return loss, y


def time_grid(t_span: torch.Tensor):
    """The (t_k, dt_k) sequence solve_euler visits, with the reference's fp32 running-sum wobble."""
    t, dt = t_span[0].clone(), t_span[1] - t_span[0]
    ts, dts = [], []
    for step in range(1, len(t_span)):
        ts.append(float(t)); dts.append(float(dt))
        t = t + dt
        if step < len(t_span) - 1:
            dt = t_span[step + 1] - t
    return ts, dts
