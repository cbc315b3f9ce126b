"""Stage-by-stage comparison of the CUDA path with the oracle (shared by the GPU tests and tests/diagnose.py).

The C ABI's debug hook (srn_debug_enable / srn_debug_fetch) keeps named intermediates of one estimator call
as channels-last (nb, rows, cols) f32; the oracle's `tap` callback yields the matching block outputs, and
`fine_reference` recomputes the first block's inner tensors in plain torch so that a failure can be pinned
to one kernel from a single GPU run.
"""
import math

import torch
import torch.nn.functional as F

from oracle import cfm_oracle as O


def rel(a, b):
    a, b = a.double(), b.double()
    d = float(b.norm())
    return float((a - b).norm() / (d if d > 0 else 1.0))


def oracle_stages(sd, batch, t):
    """name -> (B, rows, C) channels-last f32 block outputs of the oracle, plus 'v' (B, 80, T)."""
    taps = {}
    with torch.inference_mode():
        v = O.estimator_forward(sd, batch.z, batch.mask, batch.mu, torch.tensor(t), batch.spks,
                                tap=lambda n, x: taps.__setitem__(n, x.transpose(1, 2).contiguous()))
    taps["v"] = v
    return taps


def level_mask(batch, rows):
    m = batch.mask[:, 0]
    if rows != m.shape[1]:
        m = m[:, ::2]
    assert m.shape[1] == rows, (m.shape, rows)
    return m.unsqueeze(-1).float()


def fine_block(sd, prefix, tag, x, mask, temb, spks):
    """Inner tensors (channels-last) of one resnet + transformer block `prefix` whose input is x (B, C_in, S) at the
    rate of `mask` (B, 1, S): {tag}.y1/r/a1/y2/resnet/qk/P/ao/h_attn/ln3/gg, in torch fp32 from the state dict."""
    out = {}
    xm = x * mask
    y1 = F.conv1d(xm, sd[prefix + ".0.block1.block.0.weight"], sd[prefix + ".0.block1.block.0.bias"], padding=1)
    out[tag + ".y1"] = y1.transpose(1, 2)
    out[tag + ".r"] = F.conv1d(xm, sd[prefix + ".0.res_conv.weight"], sd[prefix + ".0.res_conv.bias"]).transpose(1, 2)
    h = F.mish(F.group_norm(y1, 8, sd[prefix + ".0.block1.block.1.weight"], sd[prefix + ".0.block1.block.1.bias"])) * mask
    h = h + F.linear(F.mish(temb), sd[prefix + ".0.mlp.1.weight"], sd[prefix + ".0.mlp.1.bias"]).unsqueeze(-1)
    out[tag + ".a1"] = (h * mask).transpose(1, 2)
    y2 = F.conv1d(h * mask, sd[prefix + ".0.block2.block.0.weight"], sd[prefix + ".0.block2.block.0.bias"], padding=1)
    out[tag + ".y2"] = y2.transpose(1, 2)
    res = O._resnet(sd, prefix + ".0", x, mask, temb, spks)
    out[tag + ".resnet"] = res.transpose(1, 2)
    hs = res.transpose(1, 2)
    tp = prefix + ".1.0"
    n1 = F.layer_norm(hs, (512,), sd[tp + ".norm1.weight"], sd[tp + ".norm1.bias"])
    q = F.linear(n1, sd[tp + ".attn1.to_q.weight"])
    k = F.linear(n1, sd[tp + ".attn1.to_k.weight"])
    v = F.linear(n1, sd[tp + ".attn1.to_v.weight"])
    out[tag + ".qk"] = torch.cat([q, k], dim=-1)
    B, S, _ = q.shape
    qh, kh, vh = (z.view(B, S, 4, 512).transpose(1, 2) for z in (q, k, v))
    sc = torch.matmul(qh, kh.transpose(-1, -2)) / math.sqrt(512)
    scm = sc.masked_fill(~mask[:, 0].view(B, 1, 1, S), float("-inf"))
    pr = torch.softmax(scm, dim=-1)
    out[tag + ".P"] = pr.reshape(B * 4, S, S)
    ao = torch.matmul(pr, vh).transpose(1, 2).reshape(B, S, 2048)
    out[tag + ".ao"] = ao
    h_attn = F.linear(ao, sd[tp + ".attn1.to_out.0.weight"], sd[tp + ".attn1.to_out.0.bias"]) + hs
    out[tag + ".h_attn"] = h_attn
    n3 = F.layer_norm(h_attn, (512,), sd[tp + ".norm3.weight"], sd[tp + ".norm3.bias"])
    out[tag + ".ln3"] = n3
    pj = F.linear(n3, sd[tp + ".ff.net.0.proj.weight"], sd[tp + ".ff.net.0.proj.bias"])
    hh, gg = pj.chunk(2, dim=-1)
    out[tag + ".gg"] = hh * F.gelu(gg)
    return out


def fine_reference(sd, batch, t, stages=None):
    """Inner tensors of down_blocks.0 (full rate, 242 input channels) and - given the oracle's block outputs `stages` -
    of down_blocks.1 (half rate) and up_blocks.1 (the 1024-channel skip-concat resnet), each block fed with the
    ORACLE's input so that a stage's error is its own; plus the hoisted time path (a4): the time embedding and the six
    per-resnet time biases."""
    out = {}
    with torch.inference_mode():
        mask = batch.mask
        x = torch.cat([batch.z, batch.mu], dim=1)
        out["xin"] = F.pad((x * mask).transpose(1, 2), (0, 256 - x.shape[1]))
        temb = O._time_embedding(sd, torch.tensor(t), x.shape[1])
        out["time_emb"] = temb.unsqueeze(0)
        names = ["down_blocks.0", "down_blocks.1", "mid_blocks.0", "mid_blocks.1", "up_blocks.0", "up_blocks.1"]
        out["time_bias"] = torch.cat([F.linear(F.mish(temb), sd[f"estimator.{n}.0.mlp.1.weight"], sd[f"estimator.{n}.0.mlp.1.bias"])
                                      for n in names], dim=0).unsqueeze(0)
        out.update(fine_block(sd, "estimator.down_blocks.0", "down0", x, mask, temb, batch.spks))
        if stages is not None:
            T = x.shape[-1]
            m2 = mask[:, :, ::2]
            out.update(fine_block(sd, "estimator.down_blocks.1", "down1", stages["down0.out"].transpose(1, 2), m2, temb,
                                  batch.spks))
            xin = torch.cat([stages["up0.out"].transpose(1, 2)[:, :, :T], stages["down0.tfm"].transpose(1, 2)], dim=1)
            out.update(fine_block(sd, "estimator.up_blocks.1", "up1", xin, mask, temb, batch.spks))
    return out


BLOCK_ORDER = ["down0.resnet", "down0.tfm", "down0.out", "down1.resnet", "down1.tfm", "down1.out",
               "mid0.resnet", "mid0.tfm", "mid1.resnet", "mid1.tfm", "up0.resnet", "up0.tfm", "up0.out",
               "up1.resnet", "up1.tfm", "up1.out", "final_block"]
_FINE_TAPS = ["y1", "r", "a1", "y2", "resnet", "qk", "P", "ao", "h_attn", "ln3", "gg"]
FINE_ORDER = ["xin", "time_emb", "time_bias"] + [f"{blk}.{t}" for blk in ("down0", "down1", "up1") for t in _FINE_TAPS]
# outputs that the CUDA path stores already masked (every consumer masks them anyway)
MASKED_TAPS = {"tfm", "out"}


def compare(cfm, sd, batch, t, fine=True):
    """Runs one debug forward on the GPU and returns [(name, rel_l2, note)] in execution order + v error."""
    dev = torch.device("cuda", torch.cuda.current_device())
    est = cfm.estimator
    eng = est.engine(dev)
    eng.debug_enable(True)
    try:
        v = est(batch.z.to(dev), batch.mask.to(dev), batch.mu.to(dev), torch.tensor(t), batch.spks.to(dev)).cpu()
        rows = []
        stages = oracle_stages(sd, batch, t)
        if fine:
            ref = fine_reference(sd, batch, t, stages)
            for name in FINE_ORDER:
                got = eng.debug_fetch(name)
                want = ref[name]
                if name in ("time_emb", "time_bias"):
                    rows.append((name, rel(got, want), "hoisted time path (a4)"))
                    continue
                is_p = name.endswith(".P")
                valid = level_mask(batch, want.shape[1]) if not is_p else None
                if is_p:
                    # the CUDA path keeps P~ = exp(s - rowmax) unnormalised (the row sum is divided out in the
                    # P~ V epilogue): normalise here; compare valid query rows and valid keys only
                    m = batch.mask[:, 0].float()
                    if want.shape[1] != m.shape[1]:
                        m = m[:, ::2]
                    mm = (m.unsqueeze(2) * m.unsqueeze(1)).repeat_interleave(4, dim=0)
                    got = got * mm
                    got = got / got.sum(-1, keepdim=True).clamp_min(1e-30)
                    rows.append((name, rel(got * mm, want * mm), "valid q,k, rows normalised on the host"))
                else:
                    rows.append((name, rel(got * valid, want * valid), "valid rows"))
        for name in BLOCK_ORDER:
            got = eng.debug_fetch(name)
            want = stages[name]
            want = want[:, :got.shape[1]]          # up0.out: the oracle keeps the 2*ceil(T/2) frame before trimming
            valid = level_mask(batch, got.shape[1])
            rows.append((name, rel(got * valid, want * valid), "valid rows"))
        rows.append(("v", rel(v, stages["v"]), "full padded tensor"))
    finally:
        eng.debug_enable(False)
    return rows, v, stages["v"]


def format_report(rows):
    return "\n".join(f"{n:16s} rel_l2={e:.3e}  ({note})" for n, e, note in rows)
