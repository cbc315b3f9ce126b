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


def fine_reference(sd, batch, t):
    """Inner tensors of down_blocks.0 (channels-last), computed with torch fp32 from the state dict."""
    P = "estimator.down_blocks.0"
    out = {}
    with torch.inference_mode():
        mask = batch.mask
        x = torch.cat([batch.z, batch.mu], dim=1)
        xm = x * mask
        out["xin"] = F.pad(xm.transpose(1, 2), (0, 256 - xm.shape[1]))
        temb = O._time_embedding(sd, torch.tensor(t), x.shape[1])
        y1 = F.conv1d(xm, sd[P + ".0.block1.block.0.weight"], sd[P + ".0.block1.block.0.bias"], padding=1)
        out["down0.y1"] = y1.transpose(1, 2)
        out["down0.r"] = F.conv1d(xm, sd[P + ".0.res_conv.weight"], sd[P + ".0.res_conv.bias"]).transpose(1, 2)
        h = F.mish(F.group_norm(y1, 8, sd[P + ".0.block1.block.1.weight"], sd[P + ".0.block1.block.1.bias"])) * mask
        h = h + F.linear(F.mish(temb), sd[P + ".0.mlp.1.weight"], sd[P + ".0.mlp.1.bias"]).unsqueeze(-1)
        out["down0.a1"] = (h * mask).transpose(1, 2)
        y2 = F.conv1d(h * mask, sd[P + ".0.block2.block.0.weight"], sd[P + ".0.block2.block.0.bias"], padding=1)
        out["down0.y2"] = y2.transpose(1, 2)
        res = O._resnet(sd, P + ".0", x, mask, temb, batch.spks)
        out["down0.resnet"] = res.transpose(1, 2)
        hs = res.transpose(1, 2)
        tp = P + ".1.0"
        n1 = F.layer_norm(hs, (512,), sd[tp + ".norm1.weight"], sd[tp + ".norm1.bias"])
        q = F.linear(n1, sd[tp + ".attn1.to_q.weight"])
        k = F.linear(n1, sd[tp + ".attn1.to_k.weight"])
        v = F.linear(n1, sd[tp + ".attn1.to_v.weight"])
        out["down0.qk"] = torch.cat([q, k], dim=-1)
        B, S, _ = q.shape
        qh, kh, vh = (z.view(B, S, 4, 512).transpose(1, 2) for z in (q, k, v))
        sc = torch.matmul(qh, kh.transpose(-1, -2)) / math.sqrt(512)
        out["down0.sc"] = sc.reshape(B * 4, S, S)
        scm = sc.masked_fill(~mask[:, 0].view(B, 1, 1, S), float("-inf"))
        pr = torch.softmax(scm, dim=-1)
        out["down0.P"] = pr.reshape(B * 4, S, S)
        ao = torch.matmul(pr, vh).transpose(1, 2).reshape(B, S, 2048)
        out["down0.ao"] = ao
        h_attn = F.linear(ao, sd[tp + ".attn1.to_out.0.weight"], sd[tp + ".attn1.to_out.0.bias"]) + hs
        out["down0.h_attn"] = h_attn
        n3 = F.layer_norm(h_attn, (512,), sd[tp + ".norm3.weight"], sd[tp + ".norm3.bias"])
        out["down0.ln3"] = n3
        pj = F.linear(n3, sd[tp + ".ff.net.0.proj.weight"], sd[tp + ".ff.net.0.proj.bias"])
        hh, gg = pj.chunk(2, dim=-1)
        out["down0.gg"] = hh * F.gelu(gg)
    return out


BLOCK_ORDER = ["down0.resnet", "down0.tfm", "down0.out", "down1.resnet", "down1.tfm", "down1.out",
               "mid0.resnet", "mid0.tfm", "mid1.resnet", "mid1.tfm", "up0.resnet", "up0.tfm", "up0.out",
               "up1.resnet", "up1.tfm", "up1.out", "final_block"]
FINE_ORDER = ["xin", "down0.y1", "down0.r", "down0.a1", "down0.y2", "down0.resnet", "down0.qk",
              "down0.P", "down0.ao", "down0.h_attn", "down0.ln3", "down0.gg"]
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
        if fine:
            ref = fine_reference(sd, batch, t)
            for name in FINE_ORDER:
                got = eng.debug_fetch(name)
                want = ref[name]
                valid = level_mask(batch, want.shape[1]) if name != "down0.P" else None
                if name == "down0.P":
                    # the CUDA path keeps P~ = exp(s - rowmax) unnormalised (the row sum is divided out in the
                    # P~ V epilogue): normalise here; compare valid query rows and valid keys only
                    m = batch.mask[:, 0].float()
                    mm = (m.unsqueeze(2) * m.unsqueeze(1)).repeat_interleave(4, dim=0)
                    got = got * mm
                    got = got / got.sum(-1, keepdim=True).clamp_min(1e-30)
                    rows.append((name, rel(got * mm, want * mm), "valid q,k, rows normalised on the host"))
                else:
                    rows.append((name, rel(got * valid, want * valid), "valid rows"))
        stages = oracle_stages(sd, batch, t)
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
