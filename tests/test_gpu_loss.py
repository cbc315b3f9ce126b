"""GPU parity of the CFM loss forward value (forward half): `CFM.compute_loss` / `srn_cfm_loss` and
`Decoder.forward` with one time per clip, against the goldens the reference's own `compute_loss` produced
(tests/golden/make_golden_loss.py; flow_matching.py:95-133) and against the oracle.

Tolerances: y is elementwise fp32 arithmetic in torch's rounding order -> bit exact.  The loss is a mean of squared
differences of the estimator output, so errors of the estimator largely average out: the gates are 2e-4 (TF32-class
mode; measured 3e-6 .. 2e-5) and 5e-3 (bf16 mode; measured 4e-5 .. 9e-4), see profiles/r01_loss_parity.txt.
"""
import os

import numpy as np
import pytest
import torch

from golden_util import rel_l2
from test_loss_oracle import CASES, load_case

pytestmark = pytest.mark.gpu

LOSS_TOL = {"tf32": 2e-4, "bf16": 5e-3}
STEP_TOL = {"tf32": 1e-3, "bf16": 2e-2}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda", 0)


@pytest.fixture(scope="module")
def cfms(jitter_sd):
    import serenade_b200 as S
    out = {}
    for prec in ("bf16", "tf32"):
        c = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                  decoder_attention_head_dim=512, precision=prec)
        c.load_state_dict(jitter_sd, strict=True)
        out[prec] = c
    return out


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
@pytest.mark.parametrize("name", CASES)
def test_loss_matches_reference_golden(name, prec, cfms, dev):
    g, spec, b, mask_l = load_case(name)
    cfm = cfms[prec]
    loss, y = cfm.compute_loss(torch.from_numpy(g["x1"]).to(dev), b.mask.to(dev), b.mu.to(dev), b.spks.to(dev),
                               mask_l=mask_l.to(dev) if mask_l is not None else None,
                               z=torch.from_numpy(g["z"]).to(dev), t=torch.from_numpy(g["t"]).to(dev))
    assert torch.equal(y.cpu(), torch.from_numpy(g["y"]))
    err = abs(float(loss) - float(g["loss"])) / abs(float(g["loss"]))
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"loss_{prec}_{name}.txt"), "w") as f:
        f.write(f"{name} ({prec}): loss {float(loss):.6f} reference {float(g['loss']):.6f} rel err {err:.2e}\n")
    assert err < LOSS_TOL[prec]
    # t on the host (srn_cfm_loss) and on the device (srn_cfm_loss_tdev, above) are the same computation
    loss_h, y_h = cfm.compute_loss(torch.from_numpy(g["x1"]).to(dev), b.mask.to(dev), b.mu.to(dev), b.spks.to(dev),
                                   mask_l=mask_l.to(dev) if mask_l is not None else None,
                                   z=torch.from_numpy(g["z"]).to(dev), t=torch.from_numpy(g["t"]))
    assert torch.equal(loss_h, loss) and torch.equal(y_h, y)


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
def test_estimator_with_one_time_per_clip(prec, cfms, jitter_sd, dev):
    """Decoder.forward(t = vector) == the oracle's, and == B single-clip calls each with its own scalar t (bit exact:
    the time path is the only thing that changes)."""
    import serenade_b200 as S
    from oracle import cfm_oracle as O
    cfm = cfms[prec]
    b = S.make_batch([72, 72, 72], n_steps=1, seed=31)
    t = torch.tensor([0.15, 0.5, 0.93])
    with torch.inference_mode():
        ref = O.estimator_forward(jitter_sd, b.z, b.mask, b.mu, t, b.spks)
    v = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), t, b.spks.to(dev))
    assert rel_l2(v.cpu(), ref) < STEP_TOL[prec]
    for i in range(3):
        vi = cfm.estimator(b.z[i:i + 1].to(dev), b.mask[i:i + 1].to(dev), b.mu[i:i + 1].to(dev), t[i], b.spks[i:i + 1].to(dev))
        assert torch.equal(vi[0], v[i])
    # a vector of equal times is the scalar call
    v_same = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), torch.full((3,), 0.4), b.spks.to(dev))
    v_scalar = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), torch.tensor(0.4), b.spks.to(dev))
    assert torch.equal(v_same, v_scalar)
    # times that already live on the GPU take srn_estimator_forward_tdev (no device-to-host read): same bits, and a
    # reference-style Euler loop over estimator(...) with a device-resident t reproduces the library's own solve loop
    v_dev = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), t.to(dev), b.spks.to(dev))
    assert torch.equal(v_dev, v)
    v_dev0 = cfm.estimator(b.z.to(dev), b.mask.to(dev), b.mu.to(dev), torch.tensor(0.4, device=dev), b.spks.to(dev))
    assert torch.equal(v_dev0, v_scalar)
    x = b.z.to(dev)
    mask, mu, spk = b.mask.to(dev), b.mu.to(dev), b.spks.to(dev)
    t_span = torch.linspace(0, 1, 4, device=dev)
    tt, dt = t_span[0], t_span[1] - t_span[0]
    for step in range(1, 4):                                  # flow_matching.py:79-91, all on the device
        x = x + dt * cfm.estimator(x, mask, mu, tt, spk)
        tt = tt + dt
        if step < 3:
            dt = t_span[step + 1] - tt
    want = cfm.solve_euler(b.z.to(dev), t_span=t_span, mu=mu, mask=mask, trg_spks=spk)
    # the fused Euler update rounds x + dt v once, torch twice: last-bit differences in x, which a bf16 operand
    # rounding of the next step's input can amplify - far below the per-step tolerance either way
    assert rel_l2(x.cpu(), want.cpu()) < 1e-3


def test_estimator_call_can_be_captured_in_the_callers_cuda_graph(cfms, dev):
    """With `t` on the device the single-step entry point enqueues copies and kernels only (no allocation, no host
    synchronisation, no host-side value baked into a launch), so a caller can capture it in their own CUDA graph - e.g. a
    custom solver loop - and replay it with new inputs."""
    import serenade_b200 as S
    cfm = cfms["bf16"]
    b1 = S.make_batch([72, 64], n_steps=1, seed=51)
    b2 = S.make_batch([72, 64], n_steps=1, seed=52)
    mask, spk = b1.mask.to(dev), b1.spks.to(dev)
    x, mu, t = b1.z.to(dev).clone(), b1.mu.to(dev).clone(), torch.tensor(0.3, device=dev)
    eager1 = cfm.estimator(x, mask, mu, t, spk)                     # (also builds the plan: allocation happens here)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            out = cfm.estimator(x, mask, mu, t, spk)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager1)
    x.copy_(b2.z.to(dev))
    mu.copy_(b2.mu.to(dev))
    t.fill_(0.8)
    g.replay()
    torch.cuda.synchronize()
    got = out.clone()
    eager2 = cfm.estimator(b2.z.to(dev), mask, b2.mu.to(dev), torch.tensor(0.8), spk)
    assert torch.equal(got, eager2) and not torch.equal(got, eager1)


def test_loss_draws_like_the_reference_and_rejects_bad_input(cfms, dev):
    import serenade_b200 as S
    cfm = cfms["bf16"]
    b = S.make_batch([40, 33], n_steps=1, seed=32)
    x1 = torch.randn(2, 80, 40) * b.mask
    torch.manual_seed(3)
    l1, y1 = cfm.compute_loss(x1.to(dev), b.mask.to(dev), b.mu.to(dev), b.spks.to(dev))
    torch.manual_seed(3)
    l2, y2 = cfm.forward(x1.to(dev), b.mask.to(dev), b.mu.to(dev), b.spks.to(dev))
    assert torch.equal(y1, y2) and float(l1) == float(l2)         # deterministic reductions
    assert np.isfinite(float(l1)) and float(l1) > 0
    with pytest.raises(ValueError):
        cfm.compute_loss(x1.to(dev), b.mask.to(dev), b.mu.to(dev), b.spks.to(dev), t=torch.rand(3))
    with pytest.raises(RuntimeError):
        cfm.compute_loss(x1, b.mask, b.mu, b.spks)                # CPU tensors: no CPU path
