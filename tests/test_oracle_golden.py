"""Pins oracle/cfm_oracle.py against golden vectors produced by the UNMODIFIED reference modules
(tests/golden/make_golden.py).  CPU only."""
import pytest
import torch

from oracle import cfm_oracle as O
from serenade_b200 import synth
from golden_util import CASES, load_golden, rel_l2


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name, jitter_sd, stock_sd):
    spec, v_ref, x_ref = load_golden(name)
    sd = jitter_sd if spec["jitter"] else stock_sd
    assert synth.state_dict_checksum(sd) == pytest.approx(spec["weight_checksum"], rel=1e-9)
    b = synth.make_batch(spec["lengths"], n_steps=spec["n_steps"], seed=spec["seed"])
    with torch.inference_mode():
        v = O.estimator_forward(sd, b.z, b.mask, b.mu, torch.tensor(spec["t"]), b.spks)
        x = O.solve_euler(sd, b.z, b.t_span, b.mu, b.mask, b.spks)
    assert rel_l2(v, v_ref) < 2e-5      # fp32 re-association only
    assert rel_l2(x, x_ref) < 2e-5
    # padded frames: v is exactly zero, x keeps its noise (reference decoder.py:467)
    pad = ~b.mask.expand_as(v)
    assert float(v[pad].abs().sum()) == 0.0
    assert torch.equal(x[pad], b.z[pad])


def test_oracle_matches_reference_golden_in_channels_244():
    """SerenadeNew's CFM(in_channels=244) (reference serenade_new.py:77): golden from the unmodified reference modules."""
    spec, v_ref, x_ref = load_golden("new244_mixed")
    geom = synth.DecoderGeometry(in_channels=spec["in_channels"])
    sd = synth.make_state_dict(seed=0, jitter=spec["jitter"], g=geom)
    assert synth.state_dict_checksum(sd) == pytest.approx(spec["weight_checksum"], rel=1e-9)
    b = synth.make_batch(spec["lengths"], n_steps=spec["n_steps"], seed=spec["seed"], g=geom)
    assert b.mu.shape[1] == 164
    with torch.inference_mode():
        v = O.estimator_forward(sd, b.z, b.mask, b.mu, torch.tensor(spec["t"]), b.spks)
        x = O.solve_euler(sd, b.z, b.t_span, b.mu, b.mask, b.spks)
    assert rel_l2(v, v_ref) < 2e-5
    assert rel_l2(x, x_ref) < 2e-5


def test_time_grid_matches_reference_wobble():
    # flow_matching.py:79-91: fp32 running t, dt recomputed as t_span[k+1] - t
    ts, dts = O.time_grid(torch.linspace(0, 1, 11))
    assert len(ts) == 10 and ts[0] == 0.0
    assert abs(sum(dts) - 1.0) < 1e-6
    assert len(set(dts)) > 1            # the last-ulp wobble the survey measured


def test_param_shapes_cover_ssc1():
    shapes = synth.estimator_param_shapes()
    assert len(shapes) == 192
    n = 1
    total = 0
    for s in shapes.values():
        n = 1
        for d in s:
            n *= d
        total += n
    assert total == 73_465_424          # SURVEY.md section 0 finding 3
