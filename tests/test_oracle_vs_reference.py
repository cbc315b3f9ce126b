"""When the reference tree is present (build container), run its modules directly against the oracle,
including the quirks SURVEY.md section 0 lists.  Skipped on the GPU box (no /root/reference)."""
import pytest
import torch

from oracle import cfm_oracle as O
from oracle.ref_loader import build_reference_cfm, reference_available
from serenade_b200 import synth
from golden_util import rel_l2

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref_cfm(jitter_sd):
    cfm = build_reference_cfm()
    cfm.load_state_dict(jitter_sd, strict=True)     # key-for-key, strict
    return cfm


@pytest.mark.parametrize("lengths", [[31], [40, 23], [57, 57, 12]])
def test_estimator_and_solver(ref_cfm, jitter_sd, lengths):
    b = synth.make_batch(lengths, n_steps=2, seed=3)
    with torch.inference_mode():
        t = torch.tensor(0.41)
        assert rel_l2(O.estimator_forward(jitter_sd, b.z, b.mask, b.mu, t, b.spks),
                      ref_cfm.estimator(b.z, b.mask, b.mu, t, b.spks)) < 2e-5
        assert rel_l2(O.solve_euler(jitter_sd, b.z, b.t_span, b.mu, b.mask, b.spks),
                      ref_cfm.solve_euler(b.z, b.t_span, b.mu, b.mask, b.spks)) < 2e-5


def test_padded_region_content_is_irrelevant_but_padding_amount_is_not(ref_cfm, jitter_sd):
    # SURVEY.md finding 6: GroupNorm spans T_max, so valid frames depend on how much padding there is
    b = synth.make_batch([30], n_steps=1, seed=5)
    bp = synth.make_batch([30], n_steps=1, seed=5, t_max=46)
    t = torch.tensor(0.2)
    with torch.inference_mode():
        v = O.estimator_forward(jitter_sd, b.z, b.mask, b.mu, t, b.spks)
        z2 = bp.z.clone(); z2[:, :, 30:] = 123.0           # garbage in the padded region of x
        v_pad = O.estimator_forward(jitter_sd, bp.z, bp.mask, bp.mu, t, bp.spks)
        v_pad2 = O.estimator_forward(jitter_sd, z2, bp.mask, bp.mu, t, bp.spks)
    assert rel_l2(v_pad2, v_pad) < 1e-6
    assert v_pad.shape[-1] == 46 and float(v_pad[:, :, 30:].abs().sum()) == 0.0
    assert v.shape[-1] == 30
