"""CPU: the front-half oracle (oracle/frontend_oracle.py) against the golden vectors made by the reference's own
modules (tests/golden/make_golden_frontend.py), and the drop-in state-dict contract of `serenade_b200.models.Serenade`."""
import json
import os

import numpy as np
import pytest
import torch

from golden_util import GOLDEN_DIR, rel_l2
from oracle import cfm_oracle as O
from oracle import frontend_oracle as FO
from oracle import ref_loader

FRONT_CASES = ["frontend_small", "frontend_odd_batch2", "frontend_c1"]


def load_front(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    spec = json.loads(bytes(z["spec"]).decode())
    return spec, {k: torch.from_numpy(z[k]) for k in z.files if k != "spec"}


@pytest.fixture(scope="module")
def front_sd():
    from serenade_b200 import synth
    return synth.make_frontend_state_dict(seed=0)


@pytest.mark.parametrize("name", FRONT_CASES)
def test_oracle_matches_reference_golden(name, front_sd, jitter_sd):
    from serenade_b200 import synth
    spec, ref = load_front(name)
    inp = synth.make_frontend_inputs(spec["src_len"], spec["ref_len"], seed=spec["seed"], batch=spec["batch"])
    with torch.inference_mode():
        enc = FO.content_encoder(front_sd, inp["x"])
        ref_enc = FO.content_encoder(front_sd, inp["ref_x"])
        style = FO.style_encoder(front_sd, inp["ref_logmel"])
        assert rel_l2(enc, ref["enc"]) < 2e-5
        assert rel_l2(ref_enc, ref["ref_enc"]) < 2e-5
        assert rel_l2(style, ref["style"]) < 2e-5
        if "mel" in ref:        # whole Serenade.inference: same CPU-generator noise as the reference drew
            mu = FO.assemble_mu(enc, inp["midi"], inp["lft"], ref_enc, inp["ref_midi"], inp["ref_lft"], inp["ref_logmel"])
            T = mu.shape[2]
            mask = torch.ones(1, 1, T, dtype=torch.bool)
            torch.manual_seed(spec["noise_seed"])
            z = torch.randn(1, 80, T) * 0.667
            x = O.solve_euler(jitter_sd, z, torch.linspace(0, 1, 11), mu, mask, style)
            mel = x.permute(0, 2, 1)[0, spec["ref_len"]:]
            assert rel_l2(mel, ref["mel"]) < 1e-4


def test_model_state_dict_contract():
    import serenade_b200 as S
    m = S.models.Serenade()
    sd = {("cfm_decoder." + k): v for k, v in S.make_state_dict(0, True).items()}
    sd.update(S.make_frontend_state_dict(0))
    m.load_state_dict(sd, strict=True)
    with pytest.raises(RuntimeError):
        inp = S.make_frontend_inputs(16, 12)
        m.inference(**inp)          # CPU tensors: there is no fallback path
    with pytest.raises(NotImplementedError):
        m(None)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree not present")
def test_state_dict_keys_equal_the_reference_model():
    import serenade_b200 as S
    ref = ref_loader.build_reference_serenade()
    ours = S.models.Serenade()
    rsd, osd = ref.state_dict(), ours.state_dict()
    assert list(rsd.keys()) == list(osd.keys()) or set(rsd.keys()) == set(osd.keys())
    for k in rsd:
        assert tuple(rsd[k].shape) == tuple(osd[k].shape), k
    ours.load_state_dict(rsd, strict=True)      # a reference checkpoint loads unchanged
