"""CPU: oracle/cfm_oracle.compute_loss against the goldens made by the reference's CFM.compute_loss
(tests/golden/make_golden_loss.py; flow_matching.py:95-133)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import cfm_oracle as O
from serenade_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ["loss_batch3", "loss_infill_mask", "loss_single"]


def load_case(name):
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    spec = json.loads(bytes(g["spec"]).decode())
    b = synth.make_batch(spec["lengths"], n_steps=1, seed=spec["seed"])
    mask_l = torch.from_numpy(g["mask_l"]) if g["mask_l"].size else None
    return g, spec, b, mask_l


@pytest.mark.parametrize("name", CASES)
def test_oracle_loss_matches_reference(name):
    g, spec, b, mask_l = load_case(name)
    sd = synth.make_state_dict(seed=spec["weight_seed"], jitter=True)
    assert abs(synth.state_dict_checksum(sd) - spec["weight_checksum"]) < 1e-6
    with torch.inference_mode():
        loss, y = O.compute_loss(sd, torch.from_numpy(g["x1"]), b.mask, b.mu, b.spks,
                                 torch.from_numpy(g["z"]), torch.from_numpy(g["t"]), mask_l=mask_l)
    assert torch.equal(y, torch.from_numpy(g["y"]))            # same expression, same rounding
    assert abs(float(loss) - float(g["loss"])) <= 2e-5 * abs(float(g["loss"]))
