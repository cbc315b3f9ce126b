"""CPU: the vocoder oracle against the golden vectors of the reference's own HiFiGANGenerator."""
import json
import os

import numpy as np
import pytest
import torch

from golden_util import rel_l2
from oracle import vocoder_oracle as V
from oracle.ref_loader import reference_available
from serenade_b200 import synth

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VOC_CASES = ["vocoder_default_small", "vocoder_default_batch2", "vocoder_odd_scales"]


def load_voc(name):
    z = np.load(os.path.join(HERE, name + ".npz"))
    spec = json.loads(bytes(z["spec"]).decode())
    return spec, torch.from_numpy(z["wav"])


@pytest.mark.parametrize("name", VOC_CASES)
def test_oracle_matches_reference_golden(name):
    spec, wav_ref = load_voc(name)
    sd = synth.make_vocoder_state_dict(seed=0, **spec["geometry"])
    c = synth.make_vocoder_input(spec["frames"], batch=spec["batch"], seed=spec["seed"])
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    with torch.inference_mode():
        wav = V.hifigan_forward(sd, c.transpose(1, 2)).squeeze(1)
    assert wav.shape == wav_ref.shape
    assert rel_l2(wav, wav_ref) < 2e-5          # fp32 re-association only


def test_weight_norm_folding_and_decode_normalisation():
    sd = synth.make_vocoder_state_dict(seed=1, channels=64, upsample_scales=(2, 2), resblock_kernel_sizes=(3,))
    wn = {}
    g = torch.Generator().manual_seed(0)
    for k, w in sd.items():
        if k.endswith(".weight"):
            scale = torch.rand(w.shape[0], generator=g) + 0.5
            v = w * scale.view(-1, *([1] * (w.dim() - 1)))             # any v with the same direction per dim-0 slice
            wn[k + "_v"] = v
            wn[k + "_g"] = w.flatten(1).norm(dim=1).view(-1, *([1] * (w.dim() - 1)))
        else:
            wn[k] = w
    folded = V.fold_weight_norm(wn)
    assert set(folded) == set(sd)
    for k in sd:
        assert torch.allclose(folded[k], sd[k], rtol=1e-5, atol=1e-7), k
    mel = synth.make_vocoder_input(12, batch=2, seed=5)
    mean, scale = torch.linspace(-1, 1, 80), torch.linspace(0.5, 2, 80)
    tm, ts = torch.linspace(0.2, -0.2, 80), torch.linspace(1.5, 0.7, 80)
    wav = V.vocoder_decode(sd, mel, mean, scale, tm, ts)
    with torch.inference_mode():
        direct = V.hifigan_forward(sd, (((mel * ts + tm) - mean) / scale).transpose(1, 2)).squeeze(1)
    assert wav.shape == (2, 12 * 4) and torch.equal(wav, direct)


@pytest.mark.skipif(not reference_available(), reason="/root/reference is only present in the build container")
def test_oracle_matches_reference_module_directly():
    from oracle.ref_loader import load_reference_hifigan
    geom = dict(channels=128, upsample_scales=(4, 3, 2), upsample_kernel_sizes=(8, 6, 4), resblock_kernel_sizes=(3, 5))
    gen = load_reference_hifigan().HiFiGANGenerator(resblock_dilations=[(1, 3, 5), (1, 3, 5)], **geom)
    gen.remove_weight_norm()
    gen.eval()
    sd = synth.make_vocoder_state_dict(seed=3, **geom)
    gen.load_state_dict(sd, strict=True)
    c = synth.make_vocoder_input(21, batch=2, seed=9)
    with torch.inference_mode():
        want = gen.inference_batch(c).squeeze(1)
        got = V.hifigan_forward(sd, c.transpose(1, 2)).squeeze(1)
    assert rel_l2(got, want) < 2e-5
