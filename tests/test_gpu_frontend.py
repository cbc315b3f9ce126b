"""GPU parity of the conditioning front half through the C ABI: content encoder, GST style
encoder and the whole `Serenade.inference` against the golden vectors of the reference's own modules.

Tolerances: operands are bf16 (or TF32-class) on the tensor cores with fp32 accumulation, like the decoder:
bf16 mode <= 1e-2 rel-L2 per module, TF32-class mode <= 1e-3.  The end-to-end mel is compared with the reference's
mel for the same CPU-generator noise: the decoder's own 1e-2 gate plus the bf16 error of its conditioning input.
"""
import os

import pytest
import torch

from golden_util import rel_l2
from test_frontend_oracle import FRONT_CASES, load_front

pytestmark = pytest.mark.gpu


def _model(precision):
    import serenade_b200 as S
    m = S.models.Serenade(precision=precision)
    sd = {("cfm_decoder." + k): v for k, v in S.make_state_dict(0, True).items()}
    sd.update(S.make_frontend_state_dict(0))
    m.load_state_dict(sd, strict=True)
    return m


@pytest.fixture(scope="module")
def models():
    assert torch.cuda.is_available()
    return {"bf16": _model("bf16"), "tf32": _model("tf32")}


def _report(name, text):
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", name), "a") as f:
        f.write(text + "\n")


@pytest.mark.parametrize("precision,tol", [("bf16", 1e-2), ("tf32", 1e-3)])
@pytest.mark.parametrize("name", FRONT_CASES)
def test_content_and_style_encoders_match_reference(name, precision, tol, models):
    import serenade_b200 as S
    dev = torch.device("cuda", 0)
    spec, ref = load_front(name)
    inp = S.make_frontend_inputs(spec["src_len"], spec["ref_len"], seed=spec["seed"], batch=spec["batch"])
    m = models[precision]
    enc = m.encode_content(inp["x"].to(dev)).cpu()
    ref_enc = m.encode_content(inp["ref_x"].to(dev)).cpu()
    style = m.encode_style(inp["ref_logmel"].to(dev)).cpu()
    e1, e2, e3 = rel_l2(enc, ref["enc"]), rel_l2(ref_enc, ref["ref_enc"]), rel_l2(style, ref["style"])
    _report("frontend_parity.txt", f"{name} {precision}: enc {e1:.3e} ref_enc {e2:.3e} style {e3:.3e}")
    assert e1 < tol and e2 < tol and e3 < tol


@pytest.mark.parametrize("name", [c for c in FRONT_CASES if "batch2" not in c])
def test_serenade_inference_matches_reference(name, models):
    import serenade_b200 as S
    dev = torch.device("cuda", 0)
    spec, ref = load_front(name)
    inp = {k: v.to(dev) for k, v in S.make_frontend_inputs(spec["src_len"], spec["ref_len"], seed=spec["seed"]).items()}
    for precision, tol in (("bf16", 1e-2), ("tf32", 1e-3)):
        torch.manual_seed(spec["noise_seed"])          # CFM.inference draws z on the CPU generator, like the reference
        mel = models[precision].inference(**inp).cpu()
        assert mel.shape == ref["mel"].shape
        err = rel_l2(mel, ref["mel"])
        _report("frontend_parity.txt", f"{name} {precision}: Serenade.inference mel {err:.3e}")
        assert err < tol


def test_batch_rows_are_independent(models):
    """The encoders have no cross-utterance coupling: a row of a batch equals the same row alone (bit exact)."""
    import serenade_b200 as S
    dev = torch.device("cuda", 0)
    inp = S.make_frontend_inputs(96, 64, seed=3, batch=3)
    m = models["bf16"]
    enc = m.encode_content(inp["x"].to(dev))
    sty = m.encode_style(inp["ref_logmel"].to(dev))
    for i in range(3):
        assert torch.equal(m.encode_content(inp["x"][i:i + 1].to(dev))[0], enc[i])
        assert torch.equal(m.encode_style(inp["ref_logmel"][i:i + 1].to(dev))[0], sty[i])


def test_inference_batch_equals_sequential_single_utterance_inference(models):
    """the length-bucketed batched decode returns, for every utterance, the mel that `inference` (the
    reference's batch-1 call, bin/ssc_decode.py:428) returns for it alone - same noise draws, GroupNorm statistics
    over the utterance's own frames."""
    import serenade_b200 as S
    dev = torch.device("cuda", 0)
    m = models["bf16"]
    order = ["x", "lengths", "midi", "lft", "ref_x", "ref_lengths", "ref_logmel", "ref_midi", "ref_lft"]
    shapes = [(96, 64), (131, 70), (60, 97), (200, 64), (75, 75)]
    items = []
    for k, (src, ref) in enumerate(shapes):
        inp = S.make_frontend_inputs(src, ref, seed=20 + k)
        items.append([inp[name].to(dev) for name in order])
    torch.manual_seed(7)
    solo = [m.inference(*it) for it in items]
    torch.manual_seed(7)
    batched = m.inference_batch(items, max_batch=3)
    assert m.cfm_decoder.estimator.gn_stats == "padded"      # restored
    worst = 0.0
    for (src, ref), a, b in zip(shapes, solo, batched):
        assert a.shape == b.shape == (src, 80)
        worst = max(worst, rel_l2(b.cpu(), a.cpu()))
    _report("frontend_parity.txt", f"inference_batch vs sequential inference: worst rel_l2 {worst:.3e} "
            f"(bit exact: {all(torch.equal(a, b) for a, b in zip(solo, batched))})")
    assert worst < 1e-5

