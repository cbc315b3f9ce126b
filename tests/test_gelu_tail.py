"""The GEGLU epilogue's fast GELU (serenade_b200/csrc/gemm.cuh: gelu_tail, bf16 outputs only) against the exact erf
form the reference uses (diffusers GEGLU -> F.gelu, transformer.py:120): the constants are read out of the shipped
source, evaluated in float32 exactly as the device code does (Horner, clamp at 5.5, 2^e), and the error bound stated
in the source comment and in DESIGN.md is checked.  CPU only."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _coefficients():
    src = open(os.path.join(ROOT, "serenade_b200", "csrc", "gemm.cuh")).read()
    body = src[src.index("float gelu_tail(float g)"):]
    body = body[:body.index("return")]
    lead = re.search(r"float e = fmaf\(([-0-9.e+]+)f, zc, ([-0-9.e+]+)f\)", body)
    rest = re.findall(r"e = fmaf\(e, zc, ([-0-9.e+]+)f\)", body)
    clamp = float(re.search(r"fminf\(z, ([0-9.]+)f\)", body).group(1))
    coef = [float(lead.group(1)), float(lead.group(2))] + [float(c) for c in rest]
    return np.array(coef, dtype=np.float32), np.float32(clamp)


def _gelu_tail(g, coef, clamp):
    g = g.astype(np.float32)
    z = np.abs(g)
    zc = np.minimum(z, clamp)
    e = np.full_like(zc, coef[0])
    for c in coef[1:]:
        e = (e * zc + c).astype(np.float32)
    q = np.exp2(e.astype(np.float64))
    return np.maximum(g, 0).astype(np.float64) - z.astype(np.float64) * q


def _gelu_erf(g):
    from math import erf, sqrt
    return np.array([0.5 * x * (1.0 + erf(x / sqrt(2.0))) for x in g.astype(np.float64)])


def test_gelu_tail_matches_erf_gelu():
    coef, clamp = _coefficients()
    assert len(coef) == 7 and clamp == np.float32(5.5)
    g = np.concatenate([np.linspace(-12, 12, 48001), np.array([0.0, -0.0, 1e-6, -1e-6, 30.0, -30.0, 100.0, -100.0])])
    ours, ref = _gelu_tail(g, coef, clamp), _gelu_erf(g)
    err = np.abs(ours - ref)
    assert err.max() <= 5e-6, err.max()                       # absolute, anywhere (incl. the clamped tail)
    mid = (np.abs(g) < 5.0) & (np.abs(g) > 1e-3)
    assert (err[mid] / np.abs(ref[mid])).max() <= 4e-5        # relative on either side, two orders below bf16 rounding
    assert ours[g == 0.0].tolist() == [0.0] * int((g == 0.0).sum())
