"""serenade_b200: B200-native (sm_100a) CFM mel-infilling decoder for Serenade.

Public surface (mirrors serenade/models/matcha_components of the reference):
    CFM       -- flow_matching.CFM  (inference / solve_euler / estimator)
    Decoder   -- decoder.Decoder    (the 1D-UNet estimator)
The compute lives in libserenade_b200.so (C ABI: include/serenade_b200.h); build with
`python -m serenade_b200.build`.
"""
from .cfm import CFM, Decoder  # noqa: F401
from .synth import SSC1, DecoderGeometry, make_batch, make_state_dict  # noqa: F401

__all__ = ["CFM", "Decoder", "SSC1", "DecoderGeometry", "make_batch", "make_state_dict"]
