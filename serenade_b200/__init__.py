"""serenade_b200: B200-native (sm_100a) CFM mel-infilling decoder for Serenade.

Public surface (mirrors serenade/models/matcha_components of the reference):
    CFM       -- flow_matching.CFM  (inference / solve_euler / estimator)
    Decoder   -- decoder.Decoder    (the 1D-UNet estimator)
    models.Serenade -- models.serenade.Serenade (content encoder + GST style encoder + CFM decoder, `inference`)
    vocoder.HiFiGANGenerator / vocoder.Vocoder -- serenade.vocoder (mel -> waveform)
    decode_io -- the on-disk half of bin/ssc_decode.py: feature containers, FeatsDataset, scalers, wav / lf0 writers, decode_dump
The compute lives in libserenade_b200.so (C ABI: include/serenade_b200.h); build with
`python -m serenade_b200.build`.
"""
from . import decode_io, models, vocoder  # noqa: F401
from .cfm import CFM, Decoder  # noqa: F401
from .models import Serenade  # noqa: F401
from .synth import (SSC1, DecoderGeometry, make_batch, make_frontend_inputs, make_frontend_state_dict,  # noqa: F401
                    make_state_dict, make_vocoder_input, make_vocoder_state_dict)
from .vocoder import HiFiGANGenerator, Vocoder  # noqa: F401

__all__ = ["CFM", "Decoder", "Serenade", "models", "SSC1", "DecoderGeometry", "make_batch", "make_state_dict",
           "make_frontend_state_dict", "make_frontend_inputs", "vocoder", "HiFiGANGenerator", "Vocoder",
           "make_vocoder_state_dict", "make_vocoder_input", "decode_io"]
