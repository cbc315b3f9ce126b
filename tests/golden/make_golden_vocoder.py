"""Golden vectors of the HiFi-GAN vocoder from the UNMODIFIED reference module (build container only).

    python tests/golden/make_golden_vocoder.py

Builds the reference's `HiFiGANGenerator` (serenade/vocoder/models/hifigan.py:21-161), removes weight norm as
`Vocoder.__init__` does (vocoder/vocoder.py:38), loads the synthetic state dict of serenade_b200.synth (strict) and
stores `inference_batch` outputs (hifigan.py:268-284) for seeded normalised mels.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference_hifigan  # noqa: E402
from serenade_b200 import synth  # noqa: E402

CASES = {
    # name: (frames, batch, input seed, generator kwargs)
    "vocoder_default_small": (37, 1, 31, {}),                                                   # class defaults: x256
    "vocoder_default_batch2": (64, 2, 32, {}),
    "vocoder_odd_scales": (50, 1, 33, dict(channels=256, upsample_scales=(5, 4, 3, 2), upsample_kernel_sizes=(10, 8, 6, 4))),
}


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    torch.set_num_threads(os.cpu_count())
    mod = load_reference_hifigan()
    for name, (frames, batch, seed, geom) in CASES.items():
        gen = mod.HiFiGANGenerator(**geom)
        gen.remove_weight_norm()
        gen.eval()
        sd = synth.make_vocoder_state_dict(seed=0, **geom)
        gen.load_state_dict(sd, strict=True)
        c = synth.make_vocoder_input(frames, batch=batch, seed=seed)
        with torch.inference_mode():
            wav = gen.inference_batch(c).squeeze(1)
        spec = dict(frames=frames, batch=batch, seed=seed, geometry=geom, torch=torch.__version__)
        np.savez_compressed(os.path.join(here, name + ".npz"), spec=np.frombuffer(json.dumps(spec).encode(), dtype=np.uint8),
                            wav=wav.numpy())
        print(name, tuple(wav.shape), "mean |wav| %.4f" % float(wav.abs().mean()), "max %.4f" % float(wav.abs().max()))


if __name__ == "__main__":
    main()
