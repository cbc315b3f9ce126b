"""Golden vectors of the conditioning front half, from the UNMODIFIED reference modules (build container only).

    python tests/golden/make_golden_frontend.py

Runs the reference's `Serenade.encoder` (Conv1dResnet, serenade.py:224-376), `Serenade.gst` (StyleEncoder,
modules/gst/style_encoder.py) and the whole `Serenade.inference` (serenade.py:168-221, noise from the CPU generator
seeded here) on synthetic weights / inputs from serenade_b200.synth, and stores the outputs.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_loader import build_reference_serenade  # noqa: E402
from serenade_b200 import synth  # noqa: E402

CASES = {
    # name: (src_len, ref_len, batch, input seed)
    "frontend_small": (57, 40, 1, 21),
    "frontend_odd_batch2": (131, 77, 2, 22),
    "frontend_c1": (300, 200, 1, 23),
}


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    torch.set_num_threads(os.cpu_count())
    model = build_reference_serenade()
    sd = {("cfm_decoder." + k): v for k, v in synth.make_state_dict(seed=0, jitter=True).items()}
    sd.update(synth.make_frontend_state_dict(seed=0))
    model.load_state_dict(sd, strict=True)
    for name, (ls, lr, batch, seed) in CASES.items():
        inp = synth.make_frontend_inputs(ls, lr, seed=seed, batch=batch)
        with torch.inference_mode():
            enc = model.encoder(inp["x"], inp["lengths"])
            ref_enc = model.encoder(inp["ref_x"], inp["ref_lengths"])
            style = model.gst(inp["ref_logmel"])
            out = {"enc": enc.numpy(), "ref_enc": ref_enc.numpy(), "style": style.numpy()}
            if batch == 1:      # Serenade.inference is batch-1 in the reference (serenade.py:209-219)
                torch.manual_seed(1000 + seed)      # CFM.inference draws z from the global CPU generator
                out["mel"] = model.inference(**inp).numpy()
        spec = dict(src_len=ls, ref_len=lr, batch=batch, seed=seed, noise_seed=1000 + seed, torch=torch.__version__)
        np.savez_compressed(os.path.join(here, name + ".npz"), spec=np.frombuffer(json.dumps(spec).encode(), dtype=np.uint8),
                            **out)
        print(name, {k: (v.shape, float(np.abs(v).mean())) for k, v in out.items()})


if __name__ == "__main__":
    main()
