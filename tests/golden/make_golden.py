"""Generate the committed golden vectors by running the UNMODIFIED reference modules.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
Weights and inputs come from serenade_b200.synth (numpy PCG64, reproducible anywhere); outputs come
from the reference's CFM.estimator / CFM.solve_euler (flow_matching.py:65-93, decoder.py:384-467)
loaded by oracle/ref_loader.py.  Each .npz stores the outputs, the case spec and a weight checksum.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_loader import build_reference_cfm  # noqa: E402
from serenade_b200 import synth  # noqa: E402

CASES = {
    # name: (lengths, n_steps, jitter, input seed, t for the single estimator call[, in_channels])
    "single_even": ([48], 3, True, 11, 0.3),
    "single_odd": ([37], 3, True, 12, 0.7),
    "padded_mixed": ([64, 41, 50], 4, True, 13, 0.5),
    "stock_init_c1lite": ([200], 10, False, 14, 0.0),
    # SerenadeNew's geometry: conditioning_dim 164 -> CFM(in_channels=244) (reference serenade_new.py:77)
    "new244_mixed": ([57, 40], 3, True, 15, 0.4, 244),
}


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    torch.set_num_threads(os.cpu_count())
    only = sys.argv[1:]
    cfm = None
    loaded = None
    for name, case in CASES.items():
        if only and name not in only:
            continue
        lengths, n_steps, jitter, seed, t = case[:5]
        cin = case[5] if len(case) > 5 else 242
        geom = synth.DecoderGeometry(in_channels=cin)
        if loaded != (jitter, cin):
            cfm = build_reference_cfm(in_channels=cin)
            sd = synth.make_state_dict(seed=0, jitter=jitter, g=geom)
            cfm.load_state_dict(sd, strict=True)
            loaded = (jitter, cin)
        b = synth.make_batch(lengths, n_steps=n_steps, seed=seed, g=geom)
        with torch.inference_mode():
            v = cfm.estimator(b.z, b.mask, b.mu, torch.tensor(t, dtype=torch.float32), b.spks)
            x = cfm.solve_euler(b.z, t_span=b.t_span, mu=b.mu, mask=b.mask, trg_spks=b.spks)
        spec = dict(lengths=lengths, n_steps=n_steps, jitter=jitter, seed=seed, t=t, weight_seed=0, in_channels=cin,
                    weight_checksum=synth.state_dict_checksum(sd), torch=torch.__version__)
        np.savez_compressed(os.path.join(here, name + ".npz"), v=v.numpy(), x=x.numpy(),
                            spec=np.frombuffer(json.dumps(spec).encode(), dtype=np.uint8))
        print(name, spec, float(v.norm()), float(x.norm()))


if __name__ == "__main__":
    main()
