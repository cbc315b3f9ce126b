"""Golden vectors for CFM.compute_loss (forward value) from the UNMODIFIED reference (flow_matching.py:95-133).

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_loss.py
The reference draws t = torch.rand([B, 1, 1]) and then z = torch.randn_like(x1) from the global CPU generator; the
script seeds it, calls the reference, then re-seeds and repeats the two draws to record exactly the t and z it used.
Weights and inputs come from serenade_b200.synth (numpy PCG64).  Cases: full mask; a loss mask restricted to the
infill segment as Serenade.forward builds it (serenade.py:130-132); ragged lengths.
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
    # name: (lengths, input seed, torch seed, loss mask = infill segment only)
    "loss_batch3": ([64, 41, 50], 21, 5, False),
    "loss_infill_mask": ([48, 48], 22, 6, True),
    "loss_single": ([57], 23, 7, False),
}


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    torch.set_num_threads(os.cpu_count())
    cfm = build_reference_cfm()
    sd = synth.make_state_dict(seed=0, jitter=True)
    cfm.load_state_dict(sd, strict=True)
    for name, (lengths, seed, tseed, seg) in CASES.items():
        b = synth.make_batch(lengths, n_steps=1, seed=seed)
        rng = np.random.Generator(np.random.PCG64(seed + 1000))
        x1 = torch.from_numpy(rng.standard_normal(tuple(b.z.shape), dtype=np.float32)) * b.mask
        mask = b.mask                # bool, as make_non_pad_mask gives it (serenade.py:117)
        mask_l = None
        if seg:
            mask_l = mask.clone()
            for i, L in enumerate(lengths):
                a, e = int(round(0.25 * L)), int(round(0.75 * L))
                mask_l[i, :, :a] = 0
                mask_l[i, :, e:] = 0
        torch.manual_seed(tseed)
        with torch.inference_mode():
            loss, y = cfm.compute_loss(x1, mask, b.mu, b.spks, mask_l=mask_l)
        torch.manual_seed(tseed)
        t = torch.rand([len(lengths), 1, 1], dtype=torch.float32)
        z = torch.randn_like(x1)
        spec = dict(lengths=lengths, seed=seed, torch_seed=tseed, infill_mask=seg, weight_seed=0,
                    weight_checksum=synth.state_dict_checksum(sd), torch=torch.__version__)
        np.savez_compressed(os.path.join(here, name + ".npz"), loss=np.float32(loss.item()), y=y.numpy(), t=t.numpy(),
                            z=z.numpy(), x1=x1.numpy(), mask_l=(mask_l.numpy() if mask_l is not None else np.zeros(0, dtype=bool)),
                            spec=np.frombuffer(json.dumps(spec).encode(), dtype=np.uint8))
        print(name, spec, float(loss), float(y.norm()))


if __name__ == "__main__":
    main()
