import json
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["single_even", "single_odd", "padded_mixed", "stock_init_c1lite"]


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    spec = json.loads(bytes(z["spec"]).decode())
    return spec, torch.from_numpy(z["v"]), torch.from_numpy(z["x"])


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())
