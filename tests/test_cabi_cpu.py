"""CPU-side checks of the boundary: the shared library builds/loads and exports every symbol that
include/serenade_b200.h declares; the Python front end mirrors the reference's state-dict layout and
fails loudly without a GPU (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from serenade_b200 import _lib, build
    build.build_library()
    return _lib.load_library()


def test_header_symbols_are_exported(lib):
    from serenade_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "serenade_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(srn_[a-z_]+)\s*\(", hdr)))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared
    assert lib.srn_version() >= 100


def test_create_fails_loudly_without_gpu(lib):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from serenade_b200 import _lib
    cfg = _lib.SrnConfig(242, 80, 256, 512, 4, 512, 2, 0)
    h = ctypes.c_void_p()
    rc = lib.srn_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc != 0 and lib.srn_last_error()


def test_front_end_state_dict_is_the_reference_layout(jitter_sd):
    import serenade_b200 as S
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    own = cfm.state_dict()
    assert list(own.keys()) == list(jitter_sd.keys())
    assert all(own[k].shape == jitter_sd[k].shape for k in own)
    # stock init statistics follow decoder.py:366-382
    w = own["estimator.down_blocks.0.0.block2.block.0.weight"]
    assert abs(float(w.std()) - (2.0 / (512 * 3)) ** 0.5) < 2e-3
    assert float(own["estimator.final_proj.bias"].abs().sum()) == 0.0
    missing, unexpected = cfm.load_state_dict(jitter_sd, strict=True)
    assert not missing and not unexpected


def test_reference_state_dict_keys_match_when_reference_present(jitter_sd):
    from oracle.ref_loader import build_reference_cfm, reference_available
    if not reference_available():
        pytest.skip("reference tree not present")
    ref = build_reference_cfm().state_dict()
    assert list(ref.keys()) == list(jitter_sd.keys())
    assert all(tuple(ref[k].shape) == tuple(jitter_sd[k].shape) for k in ref)


def test_no_cpu_fallback():
    import serenade_b200 as S
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    b = S.make_batch([16], n_steps=1)
    with pytest.raises(RuntimeError, match="no CPU path"):
        cfm.solve_euler(b.z, b.t_span, b.mu, b.mask, b.spks)
    with pytest.raises(RuntimeError, match="no CPU path"):      # the loss value runs on the B200 too (no backward)
        cfm.compute_loss(b.z, b.mask, b.mu, b.spks)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "serenade_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the oracle", ""), f"{f} references oracle/"


def test_header_is_plain_c_and_links_from_a_c_program(tmp_path):
    """The boundary is a C ABI: include/serenade_b200.h compiles as C (gcc -std=c99 -Wall -Werror), and a C program that takes
    the address of every declared entry point links against the in-tree library and runs its two GPU-free calls."""
    import re
    import shutil
    import subprocess
    from serenade_b200 import _lib
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "serenade_b200.h")).read()
    names = sorted(set(re.findall(r"\b(srn_[a-z0-9_]+)\s*\(", hdr)))
    assert set(names) == set(_lib.EXPORTED_SYMBOLS)
    src = tmp_path / "link_check.c"
    src.write_text('#include <stdio.h>\n#include "serenade_b200.h"\nint main(void) {\n  void* fns[] = {'
                   + ", ".join(f"(void*){n}" for n in names)
                   + '};\n  unsigned i, n = 0;\n  for (i = 0; i < sizeof fns / sizeof fns[0]; ++i) n += fns[i] != 0;\n'
                   '  printf("%u %d [%s]\\n", n, srn_version(), srn_last_error());\n  return 0;\n}\n')
    exe = tmp_path / "link_check"
    libdir = os.path.dirname(_lib.LIB_PATH)
    cmd = ["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(root, "include"), str(src),
           "-o", str(exe), "-L", libdir, "-l:" + os.path.basename(_lib.LIB_PATH), "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    count, version = out.stdout.split()[:2]
    assert int(count) == len(names) and int(version) >= 100
