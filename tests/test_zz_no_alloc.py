"""srn_cfm_loss_from_velocity_noalloc (include/serenade_b200.h): the allocation-free entry point added
- no counterpart in the reference.  Its scratch memory is a file-scope `static __device__` array
(csrc/noalloc.cu) instead of a cudaMalloc'ed buffer.

Checked against the existing entry points through the C ABI: `srn_cfm_loss` returns (loss, y) for given x1 / z / t;
`srn_estimator_forward_t` gives the velocity v for that y; the new call must turn (v, x1, z, masks) into the same loss.
The existing path is deterministic (fixed-order double reductions), so the two agree far inside the suite's bf16 loss
tolerance (tests/test_gpu_loss.py: 5e-3).  Two calls in a row with different inputs share the static scratch array, which
keeps the first call's contents: the second must not be affected by them.

This file is named so that pytest collects it LAST; it passes alone in a fresh
process and after the other tests.
"""
import ctypes as C

import pytest
import torch

LOSS_TOL = 5e-3      # the bf16 loss gate of tests/test_gpu_loss.py


def _noalloc_loss(lib, v, x1, z, mask_u8, mask_l_u8, sigma_min, out):
    """The C-ABI call itself: caller-owned inputs and output, raw pointers, the current stream."""
    from serenade_b200 import _lib
    B, Cc, T = v.shape
    st = torch.cuda.current_stream(v.device).cuda_stream
    _lib.check(lib.srn_cfm_loss_from_velocity_noalloc(v.data_ptr(), x1.data_ptr(), z.data_ptr(), mask_u8.data_ptr(),
                                                      mask_l_u8.data_ptr() if mask_l_u8 is not None else None,
                                                      B, Cc, T, C.c_float(sigma_min), out.data_ptr(), C.c_void_p(st)),
               "srn_cfm_loss_from_velocity_noalloc")
    return out


def test_noalloc_source_calls_no_allocating_routine():
    """CPU check: csrc/noalloc.cu calls no cudaMalloc* / cudaFree* / cudaHostAlloc / new / malloc, takes the scratch address
    with cudaGetSymbolAddress, and keeps the scratch in a static __device__ __align__(256) array."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "serenade_b200", "csrc",
                            "noalloc.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    assert not re.search(r"\b(cudaMalloc\w*|cudaFree\w*|cudaHostAlloc|cudaMallocAsync|malloc|calloc)\s*\(", code)
    assert not re.search(r"\bnew\b", code)
    assert re.search(r"static\s+__device__\s+__align__\(256\)", code)
    assert "cudaGetSymbolAddress" in code


@pytest.mark.gpu
def test_noalloc_loss_matches_the_existing_entry_points_twice_in_a_row(jitter_sd):
    import serenade_b200 as S
    from serenade_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a B200"
    dev = torch.device("cuda", 0)
    lib = _lib.load_library()
    cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
                decoder_attention_head_dim=512)
    cfm.load_state_dict(jitter_sd, strict=True)
    # two different cases, back to back: ragged batch with an infill mask_l, then a single clip without mask_l (the
    # reference's quirk: the sum then runs over padded frames too)
    cases = [([96, 71, 50], 3, True), ([64], 5, False)]
    outs = torch.zeros(len(cases), dtype=torch.float32, device=dev)      # caller-owned outputs, allocated up front
    want, prepared = [], []
    for lengths, seed, with_l in cases:
        b = S.make_batch(lengths, n_steps=1, seed=40 + seed)
        g = torch.Generator().manual_seed(seed)
        x1 = (torch.randn(b.z.shape, generator=g) * b.mask).to(dev)
        z = torch.randn(b.z.shape, generator=g).to(dev)
        t = torch.rand(len(lengths), generator=g)
        mask, mu, spk = b.mask.to(dev), b.mu.to(dev), b.spks.to(dev)
        mask_l = None
        if with_l:
            ml = b.mask.clone()
            for i, L in enumerate(lengths):
                ml[i, :, : L // 4] = False           # loss on the infill region only (serenade.py:134-156)
            mask_l = ml.to(dev)
        loss_ref, y = cfm.compute_loss(x1, mask, mu, spk, mask_l=mask_l, z=z, t=t)          # srn_cfm_loss
        v = cfm.estimator(y, mask, mu, t, spk)                                               # srn_estimator_forward_t
        m8 = mask[:, 0].to(torch.uint8).contiguous()
        l8 = mask_l[:, 0].to(torch.uint8).contiguous() if mask_l is not None else None
        want.append(float(loss_ref))
        prepared.append((v.contiguous(), x1.contiguous(), z.contiguous(), m8, l8))
    torch.cuda.synchronize()
    torch_bytes = torch.cuda.memory_allocated(dev)
    for i, (v, x1, z, m8, l8) in enumerate(prepared):        # the two calls in a row, nothing in between
        _noalloc_loss(lib, v, x1, z, m8, l8, cfm.sigma_min, outs[i:i + 1])
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == torch_bytes   # (nor did the binding allocate through torch)
    got = outs.cpu().tolist()
    for w, gt in zip(want, got):
        assert abs(gt - w) <= LOSS_TOL * abs(w), (gt, w)
    # and once more in the opposite order: the array now holds the second case's partial sums
    again = torch.zeros(2, dtype=torch.float32, device=dev)
    for i in (1, 0):
        v, x1, z, m8, l8 = prepared[i]
        _noalloc_loss(lib, v, x1, z, m8, l8, cfm.sigma_min, again[i:i + 1])
    assert again.cpu().tolist() == got                        # deterministic, independent of the array's old contents


@pytest.mark.gpu
def test_noalloc_loss_rejects_sizes_beyond_its_limit_and_null_arguments():
    from serenade_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a B200"
    dev = torch.device("cuda", 0)
    lib = _lib.load_library()
    small = torch.zeros(1, 1, 8, device=dev)
    m = torch.ones(1, 8, dtype=torch.uint8, device=dev)
    out = torch.zeros(1, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # B * C * T = 2^26 + 80 > SRN_NOALLOC_MAX_ELEMS (the pointers are never dereferenced: the size check comes first)
    rc = lib.srn_cfm_loss_from_velocity_noalloc(small.data_ptr(), small.data_ptr(), small.data_ptr(), m.data_ptr(), None,
                                                1, 80, (1 << 26) // 80 + 1, C.c_float(1e-4), out.data_ptr(), st)
    assert rc != 0 and b"exceeds the limit" in lib.srn_last_error()
    rc = lib.srn_cfm_loss_from_velocity_noalloc(None, small.data_ptr(), small.data_ptr(), m.data_ptr(), None, 1, 1, 8,
                                                C.c_float(1e-4), out.data_ptr(), st)
    assert rc != 0 and b"null argument" in lib.srn_last_error()
