"""Host-side multi-GPU logic on CPU: utterance sharding, length-bucketed batching and the reporting gather,
world size 2 over gloo.  The decode callable is a stand-in (the CUDA path has its own -m gpu tests); what is
checked here is that sharding covers every clip exactly once, balances the cost model, and that per-clip
results do not depend on the rank / batch a clip lands in."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from serenade_b200 import sharding
from serenade_b200.synth import estimator_flops


def _clips(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(30, 200, (n,), generator=g).tolist()
    return [dict(mu=torch.randn(162, L, generator=g), z=torch.randn(80, L, generator=g),
                 spk=torch.randn(256, generator=g)) for L in lens]


def _fake_solve(z, t_span, mu, mask, spk):
    # per-utterance, padding-independent stand-in with solve_euler's signature
    m = mask.float()
    return (z + mu[:, :80] * float(t_span[-1]) + spk[:, :80, None]) * m


def test_assignment_is_a_balanced_partition():
    lengths = [int(x) for x in torch.randint(300, 2001, (64,), generator=torch.Generator().manual_seed(1))]
    for world in (1, 2, 4, 8):
        parts = sharding.assign_clips(lengths, world)
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(64))
        loads = [sum(estimator_flops(lengths[i]) for i in p) for p in parts]
        assert max(loads) / (sum(loads) / world) < 1.10


def test_batches_respect_limits_and_sort_descending():
    lengths = [100, 90, 95, 30, 35, 200, 10]
    batches = sharding.make_batches(range(7), lengths, max_batch=3)
    assert sorted(i for b in batches for i in b) == list(range(7))
    assert all(len(b) <= 3 for b in batches)
    for b in batches:
        assert [lengths[i] for i in b] == sorted((lengths[i] for i in b), reverse=True)
    small = sharding.make_batches(range(7), lengths, max_batch=8, max_padded_frames=220)
    assert all(len(b) * lengths[b[0]] <= 220 or len(b) == 1 for b in small)


def test_sharding_properties_hold_for_arbitrary_inputs():
    """Property test (hypothesis): for any lengths, world size and limits - every clip lands on exactly one rank and in
    exactly one batch, ranks differ by at most one clip's cost, batches are sorted longest-first and respect both limits,
    and the assignment is deterministic."""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    @settings(max_examples=200, deadline=None)
    @given(lengths=st.lists(st.integers(2, 6000), min_size=0, max_size=80), world=st.integers(1, 8),
           max_batch=st.integers(1, 16), limit=st.one_of(st.none(), st.integers(6000, 40000)))
    def check(lengths, world, max_batch, limit):
        parts = sharding.assign_clips(lengths, world)
        assert len(parts) == world
        assert sorted(i for p in parts for i in p) == list(range(len(lengths)))
        assert parts == sharding.assign_clips(list(lengths), world)
        loads = [sum(estimator_flops(lengths[i]) for i in p) for p in parts]
        if lengths:       # greedy longest-first: no rank exceeds another by more than the largest single clip
            assert max(loads) - min(loads) <= max(estimator_flops(L) for L in lengths) + 1e-6
        for p in parts:
            batches = sharding.make_batches(p, lengths, max_batch=max_batch, max_padded_frames=limit)
            assert sorted(i for b in batches for i in b) == sorted(p)
            for b in batches:
                ls = [lengths[i] for i in b]
                assert 1 <= len(b) <= max_batch and ls == sorted(ls, reverse=True)
                if limit is not None:
                    assert len(b) == 1 or len(b) * ls[0] <= limit
            firsts = [lengths[b[0]] for b in batches]
            assert firsts == sorted(firsts, reverse=True)
    check()


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    clips = _clips(n)
    t_span = torch.linspace(0, 1, 4)
    local = sharding.decode_sharded(_fake_solve, clips, t_span, rank=rank, world_size=world, max_batch=4)
    merged = sharding.gather_mels(local, n)
    single = sharding.decode_sharded(_fake_solve, clips, t_span, rank=0, world_size=1, max_batch=3)
    ok = all(m is not None and torch.equal(m, single[i]) for i, m in enumerate(merged))
    q.put((rank, ok, sorted(local.keys())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_decode_matches_single_rank():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n = 11
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    keys = [k for _, _, ks in res for k in ks]
    assert sorted(keys) == list(range(n))            # every clip decoded exactly once across ranks


def test_gather_without_process_group_orders_local_results():
    merged = sharding.gather_mels({1: torch.ones(2), 0: torch.zeros(2)}, 3)
    assert merged[2] is None and float(merged[1].sum()) == 2.0
