"""Utterance sharding for multi-GPU decode (one process per GPU, no collective on the hot path).

The reference decodes one utterance per call and fans out over dump shards with separate processes
(egs/gtsinger/ssc1/run.sh:147, bin/ssc_decode.py:346-378).  Here clips are assigned to ranks by a
longest-first greedy on the estimator's cost model, grouped into length-bucketed batches, decoded with
`CFM.solve_euler`-compatible callables, and (optionally) gathered to every rank for reporting.
Only the gather touches torch.distributed; it is never inside a timed region.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from .synth import estimator_flops


def assign_clips(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-first greedy partition of clip indices over ranks, balanced on estimator FLOPs
    (cost ~ 81e6 L + 24.6e3 L^2).  Deterministic; every rank computes the same answer."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    load = [0.0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += estimator_flops(int(lengths[i]))
    return out


def make_batches(indices: Sequence[int], lengths: Sequence[int], max_batch: int = 16,
                 max_padded_frames: Optional[int] = None) -> List[List[int]]:
    """Group a rank's clips into batches of similar length (sorted descending like collaters/ssc.py:53) so that
    padding is small; a batch never exceeds `max_batch` clips or `max_padded_frames` (= B * T_max) frames."""
    if max_batch < 1:
        raise ValueError("max_batch must be >= 1")
    order = sorted(indices, key=lambda i: (-int(lengths[i]), i))
    batches: List[List[int]] = []
    cur: List[int] = []
    for i in order:
        t_max = int(lengths[cur[0]]) if cur else int(lengths[i])
        if cur and (len(cur) >= max_batch or
                    (max_padded_frames is not None and (len(cur) + 1) * t_max > max_padded_frames)):
            batches.append(cur)
            cur = []
        cur.append(i)
    if cur:
        batches.append(cur)
    return batches


def pad_batch(clips: Sequence[Dict[str, torch.Tensor]], idx: Sequence[int]):
    """Pads clips `idx` (each: mu (C_mu, L), z (80, L), spk (256,)) to one (B, ., T_max) batch + bool mask."""
    T = max(int(clips[i]["mu"].shape[-1]) for i in idx)
    B = len(idx)
    mu = torch.zeros(B, clips[idx[0]]["mu"].shape[0], T)
    z = torch.zeros(B, clips[idx[0]]["z"].shape[0], T)
    spk = torch.stack([clips[i]["spk"] for i in idx])
    mask = torch.zeros(B, 1, T, dtype=torch.bool)
    for b, i in enumerate(idx):
        L = int(clips[i]["mu"].shape[-1])
        mu[b, :, :L] = clips[i]["mu"]
        z[b, :, :L] = clips[i]["z"]
        mask[b, :, :L] = True
    return z, mu, mask, spk


def decode_sharded(solve: Callable[..., torch.Tensor], clips: Sequence[Dict[str, torch.Tensor]], t_span: torch.Tensor,
                   rank: int = 0, world_size: int = 1, max_batch: int = 16, device=None) -> Dict[int, torch.Tensor]:
    """Decodes this rank's share of `clips`.  `solve(z, t_span, mu, mask, spk) -> x` has CFM.solve_euler's
    signature.  Returns {clip index: mel (80, L)} on the CPU.  No collective."""
    lengths = [int(c["mu"].shape[-1]) for c in clips]
    mine = assign_clips(lengths, world_size)[rank]
    out: Dict[int, torch.Tensor] = {}
    for batch in make_batches(mine, lengths, max_batch=max_batch):
        z, mu, mask, spk = pad_batch(clips, batch)
        if device is not None:
            z, mu, mask, spk = z.to(device), mu.to(device), mask.to(device), spk.to(device)
        x = solve(z, t_span, mu, mask, spk)
        for b, i in enumerate(batch):
            out[i] = x[b, :, :lengths[i]].detach().cpu()
    return out


def gather_mels(local: Dict[int, torch.Tensor], n_clips: int) -> List[Optional[torch.Tensor]]:
    """Reporting-only collective: every rank receives all decoded mels, in clip order (all_gather_object over
    the default process group; NCCL or gloo).  With no process group it just orders the local results."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        parts: List[Optional[Dict[int, torch.Tensor]]] = [None] * dist.get_world_size()
        dist.all_gather_object(parts, local)
    else:
        parts = [local]
    merged: List[Optional[torch.Tensor]] = [None] * n_clips
    for p in parts:
        for i, m in (p or {}).items():
            if merged[i] is not None:
                raise RuntimeError(f"clip {i} was decoded by more than one rank")
            merged[i] = m
    return merged
