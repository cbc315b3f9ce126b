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


class _Slot:
    """Pinned host staging + device staging + events for one in-flight batch of the pipelined decode loop."""

    def __init__(self, n_x: int, n_mu: int, n_spk: int, n_mask: int, device):
        self.hx = torch.empty(n_x, dtype=torch.float32).pin_memory()
        self.hmu = torch.empty(n_mu, dtype=torch.float32).pin_memory()
        self.hspk = torch.empty(n_spk, dtype=torch.float32).pin_memory()
        self.hmask = torch.empty(n_mask, dtype=torch.uint8).pin_memory()
        self.hout = torch.empty(n_x, dtype=torch.float32).pin_memory()
        self.dx = torch.empty(n_x, dtype=torch.float32, device=device)
        self.dmu = torch.empty(n_mu, dtype=torch.float32, device=device)
        self.dspk = torch.empty(n_spk, dtype=torch.float32, device=device)
        self.dmask = torch.empty(n_mask, dtype=torch.uint8, device=device)
        self.ev_h2d = torch.cuda.Event()
        self.ev_solve = torch.cuda.Event()
        self.ev_d2h = torch.cuda.Event()
        self.batch: Optional[List[int]] = None      # clip indices whose mels sit in hout (None: slot free)
        self.shape = (0, 0)


def decode_sharded_pipelined(cfm, clips: Sequence[Dict[str, torch.Tensor]], t_span: torch.Tensor, rank: int = 0,
                             world_size: int = 1, max_batch: int = 16, device=None,
                             out: Optional[Dict[int, torch.Tensor]] = None) -> Dict[int, torch.Tensor]:
    """`decode_sharded` for a serenade_b200 `CFM`, with the host side overlapped with the GPU (the reference's decode
    loop, bin/ssc_decode.py:346-455, is strictly serial: load, H2D, decode, D2H, write).  Two slots of pinned staging
    buffers: while batch i solves, the host pads batch i+1 into the other slot and its H2D copy runs on a copy stream;
    the mel batch of i comes back as ONE D2H copy on a third stream during solve i+1.  Ordering is by CUDA events; the
    host only blocks when it needs a slot whose download has not finished.  `srn_solve_euler` itself does not block the
    host (the t_span tables stay on the device between calls).  Returns {clip index: mel (80, L)} on the CPU."""
    device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    lengths = [int(c["mu"].shape[-1]) for c in clips]
    mine = assign_clips(lengths, world_size)[rank]
    batches = make_batches(mine, lengths, max_batch=max_batch)
    out = {} if out is None else out
    if not batches:
        return out
    c_mu = int(clips[mine[0]]["mu"].shape[0])
    c_x = int(clips[mine[0]]["z"].shape[0])
    c_spk = int(clips[mine[0]]["spk"].shape[0])
    t_max = max(lengths[i] for i in mine)
    bmax = max(len(b) for b in batches)
    key = (bmax, c_x, c_mu, c_spk, t_max, str(device))
    slots = getattr(cfm, "_srn_pipeline_slots", None)
    if slots is None or slots[0] != key:
        slots = (key, [_Slot(bmax * c_x * t_max, bmax * c_mu * t_max, bmax * c_spk, bmax * t_max, device) for _ in range(2)],
                 torch.cuda.Stream(device), torch.cuda.Stream(device))
        cfm._srn_pipeline_slots = slots
    _, slot_list, h2d_stream, d2h_stream = slots
    main = torch.cuda.current_stream(device)
    ts_cpu = t_span.detach().to("cpu", torch.float32)

    def harvest(s: _Slot):
        s.ev_d2h.synchronize()
        B, T = s.shape
        mel = s.hout[:B * c_x * T].view(B, c_x, T)
        for b, i in enumerate(s.batch):
            out[i] = mel[b, :, :lengths[i]].clone()
        s.batch = None

    for k, idx in enumerate(batches):
        s = slot_list[k % 2]
        if s.batch is not None:
            harvest(s)      # (also guarantees that solve k-2 has finished reading this slot's device buffers)
        B, T = len(idx), max(lengths[i] for i in idx)
        hx = s.hx[:B * c_x * T].view(B, c_x, T)
        hmu = s.hmu[:B * c_mu * T].view(B, c_mu, T)
        hspk = s.hspk[:B * c_spk].view(B, c_spk)
        hmask = s.hmask[:B * T].view(B, T)
        for b, i in enumerate(idx):
            L = lengths[i]
            hx[b, :, :L] = clips[i]["z"]
            hmu[b, :, :L] = clips[i]["mu"]
            hspk[b] = clips[i]["spk"]
            hmask[b, :L] = 1
            if L < T:
                hx[b, :, L:] = 0
                hmu[b, :, L:] = 0
                hmask[b, L:] = 0
        dx = s.dx[:B * c_x * T].view(B, c_x, T)
        dmu = s.dmu[:B * c_mu * T].view(B, c_mu, T)
        dspk = s.dspk[:B * c_spk].view(B, c_spk)
        dmask = s.dmask[:B * T].view(B, T)
        with torch.cuda.stream(h2d_stream):
            dx.copy_(hx, non_blocking=True)
            dmu.copy_(hmu, non_blocking=True)
            dspk.copy_(hspk, non_blocking=True)
            dmask.copy_(hmask, non_blocking=True)
            s.ev_h2d.record(h2d_stream)
        main.wait_event(s.ev_h2d)
        x = cfm.solve_euler(dx, ts_cpu, dmu, dmask, dspk)
        s.ev_solve.record(main)
        with torch.cuda.stream(d2h_stream):
            d2h_stream.wait_event(s.ev_solve)
            s.hout[:B * c_x * T].view(B, c_x, T).copy_(x, non_blocking=True)
            x.record_stream(d2h_stream)
            s.ev_d2h.record(d2h_stream)
        s.batch, s.shape = list(idx), (B, T)
    for s in slot_list:
        if s.batch is not None:
            harvest(s)
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
