"""Drop-in front end for the reference's CFM decoder (the inner operator API of the hot path).

Mirrors serenade/models/matcha_components/flow_matching.py (`CFM.__init__` :10-34, `.inference` :39-63,
`.solve_euler` :65-93) and decoder.py (`Decoder.forward` :384-467): same constructor kwargs, same method
names / argument meaning, same `state_dict` keys (`estimator.*`), so a reference checkpoint loads with
`strict=True` and `Serenade.cfm_decoder` can be swapped for this class unchanged.

All arithmetic happens in libserenade_b200.so (hand-written sm_100a kernels) through the C ABI in
include/serenade_b200.h; PyTorch only owns the tensors and the stream.  There is no CPU / eager fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional, Sequence

import torch
import torch.nn as nn

from . import _lib
from .synth import DecoderGeometry, estimator_param_shapes


class _ParamNode(nn.Module):
    """Parameter container reproducing the reference module tree's names (no compute)."""


def _attach(root: nn.Module, dotted: str, p: nn.Parameter) -> None:
    parts = dotted.split(".")
    node = root
    for name in parts[:-1]:
        if name not in node._modules:
            node.add_module(name, _ParamNode())
        node = node._modules[name]
    node.register_parameter(parts[-1], p)


class _Engine:
    """One C-ABI handle (weights + plans) on one CUDA device."""

    def __init__(self, geom: DecoderGeometry, precision: int, device: torch.device):
        self.lib = _lib.load_library()
        self.device = device
        cfg = _lib.SrnConfig(geom.in_channels, geom.out_channels, geom.spk_embed_dim, geom.channels[0],
                             geom.num_heads, geom.attention_head_dim, geom.num_mid_blocks, precision)
        h = C.c_void_p()
        with torch.cuda.device(device):
            _lib.check(self.lib.srn_create(C.byref(cfg), C.byref(h)), "srn_create")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.srn_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        for key, v in sd.items():
            t = v.detach().to("cpu", torch.float32).contiguous()
            dims = (C.c_int64 * t.dim())(*t.shape)
            _lib.check(self.lib.srn_load_weight(self.h, key.encode(), C.c_void_p(t.data_ptr()), t.dim(), dims),
                       f"srn_load_weight({key})")
        with torch.cuda.device(self.device):
            _lib.check(self.lib.srn_finalize_weights(self.h), "srn_finalize_weights")

    @property
    def launches(self) -> int:
        return int(self.lib.srn_launch_count(self.h))

    def set_use_graph(self, flag: bool) -> None:
        _lib.check(self.lib.srn_set_use_graph(self.h, int(flag)), "srn_set_use_graph")

    def set_gn_valid_only(self, flag: bool) -> None:
        _lib.check(self.lib.srn_set_gn_valid_only(self.h, int(flag)), "srn_set_gn_valid_only")

    def set_option(self, name: str, value: int) -> None:
        """Run-time options of the C ABI (include/serenade_b200.h: srn_set_option), e.g. ("noshift", 0)."""
        _lib.check(self.lib.srn_set_option(self.h, name.encode(), int(value)), f"srn_set_option({name})")

    def softmax_stats(self):
        """(shift-free attention entries, all entries) of the most recent solve / estimator call (synchronises)."""
        out = (C.c_int64 * 2)()
        _lib.check(self.lib.srn_softmax_stats(self.h, out), "srn_softmax_stats")
        return int(out[0]), int(out[1])

    def profile_ops(self, B: int, T: int, reps: int = 5):
        """[(op name, kernel, avg ms)] for one estimator forward at (B, T): CUDA-event timing per op."""
        import json
        buf = C.create_string_buffer(1 << 17)
        _lib.check(self.lib.srn_profile_ops(self.h, B, T, reps, buf, len(buf)), "srn_profile_ops")
        return json.loads(buf.value.decode())

    # test hooks -----------------------------------------------------------------------------
    def debug_enable(self, flag: bool) -> None:
        _lib.check(self.lib.srn_debug_enable(self.h, int(flag)), "srn_debug_enable")

    def debug_fetch(self, name: str) -> torch.Tensor:
        """Named intermediate of the last forward as (nb, rows, cols) f32 (channels-last)."""
        n, rows, cols = C.c_int64(), C.c_int32(), C.c_int32()
        _lib.check(self.lib.srn_debug_fetch(self.h, name.encode(), None, 0, C.byref(n), C.byref(rows), C.byref(cols)),
                   f"srn_debug_fetch({name})")
        out = torch.empty(n.value, dtype=torch.float32)
        _lib.check(self.lib.srn_debug_fetch(self.h, name.encode(), C.c_void_p(out.data_ptr()), n.value, C.byref(n),
                                            C.byref(rows), C.byref(cols)), f"srn_debug_fetch({name})")
        return out.view(-1, rows.value, cols.value)


def _as_mask_u8(mask: torch.Tensor, B: int, T: int, device) -> torch.Tensor:
    m = mask
    if m.dim() == 3:
        if m.shape[1] != 1:
            raise ValueError(f"mask must be (B, 1, T), got {tuple(mask.shape)}")
        m = m[:, 0]
    if tuple(m.shape) != (B, T):
        raise ValueError(f"mask shape {tuple(mask.shape)} does not match batch {B} x {T} frames")
    # the C ABI reads "non-zero = valid": a bool or uint8 mask already on the device is passed as it is (a bool tensor
    # is one 0 / 1 byte per element), without the compare + cast kernels
    if m.device == torch.device(device) and m.is_contiguous():
        if m.dtype == torch.uint8:
            return m
        if m.dtype == torch.bool:
            return m.view(torch.uint8)
    return (m != 0).to(device=device, dtype=torch.uint8).contiguous()


def _tspan_host(t_span: torch.Tensor) -> torch.Tensor:
    """t_span as a contiguous f32 CPU tensor for the entry points that take the grid from host memory (solve_euler itself
    keeps a CUDA t_span on the device: srn_solve_euler_tdev)."""
    return t_span.detach().to("cpu", torch.float32).contiguous()


class Decoder(nn.Module):
    """`v = estimator(x, mask, mu, t, speaker_features)`; reference decoder.py:208-467."""

    def __init__(self, in_channels, out_channels, spk_embed_dim, channels=(512, 512), dropout=0.05,
                 attention_head_dim=512, n_blocks=1, num_mid_blocks=2, num_heads=4, act_fn="snake",
                 precision="bf16", gn_stats="padded", **_unused):
        super().__init__()
        if gn_stats not in ("padded", "valid"):
            raise ValueError("gn_stats must be 'padded' (the reference's padded-batch behaviour) or 'valid'")
        # 'valid': GroupNorm statistics over each utterance's own frames only, so that a padded batch reproduces the
        # reference's batch-1 decode of every utterance (srn_set_gn_valid_only); may be changed between calls
        self.gn_stats = gn_stats
        if n_blocks != 1:
            raise ValueError("only n_blocks=1 (the reference's setting) is supported")
        self.geom = DecoderGeometry(in_channels=in_channels, out_channels=out_channels, spk_embed_dim=spk_embed_dim,
                                    channels=tuple(channels), attention_head_dim=attention_head_dim,
                                    num_heads=num_heads, num_mid_blocks=num_mid_blocks)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.precision = {"bf16": _lib.PRECISION_BF16, "tf32": _lib.PRECISION_TF32}[precision]
        for key, shape in estimator_param_shapes(self.geom).items():
            _attach(self, key[len("estimator."):], nn.Parameter(torch.empty(shape), requires_grad=False))
        self.initialize_weights()
        self._engine: Optional[_Engine] = None
        self._dirty = True            # packed device copies are stale: set by load_state_dict / .to() / mark_weights_dirty()
        self._extra_state = None      # set by models.Serenade: callable -> {key: tensor} of encoder.* / gst.* weights
        self.register_load_state_dict_post_hook(lambda module, _keys: module.mark_weights_dirty())

    def initialize_weights(self):
        """Same distributions as reference decoder.py:366-382 (+ torch defaults for the ConvTranspose1d)."""
        for name, p in self.named_parameters():
            is_norm = (".block.1." in name) or (".norm1." in name) or (".norm3." in name)
            is_convT = name.startswith("up_blocks.") and ".2.conv." in name
            with torch.no_grad():
                if is_norm:
                    p.fill_(1.0 if name.endswith("weight") else 0.0)
                elif is_convT:
                    fan_in = self.geom.channels[0] * 4
                    p.uniform_(-1.0 / math.sqrt(fan_in), 1.0 / math.sqrt(fan_in))
                elif name.endswith("weight"):
                    nn.init.kaiming_normal_(p, nonlinearity="relu")
                else:
                    p.zero_()

    # ---------------------------------------------------------------- engine plumbing
    def mark_weights_dirty(self):
        """The packed / bf16 device copies are derived buffers: they are rebuilt on the next call after
        `load_state_dict` or `.to()` (both tracked here).  Call this after modifying a parameter in place by hand."""
        self._dirty = True

    def _apply(self, fn, *args, **kwargs):
        self._dirty = True
        return super()._apply(fn, *args, **kwargs)

    def engine(self, device: torch.device) -> _Engine:
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("serenade_b200 runs on a B200 (sm_100a) only; there is no CPU path. "
                               f"Got tensors on '{device}'.")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if self._engine is None or self._engine.device != device:
            if self._engine is not None:
                self._engine.close()
            self._engine = _Engine(self.geom, self.precision, device)
            self._dirty = True
        if self._dirty:
            sd = {"estimator." + k: v for k, v in self.state_dict().items()}
            if self._extra_state is not None:
                sd.update(self._extra_state())
            self._engine.load_state_dict(sd)
            self._dirty = False
        want = self.gn_stats == "valid"
        if getattr(self._engine, "_gn_valid", False) != want:
            self._engine.set_gn_valid_only(want)
            self._engine._gn_valid = want
        return self._engine

    def _check_inputs(self, x, mask, mu, spk):
        B, C_x, T = x.shape
        if C_x != self.out_channels:
            raise ValueError(f"x must have {self.out_channels} channels, got {C_x}")
        if mu.shape != (B, self.in_channels - self.out_channels, T):
            raise ValueError(f"mu must be {(B, self.in_channels - self.out_channels, T)}, got {tuple(mu.shape)}")
        if spk is None or tuple(spk.shape) != (B, self.geom.spk_embed_dim):
            raise ValueError(f"speaker_features must be {(B, self.geom.spk_embed_dim)}")
        if T < 2:
            raise ValueError("need at least 2 frames")

    def forward(self, x, mask, mu, t, speaker_features):
        self._check_inputs(x, mask, mu, speaker_features)
        eng = self.engine(x.device)
        B, _, T = x.shape
        dev = x.device
        xf = x.detach().to(torch.float32).contiguous()
        muf = mu.detach().to(device=dev, dtype=torch.float32).contiguous()
        spk = speaker_features.detach().to(device=dev, dtype=torch.float32).contiguous()
        m8 = _as_mask_u8(mask, B, T, dev)
        out = torch.empty_like(xf)
        t_dev = None
        if torch.is_tensor(t) and t.is_cuda:
            # a time that already lives on the GPU (the reference's solve_euler hands its estimator a 0-dim CUDA tensor,
            # flow_matching.py:79-86) stays there: no device-to-host read, the call does not wait for the stream
            t_dev = t.detach().reshape(-1).to(device=dev, dtype=torch.float32).contiguous()
            n_t = t_dev.numel()
        else:
            tv = t.detach().reshape(-1).to(torch.float32) if torch.is_tensor(t) else torch.tensor([float(t)])
            n_t = tv.numel()
        if n_t not in (1, B):
            raise ValueError(f"t must be a scalar or hold one time per clip ({B}), got {n_t} values")
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            if t_dev is not None:
                _lib.check(eng.lib.srn_estimator_forward_tdev(eng.h, xf.data_ptr(), muf.data_ptr(), m8.data_ptr(),
                                                              spk.data_ptr(), t_dev.data_ptr(), n_t, B, T, out.data_ptr(),
                                                              C.c_void_p(st)), "srn_estimator_forward_tdev")
            elif n_t == 1:
                _lib.check(eng.lib.srn_estimator_forward(eng.h, xf.data_ptr(), muf.data_ptr(), m8.data_ptr(),
                                                         spk.data_ptr(), C.c_float(float(tv[0])), B, T, out.data_ptr(),
                                                         C.c_void_p(st)), "srn_estimator_forward")
            else:       # one time per clip (decoder.py:401-402), as CFM.compute_loss calls it
                tarr = (C.c_float * B)(*tv.tolist())
                _lib.check(eng.lib.srn_estimator_forward_t(eng.h, xf.data_ptr(), muf.data_ptr(), m8.data_ptr(),
                                                           spk.data_ptr(), tarr, B, T, out.data_ptr(), C.c_void_p(st)),
                           "srn_estimator_forward_t")
        return out


class CFM(nn.Module):
    """Reference flow_matching.py:9-133: inference, and the forward value of `compute_loss` (no backward pass)."""

    def __init__(self, in_channels=80, out_channels=80, solver="euler", sigma_min=1e-4, spk_embed_dim=256,
                 decoder_channels=(512, 512), decoder_attention_head_dim=256, precision="bf16", gn_stats="padded",
                 shape_bucket=0):
        super().__init__()
        # shape_bucket > 0 (with gn_stats="valid" only): solve_euler pads T up to the next multiple with masked frames.
        # The engine keeps one plan (workspace, TMA maps, captured CUDA graph) per (B, T); a decode loop over utterances
        # of arbitrary lengths would build a new one - milliseconds of cudaMalloc, descriptor encoding and graph
        # instantiation - for almost every call.  In valid-frame GroupNorm mode masked frames do not change any valid
        # frame (bit for bit: tests/test_gpu_parity.py), so buckets cost nothing but the padded work.
        self.shape_bucket = int(shape_bucket)
        self.n_feats = in_channels
        self.spk_embed_dim = spk_embed_dim
        self.solver = solver
        self.sigma_min = sigma_min
        self.conditioning_shape = in_channels + out_channels
        self.out_channels = out_channels
        self.estimator = Decoder(in_channels=in_channels, out_channels=out_channels, spk_embed_dim=spk_embed_dim,
                                 channels=decoder_channels, attention_head_dim=decoder_attention_head_dim,
                                 precision=precision, gn_stats=gn_stats)

    def forward(self, x1, mask, mu, spks, mask_l=None):
        return self.compute_loss(x1, mask, mu, spks, mask_l)

    @torch.inference_mode()
    def compute_loss(self, x1, mask, mu, spks, mask_l=None, *, z=None, t=None):
        """Reference flow_matching.py:95-133, FORWARD VALUE ONLY -> (loss, y): the number `trainers/ssc.py` logs on the
        dev set.  The returned tensors carry no graph (this library has no backward pass; train with the reference).
        `t` (B,) and `z` like x1 are drawn as the reference draws them (device generator, t first) unless given."""
        B, _, T = mu.shape
        dev = mu.device
        if t is None:
            t = torch.rand([B, 1, 1], device=dev, dtype=torch.float32)
        if z is None:
            z = torch.randn_like(x1, dtype=torch.float32)
        est = self.estimator
        est._check_inputs(x1, mask, mu, spks)
        eng = est.engine(dev)
        x1f = x1.detach().to(device=dev, dtype=torch.float32).contiguous()
        zf = z.detach().to(device=dev, dtype=torch.float32).contiguous()
        muf = mu.detach().to(torch.float32).contiguous()
        spk = spks.detach().to(device=dev, dtype=torch.float32).contiguous()
        m8 = _as_mask_u8(mask, B, T, dev)
        l8 = _as_mask_u8(mask_l, B, T, dev) if mask_l is not None else None
        if t.numel() != B:
            raise ValueError(f"t must hold one time per clip ({B})")
        y = torch.empty_like(x1f)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            tail = (muf.data_ptr(), m8.data_ptr(), spk.data_ptr(), l8.data_ptr() if l8 is not None else None, B, T,
                    C.c_float(self.sigma_min), y.data_ptr(), loss.data_ptr(), C.c_void_p(st))
            if t.is_cuda:       # drawn on the device like the reference's (flow_matching.py:115): it stays there
                td = t.detach().reshape(-1).to(device=dev, dtype=torch.float32).contiguous()
                _lib.check(eng.lib.srn_cfm_loss_tdev(eng.h, x1f.data_ptr(), zf.data_ptr(), td.data_ptr(), *tail),
                           "srn_cfm_loss_tdev")
            else:
                tarr = (C.c_float * B)(*t.detach().reshape(-1).to(torch.float32).tolist())
                _lib.check(eng.lib.srn_cfm_loss(eng.h, x1f.data_ptr(), zf.data_ptr(), tarr, *tail), "srn_cfm_loss")
        return loss, y

    @torch.inference_mode()
    def inference(self, mu, mask, n_timesteps=10, temperature=0.667, spks=None):
        # noise is drawn on the CPU generator then moved, exactly like flow_matching.py:57-60
        z = torch.randn((mu.shape[0], self.out_channels, mu.shape[2])).to(mu.device) * temperature
        t_span = torch.linspace(0, 1, n_timesteps + 1)       # (host: srn_solve_euler takes the grid from host memory)
        return self.solve_euler(z, t_span=t_span, mu=mu, mask=mask, trg_spks=spks)

    @torch.inference_mode()
    def solve_euler(self, x, t_span, mu, mask, trg_spks, blend_mask=None, blend_src=None):
        """Fixed-step Euler solve; returns the final iterate and does not mutate `x` (flow_matching.py:87)."""
        est = self.estimator
        est._check_inputs(x, mask, mu, trg_spks)
        eng = est.engine(x.device)
        B, _, T = x.shape
        dev = x.device
        T_in = T
        if self.shape_bucket > 0 and est.gn_stats == "valid" and T % self.shape_bucket != 0 and blend_mask is None:
            pad = self.shape_bucket - T % self.shape_bucket
            x = torch.nn.functional.pad(x.detach().to(torch.float32), (0, pad))
            mu = torch.nn.functional.pad(mu.detach().to(device=dev, dtype=torch.float32), (0, pad))
            m2 = mask[:, 0] if mask.dim() == 3 else mask
            mask = torch.nn.functional.pad(m2.to(dev) != 0, (0, pad))
            T = T + pad
            xs = x.contiguous()
        else:
            xs = x.detach().to(torch.float32).clone().contiguous()
        muf = mu.detach().to(device=dev, dtype=torch.float32).contiguous()
        spk = trg_spks.detach().to(device=dev, dtype=torch.float32).contiguous()
        m8 = _as_mask_u8(mask, B, T, dev)
        # a grid that already lives on the GPU (the reference builds it there, flow_matching.py:61) stays there
        ts_dev = t_span.detach().reshape(-1).to(device=dev, dtype=torch.float32).contiguous() if t_span.is_cuda else None
        ts = _tspan_host(t_span) if ts_dev is None else None
        n_steps = (ts_dev if ts_dev is not None else ts).numel() - 1
        if n_steps < 1:
            raise ValueError("t_span needs at least two points")
        bm = bs = None
        if blend_mask is not None:
            bm = _as_mask_u8(blend_mask, B, T, dev)
            bs = blend_src.detach().to(device=dev, dtype=torch.float32).contiguous()
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            if ts_dev is not None:
                _lib.check(eng.lib.srn_solve_euler_tdev(eng.h, xs.data_ptr(), muf.data_ptr(), m8.data_ptr(), spk.data_ptr(),
                                                        ts_dev.data_ptr(), n_steps, B, T,
                                                        bm.data_ptr() if bm is not None else None,
                                                        bs.data_ptr() if bs is not None else None, C.c_void_p(st)),
                           "srn_solve_euler_tdev")
            else:
                ts_p = C.cast(C.c_void_p(ts.data_ptr()), C.POINTER(C.c_float))
                _lib.check(eng.lib.srn_solve_euler(eng.h, xs.data_ptr(), muf.data_ptr(), m8.data_ptr(), spk.data_ptr(),
                                                   ts_p, n_steps, B, T, bm.data_ptr() if bm is not None else None,
                                                   bs.data_ptr() if bs is not None else None, C.c_void_p(st)),
                           "srn_solve_euler")
        return xs if T == T_in else xs[:, :, :T_in].contiguous()

    @torch.inference_mode()
    def decode_host(self, z, mu, mask, spks, t_span, out=None, device=None):
        """End-to-end decode from HOST tensors (pinned for speed): H2D of z/mu/mask/spk, the Euler solve and
        the D2H of the mel all happen inside the C-ABI call `srn_solve_euler_host`."""
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        est = self.estimator
        est._check_inputs(z, mask, mu, spks)
        eng = est.engine(device)
        B, _, T = z.shape
        for tns in (z, mu, spks):
            if tns.device.type != "cpu" or tns.dtype != torch.float32 or not tns.is_contiguous():
                raise ValueError("decode_host expects contiguous float32 CPU tensors")
        m8 = _as_mask_u8(mask, B, T, "cpu")
        ts = t_span.detach().to("cpu", torch.float32).contiguous()
        if out is None:
            out = torch.empty_like(z)
        ts_p = C.cast(C.c_void_p(ts.data_ptr()), C.POINTER(C.c_float))
        with torch.cuda.device(device):
            st = torch.cuda.current_stream(device).cuda_stream
            _lib.check(eng.lib.srn_solve_euler_host(eng.h, z.data_ptr(), mu.data_ptr(), m8.data_ptr(), spks.data_ptr(),
                                                    ts_p, ts.numel() - 1, B, T, out.data_ptr(), C.c_void_p(st)),
                       "srn_solve_euler_host")
        return out
