"""Load the UNMODIFIED reference hot-path modules from /root/reference.

TEST INFRASTRUCTURE ONLY (never imported by the product package).

Only usable in the build container: /root/reference does not exist on the GPU
box.  It is used (a) by tests/golden/make_golden.py to generate the committed
golden vectors and (b) by CPU tests that pin oracle/cfm_oracle.py against the
real reference when the tree is present.

The three files flow_matching.py / decoder.py / transformer.py under
serenade/models/matcha_components/ import `diffusers`, which is an unpinned,
un-vendored dependency (reference setup.cfg:21) and is not installed here.  The
shim below restates the published semantics of the few diffusers symbols the
path touches:

  Attention(query_dim, heads, dim_head, bias=False)  -> to_q/to_k/to_v no-bias
      Linear, F.scaled_dot_product_attention with a boolean key-padding mask
      (True = attend), to_out = [Linear(bias=True), Dropout]
  GEGLU(dim_in, dim_out) -> proj = Linear(dim_in, 2*dim_out);
      h, g = chunk(2); h * gelu_erf(g)
  LoRACompatibleLinear == nn.Linear, get_activation("silu") == nn.SiLU,
  maybe_allow_in_graph == identity.
"""
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REFERENCE_ROOT = os.environ.get("SERENADE_REFERENCE_ROOT", "/root/reference")
_MC = "serenade/models/matcha_components"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, _MC, "decoder.py"))


class _Attention(nn.Module):
    """diffusers.models.attention_processor.Attention, self-attention subset."""

    def __init__(self, query_dim, heads=8, dim_head=64, dropout=0.0, bias=False,
                 cross_attention_dim=None, upcast_attention=False, **_):
        super().__init__()
        assert cross_attention_dim is None
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(query_dim, inner, bias=bias)
        self.to_v = nn.Linear(query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(dropout)])

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **_):
        b, s, _c = hidden_states.shape
        h = self.heads
        q = self.to_q(hidden_states).view(b, s, h, -1).transpose(1, 2)
        k = self.to_k(hidden_states).view(b, s, h, -1).transpose(1, 2)
        v = self.to_v(hidden_states).view(b, s, h, -1).transpose(1, 2)
        m = None
        if attention_mask is not None:
            # prepare_attention_mask: (B,S) -> repeat_interleave(heads) -> (B,H,1,S)
            m = attention_mask.repeat_interleave(h, dim=0).view(b, h, 1, s)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=m, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(b, s, -1)
        o = self.to_out[0](o)
        return self.to_out[1](o)


class _GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, x):
        h, g = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(g)


def _install_shims():
    if "diffusers" in sys.modules and not getattr(sys.modules["diffusers"], "_srn_shim", False):
        return  # a real diffusers is importable: use it
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        m.__path__ = []
        sys.modules[name] = m
        return m
    placeholder = type("_Unused", (nn.Module,), {})
    mod("diffusers", _srn_shim=True)
    mod("diffusers.models")
    mod("diffusers.models.activations", get_activation=lambda n: {"silu": nn.SiLU(), "swish": nn.SiLU(),
                                                                  "mish": nn.Mish(), "gelu": nn.GELU()}[n])
    mod("diffusers.models.attention", GEGLU=_GEGLU, GELU=placeholder, AdaLayerNorm=placeholder,
        AdaLayerNormZero=placeholder, ApproximateGELU=placeholder)
    mod("diffusers.models.attention_processor", Attention=_Attention)
    mod("diffusers.models.lora", LoRACompatibleLinear=nn.Linear)
    mod("diffusers.utils", )
    mod("diffusers.utils.torch_utils", maybe_allow_in_graph=lambda c: c)
    for pkg in ("serenade", "serenade.models", "serenade.models.matcha_components"):
        if pkg not in sys.modules:
            mod(pkg)


_cache = {}


def load_reference():
    """Returns the reference `flow_matching` module (exposes CFM)."""
    if "fm" in _cache:
        return _cache["fm"]
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")
    _install_shims()
    out = None
    for name in ("transformer", "decoder", "flow_matching"):
        dotted = f"serenade.models.matcha_components.{name}"
        spec = importlib.util.spec_from_file_location(dotted, os.path.join(REFERENCE_ROOT, _MC, name + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[dotted] = m
        spec.loader.exec_module(m)
        out = m
    _cache["fm"] = out
    return out


def build_reference_cfm(in_channels=242, out_channels=80, spk_embed_dim=256,
                        decoder_channels=(512, 512), decoder_attention_head_dim=512):
    """CFM exactly as Serenade.__init__ builds it (reference serenade.py:77-87)."""
    fm = load_reference()
    return fm.CFM(in_channels=in_channels, out_channels=out_channels, spk_embed_dim=spk_embed_dim,
                  decoder_channels=decoder_channels,
                  decoder_attention_head_dim=decoder_attention_head_dim).eval()


def load_reference_serenade():
    """Returns the reference `serenade.models.serenade` module (exposes Serenade, Conv1dResnet) with the real
    StyleEncoder / attention / masking files loaded by path (their package __init__ files pull in h5py etc.)."""
    if "srn" in _cache:
        return _cache["srn"]
    load_reference()
    def mod(name):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
    for pkg in ("serenade.utils", "serenade.modules", "serenade.modules.gst"):
        mod(pkg)
    if "typeguard" not in sys.modules:
        try:
            import typeguard  # noqa: F401
        except Exception:
            tg = types.ModuleType("typeguard")
            tg.typechecked = lambda f: f
            sys.modules["typeguard"] = tg
    out = None
    for dotted, rel in (("serenade.utils.masking", "serenade/utils/masking.py"),
                        ("serenade.modules.gst.attention", "serenade/modules/gst/attention.py"),
                        ("serenade.modules.gst.style_encoder", "serenade/modules/gst/style_encoder.py"),
                        ("serenade.models.serenade", "serenade/models/serenade.py")):
        spec = importlib.util.spec_from_file_location(dotted, os.path.join(REFERENCE_ROOT, rel))
        m = importlib.util.module_from_spec(spec)
        sys.modules[dotted] = m
        spec.loader.exec_module(m)
        out = m
    _cache["srn"] = out
    return out


def build_reference_serenade(**kw):
    """Serenade() with the ssc1 recipe's parameters (conf/serenade.yaml: all constructor defaults)."""
    return load_reference_serenade().Serenade(**kw).eval()


def load_reference_hifigan():
    """Returns the reference `serenade.vocoder.models.hifigan` module (HiFiGANGenerator) loaded by file path.
    `serenade.utils.read_hdf5` (h5py, absent here) is only used by register_stats and is stubbed."""
    if "hifigan" in _cache:
        return _cache["hifigan"]
    _install_shims()
    def mod(name, **attrs):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
        for k, v in attrs.items():
            if not hasattr(sys.modules[name], k):
                setattr(sys.modules[name], k, v)
        return sys.modules[name]
    def _no_h5(*a, **k):
        raise RuntimeError("h5py is not available in this container")
    mod("serenade")
    mod("serenade.utils", read_hdf5=_no_h5)
    mod("serenade.vocoder")
    layers = mod("serenade.vocoder.layers")
    mod("serenade.vocoder.models")
    out = None
    for dotted, rel in (("serenade.vocoder.layers.causal_conv", "serenade/vocoder/layers/causal_conv.py"),
                        ("serenade.vocoder.layers.residual_block", "serenade/vocoder/layers/residual_block.py"),
                        ("serenade.vocoder.models.hifigan", "serenade/vocoder/models/hifigan.py")):
        spec = importlib.util.spec_from_file_location(dotted, os.path.join(REFERENCE_ROOT, rel))
        m = importlib.util.module_from_spec(spec)
        sys.modules[dotted] = m
        spec.loader.exec_module(m)
        for k in dir(m):                      # the package __init__ does `from .x import *`
            if not k.startswith("_") and not hasattr(layers, k) and "layers" in dotted:
                setattr(layers, k, getattr(m, k))
        out = m
    _cache["hifigan"] = out
    return out

