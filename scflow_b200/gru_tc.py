"""ConvGRU through the fused tcgen05 kernels: scflow_convgru_fwd (both passes, one persistent launch;
the default) and scflow_convgru_pass_fwd (one pass per launch; the cross-check).

Weights stay fp32 nn.Parameters under the reference names; the packed [tap][k-block][z|r|q][32]
copy is a derived cache keyed on the parameters' version counters and storage, rebuilt after
load_state_dict / optimizer steps / .to(device) and never saved.

Operand rounding: tcgen05 kind::tf32 ignores the low 13 mantissa bits of its fp32 operands, a
truncation whose bias adds up coherently over K = 1920.  Every tensor-core operand is therefore
rounded to the NEAREST TF32 by its producer (weights when packed, x by scflow_pack_nhwc, r*h and
the MMA copy of h by the kernel's own epilogues); the gating arithmetic keeps the exact fp32 h."""
from __future__ import annotations

import os

import torch

from . import _lib
from . import ops
from .ops import _chk, _ptr, _stream, _timed


def _pack_key(mod, p):
    ps = [getattr(mod, f"conv_{g}")[p].conv for g in "zrq"]
    return tuple((t.data_ptr(), t._version) for c in ps for t in (c.weight, c.bias))


def packed_weights(mod, p: int) -> torch.Tensor:
    cache = mod.__dict__.setdefault("_scflow_wpack", {})
    key = _pack_key(mod, p)
    hit = cache.get(p)
    if hit is not None and hit[0] == key:
        return hit[1]
    lib = _lib.load()
    cz, cr, cq = (getattr(mod, f"conv_{g}")[p].conv for g in "zrq")
    ch, cx = mod.h_channels, mod.x_channels
    nfl = lib.scflow_convgru_packed_floats(ch, cx)
    if nfl == 0:
        raise _lib.ScflowError(f"ConvGRU({ch},{cx}) is not supported by the tcgen05 kernel (128/256 only)")
    ws = [_chk(c.weight.detach(), "gru weight") for c in (cz, cr, cq)]
    bs = [_chk(c.bias.detach(), "gru bias") for c in (cz, cr, cq)]
    out = torch.empty(nfl, device=ws[0].device, dtype=torch.float32)
    with _timed("scflow_convgru_pack_weights"):
        rc = lib.scflow_convgru_pack_weights(_ptr(ws[0]), _ptr(ws[1]), _ptr(ws[2]), _ptr(bs[0]), _ptr(bs[1]),
                                             _ptr(bs[2]), _ptr(out), ch, cx, p, _stream())
    _lib.check(rc, "scflow_convgru_pack_weights")
    cache[p] = (key, out)
    return out


def round_tf32(t: torch.Tensor) -> torch.Tensor:
    """Nearest-TF32 copy of a channels-last (N,H,W,C) tensor (one launch)."""
    return ops.pack_nhwc(torch.empty_like(t), 0, a=t, round_tf32=True)


def update_inplace(mod, h: torch.Tensor, h_r: torch.Tensor, x_r: torch.Tensor, rh: torch.Tensor) -> None:
    """Both passes of the SeqConv GRU, in place: h (N,H,W,128) exact state, h_r its nearest-TF32
    copy (the tensor-core operand), x_r (N,H,W,256) already rounded, rh scratch like h.
    In place is safe: a tile spans the whole convolution axis, so it reads only its own pixels.
    One persistent launch for both passes (scflow_convgru_fwd); SCFLOW_GRU_FUSED=0 keeps the
    launch-per-pass path (bit-identical results, used as the cross-check)."""
    lib = _lib.load()
    n, H, W, ch = h.shape
    cx = x_r.shape[-1]
    if os.environ.get("SCFLOW_GRU_FUSED", "1") != "0" and "SCFLOW_GRU_2CTA" not in os.environ:
        need = lib.scflow_convgru_workspace_bytes(n, H, W)
        ws = mod.__dict__.get("_scflow_gru_ws")
        if ws is None or ws.device != h.device or ws.numel() * 4 < need:
            ws = torch.empty((need + 3) // 4, device=h.device, dtype=torch.int32)
            mod.__dict__["_scflow_gru_ws"] = ws
        w0, w1 = packed_weights(mod, 0), packed_weights(mod, 1)
        with _timed("scflow_convgru_fwd"):
            rc = lib.scflow_convgru_fwd(_ptr(h), _ptr(h_r), _ptr(x_r), _ptr(w0), _ptr(w1), _ptr(rh), _ptr(ws),
                                        ws.numel() * 4, n, H, W, ch, cx, _stream())
        _lib.check(rc, "scflow_convgru_fwd")
        return
    for p in (0, 1):
        wp = packed_weights(mod, p)
        with _timed("scflow_convgru_pass_fwd"):
            rc = lib.scflow_convgru_pass_fwd(_ptr(h), _ptr(h_r), _ptr(x_r), _ptr(wp), _ptr(h), _ptr(h_r), _ptr(rh),
                                             n, H, W, ch, cx, p, _stream())
        _lib.check(rc, "scflow_convgru_pass_fwd")


def forward_nhwc(mod, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """h (N,H,W,128), x (N,H,W,256) channels-last contiguous -> new h (N,H,W,128)."""
    h, x = _chk(h, "h"), _chk(x, "x")
    out = h.clone()
    update_inplace(mod, out, round_tf32(h), round_tf32(x), torch.empty_like(h))
    return out


def forward(mod, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """NCHW-shaped in / out (reference signature); tensors already in channels_last memory format
    pass through without a copy."""
    if torch.is_grad_enabled() and (h.requires_grad or x.requires_grad):
        raise NotImplementedError("ConvGRU tcgen05 kernel is forward-only in this round")
    hn = h.permute(0, 2, 3, 1).contiguous()
    xn = x.permute(0, 2, 3, 1).contiguous()
    return forward_nhwc(mod, hn, xn).permute(0, 3, 1, 2)
