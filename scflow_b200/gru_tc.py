"""ConvGRU (SeqConv, models/decoder/raft_decoder.py:235-253) on the split-bf16 tensor-core convolution family.

Round 1 ran the gates in TF32 (one tcgen05 MMA per product): 5.7e-5 / 8.9e-5 rad of the 1e-4 rad pose budget
went to this module alone.  Here every product is three bf16 MMAs (csrc/conv_tc.cu), and a pass is two
launches of the same implicit-GEMM kernel with gate epilogues:

    zr    conv over [x | h] (384 ch)        -> z = sigmoid(.) kept in fp32, r = sigmoid(.), r*h as split planes
    gate  conv over [r*h | x] (two tensors)  -> q = tanh(.), h = (1-z) h + z q, in place (fp32) + split planes

The exact fp32 hidden state lives beside its split copy (the MMA operand of the next pass and of the heads).
Weights stay fp32 nn.Parameters under the reference names; the two packs per pass are derived caches keyed on
the parameters' version counters and storage.  Operand channel order is [x | h] (the decoder keeps x, h and the
pose head's extra features in ONE split buffer so that every consumer reads a contiguous channel range); the
filters' input channels are permuted accordingly when packed.
"""
from __future__ import annotations

import torch

from . import _lib
from . import convtc as C


def _key(mod, p):
    ps = [getattr(mod, f"conv_{g}")[p].conv for g in "zrq"]
    return tuple((t.data_ptr(), t._version) for c in ps for t in (c.weight, c.bias))


def packs(mod, p: int):
    """(zr, q) packed filters of pass p (0: (1,5), 1: (5,1))."""
    cache = mod.__dict__.setdefault("_scflow_gru_packs", {})
    key = _key(mod, p)
    hit = cache.get(p)
    if hit is not None and hit[0] == key:
        return hit[1]
    cz, cr, cq = (getattr(mod, f"conv_{g}")[p].conv for g in "zrq")
    ch, cx = mod.h_channels, mod.x_channels
    if ch % 32 or cx % 8:
        raise _lib.ScflowError(f"ConvGRU({ch},{cx}): hidden width must be a multiple of 32, x width of 8")
    with torch.no_grad():
        perm = torch.cat([torch.arange(ch, ch + cx), torch.arange(0, ch)]).to(cz.weight.device)
        w_zr = torch.cat([cz.weight, cr.weight], 0).float()[:, perm].contiguous()
        b_zr = torch.cat([cz.bias, cr.bias], 0)
        # q reads cat([r*h, x]) (raft_decoder.py:250) from where the two lie: r*h in the scratch planes, x in the
        # state buffer (two tensor maps, no concatenation): the filter keeps the reference's input-channel order
        val = (C.pack_weights(w_zr, b_zr, cz.stride, cz.padding),
               C.pack_weights(cq.weight.float().contiguous(), cq.bias, cq.stride, cq.padding))
    cache[p] = (key, val)
    return val


def packs_split(mod, p: int, cctx: int):
    """Pass p with the first `cctx` channels of x (the context features: the same in every iteration) taken out of
    the recurrent convolutions: (zr_ctx, q_ctx, zr_var, q_var).
      zr_ctx / q_ctx  the filters' context columns + the biases: run ONCE per forward (`precompute`)
      zr_var          [z | r] over [x_var | h], no bias       q_var   q over [r*h | x_var] (two tensors), no bias"""
    cache = mod.__dict__.setdefault("_scflow_gru_packs_split", {})
    key = (_key(mod, p), cctx)
    hit = cache.get(p)
    if hit is not None and hit[0] == key:
        return hit[1]
    cz, cr, cq = (getattr(mod, f"conv_{g}")[p].conv for g in "zrq")
    ch, cx = mod.h_channels, mod.x_channels
    if ch % 64 or cctx % 64 or (cx - cctx) % 8 or not 0 < cctx < cx:
        raise _lib.ScflowError(f"ConvGRU({ch},{cx}) with {cctx} context channels: widths must be multiples of 64")
    with torch.no_grad():
        dev = cz.weight.device
        w_zr = torch.cat([cz.weight, cr.weight], 0).float()
        b_zr = torch.cat([cz.bias, cr.bias], 0)
        ctx = torch.arange(ch, ch + cctx, device=dev)                         # reference order: [h | cxt | x_var]
        var = torch.arange(ch + cctx, ch + cx, device=dev)
        hh = torch.arange(0, ch, device=dev)
        wq = cq.weight.float()
        val = (C.pack_weights(w_zr[:, ctx].contiguous(), b_zr, cz.stride, cz.padding),
               C.pack_weights(wq[:, ctx].contiguous(), cq.bias, cq.stride, cq.padding),
               C.pack_weights(w_zr[:, torch.cat([var, hh])].contiguous(), None, cz.stride, cz.padding),
               C.pack_weights(wq[:, torch.cat([hh, var])].contiguous(), None, cq.stride, cq.padding))
    cache[p] = (key, val)
    return val


def precompute(mod, ctx: "C.Split"):
    """The context features' share of the six gate pre-activations (+ biases), once per forward: RAFT's context
    input does not change over the iterations, so a third of the recurrent convolutions' input channels need not be
    convolved eight times.  Returns [(P_zr, P_q, cctx)] per pass, fp32 channels-last, added in the gate epilogues."""
    out = []
    for p in (0, 1):
        zr_ctx, q_ctx, _, _ = packs_split(mod, p, ctx.c)
        out.append((C.conv(ctx, zr_ctx, want="f32"), C.conv(ctx, q_ctx, want="f32"), ctx.c))
    return out


class Scratch:
    """Per-shape work buffers of the update: z (fp32), r*h (split planes)."""

    def __init__(self, n, h, w, ch, device):
        self.z = torch.empty((n, h, w, ch), device=device, dtype=torch.float32)
        self.rh = C.alloc_split(n, h, w, ch, device)


def update_inplace(mod, xh: "C.Split", h_state: torch.Tensor, scratch: Scratch = None, pre=None) -> None:
    """Both passes, in place.  xh: split slice of cx + ch channels [x | h]; h_state (N,H,W,ch) exact fp32 state
    (its split copy is the h part of xh; both are updated).  In place is safe: the gate launch reads r*h and
    its own pixel of h only, and zr / qx of the next pass are later launches.
    `pre` = precompute(mod, xh.slice(0, cctx)): the recurrent convolutions then run over [x_var | h] / [r*h | x_var]
    only and take the context's share from `pre`."""
    n, H, W, ch = h_state.shape
    cx = xh.c - ch
    if ch != mod.h_channels or cx != mod.x_channels or not h_state.is_contiguous():
        raise _lib.ScflowError("ConvGRU: operand widths do not match the module")
    sc = scratch or Scratch(n, H, W, ch, h_state.device)
    x_part, h_part = xh.slice(0, cx), xh.slice(cx, ch)
    for p in (0, 1):
        if pre is None:
            zr, q = packs(mod, p)
            C.conv(xh, zr, out=sc.rh, epi=C.EPI_GRU_ZR, gru_h=h_state, gru_z=sc.z)
            C.conv(sc.rh, q, x2=x_part, out=h_part, epi=C.EPI_GRU_GATE, gru_h=h_state, gru_z=sc.z)
        else:
            cctx = pre[p][2]
            _, _, zr_var, q_var = packs_split(mod, p, cctx)
            C.conv(xh.slice(cctx, cx - cctx + ch), zr_var, out=sc.rh, epi=C.EPI_GRU_ZR, gru_h=h_state, gru_z=sc.z,
                   res_f32=pre[p][0])
            C.conv(sc.rh, q_var, x2=xh.slice(cctx, cx - cctx), out=h_part, epi=C.EPI_GRU_GATE, gru_h=h_state,
                   gru_z=sc.z, res_f32=pre[p][1])


def forward(mod, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """NCHW-shaped in / out (reference signature); channels-last tensors pass through without a layout copy."""
    if not h.is_cuda or h.dtype != torch.float32 or x.dtype != torch.float32:
        raise _lib.ScflowError("ConvGRU: expected CUDA fp32 tensors (scflow_b200 has no CPU path)")
    n, ch, H, W = h.shape
    cx = x.shape[1]
    h_state = h.permute(0, 2, 3, 1).contiguous().clone()
    xh = C.alloc_split(n, H, W, cx + ch, h.device)
    C.to_split(x.permute(0, 2, 3, 1), out=xh.slice(0, cx))
    C.to_split(h_state, out=xh.slice(cx, ch))
    update_inplace(mod, xh, h_state)
    return h_state.permute(0, 3, 1, 2)
