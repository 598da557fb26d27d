"""ConvGRU forward dispatch (a3, models/decoder/raft_decoder.py:235-253).

Inference runs the fused sm_100a implicit-GEMM kernel (gru_tc) and raises if it is unavailable; there is no
other backend.  With autograd recording (training, BASELINE.json configs[4]) the six gate convolutions run as
the nn.Modules they are, so that autograd supplies dgrad / wgrad; the gate arithmetic is repo kernels."""
from __future__ import annotations

import torch


def _autograd_forward(mod, h, x):
    """raft_decoder.py:245-252 under autograd: the gate convolutions are the nn.Conv2d modules (cuDNN supplies dgrad /
    wgrad), the gate arithmetic around them is one kernel forward and one backward per gate (autograd.GruResetFn /
    GruBlendFn) instead of the element-wise chain."""
    from . import autograd as ag
    for cz, cr, cq in zip(mod.conv_z, mod.conv_r, mod.conv_q):
        hx = torch.cat([h, x], dim=1)
        a_z, a_r = cz.conv(hx), cr.conv(hx)
        a_q = cq.conv(torch.cat([ag.gru_reset(a_r, h), x], dim=1))
        h = ag.gru_blend(a_z, a_q, h)
    return h


def conv_gru_forward(mod, h, x):
    if not h.is_cuda:
        from ._lib import ScflowError
        raise ScflowError("ConvGRU: expected CUDA tensors (scflow_b200 has no CPU path)")
    if torch.is_grad_enabled() and (h.requires_grad or x.requires_grad
                                    or any(p.requires_grad for p in mod.parameters())):
        return _autograd_forward(mod, h, x)
    from . import gru_tc
    return gru_tc.forward(mod, h, x)
