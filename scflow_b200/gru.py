"""ConvGRU forward dispatch (a3, models/decoder/raft_decoder.py:235-253).

Inference runs the fused sm_100a implicit-GEMM kernel (gru_tc) and raises if it is unavailable; there is no
other backend.  With autograd recording (training, BASELINE.json configs[4]) the six gate convolutions run as
the nn.Modules they are, so that autograd supplies dgrad / wgrad."""
from __future__ import annotations

import torch


def _autograd_forward(mod, h, x):
    for cz, cr, cq in zip(mod.conv_z, mod.conv_r, mod.conv_q):
        hx = torch.cat([h, x], dim=1)
        z = torch.sigmoid(cz.conv(hx))
        r = torch.sigmoid(cr.conv(hx))
        q = torch.tanh(cq.conv(torch.cat([r * h, x], dim=1)))
        h = (1 - z) * h + z * q
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
