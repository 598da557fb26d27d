"""ConvGRU forward dispatch (a3, models/decoder/raft_decoder.py:235-253).

`backend` is a build-stage switch, not a runtime fallback: "tcgen05" calls the fused sm_100a
implicit-GEMM pass kernel and raises if it is unavailable; "cudnn" is the library baseline the
kernel is measured against (what the reference itself runs on a GPU)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

import os

BACKEND = os.environ.get("SCFLOW_GRU_BACKEND", "tcgen05")


def _cudnn_forward(mod, h, x):
    for cz, cr, cq in zip(mod.conv_z, mod.conv_r, mod.conv_q):
        hx = torch.cat([h, x], dim=1)
        z = cz(hx)
        r = cr(hx)
        q = cq(torch.cat([r * h, x], dim=1))
        h = (1 - z) * h + z * q
    return h


def conv_gru_forward(mod, h, x):
    if torch.is_grad_enabled() and (h.requires_grad or x.requires_grad
                                    or any(p.requires_grad for p in mod.parameters())):
        # training: the six gate convolutions and their dgrad/wgrad run in cuDNN under autograd
        # (the tcgen05 pass kernel is inference-only; DESIGN.md "training path")
        if not h.is_cuda:
            from ._lib import ScflowError
            raise ScflowError("ConvGRU: expected CUDA tensors (scflow_b200 has no CPU path)")
        return _cudnn_forward(mod, h, x)
    if BACKEND == "tcgen05":
        from . import gru_tc
        return gru_tc.forward(mod, h, x)
    if not h.is_cuda:
        from ._lib import ScflowError
        raise ScflowError("ConvGRU: expected CUDA tensors (scflow_b200 has no CPU path)")
    return _cudnn_forward(mod, h, x)
