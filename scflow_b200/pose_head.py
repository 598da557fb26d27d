"""MultiClassPoseHead (models/head/pose_head.py:111-211).  Same parameter names/shapes; zero-initialised
output layers.  Inference: tensor-core stride-2 convolutions (csrc/conv_tc.cu) + GN/ReLU and FC kernels."""
from __future__ import annotations

import torch
import torch.nn as nn

from .modules import ConvModule
from .registry import register


@register
class MultiClassPoseHead(nn.Module):
    _feat_size = {"Basic": (32, 32), "Large": (64, 64)}

    def __init__(self, num_class: int, in_channels: int, net_type: str, norm_cfg: dict,
                 act_cfg: dict, feat_size: tuple = None, rotation_mode: str = "quaternion",
                 init_cfg=None):
        super().__init__()
        if rotation_mode != "ortho6d":
            raise NotImplementedError("only rotation_mode='ortho6d' (what every config ships)")
        assert net_type in ("Basic", "Large")
        self.num_class = num_class
        if feat_size is None:
            feat_size = self._feat_size[net_type]
        size = feat_size[0] * feat_size[1]
        convs = []
        for ch in (128, 128, 128):
            convs.append(ConvModule(in_channels, ch, 3, stride=2, padding=1, norm_cfg=norm_cfg, act_cfg=act_cfg))
            in_channels = ch
            size = int(size / 4)
        self.conv_layers = nn.Sequential(*convs)
        fc_in = 128 * size
        fcs = []
        for ch in (1024, 256):
            fcs.append(nn.Sequential(nn.Linear(fc_in, ch), nn.ReLU()))
            fc_in = ch
        self.flatten_op = nn.Flatten(1, -1)
        self.fc_layers = nn.Sequential(*fcs)
        self.rotation_mode = rotation_mode
        self.rotation_out_channels = 6
        self.rotation_pred = nn.Linear(fc_in, 6 * num_class)
        self.translation_pred = nn.Linear(fc_in, 3 * num_class)
        self.init_weights()

    def init_weights(self):
        nn.init.zeros_(self.translation_pred.weight)
        nn.init.zeros_(self.translation_pred.bias)
        nn.init.zeros_(self.rotation_pred.weight)
        with torch.no_grad():
            self.rotation_pred.bias.copy_(torch.tensor([1., 0., 0., 0., 1., 0.] * self.num_class))

    def _fc0_permuted(self, P: int) -> torch.Tensor:
        """fc0 weight with its columns re-ordered from flatten(NCHW) = c*P + p to the channels-last memory
        order p*C + c of the conv stack's output (so no flatten copy is needed); cached per version."""
        w = self.fc_layers[0][0].weight
        key = (w._version, w.data_ptr(), P)
        hit = self.__dict__.get("_scflow_fc0")
        if hit is None or hit[0] != key:
            C = w.shape[1] // P
            hit = (key, w.detach().view(w.shape[0], C, P).permute(0, 2, 1).reshape(w.shape[0], P * C).contiguous())
            self.__dict__["_scflow_fc0"] = hit
        return hit[1]

    def forward_split(self, xs, label: torch.Tensor):
        """Inference on a split-plane input (convtc.Split of cat[h | delta-flow feat | mask feat]): three
        stride-2 tensor-core convolutions, each followed by the one-launch GN+ReLU kernel, then the whole FC
        tail in three launches (csrc/pose_head.cu)."""
        from . import convtc as C, ops
        if self.fc_layers[1][0].out_features + self.fc_layers[0][0].out_features > 12288:
            raise NotImplementedError("pose head FC widths beyond the split-K kernel's workspace")
        y = xs
        for i, layer in enumerate(self.conv_layers):
            # GN+ReLU applied; split planes between the convolutions, fp32 channels-last into the FC tail
            y = layer.forward_split(y, want="split") if i + 1 < len(self.conv_layers) else layer.forward_split(y)
        n, hs, ws, c = y.shape
        return ops.pose_fc(y.reshape(n, hs * ws * c), self._fc0_permuted(hs * ws), self, label)

    def forward(self, x: torch.Tensor, label: torch.Tensor):
        if not (torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters()))):
            if not x.is_cuda or x.dtype != torch.float32:
                from ._lib import ScflowError
                raise ScflowError("MultiClassPoseHead: expected a CUDA fp32 tensor (scflow_b200 has no CPU path)")
            from . import convtc as C
            return self.forward_split(C.to_split(x.permute(0, 2, 3, 1)), label)
        # training (configs[4]): nn.Modules under autograd
        x = self.fc_layers(self.flatten_op(self.conv_layers(x)))
        dt = self.translation_pred(x).view(-1, self.num_class, 3)
        dr = self.rotation_pred(x).view(-1, self.num_class, self.rotation_out_channels)
        # pose_head.py:209-210 -- index_select(dim=1, index=label)[:, 0, :]: every crop reads the
        # class of label[0].  Reproduced as is (results must equal the reference's).
        dt = torch.index_select(dt, dim=1, index=label)[:, 0, :]
        dr = torch.index_select(dr, dim=1, index=label)[:, 0, :]
        return dr, dt
