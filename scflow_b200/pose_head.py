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

    def _fc_packs(self, hs: int, ws: int):
        """fc0 / fc1 as convolution filters for the tensor-core family: flatten(NCHW) + Linear(C*hs*ws -> n0) IS the
        hs x ws convolution (no padding) with the weight viewed as (n0, C, hs, ws); fc1 a 1x1 one (its bias and
        ReLU are applied by the tail kernel).  Cached on the parameters' version counters."""
        from . import convtc as C
        fc0, fc1 = self.fc_layers[0][0], self.fc_layers[1][0]
        key = (C._key(fc0.weight, fc0.bias, fc1.weight), hs, ws)
        hit = self.__dict__.get("_scflow_fc_packs")
        if hit is None or hit[0] != key:
            c = fc0.in_features // (hs * ws)
            with torch.no_grad():
                pk0 = C.pack_weights(fc0.weight.view(fc0.out_features, c, hs, ws), fc0.bias, 1, 0)
                pk1 = C.pack_weights(fc1.weight.view(fc1.out_features, fc1.in_features, 1, 1), None, 1, 0)
            hit = (key, (pk0, pk1))
            self.__dict__["_scflow_fc_packs"] = hit
        return hit[1]

    def forward_split(self, xs, label: torch.Tensor):
        """Inference on a split-plane input (convtc.Split of cat[h | delta-flow feat | mask feat]): three
        stride-2 tensor-core convolutions, each followed by the one-launch GN+ReLU kernel (split planes out), fc0
        and fc1 on the same convolution family (narrow 32-filter tiles so that 32 SMs share the 8 MB fc0 weight),
        and the tail kernel (bias + ReLU of fc1, the selected class's 9 predictor rows)."""
        from . import convtc as C, ops
        y = xs
        for layer in self.conv_layers:
            y = layer.forward_split(y, want="split")                     # conv -> GN+ReLU -> split planes
        n, hs, ws = y.shape
        pk0, pk1 = self._fc_packs(hs, ws)
        a0 = C.conv(y, pk0, relu=True, bn=32)                            # (n, 1, 1, n0) split planes
        a1 = C.conv(a0, pk1, want="f32", bn=32)                          # (n, 1, 1, n1) fp32, pre-activation
        return ops.pose_tail(a1.view(n, -1)[:, :pk1.cout], self, label)

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
