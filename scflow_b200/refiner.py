"""Inference-side host logic around the decoder: the feature extraction and `get_pose` of
SCFlowRefiner (models/refiner/scflow_refiner.py:84-138) and batch-sharded multi-GPU refinement
with one all-gather of the refined poses.  Renderer, data formatting, losses and
the mmengine Runner are out of scope and stay in the reference."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import os

import torch
import torch.distributed as dist
import torch.nn as nn

from .decoder import SCFlowDecoder
from .encoder import RAFTEncoder

DECODER_CFG = dict(
    net_type="Basic", num_levels=4, radius=4, iters=8,
    detach_flow=True, detach_mask=True, detach_pose=True, detach_depth_for_xy=True,
    mask_flow=False, mask_corr=False,
    pose_head_cfg=dict(type="MultiClassPoseHead", num_class=21, in_channels=224, net_type="Basic",
                       rotation_mode="ortho6d", norm_cfg=dict(type="GN", num_groups=32, requires_grad=True),
                       act_cfg=dict(type="ReLU")),
    corr_lookup_cfg=dict(align_corners=True), gru_type="SeqConv", act_cfg=dict(type="ReLU"))


def decoder_cfg(feat_size=None, **over) -> dict:
    import copy
    cfg = copy.deepcopy(DECODER_CFG)
    if feat_size is not None:
        cfg["pose_head_cfg"]["feat_size"] = tuple(feat_size)
    cfg.update(over)
    return cfg


class SCFlowRefinerCore(nn.Module):
    """encoder (shared IN), context encoder (BN), decoder -- the part of SCFlowRefiner that
    `get_pose` touches (scflow_ycbv_real.py:172-230)."""

    def __init__(self, image_size: int = 256, h_channels: int = 128, cxt_channels: int = 128,
                 decoder: Optional[dict] = None):
        super().__init__()
        self.encoder = RAFTEncoder(3, 256, net_type="Basic", norm_cfg=dict(type="IN"))
        self.context = RAFTEncoder(3, 256, net_type="Basic", norm_cfg=dict(type="BN"))
        hs = image_size // 8
        self.decoder = SCFlowDecoder(**(decoder or decoder_cfg(feat_size=(hs, hs))))
        self.h_channels, self.cxt_channels = h_channels, cxt_channels
        assert self.h_channels == self.decoder.h_channels
        assert self.cxt_channels == self.decoder.cxt_channels

    def extract_feat(self, render_images, real_images):
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            real_feat = self.encoder(real_images)                # training: nn.Module layers under autograd
            render_feat = self.encoder(render_images)
            cxt = self.context(render_images)
        else:
            # inference: tensor-core encoders; the rendered batch's space-to-depth stem input is computed once
            # and read by both the feature and the context encoder (scflow_refiner.py:84-106); the features
            # come back planar and TF32-rounded, the correlation kernel's operand format
            from .encoder import StemInput
            real_feat = self.encoder(real_images, planar_tf32=True)
            stem = StemInput(render_images)
            render_feat = self.encoder(render_images, stem=stem, planar_tf32=True)
            cxt = self.context(render_images, stem=stem)
        h_feat, cxt_feat = torch.split(cxt, [self.h_channels, self.cxt_channels], dim=1)
        return render_feat, real_feat, torch.tanh(h_feat), torch.relu(cxt_feat)

    def get_pose(self, render_images, real_images, ref_rotation, ref_translation, depth,
                 internel_k, label, init_flow=None):
        # scflow_refiner.py:129-138; called under torch.no_grad() by the test loop (inference kernels)
        # and with autograd on by `loss` (training path: scflow_b200.autograd)
        feat_render, feat_real, h_feat, cxt_feat = self.extract_feat(render_images, real_images)
        if init_flow is None:
            N, _, H, W = real_images.shape
            init_flow = feat_render.new_zeros((N, 2, H, W), dtype=torch.float32)
        return self.decoder(feat_render, feat_real, h_feat, cxt_feat, ref_rotation, ref_translation,
                            depth, internel_k, init_flow=init_flow, label=label, invalid_flow_num=0.)

    @torch.no_grad()
    def refine(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        """(N,12) packed [R row-major | t] after `decoder.iters` iterations."""
        out = self.get_pose(batch["rendered_images"], batch["real_images"], batch["R0"], batch["t0"],
                            batch["depth"], batch["K"], batch["labels"])
        return pack_pose(out[2][-1], out[3][-1])


class GraphedRefiner:
    """`core.refine` for one fixed batch signature, captured once into a CUDA graph and replayed.

    The refinement step is ~700 small launches (cuDNN convolutions plus our kernels); replaying a
    graph removes the per-launch CPU cost, which otherwise bounds the step once the kernels are
    fast.  Everything on the path is capture-safe: no host synchronisation, static shapes, all
    buffers from the graph's private pool (TMA descriptors stay valid across replays), weight
    packs / folded BN / cuDNN algorithm choices cached by the warm-up runs.  The pose all-gather
    stays outside the graph.  Re-capture (build a new object) after changing weights' storage."""

    def __init__(self, core: "SCFlowRefinerCore", example: Dict[str, torch.Tensor], warmup: int = 3,
                 own_inputs: bool = True):
        from . import ops
        if ops.STATS.timing:
            raise RuntimeError("turn ops.STATS timing off before capturing a graph")
        self.core = core
        dev = next(core.parameters()).device
        self.inputs = {k: (v.to(dev).clone() if own_inputs else v) for k, v in example.items()}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                core.refine(self.inputs)
        torch.cuda.current_stream(dev).wait_stream(side)
        launches_before = ops.STATS.count
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.poses = core.refine(self.inputs)
        self.launches_per_replay = ops.STATS.count - launches_before    # OUR kernels in the graph

    @torch.no_grad()
    def __call__(self, batch: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
        """Refine `batch` (device or pinned-host tensors, same shapes as the example); returns the
        (N,12) packed poses (a static buffer: clone it to keep it across calls)."""
        if batch is not None:
            for k, v in batch.items():
                self.inputs[k].copy_(v, non_blocking=True)
        self.graph.replay()
        from . import ops
        ops.STATS.count += self.launches_per_replay
        return self.poses


def pack_pose(R: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    return torch.cat([R.reshape(-1, 9), t.reshape(-1, 3)], dim=1).contiguous()


def unpack_pose(p: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    return p[:, :9].reshape(-1, 3, 3), p[:, 9:12]


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous batch split: rank k takes crops [lo, hi); remainders go to the first ranks."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_poses(local: torch.Tensor, total: int) -> torch.Tensor:
    """All-gather of the (n_local, 12) refined poses in rank order (= original crop order).
    The only collective on the path: 48 B per crop, enqueued on the current stream.
    Uneven shards are padded to the largest shard."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    sizes = [shard_range(total, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    buf = local.new_zeros((mx, local.shape[1]))
    buf[: local.shape[0]] = local
    out = local.new_empty((world * mx, local.shape[1]))
    dist.all_gather_into_tensor(out, buf)
    return torch.cat([out[r * mx: r * mx + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], dim=0)


class PoseGatherer:
    """The pose all-gather of a serving loop, off the compute stream.  `gather_poses` on the compute stream makes
    every rank wait for the slowest one EVERY step (measured at N = 8: 0.98 ms of a 17.5 ms step, efficiency
    0.969); here step i's poses are staged (one small copy on the compute stream) and gathered on a side stream
    while step i + 1 computes, two slots deep, so the ranks only meet when a result is read.  `submit` returns the
    slot; `result(slot)` makes the current stream wait for that gather and returns the (total, 12) poses in crop
    order; `to_host(slot, pinned)` enqueues the device-to-host read behind the gather on the side stream and
    returns an event to synchronise on.  On CPU tensors (gloo tests) everything is synchronous."""

    def __init__(self, total: int, device, width: int = 12, slots: int = 2):
        self.total, self.width = total, width
        self.world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.sizes = [shard_range(total, r, self.world) for r in range(self.world)]
        self.mx = max(hi - lo for lo, hi in self.sizes)
        self.even = all(hi - lo == self.mx for lo, hi in self.sizes)
        self.cuda = torch.device(device).type == "cuda"
        self.stage = [torch.zeros((self.mx, width), device=device) for _ in range(slots)]
        self.out = [torch.empty((self.world * self.mx, width), device=device) for _ in range(slots)]
        self.comm = torch.cuda.Stream(device=device) if self.cuda else None
        self.done = [torch.cuda.Event() if self.cuda else None for _ in range(slots)]
        self.used = [False] * slots
        self.i = 0
        self.events = None            # set to a list to collect (start, end) timing events of every gather

    def submit(self, local: torch.Tensor) -> int:
        slot = self.i % len(self.stage)
        self.i += 1
        n = local.shape[0]
        if self.cuda:
            cur = torch.cuda.current_stream(local.device)
            if self.used[slot]:
                cur.wait_event(self.done[slot])          # the gather that last read this staging slot (long done)
            self.stage[slot][:n].copy_(local)
            ready = torch.cuda.Event()
            ready.record(cur)
            with torch.cuda.stream(self.comm):
                self.comm.wait_event(ready)
                if self.events is not None:
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(self.comm)
                self._gather(slot)
                if self.events is not None:
                    b.record(self.comm)
                    self.events.append((a, b))
                self.done[slot].record(self.comm)
        else:
            self.stage[slot][:n].copy_(local)
            self._gather(slot)
        self.used[slot] = True
        return slot

    def _gather(self, slot: int) -> None:
        if self.world == 1:
            self.out[slot].copy_(self.stage[slot])
        else:
            dist.all_gather_into_tensor(self.out[slot], self.stage[slot])

    def result(self, slot: int) -> torch.Tensor:
        if self.cuda:
            torch.cuda.current_stream(self.out[slot].device).wait_event(self.done[slot])
        o = self.out[slot]
        if self.even:
            return o[: self.total]
        return torch.cat([o[r * self.mx: r * self.mx + (hi - lo)] for r, (lo, hi) in enumerate(self.sizes)], dim=0)

    def to_host(self, slot: int, pinned: torch.Tensor):
        """Device-to-host read of the gathered (padded) poses of `slot` behind its gather; returns the event (or
        None on CPU) after which `pinned` holds them."""
        if not self.cuda:
            pinned.copy_(self.out[slot])
            return None
        with torch.cuda.stream(self.comm):
            pinned.copy_(self.out[slot], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.comm)
        self.done[slot].record(self.comm)                # the slot is free again only after the read
        return ev


@torch.no_grad()
def refine_sharded(core: SCFlowRefinerCore, batch: Dict[str, torch.Tensor], micro_batch: int = 64):
    """Refine this rank's shard of `batch` (already sliced by the caller with shard_range) in
    micro-batches and return this rank's (n_local, 12) poses."""
    n = batch["depth"].shape[0]
    outs = []
    for lo in range(0, n, micro_batch):
        mb = {k: v[lo: lo + micro_batch] for k, v in batch.items()}
        outs.append(core.refine(mb))
    return torch.cat(outs, dim=0) if outs else batch["depth"].new_zeros((0, 12))
