"""Batched inference driver: the validate loop of the reference
(``run/validate.py:99-150``, ``lib/core/function.py:126-183``) without its per-batch stalls.

``StreamedInference`` feeds host batches (pinned memory) to the model with the host->device copy of
batch i+1 overlapping the forward of batch i (two device staging buffers, one copy stream, events for
the hand-over) and returns the poses in pinned host memory.  Every batch is copied in and its result
copied out -- nothing is cached across batches.

A batch is either the heatmaps themselves or the 2-D poses they are made from --
the reference's 'pred' / 'gt' heatmap sources (``lib/dataset/JointsDataset.py:207-258``): then only
8*N*J bytes per view cross PCIe and ``ops.splat_heatmaps`` renders the maps in HBM.
"""
from __future__ import annotations

from typing import Dict, Iterable, Iterator, Optional

import torch

__all__ = ["StreamedInference", "GraphedInference"]


class StreamedInference:
    def __init__(self, model, cameras: Dict, resize_transform: torch.Tensor, device, depth: int = 2,
                 splat: Optional[Dict] = None, graph: bool = False, lanes: int = 1):
        """``splat``: ``dict(sigma=, image_size=(w,h), heatmap_size=(w,h))`` -- required when batches are
        2-D poses (cfg.NETWORK.SIGMA / IMAGE_SIZE / HEATMAP_SIZE of the reference).
        ``graph``: replay the forward of every staging slot as a CUDA graph (captured on first use, see
        ``GraphedInference``); needs fixed batch shapes, heatmap batches, one sequence (``metas=None``) and the
        same ``proposals`` tensor on every call.
        ``lanes``: consecutive batches run on ``lanes`` compute streams in rotation (at most ``depth - 1``), so the
        forward of batch i+1 may start before that of batch i has finished.  Batches are independent, and a forward
        is a long tail of small launches (CenterNet on one plane per frame, NMS, the 1-D net) next to a few
        device-filling ones (the per-person projection, P2PNet): with two lanes the small kernels of one batch run in
        the shadow of the large kernels of its neighbour.  Results are identical; they are still yielded in order."""
        self.splat = splat
        self.graph = bool(graph)
        self._graphs = {}
        self.model = model.eval()
        self.cameras = cameras
        self.device = torch.device(device)
        self.rt = resize_transform.to(self.device)
        self.depth = int(depth)
        self.copy_stream = torch.cuda.Stream(self.device)
        self.lanes = max(1, min(int(lanes), self.depth - 1))
        self._lane_streams = [torch.cuda.Stream(self.device) for _ in range(self.lanes)] if self.lanes > 1 else None
        self._stage = None          # device staging buffers, allocated on the first batch
        self._copied = [torch.cuda.Event() for _ in range(self.depth)]
        self._consumed = [torch.cuda.Event() for _ in range(self.depth)]
        self._host_out = None
        self._pose_stage = None

    def _buffers(self, like: torch.Tensor):
        if self._stage is None or self._stage[0].shape != like.shape:
            self._stage = [torch.empty(like.shape, dtype=torch.float32, device=self.device) for _ in range(self.depth)]
        return self._stage

    def _upload(self, slot: int, host):
        if isinstance(host, dict):
            return self._upload_poses(slot, host)
        buf = self._buffers(host)[slot]
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._consumed[slot])     # the forward that read this slot is done
            buf.copy_(host, non_blocking=True)
            self._copied[slot].record(self.copy_stream)
        return buf

    def _upload_poses(self, slot: int, host: Dict):
        """host: {"joints": [B,V,N,J,2] fp32, "vis": [B,V,N,J] uint8 | None, "count": [B,V] int32 | None}."""
        if self.splat is None:
            raise ValueError("pose batches need StreamedInference(..., splat=dict(sigma=, image_size=, heatmap_size=))")
        if self._pose_stage is None:
            self._pose_stage = [dict() for _ in range(self.depth)]
        stage = self._pose_stage[slot]
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._consumed[slot])
            dev = {}
            for key in ("joints", "vis", "count"):
                t = host.get(key)
                if t is None:
                    dev[key] = None
                    continue
                buf = stage.get(key)
                if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
                    buf = stage[key] = torch.empty(t.shape, dtype=t.dtype, device=self.device)
                buf.copy_(t, non_blocking=True)
                dev[key] = buf
            self._copied[slot].record(self.copy_stream)
        return dev

    def _replay(self, slot: int, cur: torch.Tensor, meta: Dict, proposals):
        """Forward of staging slot ``slot`` as a CUDA graph: captured (after two eager warm-up calls) the first time
        the slot is used with this shape, replayed afterwards.  The graph reads the slot's staging buffer in place."""
        key = (slot, tuple(cur.shape), None if proposals is None else (proposals.data_ptr(), tuple(proposals.shape)))
        entry = self._graphs.get(key)
        if entry is None:
            joint_net = getattr(self.model, "joint_net", None)
            was_static = getattr(joint_net, "static_shapes", None)
            if joint_net is not None:
                joint_net.static_shapes = True

            def fwd():
                return self.model(meta=meta, input_heatmaps=cur, cameras=self.cameras, resize_transform=self.rt,
                                  proposals=proposals)[0]
            main = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(self.device)
            side.wait_stream(main)
            with torch.cuda.stream(side), torch.no_grad():
                for _ in range(2):
                    fwd()
            main.wait_stream(side)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side), torch.no_grad():
                out = fwd()
            main.wait_stream(side)
            if joint_net is not None:
                joint_net.static_shapes = was_static
            entry = self._graphs[key] = (g, out)
        entry[0].replay()
        return entry[1]

    def _heatmaps(self, cur):
        if not isinstance(cur, dict):
            return cur
        from . import ops
        return ops.splat_heatmaps(cur["joints"], cur["vis"], cur["count"], self.splat["sigma"],
                                  self.splat["image_size"], self.splat["heatmap_size"])

    def run(self, batches: Iterable, metas: Optional[Iterable] = None, proposals=None) -> Iterator[torch.Tensor]:
        """``batches``: iterable of pinned host heatmaps [B,V,J,H,W] -- or of 2-D pose dicts
        ``{"joints": [B,V,N,J,2] fp32, "vis": [B,V,N,J] uint8 | None, "count": [B,V] int32 | None}`` when the
        driver was built with ``splat=`` -- yields fused poses [B,P,J,5] on the
        host (pinned), one per batch, in order.  ``metas``: matching iterable of meta dicts (default: every
        frame in the only sequence of ``cameras``)."""
        main = torch.cuda.current_stream(self.device)
        it = iter(batches)
        mit = iter(metas) if metas is not None else None
        seq = next(iter(self.cameras.keys()))
        try:
            nxt = next(it)
        except StopIteration:
            return
        for ev in self._consumed:
            ev.record(main)
        if self._lane_streams is not None:
            for ls in self._lane_streams:
                ls.wait_stream(main)        # whatever the caller queued before run() (weights, proposals) comes first
        pending = self._upload(0, nxt)
        i = 0
        prev_out = None
        while pending is not None:
            slot = i % self.depth
            cur, cur_host = pending, nxt
            try:
                nxt = next(it)
                pending = self._upload((i + 1) % self.depth, nxt)   # overlaps the forward below
            except StopIteration:
                pending = None
            nframes = (cur_host["joints"] if isinstance(cur_host, dict) else cur_host).shape[0]
            meta = next(mit) if mit is not None else {"seq": [seq] * nframes}
            cs = main if self._lane_streams is None else self._lane_streams[i % self.lanes]
            with torch.cuda.stream(cs):
                cs.wait_event(self._copied[slot])
                if self.graph and not isinstance(cur, dict):
                    if mit is not None:
                        raise ValueError("graph=True replays one captured forward: per-batch metas are not supported")
                    out = self._replay(slot, cur, meta, proposals)
                else:
                    with torch.no_grad():
                        out = self.model(meta=meta, input_heatmaps=self._heatmaps(cur), cameras=self.cameras,
                                         resize_transform=self.rt, proposals=proposals)[0]
                self._consumed[slot].record(cs)
                if self._host_out is None or self._host_out[0].shape != out.shape:
                    self._host_out = [torch.empty(out.shape, dtype=out.dtype).pin_memory() for _ in range(self.depth)]
                host_out = self._host_out[slot]
                host_out.copy_(out, non_blocking=True)
                done = torch.cuda.Event()
                done.record(cs)
            if prev_out is not None:
                prev_out[1].synchronize()
                yield prev_out[0]
            prev_out = (host_out, done)
            i += 1
        if prev_out is not None:
            prev_out[1].synchronize()
            yield prev_out[0]
        if self._lane_streams is not None:
            for ls in self._lane_streams:
                main.wait_stream(ls)


class GraphedInference:
    """Latency mode: the whole HDN -> JLN forward of one fixed batch shape captured
    into a CUDA graph.  At small batches the forward is ~200 launches of a few microseconds each and the
    host cannot issue them as fast as the GPU retires them; a replay issues them all at once.

    The forward is capturable because it has no host synchronisation (per-person windows, masks and
    scatter all stay on the device) and every per-call constant (rig tables, frame->rig indices, folded
    conv weights, geometry) is cached after the warm-up calls.  ``seqs`` fixes which rig each frame of
    the batch uses; a different assignment needs another instance."""

    def __init__(self, model, cameras: Dict, resize_transform: torch.Tensor, device, example: torch.Tensor,
                 seqs: Optional[list] = None, proposals: Optional[torch.Tensor] = None, warmup: int = 3):
        self.model = model.eval()
        self.device = torch.device(device)
        self.cameras = cameras
        self.rt = resize_transform.to(self.device)
        B = int(example.shape[0])
        self.meta = {"seq": list(seqs) if seqs is not None else [next(iter(cameras.keys()))] * B}
        self.static_in = torch.empty(example.shape, dtype=torch.float32, device=self.device)
        self.static_in.copy_(example)
        self.static_props = None if proposals is None else proposals.to(self.device, torch.float32).clone()
        joint_net = getattr(self.model, "joint_net", None)
        was_static = getattr(joint_net, "static_shapes", None)
        if joint_net is not None:
            joint_net.static_shapes = True                  # fixed shapes, no host sync: see JointLocalizationNet
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):                 # cuDNN autotuning, weight folding, rig/geometry caches
                self._forward()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.static_out = self._forward()
        if joint_net is not None:
            joint_net.static_shapes = was_static
        self._host_out = torch.empty(self.static_out.shape, dtype=self.static_out.dtype).pin_memory()

    def _forward(self):
        return self.model(meta=self.meta, input_heatmaps=self.static_in, cameras=self.cameras, resize_transform=self.rt,
                          proposals=self.static_props)[0]

    def __call__(self, heatmaps: torch.Tensor, to_host: bool = True) -> torch.Tensor:
        """heatmaps: [B,V,J,H,W] host (pinned) or device tensor of the captured shape.  Returns the fused
        poses [B,P,J,5]: a pinned host tensor (synchronised) or, with ``to_host=False``, the graph's static
        device output (valid until the next call)."""
        if tuple(heatmaps.shape) != tuple(self.static_in.shape):
            raise ValueError(f"captured for {tuple(self.static_in.shape)}, got {tuple(heatmaps.shape)}")
        self.static_in.copy_(heatmaps, non_blocking=True)
        self.graph.replay()
        if not to_host:
            return self.static_out
        self._host_out.copy_(self.static_out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._host_out
