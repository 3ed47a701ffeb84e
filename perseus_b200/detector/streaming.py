"""Streaming (batch 1-2) latency path: camera frame in pinned host memory -> keypoints in pinned host memory,
as ONE CUDA-graph replay (H2D copy, fused ingest, backbone, head + decode, D2H copy).

Mirrors the hot loop of the reference's scripts/streaming.py:121-131 (get_frame -> model(x[None]) ->
denormalize_pixel_coordinates), minus the camera SDK and the cv2 window (out of scope).
The reference runs that loop on the CPU at ~15 ms per frame; here the per-frame host work is one memcpy into
the pinned staging buffer and one graph launch.
"""
from __future__ import annotations

import numpy as np
import torch

from .models import KeypointCNN


class StreamingDetector:
    """Fixed-shape, graph-captured `forward_frames` for B frames of (src_h, src_w, channels) uint8
    (+ float32 depth).  ZED VGA frames are 376 x 672 BGRA (streaming.py:35,68)."""

    def __init__(self, model: KeypointCNN, src_h: int = 376, src_w: int = 672, channels: int = 4, batch: int = 1,
                 bgr: bool = True, depth_divisor: float = 0.035, device: torch.device | None = None):
        if model.training:
            raise RuntimeError("StreamingDetector needs model.eval()")
        self.model = model
        self.device = device or next(model.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("StreamingDetector runs only on a B200 (no CPU fallback): move the model to cuda")
        self.B, self.bgr, self.depth_divisor = batch, bgr, depth_divisor
        self.use_depth = model.num_channels == 4
        self.host_rgb = torch.empty((batch, src_h, src_w, channels), dtype=torch.uint8).pin_memory()
        self.host_depth = torch.empty((batch, src_h, src_w), dtype=torch.float32).pin_memory() if self.use_depth else None
        self.host_px = torch.empty((batch, model.n_keypoints, 2), dtype=torch.float32).pin_memory()
        self.dev_rgb = torch.empty_like(self.host_rgb, device=self.device)
        self.dev_depth = torch.empty_like(self.host_depth, device=self.device) if self.use_depth else None
        # the detector's OWN plan: the captured graph bakes in its arena / workspace pointers, so it must not be the
        # model's shared plan (which is freed and re-created when a later forward() needs a larger batch)
        self.stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.stream(self.stream):
            self.plan = model.private_plan(self.device, batch)
        self.frames = torch.empty(self.plan.frames_bytes(batch), dtype=torch.uint8, device=self.device)
        self.graph = None
        self._capture()

    def _step(self):
        self.dev_rgb.copy_(self.host_rgb, non_blocking=True)
        if self.use_depth:
            self.dev_depth.copy_(self.host_depth, non_blocking=True)
        self.model.ingest(self.dev_rgb, self.dev_depth, bgr=self.bgr, depth_divisor=self.depth_divisor, out=self.frames,
                          plan=self.plan)
        _, px = self.model.forward_ingested(self.frames, self.B, plan=self.plan)
        self.host_px.copy_(px, non_blocking=True)

    def _capture(self):
        self.host_rgb.zero_()
        if self.use_depth:
            self.host_depth.fill_(0.3)
        with torch.cuda.stream(self.stream):
            for _ in range(3):  # warm-up: builds the plan's TMA descriptors for these pointers before capture
                self._step()
        self.stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self.stream):
            self._step()
        self.graph = g

    def infer_pinned(self) -> torch.Tensor:
        """Frames already written into `host_rgb` / `host_depth`: replay and wait. Returns `host_px` (B, K, 2)."""
        with torch.cuda.stream(self.stream):  # CUDAGraph.replay() launches on the CURRENT stream
            # load_state_dict / an optimizer step since the last frame: repack into the SAME arena, on this stream
            # (ordered before the replay; the graph's pointers stay valid)
            self.model.sync_plan_weights(self.plan)
            self.graph.replay()
        self.stream.synchronize()
        return self.host_px

    def infer(self, frame_u8: np.ndarray, depth_m: np.ndarray | None = None) -> np.ndarray:
        """frame_u8: (B, h, w, ch) or (h, w, ch) uint8 as the camera delivers it; depth_m: metres (NaN/Inf allowed)."""
        f = torch.from_numpy(np.ascontiguousarray(frame_u8)).view(self.host_rgb.shape)
        self.host_rgb.copy_(f)
        if self.use_depth:
            if depth_m is None:
                raise ValueError("an RGBD model needs a depth frame")
            self.host_depth.copy_(torch.from_numpy(np.ascontiguousarray(depth_m, dtype=np.float32)).view(self.host_depth.shape))
        return self.infer_pinned().numpy().copy()
