"""Batched end-to-end inference from HOST frames: pinned uint8 batch in -> pinned pixel keypoints out.

The reference's batched callers (perseus/detector/validate.py:109-153: DataLoader batch -> `.to(device)` -> model ->
denormalize -> `.cpu()`) move every batch host -> device -> host synchronously around the forward.  `BatchPipeline`
is the same loop as a software pipeline, so the copy of batch i+1 and the read-back of batch i-1 run under the compute
of batch i:

  copy stream(s)  : one plain cudaMemcpyAsync per submitted batch (per copy stream) into one of `slots` device buffers
  compute stream  : ONE CUDA-graph replay per batch (fused ingest -> backbone -> head + pixel decode; ~20 kernels whose
                    launches would otherwise cost ~0.1 ms of host time per step) + the 16 KB keypoint copy into a ring
                    of pinned result buffers
  host            : `submit()` never blocks on the GPU; `result(ticket)` waits for that batch's keypoints only.

The graphs bake in the pointers of a PRIVATE plan (KeypointCNN.private_plan), so later forward() calls on the module
cannot invalidate them; weight changes are picked up before every replay.
"""
from __future__ import annotations

import torch

from .models import KeypointCNN


class BatchPipeline:
    def __init__(self, model: KeypointCNN, batch: int, src_h: int | None = None, src_w: int | None = None,
                 channels: int = 3, *, bgr: bool = False, depth_divisor: float = 1.0, slots: int = 2, results: int = 4,
                 copy_streams: int = 1, device: torch.device | None = None):
        if model.training:
            raise RuntimeError("BatchPipeline needs model.eval()")
        self.model = model
        self.device = device or next(model.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("BatchPipeline runs only on a B200 (no CPU fallback): move the model to cuda")
        if channels not in (3, 4):
            raise ValueError("channels must be 3 (RGB/BGR) or 4 (RGBA/BGRA)")
        self.B = batch
        self.src_h, self.src_w = src_h or model.H, src_w or model.W
        self.channels, self.bgr, self.depth_divisor = channels, bgr, depth_divisor
        self.use_depth = model.num_channels == 4
        self.slots, self.results = slots, results
        dev = self.device
        self.compute = torch.cuda.Stream(device=dev)
        self.copies = [torch.cuda.Stream(device=dev) for _ in range(max(1, copy_streams))]
        with torch.cuda.stream(self.compute):
            self.plan = model.private_plan(dev, batch)
        shape = (batch, self.src_h, self.src_w, channels)
        self.dev_rgb = [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(slots)]
        self.dev_depth = ([torch.empty(shape[:3], dtype=torch.float32, device=dev) for _ in range(slots)]
                          if self.use_depth else [None] * slots)
        self.frames = torch.empty(self.plan.frames_bytes(batch), dtype=torch.uint8, device=dev)
        self.px_dev = [torch.empty((batch, model.n_keypoints, 2), dtype=torch.float32, device=dev) for _ in range(slots)]
        self.norm_dev = [torch.empty((batch, 2 * model.n_keypoints), dtype=torch.float32, device=dev) for _ in range(slots)]
        self.host_px = [torch.empty((batch, model.n_keypoints, 2), dtype=torch.float32).pin_memory() for _ in range(results)]
        self.ev_copied = [[torch.cuda.Event() for _ in self.copies] for _ in range(slots)]
        self.ev_free = [torch.cuda.Event() for _ in range(slots)]
        self.ev_done = [torch.cuda.Event() for _ in range(results)]
        self.graphs = [None] * slots
        self.ticket = 0
        self._capture()

    def _run(self, s: int):
        self.model.ingest(self.dev_rgb[s], self.dev_depth[s], bgr=self.bgr, depth_divisor=self.depth_divisor,
                          out=self.frames, plan=self.plan)
        return self.model.forward_ingested(self.frames, self.B, plan=self.plan, out_norm=self.norm_dev[s],
                                           out_px=self.px_dev[s])[1]

    def _capture(self):
        with torch.cuda.stream(self.compute):
            for s in range(self.slots):
                self.dev_rgb[s].zero_()
                if self.use_depth:
                    self.dev_depth[s].fill_(0.3)
            for _ in range(2):  # warm-up: builds the plan's TMA descriptors for these pointers before capture
                self._run(0)
        self.compute.synchronize()
        for s in range(self.slots):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.compute):
                self._run(s)
            self.graphs[s] = g
        for e in self.ev_free:
            e.record(self.compute)
        self.compute.synchronize()

    def submit(self, host_rgb: torch.Tensor, host_depth: torch.Tensor | None = None) -> int:
        """Enqueues one batch (pinned host tensors for an asynchronous copy; pageable memory works but serialises).
        Returns a ticket for `result`.  At most `results` tickets may be outstanding."""
        if tuple(host_rgb.shape) != tuple(self.dev_rgb[0].shape) or host_rgb.dtype != torch.uint8:
            raise ValueError(f"submit: expected uint8 {tuple(self.dev_rgb[0].shape)}, got {host_rgb.dtype} {tuple(host_rgb.shape)}")
        if self.use_depth and host_depth is None:
            raise ValueError("submit: an RGBD model needs a depth batch")
        t = self.ticket
        s, r = t % self.slots, t % self.results
        n = len(self.copies)
        rows = (self.B + n - 1) // n
        for c, cs in enumerate(self.copies):
            lo, hi = c * rows, min(self.B, (c + 1) * rows)
            with torch.cuda.stream(cs):
                cs.wait_event(self.ev_free[s])  # the previous batch in this slot has been ingested
                if hi > lo:
                    self.dev_rgb[s][lo:hi].copy_(host_rgb[lo:hi], non_blocking=True)
                    if self.use_depth:
                        self.dev_depth[s][lo:hi].copy_(host_depth[lo:hi], non_blocking=True)
                self.ev_copied[s][c].record(cs)
        with torch.cuda.stream(self.compute):
            for e in self.ev_copied[s]:
                self.compute.wait_event(e)
            self.model.sync_plan_weights(self.plan)
            self.graphs[s].replay()
            self.ev_free[s].record(self.compute)
            self.host_px[r].copy_(self.px_dev[s], non_blocking=True)
            self.ev_done[r].record(self.compute)
        self.ticket = t + 1
        return t

    def result(self, ticket: int) -> torch.Tensor:
        """Blocks until the keypoints of `ticket` are in pinned host memory; the returned (B, K, 2) float32 tensor is
        valid until `results` further batches have been submitted."""
        if not (self.ticket - self.results <= ticket < self.ticket):
            raise ValueError(f"result: ticket {ticket} is not outstanding (next ticket {self.ticket}, ring {self.results})")
        self.ev_done[ticket % self.results].synchronize()
        return self.host_px[ticket % self.results]

    def copy_only(self, host_rgb: torch.Tensor) -> None:
        """The pipeline's host -> device traffic alone (same buffers and copy streams, no compute): its ceiling."""
        s = self.ticket % self.slots
        n = len(self.copies)
        rows = (self.B + n - 1) // n
        for c, cs in enumerate(self.copies):
            lo, hi = c * rows, min(self.B, (c + 1) * rows)
            with torch.cuda.stream(cs):
                if hi > lo:
                    self.dev_rgb[s][lo:hi].copy_(host_rgb[lo:hi], non_blocking=True)
                self.ev_copied[s][c].record(cs)
        for e in self.ev_copied[s]:
            self.compute.wait_event(e)
        self.ticket += 1

    def drain(self) -> None:
        self.compute.synchronize()
        for cs in self.copies:
            cs.synchronize()
