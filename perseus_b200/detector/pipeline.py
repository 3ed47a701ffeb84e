"""Batched end-to-end inference from HOST frames: pinned uint8 batch in -> pinned pixel keypoints out.

The reference's batched callers (perseus/detector/validate.py:109-153: DataLoader batch -> `.to(device)` -> model ->
denormalize -> `.cpu()`) move every batch host -> device -> host synchronously around the forward.  `BatchPipeline`
is the same loop as a software pipeline, so the copy of batch i+1 and the read-back of batch i-1 run under the compute
of batch i:

  copy stream(s)  : one plain cudaMemcpyAsync per submitted batch (per copy stream) into one of `slots` device buffers
  ingest stream   : the fused ingest kernel of batch i+1 (HBM-bound) runs under the tensor-bound convolutions of batch i,
                    mostly in the gaps between their launches
  compute stream  : ONE CUDA-graph replay per batch (backbone -> head + pixel decode; ~11 kernels whose launches would
                    otherwise cost ~0.1 ms of host time per step) + the 16 KB keypoint copy into a ring of pinned
                    result buffers
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
                 copy_streams: int = 1, device: torch.device | None = None, post=None, graph: bool | None = None):
        """post: optional callable(px (B, K, 2) float32 device tensor) -> device tensor with a FIXED address, captured into
        every batch's CUDA graph right behind the detector (e.g. FixedLagSmoother.push: keypoints go straight into the
        smoother's measurement buffer and the newest pose comes back with `result_post`).  It must not allocate or
        synchronise.  A stateful `post` is executed once during construction (warm-up): reset its state afterwards."""
        if model.training:
            raise RuntimeError("BatchPipeline needs model.eval()")
        # graph=False enqueues the ~12 launches of a forward eagerly instead of replaying a captured graph: the eager chain
        # keeps programmatic dependent launch between the kernels (a replay serialises them: +0.02 ms per step) at the price
        # of ~60 us of host time per batch; a `post` hook always runs captured.  Default: PB200_PIPE_EAGER=1 selects eager.
        # Measured r02 (tools/gpu/r02_x.sh, two runs each): e2e 211.9 / 216.9 k frames/s replayed, 215.0 / 215.4 k eager -- no
        # difference, so the replay (no host work per batch) stays the default.
        if graph is None:
            import os
            graph = os.environ.get("PB200_PIPE_EAGER") is None
        self.use_graph = bool(graph) or post is not None
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
        self.compute = torch.cuda.Stream(device=dev, priority=-1)
        self.pre = torch.cuda.Stream(device=dev)  # ingest of the next batch
        self.d2h = torch.cuda.Stream(device=dev)  # keypoint read-back (off the compute stream: a 16 KB copy costs ~8 us of stream time)
        self.copies = [torch.cuda.Stream(device=dev) for _ in range(max(1, copy_streams))]
        with torch.cuda.stream(self.compute):
            self.plan = model.private_plan(dev, batch)
        shape = (batch, self.src_h, self.src_w, channels)
        self.dev_rgb = [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(slots)]
        self.dev_depth = ([torch.empty(shape[:3], dtype=torch.float32, device=dev) for _ in range(slots)]
                          if self.use_depth else [None] * slots)
        self.frames = [torch.empty(self.plan.frames_bytes(batch), dtype=torch.uint8, device=dev) for _ in range(slots)]
        self.px_dev = [torch.empty((batch, model.n_keypoints, 2), dtype=torch.float32, device=dev) for _ in range(slots)]
        self.norm_dev = [torch.empty((batch, 2 * model.n_keypoints), dtype=torch.float32, device=dev) for _ in range(slots)]
        self.host_px = [torch.empty((batch, model.n_keypoints, 2), dtype=torch.float32).pin_memory() for _ in range(results)]
        self.ev_copied = [[torch.cuda.Event() for _ in self.copies] for _ in range(slots)]
        self.ev_free = [torch.cuda.Event() for _ in range(slots)]      # dev_rgb[s] consumed (ingest done)
        self.ev_ingested = [torch.cuda.Event() for _ in range(slots)]  # frames[s] ready
        self.ev_fwd = [torch.cuda.Event() for _ in range(slots)]       # frames[s] consumed (forward done)
        self.ev_read = [torch.cuda.Event() for _ in range(slots)]      # px_dev[s] copied to the host
        self.ev_post_read = torch.cuda.Event()
        self.ev_done = [torch.cuda.Event() for _ in range(results)]
        self.graphs = [None] * slots
        self.ticket = 0
        self.post, self.post_dev, self.host_post = post, None, None
        self._capture()

    def _ingest(self, s: int):
        self.model.ingest(self.dev_rgb[s], self.dev_depth[s], bgr=self.bgr, depth_divisor=self.depth_divisor,
                          out=self.frames[s], plan=self.plan)

    def _run(self, s: int):
        px = self.model.forward_ingested(self.frames[s], self.B, plan=self.plan, out_norm=self.norm_dev[s],
                                         out_px=self.px_dev[s])[1]
        if self.post is not None:
            out = self.post(px)
            if self.post_dev is None:
                self.post_dev = out
                self.host_post = [torch.empty(out.shape, dtype=out.dtype).pin_memory() for _ in range(self.results)]
            elif out.data_ptr() != self.post_dev.data_ptr():
                raise RuntimeError("BatchPipeline: `post` must return the same device tensor on every call")
        return px

    def _capture(self):
        with torch.cuda.stream(self.compute):
            for s in range(self.slots):
                self.dev_rgb[s].zero_()
                if self.use_depth:
                    self.dev_depth[s].fill_(0.3)
            for s in range(self.slots):
                self._ingest(s)
            self._run(0)  # warm-up: one-off attribute / occupancy queries happen here, not under capture
        self.compute.synchronize()
        for s in range(self.slots):
            # the plan caches its TMA descriptors per (frames, workspace, batch): each slot's frames buffer differs only in
            # the stem's input pointer (a kernel argument), so re-preparing per slot is cheap and happens at capture time
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.compute):
                self._run(s)
            self.graphs[s] = g
        for e in self.ev_free + self.ev_fwd + self.ev_read + [self.ev_post_read]:
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
        with torch.cuda.stream(self.pre):
            for e in self.ev_copied[s]:
                self.pre.wait_event(e)
            self.pre.wait_event(self.ev_fwd[s])  # the previous batch in this slot's frames buffer has been consumed
            self._ingest(s)
            self.ev_free[s].record(self.pre)
            self.ev_ingested[s].record(self.pre)
        with torch.cuda.stream(self.compute):
            self.compute.wait_event(self.ev_ingested[s])
            self.model.sync_plan_weights(self.plan)
            self.compute.wait_event(self.ev_read[s])  # px_dev[s] of the batch two submits ago has been read back
            if self.post_dev is not None:
                self.compute.wait_event(self.ev_post_read)  # the shared `post` output of the previous batch has been read
            if self.use_graph:
                self.graphs[s].replay()
            else:
                self._run(s)
            self.ev_fwd[s].record(self.compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(self.ev_fwd[s])
            self.host_px[r].copy_(self.px_dev[s], non_blocking=True)
            if self.post_dev is not None:
                self.host_post[r].copy_(self.post_dev, non_blocking=True)
                self.ev_post_read.record(self.d2h)
            self.ev_done[r].record(self.d2h)
            self.ev_read[s].record(self.d2h)
        self.ticket = t + 1
        return t

    def result(self, ticket: int) -> torch.Tensor:
        """Blocks until the keypoints of `ticket` are in pinned host memory; the returned (B, K, 2) float32 tensor is
        valid until `results` further batches have been submitted."""
        if not (self.ticket - self.results <= ticket < self.ticket):
            raise ValueError(f"result: ticket {ticket} is not outstanding (next ticket {self.ticket}, ring {self.results})")
        self.ev_done[ticket % self.results].synchronize()
        return self.host_px[ticket % self.results]

    def result_post(self, ticket: int) -> torch.Tensor:
        """The `post` output of `ticket` in pinned host memory (same lifetime rule as `result`): read back right after its
        batch's graph; the next batch's graph waits for that read-back, so every ticket sees its own snapshot."""
        if self.post_dev is None:
            raise RuntimeError("result_post: the pipeline was built without `post`")
        self.result(ticket)
        return self.host_post[ticket % self.results]

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
        self.d2h.synchronize()
        self.pre.synchronize()
        for cs in self.copies:
            cs.synchronize()
