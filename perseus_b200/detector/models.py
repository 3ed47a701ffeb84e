"""Drop-in `KeypointCNN` (reference: perseus/detector/models.py:6-40).

Same constructor, attributes (`.resnet .n_keypoints .num_channels .H .W`), `state_dict()` keys
(`resnet.*`, 122 tensors) and `forward(x) -> (B, 2*n_keypoints)` contract as the reference, so the
reference's validate / streaming / train callers keep working; checkpoints saved by the reference
(`train.py:355`, optionally with DDP's `module.` prefix) load unchanged.

What differs (documented in DESIGN.md):
  * inference (`model.eval()`, as every reference inference caller does) runs hand-written sm_100a CUDA through the C ABI
    (libperseus_b200.so): frames repack -> tcgen05 implicit-GEMM backbone with BN folded -> head.
    There is NO CPU / PyTorch fallback on that path: a CPU tensor or a missing library raises.
  * construction does not download ImageNet weights (the reference does, models.py:20).
  * training-mode forward goes through the torch module tree (library autograd),
    which is what DDP wraps for the data-parallel training step (train.py:152,287).
"""
from __future__ import annotations

import ctypes

import os

import torch
from torch import nn
from torchvision import models as tv_models

from .. import _lib


def _device_index(device: torch.device) -> int:
    """torch.device('cuda') (no index) means the CURRENT device, not device 0."""
    return device.index if device.index is not None else torch.cuda.current_device()


class _Plan:
    """Owns one pb200_plan + its workspace for a (device, precision, max_batch)."""

    def __init__(self, device: torch.device, max_batch: int, C: int, H: int, W: int, n_kp: int, precision: str):
        if precision not in _lib.PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_lib.PRECISIONS)}, got {precision!r}")
        self.lib = _lib.lib()
        self.device = device
        self.max_batch = max_batch
        self.precision = precision
        handle = ctypes.c_void_p()
        _lib.check(
            self.lib.pb200_plan_create(
                ctypes.byref(handle), _device_index(device), max_batch, C, H, W, n_kp, _lib.PRECISIONS[precision]
            )
        )
        self.handle = handle
        nbytes = ctypes.c_size_t()
        _lib.check(self.lib.pb200_plan_workspace_bytes(self.handle, ctypes.byref(nbytes)))
        self.workspace = torch.empty(nbytes.value, dtype=torch.uint8, device=device)
        self.weights_key = None

    def frames_bytes(self, B: int) -> int:
        nbytes = ctypes.c_size_t()
        _lib.check(self.lib.pb200_plan_frames_bytes(self.handle, B, ctypes.byref(nbytes)))
        return nbytes.value

    def load_weights(self, state_dict: dict) -> None:
        names, tensors = [], []
        for k, v in state_dict.items():
            if not torch.is_floating_point(v):
                continue  # num_batches_tracked
            names.append(k.encode())
            tensors.append(v.detach().to(device="cpu", dtype=torch.float32).contiguous())
        n = len(names)
        ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in tensors])
        shape_arrs = [(ctypes.c_int64 * max(t.dim(), 1))(*t.shape) for t in tensors]
        shapes = (ctypes.POINTER(ctypes.c_int64) * n)(
            *[ctypes.cast(a, ctypes.POINTER(ctypes.c_int64)) for a in shape_arrs]
        )
        ndims = (ctypes.c_int * n)(*[t.dim() for t in tensors])
        cnames = (ctypes.c_char_p * n)(*names)
        with torch.cuda.device(self.device):
            _lib.check(
                self.lib.pb200_plan_load_weights(
                    self.handle, ptrs, shapes, ndims, cnames, n, _lib.current_stream_ptr()
                )
            )

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.pb200_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class KeypointCNN(nn.Module):
    """perseus keypoint CNN (ResNet-18 regressor), B200-native inference.

    Args mirror the reference (models.py:9): n_keypoints, num_channels, H, W.
    Extra keyword-only knobs: `max_batch` (initial plan capacity; grows on demand) and `precision`:
      "fp16x2" (default) tcgen05, fp16 activations x (fp16 hi + fp16 lo) weights, two MMAs per K step: the fast mode
               that meets north_star's 0.05 px against the reference on RGB (measured ~0.03 px, random-init weights);
      "bf16"   tcgen05 single pass: the highest throughput, 0.6 px off the reference on random-init weights
               (opt-in; single-pass bf16 cannot meet 0.05 px);
      "bf16x3" tcgen05 split-bf16 (3 MMAs per K step, [hi|lo] activations): ~0.006 px;
      "fp32"   SIMT FFMA: the <= 1e-3 px parity mode.
    A plan is bound to one device and one stream at a time (its workspace and cached TMA descriptors are mutable).
    """

    def __init__(
        self,
        n_keypoints: int = 8,
        num_channels: int = 3,
        H: int = 256,
        W: int = 256,
        *,
        precision: str = "fp16x2",
        max_batch: int = 256,
    ) -> None:
        super().__init__()
        # Same module tree as the reference so state_dict keys/shapes are identical
        # (models.py:20,27-32); random init instead of the ImageNet download.
        self.resnet = tv_models.resnet18(weights=None)
        self.n_keypoints = n_keypoints
        self.num_channels = num_channels
        self.H = H
        self.W = W
        if num_channels != 3:  # models.py:27-28
            self.resnet.conv1 = nn.Conv2d(num_channels, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.resnet.avgpool = nn.AdaptiveAvgPool2d((1, 1))  # models.py:31
        self.resnet.fc = nn.Linear(self.resnet.fc.in_features, 2 * n_keypoints)  # models.py:32
        self.precision = precision
        self._max_batch = max_batch
        self._plans: dict = {}
        self._tensors = None

    # ---- plan / weight management ----------------------------------------------------------
    def __getstate__(self):
        state = self.__dict__.copy()
        state["_plans"] = {}  # native handles are per-process; rebuilt lazily
        state["_tensors"] = None
        return state

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._tensors = None
        return out

    def _weights_key(self):
        if self._tensors is None:
            self._tensors = tuple(t for t in self.state_dict(keep_vars=True).values() if torch.is_floating_point(t))
        ts = self._tensors
        return (ts[0].data_ptr(), sum(t._version for t in ts))

    def refresh_weights(self) -> None:
        """Forces the folded/packed device weights to be rebuilt on the next native forward."""
        self._tensors = None
        for p in self._plans.values():
            p.weights_key = None

    def _plan(self, device: torch.device, B: int) -> _Plan:
        key = (_device_index(device), self.precision)
        plan = self._plans.get(key)
        if plan is None or plan.max_batch < B:
            cap = max(B, self._max_batch if plan is None else plan.max_batch)
            self._plans.pop(key, None)
            del plan
            plan = _Plan(device, cap, self.num_channels, self.H, self.W, self.n_keypoints, self.precision)
            self._plans[key] = plan
        wk = self._weights_key()
        if plan.weights_key != wk:
            plan.load_weights(self.state_dict())
            plan.weights_key = wk
        return plan

    # ---- checkpoints ---------------------------------------------------------------------------
    _PREFIXES = ("_orig_mod.", "module.")

    @classmethod
    def normalise_keys(cls, state_dict: dict) -> dict:
        """Reference checkpoints are `torch.save(model.state_dict())` (train.py:355): under DDP every key carries
        `module.` (train.py:152), under `torch.compile` (train.py:194-196) `_orig_mod.`, possibly both.  The reference
        callers strip `module.` themselves (validate.py:94-96) and fail on `_orig_mod.`; here both are accepted."""
        out = {}
        for k, v in state_dict.items():
            changed = True
            while changed:
                changed = False
                for pre in cls._PREFIXES:
                    if pre in k:
                        k = k.replace(pre, "")
                        changed = True
            out[k] = v
        return out

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        return super().load_state_dict(self.normalise_keys(state_dict), strict=strict, assign=assign)

    def load_checkpoint(self, path, map_location="cpu") -> "KeypointCNN":
        """validate.py:93-98 / streaming.py:113-119 in one call: `torch.load(path, weights_only=True)`, prefix strip,
        strict load, `.eval()`."""
        sd = torch.load(str(path), weights_only=True, map_location=map_location)
        self.load_state_dict(sd, strict=True)
        return self.eval()

    # ---- reference API -----------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: (B, num_channels, H, W) float, RGB in [0,1], depth in data units -> (B, 2*n_keypoints)."""
        if self.training:
            # train mode = batch-statistics BatchNorm + autograd: the data-parallel training step
            # (train.py:287) runs on the torch module tree that DDP wraps.
            return self.resnet(x)
        norm, _ = self._native_forward(x, want_pixels=False)
        return norm.to(x.dtype) if norm.dtype != x.dtype else norm

    def forward_pixels(self, x: torch.Tensor):
        """forward() plus the pixel decode (validate.py:144-153) fused in the head kernel.

        Returns (normalised (B, 2K), pixels (B, K, 2)) in float32."""
        return self._native_forward(x, want_pixels=True)

    def _native_forward(self, x: torch.Tensor, want_pixels: bool):
        squeeze = x.dim() == 3
        if squeeze:
            x = x[None]
        if x.dim() != 4 or x.shape[1] != self.num_channels or x.shape[2] != self.H or x.shape[3] != self.W:
            raise ValueError(
                f"expected (B, {self.num_channels}, {self.H}, {self.W}) input, got {tuple(x.shape)}"
            )
        if not x.is_cuda:
            raise RuntimeError(
                "perseus_b200.KeypointCNN inference runs only on a B200 (sm_100a) through "
                "libperseus_b200.so; got a CPU tensor and there is no CPU fallback — move the model and "
                "input to cuda"
            )
        B = x.shape[0]
        xf = x.detach().to(torch.float32).contiguous()
        plan = self._plan(x.device, B)
        out_norm = torch.empty((B, 2 * self.n_keypoints), dtype=torch.float32, device=x.device)
        out_px = torch.empty((B, self.n_keypoints, 2), dtype=torch.float32, device=x.device) if want_pixels else None
        with torch.cuda.device(x.device):
            _lib.check(
                plan.lib.pb200_forward_nchw_f32(
                    plan.handle, xf.data_ptr(), plan.workspace.data_ptr(), out_norm.data_ptr(),
                    _lib.ptr(out_px), B, _lib.current_stream_ptr(),
                )
            )
        if squeeze:
            out_norm = out_norm[0]
            out_px = out_px[0] if out_px is not None else None
        return out_norm, out_px

    # ---- fast path: raw camera / dataset frames ------------------------------------------------
    def ingest(
        self,
        rgb_u8: torch.Tensor,
        depth: torch.Tensor | None = None,
        *,
        bgr: bool = False,
        depth_divisor: float = 1.0,
        clip: tuple | None = None,
        cube_scale: float = 0.035,
        jitter: torch.Tensor | None = None,
        out: torch.Tensor | None = None,
        plan: "_Plan | None" = None,
    ) -> torch.Tensor:
        """uint8 (B, h, w, 3|4) [+ float32 depth (B, h, w)] -> the plan's frames buffer (uint8 view).

        Fuses scripts/streaming.py:68-80 (BGR(A)->RGB, /255, NaN/Inf->0, /0.035, centre crop) and
        the val-mode depth clip of augmentations.py:137-168 (`clip=(near, far)` in scaled metres,
        optional `jitter` (B, 2, H, W))."""
        if not rgb_u8.is_cuda:
            raise RuntimeError("ingest: frames must be cuda tensors (no CPU fallback)")
        if rgb_u8.dtype != torch.uint8 or rgb_u8.dim() != 4 or rgb_u8.shape[-1] not in (3, 4):
            raise ValueError("ingest: rgb_u8 must be uint8 (B, h, w, 3|4)")
        rgb_u8 = rgb_u8.contiguous()
        B, sh, sw, ch = rgb_u8.shape
        if depth is not None:
            if depth.shape != (B, sh, sw):
                raise ValueError(f"ingest: depth must be (B, h, w) = {(B, sh, sw)}, got {tuple(depth.shape)}")
            depth = depth.to(torch.float32).contiguous()
        if jitter is not None:
            jitter = jitter.to(torch.float32).contiguous()
            if jitter.shape != (B, 2, self.H, self.W):
                raise ValueError("ingest: jitter must be (B, 2, H, W)")
        plan = plan if plan is not None else self._plan(rgb_u8.device, B)
        nbytes = plan.frames_bytes(B)
        if out is None:
            out = torch.empty(nbytes, dtype=torch.uint8, device=rgb_u8.device)
        elif out.numel() * out.element_size() < nbytes:
            raise ValueError("ingest: out buffer too small")
        near, far = clip if clip is not None else (0.0, 0.0)
        with torch.cuda.device(rgb_u8.device):
            _lib.check(
                plan.lib.pb200_ingest(
                    plan.handle, rgb_u8.data_ptr(), ch, sw * ch, sh, sw, int(bgr), _lib.ptr(depth),
                    float(depth_divisor), int(clip is not None), float(cube_scale), float(near), float(far),
                    _lib.ptr(jitter), out.data_ptr(), B, _lib.current_stream_ptr(),
                )
            )
        return out

    def private_plan(self, device: torch.device, max_batch: int) -> "_Plan":
        """A plan (workspace + packed weights) that this module never regrows or frees: for callers that bake its
        pointers into a CUDA graph (StreamingDetector).  The caller keeps it alive and reloads the weights through
        `sync_plan_weights` when the module's parameters change."""
        plan = _Plan(device, max_batch, self.num_channels, self.H, self.W, self.n_keypoints, self.precision)
        self.sync_plan_weights(plan)
        return plan

    def sync_plan_weights(self, plan: "_Plan") -> bool:
        """(Re)packs the module's current weights into `plan` on the current stream if they changed; True if it did."""
        wk = self._weights_key()
        if plan.weights_key == wk:
            return False
        plan.load_weights(self.state_dict())
        plan.weights_key = wk
        return True

    def forward_ingested(self, frames: torch.Tensor, B: int, plan: "_Plan | None" = None, out_norm: torch.Tensor | None = None,
                         out_px: torch.Tensor | None = None):
        """Backbone + head + decode on an ingested frames buffer -> (norm (B,2K), pixels (B,K,2)); `out_norm` / `out_px`
        may be preallocated float32 cuda tensors (graph-captured callers)."""
        plan = plan if plan is not None else self._plan(frames.device, B)
        if out_norm is None:
            out_norm = torch.empty((B, 2 * self.n_keypoints), dtype=torch.float32, device=frames.device)
        if out_px is None:
            out_px = torch.empty((B, self.n_keypoints, 2), dtype=torch.float32, device=frames.device)
        with torch.cuda.device(frames.device):
            _lib.check(
                plan.lib.pb200_forward(
                    plan.handle, frames.data_ptr(), plan.workspace.data_ptr(), out_norm.data_ptr(),
                    out_px.data_ptr(), B, _lib.current_stream_ptr(),
                )
            )
        return out_norm, out_px

    def fused_rgb8_ok(self, rgb_u8: torch.Tensor, ingest_kwargs: dict | None = None) -> bool:
        """True when `forward_rgb8` applies: a tight uint8 (B, H, W, 3) cuda batch of the model's size on an RGB model in a
        16-bit tcgen05 mode, no crop and nothing but `bgr` asked of the ingest step."""
        kw = ingest_kwargs or {}
        return (self.num_channels == 3 and self.precision in ("bf16", "fp16x2") and rgb_u8.is_cuda
                and rgb_u8.dtype == torch.uint8 and rgb_u8.dim() == 4 and tuple(rgb_u8.shape[1:]) == (self.H, self.W, 3)
                and rgb_u8.is_contiguous() and rgb_u8.data_ptr() % 16 == 0
                # (the depth arguments act on the depth channel only, which an RGB model does not have)
                and set(kw) <= {"bgr", "depth_divisor", "clip", "cube_scale", "jitter"})

    def forward_rgb8(self, rgb_u8: torch.Tensor, bgr: bool = False, plan: "_Plan | None" = None,
                     out_norm: torch.Tensor | None = None, out_px: torch.Tensor | None = None):
        """uint8 (B, H, W, 3) batch -> (norm (B,2K), pixels (B,K,2)) with the ingest step fused into the stem kernel
        (pb200_forward_rgb8): bit-identical to ingest() + forward_ingested(), without the frames tensor and its launch."""
        B = rgb_u8.shape[0]
        plan = plan if plan is not None else self._plan(rgb_u8.device, B)
        if out_norm is None:
            out_norm = torch.empty((B, 2 * self.n_keypoints), dtype=torch.float32, device=rgb_u8.device)
        if out_px is None:
            out_px = torch.empty((B, self.n_keypoints, 2), dtype=torch.float32, device=rgb_u8.device)
        with torch.cuda.device(rgb_u8.device):
            _lib.check(
                plan.lib.pb200_forward_rgb8(
                    plan.handle, rgb_u8.data_ptr(), int(bool(bgr)), plan.workspace.data_ptr(), out_norm.data_ptr(),
                    out_px.data_ptr(), B, _lib.current_stream_ptr(),
                )
            )
        return out_norm, out_px

    def forward_frames(self, rgb_u8: torch.Tensor, depth: torch.Tensor | None = None, **ingest_kwargs):
        """Ingest + forward + decode: uint8 frames -> pixel keypoints (B, K, 2) float32."""
        if self.num_channels == 4 and depth is None:
            raise ValueError("forward_frames: an RGBD model needs a depth tensor")
        if self.num_channels == 3:
            depth = None
        # Opt-in (PB200_FUSED_INGEST=1): measured r02 at bs 256, the fused stem is 2-3 % FASTER than ingest + forward under the
        # power cap (1.14 vs 1.18 ms) but 8-12 % slower at full clocks (1.11-1.16 vs 1.03 ms): the conversion work lands on
        # the SMs whose pooling epilogue already bounds the stem.  It saves the 141 MB frames tensor either way.
        if os.environ.get("PB200_FUSED_INGEST") is not None and self.fused_rgb8_ok(rgb_u8, ingest_kwargs):
            return self.forward_rgb8(rgb_u8, bgr=bool(ingest_kwargs.get("bgr", False)))[1]
        frames = self.ingest(rgb_u8, depth, **ingest_kwargs)
        _, px = self.forward_ingested(frames, rgb_u8.shape[0])
        return px

    def profile_ops(self, frames: torch.Tensor, B: int):
        """Per-kernel CUDA-event times of one forward: list of (name, ms, flops, bytes, kernel)."""
        plan = self._plan(frames.device, B)
        n = plan.lib.pb200_plan_num_ops(plan.handle)
        ms = (ctypes.c_float * n)()
        out_norm = torch.empty((B, 2 * self.n_keypoints), dtype=torch.float32, device=frames.device)
        with torch.cuda.device(frames.device):
            _lib.check(
                plan.lib.pb200_forward_profile(
                    plan.handle, frames.data_ptr(), plan.workspace.data_ptr(), out_norm.data_ptr(), None, B,
                    _lib.current_stream_ptr(), ms,
                )
            )
        rows = []
        for i in range(n):
            fl, by = ctypes.c_double(), ctypes.c_double()
            _lib.check(plan.lib.pb200_plan_op_work(plan.handle, i, B, ctypes.byref(fl), ctypes.byref(by)))
            name = plan.lib.pb200_plan_op_name(plan.handle, i).decode()
            k = plan.lib.pb200_plan_op_chain_len(plan.handle, i)
            if k == 0:  # covered by the chained launch whose row is the last one appended: add this conv's work to it
                h = rows[-1]
                rows[-1] = (h[0], h[1], h[2] + fl.value, h[3] + by.value, h[4])
                continue
            if k >= 2:
                last = plan.lib.pb200_plan_op_name(plan.handle, i + k - 1).decode()
                name = f"{name.split('.')[0]}..{last.split('.')[0]} ({k} convs, one chained launch)"
            rows.append((name, float(ms[i]), fl.value, by.value, plan.lib.pb200_plan_op_kernel(plan.handle, i).decode()))
        return rows
