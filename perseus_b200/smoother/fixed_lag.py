"""Fixed-lag smoother on the device, fed straight from the detector (the camera -> keypoints ->
smoother chain of scripts/streaming.py:121-131 + perseus/smoother/factors.py:174-275).

W independent windows (one per camera / tracked cube) of T states each live in preallocated device buffers.  `push`
takes the detector's pixel keypoints (float32 (W, P, 2), the tensor `pb200_forward` writes) and enqueues, with no
allocation and no host synchronisation (so the whole thing can sit inside a CUDA graph behind the detector):

    pb200_smoother_push               drop the oldest state, append the dynamics prediction, insert the measurement
    iters x [ pb200_factors_chain     every projection / dynamics / constant-velocity factor of the window, ONE launch
              pb200_smoother_normal_equations   whitened Gauss-Newton blocks (noiseModel.Diagonal sigmas, factors.py:42,49)
              pb200_smoother_solve    block-tridiagonal Cholesky
              pb200_smoother_retract ]          Values.retract

The reference hands its factors to GTSAM's fixed-lag smoother (README.md:25; not part of the reference's sources), so
this has no reference code to be pinned against: tests check it against oracle/smoother_oracle.py and by recovering a
synthetic trajectory.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _lib
from .linearize import ChainSigmas


class FixedLagSmoother:
    def __init__(self, W: int, T: int, pts_body: torch.Tensor, K5, dt: float, *, sigmas: Optional[ChainSigmas] = None,
                 camera_pose: Optional[torch.Tensor] = None, vel_frame: str = "world", iters: int = 2, lam: float = 1e-6,
                 prior: float = 0.0, device: Optional[torch.device] = None):
        assert vel_frame in ["world", "body"], "vel_frame must be 'world' or 'body'."
        dev = device or pts_body.device
        if dev.type != "cuda":
            raise RuntimeError("FixedLagSmoother runs only on a B200 (no CPU fallback)")
        self.dev, self.W, self.T, self.P = dev, W, T, int(pts_body.shape[0])
        self.dt, self.world, self.iters, self.lam, self.prior = float(dt), int(vel_frame == "world"), iters, float(lam), float(prior)
        self.sig = sigmas or ChainSigmas()
        self._sig13 = self.sig.as_array()
        f64 = dict(dtype=torch.float64, device=dev)
        self.pts = pts_body.to(**f64).contiguous()
        self.K5 = torch.as_tensor(K5, **f64).contiguous()
        self.cam = None if camera_pose is None else camera_pose.to(**f64).contiguous()
        P, N, M = self.P, W * T, W * max(T - 1, 1)
        self.poses = torch.zeros(W, T, 12, **f64)
        self.poses[..., 0] = self.poses[..., 4] = self.poses[..., 8] = 1.0
        self.ang_vel = torch.zeros(W, T, 3, **f64)
        self.vel = torch.zeros(W, T, 3, **f64)
        self.meas = torch.zeros(W, T, P, 2, **f64)
        # factor outputs / Gauss-Newton system / solution: allocated once
        self.proj_r = torch.empty(N, P, 2, **f64)
        self.proj_J = torch.empty(N, P, 6, 2, **f64)
        self.cheir = torch.empty(N, P, dtype=torch.uint8, device=dev)
        self.dyn_r = torch.empty(M, 6, **f64)
        self.H = [torch.empty(M, c, 6, **f64) for c in (6, 3, 3, 6)]
        self.cvw_r = torch.empty(M, 3, **f64)
        self.cvv_r = torch.empty(M, 3, **f64)
        self.D = torch.empty(W, T, 12, 12, **f64)
        self.O = torch.empty(W, max(T - 1, 1), 12, 12, **f64)
        self.g = torch.empty(W, T, 12, **f64)
        self.work = torch.empty(W, T, 2, 144, **f64)
        self.delta = torch.empty(W, T, 12, **f64)
        self.status = torch.zeros(W, dtype=torch.int32, device=dev)

    def reset(self, poses: torch.Tensor, ang_vel: torch.Tensor, vel: torch.Tensor, meas: torch.Tensor) -> None:
        """Initial window contents: poses (W,T,12), ang_vel / vel (W,T,3), meas (W,T,P,2) pixel measurements."""
        self.poses.copy_(poses)
        self.ang_vel.copy_(ang_vel)
        self.vel.copy_(vel)
        self.meas.copy_(meas)

    def iterate(self) -> None:
        """One Gauss-Newton iteration on the current window contents (4 launches, no allocation)."""
        lib, s = _lib.lib(), _lib.current_stream_ptr()
        W, T, P = self.W, self.T, self.P
        multi = T > 1
        with torch.cuda.device(self.dev):
            _lib.check(lib.pb200_factors_chain(
                self.poses.data_ptr(), self.ang_vel.data_ptr(), self.vel.data_ptr(), self.meas.data_ptr(), self.pts.data_ptr(),
                self.K5.data_ptr(), _lib.ptr(self.cam), self.dt, self.world, self.proj_r.data_ptr(), self.proj_J.data_ptr(),
                self.cheir.data_ptr(), self.dyn_r.data_ptr() if multi else None, *[h.data_ptr() if multi else None for h in self.H],
                self.cvw_r.data_ptr() if multi else None, self.cvv_r.data_ptr() if multi else None, W, T, P, s))
            _lib.check(lib.pb200_smoother_normal_equations(
                self.proj_r.data_ptr(), self.proj_J.data_ptr(), self.dyn_r.data_ptr() if multi else None,
                *[h.data_ptr() if multi else None for h in self.H], self.cvw_r.data_ptr() if multi else None,
                self.cvv_r.data_ptr() if multi else None, self._sig13, self.lam, self.prior, self.D.data_ptr(),
                self.O.data_ptr(), self.g.data_ptr(), W, T, P, s))
            _lib.check(lib.pb200_smoother_solve(self.D.data_ptr(), self.O.data_ptr() if multi else None, self.g.data_ptr(),
                                                self.work.data_ptr(), self.delta.data_ptr(), self.status.data_ptr(), W, T, s))
            _lib.check(lib.pb200_smoother_retract(self.poses.data_ptr(), self.ang_vel.data_ptr(), self.vel.data_ptr(),
                                                  self.delta.data_ptr(), self.poses.data_ptr(), self.ang_vel.data_ptr(),
                                                  self.vel.data_ptr(), W * T, s))

    def push(self, keypoints_px: torch.Tensor) -> torch.Tensor:
        """Advance every window by one frame with the detector's keypoints (float32 (W, P, 2) cuda) and re-optimise.
        Returns the view of the newest poses (W, 12) (valid until the next push; same address every call)."""
        if keypoints_px.dtype != torch.float32 or tuple(keypoints_px.shape) != (self.W, self.P, 2) or not keypoints_px.is_cuda:
            raise ValueError(f"push: expected float32 cuda ({self.W}, {self.P}, 2) keypoints")
        kp = keypoints_px.contiguous()
        with torch.cuda.device(self.dev):
            _lib.check(_lib.lib().pb200_smoother_push(self.poses.data_ptr(), self.ang_vel.data_ptr(), self.vel.data_ptr(),
                                                      self.meas.data_ptr(), kp.data_ptr(), self.dt, self.world, self.W, self.T,
                                                      self.P, _lib.current_stream_ptr()))
        for _ in range(self.iters):
            self.iterate()
        return self.newest()

    def newest(self) -> torch.Tensor:
        return self.poses[:, -1]
