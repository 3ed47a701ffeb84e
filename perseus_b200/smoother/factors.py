"""Batched smoother factors on the GPU (reference: perseus/smoother/factors.py).

The reference evaluates one factor per Python callback from GTSAM's C++ optimiser (factors.py:54-61,
160, 216-225).  Here the same residuals and analytic Jacobians are evaluated for N factors per
launch by float64 SIMT kernels (csrc/factors.cu) behind the C ABI:

    pose_dynamics_batch        <- PoseDynamicsFactor.error_func        factors.py:79-142
    constant_velocity_batch    <- ConstantVelocityFactor.error_func    factors.py:160-171
    keypoint_projection_batch  <- KeypointProjectionFactor.error_func  factors.py:245-275

Conventions are GTSAM's: Pose3 tangent [rot; trans], Jacobians w.r.t. right perturbations,
Jacobian storage column-major (returned here as (N, rows, cols) tensors already transposed into
normal row/col indexing).  When `gtsam` is importable the reference's CustomFactor classes are also
provided with unchanged constructors (the callback evaluates a batch of one on the GPU); gtsam is
not installed in this image, so those classes are defined lazily.
"""
from __future__ import annotations

import torch

from .. import _lib


def _f64(t: torch.Tensor, name: str, shape_tail: tuple) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name}: cuda float64 tensor required (no CPU fallback)")
    t = t.to(torch.float64).contiguous()
    if tuple(t.shape[-len(shape_tail):]) != shape_tail:
        raise ValueError(f"{name}: trailing shape must be {shape_tail}, got {tuple(t.shape)}")
    return t


def pose12(R: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    """(N,3,3), (N,3) -> (N,12) row-major R then t: the pose record of the C ABI."""
    return torch.cat([R.reshape(*R.shape[:-2], 9), t], dim=-1).contiguous()


def pose_dynamics_batch(pose1, ang_vel1, vel1, pose2, dt: float, vel_frame: str = "world", jacobians: bool = True):
    """N PoseDynamicsFactors.  pose*: (N,12); ang_vel1, vel1: (N,3).

    Returns (error (N,6), [H0 (N,6,6), H1 (N,6,3), H2 (N,6,3), H3 (N,6,6)] or None) in the key
    order [pose1, ang_vel1, vel1, pose2] of factors.py:48."""
    assert vel_frame in ["world", "body"], "vel_frame must be 'world' or 'body'."  # factors.py:40
    pose1 = _f64(pose1, "pose1", (12,))
    pose2 = _f64(pose2, "pose2", (12,))
    ang_vel1 = _f64(ang_vel1, "ang_vel1", (3,))
    vel1 = _f64(vel1, "vel1", (3,))
    N = pose1.shape[0]
    dev = pose1.device
    resid = torch.empty((N, 6), dtype=torch.float64, device=dev)
    Hs = None
    if jacobians:
        Hs = [torch.empty((N, c, 6), dtype=torch.float64, device=dev) for c in (6, 3, 3, 6)]  # column-major
    with torch.cuda.device(dev):
        _lib.check(
            _lib.lib().pb200_factors_dynamics(
                pose1.data_ptr(), ang_vel1.data_ptr(), vel1.data_ptr(), pose2.data_ptr(), float(dt),
                int(vel_frame == "world"), resid.data_ptr(),
                *([h.data_ptr() for h in Hs] if Hs else [None] * 4), N, _lib.current_stream_ptr(),
            )
        )
    if Hs is not None:
        Hs = [h.transpose(1, 2) for h in Hs]
    return resid, Hs


def constant_velocity_batch(vel1, vel2, jacobians: bool = True):
    """N ConstantVelocityFactors: error = vel2 - vel1, H = [-I, I]."""
    vel1 = _f64(vel1, "vel1", (3,))
    vel2 = _f64(vel2, "vel2", (3,))
    N = vel1.shape[0]
    resid = torch.empty((N, 3), dtype=torch.float64, device=vel1.device)
    H0 = H1 = None
    if jacobians:
        H0 = torch.empty((N, 3, 3), dtype=torch.float64, device=vel1.device)
        H1 = torch.empty((N, 3, 3), dtype=torch.float64, device=vel1.device)
    with torch.cuda.device(vel1.device):
        _lib.check(
            _lib.lib().pb200_factors_constvel(
                vel1.data_ptr(), vel2.data_ptr(), resid.data_ptr(), _lib.ptr(H0), _lib.ptr(H1), N,
                _lib.current_stream_ptr(),
            )
        )
    return resid, ([H0, H1] if jacobians else None)


def keypoint_projection_batch(poses, K5, meas, pts_body, camera_pose=None, jacobians: bool = True):
    """N poses x P keypoints of KeypointProjectionFactor.

    poses (N,12); K5 (5,) = fx, fy, s, u0, v0 (gtsam.Cal3_S2); meas (N,P,2); pts_body (P,3);
    camera_pose (12,) or None = identity (factors.py:212).
    Returns (error (N,P,2), H (N,P,2,6) or None, cheirality (N,P) bool — where GTSAM would raise
    CheiralityException the flag is set instead)."""
    poses = _f64(poses, "poses", (12,))
    dev = poses.device
    K5 = _f64(torch.as_tensor(K5, dtype=torch.float64, device=dev), "K5", (5,))
    meas = _f64(meas, "meas", (2,))
    pts_body = _f64(pts_body, "pts_body", (3,))
    N, P = poses.shape[0], pts_body.shape[0]
    if tuple(meas.shape) != (N, P, 2):
        raise ValueError(f"meas must be (N, P, 2) = {(N, P, 2)}, got {tuple(meas.shape)}")
    cam = None if camera_pose is None else _f64(camera_pose, "camera_pose", (12,))
    resid = torch.empty((N, P, 2), dtype=torch.float64, device=dev)
    jac = torch.empty((N, P, 6, 2), dtype=torch.float64, device=dev) if jacobians else None
    cheir = torch.empty((N, P), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(
            _lib.lib().pb200_factors_projection(
                poses.data_ptr(), meas.data_ptr(), pts_body.data_ptr(), K5.data_ptr(), _lib.ptr(cam),
                resid.data_ptr(), _lib.ptr(jac), cheir.data_ptr(), N, P, _lib.current_stream_ptr(),
            )
        )
    return resid, (jac.transpose(2, 3) if jac is not None else None), cheir.bool()


def __getattr__(name):
    """PoseDynamicsFactor / ConstantVelocityFactor / KeypointProjectionFactor: the reference's
    gtsam.CustomFactor classes, available only where gtsam is importable."""
    if name in ("PoseDynamicsFactor", "ConstantVelocityFactor", "KeypointProjectionFactor"):
        from . import _gtsam_factors  # raises ImportError with a clear message if gtsam is absent

        return getattr(_gtsam_factors, name)
    raise AttributeError(name)
