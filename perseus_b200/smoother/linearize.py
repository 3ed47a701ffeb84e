"""Fixed-lag smoother step on the GPU: factor evaluation -> whitening -> Gauss-Newton normal
equations -> block-tridiagonal Cholesky solve -> retraction, for W independent windows of T states each.

The reference builds a GTSAM graph from perseus/smoother/factors.py and lets GTSAM linearise and solve it; GTSAM is not
part of the reference's sources.  Here the chain structure of the cube-tracking problem is used directly: state k =
[Pose3 tangent (rot, trans); ang_vel; vel], factors = P KeypointProjectionFactors on pose_k (factors.py:174-275), one
PoseDynamicsFactor on (pose_k, ang_vel_k, vel_k, pose_k+1) (:8-142) and ConstantVelocityFactors on consecutive angular
/ linear velocities (:145-171), each whitened by a diagonal noise model (noiseModel.Diagonal.Sigmas, :42,49).
CUDA tensors only (no CPU path); kernels: csrc/factors.cu, csrc/smoother.cu.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from .. import _lib
from .factors import constant_velocity_batch, keypoint_projection_batch, pose_dynamics_batch


@dataclass
class ChainSigmas:
    pixel: float = 1.0
    dynamics: Sequence[float] = (0.1,) * 6      # [rot; trans], tests/test_dynamics_factor.py:42
    const_ang_vel: Sequence[float] = (0.1,) * 3  # :49
    const_vel: Sequence[float] = (0.1,) * 3

    def as_array(self):
        v = [self.pixel, *self.dynamics, *self.const_ang_vel, *self.const_vel]
        assert len(v) == 13
        return (ctypes.c_double * 13)(*[float(x) for x in v])


def _hat(w: torch.Tensor) -> torch.Tensor:
    z = torch.zeros_like(w[..., 0])
    return torch.stack([z, -w[..., 2], w[..., 1], w[..., 2], z, -w[..., 0], -w[..., 1], w[..., 0], z], -1).reshape(*w.shape[:-1], 3, 3)


def retract(poses: torch.Tensor, ang_vel: torch.Tensor, vel: torch.Tensor, delta: torch.Tensor):
    """gtsam Values.retract on every state, one launch (pb200_smoother_retract): poses (...,12) * Pose3.Expmap(delta[...,:6])
    (full SE(3) exponential, tangent [rot; trans]), ang_vel + delta[...,6:9], vel + delta[...,9:12]."""
    if not poses.is_cuda:
        raise RuntimeError("retract: cuda float64 tensors required (no CPU fallback)")
    f = lambda t: t.to(torch.float64).contiguous()  # noqa: E731
    poses, ang_vel, vel, delta = f(poses), f(ang_vel), f(vel), f(delta)
    n = poses.numel() // 12
    po, ao, vo = torch.empty_like(poses), torch.empty_like(ang_vel), torch.empty_like(vel)
    with torch.cuda.device(poses.device):
        _lib.check(_lib.lib().pb200_smoother_retract(poses.data_ptr(), ang_vel.data_ptr(), vel.data_ptr(), delta.data_ptr(),
                                                     po.data_ptr(), ao.data_ptr(), vo.data_ptr(), n, _lib.current_stream_ptr()))
    return po, ao, vo


def pose_retract(pose12: torch.Tensor, xi: torch.Tensor) -> torch.Tensor:
    """Pose3.retract = pose * Expmap(xi) (full SE(3) exponential, tangent [rot; trans]); (…,12), (…,6) -> (…,12)."""
    z = torch.zeros(*xi.shape[:-1], 3, dtype=torch.float64, device=xi.device)
    d = torch.cat([xi.to(torch.float64), z, z], dim=-1)
    return retract(pose12, z, z, d)[0]


def chain_factors(poses, ang_vel, vel, meas, pts_body, K5, dt, camera_pose=None, vel_frame="world", fused=True, out=None):
    """Evaluates every factor of W chains of T states.  poses (W,T,12), ang_vel / vel (W,T,3), meas (W,T,P,2).
    Returns a dict of residuals and Jacobians in the layouts the C ABI produces (column-major Jacobians).
    fused=True: ONE launch (pb200_factors_chain); False: the three per-kind launches (same results, used as a check).
    `out`: a dict returned by an earlier call with the same shapes, whose buffers are overwritten (no allocation)."""
    W, T = poses.shape[:2]
    P = pts_body.shape[0]
    if not fused:
        pr, pJ, cheir = keypoint_projection_batch(poses.reshape(W * T, 12), K5, meas.reshape(W * T, P, 2), pts_body, camera_pose)
        res = {"proj_r": pr, "proj_J": pJ, "cheirality": cheir.reshape(W, T, P)}
        if T > 1:
            p1, p2 = poses[:, :-1].reshape(-1, 12), poses[:, 1:].reshape(-1, 12)
            dr, Hs = pose_dynamics_batch(p1, ang_vel[:, :-1].reshape(-1, 3), vel[:, :-1].reshape(-1, 3), p2, dt, vel_frame)
            cw, _ = constant_velocity_batch(ang_vel[:, :-1].reshape(-1, 3), ang_vel[:, 1:].reshape(-1, 3), jacobians=False)
            cv, _ = constant_velocity_batch(vel[:, :-1].reshape(-1, 3), vel[:, 1:].reshape(-1, 3), jacobians=False)
            res.update(dyn_r=dr, H=Hs, cvw_r=cw, cvv_r=cv)
        return res
    assert vel_frame in ["world", "body"], "vel_frame must be 'world' or 'body'."
    dev = poses.device
    f64 = dict(dtype=torch.float64, device=dev)
    poses, ang_vel, vel, meas = (t.to(torch.float64).contiguous() for t in (poses, ang_vel, vel, meas))
    pts_body = pts_body.to(torch.float64).contiguous()
    K5 = torch.as_tensor(K5, **f64).contiguous()
    cam = None if camera_pose is None else camera_pose.to(torch.float64).contiguous()
    N, M = W * T, W * (T - 1)
    if out is None:
        raw = {"proj_r": torch.empty(N, P, 2, **f64), "proj_Jc": torch.empty(N, P, 6, 2, **f64),
               "cheir_u8": torch.empty(N, P, dtype=torch.uint8, device=dev)}
        if T > 1:
            raw.update(dyn_r=torch.empty(M, 6, **f64), Hc=[torch.empty(M, c, 6, **f64) for c in (6, 3, 3, 6)],
                       cvw_r=torch.empty(M, 3, **f64), cvv_r=torch.empty(M, 3, **f64))
    else:
        raw = out["_raw"]
    Hc = raw.get("Hc", [None] * 4)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().pb200_factors_chain(
            poses.data_ptr(), ang_vel.data_ptr(), vel.data_ptr(), meas.data_ptr(), pts_body.data_ptr(), K5.data_ptr(),
            _lib.ptr(cam), float(dt), int(vel_frame == "world"), raw["proj_r"].data_ptr(), raw["proj_Jc"].data_ptr(),
            raw["cheir_u8"].data_ptr(), _lib.ptr(raw.get("dyn_r")), *[_lib.ptr(h) for h in Hc], _lib.ptr(raw.get("cvw_r")),
            _lib.ptr(raw.get("cvv_r")), W, T, P, _lib.current_stream_ptr()))
    res = {"proj_r": raw["proj_r"], "proj_J": raw["proj_Jc"].transpose(2, 3), "cheirality": raw["cheir_u8"].bool().reshape(W, T, P),
           "_raw": raw}
    if T > 1:
        res.update(dyn_r=raw["dyn_r"], H=[h.transpose(1, 2) for h in Hc], cvw_r=raw["cvw_r"], cvv_r=raw["cvv_r"])
    return res


def chain_cost(f: dict, sig: ChainSigmas) -> torch.Tensor:
    """0.5 * sum of whitened squared residuals over everything in `f` (a scalar tensor)."""
    c = (f["proj_r"] / sig.pixel).pow(2).sum()
    if "dyn_r" in f:
        dev = f["dyn_r"].device
        c = c + (f["dyn_r"] / torch.tensor(list(sig.dynamics), dtype=torch.float64, device=dev)).pow(2).sum()
        c = c + (f["cvw_r"] / torch.tensor(list(sig.const_ang_vel), dtype=torch.float64, device=dev)).pow(2).sum()
        c = c + (f["cvv_r"] / torch.tensor(list(sig.const_vel), dtype=torch.float64, device=dev)).pow(2).sum()
    return 0.5 * c


def normal_equations(f: dict, W: int, T: int, P: int, sig: ChainSigmas, lam: float = 0.0, prior: float = 0.0):
    """Whitened Gauss-Newton system of W chains: D (W,T,12,12) diagonal blocks, O (W,T-1,12,12) blocks H[k+1][k],
    g (W,T,12) = J^T W r."""
    dev = f["proj_r"].device
    D = torch.empty(W, T, 12, 12, dtype=torch.float64, device=dev)
    O = torch.empty(W, max(T - 1, 1), 12, 12, dtype=torch.float64, device=dev)
    g = torch.empty(W, T, 12, dtype=torch.float64, device=dev)
    cm = lambda h: h.transpose(1, 2)  # back to the contiguous column-major buffer the kernel wrote  # noqa: E731
    H = f.get("H")
    args = [f["proj_r"].data_ptr(), f["proj_J"].transpose(2, 3).data_ptr()]
    if T > 1:
        assert all(cm(h).is_contiguous() for h in H) and f["proj_J"].transpose(2, 3).is_contiguous()
        args += [f["dyn_r"].data_ptr(), *[cm(h).data_ptr() for h in H], f["cvw_r"].data_ptr(), f["cvv_r"].data_ptr()]
    else:
        args += [None] * 7
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().pb200_smoother_normal_equations(*args, sig.as_array(), float(lam), float(prior), D.data_ptr(),
                                                              O.data_ptr(), g.data_ptr(), W, T, P, _lib.current_stream_ptr()))
    return D, O[:, :T - 1], g


def chain_solve(D: torch.Tensor, O: torch.Tensor, g: torch.Tensor):
    """Solves H delta = -g for the block-tridiagonal H of every chain.  Returns (delta (W,T,12), status (W,) int32:
    non-zero where a pivot was not positive)."""
    W, T = D.shape[:2]
    dev = D.device
    work = torch.empty(W, T, 2, 144, dtype=torch.float64, device=dev)
    delta = torch.empty(W, T, 12, dtype=torch.float64, device=dev)
    status = torch.empty(W, dtype=torch.int32, device=dev)
    Oc = O.contiguous() if T > 1 else None
    Dc, gc = D.contiguous(), g.contiguous()  # held until after the launch (a temporary's block could be handed out again)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().pb200_smoother_solve(Dc.data_ptr(), _lib.ptr(Oc), gc.data_ptr(),
                                                   work.data_ptr(), delta.data_ptr(), status.data_ptr(), W, T,
                                                   _lib.current_stream_ptr()))
    return delta, status


def gauss_newton_step(poses, ang_vel, vel, meas, pts_body, K5, dt, sigmas: Optional[ChainSigmas] = None,
                      camera_pose=None, vel_frame: str = "world", lam: float = 0.0, prior: float = 0.0):
    """One smoother iteration over W windows: returns (new poses, new ang_vel, new vel, info dict)."""
    sig = sigmas or ChainSigmas()
    W, T = poses.shape[:2]
    P = pts_body.shape[0]
    f = chain_factors(poses, ang_vel, vel, meas, pts_body, K5, dt, camera_pose, vel_frame)
    D, O, g = normal_equations(f, W, T, P, sig, lam, prior)
    delta, status = chain_solve(D, O, g)
    new_poses, new_w, new_v = retract(poses, ang_vel, vel, delta)
    return new_poses, new_w, new_v, {
        "cost": chain_cost(f, sig), "delta": delta, "status": status, "D": D, "O": O, "g": g, "factors": f}
