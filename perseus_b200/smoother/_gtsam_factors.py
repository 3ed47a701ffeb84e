"""The reference's gtsam.CustomFactor classes (perseus/smoother/factors.py:8-275) with unchanged constructors and the
unchanged callback contract `error_func(this, v, H) -> np.ndarray` (factors.py:54-61, 160, 216-225): float64 residual
returned, Jacobians ASSIGNED into the list (`H[i] = ...`, factors.py:114-130, 166-167, 264) in Fortran order, tangent
order [rot; trans], right perturbation.

GTSAM calls the callbacks one factor at a time, with the SAME gtsam.Values for every factor of a linearisation.  A
GPU launch per callback would be slower than the reference's numpy path (one launch + copies >= 100 us vs 16-243 us),
so the factors of a kind that share their constants form a GROUP with a per-linearisation cache:

  * the first callback of a linearisation gathers the variables of EVERY factor of its group from `v`, evaluates the
    whole group with ONE batched launch (pb200_factors_*), and copies residuals + Jacobians back once;
  * every later callback of the same linearisation is a cache hit: it re-reads only its own variables from `v`, checks
    them bit for bit against what the cached batch was evaluated at, and returns its row;
  * any mismatch (new linearisation point) re-evaluates the group; factors whose keys are missing from `v` (a
    sub-graph) are skipped and evaluated alone when they are called.

Importable only where `gtsam` is importable (it is not in this image: tests/stub_gtsam provides the few classes the
callbacks touch so that this module is exercised).  The batched array API in factors.py is what the benchmark measures.
"""
from __future__ import annotations

from functools import partial
from typing import List, Optional

import numpy as np
import torch

try:
    import gtsam
except ImportError as e:  # pragma: no cover
    raise ImportError("perseus_b200.smoother: the CustomFactor classes need the `gtsam` package; "
                      "use the *_batch functions otherwise") from e

from . import factors as F


def _pose12_np(p: "gtsam.Pose3") -> np.ndarray:
    m = np.asarray(p.matrix(), dtype=np.float64)
    return np.concatenate([m[:3, :3].reshape(9), m[:3, 3]])


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("perseus_b200 factors evaluate on a B200 only (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


class _Group:
    """Factors of one kind with identical constants; one batched evaluation per linearisation point."""

    _groups: dict = {}

    def __init__(self, kind: str, const):
        self.kind, self.const = kind, const
        self.members: list = []      # per member: tuple of keys (+ per-factor constants)
        self.inputs: Optional[np.ndarray] = None  # (N, D) variables the cached batch was evaluated at (NaN = not evaluated)
        self.err: Optional[np.ndarray] = None
        self.H: Optional[list] = None
        self.flags: Optional[np.ndarray] = None
        self.evaluations = 0         # batched launches so far (observable by tests / callers)

    @classmethod
    def get(cls, kind: str, const_key, const):
        g = cls._groups.get((kind, const_key))
        if g is None:
            g = cls._groups[(kind, const_key)] = _Group(kind, const)
        return g

    @classmethod
    def reset(cls) -> None:
        """Forget every group (call between independent factor graphs)."""
        cls._groups.clear()

    def add(self, member) -> int:
        self.members.append(member)
        self.inputs = None  # shape changed
        return len(self.members) - 1

    # ---- gathering this kind's variables of one member from a gtsam.Values ----------------------------------------
    def gather(self, v, member) -> np.ndarray:
        if self.kind == "dynamics":
            k = member["keys"]
            return np.concatenate([_pose12_np(v.atPose3(k[0])), np.asarray(v.atVector(k[1]), dtype=np.float64),
                                   np.asarray(v.atVector(k[2]), dtype=np.float64), _pose12_np(v.atPose3(k[3]))])
        if self.kind == "constvel":
            k = member["keys"]
            return np.concatenate([np.asarray(v.atVector(k[0]), dtype=np.float64),
                                   np.asarray(v.atVector(k[1]), dtype=np.float64)])
        return _pose12_np(v.atPose3(member["keys"][0]))  # projection

    def lookup(self, v, idx: int, want_H: bool):
        x = self.gather(v, self.members[idx])
        hit = (self.inputs is not None and self.inputs.shape[0] == len(self.members) and np.array_equal(self.inputs[idx], x)
               and (not want_H or self.H is not None))
        if not hit:
            self.evaluate(v, idx, x, want_H)
        return idx

    def evaluate(self, v, idx: int, x_idx: np.ndarray, want_H: bool) -> None:
        n = len(self.members)
        X = np.full((n, x_idx.shape[0]), np.nan)
        X[idx] = x_idx
        for i, m in enumerate(self.members):
            if i == idx:
                continue
            try:
                X[i] = self.gather(v, m)
            except Exception:  # key not in this Values (sub-graph): evaluated when it is called itself
                pass
        present = np.flatnonzero(~np.isnan(X[:, 0]))
        dev = _device()
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
        Xp = X[present]
        # Jacobians are always computed with the batch: GTSAM asks for them on (almost) every linearisation call and a
        # second launch for a late `H is not None` caller would cost more than the extra D2H bytes
        if self.kind == "dynamics":
            dt, frame = self.const
            err, Hs = F.pose_dynamics_batch(t(Xp[:, :12]), t(Xp[:, 12:15]), t(Xp[:, 15:18]), t(Xp[:, 18:30]), dt, frame)
            flags = None
        elif self.kind == "constvel":
            err, Hs = F.constant_velocity_batch(t(Xp[:, :3]), t(Xp[:, 3:6]))
            flags = None
        else:
            K5, cam12 = self.const
            meas = np.stack([self.members[i]["meas"] for i in present])[:, None, :]
            # per-factor body points: P = 1 per launch row would waste the kernel's P loop, but the reference builds
            # one factor per (pose, corner) (factors.py:182-190); group rows by corner through a (N, 1) problem each
            pts = np.stack([self.members[i]["point"] for i in present])
            err = torch.empty((len(present), 2), dtype=torch.float64, device=dev)
            J = torch.empty((len(present), 2, 6), dtype=torch.float64, device=dev)
            fl = torch.empty((len(present),), dtype=torch.bool, device=dev)
            uniq, inv = np.unique(pts, axis=0, return_inverse=True)
            inv = inv.reshape(-1)
            for u in range(uniq.shape[0]):  # 8 cube corners -> 8 launches per linearisation, independent of N
                sel = np.flatnonzero(inv == u)
                e_u, J_u, c_u = F.keypoint_projection_batch(t(Xp[sel]), t(K5), t(meas[sel]), t(uniq[u][None]),
                                                            None if cam12 is None else t(cam12))
                st = torch.from_numpy(sel).to(dev)
                err[st], J[st], fl[st] = e_u[:, 0], J_u[:, 0], c_u[:, 0]
            Hs, flags = [J], fl.cpu().numpy()
        self.evaluations += 1
        if self.err is None or self.err.shape[0] != n:
            self.err = np.zeros((n, err.shape[-1]))
            self.H = [np.zeros((n,) + tuple(h.shape[1:])) for h in Hs]
            self.flags = np.zeros(n, dtype=bool)
        self.err[present] = err.cpu().numpy()
        for k, h in enumerate(Hs):
            self.H[k][present] = h.contiguous().cpu().numpy()
        if flags is not None:
            self.flags[present] = flags
        self.inputs = X

    def assign(self, idx: int, H) -> np.ndarray:
        if H:
            for k in range(len(self.H)):
                H[k] = np.asfortranarray(self.H[k][idx])  # GTSAM (Eigen) expects column-major float64
        return self.err[idx].copy()


class PoseDynamicsFactor(gtsam.CustomFactor):
    def __init__(self, pose1: int, ang_vel1: int, vel1: int, pose2: int, noise_model, dt: float,
                 vel_frame: str = "world") -> None:
        assert vel_frame in ["world", "body"], "vel_frame must be 'world' or 'body'."  # factors.py:40
        super().__init__(noise_model, [pose1, ang_vel1, vel1, pose2],
                         partial(self.error_func, dt=dt, vel_frame=vel_frame))
        self.dt = dt
        self.vel_frame = vel_frame
        self._group = _Group.get("dynamics", (float(dt), vel_frame), (float(dt), vel_frame))
        self._idx = self._group.add({"keys": (pose1, ang_vel1, vel1, pose2)})

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None, dt=None, vel_frame=None) -> np.ndarray:
        g = self._group
        return g.assign(g.lookup(v, self._idx, bool(H)), H)


class ConstantVelocityFactor(gtsam.CustomFactor):
    def __init__(self, vel1: int, vel2: int, noise_model) -> None:
        super().__init__(noise_model, [vel1, vel2], self.error_func)
        self._group = _Group.get("constvel", (), ())
        self._idx = self._group.add({"keys": (vel1, vel2)})

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None) -> np.ndarray:
        g = self._group
        return g.assign(g.lookup(v, self._idx, bool(H)), H)


class KeypointProjectionFactor(gtsam.CustomFactor):
    def __init__(self, body_pose: int, noise_model, camera_intrinsics, keypoint_measurement: np.ndarray,
                 point_body_frame: np.ndarray, camera_pose=None) -> None:
        super().__init__(
            noise_model, [body_pose],
            partial(self.error_func, camera_intrinsics=camera_intrinsics, keypoint_measurement=keypoint_measurement,
                    point_body_frame=point_body_frame,
                    camera_pose=camera_pose if camera_pose is not None else gtsam.Pose3()),
        )
        c = camera_intrinsics
        K5 = np.array([c.fx(), c.fy(), c.skew(), c.px(), c.py()], dtype=np.float64)
        cam12 = None if camera_pose is None else _pose12_np(camera_pose)
        key = (K5.tobytes(), None if cam12 is None else cam12.tobytes())
        self._group = _Group.get("projection", key, (K5, cam12))
        self._meas = np.asarray(keypoint_measurement, dtype=np.float64).reshape(2)
        self._idx = self._group.add({"keys": (body_pose,), "meas": self._meas,
                                     "point": np.asarray(point_body_frame, dtype=np.float64).reshape(3)})
        self.pixel = None

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None, camera_intrinsics=None,
                   keypoint_measurement=None, point_body_frame=None, camera_pose=None) -> np.ndarray:
        g = self._group
        idx = g.lookup(v, self._idx, bool(H))
        if g.flags[idx]:  # GTSAM raises CheiralityException from PinholeCamera.project (factors.py:260-261)
            raise RuntimeError("CheiralityException: point behind the camera")
        e = g.assign(idx, H)
        self.pixel = e + self._meas  # factors.py:273-274 keeps the projected pixel
        return e


def reset_groups() -> None:
    """Forget the per-kind factor groups (call before building a new, independent factor graph)."""
    _Group.reset()
