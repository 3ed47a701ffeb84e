"""The reference's gtsam.CustomFactor classes (perseus/smoother/factors.py:8-275) with unchanged
constructors; each callback evaluates a batch of one through the GPU kernels.  Importable only
where `gtsam` is installed (it is not in this image, so this module is untested here; the batched
API in factors.py is what the parity tests and the benchmark exercise)."""
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


def _pose12(p: "gtsam.Pose3") -> torch.Tensor:
    m = p.matrix()
    return torch.from_numpy(np.concatenate([m[:3, :3].reshape(9), m[:3, 3]])).cuda()[None]


def _vec(v) -> torch.Tensor:
    return torch.from_numpy(np.asarray(v, dtype=np.float64)).cuda()[None]


class PoseDynamicsFactor(gtsam.CustomFactor):
    def __init__(self, pose1: int, ang_vel1: int, vel1: int, pose2: int, noise_model, dt: float,
                 vel_frame: str = "world") -> None:
        assert vel_frame in ["world", "body"], "vel_frame must be 'world' or 'body'."
        super().__init__(noise_model, [pose1, ang_vel1, vel1, pose2],
                         partial(self.error_func, dt=dt, vel_frame=vel_frame))
        self.dt = dt
        self.vel_frame = vel_frame

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None, dt=None, vel_frame=None) -> np.ndarray:
        k = this.keys()
        err, Hs = F.pose_dynamics_batch(_pose12(v.atPose3(k[0])), _vec(v.atVector(k[1])), _vec(v.atVector(k[2])),
                                        _pose12(v.atPose3(k[3])), dt, vel_frame, jacobians=bool(H))
        if H:
            for i in range(4):
                H[i] = np.asfortranarray(Hs[i][0].cpu().numpy())
        return err[0].cpu().numpy()


class ConstantVelocityFactor(gtsam.CustomFactor):
    def __init__(self, vel1: int, vel2: int, noise_model) -> None:
        super().__init__(noise_model, [vel1, vel2], self.error_func)

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None) -> np.ndarray:
        k = this.keys()
        err, Hs = F.constant_velocity_batch(_vec(v.atVector(k[0])), _vec(v.atVector(k[1])), jacobians=bool(H))
        if H:
            H[0] = np.asfortranarray(Hs[0][0].cpu().numpy())
            H[1] = np.asfortranarray(Hs[1][0].cpu().numpy())
        return err[0].cpu().numpy()


class KeypointProjectionFactor(gtsam.CustomFactor):
    def __init__(self, body_pose: int, noise_model, camera_intrinsics, keypoint_measurement: np.ndarray,
                 point_body_frame: np.ndarray, camera_pose=None) -> None:
        super().__init__(
            noise_model, [body_pose],
            partial(self.error_func, camera_intrinsics=camera_intrinsics, keypoint_measurement=keypoint_measurement,
                    point_body_frame=point_body_frame,
                    camera_pose=camera_pose if camera_pose is not None else gtsam.Pose3()),
        )
        self.pixel = None

    def error_func(self, this, v, H: Optional[List[np.ndarray]] = None, camera_intrinsics=None,
                   keypoint_measurement=None, point_body_frame=None, camera_pose=None) -> np.ndarray:
        c = camera_intrinsics
        K5 = torch.tensor([c.fx(), c.fy(), c.skew(), c.px(), c.py()], dtype=torch.float64).cuda()
        err, J, cheir = F.keypoint_projection_batch(
            _pose12(v.atPose3(this.keys()[0])), K5, _vec(keypoint_measurement)[None], _vec(point_body_frame),
            _pose12(camera_pose)[0], jacobians=bool(H))
        if bool(cheir[0, 0]):
            raise RuntimeError("CheiralityException: point behind the camera")
        if H:
            H[0] = np.asfortranarray(J[0, 0].cpu().numpy())
        e = err[0, 0].cpu().numpy()
        self.pixel = e + np.asarray(keypoint_measurement)
        return e
