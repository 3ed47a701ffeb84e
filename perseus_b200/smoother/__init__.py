from .factors import (  # noqa: F401
    constant_velocity_batch,
    keypoint_projection_batch,
    pose_dynamics_batch,
)
