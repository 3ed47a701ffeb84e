from .factors import (  # noqa: F401
    constant_velocity_batch,
    keypoint_projection_batch,
    pose_dynamics_batch,
)
from .linearize import (  # noqa: F401,E402
    ChainSigmas,
    chain_cost,
    chain_factors,
    chain_solve,
    gauss_newton_step,
    normal_equations,
    pose_retract,
)
from .fixed_lag import FixedLagSmoother  # noqa: F401,E402
