"""not gpu: the reference's CustomFactor constructors/attributes through the stub gtsam; compute fails loudly without a GPU."""
import os
import sys

import numpy as np
import pytest
import torch

STUB = os.path.join(os.path.dirname(__file__), "stub_gtsam")


def test_factor_classes_mirror_reference_constructors_and_refuse_cpu():
    sys.path.insert(0, STUB)
    try:
        for m in ("gtsam", "perseus_b200.smoother._gtsam_factors"):
            sys.modules.pop(m, None)
        import gtsam

        from perseus_b200.smoother import factors as F

        n6 = gtsam.noiseModel.Diagonal.Sigmas(np.ones(6))
        f = F.PoseDynamicsFactor(1, 2, 3, 4, n6, 0.1)  # reference factors.py:11-20 signature, vel_frame default "world"
        assert f.keys() == [1, 2, 3, 4] and f.dt == 0.1 and f.vel_frame == "world"
        c = F.ConstantVelocityFactor(3, 7, gtsam.noiseModel.Diagonal.Sigmas(np.ones(3)))  # factors.py:151-156
        assert c.keys() == [3, 7]
        p = F.KeypointProjectionFactor(1, gtsam.noiseModel.Diagonal.Sigmas(np.ones(2)), gtsam.Cal3_S2(200, 200, 0, 128, 128),
                                       np.array([10.0, 20.0]), np.array([0.01, 0.0, 0.0]))  # factors.py:182-190
        assert p.keys() == [1] and p.pixel is None
        # factors sharing constants share one evaluation group (the per-linearisation batch)
        f2 = F.PoseDynamicsFactor(4, 5, 6, 7, n6, 0.1)
        assert f2._group is f._group and f2._idx == 1
        assert F.PoseDynamicsFactor(4, 5, 6, 7, n6, 0.2)._group is not f._group
        if not torch.cuda.is_available():
            v = gtsam.Values()
            v.insert(3, np.zeros(3))
            v.insert(7, np.ones(3))
            with pytest.raises(RuntimeError, match="no CPU fallback"):
                c.unwhitenedError(v, None)
    finally:
        sys.path.remove(STUB)
        for m in ("gtsam", "perseus_b200.smoother._gtsam_factors"):
            sys.modules.pop(m, None)
