"""-m gpu: the GTSAM-side boundary -- the reference's CustomFactor classes with unchanged
constructors, driven through `error_func(this, v, H)` the way GTSAM's optimiser calls them (factors.py:54-61,160,
216-225): one factor at a time, one Values per linearisation, `H[i] = ...` assigned in Fortran order, float64, tangent
order [rot; trans].  gtsam is not installable here; tests/stub_gtsam supplies the container classes (no geometry)."""
import os
import sys

import numpy as np
import pytest

from oracle import factors_oracle as fo

pytestmark = pytest.mark.gpu

STUB = os.path.join(os.path.dirname(__file__), "stub_gtsam")


@pytest.fixture()
def gf():
    sys.path.insert(0, STUB)
    for m in ("gtsam", "perseus_b200.smoother._gtsam_factors"):
        sys.modules.pop(m, None)
    import gtsam  # noqa: F401  (the stub)

    from perseus_b200.smoother import _gtsam_factors as mod

    mod.reset_groups()
    yield mod
    sys.path.remove(STUB)
    for m in ("gtsam", "perseus_b200.smoother._gtsam_factors"):
        sys.modules.pop(m, None)


def _scene(T=12, seed=0):
    rng = np.random.default_rng(seed)
    xi = rng.standard_normal((T, 6)) * np.array([0.4, 0.4, 0.4, 0.04, 0.04, 0.04])
    R, t = fo.pose_exp(xi)
    t[:, 2] += 0.45
    w, v = rng.standard_normal((T, 3)), rng.standard_normal((T, 3)) * 0.1
    corners = np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=np.float64) * 0.0175
    K5 = np.array([221.7, 221.7, 0.0, 128.0, 128.0])
    meas = rng.uniform(0, 255, (T, 8, 2))
    return R, t, w, v, corners, K5, meas


def _values(gtsam, R, t, w, v):
    vals = gtsam.Values()
    for k in range(R.shape[0]):
        vals.insert(1000 + k, gtsam.Pose3(gtsam.Rot3(R[k]), t[k]))
        vals.insert(2000 + k, w[k])
        vals.insert(3000 + k, v[k])
    return vals


def _call(f, vals, dims):
    """what GTSAM does: error only (H = None), then error + Jacobians into a presized list"""
    e0 = f.unwhitenedError(vals, None)
    H = [np.zeros((e0.shape[0], d), order="F") for d in dims]
    e1 = f.unwhitenedError(vals, H)
    assert np.array_equal(e0, e1) and e1.dtype == np.float64
    for h, d in zip(H, dims):
        assert h.dtype == np.float64 and h.shape == (e0.shape[0], d) and h.flags.f_contiguous, "GTSAM reads column-major"
    return e1, H


@pytest.mark.parametrize("frame", ["world", "body"])
def test_custom_factor_callbacks_match_the_oracle_and_batch_per_linearisation(gf, frame):
    import gtsam

    T, dt = 12, 0.1
    R, t, w, v, corners, K5, meas = _scene(T)
    vals = _values(gtsam, R, t, w, v)
    noise6 = gtsam.noiseModel.Diagonal.Sigmas(np.full(6, 0.1))
    cal = gtsam.Cal3_S2(*K5)
    dyn = [gf.PoseDynamicsFactor(1000 + k, 2000 + k, 3000 + k, 1001 + k, noise6, dt, frame) for k in range(T - 1)]
    cvs = [gf.ConstantVelocityFactor(3000 + k, 3001 + k, gtsam.noiseModel.Diagonal.Sigmas(np.full(3, 0.1))) for k in range(T - 1)]
    proj = [[gf.KeypointProjectionFactor(1000 + k, gtsam.noiseModel.Diagonal.Sigmas(np.ones(2)), cal, meas[k, p], corners[p])
             for p in range(8)] for k in range(T)]
    assert dyn[0].keys() == [1000, 2000, 3000, 1001] and dyn[0].dt == dt and dyn[0].vel_frame == frame  # factors.py:48-52
    with pytest.raises(AssertionError):
        gf.PoseDynamicsFactor(1, 2, 3, 4, noise6, dt, "camera")  # factors.py:40

    want_e, want_H = fo.pose_dynamics_factor(R[:-1], t[:-1], w[:-1], v[:-1], R[1:], t[1:], dt, frame)
    for k, f in enumerate(dyn):
        e, H = _call(f, vals, (6, 3, 3, 6))
        assert np.allclose(e, want_e[k], atol=1e-9)
        for i in range(4):
            assert np.allclose(H[i], want_H[i][k], atol=1e-8), f"dynamics H[{i}] of factor {k}"
    assert dyn[0]._group.evaluations == 1, "one batched launch per linearisation, not one per callback"

    for k, f in enumerate(cvs):
        e, H = _call(f, vals, (3, 3))
        assert np.array_equal(e, v[k + 1] - v[k]) and np.array_equal(H[0], -np.eye(3)) and np.array_equal(H[1], np.eye(3))
    assert cvs[0]._group.evaluations == 1

    pe, pH, cheir = fo.keypoint_projection_factor(R, t, K5, meas, corners)
    assert not cheir.any()
    for k in range(T):
        for p in range(8):
            e, H = _call(proj[k][p], vals, (6,))
            assert np.allclose(e, pe[k, p], atol=1e-9)
            assert np.allclose(H[0], pH[k, p], atol=1e-8)
            assert np.allclose(proj[k][p].pixel, pe[k, p] + meas[k, p], atol=1e-9)  # factors.py:273
    assert proj[0][0]._group.evaluations == 1

    # a new linearisation point: every group re-evaluates exactly once, results follow the new values
    R2, t2 = fo.pose_compose(R, t, *fo.pose_exp(np.full((T, 6), 1e-2)))
    vals2 = _values(gtsam, R2, t2, w, v)
    want_e2, want_H2 = fo.pose_dynamics_factor(R2[:-1], t2[:-1], w[:-1], v[:-1], R2[1:], t2[1:], dt, frame)
    for k, f in enumerate(dyn):
        e, H = _call(f, vals2, (6, 3, 3, 6))
        assert np.allclose(e, want_e2[k], atol=1e-9) and np.allclose(H[0], want_H2[0][k], atol=1e-8)
    assert dyn[0]._group.evaluations == 2
    # going back to the first point is another linearisation, not a stale cache hit
    e, _ = _call(dyn[3], vals, (6, 3, 3, 6))
    assert np.allclose(e, want_e[3], atol=1e-9) and dyn[0]._group.evaluations == 3


def test_sub_graph_values_and_cheirality(gf):
    import gtsam

    T = 6
    R, t, w, v, corners, K5, meas = _scene(T, seed=3)
    cal = gtsam.Cal3_S2(*K5)
    n2 = gtsam.noiseModel.Diagonal.Sigmas(np.ones(2))
    t[2, 2] = -0.5  # pose 2 sits behind the camera
    proj = [gf.KeypointProjectionFactor(1000 + k, n2, cal, meas[k, 0], corners[0]) for k in range(T)]
    sub = gtsam.Values()  # a Values holding only poses 0..3 (sub-graph linearisation)
    for k in range(4):
        sub.insert(1000 + k, gtsam.Pose3(gtsam.Rot3(R[k]), t[k]))
    pe, pH, cheir = fo.keypoint_projection_factor(R, t, K5, meas[:, :1], corners[:1])
    assert cheir[2, 0] and not cheir[[0, 1, 3], 0].any()
    for k in (0, 1, 3):
        e, H = _call(proj[k], sub, (6,))
        assert np.allclose(e, pe[k, 0], atol=1e-9) and np.allclose(H[0], pH[k, 0], atol=1e-8)
    with pytest.raises(RuntimeError, match="Cheirality"):
        proj[2].unwhitenedError(sub, None)
    with pytest.raises(RuntimeError, match="KeyDoesNotExist"):
        proj[5].unwhitenedError(sub, None)
