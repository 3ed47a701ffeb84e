"""Pins the float64 factor oracle (oracle/factors_oracle.py) — CPU only.

Mirrors the reference's tests/test_dynamics_factor.py: same numpy seed-0 draws (:11-21), dt (:32),
atol 1e-6 (:107,122,147), both velocity frames, and its independent definition `pypose_error`
(:57-71) restated with torch float64 + autograd (pypose is not installed here).  The projection
factor has no reference test; it is checked by central finite differences.
"""
import numpy as np
import pytest
import torch

from oracle import factors_oracle as fo

ATOL = 1e-6  # the reference tests' tolerance
DT = 1e-1  # tests/test_dynamics_factor.py:32


def reference_test_draws():
    """tests/test_dynamics_factor.py:11-21 — np.random.seed(0), draws in the same order."""
    np.random.seed(0)
    xi1 = np.random.randn(6)  # Pose3().expmap(xi): [rot; trans]
    xi2 = np.random.randn(6)
    v1 = np.random.randn(3)
    w1 = np.random.randn(3)
    v2 = np.random.randn(3)
    w2 = np.random.randn(3)
    R1, t1 = fo.pose_exp(xi1[None])
    R2, t2 = fo.pose_exp(xi2[None])
    return R1, t1, R2, t2, v1[None], w1[None], v2[None], w2[None]


# ---- torch restatement of the reference test's independent definition -------------------------
def _t_skew(w):
    z = torch.zeros((), dtype=w.dtype)
    return torch.stack(
        [torch.stack([z, -w[2], w[1]]), torch.stack([w[2], z, -w[0]]), torch.stack([-w[1], w[0], z])]
    )


def _t_se3_exp(xi_rot_first):
    """Exp of [omega; rho] -> 4x4 (matrix exponential; independent of the closed forms)."""
    w, v = xi_rot_first[:3], xi_rot_first[3:]
    M = torch.zeros(4, 4, dtype=xi_rot_first.dtype)
    M[:3, :3] = _t_skew(w)
    M[:3, 3] = v
    return torch.linalg.matrix_exp(M)


def _t_se3_log(T):
    """Log of a 4x4 -> [omega; u] via series-free closed form on torch ops (autograd-able)."""
    R, t = T[:3, :3], T[:3, 3]
    cos = ((torch.trace(R) - 1.0) / 2.0).clamp(-1 + 1e-15, 1 - 1e-15)
    th = torch.acos(cos)
    w = th / (2.0 * torch.sin(th)) * torch.stack([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    W = _t_skew(w)
    Vinv = (
        torch.eye(3, dtype=T.dtype)
        - 0.5 * W
        + (1.0 / th**2 - (1.0 + torch.cos(th)) / (2.0 * th * torch.sin(th))) * (W @ W)
    )
    return torch.cat([w, Vinv @ t])


def pypose_error_restated(T0, w0, v0, T1, dx0, dw0, dv0, dx1, vel_frame):
    """tests/test_dynamics_factor.py:57-71 in GTSAM tangent order (the test flips to/from pypose's)."""
    T0p = T0 @ _t_se3_exp(dx0)
    v0p = v0 + dv0
    w0p = w0 + dw0
    T1p = T1 @ _t_se3_exp(dx1)
    if vel_frame == "world":
        v0p = T0p[:3, :3].T @ v0p
    pred = T0p @ _t_se3_exp(DT * torch.cat([w0p, v0p]))
    return _t_se3_log(torch.linalg.inv(pred) @ T1p)


def _mat4(R, t):
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = t
    return torch.tensor(T, dtype=torch.float64)


@pytest.mark.parametrize("frame", ["world", "body"])
def test_dynamics_outputs_match_reference_test_definition(frame):
    R1, t1, R2, t2, v1, w1, _, _ = reference_test_draws()
    err, _ = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, frame, jacobians=False)
    z6, z3 = torch.zeros(6, dtype=torch.float64), torch.zeros(3, dtype=torch.float64)
    ref = pypose_error_restated(
        _mat4(R1[0], t1[0]), torch.tensor(w1[0]), torch.tensor(v1[0]), _mat4(R2[0], t2[0]), z6, z3, z3, z6, frame
    )
    assert np.allclose(err[0], ref.numpy(), atol=ATOL)
    err_j, _ = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, frame, jacobians=True)
    assert np.allclose(err_j, err, atol=1e-12)


@pytest.mark.parametrize("frame", ["world", "body"])
def test_dynamics_jacobians_match_autograd(frame):
    R1, t1, R2, t2, v1, w1, _, _ = reference_test_draws()
    _, H = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, frame)
    T0, T1 = _mat4(R1[0], t1[0]), _mat4(R2[0], t2[0])
    z6, z3 = torch.zeros(6, dtype=torch.float64), torch.zeros(3, dtype=torch.float64)
    f = lambda dx0, dw0, dv0, dx1: pypose_error_restated(  # noqa: E731
        T0, torch.tensor(w1[0]), torch.tensor(v1[0]), T1, dx0, dw0, dv0, dx1, frame
    )
    jacs = torch.autograd.functional.jacobian(f, (z6, z3, z3, z6))
    for i in range(4):
        assert np.allclose(H[i][0], jacs[i].numpy(), atol=ATOL), f"H[{i}] {frame}"


def _fd_dynamics(R1, t1, w1, v1, R2, t2, frame, h=1e-6):
    def f(d0, dw, dv, d1):
        Ra, ta = fo.pose_compose(R1, t1, *fo.pose_exp(d0[None]))
        Rb, tb = fo.pose_compose(R2, t2, *fo.pose_exp(d1[None]))
        e, _ = fo.pose_dynamics_factor(Ra, ta, w1 + dw, v1 + dv, Rb, tb, DT, frame, jacobians=False)
        return e[0]

    sizes = [6, 3, 3, 6]
    out = []
    for a, n in enumerate(sizes):
        J = np.zeros((6, n))
        for k in range(n):
            args_p = [np.zeros(s) for s in sizes]
            args_m = [np.zeros(s) for s in sizes]
            args_p[a][k] = h
            args_m[a][k] = -h
            J[:, k] = (f(*args_p) - f(*args_m)) / (2 * h)
        out.append(J)
    return out


@pytest.mark.parametrize("frame", ["world", "body"])
def test_dynamics_jacobians_finite_differences_random(frame):
    rng = np.random.default_rng(1)
    for _ in range(5):
        R1, t1 = fo.pose_exp(rng.standard_normal((1, 6)))
        R2, t2 = fo.pose_exp(rng.standard_normal((1, 6)) * 0.7)
        w1, v1 = rng.standard_normal((1, 3)), rng.standard_normal((1, 3))
        _, H = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, frame)
        fd = _fd_dynamics(R1, t1, w1, v1, R2, t2, frame)
        for i in range(4):
            assert np.allclose(H[i][0], fd[i], atol=1e-6)


def test_dynamics_small_angle_branch():
    """omega*dt below the Taylor switch and rel pose at identity must stay finite and correct."""
    R1, t1 = fo.pose_exp(np.array([[0.3, -0.2, 0.1, 1.0, 2.0, 3.0]]))
    w1 = np.zeros((1, 3))
    v1 = np.array([[0.5, -0.25, 0.125]])
    # pose2 = exact prediction -> zero residual
    vb = R1[0].T @ v1[0]
    Ri, ti = fo.pose_exp(DT * np.concatenate([w1[0], vb])[None])
    R2, t2 = fo.pose_compose(R1, t1, Ri, ti)
    err, H = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, "world")
    assert np.allclose(err, 0.0, atol=1e-12)
    assert all(np.isfinite(h).all() for h in H)
    # GTSAM's Pose3::Expmap drops the O(theta) translation terms below theta^2 = 1e-10, so a
    # finite difference has to step outside that branch (omega*dt = 1e-4) to see the slope.
    fd = _fd_dynamics(R1, t1, w1, v1, R2, t2, "world", h=1e-3)
    for i in range(4):
        assert np.allclose(H[i][0], fd[i], atol=1e-6)


def test_constant_velocity():
    _, _, _, _, v1, _, v2, _ = reference_test_draws()
    err, H = fo.constant_velocity_factor(v1, v2)
    assert np.allclose(err, v2 - v1)
    assert np.allclose(H[0][0], -np.eye(3)) and np.allclose(H[1][0], np.eye(3))


def test_projection_finite_differences():
    rng = np.random.default_rng(2)
    N, P = 4, 8
    xi = rng.standard_normal((N, 6)) * np.array([0.5, 0.5, 0.5, 0.05, 0.05, 0.05])
    Rb, tb = fo.pose_exp(xi)
    tb = tb + np.array([0.0, 0.0, 0.45])
    Rc, tc = fo.pose_exp(np.array([[0.1, -0.05, 0.02, 0.01, 0.02, -0.03]]))
    Rc, tc = Rc[0], tc[0]
    K5 = (300.0, 310.0, 0.5, 128.0, 127.0)
    corners = 0.0175 * np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=float)
    meas = rng.uniform(0, 256, (N, P, 2))
    err, H, cheir = fo.keypoint_projection_factor(Rb, tb, K5, meas, corners, Rc, tc)
    assert not cheir.any()
    h = 1e-6
    for k in range(6):
        d = np.zeros((N, 6))
        d[:, k] = h
        Rp, tp = fo.pose_compose(Rb, tb, *fo.pose_exp(d))
        Rm, tm = fo.pose_compose(Rb, tb, *fo.pose_exp(-d))
        ep, _, _ = fo.keypoint_projection_factor(Rp, tp, K5, meas, corners, Rc, tc, jacobians=False)
        em, _, _ = fo.keypoint_projection_factor(Rm, tm, K5, meas, corners, Rc, tc, jacobians=False)
        assert np.allclose(H[..., k], (ep - em) / (2 * h), atol=1e-5)
    # identity camera default
    e0, H0, _ = fo.keypoint_projection_factor(Rb, tb, K5, meas, corners)
    e1, H1, _ = fo.keypoint_projection_factor(Rb, tb, K5, meas, corners, np.eye(3), np.zeros(3))
    assert np.array_equal(e0, e1) and np.array_equal(H0, H1)


def test_projection_cheirality_flag():
    Rb, tb = np.eye(3)[None], np.array([[0.0, 0.0, -1.0]])
    _, _, cheir = fo.keypoint_projection_factor(Rb, tb, (300, 300, 0, 128, 128), np.zeros((1, 1, 2)), np.zeros((1, 3)))
    assert cheir.all()


# ---- projection factor: an independent, autograd-defined statement -------------------------------------------------
def pinhole_restated(Tb, Tc, K3, p_body, dx):
    """The reference's OTHER pinhole code path, data_generation/data_utils.py:17-66 (`get_pixel_coordinates`:
    camera_to_object = camera_pose^-1 @ object_pose, then pypose `point2pixel` = K (T p) / z), restated with 4x4
    matrices in torch float64 (pypose is absent).  It shares no formula with oracle/factors_oracle.py: homogeneous
    matrix products, a matrix exponential for the right perturbation T_body Exp(dx) (tangent [rot; trans]), and the
    3x3 intrinsics matrix instead of Cal3_S2's closed form.  Jacobians come from autograd."""
    cam_to_obj = torch.linalg.inv(Tc) @ (Tb @ _t_se3_exp(dx))
    q = cam_to_obj[:3, :3] @ p_body + cam_to_obj[:3, 3]
    uvw = K3 @ q
    return uvw[:2] / uvw[2]


@pytest.mark.parametrize("skew", [0.0, 3.5])
@pytest.mark.parametrize("with_camera_pose", [False, True])
def test_projection_matches_autograd_defined_pinhole(skew, with_camera_pose):
    rng = np.random.default_rng(5)
    N, P = 6, 8
    xi = rng.standard_normal((N, 6)) * np.array([0.6, 0.6, 0.6, 0.05, 0.05, 0.05])
    Rb, tb = fo.pose_exp(xi)
    tb[:, 2] += 0.45
    if with_camera_pose:
        Rc, tc = fo.pose_exp(np.array([[0.2, -0.1, 0.3, 0.02, -0.03, 0.05]]))
        Rc, tc = Rc[0], tc[0]
    else:
        Rc, tc = np.eye(3), np.zeros(3)
    K5 = np.array([221.7, 230.1, skew, 128.0, 120.0])
    K3 = torch.tensor([[K5[0], K5[2], K5[3]], [0.0, K5[1], K5[4]], [0.0, 0.0, 1.0]], dtype=torch.float64)
    corners = np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=np.float64) * 0.0175
    meas = rng.uniform(0, 255, (N, P, 2))
    err, H, cheir = fo.keypoint_projection_factor(Rb, tb, K5, meas, corners, Rc if with_camera_pose else None,
                                                  tc if with_camera_pose else None)
    assert not cheir.any()
    Tc = _mat4(Rc, tc)
    z6 = torch.zeros(6, dtype=torch.float64)
    for n in range(N):
        Tb = _mat4(Rb[n], tb[n])
        for p in range(P):
            pb = torch.tensor(corners[p])
            pix = pinhole_restated(Tb, Tc, K3, pb, z6)
            assert np.allclose(err[n, p], pix.numpy() - meas[n, p], atol=1e-9)
            J = torch.autograd.functional.jacobian(lambda d: pinhole_restated(Tb, Tc, K3, pb, d), z6)
            assert np.allclose(H[n, p], J.numpy(), atol=1e-8), (n, p)
