"""-m gpu: the float64 factor kernels (through the C ABI) against the numpy oracle.
Tolerances: 1e-9 abs on residuals/Jacobians at oracle-checkable sizes (the reference's own test
tolerance is 1e-6, tests/test_dynamics_factor.py:107); invariants at the BASELINE.json size."""
import numpy as np
import pytest
import torch

from oracle import factors_oracle as fo
from tests.test_oracle_factors import DT, reference_test_draws

pytestmark = pytest.mark.gpu


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _pose12(R, t):
    return _cuda(np.concatenate([R.reshape(-1, 9), t], -1))


@pytest.mark.parametrize("frame", ["world", "body"])
def test_dynamics_reference_test_draws(frame):
    from perseus_b200.smoother import pose_dynamics_batch

    R1, t1, R2, t2, v1, w1, _, _ = reference_test_draws()
    want, H = fo.pose_dynamics_factor(R1, t1, w1, v1, R2, t2, DT, frame)
    got, Hg = pose_dynamics_batch(_pose12(R1, t1), _cuda(w1), _cuda(v1), _pose12(R2, t2), DT, frame)
    assert np.allclose(got.cpu().numpy(), want, atol=1e-9)
    for i in range(4):
        assert np.allclose(Hg[i].cpu().numpy(), H[i], atol=1e-9), f"H[{i}]"
    got2, none = pose_dynamics_batch(_pose12(R1, t1), _cuda(w1), _cuda(v1), _pose12(R2, t2), DT, frame,
                                     jacobians=False)
    assert none is None and torch.equal(got, got2)


@pytest.mark.parametrize("N", [1, 127, 128, 129, 5000])
def test_dynamics_random_batches(N):
    from perseus_b200.smoother import pose_dynamics_batch

    rng = np.random.default_rng(N)
    R1, t1 = fo.pose_exp(rng.standard_normal((N, 6)))
    R2, t2 = fo.pose_exp(rng.standard_normal((N, 6)))
    w, v = rng.standard_normal((N, 3)), rng.standard_normal((N, 3))
    for frame in ("world", "body"):
        want, H = fo.pose_dynamics_factor(R1, t1, w, v, R2, t2, DT, frame)
        got, Hg = pose_dynamics_batch(_pose12(R1, t1), _cuda(w), _cuda(v), _pose12(R2, t2), DT, frame)
        assert np.abs(got.cpu().numpy() - want).max() < 1e-9
        for i in range(4):
            assert np.abs(Hg[i].cpu().numpy() - H[i]).max() < 1e-8, f"H[{i}] N={N} {frame}"


def test_dynamics_empty_and_zero_residual():
    from perseus_b200.smoother import pose_dynamics_batch

    e = torch.empty((0, 12), dtype=torch.float64, device="cuda")
    v = torch.empty((0, 3), dtype=torch.float64, device="cuda")
    r, H = pose_dynamics_batch(e, v, v, e, DT)
    assert r.shape == (0, 6) and H[0].shape == (0, 6, 6)
    # pose2 = exact Euler prediction -> zero residual (size-independent property)
    rng = np.random.default_rng(7)
    N = 10000
    R1, t1 = fo.pose_exp(rng.standard_normal((N, 6)))
    w, vb = rng.standard_normal((N, 3)), rng.standard_normal((N, 3))
    R2, t2 = fo.pose_compose(R1, t1, *fo.pose_exp(DT * np.concatenate([w, vb], -1)))
    r, _ = pose_dynamics_batch(_pose12(R1, t1), _cuda(w), _cuda(vb), _pose12(R2, t2), DT, "body")
    assert r.abs().max().item() < 1e-12


def test_constant_velocity():
    from perseus_b200.smoother import constant_velocity_batch

    rng = np.random.default_rng(0)
    v1, v2 = rng.standard_normal((1001, 3)), rng.standard_normal((1001, 3))
    r, H = constant_velocity_batch(_cuda(v1), _cuda(v2))
    assert np.array_equal(r.cpu().numpy(), v2 - v1)
    assert torch.equal(H[0].cpu(), -torch.eye(3, dtype=torch.float64).expand(1001, 3, 3))
    assert torch.equal(H[1].cpu(), torch.eye(3, dtype=torch.float64).expand(1001, 3, 3))


def _projection_problem(N, seed=0):
    rng = np.random.default_rng(seed)
    xi = rng.standard_normal((N, 6)) * np.array([0.5, 0.5, 0.5, 0.2, 0.2, 0.2])
    Rb, tb = fo.pose_exp(xi)
    tb = tb * np.array([0.3, 0.3, 0.0]) + np.stack([np.zeros(N), np.zeros(N), rng.uniform(0.3, 0.6, N)], -1)
    corners = 0.0175 * np.array(
        [[-1, -1, -1], [-1, -1, 1], [-1, 1, -1], [-1, 1, 1], [1, -1, -1], [1, -1, 1], [1, 1, -1], [1, 1, 1]], dtype=float
    )  # generate_and_label_keypoints.py:187 order
    f = 256 / (2 * np.tan(np.deg2rad(40) / 2))  # data_utils.py:53-57 convention
    K5 = np.array([f, f, 0.0, 128.0, 128.0])
    e0, _, _ = fo.keypoint_projection_factor(Rb, tb, K5, np.zeros((N, 8, 2)), corners, jacobians=False)
    meas = e0 + rng.standard_normal((N, 8, 2))
    return Rb, tb, K5, meas, corners


@pytest.mark.parametrize("N", [1, 15, 16, 17, 3000])
def test_projection_matches_oracle(N):
    from perseus_b200.smoother import keypoint_projection_batch

    Rb, tb, K5, meas, corners = _projection_problem(N, seed=N)
    Rc, tc = fo.pose_exp(np.array([[0.05, -0.02, 0.01, 0.01, 0.0, -0.02]]))
    for cam in (None, (Rc[0], tc[0])):
        want, H, cheir = fo.keypoint_projection_factor(Rb, tb, K5, meas, corners, *(cam or (None, None)))
        got, J, ch = keypoint_projection_batch(
            _pose12(Rb, tb), _cuda(K5), _cuda(meas), _cuda(corners),
            None if cam is None else _pose12(cam[0][None], cam[1][None])[0],
        )
        assert np.abs(got.cpu().numpy() - want).max() < 1e-9
        assert np.abs(J.cpu().numpy() - H).max() < 1e-8
        assert np.array_equal(ch.cpu().numpy(), cheir)
        got2, none, _ = keypoint_projection_batch(_pose12(Rb, tb), _cuda(K5), _cuda(meas), _cuda(corners),
                                                  None if cam is None else _pose12(cam[0][None], cam[1][None])[0],
                                                  jacobians=False)
        assert none is None and torch.equal(got, got2)


def test_projection_cheirality_and_full_size_linearity():
    from perseus_b200.smoother import keypoint_projection_batch

    pose = torch.tensor([[1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, -1.0]], dtype=torch.float64, device="cuda")
    _, _, ch = keypoint_projection_batch(pose, torch.tensor([300.0, 300, 0, 128, 128]).cuda(),
                                         torch.zeros(1, 1, 2, dtype=torch.float64, device="cuda"),
                                         torch.zeros(1, 3, dtype=torch.float64, device="cuda"))
    assert bool(ch.all())
    # BASELINE.json size (10k poses x 8): the residual is affine in the measurement
    N = 10000
    Rb, tb, K5, meas, corners = _projection_problem(N, seed=1)
    p, k, c = _pose12(Rb, tb), _cuda(K5), _cuda(corners)
    r0, _, _ = keypoint_projection_batch(p, k, _cuda(meas), c, jacobians=False)
    r1, _, _ = keypoint_projection_batch(p, k, _cuda(meas + 3.0), c, jacobians=False)
    assert (r0 - r1 - 3.0).abs().max().item() < 1e-9


@pytest.mark.parametrize("W,T,frame", [(1, 1, "world"), (300, 2, "world"), (7, 19, "body"), (1, 1000, "world"), (129, 3, "world")])
def test_fused_chain_launch_equals_the_three_per_kind_launches(W, T, frame):
    """pb200_factors_chain (ONE launch: projection blocks + dynamics blocks that also emit the constant-velocity
    residuals, pose1 / pose2 read from the one pose array) vs the three separate kernels: equal to 1e-12 (the same
    inlined arithmetic compiled in two kernels may contract different multiply-adds), and to the oracle to 1e-9."""
    from perseus_b200.smoother.linearize import chain_factors

    rng = np.random.default_rng(W * 1000 + T)
    P = 8
    R, t = fo.pose_exp(rng.standard_normal((W * T, 6)) * np.array([0.5, 0.5, 0.5, 0.05, 0.05, 0.05]))
    t[:, 2] += 0.45
    poses = _pose12(R, t).reshape(W, T, 12)
    w, v = _cuda(rng.standard_normal((W, T, 3))), _cuda(rng.standard_normal((W, T, 3)))
    meas = _cuda(rng.uniform(0, 255, (W, T, P, 2)))
    pts = _cuda(np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=np.float64) * 0.0175)
    K5 = _cuda(np.array([221.7, 221.7, 0.5, 128.0, 128.0]))
    a = chain_factors(poses, w, v, meas, pts, K5, 0.1, None, frame, fused=True)
    b = chain_factors(poses, w, v, meas, pts, K5, 0.1, None, frame, fused=False)
    close = lambda x, y: torch.allclose(x, y, rtol=1e-12, atol=1e-12)  # noqa: E731
    assert close(a["proj_r"], b["proj_r"]) and close(a["proj_J"], b["proj_J"])
    assert torch.equal(a["cheirality"], b["cheirality"])
    if T > 1:
        assert close(a["dyn_r"], b["dyn_r"])
        for i in range(4):
            assert close(a["H"][i], b["H"][i]), f"H[{i}]"
        assert torch.equal(a["cvw_r"], b["cvw_r"]) and torch.equal(a["cvv_r"], b["cvv_r"])
        want, _ = fo.pose_dynamics_factor(R.reshape(W, T, 3, 3)[:, :-1].reshape(-1, 3, 3), t.reshape(W, T, 3)[:, :-1].reshape(-1, 3),
                                          w.cpu().numpy()[:, :-1].reshape(-1, 3), v.cpu().numpy()[:, :-1].reshape(-1, 3),
                                          R.reshape(W, T, 3, 3)[:, 1:].reshape(-1, 3, 3), t.reshape(W, T, 3)[:, 1:].reshape(-1, 3),
                                          0.1, frame, jacobians=False)
        assert np.allclose(a["dyn_r"].cpu().numpy(), want, atol=1e-9)
    # re-using the output buffers (no allocation) gives the same result
    c = chain_factors(poses, w, v, meas, pts, K5, 0.1, None, frame, fused=True, out=a)
    assert c["proj_r"].data_ptr() == a["proj_r"].data_ptr() and close(c["proj_r"], b["proj_r"])
