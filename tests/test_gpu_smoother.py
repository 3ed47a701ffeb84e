"""Parity of the smoother linearisation kernels (csrc/smoother.cu via the C ABI) against oracle/smoother_oracle.py's
dense whitened Gauss-Newton system, + a multi-window Gauss-Newton run.  float64 throughout."""
import numpy as np
import pytest
import torch

from oracle import smoother_oracle as so
from perseus_b200.smoother import ChainSigmas, chain_solve, gauss_newton_step, pose_retract
from perseus_b200.smoother.factors import pose12

pytestmark = pytest.mark.gpu
SIG = dict(sig_px=1.0, sig_dyn=[0.1] * 6, sig_cvw=[0.1] * 3, sig_cvv=[0.1] * 3)


def _chains(W, T, seed0=0, perturb=True):
    cs, st = [], []
    for w in range(W):
        c = so.synthetic_chain(T, seed=seed0 + w)
        rng = np.random.default_rng(1000 + seed0 + w)
        d = np.concatenate([0.05 * rng.standard_normal((T, 3)), 0.005 * rng.standard_normal((T, 3)),
                            0.1 * rng.standard_normal((T, 3)), 0.02 * rng.standard_normal((T, 3))], 1) * float(perturb)
        cs.append(c)
        st.append(so.retract(c["R"], c["t"], c["ang_vel"], c["vel"], d))
    return cs, st


def _to_dev(cs, st):
    dev = torch.device("cuda", 0)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    poses = torch.stack([pose12(f(s[0]), f(s[1])) for s in st])
    return (poses, torch.stack([f(s[2]) for s in st]), torch.stack([f(s[3]) for s in st]),
            torch.stack([f(c["meas"]) for c in cs]), f(cs[0]["pts_body"]), f(cs[0]["K5"]))


@pytest.mark.parametrize("W,T,frame,lam,prior", [(3, 6, "world", 0.0, 0.0), (2, 1, "world", 1e-3, 0.0),
                                                 (4, 12, "body", 1e-4, 2.0), (1, 40, "world", 0.0, 0.5)])
def test_normal_equations_and_solve_match_dense_oracle(W, T, frame, lam, prior):
    cs, st = _chains(W, T)
    poses, w_, v_, meas, pts, K5 = _to_dev(cs, st)
    _, _, _, info = gauss_newton_step(poses, w_, v_, meas, pts, K5, cs[0]["dt"], ChainSigmas(), None, frame, lam, prior)
    torch.cuda.synchronize()
    assert int(info["status"].abs().sum()) == 0
    for w in range(W):
        R, t, av, vel = st[w]
        delta, H, g, cost = so.chain_step(R, t, av, vel, cs[w]["meas"], cs[w]["pts_body"], cs[w]["K5"], cs[w]["dt"], **SIG,
                                          vel_frame=frame, lam=lam, prior=prior)
        D, O, gg = info["D"][w].cpu().numpy(), info["O"][w].cpu().numpy(), info["g"][w].cpu().numpy()
        scale = np.abs(H).max()
        for k in range(T):
            assert np.abs(D[k] - H[12 * k:12 * k + 12, 12 * k:12 * k + 12]).max() < 1e-10 * scale
            if k + 1 < T:
                assert np.abs(O[k] - H[12 * k + 12:12 * k + 24, 12 * k:12 * k + 12]).max() < 1e-10 * scale
        assert np.abs(gg.reshape(-1) - g).max() < 1e-10 * np.abs(g).max()
        got = info["delta"][w].cpu().numpy().reshape(-1)
        assert np.abs(got - delta).max() < 1e-7 * max(1.0, np.abs(delta).max())
    assert abs(float(info["cost"]) - sum(so.chain_dense_system(*st[w], cs[w]["meas"], cs[w]["pts_body"], cs[w]["K5"],
                                                              cs[w]["dt"], **SIG, vel_frame=frame)[2] for w in range(W))) < 1e-6 * float(info["cost"])


def test_solver_on_random_spd_block_tridiagonal_systems():
    """chain_solve alone against numpy on random well-conditioned block-tridiagonal SPD systems; a non-SPD system is
    flagged, not silently 'solved'."""
    rng = np.random.default_rng(0)
    W, T = 5, 9
    n = 12 * T
    Ds, Os, gs, xs = [], [], [], []
    for _ in range(W):
        A = np.zeros((n + 12, n))
        for k in range(T):  # rows touching blocks k and k+1 -> A^T A is block tridiagonal
            A[12 * k:12 * k + 24, 12 * k:12 * k + 12] += rng.standard_normal((24, 12))
        H = A.T @ A + 0.5 * np.eye(n)
        g = rng.standard_normal(n)
        Ds.append(np.stack([H[12 * k:12 * k + 12, 12 * k:12 * k + 12] for k in range(T)]))
        Os.append(np.stack([H[12 * k + 12:12 * k + 24, 12 * k:12 * k + 12] for k in range(T - 1)]))
        gs.append(g.reshape(T, 12))
        xs.append(np.linalg.solve(H, -g))
    dev = torch.device("cuda", 0)
    D, O, g = (torch.from_numpy(np.stack(a)).to(dev) for a in (Ds, Os, gs))
    delta, status = chain_solve(D, O, g)
    assert int(status.abs().sum()) == 0
    assert np.abs(delta.cpu().numpy().reshape(W, -1) - np.stack(xs)).max() < 1e-9
    D[2, 3] = -D[2, 3]
    _, status = chain_solve(D, O, g)
    assert status.cpu().tolist() == [0, 0, 1, 0, 0]


def test_gauss_newton_over_many_windows_converges():
    W, T = 64, 10
    cs, st = _chains(W, T, seed0=50)
    poses, w_, v_, meas, pts, K5 = _to_dev(cs, st)
    costs = []
    for _ in range(5):
        poses, w_, v_, info = gauss_newton_step(poses, w_, v_, meas, pts, K5, cs[0]["dt"], lam=1e-6)
        costs.append(float(info["cost"]))
        assert int(info["status"].abs().sum()) == 0
    assert costs[-1] < 0.15 * costs[0] and costs[-1] <= costs[-2] * (1 + 1e-9)
    truth_t = np.stack([c["t"] for c in cs])
    assert np.abs(poses[..., 9:11].cpu().numpy() - truth_t[..., :2]).max() < 0.01
    # retraction keeps rotations orthonormal
    R = poses[..., :9].reshape(W, T, 3, 3)
    assert float((R @ R.transpose(-1, -2) - torch.eye(3, dtype=torch.float64, device=R.device)).abs().max()) < 1e-12
    z = torch.zeros(W, T, 6, dtype=torch.float64, device=poses.device)
    assert torch.equal(pose_retract(poses, z), poses)
