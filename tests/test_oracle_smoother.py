"""CPU checks of oracle/smoother_oracle.py: the dense Gauss-Newton system is block-tridiagonal, its gradient is the
finite-difference gradient of the whitened cost, and iterating it recovers a perturbed synthetic chain."""
import numpy as np

from oracle import factors_oracle as fo
from oracle import smoother_oracle as so

SIG = dict(sig_px=1.0, sig_dyn=[0.1] * 6, sig_cvw=[0.1] * 3, sig_cvv=[0.1] * 3)


def _perturbed(T, seed):
    c = so.synthetic_chain(T, seed=seed)
    rng = np.random.default_rng(seed + 1)
    d = np.concatenate([0.05 * rng.standard_normal((T, 3)), 0.005 * rng.standard_normal((T, 3)),
                        0.1 * rng.standard_normal((T, 3)), 0.02 * rng.standard_normal((T, 3))], 1)
    R, t, w, v = so.retract(c["R"], c["t"], c["ang_vel"], c["vel"], d)
    return c, R, t, w, v


def test_dense_system_structure_and_gradient():
    T = 5
    c, R, t, w, v = _perturbed(T, 0)
    H, g, cost = so.chain_dense_system(R, t, w, v, c["meas"], c["pts_body"], c["K5"], c["dt"], **SIG)
    assert np.allclose(H, H.T)
    for a in range(T):
        for b in range(T):
            if abs(a - b) > 1:
                assert np.all(H[12 * a:12 * a + 12, 12 * b:12 * b + 12] == 0)
    # g = d cost / d delta at 0 (right perturbations)
    eps = 1e-6
    for idx in (0, 4, 7, 10, 12 * 2 + 3, 12 * 4 + 9):
        e = np.zeros((T, 12))
        e.reshape(-1)[idx] = eps
        cp = so.chain_dense_system(*so.retract(R, t, w, v, e), c["meas"], c["pts_body"], c["K5"], c["dt"], **SIG)[2]
        cm = so.chain_dense_system(*so.retract(R, t, w, v, -e), c["meas"], c["pts_body"], c["K5"], c["dt"], **SIG)[2]
        assert abs((cp - cm) / (2 * eps) - g[idx]) < 1e-4 * max(1.0, abs(g[idx]))


def test_gauss_newton_converges_on_synthetic_chain():
    T = 8
    c, R, t, w, v = _perturbed(T, 3)
    costs = []
    for _ in range(6):
        delta, H, g, cost = so.chain_step(R, t, w, v, c["meas"], c["pts_body"], c["K5"], c["dt"], **SIG, lam=1e-6)
        costs.append(cost)
        R, t, w, v = so.retract(R, t, w, v, delta.reshape(T, 12))
    assert costs[-1] < 0.1 * costs[0] and costs[-1] < 0.5 * (T * 16 + (T - 1) * 12) and costs[-1] <= costs[-2] * (1 + 1e-6)
    # pixel noise of 1 px over 8 corners: poses recovered to a few mm / a few degrees
    dR = np.einsum("kji,kjl->kil", c["R"], R)
    ang = np.linalg.norm(fo.so3_log(dR), axis=-1)
    assert np.abs(t - c["t"])[:, :2].max() < 0.005 and np.abs(t - c["t"]).max() < 0.04 and ang.max() < 0.2  # depth is the weak direction
