"""CPU oracle for the fixed-lag smoother linearisation.  TEST INFRASTRUCTURE ONLY.

The reference has no smoother solve of its own (it hands perseus/smoother/factors.py's CustomFactors to GTSAM, whose
sources are not under /root/reference) -> PARITY UNPINNED by reference vectors.  What is restated is the textbook
definition GTSAM implements: every factor is whitened by its diagonal noise model (r / sigma, J / sigma; factors.py:42,49
pass noiseModel.Diagonal sigmas), the whitened Jacobian rows are stacked into one dense A, and the Gauss-Newton step
solves (A^T A + lambda I + prior) delta = -A^T r with a dense solver.  Factor values come from oracle/factors_oracle.py
(pinned to the reference's own test draws).
"""
from __future__ import annotations

import numpy as np

from . import factors_oracle as fo


def chain_dense_system(R, t, ang_vel, vel, meas, pts_body, K5, dt, sig_px, sig_dyn, sig_cvw, sig_cvv, vel_frame="world",
                       lam=0.0, prior=0.0):
    """One chain.  R (T,3,3), t (T,3), ang_vel / vel (T,3), meas (T,P,2).  Returns dense (H (12T,12T), g (12T,), cost)."""
    T = R.shape[0]
    n = 12 * T
    rows, rhs = [], []

    def add(blocks, r, sig):
        """blocks: list of (state index, column offset, J (m, c)); r (m,); sig (m,)"""
        m = r.shape[0]
        row = np.zeros((m, n))
        for k, off, J in blocks:
            row[:, 12 * k + off:12 * k + off + J.shape[1]] += J
        rows.append(row / np.asarray(sig)[:, None])
        rhs.append(r / np.asarray(sig))

    pr, pJ, _ = fo.keypoint_projection_factor(R, t, K5, meas, pts_body)
    for k in range(T):
        for p in range(pts_body.shape[0]):
            add([(k, 0, pJ[k, p])], pr[k, p], [sig_px, sig_px])
    if T > 1:
        dr, Hs = fo.pose_dynamics_factor(R[:-1], t[:-1], ang_vel[:-1], vel[:-1], R[1:], t[1:], dt, vel_frame)
        for k in range(T - 1):
            add([(k, 0, Hs[0][k]), (k, 6, Hs[1][k]), (k, 9, Hs[2][k]), (k + 1, 0, Hs[3][k])], dr[k], sig_dyn)
            add([(k, 6, -np.eye(3)), (k + 1, 6, np.eye(3))], ang_vel[k + 1] - ang_vel[k], sig_cvw)
            add([(k, 9, -np.eye(3)), (k + 1, 9, np.eye(3))], vel[k + 1] - vel[k], sig_cvv)
    A = np.concatenate(rows, 0)
    r = np.concatenate(rhs, 0)
    H = A.T @ A + lam * np.eye(n)
    H[:12, :12] += prior * np.eye(12)
    return H, A.T @ r, 0.5 * float(r @ r)


def chain_step(*args, **kw):
    H, g, cost = chain_dense_system(*args, **kw)
    return np.linalg.solve(H, -g), H, g, cost


def retract(R, t, ang_vel, vel, delta):
    """Pose3.retract (= compose with Expmap) + vector addition; delta (T, 12)."""
    Ri, ti = fo.pose_exp(delta[:, :6])
    Rn, tn = fo.pose_compose(R, t, Ri, ti)
    return Rn, tn, ang_vel + delta[:, 6:9], vel + delta[:, 9:12]


def synthetic_chain(T: int, P: int = 8, seed: int = 0, dt: float = 0.1, noise_px: float = 1.0):
    """A cube moving with constant body twist in front of the camera (SURVEY.md §8(d) C5 inputs): ground-truth states,
    noisy pixel measurements, intrinsics."""
    rng = np.random.default_rng(seed)
    corners = 0.0175 * np.array([[-1, -1, -1], [-1, -1, 1], [-1, 1, -1], [-1, 1, 1], [1, -1, -1], [1, -1, 1], [1, 1, -1],
                                 [1, 1, 1]], dtype=np.float64)[:P]
    f = 256 / (2 * np.tan(np.radians(60) / 2))
    K5 = np.array([f, f, 0.0, 128.0, 128.0])
    R0, t0 = fo.pose_exp(np.concatenate([0.5 * rng.standard_normal(3), [0.0, 0.0, 0.0]])[None])
    R, t = [R0[0]], [np.array([0.02 * rng.standard_normal(), 0.02 * rng.standard_normal(), rng.uniform(0.3, 0.6)])]
    w = 0.5 * rng.standard_normal(3)
    v = 0.05 * rng.standard_normal(3)  # world-frame linear velocity
    for _ in range(T - 1):
        vb = R[-1].T @ v
        Ri, ti = fo.pose_exp((dt * np.concatenate([w, vb]))[None])
        Rn, tn = fo.pose_compose(R[-1][None], t[-1][None], Ri, ti)
        R.append(Rn[0])
        t.append(tn[0])
    R, t = np.stack(R), np.stack(t)
    ang_vel, vel = np.tile(w, (T, 1)), np.tile(v, (T, 1))
    pr, _, _ = fo.keypoint_projection_factor(R, t, K5, np.zeros((T, P, 2)), corners, jacobians=False)
    meas = pr + noise_px * rng.standard_normal((T, P, 2))
    return dict(R=R, t=t, ang_vel=ang_vel, vel=vel, meas=meas, pts_body=corners, K5=K5, dt=dt)
