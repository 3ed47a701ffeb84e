"""CPU oracle for the smoother factors — TEST INFRASTRUCTURE ONLY, never imported by the product path.

Float64 numpy restatement of `perseus/smoother/factors.py` (reference):

  * PoseDynamicsFactor.error_func        factors.py:54-142
  * ConstantVelocityFactor.error_func    factors.py:160-171
  * KeypointProjectionFactor.error_func  factors.py:216-275

The reference delegates its SE(3) / camera arithmetic to the third-party dependency
`gtsam>=4.2` (pyproject.toml:21; a PyPI wheel, C++ source not under /root/reference, not
installable here: no network).  The GTSAM 4.2 algorithms it calls are restated below from
their published definitions (Pose3::Expmap/Logmap/ExpmapDerivative/LogmapDerivative,
AdjointMap, compose/between/transformFrom/transformTo Jacobians, PinholeCamera<Cal3_S2>::project),
with the conventions SURVEY.md Appendix A lists: tangent order [rot; trans], Jacobians with respect
to right perturbations T*Exp(delta).

Pinning:
  * PoseDynamicsFactor / ConstantVelocityFactor: pinned against the reference's own test
    (tests/test_dynamics_factor.py): same seed-0 draws (:11-21), dt (:32), tolerance 1e-6 (:107...),
    and the test's independent definition `pypose_error` (:57-71) which tests/test_oracle_factors.py
    restates with torch float64 autograd in place of pypose (pypose is not installed).
  * KeypointProjectionFactor: the reference has no test for it -> PARITY UNPINNED by reference
    vectors; checked by central finite differences only.

Everything is batched over a leading axis N; pose = (R (N,3,3), t (N,3)).
"""
from __future__ import annotations

import numpy as np

EPS_THETA2 = 1e-10  # small-angle switch (on theta^2), as in GTSAM's so3 / Pose3 code


# ----------------------------------------------------------------------------------------------
# SO(3)
# ----------------------------------------------------------------------------------------------
def skew(w: np.ndarray) -> np.ndarray:
    """[w]x for w (...,3) -> (...,3,3)."""
    z = np.zeros_like(w[..., 0])
    return np.stack(
        [
            np.stack([z, -w[..., 2], w[..., 1]], -1),
            np.stack([w[..., 2], z, -w[..., 0]], -1),
            np.stack([-w[..., 1], w[..., 0], z], -1),
        ],
        -2,
    )


def so3_exp(w: np.ndarray) -> np.ndarray:
    """Rot3::Expmap (Rodrigues).  w (...,3) -> (...,3,3)."""
    th2 = np.sum(w * w, -1)
    th = np.sqrt(th2)
    small = th2 < EPS_THETA2
    ths = np.where(small, 1.0, th)
    a = np.where(small, 1.0 - th2 / 6.0, np.sin(ths) / ths)
    b = np.where(small, 0.5 - th2 / 24.0, (1.0 - np.cos(ths)) / np.where(small, 1.0, th2))
    W = skew(w)
    I = np.broadcast_to(np.eye(3), W.shape)
    return I + a[..., None, None] * W + b[..., None, None] * (W @ W)


def so3_log(R: np.ndarray) -> np.ndarray:
    """Rot3::Logmap with GTSAM 4.2's SO3::Logmap branch structure.  R (...,3,3) -> (...,3).

    Branches: trace near -1 (theta near pi: largest-diagonal formula), the generic
    acos branch, and the Taylor branch for trace near 3.
    """
    R = np.asarray(R, dtype=np.float64)
    tr = R[..., 0, 0] + R[..., 1, 1] + R[..., 2, 2]
    vee = np.stack(
        [R[..., 2, 1] - R[..., 1, 2], R[..., 0, 2] - R[..., 2, 0], R[..., 1, 0] - R[..., 0, 1]], -1
    )
    tr3 = tr - 3.0
    generic = tr3 < -1e-6
    theta = np.arccos(np.clip((tr - 1.0) / 2.0, -1.0, 1.0))
    with np.errstate(divide="ignore", invalid="ignore"):
        mag_generic = theta / (2.0 * np.sin(theta))
    mag_taylor = 0.5 - tr3 / 12.0 + tr3 * tr3 / 60.0
    mag = np.where(generic, mag_generic, mag_taylor)
    out = mag[..., None] * vee
    near_pi = tr + 1.0 < 1e-3
    if np.any(near_pi):
        flat_R = R.reshape(-1, 3, 3)
        flat_out = out.reshape(-1, 3).copy()
        for i in np.flatnonzero(near_pi.reshape(-1)):
            flat_out[i] = _so3_log_near_pi(flat_R[i])
        out = flat_out.reshape(out.shape)
    return out


def _so3_log_near_pi(R: np.ndarray) -> np.ndarray:
    """theta ~ pi: pick the largest diagonal entry a, with (a,b,c) a cyclic permutation."""
    d = [R[0, 0], R[1, 1], R[2, 2]]
    if d[2] > d[1] and d[2] > d[0]:
        a, b, c = 2, 0, 1
    elif d[1] > d[0]:
        a, b, c = 1, 2, 0
    else:
        a, b, c = 0, 1, 2
    W = R[c, b] - R[b, c]
    Q1 = 2.0 + 2.0 * R[a, a]
    Q2 = R[a, b] + R[b, a]
    Q3 = R[c, a] + R[a, c]
    r = np.sqrt(Q1)
    norm = np.sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W)
    sgn = -1.0 if W < 0 else 1.0
    mag = np.pi - (2.0 * sgn * W) / norm
    scale = 0.5 * mag / r
    om = np.zeros(3)
    om[a], om[b], om[c] = Q1, Q2, Q3
    return sgn * scale * om


def so3_right_jacobian(w: np.ndarray) -> np.ndarray:
    """Jr(w) = I - (1-cos)/th^2 W + (th - sin)/th^3 W^2   (SO3::ExpmapDerivative)."""
    th2 = np.sum(w * w, -1)
    th = np.sqrt(th2)
    small = th2 < EPS_THETA2
    ths = np.where(small, 1.0, th)
    th2s = np.where(small, 1.0, th2)
    b = np.where(small, 0.5 - th2 / 24.0, (1.0 - np.cos(ths)) / th2s)
    c = np.where(small, 1.0 / 6.0 - th2 / 120.0, (ths - np.sin(ths)) / (th2s * ths))
    W = skew(w)
    I = np.broadcast_to(np.eye(3), W.shape)
    return I - b[..., None, None] * W + c[..., None, None] * (W @ W)


def so3_right_jacobian_inv(w: np.ndarray) -> np.ndarray:
    """Jr^-1(w) = I + W/2 + (1/th^2 - (1+cos)/(2 th sin)) W^2   (SO3::LogmapDerivative)."""
    th2 = np.sum(w * w, -1)
    th = np.sqrt(th2)
    small = th2 < EPS_THETA2
    ths = np.where(small, 1.0, th)
    th2s = np.where(small, 1.0, th2)
    with np.errstate(divide="ignore", invalid="ignore"):
        c = np.where(
            small,
            1.0 / 12.0 + th2 / 720.0,
            1.0 / th2s - (1.0 + np.cos(ths)) / (2.0 * ths * np.sin(ths)),
        )
    W = skew(w)
    I = np.broadcast_to(np.eye(3), W.shape)
    return I + 0.5 * W + c[..., None, None] * (W @ W)


# ----------------------------------------------------------------------------------------------
# SE(3)  (tangent xi = [omega; rho])
# ----------------------------------------------------------------------------------------------
def pose_exp(xi: np.ndarray):
    """Pose3::Expmap.  xi (...,6) -> (R, t)."""
    w, v = xi[..., :3], xi[..., 3:]
    R = so3_exp(w)
    th2 = np.sum(w * w, -1)
    small = th2 < EPS_THETA2
    th2s = np.where(small, 1.0, th2)
    wxv = np.cross(w, v)
    t_big = (wxv - np.einsum("...ij,...j->...i", R, wxv) + w * np.sum(w * v, -1, keepdims=True)) / th2s[
        ..., None
    ]
    t = np.where(small[..., None], v, t_big)
    return R, t


def pose_log(R: np.ndarray, t: np.ndarray) -> np.ndarray:
    """Pose3::Logmap.  -> (...,6) [omega; u]."""
    w = so3_log(R)
    th2 = np.sum(w * w, -1)
    th = np.sqrt(th2)
    small = th < 1e-10
    ths = np.where(small, 1.0, th)
    W = skew(w / ths[..., None])
    Wt = np.einsum("...ij,...j->...i", W, t)
    WWt = np.einsum("...ij,...j->...i", W, Wt)
    with np.errstate(divide="ignore", invalid="ignore"):
        coef = 1.0 - ths / (2.0 * np.tan(0.5 * ths))
    u = t - (0.5 * ths)[..., None] * Wt + coef[..., None] * WWt
    u = np.where(small[..., None], t, u)
    return np.concatenate([w, u], -1)


def pose_compose(Ra, ta, Rb, tb):
    return Ra @ Rb, np.einsum("...ij,...j->...i", Ra, tb) + ta


def pose_inverse(R, t):
    Rt = np.swapaxes(R, -1, -2)
    return Rt, -np.einsum("...ij,...j->...i", Rt, t)


def adjoint(R, t):
    """Pose3::AdjointMap = [[R, 0], [[t]x R, R]]."""
    A = np.zeros(R.shape[:-2] + (6, 6))
    A[..., :3, :3] = R
    A[..., 3:, 3:] = R
    A[..., 3:, :3] = skew(t) @ R
    return A


def _q_matrix(w: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Q block of Pose3::ExpmapDerivative (Barfoot eq. 102 with right-Jacobian signs = GTSAM's
    computeQforExpmapDerivative)."""
    V = skew(v)
    W = skew(w)
    WV, VW = W @ V, V @ W
    WVW = W @ V @ W
    th2 = np.sum(w * w, -1)
    th = np.sqrt(th2)
    small = th2 < EPS_THETA2
    ths = np.where(small, 1.0, th)
    s, c = np.sin(ths), np.cos(ths)
    th3, th4, th5 = ths**3, ths**4, ths**5
    a = np.where(small, 1.0 / 6.0, (ths - s) / th3)
    b = np.where(small, 1.0 / 24.0, (1.0 - 0.5 * ths * ths - c) / th4)
    # GTSAM: -0.5*((1 - th^2/2 - cos)/th^4 - 3 (th - sin - th^3/6)/th^5)
    cc = np.where(small, 1.0 / 120.0, -0.5 * (b - 3.0 * (ths - s - th3 / 6.0) / th5))
    # note: for the small-angle branch GTSAM uses -V/2 + 1/6(...) - 1/24(...) + 1/120(...)
    b_signed = np.where(small, -1.0 / 24.0, b)
    Q = (
        -0.5 * V
        + a[..., None, None] * (WV + VW - WVW)
        + b_signed[..., None, None] * (W @ WV + VW @ W - 3.0 * WVW)
        + cc[..., None, None] * (WVW @ W + W @ WVW)
    )
    return Q


def pose_expmap_derivative(xi: np.ndarray) -> np.ndarray:
    """Pose3::ExpmapDerivative(xi) = [[Jr, 0], [Q, Jr]]."""
    w, v = xi[..., :3], xi[..., 3:]
    Jr = so3_right_jacobian(w)
    Q = _q_matrix(w, v)
    J = np.zeros(xi.shape[:-1] + (6, 6))
    J[..., :3, :3] = Jr
    J[..., 3:, 3:] = Jr
    J[..., 3:, :3] = Q
    return J


def pose_logmap_derivative(R: np.ndarray, t: np.ndarray) -> np.ndarray:
    """Pose3::LogmapDerivative(T) = [[Jr^-1, 0], [-Jr^-1 Q Jr^-1, Jr^-1]] at xi = Log(T)."""
    xi = pose_log(R, t)
    w, v = xi[..., :3], xi[..., 3:]
    Ji = so3_right_jacobian_inv(w)
    Q = _q_matrix(w, v)
    J = np.zeros(xi.shape[:-1] + (6, 6))
    J[..., :3, :3] = Ji
    J[..., 3:, 3:] = Ji
    J[..., 3:, :3] = -Ji @ Q @ Ji
    return J


# ----------------------------------------------------------------------------------------------
# factors
# ----------------------------------------------------------------------------------------------
def pose_dynamics_factor(R1, t1, ang_vel1, vel1, R2, t2, dt: float, vel_frame: str = "world", jacobians=True):
    """PoseDynamicsFactor.error_func (factors.py:79-142), batched.

    Returns (error (N,6), [H0 (N,6,6), H1 (N,6,3), H2 (N,6,3), H3 (N,6,6)] or None).
    """
    assert vel_frame in ("world", "body")
    R1, t1, R2, t2 = (np.asarray(a, dtype=np.float64) for a in (R1, t1, R2, t2))
    ang_vel1 = np.asarray(ang_vel1, dtype=np.float64)
    vel1 = np.asarray(vel1, dtype=np.float64)
    R1T = np.swapaxes(R1, -1, -2)
    # factors.py:97-101 / :133-134 — rotate the world-frame velocity into the body frame
    vb = np.einsum("...ij,...j->...i", R1T, vel1) if vel_frame == "world" else vel1
    xi = dt * np.concatenate([ang_vel1, vb], -1)  # factors.py:104
    Ri, ti = pose_exp(xi)
    Rp, tp = pose_compose(R1, t1, Ri, ti)  # factors.py:105
    Rpi, tpi = pose_inverse(Rp, tp)
    Rr, tr = pose_compose(Rpi, tpi, R2, t2)  # between, factors.py:108
    err = pose_log(Rr, tr)  # factors.py:109
    if not jacobians:
        return err, None
    dlog = pose_logmap_derivative(Rr, tr)  # factors.py:112
    Rri, tri = pose_inverse(Rr, tr)
    drel_dpred = -adjoint(Rri, tri)  # between: H1 = -Ad(rel^-1)
    Rii, tii = pose_inverse(Ri, ti)
    dpred_dx0 = adjoint(Rii, tii)  # compose: H1 = Ad(inc^-1); H2 = I
    J_exp = pose_expmap_derivative(xi)
    A = dlog @ drel_dpred
    H0 = A @ dpred_dx0  # factors.py:114
    D = dt * (A @ J_exp)  # factors.py:117 (dpred_dtwist = I)
    H1 = D[..., :, :3].copy()  # factors.py:118
    if vel_frame == "world":
        # transformTo(v) on Pose3(R1, 0): d/dpose[:, :3] = [v_b]x ; d/dv = R1^T   (factors.py:122,125)
        H0 = H0.copy()
        H0[..., :, :3] += D[..., :, 3:] @ skew(vb)
        H2 = D[..., :, 3:] @ R1T
    else:
        H2 = D[..., :, 3:].copy()  # factors.py:128
    H3 = dlog.copy()  # factors.py:130 (drel_dpose2 = I)
    return err, [H0, H1, H2, H3]


def constant_velocity_factor(vel1, vel2, jacobians=True):
    """ConstantVelocityFactor.error_func (factors.py:160-171)."""
    vel1 = np.asarray(vel1, dtype=np.float64)
    vel2 = np.asarray(vel2, dtype=np.float64)
    err = vel2 - vel1
    if not jacobians:
        return err, None
    I = np.broadcast_to(np.eye(3), vel1.shape[:-1] + (3, 3))
    return err, [-I.copy(), I.copy()]


def keypoint_projection_factor(Rb, tb, K5, meas, pts_body, Rc=None, tc=None, jacobians=True):
    """KeypointProjectionFactor.error_func (factors.py:245-275), batched over N poses x P points.

    Rb (N,3,3), tb (N,3); K5 = (fx, fy, s, u0, v0); meas (N,P,2); pts_body (P,3);
    camera pose (Rc, tc) in the world (identity if None).
    Returns (error (N,P,2), H (N,P,2,6) or None, cheirality (N,P) bool).
    """
    Rb = np.asarray(Rb, dtype=np.float64)
    tb = np.asarray(tb, dtype=np.float64)
    meas = np.asarray(meas, dtype=np.float64)
    pts_body = np.asarray(pts_body, dtype=np.float64)
    fx, fy, s, u0, v0 = (float(k) for k in K5)
    if Rc is None:
        Rc, tc = np.eye(3), np.zeros(3)
    Rc = np.asarray(Rc, dtype=np.float64)
    tc = np.asarray(tc, dtype=np.float64)
    # body_pose.transformFrom(p_b)  (factors.py:257/269)
    pw = np.einsum("nij,pj->npi", Rb, pts_body) + tb[:, None, :]
    # PinholeCamera::project: q = camera_pose.transformTo(p_w)
    q = np.einsum("ji,npj->npi", Rc, pw - tc)
    cheir = q[..., 2] <= 0.0
    x = q[..., 0] / q[..., 2]
    y = q[..., 1] / q[..., 2]
    pix = np.stack([fx * x + s * y + u0, fy * y + v0], -1)
    err = pix - meas  # factors.py:265/272
    if not jacobians:
        return err, None, cheir
    N, P = meas.shape[:2]
    iz = 1.0 / q[..., 2]
    Dpn = np.zeros((N, P, 2, 3))
    Dpn[..., 0, 0] = iz
    Dpn[..., 0, 2] = -x * iz
    Dpn[..., 1, 1] = iz
    Dpn[..., 1, 2] = -y * iz
    Dcal = np.array([[fx, s], [0.0, fy]])
    dproj_dpoint = Dcal @ Dpn @ Rc.T  # (N,P,2,3)
    # transformFrom: d/dpose = [ -R [p]x , R ]
    Rp = -np.einsum("nij,pjk->npik", Rb, skew(pts_body))
    dpc_dpose = np.concatenate([Rp, np.broadcast_to(Rb[:, None], (N, P, 3, 3))], -1)  # (N,P,3,6)
    H = dproj_dpoint @ dpc_dpose  # factors.py:264
    return err, H, cheir
