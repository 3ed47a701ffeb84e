"""The expressions the CUDA factor kernels evaluate (perseus_b200/csrc/factors_math.h), compiled
for the host by tests/host_shim.cpp, against the numpy oracle — CPU only, no GPU needed.
The GPU run of the same arithmetic is tests/test_gpu_factors.py."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import factors_oracle as fo

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def shim(tmp_path_factory):
    out = tmp_path_factory.mktemp("shim") / "host_shim.so"
    subprocess.check_call(
        ["g++", "-O2", "-shared", "-fPIC", "-std=c++17", "-o", str(out), os.path.join(HERE, "host_shim.cpp")]
    )
    return ctypes.CDLL(str(out))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def pose12(R, t):
    return np.ascontiguousarray(np.concatenate([R.reshape(-1, 9), t], -1))


def random_problem(N, seed, rel_scale=1.0):
    rng = np.random.default_rng(seed)
    R1, t1 = fo.pose_exp(rng.standard_normal((N, 6)))
    R2, t2 = fo.pose_exp(rng.standard_normal((N, 6)) * rel_scale)
    return R1, t1, R2, t2, rng.standard_normal((N, 3)), rng.standard_normal((N, 3))


@pytest.mark.parametrize("frame", ["world", "body"])
def test_dynamics_matches_oracle(shim, frame):
    N = 2000
    R1, t1, R2, t2, w, v = random_problem(N, 0)
    err, H = fo.pose_dynamics_factor(R1, t1, w, v, R2, t2, 0.1, frame)
    p1, p2 = pose12(R1, t1), pose12(R2, t2)
    res = np.zeros((N, 6))
    H0, H1, H2, H3 = np.zeros((N, 36)), np.zeros((N, 18)), np.zeros((N, 18)), np.zeros((N, 36))
    shim.shim_dynamics(_p(p1), _p(w), _p(v), _p(p2), ctypes.c_double(0.1), int(frame == "world"), 1,
                       _p(res), _p(H0), _p(H1), _p(H2), _p(H3), ctypes.c_int64(N))
    assert np.abs(res - err).max() < 1e-11
    # column-major 6x6 / 6x3
    assert np.abs(H0.reshape(N, 6, 6).transpose(0, 2, 1) - H[0]).max() < 1e-9
    assert np.abs(H1.reshape(N, 3, 6).transpose(0, 2, 1) - H[1]).max() < 1e-9
    assert np.abs(H2.reshape(N, 3, 6).transpose(0, 2, 1) - H[2]).max() < 1e-9
    assert np.abs(H3.reshape(N, 6, 6).transpose(0, 2, 1) - H[3]).max() < 1e-9
    res2 = np.zeros((N, 6))
    shim.shim_dynamics(_p(p1), _p(w), _p(v), _p(p2), ctypes.c_double(0.1), int(frame == "world"), 0,
                       _p(res2), None, None, None, None, ctypes.c_int64(N))
    assert np.array_equal(res, res2)


@pytest.mark.parametrize("frame", ["world", "body"])
@pytest.mark.parametrize("rel_scale", [1.0, 1e-9])
def test_block_triangular_dynamics_equals_the_dense_6x6_form(shim, frame, rel_scale):
    """dynamics_factor multiplies the [[A,0],[B,A]] Jacobians blockwise (factors_math.h); the dense 6x6 form multiplies the
    structural zeros out and builds the coefficient matrices (Jr, Jr^-1, Q) as GTSAM writes them, from products of skew
    matrices, where dynamics_factor uses their outer-product closed forms: equal up to rounding, at generic poses and in the
    small-angle branches."""
    N = 3000
    R1, t1, R2, t2, w, v = random_problem(N, 3)
    if rel_scale < 1.0:  # x2 ~ x1 * Exp(dt * twist): the logarithm's and the Jacobians' small-angle branches
        w, v = w * rel_scale, v * rel_scale
        R2, t2 = R1.copy(), t1.copy()
    p1, p2 = pose12(R1, t1), pose12(R2, t2)
    out = []
    for dense in (False, True):
        res = np.zeros((N, 6))
        H0, H1, H2, H3 = np.zeros((N, 36)), np.zeros((N, 18)), np.zeros((N, 18)), np.zeros((N, 36))
        if dense:
            shim.shim_dynamics_dense(_p(p1), _p(w), _p(v), _p(p2), ctypes.c_double(0.1), int(frame == "world"),
                                     _p(res), _p(H0), _p(H1), _p(H2), _p(H3), ctypes.c_int64(N))
        else:
            shim.shim_dynamics(_p(p1), _p(w), _p(v), _p(p2), ctypes.c_double(0.1), int(frame == "world"), 1,
                               _p(res), _p(H0), _p(H1), _p(H2), _p(H3), ctypes.c_int64(N))
        out.append((res, H0, H1, H2, H3))
    for a, b in zip(*out):
        assert np.isfinite(a).all()
        scale = max(1.0, float(np.abs(b).max()))
        assert np.abs(a - b).max() <= 2e-12 * scale, float(np.abs(a - b).max())


def test_so3_log_branches(shim):
    rng = np.random.default_rng(3)
    ax = rng.standard_normal((64, 3))
    ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    for ang in [0.0, 1e-9, 1e-4, 0.5, 3.0, np.pi - 1e-2, np.pi - 1e-5, np.pi - 1e-9]:
        R = fo.so3_exp(ax * ang)
        want = fo.so3_log(R)
        got = np.zeros_like(want)
        for i in range(len(R)):
            Ri = np.ascontiguousarray(R[i])
            shim.shim_so3_log(_p(Ri), _p(got[i]))
        assert np.abs(got - want).max() < 1e-12, ang


def test_projection_matches_oracle(shim):
    rng = np.random.default_rng(4)
    N, P = 500, 8
    Rb, tb = fo.pose_exp(rng.standard_normal((N, 6)) * np.array([0.5, 0.5, 0.5, 0.05, 0.05, 0.05]))
    tb = tb + np.array([0.0, 0.0, 0.45])
    Rc, tc = fo.pose_exp(np.array([[0.1, -0.05, 0.02, 0.01, 0.02, -0.03]]))
    K5 = np.array([300.0, 310.0, 0.5, 128.0, 127.0])
    corners = 0.0175 * np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=float)
    meas = rng.uniform(0, 256, (N, P, 2))
    for cam in (None, pose12(Rc, tc)[0].copy()):
        err, H, cheir = fo.keypoint_projection_factor(
            Rb, tb, K5, meas, corners, None if cam is None else Rc[0], None if cam is None else tc[0]
        )
        res = np.zeros((N, P, 2))
        J = np.zeros((N, P, 12))
        ch = np.zeros((N, P), dtype=np.uint8)
        pp = pose12(Rb, tb)
        shim.shim_projection(_p(pp), _p(meas), _p(corners), _p(K5), None if cam is None else _p(cam), 1,
                             _p(res), _p(J), _p(ch), ctypes.c_int64(N), P)
        assert np.abs(res - err).max() < 1e-9
        assert np.abs(J.reshape(N, P, 6, 2).transpose(0, 1, 3, 2) - H).max() < 1e-8
        assert np.array_equal(ch.astype(bool), cheir)
