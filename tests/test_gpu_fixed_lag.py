"""-m gpu: the camera -> keypoints -> smoother chain: retraction and window-advance kernels against
numpy, the allocation-free FixedLagSmoother against the dense oracle and on a synthetic trajectory, and the detector +
smoother captured in ONE CUDA graph per frame (BatchPipeline(post=smoother.push)).  No reference code exists for the
smoother itself (GTSAM does it in the reference), so its parity stays "unpinned": checks are against
oracle/smoother_oracle.py and by recovering a known trajectory."""
import numpy as np
import pytest
import torch

from oracle import factors_oracle as fo
from oracle import smoother_oracle as so
from perseus_b200.smoother import ChainSigmas, FixedLagSmoother
from perseus_b200.smoother.factors import pose12
from perseus_b200.smoother.linearize import retract
from tests import gpu_util as gu

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_retract_kernel_matches_oracle():
    rng = np.random.default_rng(0)
    n = 1000
    R, t = fo.pose_exp(rng.standard_normal((n, 6)))
    w, v = rng.standard_normal((n, 3)), rng.standard_normal((n, 3))
    d = rng.standard_normal((n, 12)) * 0.3
    d[:10, :3] = 1e-9  # the small-angle branch
    po, ao, vo = retract(pose12(_dev(R), _dev(t)), _dev(w), _dev(v), _dev(d))
    Rw, tw, ww, vw = so.retract(R, t, w, v, d)
    assert np.abs(po.cpu().numpy() - np.concatenate([Rw.reshape(n, 9), tw], 1)).max() < 1e-12
    assert np.array_equal(ao.cpu().numpy(), ww) and np.array_equal(vo.cpu().numpy(), vw)


@pytest.mark.parametrize("W,T,frame", [(1, 1, "world"), (3, 2, "body"), (5, 10, "world")])
def test_window_push_kernel(W, T, frame):
    rng = np.random.default_rng(T)
    P, dt = 8, 0.1
    pts = np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=np.float64) * 0.0175
    sm = FixedLagSmoother(W, T, _dev(pts), [221.7, 221.7, 0, 128, 128], dt, vel_frame=frame, iters=0)
    R, t = fo.pose_exp(rng.standard_normal((W * T, 6)) * 0.3)
    w, v = rng.standard_normal((W, T, 3)), rng.standard_normal((W, T, 3))
    meas = rng.uniform(0, 255, (W, T, P, 2))
    sm.reset(pose12(_dev(R), _dev(t)).reshape(W, T, 12), _dev(w), _dev(v), _dev(meas))
    kp = torch.rand(W, P, 2, device="cuda") * 255
    newest = sm.push(kp)
    poses = np.concatenate([R.reshape(W, T, 9), t.reshape(W, T, 3)], -1)
    # survivors moved one slot earlier
    assert np.array_equal(sm.poses[:, :-1].cpu().numpy(), poses[:, 1:])
    assert np.array_equal(sm.ang_vel[:, :-1].cpu().numpy(), w[:, 1:]) and np.array_equal(sm.vel[:, :-1].cpu().numpy(), v[:, 1:])
    assert np.array_equal(sm.meas[:, :-1].cpu().numpy(), meas[:, 1:])
    # newest = the dynamics factor's prediction from the old newest state; velocities carried over; measurement widened
    Rl, tl = R.reshape(W, T, 3, 3)[:, -1], t.reshape(W, T, 3)[:, -1]
    vb = np.einsum("wji,wj->wi", Rl, v[:, -1]) if frame == "world" else v[:, -1]
    Ri, ti = fo.pose_exp(dt * np.concatenate([w[:, -1], vb], -1))
    Rn, tn = fo.pose_compose(Rl, tl, Ri, ti)
    assert np.abs(newest.cpu().numpy() - np.concatenate([Rn.reshape(W, 9), tn], -1)).max() < 1e-12
    assert np.array_equal(sm.ang_vel[:, -1].cpu().numpy(), w[:, -1]) and np.array_equal(sm.vel[:, -1].cpu().numpy(), v[:, -1])
    assert np.array_equal(sm.meas[:, -1].cpu().numpy(), kp.double().cpu().numpy())


def test_fixed_lag_iteration_matches_dense_oracle():
    W, T = 3, 8
    cs = [so.synthetic_chain(T, seed=w) for w in range(W)]
    rng = np.random.default_rng(7)
    st = [so.retract(c["R"], c["t"], c["ang_vel"], c["vel"],
                     np.concatenate([0.05 * rng.standard_normal((T, 3)), 0.005 * rng.standard_normal((T, 3)),
                                     0.1 * rng.standard_normal((T, 3)), 0.02 * rng.standard_normal((T, 3))], 1)) for c in cs]
    sm = FixedLagSmoother(W, T, _dev(cs[0]["pts_body"]), cs[0]["K5"], cs[0]["dt"], sigmas=ChainSigmas(), iters=1, lam=1e-6)
    sm.reset(torch.stack([pose12(_dev(s[0]), _dev(s[1])) for s in st]), torch.stack([_dev(s[2]) for s in st]),
             torch.stack([_dev(s[3]) for s in st]), torch.stack([_dev(c["meas"]) for c in cs]))
    sm.iterate()
    torch.cuda.synchronize()
    assert int(sm.status.abs().sum()) == 0
    for w in range(W):
        delta, _, _, _ = so.chain_step(*st[w], cs[w]["meas"], cs[w]["pts_body"], cs[w]["K5"], cs[w]["dt"], sig_px=1.0,
                                       sig_dyn=[0.1] * 6, sig_cvw=[0.1] * 3, sig_cvv=[0.1] * 3, lam=1e-6)
        assert np.abs(sm.delta[w].cpu().numpy().reshape(-1) - delta).max() < 1e-7 * max(1.0, np.abs(delta).max())
        Rw, tw, ww, vw = so.retract(*st[w], delta.reshape(T, 12))
        assert np.abs(sm.poses[w].cpu().numpy() - np.concatenate([Rw.reshape(T, 9), tw], 1)).max() < 1e-7


def test_fixed_lag_tracks_a_synthetic_trajectory():
    """Feed noisy projections of a known constant-twist trajectory frame by frame: the newest pose stays on the truth."""
    T, n_frames = 10, 30
    c = so.synthetic_chain(n_frames + T, seed=3, noise_px=0.2)
    meas = c["meas"]
    sm = FixedLagSmoother(1, T, _dev(c["pts_body"]), c["K5"], c["dt"], iters=2, lam=1e-6)
    sm.reset(pose12(_dev(c["R"][:T]), _dev(c["t"][:T]))[None], _dev(c["ang_vel"][:T])[None], _dev(c["vel"][:T])[None],
             _dev(meas[:T])[None])
    for k in range(T, T + n_frames):
        newest = sm.push(_dev(meas[k].astype(np.float32))[None])
        got = newest[0].cpu().numpy()
        # a 35 mm cube at ~0.45 m seen with 0.2 px noise: lateral position to a few mm, depth to a few cm
        assert np.abs(got[9:11] - c["t"][k][:2]).max() < 5e-3, f"frame {k}: lateral error {np.abs(got[9:11] - c['t'][k][:2]).max():.4f} m"
        assert abs(got[11] - c["t"][k][2]) < 4e-2, f"frame {k}: depth error {abs(got[11] - c['t'][k][2]):.4f} m"
        assert np.abs(got[:9].reshape(3, 3) @ c["R"][k].T - np.eye(3)).max() < 0.15
    assert int(sm.status.abs().sum()) == 0


def test_detector_and_smoother_in_one_graph_per_frame():
    """BatchPipeline(post=smoother.push): H2D -> ingest -> backbone -> head/decode -> window advance -> Gauss-Newton x2,
    one CUDA-graph replay per camera frame; equals the eager chain (detector, then smoother.push on its keypoints)."""
    from perseus_b200.detector import BatchPipeline

    model, _ = gu.native_model(3, "fp16x2")
    B, T = 2, 10  # two cameras / cubes, window of 10
    c = so.synthetic_chain(T, seed=5)
    mk = lambda: FixedLagSmoother(B, T, _dev(c["pts_body"]), c["K5"], c["dt"], iters=2, lam=1e-3, prior=1.0)  # noqa: E731
    init = (pose12(_dev(c["R"]), _dev(c["t"]))[None].repeat(B, 1, 1), _dev(c["ang_vel"])[None].repeat(B, 1, 1),
            _dev(c["vel"])[None].repeat(B, 1, 1), _dev(c["meas"])[None].repeat(B, 1, 1, 1))
    sm_graph, sm_eager = mk(), mk()
    pipe = BatchPipeline(model, B, 376, 672, 4, bgr=True, post=sm_graph.push)
    sm_graph.reset(*init)
    sm_eager.reset(*init)
    for i in range(5):
        frame = gu.synth_u8_frames(B, 376, 672, 4, seed=60 + i).pin_memory()
        t = pipe.submit(frame)
        pose = pipe.result_post(t).clone()
        px = pipe.result(t)
        want_px = model.forward_frames(frame.cuda(), bgr=True)
        assert torch.equal(px, want_px.cpu())
        want_pose = sm_eager.push(want_px).cpu()
        assert torch.allclose(pose, want_pose, rtol=0, atol=1e-12), f"frame {i}"
