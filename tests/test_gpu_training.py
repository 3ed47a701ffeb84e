"""-m gpu: training-step operators (pb200_train_*) vs torch autograd on the same bf16 operands.

Tolerances: tensor-core GEMMs on identical bf16 operands differ from the fp64 result only by fp32 accumulation
(relative ~1e-4 of the sum of |terms|); outputs stored in bf16 add 2^-8 relative.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _lib():
    from perseus_b200 import _lib as L

    return L


def _stream():
    return _lib().current_stream_ptr()


WGRAD_CASES = [
    # B, H, W, Cin, Cout, R, stride, pad
    (3, 64, 64, 64, 64, 3, 1, 1),     # layer1: 9 units -> 5 pairs (odd tail), N = 64
    (4, 64, 64, 64, 128, 3, 2, 1),    # layer2.0.conv1: stride 2 parity views
    (4, 64, 64, 64, 128, 1, 2, 0),    # layer2.0.downsample: 1 unit, half-empty pair
    (5, 32, 32, 128, 128, 3, 1, 1),
    (6, 16, 16, 256, 256, 3, 1, 1),   # N = 256
    (9, 8, 8, 512, 512, 3, 1, 1),     # image smaller than a k-block (2 images per tile, ragged batch)
    (3, 16, 16, 256, 512, 3, 2, 1),
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=[str(c) for c in WGRAD_CASES])
def test_wgrad_tcgen05_vs_autograd(case):
    L = _lib()
    B, H, W, Cin, Cout, R, stride, pad = case
    g = torch.Generator().manual_seed(sum(case))
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    x = torch.randn(B, H, W, Cin, generator=g).bfloat16()
    dz = torch.randn(B, Ho, Wo, Cout, generator=g).bfloat16()
    want = torch.nn.grad.conv2d_weight(x.double().permute(0, 3, 1, 2), (Cout, Cin, R, R),
                                       dz.double().permute(0, 3, 1, 2), stride=stride, padding=pad)
    xd, dzd = x.cuda(), dz.cuda()
    scratch = torch.zeros(R * R, Cin, Cout, dtype=torch.float32, device="cuda")
    dw = torch.full((Cout, Cin, R, R), float("nan"), dtype=torch.float32, device="cuda")
    L.check(L.lib().pb200_train_wgrad(xd.data_ptr(), dzd.data_ptr(), B, H, W, Cin, Cout, R, stride, pad,
                                      scratch.data_ptr(), _stream()))
    L.check(L.lib().pb200_train_unpack_wgrad(scratch.data_ptr(), Cout, Cin, R, Cin, dw.data_ptr(), _stream()))
    torch.cuda.synchronize()
    got = dw.double().cpu()
    assert torch.isfinite(got).all()
    err = (got - want).abs()
    tol = 2e-3 * (B * Ho * Wo) ** 0.5 + 1e-4 * want.abs()
    assert (err <= tol).all(), f"max err {float(err.max()):.4f} (|want| max {float(want.abs().max()):.2f}); " \
                               f"bad {int((err > tol).sum())}/{err.numel()} first {(err > tol).nonzero()[:4].tolist()}"


DGRAD_CASES = [
    (2, 64, 64, 64, 64, 3),
    (3, 16, 16, 256, 256, 3),
    (5, 8, 8, 512, 512, 3),
]


@pytest.mark.parametrize("case", DGRAD_CASES, ids=[str(c) for c in DGRAD_CASES])
def test_dgrad_stride1_is_a_conv_with_the_flipped_transposed_filter(case):
    """data gradient of a 3x3/1 conv = pb200_train_conv on dz with weights packed in mode 1."""
    L = _lib()
    B, H, W, Cin, Cout, R = case
    g = torch.Generator().manual_seed(sum(case))
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5)
    dz = torch.randn(B, H, W, Cout, generator=g).bfloat16()
    wb = w.bfloat16().double()
    want = torch.nn.grad.conv2d_input((B, Cin, H, W), wb, dz.double().permute(0, 3, 1, 2), stride=1, padding=1)
    want = want.permute(0, 2, 3, 1)
    wd = w.cuda().contiguous()
    packed = torch.empty(Cin * R * R * Cout, dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_pack_weights(wd.data_ptr(), Cout, Cin, R, 1, packed.data_ptr(), _stream()))
    zero_bias = torch.zeros(512, dtype=torch.float32, device="cuda")
    out = torch.full((B, H, W, Cin), float("nan"), dtype=torch.bfloat16, device="cuda")
    dzd = dz.cuda()
    # dgrad: input channels = Cout, output channels = Cin
    L.check(L.lib().pb200_train_conv(dzd.data_ptr(), B, H, W, Cout, packed.data_ptr(), zero_bias.data_ptr(), Cin, R, 1, 1,
                                     None, 0, out.data_ptr(), None, _stream()))
    torch.cuda.synchronize()
    got = out.double().cpu()
    assert torch.isfinite(got).all()
    err = (got - want).abs()
    tol = 2.0 ** -7 * want.abs() + 1e-2
    assert (err <= tol).all(), f"max err {float(err.max()):.4f}"


# ---- memory-bound training kernels ------------------------------------------------------------------------------
@pytest.mark.parametrize("M,C,use_res,relu", [(4096, 64, False, True), (1000, 128, True, True), (257, 512, False, False)])
def test_bn_forward_and_backward_vs_autograd(M, C, use_res, relu):
    L = _lib()
    g = torch.Generator().manual_seed(M + C)
    z = (torch.randn(M, C, generator=g) * 1.5 + 0.3).bfloat16()
    res = torch.randn(M, C, generator=g).bfloat16() if use_res else None
    gamma, beta = torch.rand(C, generator=g) + 0.5, torch.randn(C, generator=g) * 0.1
    rm, rv = torch.randn(C, generator=g) * 0.1, torch.rand(C, generator=g) + 0.5
    dy = torch.randn(M, C, generator=g).bfloat16()
    # reference in fp64 on the same bf16 inputs
    zr = z.double().requires_grad_(True)
    gr, br = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    rm_ref, rv_ref = rm.double().clone(), rv.double().clone()
    pre = F.batch_norm(zr, rm_ref, rv_ref, gr, br, training=True, momentum=0.1, eps=1e-5)
    if use_res:
        pre = pre + res.double()
    y_ref = pre.relu() if relu else pre
    zd, dyd = z.cuda(), dy.cuda()
    resd = res.cuda() if use_res else None
    gd, bd, rmd, rvd = gamma.cuda(), beta.cuda(), rm.cuda(), rv.cuda()
    sums = torch.zeros(2 * C, dtype=torch.float64, device="cuda")
    mean, invstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    y = torch.full((M, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_bn_stats(zd.data_ptr(), M, C, sums.data_ptr(), _stream()))
    L.check(L.lib().pb200_train_bn_apply(zd.data_ptr(), sums.data_ptr(), gd.data_ptr(), bd.data_ptr(), rmd.data_ptr(),
                                         rvd.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(),
                                         None if resd is None else resd.data_ptr(), int(relu), y.data_ptr(), M, C, _stream()))
    torch.cuda.synchronize()
    assert torch.allclose(rmd.double().cpu(), rm_ref, atol=1e-5) and torch.allclose(rvd.double().cpu(), rv_ref, rtol=1e-5, atol=1e-5)
    err = (y.double().cpu() - y_ref.detach()).abs()
    assert (err <= 2.0 ** -7 * y_ref.detach().abs() + 1e-3).all(), float(err.max())
    # backward: the kernel masks with ITS OWN stored output, as the trainer does
    y_mask = y if relu else None
    g_masked = dy.double() * ((y.double().cpu() > 0) if relu else 1.0)
    pre.backward(g_masked)
    dz = torch.full((M, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    g_out = torch.full((M, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    dgam, dbet = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    L.check(L.lib().pb200_train_bn_backward(dyd.data_ptr(), None if y_mask is None else y_mask.data_ptr(), zd.data_ptr(),
                                            mean.data_ptr(), invstd.data_ptr(), gd.data_ptr(), None, sums.data_ptr(), M, C,
                                            dz.data_ptr(), g_out.data_ptr(), None, 0, 0, dgam.data_ptr(), dbet.data_ptr(),
                                            _stream()))
    torch.cuda.synchronize()
    if relu and not use_res:  # the mask recomputed from z must be the mask read from y
        dz2 = torch.full((M, C), float("nan"), dtype=torch.bfloat16, device="cuda")
        L.check(L.lib().pb200_train_bn_backward(dyd.data_ptr(), None, zd.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                gd.data_ptr(), bd.data_ptr(), sums.data_ptr(), M, C, dz2.data_ptr(), None, None,
                                                0, 0, dgam.data_ptr(), dbet.data_ptr(), _stream()))
        torch.cuda.synchronize()
        d2 = (dz2.double() - dz.double()).abs()  # the two calls reduce with different atomic orders: 1 bf16 ulp
        assert (d2 <= 2.0 ** -7 * dz.double().abs() + 1e-5).all() and torch.equal(dz2 == 0, dz == 0)
    assert torch.equal(g_out.double().cpu(), g_masked.bfloat16().double())
    e = (dz.double().cpu() - zr.grad).abs()
    assert (e <= 2.0 ** -7 * zr.grad.abs() + 2e-3).all(), float(e.max())
    assert torch.allclose(dgam.double().cpu(), gr.grad, rtol=2e-3, atol=2e-2)
    assert torch.allclose(dbet.double().cpu(), br.grad, rtol=2e-3, atol=2e-2)


def test_bn_backward_zero_stuffed_copy():
    L = _lib()
    B, Ho, Wo, C = 2, 4, 8, 64
    M = B * Ho * Wo
    g = torch.Generator().manual_seed(1)
    z, dy = torch.randn(M, C, generator=g).bfloat16().cuda(), torch.randn(M, C, generator=g).bfloat16().cuda()
    mean, invstd = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    gamma = torch.ones(C, device="cuda")
    sums = torch.zeros(2 * C, dtype=torch.float64, device="cuda")
    dz = torch.empty(M, C, dtype=torch.bfloat16, device="cuda")
    stuffed = torch.full((B, 2 * Ho, 2 * Wo, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    L.check(L.lib().pb200_train_bn_backward(dy.data_ptr(), None, z.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                            gamma.data_ptr(), None, sums.data_ptr(), M, C, dz.data_ptr(), None, stuffed.data_ptr(),
                                            Ho, Wo, dg.data_ptr(), db.data_ptr(), _stream()))
    torch.cuda.synchronize()
    want = torch.zeros_like(stuffed)
    want[:, ::2, ::2] = dz.view(B, Ho, Wo, C)
    assert torch.equal(stuffed, want)


def test_maxpool_forward_backward_vs_autograd():
    L = _lib()
    B, H, W, C = 3, 16, 32, 64
    g = torch.Generator().manual_seed(2)
    x = torch.randn(B, H, W, C, generator=g).relu().bfloat16()  # many exact ties at 0 and some at equal bf16 values
    dout = torch.randn(B, H // 2, W // 2, C, generator=g).bfloat16()
    xr = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 2, 1)
    yr.backward(dout.double().permute(0, 3, 1, 2))
    xd, dd = x.cuda(), dout.cuda()
    out = torch.empty(B, H // 2, W // 2, C, dtype=torch.bfloat16, device="cuda")
    idx = torch.empty(B * (H // 2) * (W // 2) * C, dtype=torch.uint8, device="cuda")
    din = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_maxpool(xd.data_ptr(), out.data_ptr(), idx.data_ptr(), B, H, W, C, _stream()))
    L.check(L.lib().pb200_train_maxpool_backward(dd.data_ptr(), idx.data_ptr(), din.data_ptr(), B, H, W, C, _stream()))
    torch.cuda.synchronize()
    assert torch.equal(out.double().cpu(), yr.detach().permute(0, 2, 3, 1))
    want = xr.grad.permute(0, 2, 3, 1)
    e = (din.double().cpu() - want).abs()
    assert (e <= 2.0 ** -7 * want.abs() + 1e-6).all(), float(e.max())


def test_adamw_clip_step_vs_torch():
    L = _lib()
    n = 100_003
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(n, generator=g)
    p = torch.nn.Parameter(p0.clone().cuda())
    opt = torch.optim.AdamW([p], lr=1e-3)
    mine_p = p0.clone().cuda()
    m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    sq = torch.zeros(1, dtype=torch.float64, device="cuda")
    for step in range(1, 4):
        grad = (torch.randn(n, generator=g) * (3.0 if step == 1 else 0.001)).cuda()  # clipped, then not clipped
        p.grad = grad.clone()
        torch.nn.utils.clip_grad_norm_([p], 1.0)
        opt.step()
        L.check(L.lib().pb200_train_adamw_step(mine_p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), n, 1e-3, 0.9, 0.999,
                                               1e-8, 1e-2, step, 1.0, 1.0, sq.data_ptr(), _stream()))
        torch.cuda.synchronize()
        assert torch.allclose(mine_p, p.detach(), rtol=1e-5, atol=1e-6), float((mine_p - p.detach()).abs().max())


# ---- the whole step ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("C", [4, 3])
def test_native_train_step_matches_autograd(C):
    """forward + SmoothL1 + backward + clip + AdamW through pb200_train_* vs the torch module tree under bf16
    autocast (what stage A / the reference's train.py:259-304 runs) and vs the same in fp32."""
    import copy

    from perseus_b200.detector import KeypointCNN
    from perseus_b200.detector.training import NativeTrainer

    torch.manual_seed(0)
    B = 8
    model = KeypointCNN(num_channels=C).cuda()
    ref = copy.deepcopy(model.resnet).float().train()
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.rand(B, C, 256, 256, device="cuda", generator=g)
    if C == 4:
        x[:, 3] = x[:, 3] * 8 + 3  # depth in data units
    tgt = torch.rand(B, 8, 2, device="cuda", generator=g) * 2 - 1
    tr = NativeTrainer(model, B)
    pred = tr.forward(x).clone()
    loss = float(tr.backward(tgt))

    def run_ref(autocast):
        r = copy.deepcopy(ref)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            out = r(x).float()
            ls = F.smooth_l1_loss(out.reshape(-1, 8, 2), tgt, beta=1.0)
        ls.backward()
        return r, out.detach(), float(ls)

    r32, out32, loss32 = run_ref(False)
    r16, out16, loss16 = run_ref(True)
    e_pred = float((pred - out32).abs().max())
    e_pred16 = float((out16 - out32).abs().max())
    print(f"[train C={C}] pred err vs fp32 {e_pred:.4f} (autocast's own {e_pred16:.4f}); loss {loss:.5f} / {loss32:.5f} / {loss16:.5f}")
    assert e_pred <= 3 * e_pred16 + 0.02
    assert abs(loss - loss32) <= 0.02 * abs(loss32) + 3 * abs(loss16 - loss32)
    worst = 0.0
    for (name, p32), (_, p16) in zip(r32.named_parameters(), r16.named_parameters()):
        mine = tr._g(dict(model.resnet.named_parameters())[name])
        den = float(p32.grad.norm()) + 1e-12
        rel = float((mine - p32.grad).norm()) / den
        rel16 = float((p16.grad.float() - p32.grad).norm()) / den
        if rel > worst:
            worst, worst_name, worst16 = rel, name, rel16
        assert rel <= 3 * rel16 + 0.05, f"{name}: rel err {rel:.3f} vs fp32 autograd (autocast's own {rel16:.3f})"
    print(f"[train C={C}] worst per-tensor gradient rel err {worst:.3f} ({worst_name}; autocast's own {worst16:.3f})")
    # optimizer: same update as clip_grad_norm_ + AdamW on OUR gradients
    params = [p for p in model.parameters()]
    before = [p.detach().clone() for p in params]
    grads = [tr._g(p).clone() for p in params]
    tr.optimizer_step()
    torch.cuda.synchronize()
    shadow = [torch.nn.Parameter(b.clone()) for b in before]
    for s, gr in zip(shadow, grads):
        s.grad = gr
    torch.nn.utils.clip_grad_norm_(shadow, 1.0)
    torch.optim.AdamW(shadow, lr=1e-3).step()
    for p, s in zip(params, shadow):
        assert torch.allclose(p.detach(), s.detach(), rtol=1e-5, atol=1e-6)
    # BN running statistics follow nn.BatchNorm2d
    for (n1, b1), (n2, b2) in zip(model.resnet.named_buffers(), r32.named_buffers()):
        if n1.endswith("num_batches_tracked"):
            assert int(b1) == int(b2)
        else:
            assert torch.allclose(b1, b2, rtol=2e-2, atol=2e-2), n1
    # the eval-mode native forward sees the updated weights
    model.eval()
    with torch.no_grad():
        y = model(x)
    assert torch.isfinite(y).all()


def test_graph_replayed_steps_match_eager_steps():
    """step() captures forward + backward into one CUDA graph after two eager steps.  The split-K wgrad and the BN
    statistics use atomics, so two runs differ by rounding noise that Adam's sign-like first updates amplify; the
    graphed trainer must stay as close to an eager trainer as a second eager trainer does."""
    import copy

    from perseus_b200.detector import KeypointCNN
    from perseus_b200.detector.training import NativeTrainer

    torch.manual_seed(0)
    B = 4
    models = [KeypointCNN(num_channels=4).cuda()]
    models += [copy.deepcopy(models[0]) for _ in range(2)]
    trainers = [NativeTrainer(models[0], B, use_graph=False), NativeTrainer(models[1], B, use_graph=False),
                NativeTrainer(models[2], B, use_graph=True)]
    g = torch.Generator(device="cuda").manual_seed(5)
    for i in range(5):
        x = torch.rand(B, 4, 256, 256, device="cuda", generator=g)
        y = torch.rand(B, 8, 2, device="cuda", generator=g) * 2 - 1
        ls = [float(t.step(x, y)) for t in trainers]
        print(f"[graph] step {i}: losses {ls}")
        assert abs(ls[0] - ls[2]) <= 5 * abs(ls[0] - ls[1]) + 3e-2 * abs(ls[0]) + 1e-4, ls
    assert trainers[2]._graph is not None
    worst = 0.0
    for (n, a), (_, b), (_, c) in zip(*[m.named_parameters() for m in models]):
        noise = float((a - b).abs().mean())
        d = float((a - c).abs().mean())
        worst = max(worst, d / (noise + 1e-7))
        assert d <= 3 * noise + 1e-4, (n, d, noise)
    print(f"[graph] worst graph-vs-eager / eager-vs-eager mean |dw| ratio {worst:.2f}")


STAT_CASES = [
    # B, H, W, Cin, Cout, R, stride   (Cout 64 -> BLOCK_N 64 path; big M -> CTA-pair kernel; ragged M tile)
    (2, 64, 64, 64, 64, 3, 1),
    (3, 16, 16, 128, 256, 3, 1),
    (5, 8, 8, 512, 512, 3, 1),
    (40, 32, 32, 128, 128, 3, 1),
    (4, 32, 32, 64, 128, 1, 2),
    (3, 8, 8, 256, 512, 3, 1),
]


@pytest.mark.parametrize("case", STAT_CASES, ids=[str(c) for c in STAT_CASES])
def test_conv_epilogue_batch_statistics(case):
    """pb200_train_conv's fused per-channel sum / sum of squares == the statistics of the bf16 tensor it wrote."""
    L = _lib()
    B, H, W, Cin, Cout, R, stride = case
    g = torch.Generator().manual_seed(sum(case))
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5).cuda().contiguous()
    x = torch.randn(B, H, W, Cin, generator=g).bfloat16().cuda()
    packed = torch.empty(Cin * R * R * Cout, dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_pack_weights(w.data_ptr(), Cout, Cin, R, 0, packed.data_ptr(), _stream()))
    zero_bias = torch.zeros(512, dtype=torch.float32, device="cuda")
    Ho, Wo = (H + 2 * (R // 2) - R) // stride + 1, (W + 2 * (R // 2) - R) // stride + 1
    out = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    stats = torch.zeros(2 * Cout, dtype=torch.float64, device="cuda")
    L.check(L.lib().pb200_train_conv(x.data_ptr(), B, H, W, Cin, packed.data_ptr(), zero_bias.data_ptr(), Cout, R, stride, R // 2,
                                     None, 0, out.data_ptr(), stats.data_ptr(), _stream()))
    torch.cuda.synchronize()
    z = out.double().reshape(-1, Cout)
    assert torch.isfinite(z).all()
    want = torch.cat([z.sum(0), (z * z).sum(0)])
    scale = torch.cat([z.abs().sum(0), (z * z).sum(0)]) + 1e-6
    assert float(((stats - want).abs() / scale).max()) < 1e-5
    # and the separate pass agrees with it
    sums = torch.zeros(2 * Cout, dtype=torch.float64, device="cuda")
    L.check(L.lib().pb200_train_bn_stats(out.data_ptr(), z.shape[0], Cout, sums.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert float(((stats - sums).abs() / scale).max()) < 1e-5


DGRAD_S2_CASES = [
    # B, Ho, Wo, Cin (forward input channels), Cout, R
    (3, 32, 32, 64, 128, 3),    # layer2.0.conv1: N = 64 tiles
    (2, 16, 16, 128, 256, 3),   # layer3.0.conv1
    (5, 8, 8, 256, 512, 3),     # layer4.0.conv1: two images per tile, ragged tail
    (3, 32, 32, 64, 128, 1),    # layer2.0.downsample
    (3, 8, 8, 256, 512, 1),
]


@pytest.mark.parametrize("with_res", [False, True])
@pytest.mark.parametrize("case", DGRAD_S2_CASES, ids=[str(c) for c in DGRAD_S2_CASES])
def test_dgrad_stride2_by_output_parity(case, with_res):
    """pb200_train_dgrad_s2 == the exact input gradient of a stride-2 conv (fp64 on the same bf16 operands)."""
    L = _lib()
    B, Ho, Wo, Cin, Cout, R = case
    H, W = 2 * Ho, 2 * Wo
    g = torch.Generator().manual_seed(sum(case) + with_res)
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5)
    dz = torch.randn(B, Ho, Wo, Cout, generator=g).bfloat16()
    want = torch.nn.grad.conv2d_input((B, Cin, H, W), w.bfloat16().double(), dz.double().permute(0, 3, 1, 2), stride=2,
                                      padding=R // 2).permute(0, 2, 3, 1)
    res = torch.randn(B, H, W, Cin, generator=g).bfloat16() if with_res else None
    if with_res:
        want = want.clone()
        want[:, 0::2, 0::2] += res.double()[:, 0::2, 0::2]  # the projection's gradient lives on the parity-(0,0) pixels
    packed = torch.empty(Cin * R * R * Cout, dtype=torch.bfloat16, device="cuda")
    wd = w.cuda().contiguous()
    L.check(L.lib().pb200_train_pack_weights(wd.data_ptr(), Cout, Cin, R, 1, packed.data_ptr(), _stream()))
    zero_bias = torch.zeros(512, dtype=torch.float32, device="cuda")
    out = torch.full((B, H, W, Cin), float("nan"), dtype=torch.bfloat16, device="cuda")
    dz_d, res_d = dz.cuda(), None if res is None else res.cuda()  # held: a temporary's block could be reused before the launch
    L.check(L.lib().pb200_train_dgrad_s2(dz_d.data_ptr(), B, Ho, Wo, Cout, packed.data_ptr(), zero_bias.data_ptr(), Cin, R,
                                         None if res_d is None else res_d.data_ptr(), out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    got = out.double().cpu()
    if R == 1:  # a 1x1/2 layer touches the parity-(0,0) pixels only; the others are left as they were
        assert torch.isnan(got[:, 1::2]).all() and torch.isnan(got[:, :, 1::2]).all()
        got, want = got[:, 0::2, 0::2], want[:, 0::2, 0::2]
    assert torch.isfinite(got).all()
    err = (got - want).abs()
    assert (err <= 2.0 ** -7 * want.abs() + 1e-2).all(), float(err.max())


def test_strided_and_statistics_epilogues_stay_inside_their_tensors():
    """Sentinel arenas around the outputs of the per-parity stride-2 data gradient and of the statistics-fusing conv."""
    L = _lib()
    B, Ho, Wo, Cin, Cout, R = 5, 8, 8, 256, 512, 3
    g = torch.Generator().manual_seed(1)
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5).cuda().contiguous()
    dz = torch.randn(B, Ho, Wo, Cout, generator=g).bfloat16().cuda()
    packed = torch.empty(Cin * R * R * Cout, dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_pack_weights(w.data_ptr(), Cout, Cin, R, 1, packed.data_ptr(), _stream()))
    zero_bias = torch.zeros(512, dtype=torch.float32, device="cuda")
    n, guard = B * 2 * Ho * 2 * Wo * Cin, 8192
    arena = torch.full((guard + n + guard,), 3.0, dtype=torch.bfloat16, device="cuda")
    out = arena[guard:guard + n]
    L.check(L.lib().pb200_train_dgrad_s2(dz.data_ptr(), B, Ho, Wo, Cout, packed.data_ptr(), zero_bias.data_ptr(), Cin, R, None,
                                         out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert bool((arena[:guard] == 3.0).all()) and bool((arena[guard + n:] == 3.0).all()) and torch.isfinite(out.float()).all()
    # statistics-fusing forward conv: output and the 2*Cout doubles
    x = torch.randn(B, Ho, Wo, Cin, generator=g).bfloat16().cuda()
    L.check(L.lib().pb200_train_pack_weights(w.data_ptr(), Cout, Cin, R, 0, packed.data_ptr(), _stream()))
    n2 = B * Ho * Wo * Cout
    arena2 = torch.full((guard + n2 + guard,), 3.0, dtype=torch.bfloat16, device="cuda")
    sarena = torch.full((64 + 2 * Cout + 64,), -1.0, dtype=torch.float64, device="cuda")
    sarena[64:64 + 2 * Cout] = 0
    L.check(L.lib().pb200_train_conv(x.data_ptr(), B, Ho, Wo, Cin, packed.data_ptr(), zero_bias.data_ptr(), Cout, R, 1, 1, None, 0,
                                     arena2[guard:guard + n2].data_ptr(), sarena[64:].data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert bool((arena2[:guard] == 3.0).all()) and bool((arena2[guard + n2:] == 3.0).all())
    assert bool((sarena[:64] == -1).all()) and bool((sarena[64 + 2 * Cout:] == -1).all())


def test_pack_all_equals_the_single_packs():
    """pb200_train_pack_all (tile-staged, one launch for every conv) == pb200_train_pack_weights / _stem_pack per conv."""
    L = _lib()
    g = torch.Generator().manual_seed(5)
    specs = [(64, 4, 7, True), (64, 64, 3, False), (128, 64, 3, False), (128, 64, 1, False), (512, 256, 3, False)]
    rows, keep, start = [], [], 0
    for Cout, Cin, R, stem in specs:
        w = torch.randn(Cout, Cin, R, R, generator=g).cuda().contiguous()
        n = 64 * 256 if stem else Cout * Cin * R * R
        fwd = torch.full((n,), float("nan"), dtype=torch.bfloat16, device="cuda")
        dgr = None if stem else torch.full((n,), float("nan"), dtype=torch.bfloat16, device="cuda")
        tiles = 64 if stem else (Cout // 32) * (Cin // 32)
        rows.append([w.data_ptr(), fwd.data_ptr(), 0 if stem else dgr.data_ptr(), Cout, Cin, R, start, 1 if stem else 0])
        start += tiles
        keep.append((w, fwd, dgr))
    table = torch.tensor(rows, dtype=torch.int64, device="cuda")
    L.check(L.lib().pb200_train_pack_all(table.data_ptr(), len(specs), start, _stream()))
    torch.cuda.synchronize()
    for (Cout, Cin, R, stem), (w, fwd, dgr) in zip(specs, keep):
        ref = torch.empty_like(fwd)
        if stem:
            L.check(L.lib().pb200_train_stem_pack(w.data_ptr(), Cin, ref.data_ptr(), _stream()))
        else:
            L.check(L.lib().pb200_train_pack_weights(w.data_ptr(), Cout, Cin, R, 0, ref.data_ptr(), _stream()))
        torch.cuda.synchronize()
        assert torch.equal(fwd.view(torch.int16), ref.view(torch.int16)), (Cout, Cin, R, "forward operand")
        if dgr is not None:
            L.check(L.lib().pb200_train_pack_weights(w.data_ptr(), Cout, Cin, R, 1, ref.data_ptr(), _stream()))
            torch.cuda.synchronize()
            assert torch.equal(dgr.view(torch.int16), ref.view(torch.int16)), (Cout, Cin, R, "data-gradient operand")
