"""pb200_plasma_static: the plasma fractal with its scratch in a static device array (no allocation inside the call).

No counterpart in the reference; added (callers that cannot call cudaMalloc).  The
new entry point must reproduce pb200_plasma -- same kernel, same arithmetic: bit-equal -- on two calls in a row (the shared
scratch keeps the previous call's contents), must reject problems beyond PB200_PLASMA_STATIC_MAX_FLOATS, and must not
touch the CUDA allocator.  Named test_zz_* so that pytest collects it last.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 16 * 257 * 257  # PB200_PLASMA_STATIC_MAX_FLOATS (include/perseus_b200.h)


def _setup():
    from perseus_b200 import _lib

    if not torch.cuda.is_available():
        pytest.skip("needs a B200")
    return _lib, _lib.lib(), torch.device("cuda", torch.cuda.current_device())


def _with_caller_scratch(_lib, lib, dev, rough, seed, B, H, W):
    S = lib.pb200_plasma_side(H, W)
    scratch = torch.full((B, S, S), float("nan"), dtype=torch.float32, device=dev)
    out = torch.full((B, H, W), float("nan"), dtype=torch.float32, device=dev)
    _lib.check(lib.pb200_plasma(rough.data_ptr(), ctypes.c_uint64(seed), scratch.data_ptr(), out.data_ptr(), B, H, W,
                                _lib.current_stream_ptr()))
    return out


def test_plasma_static_equals_plasma_on_two_calls_in_a_row_without_allocating():
    _lib, lib, dev = _setup()
    cases = [(16, 256, 256, 11), (5, 37, 70, 12)]  # the maximum (sixteen 257^2 grids), then a smaller ragged one
    inputs, want = [], []
    for B, H, W, seed in cases:
        g = torch.Generator().manual_seed(seed)
        rough = (0.1 + 0.6 * torch.rand(B, generator=g)).to(dev)
        inputs.append(rough)
        want.append(_with_caller_scratch(_lib, lib, dev, rough, seed, B, H, W))
    outs = [torch.full((B, H, W), float("nan"), dtype=torch.float32, device=dev) for B, H, W, _ in cases]
    # one untimed call first: the first call on a device resolves the static array's address
    _lib.check(lib.pb200_plasma_static(inputs[1].data_ptr(), ctypes.c_uint64(99), outs[1].data_ptr(), 5, 37, 70,
                                       _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    for rep in range(2):  # two rounds back to back: every call starts from the previous call's scratch contents
        for out in outs:
            out.fill_(float("nan"))
        torch.cuda.synchronize()
        # the allocator counters bracket ONLY the entry-point calls (the comparisons below allocate temporaries)
        allocs0 = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
        free0 = torch.cuda.mem_get_info(dev)[0]
        for (B, H, W, seed), rough, out in zip(cases, inputs, outs):
            _lib.check(lib.pb200_plasma_static(rough.data_ptr(), ctypes.c_uint64(seed), out.data_ptr(), B, H, W,
                                               _lib.current_stream_ptr()))
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == allocs0, "torch allocations inside the calls"
        assert torch.cuda.mem_get_info(dev)[0] >= free0, "device memory was taken inside the calls"
        for out, ref in zip(outs, want):
            assert torch.isfinite(out).all()
            assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0
            assert torch.equal(out, ref), f"round {rep}: max diff {float((out - ref).abs().max()):.3e}"


def test_plasma_static_rejects_sizes_beyond_its_scratch():
    _lib, lib, dev = _setup()
    rough = torch.full((17,), 0.5, dtype=torch.float32, device=dev)
    out = torch.full((17, 256, 256), 7.5, dtype=torch.float32, device=dev)
    S = lib.pb200_plasma_side(256, 256)
    assert 16 * S * S == MAX_FLOATS and 17 * S * S > MAX_FLOATS
    rc = lib.pb200_plasma_static(rough.data_ptr(), ctypes.c_uint64(1), out.data_ptr(), 17, 256, 256, _lib.current_stream_ptr())
    assert rc != 0
    msg = lib.pb200_last_error().decode()
    assert "plasma_static" in msg and str(MAX_FLOATS) in msg, msg
    # one image, but a grid larger than the whole array: S = 2049, 2049^2 > 16 * 257^2
    rc = lib.pb200_plasma_static(rough.data_ptr(), ctypes.c_uint64(1), out.data_ptr(), 1, 1026, 600, _lib.current_stream_ptr())
    assert rc != 0
    torch.cuda.synchronize()
    assert bool((out == 7.5).all()), "a rejected call must not launch"
    # B = 0 is accepted and does nothing
    _lib.check(lib.pb200_plasma_static(rough.data_ptr(), ctypes.c_uint64(1), out.data_ptr(), 0, 256, 256, _lib.current_stream_ptr()))
    assert np.isfinite(out.cpu().numpy()).all()
