"""Runs the BatchNorm training kernels on stem- and layer1-sized tensors (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from perseus_b200 import _lib as L

lib = L.lib()
st = L.current_stream_ptr()
for M, C in ((256 * 128 * 128, 64), (256 * 64 * 64, 64), (256 * 16 * 16, 256)):
    z = torch.randn(M, C, device="cuda").bfloat16()
    y = torch.empty_like(z)
    dy = torch.randn(M, C, device="cuda").bfloat16()
    dz = torch.empty_like(z)
    g = torch.empty_like(z)
    gamma, beta = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    mean, invstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    sums = torch.zeros(2 * C, dtype=torch.float64, device="cuda")
    for it in range(2):
        L.check(lib.pb200_train_bn_stats(z.data_ptr(), M, C, sums.data_ptr(), st))
        L.check(lib.pb200_train_bn_apply(z.data_ptr(), sums.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                         0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), None, 1, y.data_ptr(), M, C, st))
        L.check(lib.pb200_train_bn_backward(dy.data_ptr(), None, z.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                                            beta.data_ptr(), sums.data_ptr(), M, C, dz.data_ptr(), None, None, 0, 0, dg.data_ptr(),
                                            db.data_ptr(), st))
        L.check(lib.pb200_train_bn_backward(dy.data_ptr(), y.data_ptr(), z.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                                            None, sums.data_ptr(), M, C, dz.data_ptr(), g.data_ptr(), None, 0, 0, dg.data_ptr(),
                                            db.data_ptr(), st))
    torch.cuda.synchronize()
    del z, y, dy, dz, g
print("ok")
