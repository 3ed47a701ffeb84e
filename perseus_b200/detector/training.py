"""Native data-parallel training step of the keypoint detector (reference: perseus/detector/train.py:259-304).

One `NativeTrainer.step(x, target)` = forward with batch-statistics BatchNorm -> SmoothL1Loss(beta=1) -> backward ->
(NCCL all-reduce of the flat gradient when a process group is given) -> clip_grad_norm_(1.0) -> AdamW, all on
hand-written sm_100a kernels through the C ABI (`pb200_train_*`):

  * convolutions: forward and data gradient on the tcgen05 implicit-GEMM kernels (the data gradient is the same
    kernel fed with the flipped/transposed filter; stride-2 layers convolve a zero-stuffed gradient), weight gradient
    on the MN-major tcgen05 GEMM of csrc/wgrad_umma.cu; conv1 is an explicit im2col GEMM;
  * BatchNorm statistics / apply / backward, ReLU masks, max-pool, head, loss and the fused optimizer: memory-bound
    SIMT kernels of csrc/train.cu.

Activations are NHWC bf16 (the reference trains under fp16 autocast, train.py:259-261; north_star asks for bf16),
parameters, gradients and optimizer state are fp32 in ONE flat buffer each, so the all-reduce is a single NCCL call
over NVLink and the optimizer a single kernel.  The module's parameters are re-pointed at views of the flat buffer:
`state_dict()`, checkpoints and the eval-mode native forward keep working on the trained weights.

There is no autograd and no PyTorch fallback on this path; torch is used for memory and `torch.distributed`.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from .. import _lib


class _Conv:
    """One conv + its BatchNorm: parameter views, packed operands and saved tensors of a step."""

    def __init__(self, name, conv, bn, Hin, Win, stem=False):
        self.name = name
        self.conv, self.bn = conv, bn
        self.Cout, self.Cin_real = conv.weight.shape[0], conv.weight.shape[1]
        self.R = conv.weight.shape[2]
        self.stride, self.pad = conv.stride[0], conv.padding[0]
        self.Hin, self.Win = Hin, Win
        self.Ho = (Hin + 2 * self.pad - self.R) // self.stride + 1
        self.Wo = (Win + 2 * self.pad - self.R) // self.stride + 1
        self.stem = stem
        self.Cin = 256 if stem else self.Cin_real  # the stem runs as a 1x1 conv over its im2col matrix (K = 256)


class NativeTrainer:
    def __init__(self, model, batch: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, max_norm: float = 1.0, l1_beta: float = 1.0, process_group=None,
                 use_graph: bool = True, broadcast_buffers: bool = True, overlap_allreduce: bool = True):
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise RuntimeError("NativeTrainer: the model must be on a B200 (cuda); there is no CPU training path")
        if model.H != 256 or model.W != 256:
            raise ValueError("NativeTrainer: the tcgen05 kernels are tiled for 256x256 frames")
        self.lib = _lib.lib()
        self.model = model
        self.dev = p0.device
        self.B = batch
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.max_norm, self.l1_beta = max_norm, l1_beta
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if (process_group is not None or dist.is_initialized()) else 1
        self.step_count = 0
        self.use_graph = use_graph   # forward + backward captured once into a CUDA graph (~250 launches per step)
        self._graph, self._eager_steps = None, 0
        # DistributedDataParallel semantics mirrored (train.py:152, default arguments):
        #  * broadcast_buffers=True: before every forward rank 0's BatchNorm running statistics replace every rank's
        #    (ONE broadcast of the flat 9,600-element buffer); batch statistics stay per rank (no SyncBN in the reference),
        #    so the running statistics follow rank 0's batches and rank 0's are what a checkpoint holds;
        #  * gradient all-reduce overlapped with backward: the flat gradient is reduced in two buckets, [layer4, fc]
        #    (80 % of the bytes, complete after the first two blocks of backward) on NCCL's stream under the rest of
        #    backward, then [conv1 .. layer3].
        self.broadcast_buffers = broadcast_buffers
        self.overlap_allreduce = overlap_allreduce and self.world > 1
        self._tail_work = None
        self._graph2 = None
        self.verify = None   # set to a dict by verify_next_step(): filled with the cross-rank checks of the next step
        self._flatten()
        self._build_graph()
        self._allocate()

    # ---- flat parameter / gradient / optimizer-state buffers ------------------------------------------------------
    def _flatten(self):
        params = list(self.model.parameters())
        n = sum(p.numel() for p in params)
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.flat_p = torch.empty(n, **f32)
        self.flat_g = torch.zeros(n, **f32)
        self.flat_m = torch.zeros(n, **f32)
        self.flat_v = torch.zeros(n, **f32)
        self.grad_of = {}
        off = 0
        for p in params:
            k = p.numel()
            self.flat_p[off:off + k].copy_(p.detach().reshape(-1).float())
            p.data = self.flat_p[off:off + k].view(p.shape)
            gview = self.flat_g[off:off + k].view(p.shape)
            p.grad = gview
            self.grad_of[id(p)] = gview
            off += k
        self.n_params = n
        # first element of the [layer4, fc] gradient bucket (model.parameters() order: conv1, bn1, layer1 .. layer4, fc)
        first4 = next(self.model.resnet.layer4.parameters())
        self.tail_off = 0
        for p in params:
            if p is first4:
                break
            self.tail_off += p.numel()
        # BatchNorm running statistics in ONE flat buffer (DDP broadcasts the module's buffers before each forward)
        bufs = [b for nme, b in self.model.named_buffers() if not nme.endswith("num_batches_tracked")]
        self.flat_bn = torch.empty(sum(b.numel() for b in bufs), **f32)
        off = 0
        for b in bufs:
            k = b.numel()
            self.flat_bn[off:off + k].copy_(b.detach().reshape(-1).float())
            b.data = self.flat_bn[off:off + k].view(b.shape)
            off += k
        if self.world > 1:  # DDP's initial parameter / buffer broadcast (train.py:152)
            dist.broadcast(self.flat_p, src=0, group=self.pg)
            dist.broadcast(self.flat_bn, src=0, group=self.pg)

    def _g(self, p):
        return self.grad_of[id(p)]

    # ---- layer graph (torchvision resnet18, resnet.py:266-282) ----------------------------------------------------
    def _build_graph(self):
        r = self.model.resnet
        H = self.model.H
        self.stem = _Conv("conv1", r.conv1, r.bn1, H, H, stem=True)
        self.blocks = []
        h = H // 4
        for li, layer in enumerate((r.layer1, r.layer2, r.layer3, r.layer4)):
            for bi, blk in enumerate(layer):
                c1 = _Conv(f"layer{li + 1}.{bi}.conv1", blk.conv1, blk.bn1, h, h)
                c2 = _Conv(f"layer{li + 1}.{bi}.conv2", blk.conv2, blk.bn2, c1.Ho, c1.Wo)
                ds = None
                if blk.downsample is not None:
                    ds = _Conv(f"layer{li + 1}.{bi}.downsample", blk.downsample[0], blk.downsample[1], h, h)
                self.blocks.append((c1, c2, ds))
                h = c1.Ho
        self.convs = [self.stem] + [c for b in self.blocks for c in b if c is not None]

    def _allocate(self):
        B, dev = self.B, self.dev
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        H = self.model.H
        self.A = torch.empty(B * (H // 2) * (H // 2), 256, **bf)  # conv1 im2col
        max_act = 0
        for c in self.convs:
            M = B * c.Ho * c.Wo
            c.M = M
            c.z = torch.empty(M, c.Cout, **bf)
            c.y = torch.empty(M, c.Cout, **bf)
            c.mean = torch.empty(c.Cout, **f32)
            c.invstd = torch.empty(c.Cout, **f32)
            c.w_fwd = torch.empty(c.Cout * c.R * c.R * c.Cin, **bf)
            c.w_dgrad = None if c.stem else torch.empty(c.Cout * c.R * c.R * c.Cin, **bf)
            max_act = max(max_act, M * c.Cout)
        s = self.stem
        self.pool = torch.empty(B * (s.Ho // 2) * (s.Wo // 2), 64, **bf)
        self.pool_idx = torch.empty(B * (s.Ho // 2) * (s.Wo // 2) * 64, dtype=torch.uint8, device=dev)
        # backward work buffers: gradients w.r.t. activations (ping-pong), conv outputs, identity branch, zero-stuffed
        act1 = max(c.M * c.Cout for c in self.convs if not c.stem)   # largest non-stem activation (layer1)
        self.dY = [torch.empty(act1, **bf) for _ in range(2)]
        self.dT = torch.empty(act1, **bf)
        self.dz = torch.empty(act1, **bf)
        self.dzd = torch.empty(act1, **bf)
        self.g_idn = torch.empty(act1, **bf)
        self.tmp = torch.empty(act1, **bf)
        # stride-2 data gradients: per-parity sub-convolutions (pb200_train_dgrad_s2); PB200_STUFFED_DGRAD=1 keeps the
        # older zero-stuffed formulation (4x the MACs, two extra tensors) for A/B measurements
        self.parity_dgrad = os.environ.get("PB200_STUFFED_DGRAD") is None
        stuffed = 8 if self.parity_dgrad else max([4 * c.M * c.Cout for c in self.convs if c.stride == 2 and not c.stem] + [8])
        self.stuffed = [torch.empty(stuffed, **bf) for _ in range(2)]
        self.dy0 = torch.empty(s.M * 64, **bf)   # gradient w.r.t. the stem activation (after max-pool backward)
        self.dz0 = torch.empty(s.M * 64, **bf)
        self.scratch = torch.zeros(9 * 512 * 512, **f32)   # wgrad partial sums [R*R][Cin][Cout]
        self.sums = torch.zeros(2 * 512, dtype=torch.float64, device=dev)
        # forward batch statistics, filled by the conv epilogues (two: a block's conv2 and its projection are in flight together)
        self.fstats = [torch.zeros(2 * 512, dtype=torch.float64, device=dev) for _ in range(2)]
        self.fused_stats = os.environ.get("PB200_NO_FUSED_STATS") is None
        self.sqnorm = torch.zeros(1, dtype=torch.float64, device=dev)
        self.zero_bias = torch.zeros(512, **f32)
        nk = 2 * self.model.n_keypoints
        self.feat = torch.empty(B, 512, **f32)
        self.pred = torch.empty(B, nk, **f32)
        self.dpred = torch.empty(B, nk, **f32)
        self.loss = torch.zeros(1, **f32)
        self.x_static = torch.zeros(B, self.model.num_channels, H, H, **f32)
        self.t_static = torch.zeros(B, nk, **f32)

    # ---- thin wrappers over the C ABI --------------------------------------------------------------------------
    def _s(self):
        return _lib.current_stream_ptr()

    _pack_table = None
    _prof = None  # when a list: (label, start event, end event) per wrapper call (see profile_step)

    def _mark(self, label):
        if self._prof is None:
            return None
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
        return (label, e0)

    def _done(self, tok):
        if tok is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            self._prof.append((tok[0], tok[1], e1))

    def profile_step(self, x, target):
        """One step with CUDA events around every operator class -> {label: ms} (adds event overhead; not a bench)."""
        self._prof = []
        with torch.cuda.device(self.dev):
            self.forward(x)
            self.backward(target)
            self.optimizer_step()
            torch.cuda.synchronize()
        out = {}
        for label, e0, e1 in self._prof:
            out[label] = out.get(label, 0.0) + e0.elapsed_time(e1)
        self._prof = None
        return out

    def _conv(self, x, B, H, W, Cin, w, Cout, R, stride, pad, out, residual=None, relu=0, stats=None):
        """stats: double [2][Cout] that the conv epilogue fills with the batch statistics of `out` (zeroed here)"""
        if stats is not None:
            stats.zero_()
        _lib.check(self.lib.pb200_train_conv(x.data_ptr(), B, H, W, Cin, w.data_ptr(), self.zero_bias.data_ptr(), Cout, R,
                                             stride, pad, _lib.ptr(residual), relu, out.data_ptr(), _lib.ptr(stats), self._s()))

    def _bn_forward(self, c, residual=None, relu=1, stats=None):
        """stats: the buffer the conv epilogue filled (fused statistics); None = a separate statistics pass over z"""
        bn = c.bn
        if stats is None:
            stats = self.sums
            _lib.check(self.lib.pb200_train_bn_stats(c.z.data_ptr(), c.M, c.Cout, stats.data_ptr(), self._s()))
        _lib.check(self.lib.pb200_train_bn_apply(
            c.z.data_ptr(), stats.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
            bn.running_var.data_ptr(), float(bn.momentum), float(bn.eps), c.mean.data_ptr(), c.invstd.data_ptr(),
            _lib.ptr(residual), relu, c.y.data_ptr(), c.M, c.Cout, self._s()))
        bn.num_batches_tracked += 1

    def _bn_backward(self, c, dy, mask, dz, g_out=None, stuffed=None):
        """mask: the layer output y (ReLU mask read from it), "z" (no residual input: mask recomputed from z), or None"""
        bn = c.bn
        from_z = isinstance(mask, str)
        _lib.check(self.lib.pb200_train_bn_backward(
            dy.data_ptr(), None if from_z else _lib.ptr(mask), c.z.data_ptr(), c.mean.data_ptr(), c.invstd.data_ptr(),
            bn.weight.data_ptr(), bn.bias.data_ptr() if from_z else None, self.sums.data_ptr(), c.M, c.Cout, dz.data_ptr(), _lib.ptr(g_out), _lib.ptr(stuffed), c.Ho, c.Wo,
            self._g(bn.weight).data_ptr(), self._g(bn.bias).data_ptr(), self._s()))

    def _wgrad(self, c, x, dz):
        n = c.R * c.R * c.Cin * c.Cout
        self.scratch[:n].zero_()
        if c.stem:  # 1x1 "conv" over the im2col matrix
            H = self.model.H // 2
            _lib.check(self.lib.pb200_train_wgrad(x.data_ptr(), dz.data_ptr(), self.B, H, H, 256, 64, 1, 1, 0,
                                                  self.scratch.data_ptr(), self._s()))
            _lib.check(self.lib.pb200_train_stem_unpack(self.scratch.data_ptr(), c.Cin_real,
                                                        self._g(c.conv.weight).data_ptr(), self._s()))
            return
        _lib.check(self.lib.pb200_train_wgrad(x.data_ptr(), dz.data_ptr(), self.B, c.Hin, c.Win, c.Cin, c.Cout, c.R, c.stride,
                                              c.pad, self.scratch.data_ptr(), self._s()))
        _lib.check(self.lib.pb200_train_unpack_wgrad(self.scratch.data_ptr(), c.Cout, c.Cin, c.R, c.Cin,
                                                     self._g(c.conv.weight).data_ptr(), self._s()))

    def _dgrad_s2(self, c, dz, out, residual00=None):
        """gradient w.r.t. the input of a stride-2 conv: one small stride-1 conv of dz per output parity, written to the
        strided pixels; a 1x1 layer only has the parity-(0,0) pixels (residual00 is added there)"""
        _lib.check(self.lib.pb200_train_dgrad_s2(dz.data_ptr(), self.B, c.Ho, c.Wo, c.Cout, c.w_dgrad.data_ptr(),
                                                 self.zero_bias.data_ptr(), c.Cin, c.R, _lib.ptr(residual00), out.data_ptr(),
                                                 self._s()))

    def _dgrad(self, c, dz_or_stuffed, out, residual=None):
        """gradient w.r.t. the conv input: stride-1 conv of dz (zero-stuffed for stride-2 layers) with the flipped filter"""
        _lib.check(self.lib.pb200_train_conv(dz_or_stuffed.data_ptr(), self.B, c.Hin, c.Win, c.Cout, c.w_dgrad.data_ptr(),
                                             self.zero_bias.data_ptr(), c.Cin, c.R, 1, c.R // 2, _lib.ptr(residual), 0,
                                             out.data_ptr(), None, self._s()))

    def _pack_weights(self):
        if self._pack_table is None:
            rows, start = [], 0
            for c in self.convs:
                # tiles of the pack kernel: 256 elements of conv1's [64][256] operand, else 32 x 32 channels x R*R taps
                n = 64 if c.stem else (c.Cout // 32) * (c.Cin // 32)
                rows.append([c.conv.weight.data_ptr(), c.w_fwd.data_ptr(), 0 if c.stem else c.w_dgrad.data_ptr(), c.Cout,
                             c.Cin_real if c.stem else c.Cin, c.R, start, 1 if c.stem else 0])
                start += n
            self._pack_table = torch.tensor(rows, dtype=torch.int64, device=self.dev)
            self._pack_total = start
        _lib.check(self.lib.pb200_train_pack_all(self._pack_table.data_ptr(), len(self.convs), self._pack_total, self._s()))

    # ---- the step ---------------------------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: (B, C, 256, 256) float32 cuda, NCHW as train.py:266-268 builds it -> predictions (B, 2K) float32."""
        B, m = self.B, self.model
        if tuple(x.shape) != (B, m.num_channels, m.H, m.W) or not x.is_cuda:
            raise ValueError(f"NativeTrainer.forward: expected a cuda ({B}, {m.num_channels}, {m.H}, {m.W}) batch")
        x = x.detach().to(torch.float32).contiguous()
        self._pack_weights()
        s = self.stem
        _lib.check(self.lib.pb200_train_stem_im2col(x.data_ptr(), B, m.num_channels, m.H, m.W, self.A.data_ptr(), self._s()))
        f0, f1 = (self.fstats if self.fused_stats else (None, None))
        self._conv(self.A, B, s.Ho, s.Wo, 256, s.w_fwd, 64, 1, 1, 0, s.z, stats=f0)
        self._bn_forward(s, relu=1, stats=f0)
        _lib.check(self.lib.pb200_train_maxpool(s.y.data_ptr(), self.pool.data_ptr(), self.pool_idx.data_ptr(), B, s.Ho, s.Wo,
                                                64, self._s()))
        cur = self.pool
        for c1, c2, ds in self.blocks:
            self._conv(cur, B, c1.Hin, c1.Win, c1.Cin, c1.w_fwd, c1.Cout, c1.R, c1.stride, c1.pad, c1.z, stats=f0)
            self._bn_forward(c1, relu=1, stats=f0)
            self._conv(c1.y, B, c2.Hin, c2.Win, c2.Cin, c2.w_fwd, c2.Cout, c2.R, 1, c2.pad, c2.z, stats=f0)
            idn = cur
            if ds is not None:
                self._conv(cur, B, ds.Hin, ds.Win, ds.Cin, ds.w_fwd, ds.Cout, 1, ds.stride, 0, ds.z, stats=f1)
                self._bn_forward(ds, relu=0, stats=f1)
                idn = ds.y
            self._bn_forward(c2, residual=idn, relu=1, stats=f0)
            c1.x_in = cur
            cur = c2.y
        last = self.blocks[-1][1]
        fc = m.resnet.fc
        _lib.check(self.lib.pb200_train_head_forward(cur.data_ptr(), fc.weight.data_ptr(), fc.bias.data_ptr(), self.feat.data_ptr(),
                                                     self.pred.data_ptr(), B, last.Ho * last.Wo, last.Cout, fc.weight.shape[0],
                                                     self._s()))
        return self.pred

    def backward(self, target: torch.Tensor) -> torch.Tensor:
        """SmoothL1Loss(pred, target) (train.py:119,290) and the gradient of every parameter into the flat buffer."""
        loss = self.backward_tail(target)
        self._after_tail()
        self.backward_rest()
        return loss

    def _after_tail(self):
        """the [layer4, fc] gradients are complete on the current stream: start their all-reduce under the rest of backward"""
        if self.overlap_allreduce and self.tail_off > 0:
            if self.verify is not None:
                self.verify["_tail_local"] = self.flat_g[self.tail_off:self.tail_off + self.verify["n"]].clone()
            self._tail_work = dist.all_reduce(self.flat_g[self.tail_off:], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)

    def backward_tail(self, target: torch.Tensor) -> torch.Tensor:
        """loss, head and layer4: everything whose parameter gradients live in flat_g[tail_off:]"""
        B, m = self.B, self.model
        fc = m.resnet.fc
        n_out = fc.weight.shape[0]
        t = target.detach().to(torch.float32).reshape(B, n_out).contiguous()
        _lib.check(self.lib.pb200_train_smooth_l1(self.pred.data_ptr(), t.data_ptr(), B * n_out, float(self.l1_beta),
                                                  self.loss.data_ptr(), self.dpred.data_ptr(), self._s()))
        last = self.blocks[-1][1]
        dY, other = self.dY
        _lib.check(self.lib.pb200_train_head_backward(self.dpred.data_ptr(), self.feat.data_ptr(), fc.weight.data_ptr(), B,
                                                      last.Ho * last.Wo, last.Cout, n_out, self._g(fc.weight).data_ptr(),
                                                      self._g(fc.bias).data_ptr(), dY.data_ptr(), self._s()))
        self._dY = (dY, other)
        self._backward_blocks(self.blocks[6:])
        return self.loss

    def backward_rest(self) -> None:
        """layer3 .. layer1 and the stem"""
        self._backward_blocks(self.blocks[:6])
        dY, _ = self._dY
        s = self.stem
        _lib.check(self.lib.pb200_train_maxpool_backward(dY.data_ptr(), self.pool_idx.data_ptr(), self.dy0.data_ptr(), self.B, s.Ho,
                                                         s.Wo, 64, self._s()))
        self._bn_backward(s, self.dy0, "z", self.dz0)
        self._wgrad(s, self.A, self.dz0)

    def _backward_blocks(self, blocks) -> None:
        dY, other = self._dY
        for c1, c2, ds in reversed(blocks):
            # out = relu(bn2(z2) + idn): g = dY * [out > 0] feeds bn2's backward and the identity branch
            self._bn_backward(c2, dY, c2.y, self.dz, g_out=self.g_idn)
            self._wgrad(c2, c1.y, self.dz)
            self._dgrad(c2, self.dz, self.dT)                      # gradient w.r.t. y1 (before its ReLU mask)
            st = c1.stride == 2
            par = self.parity_dgrad and st and ds is not None and ds.stride == 2
            self._bn_backward(c1, self.dT, "z", self.dz, stuffed=self.stuffed[0] if (st and not par) else None)
            self._wgrad(c1, c1.x_in, self.dz)
            if par:
                self._bn_backward(ds, self.g_idn, None, self.dzd)
                self._wgrad(ds, c1.x_in, self.dzd)
                self._dgrad_s2(ds, self.dzd, self.tmp)                       # parity-(0,0) pixels of tmp only
                self._dgrad_s2(c1, self.dz, other, residual00=self.tmp)      # all four parities; + tmp at (0,0)
            elif ds is not None:
                self._bn_backward(ds, self.g_idn, None, self.dzd, stuffed=self.stuffed[1] if ds.stride == 2 else None)
                self._wgrad(ds, c1.x_in, self.dzd)
                self._dgrad(ds, self.stuffed[1] if ds.stride == 2 else self.dzd, self.tmp)
                self._dgrad(c1, self.stuffed[0] if st else self.dz, other, residual=self.tmp)
            else:
                self._dgrad(c1, self.dz, other, residual=self.g_idn)
            dY, other = other, dY
        self._dY = (dY, other)

    def optimizer_step(self):
        """gradient all-reduce (NCCL, SUM; 1/world is folded into the optimizer kernel) + clip_grad_norm_ + AdamW
        (train.py:300-303).  With overlap the [layer4, fc] bucket has been in flight since backward_tail()."""
        if self.world > 1:
            vf = self.verify
            if self._tail_work is not None:
                if vf is not None:
                    vf["_head_local"] = self.flat_g[:vf["n"]].clone()
                self._tail_work.wait()  # the current stream waits for NCCL's
                self._tail_work = None
                dist.all_reduce(self.flat_g[:self.tail_off], op=dist.ReduceOp.SUM, group=self.pg)
            else:
                if vf is not None:
                    vf["_head_local"] = self.flat_g[:vf["n"]].clone()
                    vf["_tail_local"] = self.flat_g[self.tail_off:self.tail_off + vf["n"]].clone()
                dist.all_reduce(self.flat_g, op=dist.ReduceOp.SUM, group=self.pg)
            if vf is not None:
                self._finish_verify()
        self.step_count += 1
        _lib.check(self.lib.pb200_train_adamw_step(
            self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_m.data_ptr(), self.flat_v.data_ptr(), self.n_params,
            float(self.lr), float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.wd), self.step_count,
            1.0 / self.world, float(self.max_norm), self.sqnorm.data_ptr(), self._s()))
        self.model.refresh_weights()  # the eval-mode native forward re-folds BN from the updated parameters

    # ---- cross-rank checks (what DDP guarantees, train.py:130,152,300) ---------------------------------------------
    def verify_next_step(self, n: int = 1 << 20) -> dict:
        """Arms the checks for the NEXT step(); the returned dict is filled in by that step:
          grad_allreduce_max_rel_err : |all-reduced gradient - sum of the all-gathered per-rank gradients| over the first n
                                       elements of each bucket, relative to the largest entry
          params_identical           : bit-identical flat parameters on every rank after the optimizer step
          bn_buffers_follow_rank0    : every rank entered the step with rank 0's BatchNorm running statistics"""
        self.verify = {"n": int(min(n, self.tail_off if self.tail_off > 0 else self.n_params, self.n_params - self.tail_off))}
        return self.verify

    def _gather(self, t: torch.Tensor) -> list:
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t.contiguous(), group=self.pg)
        return out

    def _finish_verify(self) -> None:
        vf, n = self.verify, self.verify["n"]
        worst = 0.0
        for key, off in (("_head_local", 0), ("_tail_local", self.tail_off)):
            parts = self._gather(vf.pop(key))
            want = torch.stack(parts).double().sum(0)
            got = self.flat_g[off:off + n].double()
            worst = max(worst, float((got - want).abs().max() / want.abs().max().clamp_min(1e-30)))
        vf["grad_allreduce_max_rel_err"] = worst
        vf["_pending_params"] = True

    def _finish_verify_params(self) -> None:
        vf = self.verify
        # a checksum cannot prove bit equality; compare the whole buffer against rank 0's
        ref = self.flat_p.clone()
        dist.broadcast(ref, src=0, group=self.pg)
        same = torch.tensor([int(torch.equal(ref, self.flat_p))], device=self.dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN, group=self.pg)
        vf["params_identical"] = bool(same.item())
        vf.pop("_pending_params")
        vf["world"] = self.world
        self.verify = None

    def step(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """One training step; returns the (device) loss of this batch."""
        with torch.cuda.device(self.dev):
            if self.world > 1 and self.broadcast_buffers:
                dist.broadcast(self.flat_bn, src=0, group=self.pg)  # DDP's per-forward buffer broadcast (train.py:152)
                if self.verify is not None:
                    ref = self.flat_bn.clone()
                    dist.broadcast(ref, src=0, group=self.pg)
                    same = torch.tensor([int(torch.equal(ref, self.flat_bn))], device=self.dev)
                    dist.all_reduce(same, op=dist.ReduceOp.MIN, group=self.pg)
                    self.verify["bn_buffers_follow_rank0"] = bool(same.item())
            if not self.use_graph:
                self.forward(x)
                loss = self.backward(target)
            else:
                self.x_static.copy_(x, non_blocking=True)
                self.t_static.copy_(target.reshape(self.B, -1), non_blocking=True)
                if self._graph is not None:
                    self._graph.replay()
                    if self._graph2 is not None:  # the [layer4, fc] all-reduce runs under the second half of backward
                        self._after_tail()
                        self._graph2.replay()
                elif self._eager_steps < 2:  # first steps eagerly: one-off attribute / occupancy queries happen here
                    self.forward(self.x_static)
                    self.backward(self.t_static)
                    self._eager_steps += 1
                elif self.overlap_allreduce:
                    g1, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g1):
                        self.forward(self.x_static)
                        self.backward_tail(self.t_static)
                    with torch.cuda.graph(g2, pool=g1.pool()):
                        self.backward_rest()
                    self._graph, self._graph2 = g1, g2
                    g1.replay()
                    self._after_tail()
                    g2.replay()
                else:
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        self.forward(self.x_static)
                        self.backward(self.t_static)
                    self._graph = graph
                    graph.replay()
                loss = self.loss
            self.optimizer_step()
            if self.verify is not None and self.verify.get("_pending_params"):
                self._finish_verify_params()
        return loss


def _profiled(label_fn, fn):
    def wrapper(self, *a, **k):
        tok = self._mark(label_fn(*a, **k))
        out = fn(self, *a, **k)
        self._done(tok)
        return out
    wrapper.__doc__ = fn.__doc__
    return wrapper


def _layer_class(c):
    return "stem" if c.stem else c.name.split(".")[0]


NativeTrainer._conv = _profiled(lambda x, B, H, W, Cin, *a, **k: f"conv_fwd[{'stem' if Cin == 256 and H == 128 else str(H)}]",
                                NativeTrainer._conv)
NativeTrainer._bn_forward = _profiled(lambda c, **k: f"bn_fwd[{_layer_class(c)}]", NativeTrainer._bn_forward)
NativeTrainer._bn_backward = _profiled(lambda c, *a, **k: f"bn_bwd[{_layer_class(c)}]", NativeTrainer._bn_backward)
NativeTrainer._wgrad = _profiled(lambda c, *a, **k: f"wgrad[{_layer_class(c)}]", NativeTrainer._wgrad)
NativeTrainer._dgrad = _profiled(lambda c, *a, **k: f"dgrad[{_layer_class(c)}]", NativeTrainer._dgrad)
NativeTrainer._pack_weights = _profiled(lambda: "pack_weights", NativeTrainer._pack_weights)
