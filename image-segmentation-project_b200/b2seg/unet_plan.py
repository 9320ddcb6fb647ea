"""Static execution plan of one UNetWithBackbone training / inference step on libb2seg.

The reference runs `model(images)` / `loss.backward()` (train.py:43-48) through ~400 ATen
calls per step.  Here the network (advanced_models.py:264-357 + torchvision BasicBlock /
Bottleneck) is lowered ONCE per (batch, height, width, mode) into two flat lists of planned
libb2seg ops -- forward and backward -- over statically allocated NHWC bf16 buffers:

  * every Conv2d / ConvTranspose2d and its data / weight gradient is a tcgen05 op (convops),
  * BatchNorm statistics come out of the conv epilogue; normalise+ReLU(+residual) is one pass,
  * `torch.cat((skip, up), 1)` never happens: both producers write channel slices of one buffer,
  * parameter gradients land in one flat fp32 buffer (one NCCL all-reduce in data-parallel runs).

Replaying a plan is launch-only, so a whole step is CUDA-graph capturable.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L
from . import convops as co
from . import ewops as eo
from .convops import View

BF16 = torch.bfloat16
F32 = torch.float32


class Act:
    """An activation: value view, gradient view (or None) and whether the gradient view has
    been written yet (the first producer stores, later ones accumulate)."""

    __slots__ = ("v", "g", "gw")

    def __init__(self, v: View, g: View | None = None):
        self.v, self.g, self.gw = v, g, False

    @property
    def P(self):
        return self.v.n * self.v.h * self.v.w

    def take_grad(self) -> bool:
        """Returns the accumulate flag for the next writer of this gradient."""
        acc = self.gw
        self.gw = True
        return acc


class BNRef:
    """One BatchNorm2d of the plan: its parameter prefix, the [4][C] constants buffer (scale,
    shift, mean, invstd) the backward pass reads, and -- in training mode, until a consumer has
    folded them -- the conv epilogue's partial statistics."""

    __slots__ = ("prefix", "C", "bnp", "stats", "rows", "count", "ready")

    def __init__(self, prefix, C_, bnp, stats, rows, count, ready):
        self.prefix, self.C, self.bnp, self.stats = prefix, C_, bnp, stats
        self.rows, self.count, self.ready = rows, count, ready


_BN_FUSED_FWD_DEFAULT = "2"
_BN_FIXED_BWD_DEFAULT = "0"
_FX_CAP = 1 << 18  # int64 words of fixed-point BatchNorm accumulators per plan (2 x the BatchNorm channels)


def _fused_bn_fwd():
    """EW_BN_APPLY_FUSED: no finalize launch, the consumer of a BatchNorm derives scale / shift itself.
    B2SEG_BN_FUSED_FWD = 0: finalize kernel (fp32 partial rows, one per CTA of the conv);
    1: the consumer folds those <= 296 x 2C floats in every block (costs more than the launch it saves,
       DESIGN 3.7);
    2: the conv epilogues add their partial sums as fixed-point integers into ONE int64 row with atomics
       (exact and order-independent, so still bit-reproducible) and the consumer reads 2C words."""
    import os
    return os.environ.get("B2SEG_BN_FUSED_FWD", _BN_FUSED_FWD_DEFAULT)


def _fixed_bn_bwd():
    """B2SEG_BN_FIXED_BWD=1: the BatchNorm-backward reduce adds its block sums (sum dz, sum dz*y) as fixed-point
    integers (x 2^40) into one int64 row with atomics and the apply kernel derives its per-channel constants (and
    block 0 dgamma / dbeta) from it -- no finalize launch.  Exact and order-independent like the forward
    statistics; range |block sum| < 2^22, resolution 9e-13."""
    import os
    return os.environ.get("B2SEG_BN_FIXED_BWD", _BN_FIXED_BWD_DEFAULT) == "1"


def _fused_bn_bwd_max_bytes():
    """EW_BN_BWD_FUSED is planned for BatchNorm layers whose pre-BN tensor is at most this many
    bytes (B2SEG_BN_FUSED_BWD_MB; 0 = never, the default).  It won up to 12 MB while the reduce / apply
    kernels ran as several waves of blocks; with one-resident-wave grids and batched loads the three-kernel
    sequence is faster at every size measured (64 x 256^2: 7.74 vs 7.80 ms/step; 4 x 256^2: 3.67 vs 3.75):
    a device-wide barrier costs more than two programmatic dependent launches."""
    import os
    return float(os.environ.get("B2SEG_BN_FUSED_BWD_MB", "0")) * 2 ** 20


def _grad_group(name: str) -> int:
    """Order in which the backward pass finishes parameter gradients: 0 = decoder, attention gates
    and head, 1..3 = enc4..enc2, 4 = enc1 and the stem."""
    for g, pre in enumerate(("enc4.", "enc3.", "enc2."), start=1):
        if name.startswith(pre):
            return g
    return 4 if name.startswith(("enc1.", "input_conv.", "bn1.")) else 0


class UNetPlan:
    def __init__(self, model, N, H, W, training, need_grad):
        if H % 32 or W % 32:
            raise ValueError(
                f"UNetWithBackbone (B200 path) needs H and W to be multiples of 32, got {H}x{W}; "
                "the reference's crop logic (advanced_models.py:317-347) only accepts such sizes too")
        self.N, self.H, self.W = N, H, W
        self.training, self.need_grad = bool(training), bool(need_grad)
        self.dev = next(model.parameters()).device
        self.params = dict(model.named_parameters())
        self.buffers = dict(model.named_buffers())
        self.backbone = model.backbone_name
        self.attn = bool(model.use_attention)
        self.fwd_ops, self.bwd_ops, self.tape = [], [], []
        self._gview_log, self.buckets = None, []
        self.stored_g = {}  # name -> View of the gradient w.r.t. that activation (after backward)
        self.stored = {}  # name -> View / fp32 tensor of every activation kept in HBM (parity tests)
        self.keep = []
        self.generation = 0
        self.nbytes = 0
        for n, p in self.params.items():
            if p.dtype != F32 or not p.is_contiguous() or p.device != self.dev:
                raise L.B2SegError(f"parameter {n} must be a contiguous fp32 tensor on {self.dev}")
        self.param_ptrs = tuple(p.data_ptr() for p in self.params.values())
        # flat gradient buffer, one view per parameter (64-element aligned)
        # The buffer is laid out in the order the backward pass FINISHES gradients (decoder, gates and
        # head first, then enc4 .. enc1 and the stem), so that the slice of each group is contiguous
        # and a data-parallel run can all-reduce it while the rest of the backward still runs.
        if self.need_grad:
            off, self.goff, self.ggroup = 0, {}, {}
            order = sorted(self.params, key=lambda n: (_grad_group(n), list(self.params).index(n)))
            self.group_range = {}
            for n in order:
                p, g = self.params[n], _grad_group(n)
                self.goff[n], self.ggroup[n] = off, g
                lo, _ = self.group_range.get(g, (off, off))
                off += (p.numel() + 63) // 64 * 64
                self.group_range[g] = (lo, off)
            self.gflat = self._alloc((off,), F32, zero=True)
            self._gview_log, self.bwd_segs, self.buckets = [], None, None
        # bf16 GEMM-operand copies of the conv weights
        tot = sum(2 * ((p.numel() + 63) // 64 * 64) for p in self.params.values() if p.dim() == 4)
        self.packbuf = self._alloc((tot + 64,), BF16)
        self.pack_off, self.pack_jobs, self.pack_cache = 0, [], {}
        self._ws_req = 0
        # shared block-partial scratch (BatchNorm backward, column sums, 1-channel convs): every
        # use is consumed by the op that follows it on the stream
        self.part = self._alloc((L.EW_MAX_PARTIAL_ROWS * 2 * 2048 + 1024,), F32)
        self.part_side = None
        self.x_in = self._alloc((N, H, W), F32)
        self.logits = self._alloc((N, 1, H, W), F32)
        if self.need_grad:
            self.dlogits = self._alloc((N, 1, H, W), F32)
        self._build()

    # ------------------------------------------------------------------ memory ----
    def _alloc(self, shape, dtype, zero=False):
        t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.dev)
        self.nbytes += t.numel() * t.element_size()
        self.keep.append(t)
        return t

    def _fx_alloc(self, n):
        """n int64 words of the plan's fixed-point accumulator arena (zeroed by ONE fill at the head of
        every forward pass, `_fx_fill_op`)."""
        if getattr(self, "_fx", None) is None:
            self._fx = self._alloc((_FX_CAP,), torch.int64, zero=True)
            self._fx_used = 0
        o = self._fx_used
        self._fx_used = o + ((n + 1) & ~1)  # 16-byte granules
        if self._fx_used > _FX_CAP:
            raise RuntimeError("fixed-point BatchNorm accumulator arena exhausted")
        return self._fx[o:o + n]

    def _fxb_alloc(self, n):
        """The same for the backward pass (its own arena and fill, so that a second backward after one forward
        starts from zero too)."""
        if getattr(self, "_fxb", None) is None:
            self._fxb = self._alloc((_FX_CAP,), torch.int64, zero=True)
            self._fxb_used = 0
        o = self._fxb_used
        self._fxb_used = o + ((n + 1) & ~1)
        if self._fxb_used > _FX_CAP:
            raise RuntimeError("fixed-point BatchNorm accumulator arena exhausted")
        return self._fxb[o:o + n]

    def act(self, n, h, w, c, grad=None):
        """New dense activation (and gradient buffer when a backward pass is planned)."""
        v = View(self._alloc((n, h, w, c), BF16))
        g = None
        if self.need_grad and grad is not False:
            g = View(self._alloc((n, h, w, c), BF16))
        return Act(v, g)

    def gview(self, name):
        p = self.params[name]
        o = self.goff[name]
        if self._gview_log is not None:
            self._gview_log.append(name)
        return self.gflat[o:o + p.numel()].view(p.shape)

    def packed(self, name, layout):
        """bf16 operand of parameter `name`: 'fprop' / 'dgrad' for Conv2d (OIHW), 'Tfprop' /
        'Tdgrad' for ConvTranspose2d (IOHW).  Filled by the PACK op at the head of the plan."""
        key = (name, layout)
        if key in self.pack_cache:
            return self.pack_cache[key]
        p = self.params[name]
        d0, d1, kh, kw = p.shape
        K = kh * kw
        if layout == "fprop":      # out[t][o][i] = W[o][i][t]
            T, A, B, sa, sb = K, d0, d1, d1 * K, K
        elif layout == "dgrad":    # out[t][i][o]
            T, A, B, sa, sb = K, d1, d0, K, d1 * K
        elif layout == "Tfprop":   # W[i][o][q] -> out[q][o][i]  (== [1][4*O][I])
            T, A, B, sa, sb = K, d1, d0, K, d1 * K
        elif layout == "Tdgrad":   # out[q][i][o]
            T, A, B, sa, sb = K, d0, d1, d1 * K, K
        else:
            raise KeyError(layout)
        n = T * A * B
        out = self.packbuf[self.pack_off:self.pack_off + n]
        self.pack_off += (n + 63) // 64 * 64
        self.pack_jobs.append((p, out, T, A, B, sa, sb, 1))
        if layout == "Tfprop":
            t = out.view(1, T * A, B)
        else:
            t = out.view(T, A, B)
        self.pack_cache[key] = t
        return t

    # ------------------------------------------------------------- op emitters ----
    def F(self, op):
        self.fwd_ops.append(op)

    def B(self, op):
        self.bwd_ops.append(op)

    def wgrad_ws(self, a_ch, b_ch, ntaps, n, h, w):
        """Register a weight-gradient GEMM's split-K workspace need (allocated before the tape runs)."""
        if self.need_grad:
            self._ws_req = max(self._ws_req, co.wgrad_workspace_bytes(a_ch, b_ch, ntaps, n, h, w))

    def bn_finalize(self, prefix, C_, rows, count, partial, defer=False) -> BNRef:
        """BatchNorm statistics -> BNRef.  Eval mode (running statistics) and `defer=False` emit the
        finalize kernel here; in training mode with `defer=True` the ONE consumer of the statistics
        (`bn_apply`) folds the partial rows itself, updates the running statistics and publishes
        bnp [4][C] (scale, shift, mean, invstd) -- one launch less per BatchNorm."""
        if self.training and count <= 1:
            raise ValueError(f"Expected more than 1 value per channel when training ({prefix})")
        bnp = self._alloc((4, C_), F32)
        if defer and self.training and C_ % 8 == 0 and (rows < 0 or _fused_bn_fwd() == "1"):
            return BNRef(prefix, C_, bnp, partial, rows, count, False)
        assert rows >= 0, "fixed-point statistics have no finalize kernel"
        self._emit_finalize(prefix, C_, rows, count, partial, bnp)
        return BNRef(prefix, C_, bnp, partial, rows, count, True)

    def _emit_finalize(self, prefix, C_, rows, count, partial, bnp):
        self.F(eo.ew(L.EW_BN_FINALIZE,
                     p=[partial if self.training else None, self.params[prefix + ".weight"],
                        self.params[prefix + ".bias"], self.buffers[prefix + ".running_mean"],
                        self.buffers[prefix + ".running_var"],
                        self.buffers[prefix + ".num_batches_tracked"], bnp],
                     i=[rows, C_, count, 1 if self.training else 0], f=[1e-5, 0.1]))

    def _bn_ptrs(self, bn: BNRef):
        pr = bn.prefix
        return [bn.stats, self.params[pr + ".weight"], self.params[pr + ".bias"],
                self.buffers[pr + ".running_mean"], self.buffers[pr + ".running_var"],
                self.buffers[pr + ".num_batches_tracked"], bn.bnp]

    def bn_apply(self, y: View, bn: BNRef, out: View, relu, res: View | None = None, rbn: BNRef | None = None):
        """out = act(BN(y) (+ res | + BN_r(res))).  Statistics still pending on `bn` (and `rbn`) are
        folded inside the same launch (EW_BN_APPLY_FUSED)."""
        P = y.n * y.h * y.w
        if not bn.ready:
            if rbn is not None and rbn.ready:
                raise AssertionError("a pending BatchNorm cannot take a finalized residual BatchNorm")
            pp = [y, out, res] + self._bn_ptrs(bn) + (self._bn_ptrs(rbn) if rbn is not None else [])
            self.F(eo.ew(L.EW_BN_APPLY_FUSED, p=pp,
                         i=[P, y.c, y.sw, out.sw, res.sw if res else 0, 1 if relu else 0, bn.rows, bn.count,
                            rbn.rows if rbn is not None else 0], f=[1e-5, 0.1]))
            bn.ready = True
            if rbn is not None:
                rbn.ready = True
            return
        if rbn is not None and not rbn.ready:
            self._emit_finalize(rbn.prefix, rbn.C, rbn.rows, rbn.count, rbn.stats, rbn.bnp)
            rbn.ready = True
        self.F(eo.ew(L.EW_BN_APPLY, p=[y, bn.bnp, out, res, rbn.bnp if rbn is not None else None],
                     i=[P, y.c, y.sw, out.sw, res.sw if res else 0, 1 if relu else 0]))

    def conv_bn(self, x: View, wname, bnprefix, k, stride, cout, biasname=None):
        """Conv2d (+bias) with BatchNorm partial statistics in the epilogue, then the statistics
        finalize.  Returns (y pre-BN, bnp)."""
        N, ho, wo = x.n, x.h // stride, x.w // stride
        y = View(self._alloc((N, ho, wo, cout), BF16))
        rows = co.stat_rows(N, ho, wo)
        if self.training and cout % 8 == 0 and _fused_bn_fwd() == "2":
            rows, stats = -1, self._fx_alloc(2 * cout)
        else:
            stats = self._alloc((rows, 2, cout), F32) if self.training else None
        bias = self.params[biasname] if biasname else None
        self.F(co.conv_fprop(x, self.packed(wname, "fprop"), k, stride, y, bias=bias, stats=stats))
        self.wgrad_ws(x.c, cout, k * k, N, ho, wo)
        return y, self.bn_finalize(bnprefix, cout, rows, N * ho * wo, stats, defer=True)

    def bn_backward(self, prefix, da: View, mask, y: View, bn, dy: View, da2=None, dz=None,
                    dz_acc=False, dbias=None):
        """dz = (da [+ da2]) * [mask > 0];  dy = BatchNorm-backward(dz);  dgamma, dbeta (, dbias).
        One persistent launch (EW_BN_BWD_FUSED: reduce, device-wide barrier, distributed fold,
        barrier, apply) for layers up to _fused_bn_bwd_max_bytes(); the reduce / finalize / apply
        sequence above that."""
        bnp = bn.bnp if isinstance(bn, BNRef) else bn
        P, C_ = y.n * y.h * y.w, y.c
        rows = eo.partial_rows(L.EW_BN_BWD_REDUCE, P, C_)
        self_mask = isinstance(mask, str)  # "self": a = relu(BN(y)), the mask is recomputed from y
        if self_mask:
            assert mask == "self"
            mask = None
        pp = [da, da2, mask, y, bnp]
        ii = [P, C_, da.sw, da2.sw if da2 else 0, mask.sw if mask else 0, y.sw]
        if P * C_ * 2 <= _fused_bn_bwd_max_bytes():
            self.B(eo.ew(L.EW_BN_BWD_FUSED,
                         p=pp + [self.part, dy, dz, self.params[prefix + ".weight"], self.gview(prefix + ".weight"),
                                 self.gview(prefix + ".bias"), dbias],
                         i=ii + [dy.sw, dz.sw if dz else 0, 2 if dz_acc else 1, 1 if self_mask else 0,
                                 1 if self.training else 0, 0]))
            return
        if _fixed_bn_bwd() and C_ % 8 == 0 and 256 % (C_ // 8) == 0:
            acc = self._fxb_alloc(2 * C_)
            self.B(eo.ew(L.EW_BN_BWD_REDUCE, p=pp + [acc], i=ii + [1, 0, 0, 1 if self_mask else 0]))  # i6: fixed
            self.B(eo.ew(L.EW_BN_BWD_APPLY,
                         p=pp + [acc, dy, dz, self.params[prefix + ".weight"], self.gview(prefix + ".weight"),
                                 self.gview(prefix + ".bias"), dbias],
                         i=ii + [dy.sw, dz.sw if dz else 0, 2 if dz_acc else 1, 1 if self_mask else 0,
                                 1 | (2 if self.training else 0)]))
            return
        tail = [0, 0, 0, 1 if self_mask else 0]  # i6..i8 unused by REDUCE, i9 = self-mask flag
        self.B(eo.ew(L.EW_BN_BWD_REDUCE, p=pp + [self.part], i=ii + tail))
        coef = self._alloc((3, C_), F32)
        self.B(eo.ew(L.EW_BN_BWD_FINALIZE,
                     p=[self.part, self.params[prefix + ".weight"], bnp, self.gview(prefix + ".weight"),
                        self.gview(prefix + ".bias"), coef, dbias],
                     i=[rows, C_, P, 0, 1 if self.training else 0, 1]))  # i5: partial holds sum dz*y
        self.B(eo.ew(L.EW_BN_BWD_APPLY, p=pp + [coef, dy, dz],
                     i=ii + [dy.sw, dz.sw if dz else 0, 2 if dz_acc else 1, 1 if self_mask else 0]))

    def conv_backward(self, x: Act, dy: View, wname, k, stride, want_dx=True):
        """Weight gradient into the flat buffer and (accumulating) data gradient into x.g."""
        self.B(co.conv_wgrad(x.v, dy, k, stride, self.gview(wname), self.ws))
        if want_dx and x.g is not None:
            self.B(co.conv_dgrad(dy, self.packed(wname, "dgrad"), k, stride, x.g, accumulate=x.take_grad()))

    def colsum_to(self, x: View, gname):
        """grad[gname][c] = sum over pixels of x[:, c] (ConvTranspose2d bias gradient)."""
        P = x.n * x.h * x.w
        rows = eo.partial_rows(L.EW_COLSUM, P, x.c)
        # a parameter gradient: both ops ride the side branch with the weight gradients (which run
        # in order among themselves), on partial rows of their own
        if self.part_side is None:
            self.part_side = self._alloc((L.EW_MAX_PARTIAL_ROWS * 2048 + 1024,), F32)
        for op in (eo.ew(L.EW_COLSUM, p=[x, self.part_side], i=[P, x.c, x.sw]),
                   eo.ew(L.EW_COLSUM_FINALIZE, p=[self.part_side, self.gview(gname)], i=[rows, x.c, 0, x.c])):
            op.side = True
            self.B(op)

    # ------------------------------------------------------------ network walk ----
    def _build(self):
        N, H, W, ng = self.N, self.H, self.W, self.need_grad
        P = self.params
        r50 = self.backbone == "resnet50"
        widths = (256, 512, 1024, 2048) if r50 else (64, 128, 256, 512)
        self.up_out = {4: widths[2], 3: widths[1], 2: widths[0], 1: 64 if r50 else 32}
        self.skip_c = {4: widths[2], 3: widths[1], 2: widths[0], 1: 64}
        # concat buffers [skip | up] of decoder stage k (resolution H / 2^k) and their gradients
        self.cat, self.dcat = {}, {}
        for k in (4, 3, 2, 1):
            shape = (N, H >> k, W >> k, self.skip_c[k] + self.up_out[k])
            self.cat[k] = self._alloc(shape, BF16)
            self.dcat[k] = self._alloc(shape, BF16) if ng else None

        def skip_act(k):
            """Encoder output feeding decoder stage k: lives inside the concat buffer unless an
            attention gate rescales it on the way (advanced_models.py:287-288)."""
            if self.attn:
                return self.act(N, H >> k, W >> k, self.skip_c[k])
            return Act(View(self.cat[k], 0, self.skip_c[k]),
                       View(self.dcat[k], 0, self.skip_c[k]) if ng else None)

        # ---- stem: input_conv 7x7 s2 + bn1 + relu (:268-270), maxpool 3x3 s2 (:272)
        h1, w1 = H // 2, W // 2
        x1 = skip_act(1)
        y0 = View(self._alloc((N, h1, w1, 64), BF16))
        rows0 = eo.stem_stat_rows(N, H, W)
        st0 = self._alloc((rows0, 2, 64), F32) if self.training else None
        self.F(eo.ew(L.EW_STEM_FWD, p=[self.x_in, P["input_conv.weight"], y0, st0], i=[N, H, W]))
        bnp0 = self.bn_finalize("bn1", 64, rows0, N * h1 * w1, st0, defer=True)
        self.bn_apply(y0, bnp0, x1.v, True)
        self.stored.update({"stem.y": y0, "x1": x1.v})
        h0 = self.act(N, h1 // 2, w1 // 2, 64)
        idx = self._alloc((N, h1 // 2, w1 // 2, 64), torch.uint8)
        self.F(eo.ew(L.EW_MAXPOOL_FWD, p=[x1.v, h0.v, idx], i=[N, h1, w1, 64, x1.v.sw, h0.v.sw]))

        def bwd_stem():
            dy0 = View(self._alloc((N, h1, w1, 64), BF16))
            self.bn_backward("bn1", x1.g, "self", y0, bnp0, dy0)
            self.stored_g.update({"stem.y": dy0, "x1": x1.g})
            rows = eo.stem_wgrad_rows(N, H, W)
            wsS = self._alloc((rows, 3136), F32)
            self.B(eo.ew(L.EW_STEM_WGRAD, p=[self.x_in, dy0, wsS], i=[N, H, W]))
            self.B(eo.ew(L.EW_COLSUM_FINALIZE, p=[wsS, self.gview("input_conv.weight")],
                         i=[rows, 3136, 0, 3136]))

        def bwd_pool():
            self.B(eo.ew(L.EW_MAXPOOL_BWD, p=[h0.g, idx, x1.g],
                         i=[N, h1, w1, 64, h0.g.sw, x1.g.sw, 1 if x1.take_grad() else 0]))

        self.tape += [bwd_stem, bwd_pool]
        self.named = {"x1": x1}

        # ---- encoder stages (torchvision layer1..4 wired at advanced_models.py:84-87,114-117)
        depths = (3, 4, 6, 3)
        h = h0
        skips = {1: x1}
        for li in range(1, 5):
            for bi in range(depths[li - 1]):
                stride = 2 if (li > 1 and bi == 0) else 1
                last = bi == depths[li - 1] - 1
                out = skip_act(li + 1) if (last and li < 4) else None
                if r50:
                    h = self.bottleneck(h, f"enc{li}.{bi}", stride, widths[li - 1] // 4, out)
                else:
                    h = self.basic_block(h, f"enc{li}.{bi}", stride, widths[li - 1], out)
            if li < 4:
                skips[li + 1] = h
            self.named[f"x{li + 1}"] = h
        # ---- decoder (:284-334)
        d = h
        for k in (4, 3, 2, 1):
            d = self.decoder_stage(k, d, skips[k])
            self.named[f"d{k}"] = d
        # ---- upconv0 + conv_final (:337-350)
        d0 = self.act(N, H, W, 16)
        self.F(co.convT_fprop(d.v, self.packed("upconv0.weight", "Tfprop"), d0.v, bias=P["upconv0.bias"]))
        self.wgrad_ws(16, d.v.c, 4, N, d.v.h, d.v.w)
        if P["conv_final.weight"].shape[0] != 1:
            raise NotImplementedError("the B200 path implements n_classes=1 (binary masks), as every "
                                      "experiment of the reference uses")
        NP = N * H * W
        self.F(eo.ew(L.EW_DOT1_FWD, p=[d0.v, P["conv_final.weight"], P["conv_final.bias"], self.logits, None],
                     i=[NP, 16, 16]))
        d1 = d
        self.stored.update({"d0": d0.v, "logits": self.logits})

        def bwd_up0():
            self.upconv_backward("upconv0", d1, d0.g)

        def bwd_final():
            self.B(eo.ew(L.EW_DOT1_BWD, p=[self.dlogits, d0.v, P["conv_final.weight"], d0.g, self.part],
                         i=[NP, 16, 16, 16]))
            rows = eo.partial_rows(L.EW_DOT1_BWD, NP, 16)
            self.B(eo.ew(L.EW_COLSUM_FINALIZE, p=[self.part, self.gview("conv_final.weight")], i=[rows, 16, 0, 17]))
            self.B(eo.ew(L.EW_COLSUM_FINALIZE, p=[self.part[16:], self.gview("conv_final.bias")], i=[rows, 1, 0, 17]))
            d0.gw = True
            self.stored_g.update({"d0": d0.g, "logits": self.dlogits})

        self.tape += [bwd_up0, bwd_final]
        if ng:
            self.ws = self._alloc((max(self._ws_req, 1024) // 4 + 64,), F32)
            if _fixed_bn_bwd():  # first backward op: zero the backward accumulators (whole arena: 2 MB, ~1 us)
                self._fxb_alloc(0)
                self.B(eo.ew(L.EW_FILL, p=[self._fxb], i=[_FX_CAP * 8, 0]))
            done = {}  # gradient group -> index one past the last backward op that writes into it
            for fn in reversed(self.tape):
                self._gview_log = []
                fn()
                for n in self._gview_log:
                    done[self.ggroup[n]] = len(self.bwd_ops)
            self._gview_log = None
            # buckets in group order; a group can go out once every earlier group is final too
            self.buckets, cut = [], 0
            for g in sorted(self.group_range):
                cut = max(cut, done.get(g, 0))
                lo, hi = self.group_range[g]
                if self.buckets and self.buckets[-1][0] == cut:
                    self.buckets[-1] = (cut, self.buckets[-1][1], hi)
                else:
                    self.buckets.append((cut, lo, hi))
            self.buckets[-1] = (len(self.bwd_ops),) + self.buckets[-1][1:]
        self.tape = None
        self._finish_pack()
        # the weight repack (fp32 parameters -> bf16 GEMM operands) overlaps the stem, which reads the
        # fp32 stem weights directly; it joins the chain before the first tensor-core convolution
        self.fwd = co.Seq(self.fwd_ops, side_kinds=(f"ew{L.EW_PACK}",), join_before_kinds=("conv",))
        # weight-gradient GEMMs read a forward activation and a gradient that is final when they are
        # appended, and write only their slice of gflat: they float on the graph's side branch
        self.bwd = co.Seq(self.bwd_ops, side_kinds=("wgrad",)) if ng else None
        self.launches_fwd = self.fwd.launches
        self.launches_bwd = self.bwd.launches if ng else 0

    def _finish_pack(self):
        jobs = (L.PackJob * len(self.pack_jobs))()
        start = 0
        for j, (p, out, T, A, B_, sa, sb, st) in enumerate(self.pack_jobs):
            jobs[j] = L.PackJob(p.data_ptr(), out.data_ptr(), T, A, B_, 0, sa, sb, st, start)
            start += ((A + 31) // 32) * ((B_ + 31) // 32)  # job.start = first 32x32xT tile
        raw = torch.frombuffer(bytearray(bytes(jobs)), dtype=torch.uint8).to(self.dev)
        self.keep.append(raw)
        self.pack_op = eo.ew(L.EW_PACK, p=[raw], i=[len(self.pack_jobs), start])
        self.fwd_ops.insert(0, self.pack_op)
        if getattr(self, "_fx", None) is not None:  # fixed-point BatchNorm accumulators start every pass at zero
            self.fwd_ops.insert(0, eo.ew(L.EW_FILL, p=[self._fx], i=[self._fx_used * 8, 0]))

    def upconv_backward(self, name, dprev: Act, dy: View):
        """ConvTranspose2d(k=2,s=2) backward: weight, bias and (accumulating) input gradient."""
        self.B(co.convT_wgrad(dprev.v, dy, self.gview(name + ".weight"), self.ws))
        self.colsum_to(dy, name + ".bias")
        if dprev.g is not None:
            self.B(co.convT_dgrad(dy, self.packed(name + ".weight", "Tdgrad"), dprev.g,
                                  accumulate=dprev.take_grad()))

    def _like(self, v: View):
        return View(self._alloc((v.n, v.h, v.w, v.c), BF16))

    def basic_block(self, hin: Act, pre, stride, cout, out: Act | None):
        """torchvision BasicBlock: conv3x3-bn-relu-conv3x3-bn (+1x1 downsample-bn) += identity, relu."""
        N = self.N
        y1, b1 = self.conv_bn(hin.v, pre + ".conv1.weight", pre + ".bn1", 3, stride, cout)
        a1 = self.act(N, y1.h, y1.w, cout)
        self.bn_apply(y1, b1, a1.v, True)
        y2, b2 = self.conv_bn(a1.v, pre + ".conv2.weight", pre + ".bn2", 3, 1, cout)
        out = out or self.act(N, y2.h, y2.w, cout)
        ds = (pre + ".downsample.0.weight") in self.params
        if ds:
            yd, bd = self.conv_bn(hin.v, pre + ".downsample.0.weight", pre + ".downsample.1", 1, stride, cout)
            self.bn_apply(y2, b2, out.v, True, res=yd, rbn=bd)
        else:
            self.bn_apply(y2, b2, out.v, True, res=hin.v)
        self.stored.update({pre + ".y1": y1, pre + ".a1": a1.v, pre + ".y2": y2, pre + ".out": out.v})
        if ds:
            self.stored[pre + ".yd"] = yd

        def bwd():
            dy2 = self._like(y2)
            if ds:
                dzb = self._like(y2)
                self.bn_backward(pre + ".bn2", out.g, out.v, y2, b2, dy2, dz=dzb)
            else:
                self.bn_backward(pre + ".bn2", out.g, out.v, y2, b2, dy2, dz=hin.g, dz_acc=hin.take_grad())
            self.conv_backward(a1, dy2, pre + ".conv2.weight", 3, 1)
            dy1 = self._like(y1)
            self.bn_backward(pre + ".bn1", a1.g, "self", y1, b1, dy1)
            self.conv_backward(hin, dy1, pre + ".conv1.weight", 3, stride)
            if ds:
                dyd = self._like(yd)
                self.bn_backward(pre + ".downsample.1", dzb, None, yd, bd, dyd)
                self.conv_backward(hin, dyd, pre + ".downsample.0.weight", 1, stride)
                self.stored_g[pre + ".yd"] = dyd
            self.stored_g.update({pre + ".y1": dy1, pre + ".a1": a1.g, pre + ".y2": dy2, pre + ".out": out.g})

        self.tape.append(bwd)
        return out

    def bottleneck(self, hin: Act, pre, stride, planes, out: Act | None):
        """torchvision Bottleneck v1.5: 1x1 - 3x3(stride) - 1x1(x4), each with BN, residual, relu."""
        N, cout = self.N, planes * 4
        y1, b1 = self.conv_bn(hin.v, pre + ".conv1.weight", pre + ".bn1", 1, 1, planes)
        a1 = self.act(N, y1.h, y1.w, planes)
        self.bn_apply(y1, b1, a1.v, True)
        y2, b2 = self.conv_bn(a1.v, pre + ".conv2.weight", pre + ".bn2", 3, stride, planes)
        a2 = self.act(N, y2.h, y2.w, planes)
        self.bn_apply(y2, b2, a2.v, True)
        y3, b3 = self.conv_bn(a2.v, pre + ".conv3.weight", pre + ".bn3", 1, 1, cout)
        out = out or self.act(N, y3.h, y3.w, cout)
        ds = (pre + ".downsample.0.weight") in self.params
        if ds:
            yd, bd = self.conv_bn(hin.v, pre + ".downsample.0.weight", pre + ".downsample.1", 1, stride, cout)
            self.bn_apply(y3, b3, out.v, True, res=yd, rbn=bd)
        else:
            self.bn_apply(y3, b3, out.v, True, res=hin.v)
        self.stored.update({pre + ".y1": y1, pre + ".a1": a1.v, pre + ".y2": y2, pre + ".a2": a2.v,
                            pre + ".y3": y3, pre + ".out": out.v})
        if ds:
            self.stored[pre + ".yd"] = yd

        def bwd():
            dy3 = self._like(y3)
            if ds:
                dzb = self._like(y3)
                self.bn_backward(pre + ".bn3", out.g, out.v, y3, b3, dy3, dz=dzb)
            else:
                self.bn_backward(pre + ".bn3", out.g, out.v, y3, b3, dy3, dz=hin.g, dz_acc=hin.take_grad())
            self.conv_backward(a2, dy3, pre + ".conv3.weight", 1, 1)
            dy2 = self._like(y2)
            self.bn_backward(pre + ".bn2", a2.g, "self", y2, b2, dy2)
            self.conv_backward(a1, dy2, pre + ".conv2.weight", 3, stride)
            dy1 = self._like(y1)
            self.bn_backward(pre + ".bn1", a1.g, "self", y1, b1, dy1)
            self.conv_backward(hin, dy1, pre + ".conv1.weight", 1, 1)
            if ds:
                dyd = self._like(yd)
                self.bn_backward(pre + ".downsample.1", dzb, None, yd, bd, dyd)
                self.conv_backward(hin, dyd, pre + ".downsample.0.weight", 1, stride)
                self.stored_g[pre + ".yd"] = dyd
            self.stored_g.update({pre + ".y1": dy1, pre + ".a1": a1.g, pre + ".y2": dy2, pre + ".a2": a2.g,
                                  pre + ".y3": dy3, pre + ".out": out.g})

        self.tape.append(bwd)
        return out

    def decoder_stage(self, k, dprev: Act, skip: Act):
        """upconv_k -> [attention gate on the skip] -> cat(skip, up) -> conv-bn-relu x2 ->
        [channel attention]  (advanced_models.py:284-334).  The concat is two channel slices."""
        N, ng = self.N, self.need_grad
        Cs, Cu = self.skip_c[k], self.up_out[k]
        h, w = self.H >> k, self.W >> k
        catA = Act(View(self.cat[k]), View(self.dcat[k]) if ng else None)
        up = Act(View(self.cat[k], Cs, Cu), View(self.dcat[k], Cs, Cu) if ng else None)
        sv = View(self.cat[k], 0, Cs)
        dsv = View(self.dcat[k], 0, Cs) if ng else None
        un = f"upconv{k}"
        self.F(co.convT_fprop(dprev.v, self.packed(un + ".weight", "Tfprop"), up.v, bias=self.params[un + ".bias"]))
        self.wgrad_ws(Cu, dprev.v.c, 4, N, dprev.v.h, dprev.v.w)
        self.tape.append(lambda: self.upconv_backward(un, dprev, up.g))
        if self.attn:
            self.gate(k, up, skip, sv, dsv)
        p = f"decoder{k}"
        ya, ba = self.conv_bn(catA.v, p + ".0.weight", p + ".1", 3, 1, Cu, biasname=p + ".0.bias")
        a = self.act(N, h, w, Cu)
        self.bn_apply(ya, ba, a.v, True)
        yb, bb = self.conv_bn(a.v, p + ".3.weight", p + ".4", 3, 1, Cu, biasname=p + ".3.bias")
        b = self.act(N, h, w, Cu)
        self.bn_apply(yb, bb, b.v, True)
        self.stored.update({f"up{k}": up.v, f"dec{k}.ya": ya, f"dec{k}.a": a.v, f"dec{k}.yb": yb, f"dec{k}.b": b.v})

        def bwd_dec():
            dyb = self._like(yb)
            self.bn_backward(p + ".4", b.g, "self", yb, bb, dyb, dbias=self.gview(p + ".3.bias"))
            self.conv_backward(a, dyb, p + ".3.weight", 3, 1)
            dya = self._like(ya)
            self.bn_backward(p + ".1", a.g, "self", ya, ba, dya, dbias=self.gview(p + ".0.bias"))
            self.conv_backward(catA, dya, p + ".0.weight", 3, 1)  # first writer of the whole dcat
            up.gw = True
            if not self.attn:
                skip.gw = True  # skip.g IS the skip slice of dcat
            self.stored_g.update({f"up{k}": up.g, f"dec{k}.ya": dya, f"dec{k}.a": a.g, f"dec{k}.yb": dyb,
                                  f"dec{k}.b": b.g})

        self.tape.append(bwd_dec)
        if not self.attn:
            return b
        out = self.act(N, h, w, Cu)
        self.channel_attention(k, b, out)
        self.stored[f"se{k}.out"] = out.v
        return out

    def gate(self, k, up: Act, skip: Act, outS: View, doutS: View | None):
        """AttentionGate.forward (advanced_models.py:28-40): x * sigmoid(BN(psi(relu(BN(Wg g) + BN(Wx x)))))."""
        N, P = self.N, self.params
        a = f"attention{k}"
        fint = P[a + ".W_g.0.weight"].shape[0]
        hh, ww = up.v.h, up.v.w
        NP = N * hh * ww
        yg, bg = self.conv_bn(up.v, a + ".W_g.0.weight", a + ".W_g.1", 1, 1, fint, biasname=a + ".W_g.0.bias")
        ys, bs = self.conv_bn(skip.v, a + ".W_x.0.weight", a + ".W_x.1", 1, 1, fint, biasname=a + ".W_x.0.bias")
        pin = self.act(N, hh, ww, fint)
        self.bn_apply(yg, bg, pin.v, True, res=ys, rbn=bs)
        pre = self._alloc((NP,), F32)
        psi = self._alloc((NP,), F32)
        rows = eo.partial_rows(L.EW_DOT1_FWD, NP, fint)
        self.F(eo.ew(L.EW_DOT1_FWD, p=[pin.v, P[a + ".psi.0.weight"], P[a + ".psi.0.bias"], pre,
                                       self.part if self.training else None], i=[NP, fint, pin.v.sw]))
        bp = self.bn_finalize(a + ".psi.1", 1, rows, NP, self.part).bnp
        self.F(eo.ew(L.EW_GATE_APPLY, p=[pre, bp, skip.v, outS, psi], i=[NP, skip.v.c, skip.v.sw, outS.sw]))
        self.stored.update({f"att{k}.yg": yg, f"att{k}.ys": ys, f"att{k}.pin": pin.v,
                            f"att{k}.pre": pre.view(N, 1, hh, ww), f"att{k}.out": outS})

        def bwd():
            dz = self._alloc((NP,), F32)
            rows2 = eo.partial_rows(L.EW_GATE_BWD, NP, skip.v.c)
            self.B(eo.ew(L.EW_GATE_BWD, p=[doutS, skip.v, psi, pre, bp, dz, skip.g, self.part],
                         i=[NP, skip.v.c, doutS.sw, skip.v.sw, skip.g.sw, 1 if skip.take_grad() else 0]))
            coef = self._alloc((3, 1), F32)
            self.B(eo.ew(L.EW_BN_BWD_FINALIZE,
                         p=[self.part, P[a + ".psi.1.weight"], bp, self.gview(a + ".psi.1.weight"),
                            self.gview(a + ".psi.1.bias"), coef, self.gview(a + ".psi.0.bias")],
                         i=[rows2, 1, NP, 0, 1 if self.training else 0]))
            self.B(eo.ew(L.EW_DOT1_BWD, p=[dz, pin.v, P[a + ".psi.0.weight"], pin.g, self.part, coef, pre, bp],
                         i=[NP, fint, pin.v.sw, pin.g.sw]))
            rows3 = eo.partial_rows(L.EW_DOT1_BWD, NP, fint)
            self.B(eo.ew(L.EW_COLSUM_FINALIZE, p=[self.part, self.gview(a + ".psi.0.weight")],
                         i=[rows3, fint, 0, fint + 1]))
            pin.gw = True
            dyg, dys = self._like(yg), self._like(ys)
            self.bn_backward(a + ".W_g.1", pin.g, pin.v, yg, bg, dyg, dbias=self.gview(a + ".W_g.0.bias"))
            self.bn_backward(a + ".W_x.1", pin.g, pin.v, ys, bs, dys, dbias=self.gview(a + ".W_x.0.bias"))
            self.conv_backward(up, dyg, a + ".W_g.0.weight", 1, 1)
            self.conv_backward(skip, dys, a + ".W_x.0.weight", 1, 1)
            self.stored_g.update({f"att{k}.yg": dyg, f"att{k}.ys": dys, f"att{k}.pin": pin.g, f"att{k}.out": doutS})

        self.tape.append(bwd)

    def channel_attention(self, k, b: Act, out: Act):
        """ChannelAttention.forward (advanced_models.py:57-61)."""
        N, P = self.N, self.params
        C_, hw = b.v.c, b.v.h * b.v.w
        ch = C_ // 16
        n1, n2 = f"ch_attention{k}.fc.0.weight", f"ch_attention{k}.fc.2.weight"
        rows = eo.partial_rows(L.EW_SE_POOL, hw, C_)
        part = self._alloc((N, rows, 3, C_), F32)
        saved = self._alloc((N, 4, C_), F32)
        hidden = self._alloc((N, 2, ch), F32)
        self.F(eo.ew(L.EW_SE_POOL, p=[b.v, part], i=[N, hw, C_, b.v.sw]))
        self.F(eo.ew(L.EW_SE_MLP, p=[part, P[n1], P[n2], saved, hidden], i=[N, hw, C_, ch, rows]))
        self.F(eo.ew(L.EW_SE_SCALE, p=[b.v, saved, out.v], i=[N, hw, C_, b.v.sw, out.v.sw]))

        def bwd():
            dpool = self._alloc((N, 2, C_), F32)
            self.B(eo.ew(L.EW_SE_BWD_REDUCE, p=[out.g, b.v, part], i=[N, hw, C_, out.g.sw, b.v.sw]))
            self.B(eo.ew(L.EW_SE_MLP_BWD, p=[part, P[n1], P[n2], saved, hidden, self.gview(n1),
                                             self.gview(n2), dpool], i=[N, hw, C_, ch, rows, 0]))
            self.B(eo.ew(L.EW_SE_BWD_APPLY, p=[out.g, saved, dpool, b.g], i=[N, hw, C_, out.g.sw, b.g.sw]))
            b.gw = True
            self.stored_g[f"se{k}.out"] = out.g

        self.tape.append(bwd)

    def stored_nchw(self, grads=False):
        """Host copies (NCHW fp32) of every stored activation (or, with grads=True, of the
        gradient w.r.t. it after a backward pass), keyed like oracle.unet_functional's storage
        points."""
        out = {}
        for name, v in (self.stored_g if grads else self.stored).items():
            if torch.is_tensor(v):
                out[name] = v.detach().float().cpu().clone()
            else:
                out[name] = v.torch().float().cpu().permute(0, 3, 1, 2).contiguous()
        return out

    # ---------------------------------------------------------------- execution ----
    def params_moved(self) -> bool:
        return any(p.data_ptr() != q for p, q in zip(self.params.values(), self.param_ptrs))

    def run_forward(self, x: torch.Tensor) -> torch.Tensor:
        self.x_in.copy_(x.reshape(self.N, self.H, self.W))
        self.fwd.run()
        self.generation += 1
        return self.logits

    def run_backward(self, dlogits: torch.Tensor, on_bucket=None):
        """Replay the backward pass.  With `on_bucket` (data-parallel runs) it is replayed as one CUDA
        graph per gradient bucket, and on_bucket(slice of gflat) is called as soon as the segment
        that completes that slice has been enqueued -- the caller starts its all-reduce there, which
        then overlaps the remaining segments."""
        self.dlogits.copy_(dlogits.reshape(self.dlogits.shape))
        if on_bucket is None:
            self.bwd.run()
            return
        if self.bwd_segs is None:
            self.bwd_segs, start = [], 0
            for cut, lo, hi in self.buckets:
                self.bwd_segs.append((co.Seq(self.bwd_ops[start:cut], side_kinds=("wgrad",)), lo, hi))
                start = cut
        for seg, lo, hi in self.bwd_segs:
            seg.run()
            on_bucket(self.gflat[lo:hi])
