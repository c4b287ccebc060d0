"""Native LeNet runner: the whole training step as a short fixed sequence of
hand-written sm_100a kernels over flat arenas, replayed from a CUDA graph.

Default step for the reference LeNet = 11 launches, 7 of them on the main stream's dependency chain; reference: ~50
ATen/cuDNN/cuBLAS kernels + 10 NCCL all-reduces + 10 divides + the optimizer's elementwise kernels, train.py:45-51:

    main stream                                                      other streams (overlapped)
    ---------------------------------------------------------------  ------------------------------------------------
    conv1 fwd   (normalise+upsample+conv+act+pool)   CUDA cores      tcw_pack_weights (bf16 panels)
    conv2 fwd   (conv+act+pool)                      tcgen05 bf16
    head_fused  (conv3+FC.0+FC.2+softmax+loss+dlogits+dz1+dz3) fp32, weights by TMA
    tc_gemm     {dW FC.2, dW FC.0, dW conv3, dP2}    tcgen05 tf32    post_loss (loss -> host's mapped array)
    conv2 dgrad (pool/act bwd fused)                 tcgen05 bf16    reduce_update_ll, dense buckets: push to peers + wait +
    conv1 wgrad                                      tcgen05 bf16      rank-order sum + average + optimizer;  conv2 wgrad
    reduce_update_ll, conv bucket (the step's last launch: partial rows + push + wait + sum + average + optimizer)

Which kernel runs a layer is a measured choice (``dtype="auto"``, ``conv="auto"``; profiles/lenet_tc_vs_simt.md): conv1
(K = 25) is faster on CUDA cores, everything from conv2 on is faster on the tensor cores; ``dtype="fp32", conv="simt"``
keeps every layer on the exact fp32 CUDA-core kernels; the 4x-wide model runs 14 launches, all GEMM-shaped work on
tcgen05, with one fused handshake reduce+update per bucket.  Every launch of the main chain uses programmatic dependent
launch, so the chain's scheduling gaps overlap (csrc/common.cuh).

The gradient exchange is the fused P2P kernel pair of csrc/ll_reduce.cu (push protocol) or csrc/reduce_update.cu
(handshake per bucket); with world size 1 both degenerate to the optimizer update.  ``state_dict`` keys/shapes are the
reference's because the parameters are views of an eager :class:`LeNet` re-pointed into the arena.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from ..data.synthetic import MNIST_SPEC, ImageSpec
from ..ops import native
from ..optim.flat import FlatOptimizer, OptHyper
from ..parallel.arena import FlatArena, lenet_groups
from ..parallel.p2p import P2PComm
from .lenet import LeNet, LeNetConfig

_LOSS_MODES = {"ref": 0, "xent": 1}


class NativeLeNet:
    """Buffers + kernel launch sequences for one model replica on one GPU."""

    def __init__(self, cfg: LeNetConfig, batch: int, device: torch.device, *, loss_mode: str = "ref",
                 spec: ImageSpec = MNIST_SPEC, module: Optional[LeNet] = None, tc: bool = False,
                 conv_tc=None, head_fused: Optional[bool] = None):
        """``tc=True`` routes the five GEMM-shaped layers (conv3, FC.0 and their backward, FC.2's weight
        gradient) through the tcgen05/TMEM/TMA tf32 kernel; ``tc=False`` keeps them in exact fp32 on CUDA cores.
        ``conv_tc`` does the same for the two 5x5 convolutions (window-tile bf16 tcgen05 kernels): ``True`` / ``False``
        for all of them, or a triple ``(conv1 forward, conv2 forward, convolution backward)`` — the measured-fastest
        mix for the reference LeNet is ``(False, True, True)`` (profiles/lenet_tc_vs_simt.md).  In tc mode the GEMM
        operand buffers carry one extra column of ones (the bias gradients fall out of the weight-gradient GEMMs) and
        are padded to a 16-byte row pitch for TMA."""
        native.load()
        self.cfg, self.B, self.device, self.spec = cfg, int(batch), torch.device(device), spec
        self.tc = bool(tc)
        # convolutions: window-tile tcgen05 kernels (csrc/tc_win.cu, tc_wgrad.cu) or CUDA cores (conv_fwd.cu / conv_bwd.cu).
        # The fused CUDA-core backward only exists for the reference geometry; the wide model needs tc.
        if conv_tc is None:
            conv_tc = cfg.width_mult > 1
        if isinstance(conv_tc, (tuple, list)):
            self.conv1_tc, self.conv2_tc, self.convbwd_tc = (bool(v) for v in conv_tc)
        else:
            self.conv1_tc = self.conv2_tc = self.convbwd_tc = bool(conv_tc)
        self.conv_tc = self.conv1_tc or self.conv2_tc or self.convbwd_tc      # any window-tile kernel in the step
        if cfg.width_mult > 1 and not self.convbwd_tc:
            raise ValueError("the wide model's convolution backward runs on the tensor-core kernels only")
        self.loss_mode = _LOSS_MODES[loss_mode]
        # conv3 + FC.0 + FC.2 + loss + the row-local backward in ONE fp32 launch (csrc/head_fused.cu) where the layer
        # widths allow it (the reference geometry); the dense layers of the wide model stay separate GEMM launches
        fits = native.head_fused_supported(cfg.c2 * 25, cfg.c3, cfg.f1, cfg.num_classes)
        self.head_fused = fits if head_fused is None else (bool(head_fused) and fits)
        self.module = module if module is not None else LeNet(cfg)
        self.module.to(self.device)
        names = [n for n, _ in self.module.named_parameters()]
        self.arena = FlatArena(self.module, lenet_groups(names), self.device)
        L = self.arena.layout
        self.w = {n: L.view(self.arena.param, n) for n in names}
        self.g = {n: L.view(self.arena.grad, n) for n in names}
        self.bucket_fc, self.bucket_c3, self.bucket_conv = L.buckets[0], L.buckets[1], L.buckets[2]
        B, c = self.B, cfg
        f32 = dict(dtype=torch.float32, device=self.device)
        u8 = dict(dtype=torch.uint8, device=self.device)
        self.p1 = torch.empty(B, c.c1, 14, 14, **f32)
        self.idx1 = torch.empty(B, c.c1, 14, 14, **u8)
        k3 = c.c2 * 25
        if self.tc:
            self.ld_p2, self.ld_a3, self.ld_h1 = k3 + 4, c.c3 + 4, c.f1 + 4
            self.ld_dl = (c.num_classes + 3) // 4 * 4
            self.p2 = torch.zeros(B, self.ld_p2, **f32)
            self.a3 = torch.zeros(B, self.ld_a3, **f32)
            self.h1 = torch.zeros(B, self.ld_h1, **f32)
            self.p2[:, k3] = 1.0            # ones columns: dW GEMMs then also produce db
            self.a3[:, c.c3] = 1.0
            self.h1[:, c.f1] = 1.0
            self.dlogits = torch.zeros(B, self.ld_dl, **f32)
        else:
            self.ld_p2, self.ld_a3, self.ld_h1, self.ld_dl = 0, 0, 0, 0
            self.p2 = torch.empty(B, c.c2, 5, 5, **f32)
            self.a3 = torch.empty(B, c.c3, **f32)
            self.h1 = torch.empty(B, c.f1, **f32)
            self.dlogits = torch.empty(B, c.num_classes, **f32)
        self.idx2 = torch.empty(B, c.c2, 5, 5, **u8)
        self.probs = torch.empty(B, c.num_classes, **f32)
        self.loss_rows = torch.empty(B, **f32)
        self.loss = torch.zeros(1, **f32)
        self.dz1 = torch.empty(B, c.f1, **f32)
        self.dz3 = torch.empty(B, c.c3, **f32)
        self.dp2 = torch.empty(B, c.c2 * 25, **f32)
        self.ticket = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.eval_acc = torch.zeros(2, dtype=torch.float64, device=self.device)
        self.g1 = torch.empty(B, c.c1, 14, 14, **f32) if self.convbwd_tc else None
        self._num_partials = native.tcw_conv_wgrad_num_partials if self.convbwd_tc else native.conv_bwd_num_partials
        # per-tap weight panels of the tensor-core convolutions, re-packed at the start of every forward
        self.wpack = torch.empty(native.tcw_pack_floats(self.dims(B)), **f32) if self.conv_tc else None
        self.n_partials = self._num_partials(self.dims(B))   # rows allocated (full batch)
        self.cur_partials = self.n_partials    # rows the last backward actually wrote (fewer on a ragged batch)
        self.partial_stride = self.bucket_conv.numel
        # scratch: one partial-gradient row per conv-backward CTA (zeroed once: padding columns stay 0)
        self.partials = torch.zeros(self.n_partials, self.partial_stride, **f32)
        s0 = self.bucket_conv.start
        sl = L.slots
        self.conv_offsets = tuple(sl[k].offset - s0 for k in
                                  ("ConvNet.1.weight", "ConvNet.1.bias", "ConvNet.4.weight", "ConvNet.4.bias"))
        self.launches_per_step = 0
        self.gemm_ws = native.GemmWorkspace(self.device) if self.tc else None
        # second stream for work that does not sit on the step's dependency chain: the weight pack beside the first
        # layer, conv2's weight gradient beside conv2's data gradient
        self.pack_stream = torch.cuda.Stream(device=self.device) if self.conv_tc else None
        self.parallel_wgrad = True
        self.after_first_kernel = None  # optional callable run right after the first launch of a forward (see StepPump)

    def dims(self, b: int) -> native.Dims:
        c = self.cfg
        return native.Dims(b, c.in_ch, c.in_hw, c.c1, c.c2, c.c3, c.f1, c.num_classes, self.ld_p2)

    # ---- forward ----------------------------------------------------------------
    def forward(self, x: torch.Tensor, y: torch.Tensor, train: bool) -> None:
        b = x.shape[0]
        assert b <= self.B and x.is_contiguous() and y.dtype == torch.int64
        d, w = self.dims(b), self.w
        need_pack = self.conv1_tc or self.conv2_tc or (self.convbwd_tc and train)
        fork = None
        if need_pack and not self.conv1_tc:
            # the first layer runs on CUDA cores and needs no packed panels: pack beside it
            main = torch.cuda.current_stream(self.device)
            self.pack_stream.wait_stream(main)
            with torch.cuda.stream(self.pack_stream):
                native.tcw_pack_weights(w["ConvNet.1.weight"], w["ConvNet.4.weight"], self.wpack, d)
            fork = main
        elif need_pack:
            native.tcw_pack_weights(w["ConvNet.1.weight"], w["ConvNet.4.weight"], self.wpack, d)
        if self.conv1_tc:
            native.tcw_conv1_fwd(x, self.wpack, w["ConvNet.1.bias"], self.p1, self.idx1 if train else None,
                                 d, self.spec.mean, self.spec.std)
        else:
            native.conv1_fwd(x, w["ConvNet.1.weight"], w["ConvNet.1.bias"], self.p1, self.idx1 if train else None,
                             d, self.spec.mean, self.spec.std)
        if self.after_first_kernel is not None:
            self.after_first_kernel()
        if fork is not None:
            fork.wait_stream(self.pack_stream)
        if self.conv2_tc:
            native.tcw_conv2_fwd(self.p1, self.wpack, w["ConvNet.4.bias"], self.p2, self.idx2 if train else None, d)
        else:
            native.conv2_fwd(self.p1, w["ConvNet.4.weight"], w["ConvNet.4.bias"], self.p2, self.idx2 if train else None, d)
        c = self.cfg
        k3 = c.c2 * 25
        if self.head_fused:
            native.head_fused(self.p2, self.ld_p2, w["ConvNet.7.weight"], w["ConvNet.7.bias"], w["FC.0.weight"],
                              w["FC.0.bias"], w["FC.2.weight"], w["FC.2.bias"], y, self.loss_mode, self.a3, self.ld_a3,
                              self.h1, self.ld_h1, self.probs, self.loss_rows, self.loss,
                              self.dlogits if train else None, self.ld_dl, self.dz1 if train else None,
                              self.dz3 if train else None, None if train else self.eval_acc, self.ticket, b)
            return
        if self.tc:
            T = native.tc_problem
            native.tc_gemm_group([T(self.p2, self.ld_p2, False, w["ConvNet.7.weight"], k3, False, b, c.c3, k3,
                                    bias=w["ConvNet.7.bias"], act=1, out=self.a3, ld_out=self.ld_a3)], self.gemm_ws)
            native.tc_gemm_group([T(self.a3, self.ld_a3, False, w["FC.0.weight"], c.c3, False, b, c.f1, c.c3,
                                    bias=w["FC.0.bias"], act=1, out=self.h1, ld_out=self.ld_h1)], self.gemm_ws)
        else:
            native.linear_fwd(self.p2[:b].view(b, -1), w["ConvNet.7.weight"].view(c.c3, -1),
                              w["ConvNet.7.bias"], self.a3[:b], 1)
            native.linear_fwd(self.a3[:b], w["FC.0.weight"], w["FC.0.bias"], self.h1[:b], 1)
        native.head_tail(self.h1[:b], w["FC.2.weight"], w["FC.2.bias"], y, self.loss_mode, self.probs[:b],
                         self.loss_rows[:b], self.loss, self.dlogits[:b] if train else None,
                         self.dz1[:b] if train else None, None if train else self.eval_acc, self.ticket,
                         self.ld_h1, self.ld_dl)

    # ---- backward, in three launches; a callback fires as each bucket's gradients complete ----
    def backward(self, x: torch.Tensor, on_bucket=None) -> None:
        b, c, w, g = x.shape[0], self.cfg, self.w, self.g
        d = self.dims(b)
        k3 = c.c2 * 25
        if self.head_fused:
            # dz3 came out of the fused head: the four remaining products depend on its outputs only -> one launch
            if self.tc:
                T, nc = native.tc_problem, c.num_classes
                native.tc_gemm_group([
                    T(self.dlogits, self.ld_dl, True, self.h1, self.ld_h1, True, nc, c.f1 + 1, b,
                      out=g["FC.2.weight"], ld_out=c.f1, colext=g["FC.2.bias"]),
                    T(self.dz1, c.f1, True, self.a3, self.ld_a3, True, c.f1, c.c3 + 1, b,
                      out=g["FC.0.weight"], ld_out=c.c3, colext=g["FC.0.bias"]),
                    T(self.dz3, c.c3, True, self.p2, self.ld_p2, True, c.c3, k3 + 1, b,
                      out=g["ConvNet.7.weight"], ld_out=k3, colext=g["ConvNet.7.bias"]),
                    T(self.dz3, c.c3, False, w["ConvNet.7.weight"], k3, True, b, k3, c.c3, out=self.dp2, ld_out=k3),
                ], self.gemm_ws)
            else:
                native.linear_bwd_group([
                    native.bwd_problem(1, self.dlogits, self.h1, None, g["FC.2.weight"], g["FC.2.bias"], b, c.num_classes, c.f1),
                    native.bwd_problem(1, self.dz1, self.a3, None, g["FC.0.weight"], g["FC.0.bias"], b, c.f1, c.c3),
                    native.bwd_problem(1, self.dz3, self.p2, None, g["ConvNet.7.weight"], g["ConvNet.7.bias"], b, c.c3, k3),
                    native.bwd_problem(0, self.dz3, w["ConvNet.7.weight"], None, self.dp2, None, b, c.c3, k3),
                ])
            if on_bucket:
                on_bucket(0)
                on_bucket(1)
        elif self.tc:
            T, nc = native.tc_problem, c.num_classes
            native.tc_gemm_group([
                # dW fc2 = dlogits^T h1 (+db through h1's ones column)
                T(self.dlogits, self.ld_dl, True, self.h1, self.ld_h1, True, nc, c.f1 + 1, b,
                  out=g["FC.2.weight"], ld_out=c.f1, colext=g["FC.2.bias"]),
                # dW fc1 = dz1^T a3 (+db)
                T(self.dz1, c.f1, True, self.a3, self.ld_a3, True, c.f1, c.c3 + 1, b,
                  out=g["FC.0.weight"], ld_out=c.c3, colext=g["FC.0.bias"]),
                # dz3 = (dz1 Wfc1) * leaky'(a3)
                T(self.dz1, c.f1, False, w["FC.0.weight"], c.c3, True, b, c.c3, c.f1,
                  mask=self.a3, ld_mask=self.ld_a3, out=self.dz3, ld_out=c.c3),
            ], self.gemm_ws)
            if on_bucket:
                on_bucket(0)
            native.tc_gemm_group([
                # dW conv3 = dz3^T p2 (+db)
                T(self.dz3, c.c3, True, self.p2, self.ld_p2, True, c.c3, k3 + 1, b,
                  out=g["ConvNet.7.weight"], ld_out=k3, colext=g["ConvNet.7.bias"]),
                # dp2 = dz3 W3 (pool/activation masks are applied inside conv_bwd)
                T(self.dz3, c.c3, False, w["ConvNet.7.weight"], k3, True, b, k3, c.c3, out=self.dp2, ld_out=k3),
            ], self.gemm_ws)
            if on_bucket:
                on_bucket(1)
        else:
            native.linear_bwd_group([
                native.bwd_problem(1, self.dlogits, self.h1, None, g["FC.2.weight"], g["FC.2.bias"], b, c.num_classes, c.f1),
                native.bwd_problem(1, self.dz1, self.a3, None, g["FC.0.weight"], g["FC.0.bias"], b, c.f1, c.c3),
                native.bwd_problem(0, self.dz1, w["FC.0.weight"], self.a3, self.dz3, None, b, c.f1, c.c3),
            ])
            if on_bucket:
                on_bucket(0)
            native.linear_bwd_group([
                native.bwd_problem(1, self.dz3, self.p2, None, g["ConvNet.7.weight"], g["ConvNet.7.bias"], b, c.c3, k3),
                native.bwd_problem(0, self.dz3, w["ConvNet.7.weight"], None, self.dp2, None, b, c.c3, k3),
            ])
            if on_bucket:
                on_bucket(1)
        ow1, ob1, ow2, ob2 = self.conv_offsets
        self.cur_partials = self._num_partials(d)
        if self.convbwd_tc:
            ld_p2 = self.ld_p2 or k3
            # conv2's weight gradient needs dp2 only; conv1's needs conv2's data gradient g1: the former runs beside
            # the dgrad on the second stream (disjoint columns of the partial rows)
            main = torch.cuda.current_stream(self.device)
            par = self.parallel_wgrad and self.pack_stream is not None
            if par:
                self.pack_stream.wait_stream(main)
                with torch.cuda.stream(self.pack_stream):
                    native.tcw_conv_wgrad(2, self.p1, 0.0, 1.0, self.dp2, self.p2, ld_p2, self.idx2, self.partials,
                                          self.partial_stride, ow2, ob2, d)
            native.tcw_conv2_dgrad(self.dp2, self.p2, self.idx2, self.wpack, self.p1, self.g1, d)
            if not par:
                native.tcw_conv_wgrad(2, self.p1, 0.0, 1.0, self.dp2, self.p2, ld_p2, self.idx2, self.partials,
                                      self.partial_stride, ow2, ob2, d)
            native.tcw_conv_wgrad(1, x, self.spec.mean, self.spec.std, self.g1, None, 0, self.idx1, self.partials,
                                  self.partial_stride, ow1, ob1, d)
            if par:
                main.wait_stream(self.pack_stream)
        else:
            native.conv_bwd(x, self.spec.mean, self.spec.std, self.p1, self.idx1, self.p2, self.idx2, self.dp2,
                            w["ConvNet.4.weight"], self.partials, self.partial_stride, ow1, ob1, ow2, ob2, d)
        if on_bucket:
            on_bucket(2)


class NativeStepper:
    """Engine-facing stepper: ``train_step`` / ``eval_batch`` / ``eval_result`` on native kernels,
    gradient reduction + optimizer fused in the P2P kernel, step replayed from CUDA graphs."""

    name = "native"
    MAX_GRAPHS = 8

    def __init__(self, ctx, cfg: LeNetConfig, hyper: OptHyper, *, batch: int, loss_mode: str = "ref",
                 spec: ImageSpec = MNIST_SPEC, algo: str = "auto", graph: bool = True, overlap: bool = True,
                 dtype: str = "auto", module: Optional[LeNet] = None, conv: str = "auto", wire: str = "auto",
                 timeout_ms: int = 0):
        """``dtype``: ``fp32`` = every layer on the exact fp32 CUDA-core kernels; ``tf32`` = dense layers on the tcgen05
        tf32 GEMM; ``bf16`` = additionally the bf16 gradient wire format; ``auto`` = the measured-fastest mix for the
        model (profiles/lenet_tc_vs_simt.md).  ``conv``: ``simt`` / ``tc`` / ``hybrid`` (first layer on CUDA cores,
        second layer forward and the whole convolution backward on the bf16 window-tile tcgen05 kernels) / ``auto``.
        ``algo``: see :class:`b200dp.parallel.p2p.P2PComm`.  ``wire``: ``fp32`` / ``bf16`` gradient lines of the push
        protocol (``auto``: bf16 only with ``dtype="bf16"``).  ``timeout_ms``: bound on every wait for a peer."""
        self.ctx, self.cfg = ctx, cfg
        self.device = ctx.device
        if dtype not in ("auto", "fp32", "tf32", "bf16"):
            raise ValueError(f"unknown dtype {dtype!r}")
        if conv not in ("auto", "simt", "tc", "hybrid"):
            raise ValueError(f"unknown conv mode {conv!r}")
        self.requested_dtype = dtype
        wide = cfg.width_mult > 1
        if conv == "auto":
            # measured on B200 with the current kernels (profiles/lenet_tc_vs_simt.md): the first layer (K = 25) is
            # faster on CUDA cores, everything from the second layer on is faster on the tensor cores
            conv = "tc" if wide else ("simt" if dtype == "fp32" else "hybrid")
        if dtype == "auto":
            # dense layers: with the fused fp32 head (conv3 + FC.0 + FC.2 forward, one launch) the only GEMM launch left
            # is the backward group {dW FC.2, dW FC.0, dW conv3, dP2}, which is faster on tcgen05 (tf32) at both sizes
            dtype = "tf32"
        self.dtype = "fp32" if dtype == "fp32" else "tf32"       # GEMM operand type
        self.conv_mode = conv
        conv_tc = {"simt": False, "tc": True, "hybrid": (False, True, True)}[conv]
        if wire == "auto":
            wire = "bf16" if self.requested_dtype == "bf16" else "fp32"
        self.wire = wire
        self.net = NativeLeNet(cfg, batch, self.device, loss_mode=loss_mode, spec=spec, module=module,
                               tc=self.dtype != "fp32", conv_tc=conv_tc)
        self.model = self.net.module
        self.arena = self.net.arena
        self.optimizer = FlatOptimizer(self.arena.param, self.arena.grad, hyper)
        self.world = ctx.world_size
        self.comm = (P2PComm(self.arena.layout.total, ctx.rank, self.world, algo, wire=wire, timeout_ms=timeout_ms)
                     if self.world > 1 else None)
        # One launch ends the step (wait for the peers' pushed gradients, sum, average, update the whole arena) when the
        # model is in the latency regime; big arenas keep one handshake launch per bucket (csrc/reduce_update.cu).
        from ..parallel.p2p import ll_max_bytes
        small = self.arena.layout.total * 4 <= ll_max_bytes(self.world)
        self.fused_final = (self.comm.push if self.comm is not None else (small and algo in ("auto", "push")))
        self.use_graph, self.overlap = graph, overlap
        self.side = torch.cuda.Stream(device=self.device) if overlap else None
        self.hyper_dev = torch.zeros(8, dtype=torch.float32, device=self.device)
        self.hyper_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self._hyper_cache: Optional[Tuple] = None
        self.step_counter = torch.zeros(4, dtype=torch.int32, device=self.device)
        # {mapped pinned loss array, generation at its index 0, its length}: see b200dp_reduce_update
        # loss_sink[3]: address of a mapped host word; the step's last kernel copies the sticky error word there once it is
        # set, so that error_flag() is a plain memory read
        self._err_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._err_np = self._err_host.numpy()
        self.loss_sink = torch.tensor([0, 0, 0, self._err_host.data_ptr()], dtype=torch.int64).to(self.device)
        self._graphs: Dict[Tuple[int, int, int], torch.cuda.CUDAGraph] = {}
        self._static_x: Optional[torch.Tensor] = None
        self._static_y: Optional[torch.Tensor] = None
        self._warm = 0
        self._comm_off = False        # measurement aid: see set_comm_enabled
        self._loss_posted = False
        self._eval_graphs, self._eval_warm = 0, False
        self._pumps = []
        hf = self.net.head_fused
        n = (3 if hf else 5) + (1 if self.net.conv_tc else 0) + (1 if hf else 2) + (3 if self.net.convbwd_tc else 1)
        pushes = 1 if hf else (0 if self.world == 1 else 2)      # fused head: one dense push+wait+update launch
        n += (1 + pushes) if self.fused_final else 3                                       # reduce/update (+ pushes)
        n += 1 if overlap else 0                                                            # post_loss beside the backward
        self.launches_per_step = n            # forward + GEMM backward + conv backward + gradient exchange/update

    # ---- optimizer scalars live on the device so graph replays see scheduler changes ----
    def _sync_hyper(self) -> None:
        vec = self.optimizer.hyper_vector()
        if vec != self._hyper_cache:
            self.hyper_host[:5] = torch.tensor(vec, dtype=torch.float32)
            self.hyper_dev.copy_(self.hyper_host, non_blocking=True)
            self._hyper_cache = vec

    def _comm_struct(self, nbytes: int, overlapped: bool):
        if self.comm is None or self._comm_off:
            return None
        return self.comm.struct_for(nbytes, overlapped=overlapped)

    def _reduce_update(self, bucket_index: int) -> None:
        """Handshake protocol: one fused reduce + average + update launch per bucket."""
        net, opt = self.net, self.optimizer
        b = self.arena.layout.buckets[bucket_index]
        is_conv = bucket_index == 2
        native.reduce_update(self.arena.param, self.arena.grad, opt.state0, opt.state1, None, b.start, b.numel,
                             net.partials if is_conv else None, net.cur_partials if is_conv else 0,
                             net.partial_stride if is_conv else 0, opt.hyper.code(), self.hyper_dev,
                             self.step_counter, bucket_index == 2,
                             self._comm_struct(b.numel * 4, self.overlap and bucket_index != 2),
                             None if self._loss_posted else net.loss, self.loss_sink)

    def _push(self, bucket_index: int) -> None:
        """Push protocol: send a finished bucket to every peer, wait for nobody."""
        cs = self._comm_struct(0, False)
        if cs is not None:
            b = self.arena.layout.buckets[bucket_index]
            native.grad_push(self.arena.grad, b.start, b.numel, self.step_counter, cs)

    def _final_update(self, pushed: bool, conv_only: bool = False) -> None:
        """Push protocol: the step's last launch — sums the convolution backward's partial rows, pushes that bucket,
        waits for the peers' lines, sums in rank order, averages, updates.  Over the whole arena, or (``conv_only``) over
        the convolution bucket alone when the dense buckets had their own launch beside the convolution backward."""
        net, opt = self.net, self.optimizer
        conv = self.arena.layout.buckets[2]
        start, count = (conv.start, conv.numel) if conv_only else (0, self.arena.layout.total)
        native.reduce_update_ll(self.arena.param, self.arena.grad, opt.state0, opt.state1, None, start, count,
                                net.partials, net.cur_partials, net.partial_stride, conv.start, conv.numel, not pushed,
                                opt.hyper.code(), self.hyper_dev, self.step_counter, self._comm_struct(0, False),
                                None if self._loss_posted else net.loss, self.loss_sink, last=True)

    def _dense_update(self, start: int, count: int) -> None:
        """Push protocol, dense buckets (FC.*, conv3: 95 % of the gradient): ONE launch pushes them to every peer, waits
        for the peers' lines and updates — on the side stream, beside the convolution backward, so that the step's last
        launch only has the 10 KB convolution bucket left (at 8 GPUs gathering 7 peers' lines for the whole arena in the
        last launch cost ~3 us of exposed time)."""
        opt = self.optimizer
        native.reduce_update_ll(self.arena.param, self.arena.grad, opt.state0, opt.state1, None, start, count, None, 0, 0,
                                0, 0, True, opt.hyper.code(), self.hyper_dev, self.step_counter,
                                self._comm_struct(0, False), None, None, last=False)

    def _step_body(self, x: torch.Tensor, y: torch.Tensor) -> None:
        main = torch.cuda.current_stream(self.device)
        native.nvtx_push("b200dp.forward")
        self.net.forward(x, y, train=True)
        native.nvtx_pop()
        side = self.side
        # the loss goes to the host from a one-thread kernel beside the backward pass (see b200dp_post_loss)
        self._loss_posted = side is not None
        if side is not None:
            side.wait_stream(main)
            with torch.cuda.stream(side):
                native.post_loss(self.net.loss, self.step_counter, self.loss_sink)

        bk = self.arena.layout.buckets
        both = self.net.head_fused and bk[1].stop == bk[0].start     # the two dense buckets finish in one launch

        def on_bucket(i: int) -> None:
            if self.fused_final:
                last = i == 2
                if both and i == 0:
                    return                           # sent together with bucket 1 (adjacent in the arena)
                if side is None:                     # no overlap: everything after backward, in one launch
                    if last:
                        self._final_update(pushed=False)
                    return
                if last:
                    # the step's final launch runs on the MAIN stream, right behind the last backward kernel (programmatic
                    # dependent launch: its CTAs are resident when that kernel drains), after joining the side stream: the
                    # pushes and the loss post read the generation this launch advances
                    main.wait_stream(side)
                    self._final_update(pushed=True, conv_only=both)
                    return
                if not both and (self.comm is None or self._comm_off):
                    return                           # nothing to send at world size 1
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    if both:                         # push + wait + update of the dense buckets, beside the conv backward
                        self._dense_update(bk[1].start, bk[0].stop - bk[1].start)
                    else:
                        self._push(i)
                return
            if side is None:
                self._reduce_update(i)
                return
            side.wait_stream(main)
            with torch.cuda.stream(side):
                self._reduce_update(i)

        native.nvtx_push("b200dp.backward+exchange+update")
        self.net.backward(x, on_bucket)
        if side is not None:
            main.wait_stream(side)
        native.nvtx_pop()

    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self._sync_hyper()
        self.optimizer.step_count += 1
        b = x.shape[0]
        if not self.use_graph or b != self.net.B:
            self._step_body(x, y)
            return self.net.loss
        if self._warm < 2:                       # lazy module loading etc. must happen outside capture
            self._step_body(x, y)
            self._warm += 1
            return self.net.loss
        key = (x.data_ptr(), y.data_ptr(), b)
        gph = self._graphs.get(key)
        if gph is None:
            if len(self._graphs) - self._eval_graphs >= self.MAX_GRAPHS:
                return self._replay_static(x, y)
            gph = self._capture(x, y)
            self._graphs[key] = gph
        gph.replay()
        return self.net.loss

    def _capture(self, x: torch.Tensor, y: torch.Tensor) -> torch.cuda.CUDAGraph:
        torch.cuda.synchronize(self.device)
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, capture_error_mode="thread_local"):
            self._step_body(x, y)
        return gph

    def _replay_static(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self._static_x is None:
            self._static_x, self._static_y = torch.empty_like(x), torch.empty_like(y)
            self._graphs[("static",)] = self._capture(self._static_x, self._static_y)
        self._static_x.copy_(x)
        self._static_y.copy_(y)
        self._graphs[("static",)].replay()
        return self.net.loss

    def make_pump(self, images: torch.Tensor, targets: torch.Tensor, batch: int, depth: int = 6):
        """A native step pump (csrc/pump.cu) over the host arrays ``images`` (uint8) / ``targets``: queues whole
        runs of steps — sample gather into a pinned ring, H2D copy, this stepper's step, loss read-back — from
        a C++ loop.  See :class:`b200dp.data.pump.StepPump`."""
        from ..data.pump import StepPump
        pump = StepPump(self, images, targets, batch, depth=depth)
        self._pumps.append(pump)
        return pump

    def set_comm_enabled(self, on: bool) -> None:
        """Measurement aid for the exposed-communication number: with ``on=False`` the reduce+update
        launches skip the cross-rank exchange (each rank applies its local gradient), everything else
        stays the same.  Replicas diverge while it is off — use only at the end of a benchmark."""
        self._comm_off = not on
        self._graphs.clear()
        self._eval_graphs = 0
        for p in self._pumps:
            p.invalidate()
        self._static_x = self._static_y = None

    # ---- eval ----------------------------------------------------------------------
    def eval_batch(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Forward + device-side metric accumulation (train.py:60-72 without the two host syncs per batch).  Full
        batches replay a CUDA graph per input buffer (the loader's device ring has three), like the training step."""
        native.nvtx_push("b200dp.eval_forward")
        key = ("eval", x.data_ptr(), y.data_ptr(), x.shape[0])
        if not (self.use_graph and x.shape[0] == self.net.B and self._warm >= 2):
            self.net.forward(x, y, train=False)
        else:
            gph = self._graphs.get(key)
            if gph is None and self._eval_graphs < self.MAX_GRAPHS:
                if not self._eval_warm:
                    self.net.forward(x, y, train=False)          # lazy set-up outside capture, once
                    self._eval_warm = True
                    native.nvtx_pop()
                    return
                torch.cuda.synchronize(self.device)
                gph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gph, capture_error_mode="thread_local"):
                    self.net.forward(x, y, train=False)
                self._graphs[key] = gph
                self._eval_graphs += 1
            if gph is None:
                self.net.forward(x, y, train=False)
            else:
                gph.replay()
        native.nvtx_pop()

    def eval_result(self, reset: bool = True) -> Tuple[float, int]:
        vals = self.net.eval_acc.tolist()        # the one D2H of the epoch
        if reset:
            self.net.eval_acc.zero_()
        return float(vals[0]), int(vals[1])

    # ---- bookkeeping -----------------------------------------------------------------
    def params_changed(self) -> None:
        """Call after loading a checkpoint (collective at world > 1): re-seed the device generation counter and
        forget the generation numbers the peers have seen."""
        self.step_counter[0] = self.optimizer.step_count
        self._hyper_cache = None
        if self.comm is not None:
            self.comm.reset()

    def error_flag(self) -> int:
        """The device's sticky error word, read WITHOUT waiting for the queued steps and without a copy: the last kernel
        of the step that raised it (and of every later step) stores it into a mapped host word (loss_sink[3]; see
        ru_finish_step in csrc/opt_math.cuh).  A `.item()` on the compute stream at every log point drained the step
        pump's queue — one pipeline bubble per log interval — and a 4-byte copy on a stream of its own still cost 30-60 us
        of host time per check, the last one of an epoch with the GPU idle.  The word is sticky, so a check that runs
        while the failing step is still in flight only delays the report by a log interval."""
        return int(self._err_np[0])

    def check_error(self) -> None:
        """Raise if a peer did not deliver (synchronises with the compute stream first: for the end of a run / tests)."""
        torch.cuda.current_stream(self.device).synchronize()
        if self.error_flag() != 0:
            raise RuntimeError("a peer rank did not deliver its gradients within the time bound (dead or stalled "
                               "rank?); this rank stopped updating its parameters at that step")

    @torch.no_grad()
    def replica_divergence(self) -> float:
        if self.world == 1:
            return 0.0
        ref = self.arena.param.clone()
        dist.broadcast(ref, src=0)
        d = (self.arena.param - ref).abs().max().reshape(1)
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        return float(d.item())

    def close(self) -> None:
        for p in self._pumps:
            p.close()
        self._pumps.clear()
        self._graphs.clear()
        if self.comm is not None:
            self.comm.close()
            self.comm = None
