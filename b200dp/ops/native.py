"""ctypes bindings to the in-tree CUDA library (include/b200dp.h, csrc/*.cu).

The library is plain C ABI + raw device pointers, so the binding is a thin
layer: ``tensor.data_ptr()`` in, ``torch.cuda.current_stream().cuda_stream``
as the launch stream (so launches are captured by ``torch.cuda.graph``).
Nothing here falls back to torch ops: if the library is missing on a GPU box
the import raises (the driver checks that the native code is what ran).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_native", "libb200dp.so")
_lib = None


class Dims(C.Structure):
    _fields_ = [("batch", C.c_int), ("in_ch", C.c_int), ("in_hw", C.c_int), ("c1", C.c_int),
                ("c2", C.c_int), ("c3", C.c_int), ("f1", C.c_int), ("ncls", C.c_int), ("ld_p2", C.c_int)]


class BwdProblem(C.Structure):
    _fields_ = [("kind", C.c_int), ("dy", C.c_void_p), ("w_or_x", C.c_void_p), ("act_prev", C.c_void_p),
                ("out", C.c_void_p), ("db", C.c_void_p), ("M", C.c_int), ("N", C.c_int), ("K", C.c_int)]


class TcProblem(C.Structure):
    _fields_ = [("a", C.c_void_p), ("lda", C.c_int), ("a_mn_major", C.c_int),
                ("b", C.c_void_p), ("ldb", C.c_int), ("b_mn_major", C.c_int),
                ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
                ("bias", C.c_void_p), ("act", C.c_int),
                ("mask", C.c_void_p), ("ld_mask", C.c_int),
                ("out", C.c_void_p), ("ld_out", C.c_int),
                ("colext", C.c_void_p)]


class Comm(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("algo", C.c_int),
                ("stage", C.c_void_p * 8), ("flags", C.c_void_p * 8), ("mc_stage", C.c_void_p),
                ("arena_len", C.c_int), ("max_slices", C.c_int), ("max_ctas", C.c_int),
                ("ll", C.c_void_p * 8), ("wire", C.c_int), ("timeout_ms", C.c_uint32)]


class PumpDesc(C.Structure):
    _fields_ = [("images", C.c_void_p), ("targets", C.c_void_p), ("image_bytes", C.c_int64),
                ("batch", C.c_int), ("depth", C.c_int),
                ("host_x", C.c_void_p * 8), ("host_y", C.c_void_p * 8), ("dev_x", C.c_void_p * 8),
                ("dev_y", C.c_void_p * 8), ("graph_exec", C.c_void_p * 8)]


def lib_path() -> str:
    return _LIB_PATH


def load(build_if_missing: bool = True):
    """dlopen the library (building it first if the sources are newer / it is absent)."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from .. import build as _build
        try:
            _build.build()
        except Exception as e:  # no nvcc on this box: use what was shipped
            if not os.path.exists(_LIB_PATH):
                raise RuntimeError(f"libb200dp.so is missing and could not be built: {e}") from e
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(f"{_LIB_PATH} not found; run `python -m b200dp.build`")
    lib = C.CDLL(_LIB_PATH)
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    lib.b200dp_last_error.restype = C.c_char_p
    lib.b200dp_conv1_fwd.argtypes = [vp, i, f, f, vp, vp, vp, vp, Dims, vp]
    lib.b200dp_conv2_fwd.argtypes = [vp, vp, vp, vp, vp, Dims, vp]
    lib.b200dp_linear_fwd.argtypes = [vp, vp, vp, vp, i, i, i, i, vp]
    lib.b200dp_head_tail.argtypes = [vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, vp]
    lib.b200dp_head_fused_supported.argtypes = [i, i, i, i]
    lib.b200dp_head_fused_set_debug.argtypes = [vp]
    lib.b200dp_head_fused_set_debug.restype = None
    lib.b200dp_head_fused_set_tma.argtypes = [i]
    lib.b200dp_head_fused_set_tma.restype = None
    lib.b200dp_head_fused.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, vp, i, vp, i, vp, i, vp, vp, vp, vp, i, vp, vp, vp, vp,
                                      i, i, i, i, i, vp]
    lib.b200dp_linear_bwd_group.argtypes = [C.POINTER(BwdProblem), i, vp]
    lib.b200dp_tc_gemm_group.argtypes = [C.POINTER(TcProblem), i, vp]
    lib.b200dp_tc_gemm_group_splitk.argtypes = [C.POINTER(TcProblem), i, vp, C.c_long, vp, i, vp]
    lib.b200dp_tc_gemm_ws_floats.argtypes = [C.POINTER(TcProblem), i, C.POINTER(C.c_int)]
    lib.b200dp_tc_gemm_ws_floats.restype = C.c_long
    lib.b200dp_conv_bwd_num_partials.argtypes = [Dims]
    lib.b200dp_tcw_pack_floats.argtypes = [Dims]
    lib.b200dp_tcw_pack_floats.restype = C.c_long
    lib.b200dp_tcw_pack_weights.argtypes = [vp, vp, vp, Dims, vp]
    lib.b200dp_tcw_conv1_fwd.argtypes = [vp, i, f, f, vp, vp, vp, vp, Dims, vp]
    lib.b200dp_tcw_conv2_fwd.argtypes = [vp, vp, vp, vp, vp, Dims, vp]
    lib.b200dp_tcw_conv2_dgrad.argtypes = [vp, vp, vp, vp, vp, vp, Dims, vp]
    lib.b200dp_tcw_set_debug.argtypes = [vp]
    lib.b200dp_tcw_conv_wgrad_num_partials.argtypes = [Dims]
    lib.b200dp_tcw_conv_wgrad.argtypes = [i, vp, i, f, f, vp, vp, i, vp, vp, i, i, i, Dims, vp]
    lib.b200dp_tcw_set_debug.restype = None
    lib.b200dp_conv_bwd.argtypes = [vp, i, f, f, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, Dims, vp]
    lib.b200dp_reduce_update.argtypes = [vp, vp, vp, vp, vp, i, i, vp, i, i, i, vp, vp, i, C.POINTER(Comm), vp, vp, vp]
    lib.b200dp_grad_push.argtypes = [vp, i, i, vp, C.POINTER(Comm), vp]
    lib.b200dp_reduce_update_ll.argtypes = [vp, vp, vp, vp, vp, i, i, vp, i, i, i, i, i, i, i, vp, vp, C.POINTER(Comm), vp, vp, vp]
    lib.b200dp_post_loss.argtypes = [vp, vp, vp, vp]
    lib.b200dp_symm_alloc.argtypes = [C.POINTER(vp), C.c_uint64]
    lib.b200dp_symm_free.argtypes = [vp]
    lib.b200dp_symm_zero.argtypes = [vp, C.c_uint64]
    lib.b200dp_symm_export.argtypes = [vp, C.c_char_p]
    lib.b200dp_symm_import.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.b200dp_symm_close.argtypes = [vp]
    lib.b200dp_fetch_slot.argtypes = [vp, vp, C.c_int64, vp]
    lib.b200dp_pump_create.argtypes = [C.POINTER(vp), C.POINTER(PumpDesc)]
    lib.b200dp_pump_run.argtypes = [vp, vp, i, vp, C.c_uint64, C.c_uint64, vp]
    lib.b200dp_pump_pregather.argtypes = [vp, vp, i]
    lib.b200dp_pump_wait_first.argtypes = [vp]
    lib.b200dp_pump_drain.argtypes = [vp]
    lib.b200dp_pump_destroy.argtypes = [vp]
    lib.b200dp_nvtx_push.argtypes = [C.c_char_p]
    lib.b200dp_nvtx_push.restype = None
    lib.b200dp_nvtx_pop.restype = None
    _lib = lib
    return lib


def is_loaded() -> bool:
    return _lib is not None


class NativeError(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200dp_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed (cuda error {rc}): {msg}")


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _f32(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda):
            raise TypeError("expected contiguous CUDA float32 tensors")


# ---- forward -----------------------------------------------------------------
def conv1_fwd(x, w, b, p1, idx1, dims: Dims, mean: float, std: float) -> None:
    _f32(w, b, p1)
    check(load().b200dp_conv1_fwd(_p(x), int(x.dtype == torch.uint8), mean, std, _p(w), _p(b), _p(p1), _p(idx1),
                                  dims, _stream()), "conv1_fwd")


def conv2_fwd(p1, w, b, p2, idx2, dims: Dims) -> None:
    _f32(p1, w, b, p2)
    check(load().b200dp_conv2_fwd(_p(p1), _p(w), _p(b), _p(p2), _p(idx2), dims, _stream()), "conv2_fwd")


# ---- window-tile tcgen05 convolutions (csrc/tc_win.cu) ----
def tcw_pack_floats(dims: Dims) -> int:
    n = int(load().b200dp_tcw_pack_floats(dims))
    if n < 0:
        raise NativeError("unsupported geometry for the window-tile convolutions")
    return n


def tcw_pack_weights(w1, w2, wpack, dims: Dims) -> None:
    _f32(w1, w2, wpack)
    check(load().b200dp_tcw_pack_weights(_p(w1), _p(w2), _p(wpack), dims, _stream()), "tcw_pack_weights")


def tcw_conv1_fwd(x, wpack, b, p1, idx1, dims: Dims, mean: float, std: float) -> None:
    check(load().b200dp_tcw_conv1_fwd(_p(x), int(x.dtype == torch.uint8), mean, std, _p(wpack), _p(b), _p(p1),
                                      _p(idx1), dims, _stream()), "tcw_conv1_fwd")


def tcw_conv2_fwd(p1, wpack, b, p2, idx2, dims: Dims) -> None:
    check(load().b200dp_tcw_conv2_fwd(_p(p1), _p(wpack), _p(b), _p(p2), _p(idx2), dims, _stream()), "tcw_conv2_fwd")


def tcw_conv2_dgrad(dp2, p2, idx2, wpack, p1, g1, dims: Dims) -> None:
    check(load().b200dp_tcw_conv2_dgrad(_p(dp2), _p(p2), _p(idx2), _p(wpack), _p(p1), _p(g1), dims, _stream()),
          "tcw_conv2_dgrad")


def tcw_conv_wgrad_num_partials(dims: Dims) -> int:
    return int(load().b200dp_tcw_conv_wgrad_num_partials(dims))


def tcw_conv_wgrad(layer: int, x, mean, std, g, act, ld_act: int, idx, partials, stride: int, off_w: int, off_b: int,
                   dims: Dims) -> None:
    check(load().b200dp_tcw_conv_wgrad(layer, _p(x), int(x.dtype == torch.uint8), mean, std, _p(g), _p(act), ld_act,
                                       _p(idx), _p(partials), stride, off_w, off_b, dims, _stream()), "tcw_conv_wgrad")


def linear_fwd(x, w, b, y, act: int) -> None:
    _f32(x, w, b, y)
    M, K = x.shape[0], x.numel() // x.shape[0]
    N = w.shape[0]
    check(load().b200dp_linear_fwd(_p(x), _p(w), _p(b), _p(y), M, N, K, act, _stream()), "linear_fwd")


def head_tail(h1, w, b, target, loss_mode: int, probs, loss_rows, loss_out, dlogits, dz1, eval_acc, ticket,
              ld_h1: int = 0, ld_dlogits: int = 0) -> None:
    """``h1`` is [B][ld_h1] (ld 0 = dense F1 columns); F1 is taken from ``w``."""
    B, F1 = h1.shape[0], w.shape[1]
    check(load().b200dp_head_tail(_p(h1), _p(w), _p(b), _p(target), loss_mode, _p(probs), _p(loss_rows),
                                  _p(loss_out), _p(dlogits), _p(dz1), _p(eval_acc), _p(ticket), B, F1,
                                  w.shape[0], ld_h1, ld_dlogits, _stream()), "head_tail")


def head_fused_supported(K3: int, C3: int, F1: int, ncls: int) -> bool:
    return bool(load().b200dp_head_fused_supported(K3, C3, F1, ncls))


def head_fused(p2, ld_p2: int, w3, b3, w1, b1, w2, b2, target, loss_mode: int, a3, ld_a3: int, h1, ld_h1: int, probs,
               loss_rows, loss_out, dlogits, ld_dlogits: int, dz1, dz3, eval_acc, ticket, B: int) -> None:
    """conv3 + FC.0 + FC.2 + softmax + loss (+ dlogits, dz1, dz3 when training) in one launch (csrc/head_fused.cu)."""
    C3, K3 = w3.shape[0], w3.numel() // w3.shape[0]
    F1, ncls = w1.shape[0], w2.shape[0]
    check(load().b200dp_head_fused(_p(p2), ld_p2, _p(w3), _p(b3), _p(w1), _p(b1), _p(w2), _p(b2), _p(target), loss_mode,
                                   _p(a3), ld_a3, _p(h1), ld_h1, _p(probs), _p(loss_rows), _p(loss_out), _p(dlogits),
                                   ld_dlogits, _p(dz1), _p(dz3), _p(eval_acc), _p(ticket), B, K3, C3, F1, ncls, _stream()),
          "head_fused")


# ---- backward ----------------------------------------------------------------
def bwd_problem(kind: int, dy, w_or_x, act_prev, out, db, M: int, N: int, K: int) -> BwdProblem:
    return BwdProblem(kind, _p(dy), _p(w_or_x), _p(act_prev), _p(out), _p(db), M, N, K)


def linear_bwd_group(problems: Sequence[BwdProblem]) -> None:
    arr = (BwdProblem * len(problems))(*problems)
    check(load().b200dp_linear_bwd_group(arr, len(problems), _stream()), "linear_bwd_group")


def tc_problem(a, lda: int, a_mn: bool, b, ldb: int, b_mn: bool, M: int, N: int, K: int, *, bias=None, act: int = 0,
               mask=None, ld_mask: int = 0, out=None, ld_out: int = 0, colext=None) -> TcProblem:
    """One tensor-core GEMM problem: out[M][N] = epi(sum_k A(m,k) B(n,k)); see include/b200dp.h."""
    return TcProblem(_p(a), lda, int(a_mn), _p(b), ldb, int(b_mn), M, N, K, _p(bias), act, _p(mask), ld_mask,
                     _p(out), ld_out, _p(colext))


class GemmWorkspace:
    """Split-K scratch of the tensor-core GEMM launches of one stream: partial tiles + per-tile tickets.  Grows on
    demand (outside CUDA-graph capture: run the launch sequence once eagerly first)."""

    def __init__(self, device):
        self.device = device
        self.ws = torch.empty(0, dtype=torch.float32, device=device)
        self.tickets = torch.zeros(0, dtype=torch.int32, device=device)

    def reserve(self, floats: int, tickets: int) -> None:
        if floats > self.ws.numel():
            self.ws = torch.empty(floats, dtype=torch.float32, device=self.device)
        if tickets > self.tickets.numel():
            self.tickets = torch.zeros(tickets, dtype=torch.int32, device=self.device)


def tc_gemm_group(problems: Sequence[TcProblem], workspace: Optional[GemmWorkspace] = None) -> None:
    arr = (TcProblem * len(problems))(*problems)
    if workspace is None:
        check(load().b200dp_tc_gemm_group(arr, len(problems), _stream()), "tc_gemm_group")
        return
    nt = C.c_int(0)
    need = int(load().b200dp_tc_gemm_ws_floats(arr, len(problems), C.byref(nt)))
    workspace.reserve(need, nt.value)
    check(load().b200dp_tc_gemm_group_splitk(arr, len(problems), _p(workspace.ws) if need else None, workspace.ws.numel(),
                                             _p(workspace.tickets) if nt.value else None, workspace.tickets.numel(),
                                             _stream()), "tc_gemm_group_splitk")


def conv_bwd_num_partials(dims: Dims) -> int:
    return int(load().b200dp_conv_bwd_num_partials(dims))


def conv_bwd(x, mean, std, p1, idx1, p2, idx2, dp2, w2, partials, stride, off_w1, off_b1, off_w2, off_b2,
             dims: Dims) -> None:
    check(load().b200dp_conv_bwd(_p(x), int(x.dtype == torch.uint8), mean, std, _p(p1), _p(idx1), _p(p2),
                                 _p(idx2), _p(dp2), _p(w2), _p(partials), stride, off_w1, off_b1, off_w2,
                                 off_b2, dims, _stream()), "conv_bwd")


# ---- reduce + update -----------------------------------------------------------
def reduce_update(param, grad, state0, state1, shadow, start: int, count: int, partials, n_partials: int,
                  partial_stride: int, opt_kind: int, hyper, step_counter, last_bucket: bool,
                  comm: Optional[Comm], loss_src=None, loss_sink=None) -> None:
    check(load().b200dp_reduce_update(_p(param), _p(grad), _p(state0), _p(state1), _p(shadow), start, count,
                                      _p(partials), n_partials, partial_stride, opt_kind, _p(hyper),
                                      _p(step_counter), int(last_bucket),
                                      C.byref(comm) if comm is not None else None, _p(loss_src), _p(loss_sink),
                                      _stream()), "reduce_update")


def grad_push(grad, start: int, count: int, step_counter, comm: Comm) -> None:
    """Fire-and-forget: write grad[start, start+count) into every peer's receive area (csrc/ll_reduce.cu)."""
    check(load().b200dp_grad_push(_p(grad), start, count, _p(step_counter), C.byref(comm), _stream()), "grad_push")


def reduce_update_ll(param, grad, state0, state1, shadow, start: int, count: int, partials, n_partials: int,
                     partial_stride: int, p_start: int, p_count: int, push_rest: bool, opt_kind: int, hyper,
                     step_counter, comm: Optional[Comm], loss_src=None, loss_sink=None, last: bool = True) -> None:
    """The step's single wait + rank-order sum + average + optimizer launch of the push protocol."""
    check(load().b200dp_reduce_update_ll(_p(param), _p(grad), _p(state0), _p(state1), _p(shadow), start, count,
                                         _p(partials), n_partials, partial_stride, p_start, p_count, int(push_rest),
                                         int(last), opt_kind, _p(hyper), _p(step_counter),
                                         C.byref(comm) if comm is not None else None, _p(loss_src), _p(loss_sink),
                                         _stream()), "reduce_update_ll")


def post_loss(loss_src, step_counter, loss_sink) -> None:
    """One-thread kernel: the step's loss into the host's mapped array (see include/b200dp.h)."""
    check(load().b200dp_post_loss(_p(loss_src), _p(step_counter), _p(loss_sink), _stream()), "post_loss")


# ---- symmetric memory ------------------------------------------------------------
def symm_alloc(nbytes: int) -> int:
    ptr = C.c_void_p()
    check(load().b200dp_symm_alloc(C.byref(ptr), nbytes), "symm_alloc")
    return ptr.value


def symm_zero(ptr: int, nbytes: int) -> None:
    check(load().b200dp_symm_zero(ptr, nbytes), "symm_zero")


def symm_free(ptr: int) -> None:
    check(load().b200dp_symm_free(ptr), "symm_free")


def symm_export(ptr: int) -> bytes:
    buf = C.create_string_buffer(64)
    check(load().b200dp_symm_export(ptr, buf), "symm_export")
    return buf.raw


def symm_import(handle: bytes) -> int:
    ptr = C.c_void_p()
    check(load().b200dp_symm_import(handle, C.byref(ptr)), "symm_import")
    return ptr.value


def symm_close(ptr: int) -> None:
    check(load().b200dp_symm_close(ptr), "symm_close")


# ---- native step pump ------------------------------------------------------------
def fetch_slot(dst: torch.Tensor, src_pinned: torch.Tensor) -> None:
    """Upload a pinned host slot into a device slot with a small kernel (the prefetch branch of a step graph)."""
    nbytes = src_pinned.numel() * src_pinned.element_size()
    assert dst.is_cuda and src_pinned.is_pinned() and dst.numel() * dst.element_size() == nbytes
    check(load().b200dp_fetch_slot(dst.data_ptr(), src_pinned.data_ptr(), nbytes, _stream()), "fetch_slot")


def pump_create(desc: PumpDesc) -> int:
    h = C.c_void_p()
    check(load().b200dp_pump_create(C.byref(h), C.byref(desc)), "pump_create")
    return h.value


def pump_run(handle: int, order: torch.Tensor, n_steps: int, order_offset: int = 0, sink_dev: int = 0,
             loss_ring_host: int = 0, gen0: int = 0) -> None:
    """``order``: contiguous CPU int64 sample indices; n_steps*batch of them are read from element ``order_offset`` on.
    ``sink_dev``: device address of the stepper's loss-sink descriptor (0: leave it alone), ``loss_ring_host``: pinned
    float32 array that receives the run's losses, ``gen0``: generation of the run's first step."""
    assert order.dtype == torch.int64 and order.is_contiguous() and not order.is_cuda
    assert order.numel() - order_offset >= n_steps
    check(load().b200dp_pump_run(handle, order.data_ptr() + 8 * order_offset, n_steps, sink_dev or None, loss_ring_host,
                                 gen0, _stream()), "pump_run")


def pump_pregather(handle: int, order: torch.Tensor, n_steps: int) -> None:
    """Gather the first batches of the next run into their ring slots now (see b200dp_pump_pregather)."""
    assert order.dtype == torch.int64 and order.is_contiguous() and not order.is_cuda
    check(load().b200dp_pump_pregather(handle, order.data_ptr(), n_steps), "pump_pregather")


def pump_wait_first(handle: int) -> None:
    check(load().b200dp_pump_wait_first(handle), "pump_wait_first")


def pump_drain(handle: int) -> None:
    check(load().b200dp_pump_drain(handle), "pump_drain")


def pump_destroy(handle: int) -> None:
    check(load().b200dp_pump_destroy(handle), "pump_destroy")


def nvtx_push(name: str) -> None:
    load().b200dp_nvtx_push(name.encode())


def nvtx_pop() -> None:
    load().b200dp_nvtx_pop()
