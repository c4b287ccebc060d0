"""Native step pump front end (csrc/pump.cu): an epoch's host work in a C++ loop.

The reference's hot loop (train.py:44-57) pays, per batch, a DataLoader ``__next__``, two synchronous
``.cuda()`` copies and the Python of the step itself.  Here one ctypes call queues a whole run of
steps: a C++ loop gathers each step's shuffled samples of this rank's shard into a pinned ring slot,
launches one CUDA graph (H2D copies of that slot + the entire training step) and queues the 4-byte
loss read-back into a pinned array.  Python only sees chunks of steps (one chunk per log line).

:class:`StepPump` owns the pinned/device rings and the per-slot graphs of one
:class:`~b200dp.models.fused.NativeStepper`.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import native


class StepPump:
    """``run(order, n_steps)`` queues ``n_steps`` full-batch training steps; sample ``j`` of step ``k`` is
    base row ``order[k*batch + j]``.  Ragged batches do not go through the pump (the graphs are
    captured for the full batch): feed them to ``stepper.train_step``."""

    def __init__(self, stepper, images: torch.Tensor, targets: torch.Tensor, batch: int, *, depth: int = 3):
        assert not images.is_cuda and images.is_contiguous() and images.dtype == torch.uint8, \
            "the pump streams raw uint8 samples (the first kernel normalises on the device)"
        assert targets.dtype == torch.int64 and targets.is_contiguous()
        assert 2 <= depth <= 8
        self.stepper, self.images, self.targets = stepper, images, targets
        self.batch, self.depth = int(batch), int(depth)
        assert self.batch == stepper.net.B, "the pump runs the stepper's full batch"
        dev = stepper.device
        shape = (self.batch,) + tuple(images.shape[1:])
        self.host_x = [torch.empty(shape, dtype=torch.uint8).pin_memory() for _ in range(depth)]
        self.host_y = [torch.empty(self.batch, dtype=torch.int64).pin_memory() for _ in range(depth)]
        self.dev_x = [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(depth)]
        self.dev_y = [torch.empty(self.batch, dtype=torch.int64, device=dev) for _ in range(depth)]
        self.h2d_bytes_per_step = self.host_x[0].numel() + 8 * self.batch
        self._graphs = []
        self._handle: Optional[int] = None
        self._loss_keepalive = None

    # graphs are captured lazily: the stepper wants two eager steps first (lazy module loading)
    def _ensure_graphs(self) -> None:
        if self._handle is not None:
            return
        st = self.stepper
        assert st._warm >= 2, "run two eager steps before capturing"
        torch.cuda.synchronize(st.device)
        for s in range(self.depth):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                self.dev_x[s].copy_(self.host_x[s], non_blocking=True)     # H2D memcpy nodes
                self.dev_y[s].copy_(self.host_y[s], non_blocking=True)
                st._step_body(self.dev_x[s], self.dev_y[s])
            self._graphs.append(g)
        d = native.PumpDesc()
        d.images, d.targets = self.images.data_ptr(), self.targets.data_ptr()
        d.image_bytes = self.images[0].numel()
        d.batch, d.depth = self.batch, self.depth
        for s in range(self.depth):
            d.host_x[s], d.host_y[s] = self.host_x[s].data_ptr(), self.host_y[s].data_ptr()
            d.graph_exec[s] = int(self._graphs[s].raw_cuda_graph_exec())
        d.loss_dev = st.net.loss.data_ptr()
        self._handle = native.pump_create(d)

    def invalidate(self) -> None:
        """Drop the captured graphs (the stepper's launch sequence changed)."""
        self.close()

    def _eager_step(self, idx: torch.Tensor) -> torch.Tensor:
        torch.index_select(self.images, 0, idx, out=self.host_x[0])
        torch.index_select(self.targets, 0, idx, out=self.host_y[0])
        self.dev_x[0].copy_(self.host_x[0], non_blocking=True)
        self.dev_y[0].copy_(self.host_y[0], non_blocking=True)
        loss = self.stepper.train_step(self.dev_x[0], self.dev_y[0])
        torch.cuda.current_stream(self.stepper.device).synchronize()   # host_x[0] is reused right away
        return loss

    def run(self, order: torch.Tensor, n_steps: int, loss_host: Optional[torch.Tensor] = None) -> None:
        """Queue ``n_steps`` steps (returns before they finish).  ``loss_host``: pinned float32 [>= n_steps]."""
        st = self.stepper
        k0 = 0
        while st._warm < 2 and k0 < n_steps:            # the stepper's first two steps run eagerly
            loss = self._eager_step(order[k0 * self.batch:(k0 + 1) * self.batch])
            if loss_host is not None:
                loss_host[k0:k0 + 1].copy_(loss)
            k0 += 1
        if k0 == n_steps:
            return
        self._ensure_graphs()
        st._sync_hyper()
        n = n_steps - k0
        st.optimizer.step_count += n
        self._loss_keepalive = loss_host
        native.pump_run(self._handle, order[k0 * self.batch:], n, None if loss_host is None else loss_host[k0:])

    def wait_first(self) -> None:
        if self._handle is not None:
            native.pump_wait_first(self._handle)

    def drain(self) -> None:
        if self._handle is not None:
            native.pump_drain(self._handle)

    def close(self) -> None:
        if self._handle is not None:
            native.pump_destroy(self._handle)
            self._handle = None
        self._graphs.clear()
