"""Native step pump front end (csrc/pump.cu): an epoch's host work in a C++ loop.

The reference's hot loop (train.py:44-57) pays, per batch, a DataLoader ``__next__``, two synchronous
``.cuda()`` copies and the Python of the step itself.  Here one ctypes call queues a whole run of
steps: a C++ loop gathers each step's shuffled samples of this rank's shard into a pinned ring slot,
launches one CUDA graph per step — the entire training step, with the upload of the NEXT step's slot as
a parallel branch — and every step's loss lands in a pinned array, stored there by the step's last
kernel through mapped memory.  Python only sees chunks of steps (one chunk per log line).

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

    def __init__(self, stepper, images: torch.Tensor, targets: torch.Tensor, batch: int, *, depth: int = 6):
        assert not images.is_cuda and images.is_contiguous() and images.dtype in (torch.uint8, torch.float32), \
            "the pump streams raw uint8 samples (the first kernel normalises on the device) or float32 samples that " \
            "are already transformed"
        assert targets.dtype == torch.int64 and targets.is_contiguous()
        assert 2 <= depth <= 8
        self.stepper, self.images, self.targets = stepper, images, targets
        self.batch, self.depth = int(batch), int(depth)
        assert self.batch == stepper.net.B, "the pump runs the stepper's full batch"
        dev = stepper.device
        shape = (self.batch,) + tuple(images.shape[1:])
        # a ring slot is ONE allocation, [images | pad to 16 B | labels], on the host (pinned) and on the device: a step's
        # upload is a single copy node in its graph (two nodes cost ~2 us more per step)
        x_bytes = self.batch * images[0].numel() * images.element_size()
        y_off = (x_bytes + 15) // 16 * 16
        slot_bytes = (y_off + 8 * self.batch + 15) // 16 * 16
        self.host_slot = [torch.empty(slot_bytes, dtype=torch.uint8).pin_memory() for _ in range(depth)]
        self.dev_slot = [torch.empty(slot_bytes, dtype=torch.uint8, device=dev) for _ in range(depth)]
        view_x = lambda t: t[:x_bytes].view(images.dtype).view(shape)        # noqa: E731
        view_y = lambda t: t[y_off:y_off + 8 * self.batch].view(torch.int64)   # noqa: E731
        self.host_x, self.host_y = [view_x(t) for t in self.host_slot], [view_y(t) for t in self.host_slot]
        self.dev_x, self.dev_y = [view_x(t) for t in self.dev_slot], [view_y(t) for t in self.dev_slot]
        self.h2d_bytes_per_step = x_bytes + 8 * self.batch
        self._graphs = []
        self._handle: Optional[int] = None
        self._loss_keepalive = None
        self._sink_ptr = stepper.loss_sink.data_ptr()

    # graphs are captured lazily: the stepper wants two eager steps first (lazy module loading)
    def _ensure_graphs(self) -> None:
        if self._handle is not None:
            return
        st = self.stepper
        assert st._warm >= 2, "run two eager steps before capturing"
        torch.cuda.synchronize(st.device)
        copy_stream = torch.cuda.Stream(device=st.device)
        for s in range(self.depth):
            nxt = (s + 1) % self.depth
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                main = torch.cuda.current_stream(st.device)

                def prefetch(nxt=nxt):
                    # branch 1: upload the NEXT step's batch, by a small kernel that reads the pinned slot over PCIe (a
                    # memcpy node here cost ~4 us per step); forked after the step's first kernel
                    copy_stream.wait_stream(main)
                    with torch.cuda.stream(copy_stream):
                        native.fetch_slot(self.dev_slot[nxt], self.host_slot[nxt])   # images + labels, one small kernel

                st.net.after_first_kernel = prefetch
                try:
                    st._step_body(self.dev_x[s], self.dev_y[s])   # branch 2: this step
                finally:
                    st.net.after_first_kernel = None
                main.wait_stream(copy_stream)
            self._graphs.append(g)
        d = native.PumpDesc()
        d.images, d.targets = self.images.data_ptr(), self.targets.data_ptr()
        d.image_bytes = self.images[0].numel() * self.images.element_size()
        d.batch, d.depth = self.batch, self.depth
        for s in range(self.depth):
            d.host_x[s], d.host_y[s] = self.host_x[s].data_ptr(), self.host_y[s].data_ptr()
            d.dev_x[s], d.dev_y[s] = self.dev_x[s].data_ptr(), self.dev_y[s].data_ptr()
            d.graph_exec[s] = int(self._graphs[s].raw_cuda_graph_exec())
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
        """Queue ``n_steps`` steps (returns before they finish).  ``loss_host``: pinned float32 [>= n_steps];
        element k receives step k's loss, stored by the step's last kernel (mapped memory, no copy command)."""
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
        # the run tells the device where its losses go: {array, generation of its element 0, length} (a one-thread kernel
        # queued by the C++ side ahead of the run's first graph)
        ring = 0
        if loss_host is not None:
            assert loss_host.dtype == torch.float32 and not loss_host.is_cuda and loss_host.numel() >= n_steps
            ring = loss_host.data_ptr() + 4 * k0
        gen0 = st.optimizer.step_count + 1
        st.optimizer.step_count += n
        self._loss_keepalive = loss_host
        native.pump_run(self._handle, order, n, k0 * self.batch, self._sink_ptr, ring, gen0)

    def pregather(self, order: torch.Tensor, n_steps: int) -> None:
        """Host-side prefetch across epochs: put the first batches of the NEXT run (sample order ``order``) into their
        pinned ring slots now, while the GPU works through the steps that are still queued.  The next ``run`` checks the
        indices and gathers as usual if they differ."""
        if self._handle is not None and n_steps > 0 and self.stepper._warm >= 2:
            native.pump_pregather(self._handle, order, n_steps)

    def wait_first(self) -> None:
        if self._handle is not None:
            native.pump_wait_first(self._handle)

    def drain(self) -> None:
        if self._handle is not None:
            native.pump_drain(self._handle)
            torch.cuda.current_stream(self.stepper.device).synchronize()

    def close(self) -> None:
        if self._handle is not None:
            native.pump_destroy(self._handle)
            self._handle = None
            self.stepper.loss_sink[:3].zero_()      # element 3 (the error mirror's address) stays
        self._graphs.clear()
