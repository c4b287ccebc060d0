"""Symmetric staging memory for the fused P2P all-reduce + update kernel.

Replaces the NCCL communicator the reference obtains from
``dist.init_process_group('nccl')`` (train.py:26-31).  torch.distributed is the
bootstrap channel only; after set-up ``csrc/reduce_update.cu`` reads and writes
peer memory directly over NVLink.

Two transports provide the peer mappings:

``ipc``   our own plumbing: ``cudaMalloc`` + CUDA IPC handles (64 opaque bytes
          each, all-gathered once).  No multicast, so algorithms 0 (one-shot)
          and 1 (two-shot).
``symm``  ``torch.distributed._symmetric_memory`` (cuMem VMM + multicast
          object): additionally yields the NVSwitch multicast alias needed by
          algorithm 2 (NVLS ``multimem.ld_reduce`` / ``multimem.st``).

Block layout per rank (floats / words), identical on every rank:
    stage : [2 parities][arena_len] published slices, then [arena_len] landing
            area for reduced slices (two-shot / NVLS)
    flags : [8 source ranks][max_slices] "slice published", then [max_slices]
            "reduced slice delivered"; generation numbers that only ever grow
    ll    : (arenas up to ``LL_MAX_ARENA_BYTES``) receive area of the push
            protocol, [2 parities][8 source ranks][arena_len x 2 words] of
            flagged 8-byte lines (csrc/ll_reduce.cu)

Algorithms (``algo``): ``push`` — every rank writes its gradient straight into
every peer's receive area as soon as a bucket exists and the step ends with ONE
wait+sum+update launch (latency regime; whole-model gradients up to ~1 MB);
``oneshot`` / ``twoshot`` / ``nvls`` — the per-bucket flag-handshake kernels of
csrc/reduce_update.cu (bandwidth regime).  ``auto`` = push when the receive
area exists, else per bucket by size.
"""
from __future__ import annotations

import warnings
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native

FLAG_GRAIN = 32   # csrc/reduce_update.cu RU_FLAG_GRAIN
MAX_WORLD = 8
ALGO_CODES = {"oneshot": 0, "twoshot": 1, "nvls": 2}
WIRE_CODES = {"fp32": 0, "bf16": 1}
# The push protocol moves 2 x (world-1) x arena bytes out of every rank per step (fp32 wire: every 4 payload bytes
# travel with a 4-byte flag) and keeps a receive area of 128 bytes per arena float on every rank; beyond ~1 MB the
# handshake kernels, which move (world-1)/world .. (world-1) x arena bytes, are the better trade.
LL_MAX_ARENA_BYTES = 1 << 20


def ll_max_bytes(world: int) -> int:
    """Largest arena the push protocol is chosen for by ``algo="auto"``.  Measured cross-over against the one-shot
    handshake kernel as ISOLATED launches (profiles/allreduce_sweep.md): 2 ranks 10-12 us up to 1 MB (one-shot 12-23);
    8 ranks 11-13 us up to 64 KB, level at 256 KB (21.5 vs 20.5), behind at 1 MB (63 vs 35: 14 MB leave every rank).
    Inside a training step the early buckets' pushes overlap the backward pass, which favours the push protocol further."""
    return max(256 << 10, (2 << 20) // max(world, 1))
# Measured cross-over (profiles/allreduce_sweep.md).  One-shot pulls (world-1) x bytes into every rank but needs a
# single handshake; two-shot / NVLS move 2(world-1)/world x bytes with two.  With 2 ranks the byte counts are
# equal, so one-shot always wins; at 8 ranks it wins up to ~2 MB (1 MB: 35 vs 39 us, 4 MB: 70 vs 45 us NVLS).
def oneshot_max_bytes(world: int) -> int:
    return (1 << 62) if world <= 2 else (16 << 20) // world


class P2PComm:
    def __init__(self, arena_len: int, rank: int, world: int, algo: str = "auto", transport: str = "auto",
                 wire: str = "fp32", timeout_ms: int = 0):
        if world > MAX_WORLD:
            raise ValueError(f"world size {world} > {MAX_WORLD} (single NVSwitch node)")
        if algo not in ("auto", "push", "oneshot", "twoshot", "nvls"):
            raise ValueError(f"unknown all-reduce algorithm {algo!r}")
        if wire not in WIRE_CODES:
            raise ValueError(f"unknown gradient wire format {wire!r}")
        self.rank, self.world, self.arena_len, self.algo = rank, world, arena_len, algo
        self.wire, self.timeout_ms = wire, int(timeout_ms)
        self.max_slices = (arena_len + FLAG_GRAIN - 1) // FLAG_GRAIN
        self.stage_bytes = 3 * arena_len * 4
        self.flag_off = (self.stage_bytes + 255) // 256 * 256
        flags_end = self.flag_off + (MAX_WORLD + 1) * self.max_slices * 4
        # receive area of the push protocol (only for arenas in the latency regime, or when asked for)
        self.ll_enabled = algo == "push" or (algo == "auto" and arena_len * 4 <= ll_max_bytes(world))
        self.ll_off = (flags_end + 255) // 256 * 256
        self.ll_bytes = 2 * MAX_WORLD * arena_len * 2 * 4 if self.ll_enabled else 0
        self.total_bytes = self.ll_off + self.ll_bytes
        self.local_ptr: Optional[int] = None
        self.peer_ptrs: List[Optional[int]] = [None] * world
        self.multicast_ptr = 0
        self.transport = "none"
        self._symm_tensor = None
        self._symm_handle = None
        self.struct = native.Comm()
        self.struct.world, self.struct.rank, self.struct.algo = world, rank, 0
        self.struct.arena_len, self.struct.max_slices = arena_len, self.max_slices
        self.struct.mc_stage = None
        self._by_algo = {}
        if world > 1:
            # CUDA IPC is the default; torch symmetric memory is only needed (and only tried) for the
            # multicast alias that NVLS requires
            big = not self.ll_enabled and arena_len * 4 > oneshot_max_bytes(world)   # some bucket may want the multicast (NVLS) path
            want_symm = transport == "symm" or (transport == "auto" and (algo == "nvls" or (algo == "auto" and big)))
            if not (want_symm and self._setup_symm()):
                self._setup_ipc()
            for r in range(world):
                self.struct.stage[r] = self.peer_ptrs[r]
                self.struct.flags[r] = self.peer_ptrs[r] + self.flag_off
                self.struct.ll[r] = self.peer_ptrs[r] + self.ll_off if self.ll_enabled else None
            self.struct.wire, self.struct.timeout_ms = WIRE_CODES[wire], self.timeout_ms
            self.struct.mc_stage = self.multicast_ptr or None
            if algo == "nvls" and not self.multicast_ptr:
                warnings.warn("NVLS multicast is not available on this system; using the two-shot P2P algorithm")
            dist.barrier()   # nobody launches before every mapping exists

    # -- transports ---------------------------------------------------------------
    def _setup_ipc(self) -> None:
        self.local_ptr = native.symm_alloc(self.total_bytes)
        handle = native.symm_export(self.local_ptr)
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, handle)
        for r in range(self.world):
            self.peer_ptrs[r] = self.local_ptr if r == self.rank else native.symm_import(handles[r])
        self.transport = "ipc"

    def _setup_symm(self) -> bool:
        ok = 1
        try:
            import torch.distributed._symmetric_memory as symm_mem
            t = symm_mem.empty(self.total_bytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
            t.zero_()
            torch.cuda.synchronize()
            hdl = symm_mem.rendezvous(t, dist.group.WORLD)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        except Exception as e:  # no VMM / fabric support in this container: fall back to CUDA IPC
            warnings.warn(f"torch symmetric memory unavailable ({type(e).__name__}: {e}); using CUDA IPC")
            ok = 0
        flag = torch.tensor([ok], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)      # all ranks must agree on the transport
        if int(flag.item()) == 0:
            return False
        self._symm_tensor, self._symm_handle = t, hdl
        self.peer_ptrs = ptrs
        self.local_ptr = ptrs[self.rank]
        self.multicast_ptr = mc
        self.transport = "symm"
        return True

    # -- per-bucket algorithm choice ---------------------------------------------------
    OVERLAP_CTAS = 64      # CTAs of a reduce+update that runs beside backward kernels.  Measured on the wide model at 2 GPUs
                           # (3 MB conv3 bucket; step us / exposed us): 16 CTAs 319 / 153, 32: 237 / 71, 64: 213 / 48,
                           # 128: 212 / 46, whole GPU: slower again (spinning CTAs hold the SMs backward needs).  The
                           # buckets are reduced one after the other, so a slow middle bucket delays the last one.

    def struct_for(self, bucket_bytes: int, overlapped: bool = False) -> "native.Comm":
        """The ``b200dp_comm`` to pass for a bucket of this size (algorithm chosen per bucket).  ``overlapped``:
        the launch shares the GPU with the rest of backward, so it is kept to a few CTAs (``max_ctas``)."""
        algo = self.algo
        if algo in ("auto", "push"):
            algo = "oneshot" if bucket_bytes <= oneshot_max_bytes(self.world) else ("nvls" if self.multicast_ptr else "twoshot")
        if algo == "nvls" and not self.multicast_ptr:
            algo = "twoshot"
        key = (algo, bool(overlapped))
        st = self._by_algo.get(key)
        if st is None:
            st = native.Comm()
            st.max_ctas = self.OVERLAP_CTAS if overlapped else 0
            st.world, st.rank, st.algo = self.world, self.rank, ALGO_CODES[algo]
            st.arena_len, st.max_slices = self.arena_len, self.max_slices
            for r in range(self.world):
                st.stage[r] = self.struct.stage[r]
                st.flags[r] = self.struct.flags[r]
                st.ll[r] = self.struct.ll[r]
            st.wire, st.timeout_ms = self.struct.wire, self.struct.timeout_ms
            st.mc_stage = self.multicast_ptr or None
            self._by_algo[key] = st
        return st

    @property
    def push(self) -> bool:
        """True when the step should run the push protocol (``b200dp_grad_push`` per bucket + one
        ``b200dp_reduce_update_ll``) instead of one handshake launch per bucket."""
        return self.ll_enabled and self.algo in ("auto", "push")

    def reset(self) -> None:
        """Forget every generation number seen so far (collective).  The protocols rely on generation numbers that
        only grow; a caller that moves the step counter BACK (loading a checkpoint) must clear the flags and the
        receive area first, or stale lines of the abandoned future would be taken for fresh ones."""
        if self.local_ptr is None:
            return
        torch.cuda.synchronize()
        dist.barrier()
        native.symm_zero(self.local_ptr, self.total_bytes)
        torch.cuda.synchronize()
        dist.barrier()

    def close(self) -> None:
        if self.local_ptr is None:
            return
        torch.cuda.synchronize()
        if dist.is_initialized():
            try:
                dist.barrier()   # peers must stop touching our block before it goes away
            except Exception:
                pass
        if self.transport == "ipc":
            for r, p in enumerate(self.peer_ptrs):
                if p is not None and r != self.rank:
                    native.symm_close(p)
            native.symm_free(self.local_ptr)
        self._symm_tensor = self._symm_handle = None
        self.local_ptr = None
        self.peer_ptrs = [None] * self.world
