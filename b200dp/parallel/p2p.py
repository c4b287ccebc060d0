"""Symmetric staging memory for the fused P2P all-reduce + update kernel.

Replaces the NCCL communicator the reference obtains from
``dist.init_process_group('nccl')`` (train.py:26-31).  torch.distributed is the
bootstrap channel only: each rank allocates one block (staging buffers for two
step parities + flag slots), exports a 64-byte CUDA IPC handle, the handles are
all-gathered once, every rank maps every peer's block.  From then on
``csrc/reduce_update.cu`` reads and writes peer memory directly over NVLink.

Block layout per rank (bytes):
    [0, 2*arena_len*4)                      staging, parity 0 then parity 1 (fp32)
    [flag_off, flag_off + 8*max_slices*4)   flags[src_rank][slice] (uint32 generation)
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import native

FLAG_GRAIN = 32   # csrc/reduce_update.cu RU_FLAG_GRAIN
MAX_WORLD = 8


class P2PComm:
    def __init__(self, arena_len: int, rank: int, world: int, algo: str = "auto"):
        if world > MAX_WORLD:
            raise ValueError(f"world size {world} > {MAX_WORLD} (single NVSwitch node)")
        self.rank, self.world, self.arena_len = rank, world, arena_len
        self.max_slices = (arena_len + FLAG_GRAIN - 1) // FLAG_GRAIN
        self.stage_bytes = 2 * arena_len * 4
        self.flag_off = (self.stage_bytes + 255) // 256 * 256
        self.total_bytes = self.flag_off + MAX_WORLD * self.max_slices * 4
        self.algo = {"auto": 0, "oneshot": 0, "twoshot": 1, "nvls": 2}[algo]
        self.local_ptr: Optional[int] = None
        self.peer_ptrs: List[Optional[int]] = [None] * world
        self.struct = native.Comm()
        self.struct.world, self.struct.rank, self.struct.algo = world, rank, 0 if algo == "auto" else self.algo
        self.struct.arena_len, self.struct.max_slices = arena_len, self.max_slices
        self.struct.mc_stage = None
        if world > 1:
            self._setup()

    def _setup(self) -> None:
        self.local_ptr = native.symm_alloc(self.total_bytes)
        handle = native.symm_export(self.local_ptr)
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, handle)
        for r in range(self.world):
            self.peer_ptrs[r] = self.local_ptr if r == self.rank else native.symm_import(handles[r])
        for r in range(self.world):
            self.struct.stage[r] = self.peer_ptrs[r]
            self.struct.flags[r] = self.peer_ptrs[r] + self.flag_off
        dist.barrier()   # nobody launches before every mapping exists

    def close(self) -> None:
        if self.local_ptr is None:
            return
        torch.cuda.synchronize()
        if dist.is_initialized():
            try:
                dist.barrier()   # peers must stop touching our block before it goes away
            except Exception:
                pass
        for r, p in enumerate(self.peer_ptrs):
            if p is not None and r != self.rank:
                native.symm_close(p)
        native.symm_free(self.local_ptr)
        self.local_ptr = None
        self.peer_ptrs = [None] * self.world
