#!/usr/bin/env python
"""BASELINE.json configs[4]: gradient all-reduce + SGD update, 4 KB - 256 MB, fused P2P kernel vs the
reference's NCCL path (dist.all_reduce + '/= size' + optimizer.step), device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py

Roofline: bytes that must cross NVLink per rank / peer bandwidth (770 GB/s per direction: the peer-copy rate measured
on this pool's B200s, /opt/skills/guides/B200_PROFILING.md; nominal 900): one-shot pulls (N-1) x bytes in; two-shot /
NVLS pull (N-1)/N x bytes in and push the same out; the push protocol (<= 1 MB) sends 2 (N-1) x bytes out (every 4
payload bytes travel with a 4-byte flag).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200dp.ops import native                      # noqa: E402
from b200dp.optim.flat import OPT_SGD              # noqa: E402
from b200dp.parallel.p2p import P2PComm            # noqa: E402

LINK_GBS = 770.0


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3   # us


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    max_bytes = int(os.environ.get("SWEEP_MAX_BYTES", 256 << 20))
    n_max = max_bytes // 4
    native.load()
    comm = P2PComm(n_max, rank, world, algo=os.environ.get("SWEEP_TRANSPORT_ALGO", "nvls"))
    push_max = min(max_bytes, 1 << 20)                      # the push protocol's regime
    comm_ll = P2PComm(push_max // 4, rank, world, algo="push")
    param = torch.zeros(n_max, device=dev)
    grad = torch.randn(n_max, device=dev) * 1e-3
    s0, s1 = torch.zeros(n_max, device=dev), torch.zeros(1, device=dev)
    hyper = torch.tensor([0.01, 0.9, 0, 0, 0, 0, 0, 0], dtype=torch.float32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    rows = []
    size = 4096
    while size <= max_bytes:
        n = size // 4
        n = (n + 127) // 128 * 128
        iters = 200 if size <= (1 << 20) else (50 if size <= (32 << 20) else 15)
        rec = {"bytes": n * 4, "world": world}
        # ---- reference path: NCCL all-reduce, divide, optimizer step (train.py:34-39, 51) ----
        p_ref = torch.nn.Parameter(torch.zeros(n, device=dev))
        p_ref.grad = grad[:n].clone()
        opt = torch.optim.SGD([p_ref], lr=0.01, momentum=0.9)

        def nccl_step():
            dist.all_reduce(p_ref.grad.data, op=dist.ReduceOp.SUM)
            p_ref.grad.data /= float(world)
            opt.step()
        rec["nccl_us"] = timed(nccl_step, iters)
        for algo in ("oneshot", "twoshot", "nvls"):
            if algo == "nvls" and not comm.multicast_ptr:
                continue
            if algo == "oneshot" and size > (64 << 20):
                continue
            comm.algo = algo
            st = comm.struct_for(n * 4)

            def fused_step():
                native.reduce_update(param, grad, s0, s1, None, 0, n, None, 0, 0, OPT_SGD, hyper, counter, True, st)
            rec[f"{algo}_us"] = timed(fused_step, iters)
            wire = (world - 1) * n * 4 if algo == "oneshot" else (world - 1) / world * n * 4
            rec[f"{algo}_roofline_frac"] = (wire / (LINK_GBS * 1e9)) / (rec[f"{algo}_us"] * 1e-6)
        if size <= push_max:
            st_ll = comm_ll.struct_for(0)

            def push_step():   # one launch: push + wait + rank-order sum + average + momentum update
                native.reduce_update_ll(param, grad, s0, s1, None, 0, n, None, 0, 0, 0, 0, True, OPT_SGD, hyper, counter, st_ll, last=True)
            rec["push_us"] = timed(push_step, iters)
            rec["push_roofline_frac"] = (2 * (world - 1) * n * 4 / (LINK_GBS * 1e9)) / (rec["push_us"] * 1e-6)
        if int(counter[2]) != 0:
            raise RuntimeError("peer timeout flag set")
        rows.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        del p_ref, opt
        size *= 4
    if rank == 0:
        out = os.environ.get("SWEEP_OUT", "gpurun_out/allreduce_sweep_n%d.json" % world)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        json.dump({"transport": comm.transport, "multicast": bool(comm.multicast_ptr), "rows": rows}, open(out, "w"), indent=1)
    comm_ll.close()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
