"""Where does the time go inside one tcgen05 GEMM CTA?  Prints clock64 deltas (cycles)."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.ops import native

lib = native.load()
lib.b200dp_tc_gemm_group_timed.argtypes = [C.POINTER(native.TcProblem), C.c_int, C.c_void_p, C.c_void_p]
dev = "cuda"
for (M, N, K, amn, bmn) in [(256, 120, 400, False, False), (256, 480, 1600, False, False), (256, 336, 480, False, False),
                            (480, 1601, 256, True, True), (256, 1600, 480, False, True)]:
    a = torch.randn((K, M + 4) if amn else (M, K + 4), device=dev)
    b = torch.randn((K, (N + 3) // 4 * 4) if bmn else (N, K + 4), device=dev)
    out = torch.empty(M, N, device=dev)
    bias = torch.randn(N, device=dev)
    dbg = torch.zeros(32, dtype=torch.int64, device=dev)
    prob = native.tc_problem(a, a.shape[1], amn, b, b.shape[1], bmn, M, N, K, bias=bias, act=1, out=out, ld_out=N)
    arr = (native.TcProblem * 1)(prob)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(3):
        e0.record()
        rc = lib.b200dp_tc_gemm_group_timed(arr, 1, dbg.data_ptr(), torch.cuda.current_stream().cuda_stream)
        e1.record()
        torch.cuda.synchronize()
    ws = native.GemmWorkspace(torch.device(dev))
    def split_call():
        native.tc_gemm_group([prob], ws)
    for _ in range(3):
        split_call()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        split_call()
    e1.record()
    torch.cuda.synchronize()
    t_split = e0.elapsed_time(e1) * 100
    e0.record()
    for _ in range(10):
        native.tc_gemm_group([prob])
    e1.record()
    torch.cuda.synchronize()
    t_plain = e0.elapsed_time(e1) * 100
    lib.b200dp_tc_gemm_set_debug.argtypes = [C.c_void_p]; lib.b200dp_tc_gemm_set_debug.restype = None
    dbg2 = torch.zeros(32, dtype=torch.int64, device=dev)
    lib.b200dp_tc_gemm_set_debug(dbg2.data_ptr()); split_call(); torch.cuda.synchronize(); lib.b200dp_tc_gemm_set_debug(None)
    u = dbg2.tolist()
    print(f"   split/retiled CTA 0: setup {u[1]-u[0]} acc_done {u[2]-u[0]} staged {u[5]-u[0]} epi_done {u[3]-u[0]} end {u[4]-u[0]} landed {[x-u[0] for x in u[16:24] if x]}")
    print(f"   back-to-back x10: plain {t_plain:.1f} us/launch, split-K {t_split:.1f} us/launch (ws floats {ws.ws.numel()})")
    t = dbg.tolist()
    t0 = t[0]
    print(f"M{M} N{N} K{K} amn={amn} bmn={bmn}: event {e0.elapsed_time(e1)*1e3:.1f} us | setup {t[1]-t0} acc_done {t[2]-t0} "
          f"staged {t[5]-t0} epi_done {t[3]-t0} end {t[4]-t0} | issued {[x-t0 for x in t[8:16] if x]} | landed {[x-t0 for x in t[16:24] if x]}")
