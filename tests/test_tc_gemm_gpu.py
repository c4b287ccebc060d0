"""tcgen05/TMEM/TMA GEMM (tf32) vs fp64 torch, all four operand-majorness combinations."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(M, N, K, a_mn, b_mn, *, bias=False, act=0, mask=False, colext=False, pad=0, splitk=False):
    from b200dp.ops import native
    native.load()
    dev = "cuda"
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K + 2 * a_mn + b_mn)
    A = torch.randn(M, K, generator=g).to(dev)           # logical A(m,k)
    Bm = torch.randn(N, K, generator=g).to(dev)          # logical B(n,k)
    def store(X, mn_major):
        rows, cols = (X.shape[1], X.shape[0]) if mn_major else X.shape
        ld = (cols + pad + 3) // 4 * 4
        buf = torch.full((rows, ld), 7.0, device=dev)    # poison in the padding: must never be read into the result
        buf[:, :cols] = X.t() if mn_major else X
        return buf, ld
    a_buf, lda = store(A, a_mn)
    b_buf, ldb = store(Bm, b_mn)
    bias_t = torch.randn(N, device=dev) if bias else None
    mask_t = torch.randn(M, N, device=dev) if mask else None
    n_out = N - 1 if colext else N
    out = torch.full((M, n_out + 3), -5.0, device=dev)
    col = torch.full((M,), -5.0, device=dev) if colext else None
    ws = native.GemmWorkspace(torch.device(dev)) if splitk else None
    for _ in range(2 if splitk else 1):                  # twice: the tickets must come back to zero
        native.tc_gemm_group([native.tc_problem(a_buf, lda, a_mn, b_buf, ldb, b_mn, M, N, K, bias=bias_t, act=act,
                                                mask=mask_t, ld_mask=N, out=out, ld_out=n_out + 3, colext=col)], ws)
    torch.cuda.synchronize()
    if splitk:
        assert int(ws.tickets.abs().sum()) == 0
    ref = A.double() @ Bm.double().t()
    full = ref.clone()
    if bias:
        full = full + bias_t.double()
    if act:
        full = torch.nn.functional.leaky_relu(full, 0.2)
    if mask:
        full = full * torch.where(mask_t > 0, 1.0, 0.2).double()
    tol = 6e-3 * (K ** 0.5)                              # tf32 operands are truncated to 10 mantissa bits
    assert (out[:, :n_out].double() - full[:, :n_out]).abs().max() < tol
    assert (out[:, n_out:] == -5.0).all()                # nothing written outside the logical tile
    if colext:
        assert (col.double() - ref[:, N - 1]).abs().max() < tol


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 120, 400), (120, 401, 256), (37, 84, 120), (10, 85, 256)])
def test_tc_gemm_majorness(a_mn, b_mn, M, N, K):
    _run(M, N, K, a_mn, b_mn)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(256, 480, 1600), (256, 120, 400), (120, 401, 256), (37, 84, 120), (480, 1601, 256)])
def test_tc_gemm_splitk(a_mn, b_mn, M, N, K):
    """Re-tiled + split-K launch (workspace given): same results, deterministic, tickets reset."""
    _run(M, N, K, a_mn, b_mn, splitk=True, bias=True, act=1)


def test_tc_gemm_splitk_deterministic():
    from b200dp.ops import native
    native.load()
    dev = torch.device("cuda")
    torch.manual_seed(3)
    A, Bm = torch.randn(256, 1600, device=dev), torch.randn(480, 1600, device=dev)
    ws = native.GemmWorkspace(dev)
    outs = []
    for _ in range(3):
        out = torch.empty(256, 480, device=dev)
        native.tc_gemm_group([native.tc_problem(A, 1600, False, Bm, 1600, False, 256, 480, 1600, out=out, ld_out=480)], ws)
        torch.cuda.synchronize()
        outs.append(out)
    assert ws.ws.numel() > 0                              # this shape does get split
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_tc_gemm_epilogues():
    _run(256, 480, 1600, False, False, bias=True, act=1, pad=4, splitk=True)
    _run(120, 401, 256, True, True, colext=True, splitk=True)
    _run(256, 120, 84, False, True, mask=True, splitk=True)
    _run(256, 120, 400, False, False, bias=True, act=1)
    _run(256, 120, 84, False, True, mask=True)
    _run(120, 401, 256, True, True, colext=True)
    _run(256, 480, 1600, False, False, bias=True, act=1, pad=4)


def test_tc_gemm_group_of_three():
    from b200dp.ops import native
    native.load()
    dev = "cuda"
    torch.manual_seed(0)
    B_, F1, C3 = 256, 84, 120
    dz1, a3, w = torch.randn(B_, F1, device=dev), torch.randn(B_, C3 + 4, device=dev), torch.randn(F1, C3, device=dev)
    a3[:, C3] = 1.0
    dw, db, dz3 = torch.empty(F1, C3, device=dev), torch.empty(F1, device=dev), torch.empty(B_, C3, device=dev)
    native.tc_gemm_group([
        native.tc_problem(dz1, F1, True, a3, C3 + 4, True, F1, C3 + 1, B_, out=dw, ld_out=C3, colext=db),
        native.tc_problem(dz1, F1, False, w, C3, True, B_, C3, F1, mask=a3, ld_mask=C3 + 4, out=dz3, ld_out=C3),
    ])
    torch.cuda.synchronize()
    tol = 6e-3 * 16
    assert (dw.double() - dz1.double().t() @ a3[:, :C3].double()).abs().max() < tol
    assert (db.double() - dz1.double().sum(0)).abs().max() < tol
    ref = (dz1.double() @ w.double()) * torch.where(a3[:, :C3] > 0, 1.0, 0.2).double()
    assert (dz3.double() - ref).abs().max() < tol
