"""Kernel tier (1xB200): every native kernel vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

from b200dp.models.lenet import CIFAR_WIDE_LENET, MNIST_LENET, LeNet, LeNetConfig, loss_fn

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact_fp32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield


def _native():
    from b200dp.ops import native
    native.load()
    return native


def _dims(native, cfg: LeNetConfig, b: int):
    return native.Dims(b, cfg.in_ch, cfg.in_hw, cfg.c1, cfg.c2, cfg.c3, cfg.f1, cfg.num_classes)


def _pool_ref(z):
    """leaky + 2x2 maxpool with argmax as 0..3 inside the window (dy*2+dx)."""
    a = F.leaky_relu(z, 0.2)
    p, flat = F.max_pool2d(a, 2, 2, return_indices=True)
    W = a.shape[-1]
    yy, xx = flat // W, flat % W
    return p, ((yy % 2) * 2 + (xx % 2)).to(torch.uint8)


@pytest.mark.parametrize("cfg,u8", [(MNIST_LENET, True), (MNIST_LENET, False), (CIFAR_WIDE_LENET, True)])
@pytest.mark.parametrize("b", [1, 37, 256])
def test_conv1_fwd(cfg, u8, b):
    native = _native()
    torch.manual_seed(0)
    dev = "cuda"
    m = LeNet(cfg).to(dev)
    md = LeNet(cfg).to(dev).double()          # fp64 oracle (cuDNN fp32 may silently use TF32)
    md.load_state_dict(m.state_dict())
    mean, std = 0.1307, 0.3081
    if u8:
        x = torch.randint(0, 256, (b, cfg.in_ch, cfg.in_hw, cfg.in_hw), dtype=torch.uint8, device=dev)
        xf = (x.double().div(255) - mean) / std
    else:
        x = torch.randn(b, cfg.in_ch, cfg.in_hw, cfg.in_hw, device=dev)
        xf = x.double()
    with torch.no_grad():
        z = md.ConvNet[1](md.ConvNet[0](xf))
    p_ref, i_ref = _pool_ref(z)
    p_ref = p_ref.float()
    p1 = torch.empty(b, cfg.c1, 14, 14, device=dev)
    idx = torch.empty(b, cfg.c1, 14, 14, dtype=torch.uint8, device=dev)
    native.conv1_fwd(x, m.ConvNet[1].weight.detach().contiguous(), m.ConvNet[1].bias.detach(), p1, idx,
                     _dims(native, cfg, b), mean, std)
    torch.cuda.synchronize()
    assert torch.allclose(p1, p_ref, atol=2e-5, rtol=1e-5), (p1 - p_ref).abs().max()
    # argmax may legitimately differ only where two window entries tie to round-off
    assert (idx != i_ref).float().mean() < 1e-3


@pytest.mark.parametrize("cfg", [MNIST_LENET, CIFAR_WIDE_LENET])
@pytest.mark.parametrize("b", [1, 37, 256])
def test_conv2_fwd(cfg, b):
    native = _native()
    torch.manual_seed(1)
    dev = "cuda"
    m = LeNet(cfg).to(dev)
    md = LeNet(cfg).to(dev).double()          # fp64 oracle
    md.load_state_dict(m.state_dict())
    p1 = torch.randn(b, cfg.c1, 14, 14, device=dev)
    with torch.no_grad():
        z = md.ConvNet[4](p1.double())
    p_ref, i_ref = _pool_ref(z)
    p_ref = p_ref.float()
    p2 = torch.empty(b, cfg.c2, 5, 5, device=dev)
    idx = torch.empty(b, cfg.c2, 5, 5, dtype=torch.uint8, device=dev)
    native.conv2_fwd(p1, m.ConvNet[4].weight.detach().contiguous(), m.ConvNet[4].bias.detach(), p2, idx,
                     _dims(native, cfg, b))
    torch.cuda.synchronize()
    assert torch.allclose(p2, p_ref, atol=1e-4, rtol=1e-4), (p2 - p_ref).abs().max()
    assert (idx != i_ref).float().mean() < 1e-3


@pytest.mark.parametrize("M,N,K", [(256, 120, 400), (256, 84, 120), (37, 480, 1600), (5, 10, 84), (256, 336, 480)])
@pytest.mark.parametrize("act", [0, 1])
def test_linear_fwd(M, N, K, act):
    native = _native()
    torch.manual_seed(2)
    x, w, b = (torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda") / K ** 0.5,
               torch.randn(N, device="cuda"))
    y = torch.empty(M, N, device="cuda")
    native.linear_fwd(x, w, b, y, act)
    ref = F.linear(x.double(), w.double(), b.double())
    ref = F.leaky_relu(ref, 0.2) if act else ref
    assert torch.allclose(y.double(), ref, atol=1e-4, rtol=1e-4), (y.double() - ref).abs().max()


@pytest.mark.parametrize("mode", ["ref", "xent"])
@pytest.mark.parametrize("B,F1", [(256, 84), (37, 84), (64, 336)])
def test_head_tail(mode, B, F1):
    native = _native()
    torch.manual_seed(3)
    dev = "cuda"
    z1 = torch.randn(B, F1, device=dev, dtype=torch.float64, requires_grad=True)   # fc1 pre-activation
    w = (torch.randn(10, F1, device=dev, dtype=torch.float64) / F1 ** 0.5)
    bias = torch.randn(10, device=dev, dtype=torch.float64)
    t = torch.randint(0, 10, (B,), device=dev)
    h1 = F.leaky_relu(z1, 0.2)
    logits = F.linear(h1, w, bias)
    logits.retain_grad()
    probs = logits.softmax(-1)
    loss = loss_fn(probs if mode == "ref" else logits, t, mode)
    loss.backward()

    f = lambda a: a.detach().float().contiguous()  # noqa: E731
    o_probs, o_rows, o_loss = torch.empty(B, 10, device=dev), torch.empty(B, device=dev), torch.zeros(1, device=dev)
    o_dl, o_dz1 = torch.empty(B, 10, device=dev), torch.empty(B, F1, device=dev)
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)
    for _ in range(2):   # twice: the ticket must re-arm itself
        native.head_tail(f(h1), f(w), f(bias), t, 0 if mode == "ref" else 1, o_probs, o_rows, o_loss, o_dl, o_dz1,
                         None, ticket)
    torch.cuda.synchronize()
    assert torch.allclose(o_probs.double(), probs, atol=1e-6)
    assert abs(float(o_loss) - float(loss.detach())) < 1e-6
    assert torch.allclose(o_dl.double(), logits.grad, atol=1e-8, rtol=1e-4)
    assert torch.allclose(o_dz1.double(), z1.grad, atol=1e-8, rtol=1e-4)
    # eval counters
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    native.head_tail(f(h1), f(w), f(bias), t, 0 if mode == "ref" else 1, o_probs, o_rows, o_loss, None, None,
                     acc, ticket)
    ref_sum = loss_fn(probs if mode == "ref" else logits, t, mode, reduction="sum")
    assert abs(float(acc[0]) - float(ref_sum)) < 1e-4
    assert int(acc[1]) == int(probs.argmax(1).eq(t).sum())


@pytest.fixture(params=[1, 0], ids=["tma", "cp_async"])
def head_weight_path(request):
    """Both weight streams of the fused head: TMA tensor maps (default) and the cp.async ring."""
    lib = _native().load()
    lib.b200dp_head_fused_set_tma(request.param)
    yield request.param
    lib.b200dp_head_fused_set_tma(1)


@pytest.mark.parametrize("mode", ["ref", "xent"])
@pytest.mark.parametrize("B,K3,C3,F1,pitched", [(256, 400, 120, 84, False), (37, 400, 120, 84, True), (64, 152, 128, 128, True),
                                                  (9, 33, 17, 5, False)])
def test_head_fused(mode, B, K3, C3, F1, pitched, head_weight_path):
    """conv3 + FC.0 + FC.2 + softmax + loss + dlogits + dz1 + dz3 in one launch against fp64 autograd (cnn.py:16-23,
    train.py:48-49), dense and pitched buffers, ragged row blocks, eval counters."""
    native = _native()
    torch.manual_seed(5)
    dev = "cuda"
    d64 = dict(device=dev, dtype=torch.float64)
    p2 = torch.randn(B, K3, **d64)
    w3, b3 = torch.randn(C3, K3, **d64) / K3 ** 0.5, torch.randn(C3, **d64) * 0.1
    w1, b1 = torch.randn(F1, C3, **d64) / C3 ** 0.5, torch.randn(F1, **d64) * 0.1
    w2, b2 = torch.randn(10, F1, **d64) / F1 ** 0.5, torch.randn(10, **d64) * 0.1
    t = torch.randint(0, 10, (B,), device=dev)
    z3 = F.linear(p2, w3, b3).requires_grad_(True)
    z3.retain_grad()
    a3 = F.leaky_relu(z3, 0.2)
    z1 = F.linear(a3, w1, b1)
    z1.retain_grad()
    h1 = F.leaky_relu(z1, 0.2)
    logits = F.linear(h1, w2, b2)
    logits.retain_grad()
    probs = logits.softmax(-1)
    loss = loss_fn(probs if mode == "ref" else logits, t, mode)
    loss.backward()

    f = lambda a: a.detach().float().contiguous()  # noqa: E731
    pad = 4 if pitched else 0
    ld_p2, ld_a3, ld_h1, ld_dl = K3 + pad, C3 + pad, F1 + pad, (12 if pitched else 10)
    poison = float("nan")
    p2f = torch.full((B, ld_p2), poison, device=dev)
    p2f[:, :K3] = f(p2)
    o_a3, o_h1 = torch.full((B, ld_a3), -5.0, device=dev), torch.full((B, ld_h1), -5.0, device=dev)
    o_dl = torch.full((B, ld_dl), -7.0, device=dev)
    o_probs, o_rows, o_loss = torch.empty(B, 10, device=dev), torch.empty(B, device=dev), torch.zeros(1, device=dev)
    o_dz1, o_dz3 = torch.empty(B, F1, device=dev), torch.empty(B, C3, device=dev)
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)
    lm = 0 if mode == "ref" else 1
    for _ in range(2):   # twice: the ticket must re-arm itself
        native.head_fused(p2f, ld_p2 if pitched else 0, f(w3), f(b3), f(w1), f(b1), f(w2), f(b2), t, lm, o_a3,
                          ld_a3 if pitched else 0, o_h1, ld_h1 if pitched else 0, o_probs, o_rows, o_loss, o_dl,
                          ld_dl if pitched else 0, o_dz1, o_dz3, None, ticket, B)
    torch.cuda.synchronize()
    assert torch.allclose(o_a3[:, :C3].double(), a3, atol=2e-5, rtol=1e-5)
    assert torch.allclose(o_h1[:, :F1].double(), h1, atol=2e-5, rtol=1e-5)
    if pitched:   # padding columns are never written
        assert (o_a3[:, C3:] == -5.0).all() and (o_h1[:, F1:] == -5.0).all() and (o_dl[:, 10:] == -7.0).all()
    assert torch.allclose(o_probs.double(), probs, atol=1e-5)
    assert abs(float(o_loss) - float(loss.detach())) < 1e-5
    assert torch.allclose(o_dl[:, :10].double(), logits.grad, atol=1e-7, rtol=1e-3)
    assert torch.allclose(o_dz1.double(), z1.grad, atol=1e-7, rtol=1e-3)
    assert torch.allclose(o_dz3.double(), z3.grad, atol=1e-7, rtol=1e-3)
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    native.head_fused(p2f, ld_p2 if pitched else 0, f(w3), f(b3), f(w1), f(b1), f(w2), f(b2), t, lm, o_a3,
                      ld_a3 if pitched else 0, o_h1, ld_h1 if pitched else 0, o_probs, o_rows, o_loss, None, 0, None, None,
                      acc, ticket, B)
    ref_sum = loss_fn(probs if mode == "ref" else logits, t, mode, reduction="sum")
    assert abs(float(acc[0]) - float(ref_sum.detach())) < 1e-3
    assert int(acc[1]) == int(probs.argmax(1).eq(t).sum())


def test_linear_bwd_group():
    native = _native()
    torch.manual_seed(4)
    dev = "cuda"
    B, N, K = 256, 120, 400
    x = torch.randn(B, K, device=dev)
    w = torch.randn(N, K, device=dev) / K ** 0.5
    dy = torch.randn(B, N, device=dev)
    act_prev = torch.randn(B, K, device=dev)
    dx, dw, db = torch.empty(B, K, device=dev), torch.empty(N, K, device=dev), torch.empty(N, device=dev)
    dx2 = torch.empty(B, K, device=dev)
    native.linear_bwd_group([
        native.bwd_problem(0, dy, w, act_prev, dx, None, B, N, K),
        native.bwd_problem(1, dy, x, None, dw, db, B, N, K),
        native.bwd_problem(0, dy, w, None, dx2, None, B, N, K),
    ])
    torch.cuda.synchronize()
    ref_dx2 = dy.double() @ w.double()
    ref_dx = ref_dx2 * torch.where(act_prev > 0, 1.0, 0.2).double()
    assert torch.allclose(dx.double(), ref_dx, atol=1e-4, rtol=1e-4)
    assert torch.allclose(dx2.double(), ref_dx2, atol=1e-4, rtol=1e-4)
    assert torch.allclose(dw.double(), dy.double().t() @ x.double(), atol=1e-3, rtol=1e-4)
    assert torch.allclose(db.double(), dy.double().sum(0), atol=1e-4, rtol=1e-4)


def _eager_grads(model, x_f32, y, mode):
    model.zero_grad(set_to_none=True)
    out = model(x_f32) if mode == "ref" else model.logits(x_f32)
    loss = loss_fn(out, y, mode)
    loss.backward()
    return loss.detach(), {n: p.grad.detach().clone() for n, p in model.named_parameters()}


@pytest.mark.parametrize("tc", [False, True])
@pytest.mark.parametrize("mode", ["ref", "xent"])
@pytest.mark.parametrize("b,u8", [(256, True), (37, False), (1, True)])
def test_forward_backward_matches_autograd(mode, b, u8, tc):
    """All ten gradients of one fwd+bwd through the native kernels vs torch autograd."""
    from b200dp.data.synthetic import MNIST_SPEC
    from b200dp.models.fused import NativeLeNet
    torch.manual_seed(5)
    dev = torch.device("cuda")
    ref = LeNet().to(dev).double()
    net = NativeLeNet(MNIST_LENET, 256, dev, loss_mode=mode, tc=tc)
    with torch.no_grad():
        for (n, p), (_, q) in zip(net.module.named_parameters(), ref.named_parameters()):
            q.copy_(p.double())
    if u8:
        x = torch.randint(0, 256, (b, 1, 28, 28), dtype=torch.uint8, device=dev)
        xf = (x.double().div(255) - MNIST_SPEC.mean) / MNIST_SPEC.std
    else:
        x = torch.randn(b, 1, 28, 28, device=dev)
        xf = x.double()
    y = torch.randint(0, 10, (b,), device=dev)
    loss_ref, g_ref = _eager_grads(ref, xf, y, mode)
    net.forward(x, y, train=True)
    net.backward(x)
    torch.cuda.synchronize()
    assert abs(float(net.loss) - float(loss_ref)) < (2e-3 if tc else 1e-5)
    got = {n: net.g[n].clone() for n in net.g}
    conv = net.partials[: net.cur_partials].double().sum(0)
    for name, off in zip(("ConvNet.1.weight", "ConvNet.1.bias", "ConvNet.4.weight", "ConvNet.4.bias"), net.conv_offsets):
        got[name] = conv[off: off + got[name].numel()].view_as(got[name]).float()
    for n, gr in g_ref.items():
        scale = gr.abs().max().clamp_min(1e-12)
        err = (got[n].double() - gr).abs().max() / scale
        assert err < (8e-2 if tc else 2e-4), (n, float(err))      # tf32 tensor-core GEMMs vs exact fp32


@pytest.mark.parametrize("kind", ["adadelta", "sgd", "adam"])
def test_reduce_update_matches_flat_optimizer(kind):
    native = _native()
    from b200dp.optim.flat import FlatOptimizer, OptHyper
    torch.manual_seed(6)
    dev = "cuda"
    n = 4096 + 128
    p0 = torch.randn(n, device=dev)
    hyper = OptHyper(kind, lr=0.05, weight_decay=0.01 if kind == "sgd" else 0.0)
    ref = FlatOptimizer(p0.clone(), torch.zeros(n, device=dev), hyper)
    p, g = p0.clone(), torch.zeros(n, device=dev)
    s0, s1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    h = ref.hyper
    vec = {"adadelta": [0.05, h.rho, h.eps, 0, h.weight_decay], "sgd": [0.05, h.momentum, 0, 0, h.weight_decay],
           "adam": [0.05, h.beta1, h.beta2, h.eps, h.weight_decay]}[kind]
    hv = torch.tensor(vec + [0, 0, 0], dtype=torch.float32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    for step in range(4):
        grad = torch.randn(n, device=dev) * 0.1
        ref.grad.copy_(grad)
        ref.step_torch()
        g.copy_(grad)
        native.reduce_update(p, g, s0, s1, None, 0, 4096, None, 0, 0, hyper.code(), hv, counter, False, None)
        native.reduce_update(p, g, s0, s1, None, 4096, 128, None, 0, 0, hyper.code(), hv, counter, True, None)
        torch.cuda.synchronize()
        assert int(counter[0]) == step + 1
        assert torch.allclose(p, ref.param, atol=1e-6, rtol=1e-5), (kind, step, (p - ref.param).abs().max())
        assert torch.allclose(s0, ref.state0, atol=1e-7, rtol=1e-5)


@pytest.mark.parametrize("kind", ["adadelta", "sgd", "adam"])
def test_reduce_update_ll_world1_matches_flat_optimizer(kind):
    """The push protocol's final kernel without peers: whole-arena update in one launch, a sub-range fed by partial
    rows (fixed-order sum), bf16 shadow weights, generation counter — against the torch flat optimizer."""
    native = _native()
    from b200dp.optim.flat import FlatOptimizer, OptHyper
    torch.manual_seed(6)
    dev = "cuda"
    n, p_start, p_count, rows = 4096 + 640, 1024, 512, 37
    p0 = torch.randn(n, device=dev)
    hyper = OptHyper(kind, lr=0.05, weight_decay=0.01 if kind == "sgd" else 0.0)
    ref = FlatOptimizer(p0.clone(), torch.zeros(n, device=dev), hyper)
    p, g = p0.clone(), torch.zeros(n, device=dev)
    s0, s1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    hv = torch.tensor(list(ref.hyper_vector()) + [0, 0, 0], dtype=torch.float32, device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    for step in range(4):
        grad = torch.randn(n, device=dev) * 0.1
        partials = torch.randn(rows, p_count, device=dev) * 0.02
        grad[p_start:p_start + p_count] = partials.double().sum(0).float()
        ref.grad.copy_(grad)
        ref.step_torch()
        g.copy_(grad)
        g[p_start:p_start + p_count] = float("nan")          # that range must come from the rows, not from the arena
        native.reduce_update_ll(p, g, s0, s1, shadow, 0, n, partials, rows, p_count, p_start, p_count, False,
                                hyper.code(), hv, counter, None)
        torch.cuda.synchronize()
        assert int(counter[0]) == step + 1 and int(counter[2]) == 0
        assert torch.allclose(p, ref.param, atol=2e-6, rtol=1e-5), (kind, step, (p - ref.param).abs().max())
        assert torch.allclose(s0, ref.state0, atol=1e-6, rtol=1e-5)
        assert torch.allclose(g, grad, atol=1e-6)            # the averaged gradient is written back (train.py:39)
        assert torch.equal(shadow, p.to(torch.bfloat16))     # round-to-nearest-even copy for tensor-core consumers


def test_post_loss_and_fetch_slot():
    """The two small kernels of the step pump: the loss into a mapped pinned array at slot (generation - base), and
    the upload of a pinned ring slot by a kernel that reads host memory through its mapped address."""
    native = _native()
    dev = "cuda"
    ring = torch.full((8,), -1.0).pin_memory()
    loss = torch.tensor([0.625], device=dev)
    counter = torch.zeros(4, dtype=torch.int32, device=dev)
    counter[0] = 41                                              # 41 steps done: the running step has generation 42
    sink = torch.tensor([ring.data_ptr(), 40, ring.numel(), 0], dtype=torch.int64, device=dev)
    native.post_loss(loss, counter, sink)
    sink_off = torch.tensor([ring.data_ptr(), 50, ring.numel(), 0], dtype=torch.int64, device=dev)
    native.post_loss(loss, counter, sink_off)                    # slot would be negative: nothing is written
    torch.cuda.synchronize()
    assert ring[2] == 0.625 and (ring[:2] == -1).all() and (ring[3:] == -1).all()
    src = torch.randint(0, 256, (200704 + 2048,), dtype=torch.uint8).pin_memory()
    dst = torch.zeros_like(src, device=dev)
    native.fetch_slot(dst, src)
    torch.cuda.synchronize()
    assert torch.equal(dst.cpu(), src)


def test_reduce_update_partials_deterministic():
    native = _native()
    torch.manual_seed(7)
    dev = "cuda"
    n, rows = 2688, 128
    partials = torch.randn(rows, n, device=dev)
    outs = []
    for _ in range(2):
        p, g = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        s0, s1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        hv = torch.tensor([1.0, 0.0, 0, 0, 0, 0, 0, 0], dtype=torch.float32, device=dev)   # plain SGD, lr 1
        counter = torch.zeros(4, dtype=torch.int32, device=dev)
        native.reduce_update(p, g, s0, s1, None, 0, n, partials, rows, n, 1, hv, counter, True, None)
        torch.cuda.synchronize()
        outs.append((p.clone(), g.clone()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.allclose(outs[0][1].double(), partials.double().sum(0), atol=1e-4)
    assert torch.allclose(outs[0][0], -outs[0][1])


@pytest.mark.parametrize("dtype", ["fp32", "tf32"])
@pytest.mark.parametrize("opt", ["adadelta", "sgd", "adam"])
@pytest.mark.parametrize("graph", [False, True])
def test_train_steps_match_eager(opt, graph, dtype):
    """N full steps (fwd, bwd, fused update; optionally CUDA-graph replay) vs the eager stepper."""
    from b200dp.engine import EagerStepper
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    torch.manual_seed(8)
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    lr = {"adadelta": 0.1, "sgd": 0.05, "adam": 1e-3}[opt]
    eager = EagerStepper(ctx, MNIST_LENET, OptHyper(opt, lr=lr), loss_mode="xent")
    native_s = NativeStepper(ctx, MNIST_LENET, OptHyper(opt, lr=lr), batch=64, loss_mode="xent", graph=graph,
                             dtype=dtype)
    native_s.model.load_state_dict(eager.model.state_dict())
    loose = 10.0 if dtype == "tf32" else 1.0
    xs = [torch.randint(0, 256, (64, 1, 28, 28), dtype=torch.uint8, device=dev) for _ in range(3)]
    ys = [torch.randint(0, 10, (64,), device=dev) for _ in range(3)]
    for i in range(9):
        le = float(eager.train_step(xs[i % 3], ys[i % 3]))
        ln = float(native_s.train_step(xs[i % 3], ys[i % 3]))
        assert abs(le - ln) < 5e-4 * loose * max(1.0, abs(le)), (i, le, ln)
    if dtype == "fp32" or opt == "sgd":   # adaptive optimizers amplify tf32 round-off in tiny gradients
        for (n, a), (_, b) in zip(eager.model.state_dict().items(), native_s.model.state_dict().items()):
            assert torch.allclose(a, b, atol=5e-4 * loose, rtol=5e-3 * loose), (n, (a - b).abs().max())
    # eval path
    eager.eval_batch(xs[0], ys[0]); native_s.eval_batch(xs[0], ys[0])
    (l0, c0), (l1, c1) = eager.eval_result(), native_s.eval_result()
    assert abs(l0 - l1) < 2e-4 * loose * max(1.0, abs(l0)) and abs(c0 - c1) <= 1


def test_ragged_batch_after_full_batch():
    """A short last batch must not pick up partial-gradient rows left by the previous full batch."""
    from b200dp.engine import EagerStepper
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    torch.manual_seed(9)
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    eager = EagerStepper(ctx, MNIST_LENET, OptHyper("sgd", lr=0.05), loss_mode="xent")
    nat = NativeStepper(ctx, MNIST_LENET, OptHyper("sgd", lr=0.05), batch=64, loss_mode="xent", graph=True,
                        dtype="fp32")
    nat.model.load_state_dict(eager.model.state_dict())
    for b in (64, 64, 64, 64, 23, 64, 5):
        x = torch.randint(0, 256, (b, 1, 28, 28), dtype=torch.uint8, device=dev)
        y = torch.randint(0, 10, (b,), device=dev)
        le, ln = float(eager.train_step(x, y)), float(nat.train_step(x, y))
        assert abs(le - ln) < 5e-4 * max(1.0, abs(le)), (b, le, ln)
    for (n, a), (_, c) in zip(eager.model.state_dict().items(), nat.model.state_dict().items()):
        assert torch.allclose(a, c, atol=5e-4, rtol=5e-3), (n, (a - c).abs().max())


@pytest.mark.parametrize("cfg_name", ["mnist_tc_conv", "wide"])
def test_tc_conv_training_matches_eager(cfg_name):
    """Full steps with the tcgen05 convolutions (and GEMMs): the wide CIFAR-shaped model, and the
    reference model with conv='tc'."""
    from b200dp.data.synthetic import CIFAR_SPEC, MNIST_SPEC
    from b200dp.engine import EagerStepper
    from b200dp.models.fused import NativeStepper
    from b200dp.optim.flat import OptHyper
    from b200dp.parallel.launch import RankContext
    torch.manual_seed(10)
    dev = torch.device("cuda", 0)
    ctx = RankContext(0, 1, 0, dev, "nccl")
    cfg, spec = (CIFAR_WIDE_LENET, CIFAR_SPEC) if cfg_name == "wide" else (MNIST_LENET, MNIST_SPEC)
    eager = EagerStepper(ctx, cfg, OptHyper("sgd", lr=0.02), loss_mode="xent", spec=spec)
    nat = NativeStepper(ctx, cfg, OptHyper("sgd", lr=0.02), batch=64, loss_mode="xent", spec=spec, graph=True,
                        dtype="tf32", conv="tc")
    nat.model.load_state_dict(eager.model.state_dict())
    shape = (64, cfg.in_ch, cfg.in_hw, cfg.in_hw)
    xs = [torch.randint(0, 256, shape, dtype=torch.uint8, device=dev) for _ in range(2)]
    ys = [torch.randint(0, 10, (64,), device=dev) for _ in range(2)]
    for i in range(6):
        le, ln = float(eager.train_step(xs[i % 2], ys[i % 2])), float(nat.train_step(xs[i % 2], ys[i % 2]))
        assert abs(le - ln) < 1e-2 * max(1.0, abs(le)), (i, le, ln)
    for (n, a), (_, b) in zip(eager.model.state_dict().items(), nat.model.state_dict().items()):
        assert torch.allclose(a, b, atol=5e-3, rtol=5e-2), (n, (a - b).abs().max())
