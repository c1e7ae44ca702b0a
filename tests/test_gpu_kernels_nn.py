"""Numerics of the fused memory-bound kernels vs plain PyTorch fp32 references (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bn_ref(x, gamma, beta, res, relu, eps=1e-5):
    xf = x.float()
    mean = xf.mean(0)
    var = xf.var(0, unbiased=False)
    xh = (xf - mean) * torch.rsqrt(var + eps)
    z = xh * gamma + beta
    if res is not None:
        z = z + res.float()
    return (torch.relu(z) if relu else z), mean, var


@pytest.mark.parametrize("M,C,relu,use_res", [
    (4096, 64, True, False), (3000, 256, True, True), (777, 2048, False, False),
    (12544, 512, True, True), (1000, 200, True, False), (50000, 64, True, True),
])
def test_bn_fwd_bwd(cuda, M, C, relu, use_res):
    from kubedl_b200.ops import nn as K
    torch.manual_seed(0)
    x = (torch.randn(M, C, device=cuda) * 1.5 + 0.3).bfloat16()
    res = torch.randn(M, C, device=cuda).bfloat16() if use_res else None
    gamma = torch.rand(C, device=cuda) + 0.5
    beta = torch.randn(C, device=cuda) * 0.1
    rm, rv = torch.zeros(C, device=cuda), torch.ones(C, device=cuda)
    y, save = K.bn_fwd(x, gamma, beta, rm, rv, res=res, relu=relu)
    yr, mean, var = _bn_ref(x, gamma, beta, res, relu)
    torch.testing.assert_close(y.float(), yr, atol=3e-2, rtol=2e-2)
    torch.testing.assert_close(save[0], mean, atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(save[1], torch.rsqrt(var + 1e-5), atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(rm, 0.1 * mean, atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(rv, 0.9 + 0.1 * var * M / (M - 1), atol=1e-3, rtol=1e-3)
    # second call must give the same answer (workspace counters self-reset)
    y2, _ = K.bn_fwd(x, gamma, beta, None, None, res=res, relu=relu)
    assert torch.equal(y, y2)

    # backward against autograd on the fp32 reference
    dy = torch.randn(M, C, device=cuda).bfloat16()
    xf = x.float().requires_grad_(True)
    g = gamma.clone().requires_grad_(True)
    b = beta.clone().requires_grad_(True)
    rf = res.float().requires_grad_(True) if use_res else None
    yy, _, _ = _bn_ref(xf, g, b, rf, relu)
    # mask from OUR bf16 output so both sides agree on which units are active
    mask = (y.float() > 0).float() if relu else torch.ones_like(yy)
    z = (xf - xf.mean(0)) * torch.rsqrt(xf.var(0, unbiased=False) + 1e-5) * g + b
    if use_res:
        z = z + rf
    (z * mask * dy.float()).sum().backward()
    dgamma = torch.empty(C, device=cuda, dtype=torch.bfloat16)
    dbeta = torch.empty(C, device=cuda, dtype=torch.bfloat16)
    dres = torch.empty_like(dy) if use_res else None
    dx = K.bn_bwd(dy, x, y, gamma, save, relu, dgamma, dbeta, dres=dres)
    scale = xf.grad.abs().max().item()
    torch.testing.assert_close(dx.float(), xf.grad, atol=2e-2 * max(scale, 1.0), rtol=3e-2)
    torch.testing.assert_close(dgamma.float(), g.grad, atol=0.02 * g.grad.abs().max().item() + 0.05, rtol=2e-2)
    torch.testing.assert_close(dbeta.float(), b.grad, atol=0.02 * b.grad.abs().max().item() + 0.05, rtol=2e-2)
    if use_res:
        torch.testing.assert_close(dres.float(), rf.grad, atol=1e-2, rtol=1e-2)
        # aliasing dres with dy is allowed
        dy2 = dy.clone()
        dx2 = K.bn_bwd(dy2, x, y, gamma, save, relu, dgamma, dbeta, dres=dy2)
        assert torch.equal(dx2, dx) and torch.equal(dy2, dres)
        if relu:
            # the forward pass can leave a 1-bit-per-element mask; backward from it == backward from y
            mask = torch.empty(M * C // 8, dtype=torch.uint8, device=cuda)
            y3, _ = K.bn_fwd(x, gamma, beta, None, None, res=res, relu=relu, relu_mask=mask)
            assert torch.equal(y3, y)
            dg4, db4 = torch.empty_like(dgamma), torch.empty_like(dbeta)
            dres4 = torch.empty_like(dy)
            dx4 = K.bn_bwd(dy, x, None, gamma, save, relu, dg4, db4, dres=dres4, relu_mask=mask)
            assert torch.equal(dx4, dx) and torch.equal(dres4, dres)
            assert torch.equal(dg4, dgamma) and torch.equal(db4, dbeta)
    elif relu:
        # without a residual the ReLU mask can be recomputed from x: same result, y never read
        dg3, db3 = torch.empty_like(dgamma), torch.empty_like(dbeta)
        dx3 = K.bn_bwd(dy, x, None, gamma, save, relu, dg3, db3, beta=beta)
        frac = float((dx3 != dx).float().mean())
        assert frac < 1e-3, frac          # only units whose pre-activation rounds to +-0 may differ
        torch.testing.assert_close(dg3.float(), dgamma.float(), atol=0.05, rtol=2e-2)
        torch.testing.assert_close(db3.float(), dbeta.float(), atol=0.05, rtol=2e-2)


def test_softmax_xent(cuda):
    from kubedl_b200.ops import nn as K
    torch.manual_seed(1)
    B, V = 256, 1000
    logits = (torch.randn(B, V, device=cuda) * 3).bfloat16()
    labels = torch.randint(0, V, (B,), device=cuda)
    loss, dl = K.softmax_xent(logits, labels)
    lf = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lf, labels)
    ref.backward()
    torch.testing.assert_close(loss[0], ref.detach(), atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(dl.float(), lf.grad, atol=2e-5, rtol=2e-2)
    # fp32 logits, padded leading dimension
    lp = torch.zeros(B, 1008, device=cuda)
    lp[:, :V] = logits.float()
    loss2, dl2 = K.softmax_xent(lp, labels, V=V)
    torch.testing.assert_close(loss2[0], ref.detach(), atol=1e-3, rtol=1e-3)
    assert dl2[:, V:].abs().max().item() == 0


def test_sgd_momentum(cuda):
    from kubedl_b200.ops import nn as K
    torch.manual_seed(2)
    n, wdc = 100003, 60000
    p32 = torch.randn(n, device=cuda)
    mom = torch.randn(n, device=cuda) * 0.1
    g = torch.randn(n, device=cuda).bfloat16()
    p16 = torch.empty(n, device=cuda, dtype=torch.bfloat16)
    pr, mr = p32.clone(), mom.clone()
    K.sgd_momentum(p32, mom, g, p16, lr=0.1, momentum=0.9, weight_decay=1e-2, wd_count=wdc)
    wd = torch.zeros(n, device=cuda)
    wd[:wdc] = 1e-2
    gg = g.float() + wd * pr
    mr = 0.9 * mr + gg
    pr = pr - 0.1 * mr
    torch.testing.assert_close(mom, mr, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(p32, pr, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(p16.float(), pr, atol=1e-2, rtol=1e-2)


def test_pools_and_add(cuda):
    from kubedl_b200.ops import nn as K
    torch.manual_seed(3)
    x = torch.randn(4, 30, 28, 64, device=cuda).bfloat16()
    y, idx = K.maxpool3x3s2_fwd(x)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = torch.nn.functional.max_pool2d(xr, 3, 2, 1)
    torch.testing.assert_close(y.float(), yr.permute(0, 2, 3, 1), atol=0, rtol=0)
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = K.maxpool3x3s2_bwd(dy, idx, x.shape)
    torch.testing.assert_close(dx.float(), xr.grad.permute(0, 2, 3, 1), atol=2e-2, rtol=2e-2)

    # odd sizes take the generic backward kernel, even sizes the 2x2-block one: both against autograd
    x2 = torch.randn(2, 13, 15, 64, device=cuda).bfloat16()
    y2, idx2 = K.maxpool3x3s2_fwd(x2)
    xr2 = x2.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr2 = torch.nn.functional.max_pool2d(xr2, 3, 2, 1)
    torch.testing.assert_close(y2.float(), yr2.permute(0, 2, 3, 1), atol=0, rtol=0)
    dy2 = torch.randn_like(y2)
    yr2.backward(dy2.float().permute(0, 3, 1, 2))
    dx2 = K.maxpool3x3s2_bwd(dy2, idx2, x2.shape)
    torch.testing.assert_close(dx2.float(), xr2.grad.permute(0, 2, 3, 1), atol=2e-2, rtol=2e-2)

    h = torch.randn(8, 7, 7, 2048, device=cuda).bfloat16()
    p = K.avgpool_fwd(h)
    torch.testing.assert_close(p.float(), h.float().mean((1, 2)), atol=1e-2, rtol=1e-2)
    dp = torch.randn(8, 2048, device=cuda).bfloat16()
    dh = K.avgpool_bwd(dp, h.shape)
    torch.testing.assert_close(dh.float(), (dp.float() / 49)[:, None, None, :].expand(8, 7, 7, 2048),
                               atol=1e-3, rtol=1e-2)
    a, b = torch.randn(4096, device=cuda).bfloat16(), torch.randn(4096, device=cuda).bfloat16()
    ref = (a.float() + b.float()).bfloat16()
    K.add_(a, b)
    assert torch.equal(a, ref)


def test_u8_normalize(cuda):
    from kubedl_b200.ops import nn as K
    img = torch.randint(0, 256, (2, 16, 16, 3), dtype=torch.uint8, device=cuda)
    out = K.u8_normalize(img, cpad=4)
    mean = torch.tensor([0.485, 0.456, 0.406], device=cuda)
    std = torch.tensor([0.229, 0.224, 0.225], device=cuda)
    ref = (img.float() / 255 - mean) / std
    torch.testing.assert_close(out[..., :3].float(), ref, atol=2e-2, rtol=1e-2)
    assert out[..., 3].abs().max().item() == 0


@pytest.mark.parametrize("W,n,dt_in,dt_out", [
    (2, 8 * 1000, torch.bfloat16, torch.bfloat16), (4, 8 * 100000, torch.bfloat16, torch.bfloat16),
    (8, 8 * 70001, torch.float32, torch.bfloat16), (8, 8 * 4096, torch.float32, torch.float32),
])
def test_allreduce_protocol_emulated(cuda, W, n, dt_in, dt_out):
    """All W ranks of the P2P all-reduce inside one cooperative launch on one GPU: exercises the
    flag protocol, tile ownership, fused scale+cast, and replay (epochs advance on device)."""
    from kubedl_b200.parallel.collectives import emulated_allreduce
    torch.manual_seed(4)
    src = [torch.randn(n, device=cuda).to(dt_in) for _ in range(W)]
    ref = torch.stack([s.float() for s in src]).sum(0) / W
    bufs = [s.clone() for s in src]
    run = emulated_allreduce(bufs, out_dtype=dt_out)
    outs = run()
    torch.cuda.synchronize()
    tol = 1e-5 if dt_out == torch.float32 else 2e-2
    for o in outs:
        torch.testing.assert_close(o.float(), ref, atol=tol, rtol=tol)
        assert torch.equal(o, outs[0])  # bitwise identical on every rank
    # replay twice more on fresh data: barrier epochs must keep working
    for it in range(2):
        for b, s in zip(bufs, src):
            b.copy_(s * (it + 2))
        outs = run()
        torch.cuda.synchronize()
        for o in outs:
            torch.testing.assert_close(o.float(), ref * (it + 2), atol=tol * (it + 2), rtol=tol)
    assert int(run.state["status"][0].item()) == 0


@pytest.mark.parametrize("W", [2, 4, 8])
def test_symm_barrier_protocol_emulated(cuda, W):
    """`kdl_symm_barrier` with all W ranks inside one cooperative launch: returns (no rank is left waiting),
    advances every rank's device-side epoch by one per call, and replays from a CUDA graph."""
    from kubedl_b200.parallel.collectives import FLAG_CHANNELS, emulated_barrier
    from kubedl_b200.ops import lib as L
    run = emulated_barrier(W, cuda)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    # epoch words: [rank][channel][block] with KDL_AR_MAX_BLOCKS blocks per channel (csrc/allreduce.cu epoch_slot)
    fb = L.lib().kdl_allreduce_flags_bytes(FLAG_CHANNELS) // 4
    max_blocks = fb // FLAG_CHANNELS // 8                # flags are [channel][KDL_MAX_WORLD = 8][block]
    ep_all = run.state["epochs"]
    ep = ep_all[: W * FLAG_CHANNELS * max_blocks].view(W, FLAG_CHANNELS, max_blocks)
    assert ep[:, FLAG_CHANNELS - 1, 0].tolist() == [3] * W
    assert int(ep_all.sum().item()) == 3 * W             # nothing else was touched
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    for _ in range(4):
        g.replay()
    torch.cuda.synchronize()
    assert ep[:, FLAG_CHANNELS - 1, 0].tolist() == [7] * W      # 3 eager + 4 replays (the capture itself runs nothing)
    assert int(run.state["status"][0].item()) == 0


def test_bn_relu_maxpool_fused_matches_separate_passes(cuda):
    """BN apply + ReLU + 3x3/s2 max-pool in one pass (statistics from the producer's epilogue) == bn_fwd then
    maxpool; argmax indices may differ only on exact ties."""
    from kubedl_b200.ops import nn as K
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(4)
    N, H, C, Kd = 3, 16, 64, 64
    a = torch.randn(N * H * H, Kd, device=cuda).bfloat16()
    w = torch.randn(C, Kd, device=cuda).bfloat16()
    x2, stats = g.nt_stats(a, w)
    x = x2.view(N, H, H, C)
    gamma = torch.rand(C, device=cuda) + 0.5
    beta = torch.randn(C, device=cuda) * 0.1
    save1 = torch.empty(2, C, device=cuda)
    y, _ = K.bn_fwd(x, gamma, beta, None, None, relu=True, save=save1, stats=stats)
    p_ref, _ = K.maxpool3x3s2_fwd(y)
    save2 = torch.empty(2, C, device=cuda)
    p, idx = K.bn_relu_maxpool_fwd(x, gamma, beta, None, None, save2, stats)
    assert torch.equal(save1, save2)
    assert torch.equal(p, p_ref)
    # the saved argmax really points at the maximum
    dy = torch.ones_like(p)
    dx = K.maxpool3x3s2_bwd(dy, idx, x.shape)
    assert float(dx.float().sum()) == float(p.numel())


def test_bn_bwd_pooled_matches_separate_passes(cuda):
    """max-pool backward gathered inside the BatchNorm backward == maxpool_bwd followed by bn_bwd."""
    from kubedl_b200.ops import nn as K
    torch.manual_seed(5)
    N, H, C = 3, 16, 64
    x = torch.randn(N, H, H, C, device=cuda).bfloat16()
    gamma = torch.rand(C, device=cuda) + 0.5
    beta = torch.randn(C, device=cuda) * 0.1
    y, save = K.bn_fwd(x, gamma, beta, None, None, relu=True)
    p, idx = K.maxpool3x3s2_fwd(y)
    dp = torch.randn_like(p)
    da = K.maxpool3x3s2_bwd(dp, idx, x.shape)
    dg1, db1 = torch.empty(C, device=cuda, dtype=torch.bfloat16), torch.empty(C, device=cuda, dtype=torch.bfloat16)
    dx1 = K.bn_bwd(da, x, None, gamma, save, True, dg1, db1, beta=beta)
    dg2, db2 = torch.empty_like(dg1), torch.empty_like(db1)
    dx2 = K.bn_bwd_pooled(dp, idx, x, gamma, beta, save, dg2, db2)
    assert torch.equal(dx1, dx2) and torch.equal(dg1, dg2) and torch.equal(db1, db2)
