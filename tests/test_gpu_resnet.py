"""End-to-end numerics of the explicit ResNet-50 program vs torchvision in fp32 (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _load_into_torchvision(ours, tv):
    sd = tv.state_dict()
    a = ours.arena
    for sp in a.layout:
        v = a.view(a.p32, sp.name)
        if sp.name.endswith(".weight") or sp.name.endswith(".bias"):
            key = sp.name
        else:
            key = sp.name + ".weight"
        if v.dim() == 4:  # ours [Cout,kh,kw,Cin] -> torch [Cout,Cin,kh,kw]
            v = v.permute(0, 3, 1, 2)
        assert sd[key].shape == v.shape, (key, sd[key].shape, v.shape)
        sd[key].copy_(v)
    tv.load_state_dict(sd)


def _cos(a, b):
    a, b = a.flatten().float(), b.flatten().float()
    return float((a @ b) / (a.norm() * b.norm() + 1e-12))


@pytest.mark.parametrize("use_gemm", [False, True])
def test_resnet50_step_matches_torchvision(cuda, use_gemm):
    import torchvision
    from kubedl_b200.models.resnet import ResNet50
    from kubedl_b200.ops import nn as K
    gemm = None
    if use_gemm:
        G = pytest.importorskip("kubedl_b200.ops.gemm")
        gemm = G.Gemm(cuda)
    torch.manual_seed(0)
    net = ResNet50(cuda, gemm=gemm, seed=3)
    tv = torchvision.models.resnet50().to(cuda).float().train()
    _load_into_torchvision(net, tv)
    B, S = 16, 96
    img = torch.randint(0, 256, (B, S, S, 3), dtype=torch.uint8, device=cuda)
    lab = torch.randint(0, 1000, (B,), device=cuda)

    x = K.u8_normalize(img)
    # tcgen05 path: the stem takes the raw uint8 image (fused normalise + im2col + GEMM)
    logits = net.forward(img if use_gemm else x, train=True)
    loss, dlogits = K.softmax_xent(logits, lab, V=1000)
    net.backward(dlogits)
    torch.cuda.synchronize()

    xr = x.float().permute(0, 3, 1, 2).contiguous()
    lr = tv(xr)
    loss_r = torch.nn.functional.cross_entropy(lr, lab)
    loss_r.backward()

    # second reference: the same torchvision model under bf16 autocast shows how much of the
    # distance to fp32 is plain bf16 rounding through 53 batch-normalised layers (random-init
    # network, tiny batch statistics: the noise compounds); we must be about as close as it is.
    ref_grads = {n: p.grad.clone() for n, p in tv.named_parameters()}
    tv.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        l_amp = tv(xr)
        loss_amp = torch.nn.functional.cross_entropy(l_amp.float(), lab)
    loss_amp.backward()
    amp_grads = {n: p.grad.clone() for n, p in tv.named_parameters()}
    amp_cos = _cos(l_amp, lr)
    assert torch.isfinite(logits.float()).all()
    assert _cos(logits, lr) > min(0.99, amp_cos - 0.03), (_cos(logits, lr), amp_cos)
    assert abs(float(loss[0]) - float(loss_r)) < 0.05 * max(1.0, float(loss_r))
    g = net.arena
    names = ["fc.weight", "layer4.2.conv3", "layer3.0.downsample.0", "layer2.1.conv2",
             "layer1.0.conv1", "layer3.2.bn2.weight", "conv1"]
    for name in names:
        key = name if name.endswith((".weight", ".bias")) else name + ".weight"
        ours = g.grad(name)
        if ours.dim() == 4:
            ours = ours.permute(0, 3, 1, 2)
        c_ours = _cos(ours, ref_grads[key])
        c_amp = _cos(amp_grads[key], ref_grads[key])
        assert c_ours > min(0.97, c_amp - 0.12), f"{name}: ours {c_ours:.3f} vs autocast {c_amp:.3f}"
    # one optimizer step moves the bf16 compute copy and keeps it equal to the rounded master
    before = g.p16.clone()
    net.sgd_step(0.1)
    assert not torch.equal(before, g.p16)
    assert torch.equal(g.p16, g.p32.bfloat16())


def test_trainer_graph_replay_matches_eager(cuda):
    """CUDA-graph replay must produce the same losses as eager launches of the same program."""
    from kubedl_b200.worker.ddp_trainer import DDPTrainer
    losses = {}
    for use_graph in (False, True):
        torch.manual_seed(0)
        tr = DDPTrainer(batch_per_gpu=8, use_graph=use_graph, image_size=64, seed=5, use_gemm=False)
        g = torch.Generator().manual_seed(7)
        img = torch.randint(0, 256, (8, 64, 64, 3), dtype=torch.uint8, generator=g).pin_memory()
        lab = torch.randint(0, 1000, (8,), dtype=torch.int64, generator=g).pin_memory()
        tr.images[0].copy_(img); tr.images[1].copy_(img)
        tr.labels[0].copy_(lab); tr.labels[1].copy_(lab)
        tr.warmup(3)
        out = [tr.train_step(img, lab, sync=True) for _ in range(4)]
        losses[use_graph] = out
        assert all(l == l for l in out)
    # same number of optimizer steps on the same data -> same trajectory (cuDNN algo choice may
    # differ between instances, so allow bf16-level noise)
    for a, b in zip(losses[False], losses[True]):
        assert abs(a - b) < 0.15 * max(1.0, abs(a)), (losses[False], losses[True])


@pytest.mark.parametrize("block,shape", [((0, 0), (8, 28, 28, 64)), ((1, 1), (8, 14, 14, 512)),
                                         ((2, 0), (8, 14, 14, 512))])
@pytest.mark.parametrize("use_gemm", [False, True, "own_conv"])
def test_bottleneck_block_fwd_bwd(cuda, block, shape, use_gemm, monkeypatch):
    """One bottleneck in isolation (errors do not compound): outputs, input gradient and every
    parameter gradient against the fp32 torchvision block with identical weights."""
    import torchvision
    from kubedl_b200.models.resnet import ResNet50
    gemm = None
    if use_gemm:
        G = pytest.importorskip("kubedl_b200.ops.gemm")
        gemm = G.Gemm(cuda)
    # 3x3 convolutions on our kernels (N7) or on cuDNN
    monkeypatch.setenv("KUBEDL_OWN_CONV", "1" if use_gemm == "own_conv" else "0")
    net = ResNet50(cuda, gemm=gemm, seed=11)
    assert net.own_conv == (use_gemm == "own_conv")
    tv = torchvision.models.resnet50().to(cuda).float().train()
    _load_into_torchvision(net, tv)
    li, bi = block
    idx = sum(n for _, n, _ in ResNet50.LAYERS[:li]) + bi
    ours = net.blocks[idx]
    ref = getattr(tv, f"layer{li + 1}")[bi]
    torch.manual_seed(5)
    x = torch.randn(*shape, device=cuda).bfloat16()
    y = ours.fwd(x, train=True)
    xr = x.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = ref(xr)
    assert _cos(y, yr.permute(0, 2, 3, 1)) > 0.999
    dy = (torch.randn_like(y.float()) * (y > 0)).bfloat16().contiguous()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = ours.bwd(dy.clone())
    assert _cos(dx, xr.grad.permute(0, 2, 3, 1)) > 0.995, _cos(dx, xr.grad.permute(0, 2, 3, 1))
    for n, p in ref.named_parameters():
        name = f"layer{li + 1}.{bi}.{n}"
        aname = name[:-len(".weight")] if (n.endswith(".weight") and p.dim() == 4) else name
        go = net.arena.grad(aname)
        if go.dim() == 4:
            go = go.permute(0, 3, 1, 2)
        c = _cos(go, p.grad)
        assert c > 0.99, f"{name}: cosine {c}"


def test_stem_im2col_gemm_matches_conv2d(cuda):
    """uint8 image -> fused normalise+im2col -> tcgen05 GEMM  ==  conv2d(normalised image, 7x7/s2/p3);
    and the wgrad GEMM over the same patches == conv2d weight gradient."""
    from kubedl_b200.ops import nn as K
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(0)
    N, S = 4, 64
    img = torch.randint(0, 256, (N, S, S, 3), dtype=torch.uint8, device=cuda)
    w = (torch.randn(64, 7, 7, 3, device=cuda) * 0.05).bfloat16()
    patches = K.stem_im2col(img)
    assert patches.shape == (N * 32 * 32, K.STEM_KP)
    assert float(patches.view(-1, 7, K.STEM_KW)[:, :, 21:].abs().max()) == 0.0
    wp = K.stem_weight_padded(w)
    y = g.nt(patches, wp).view(N, 32, 32, 64)
    xn = K.u8_normalize(img).float().permute(0, 3, 1, 2)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xn, wr, stride=2, padding=3)
    torch.testing.assert_close(y.float(), yr.permute(0, 2, 3, 1), atol=5e-2, rtol=3e-2)
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    gw = K.stem_weight_unpad(g.tn(dy.view(-1, 64), patches)).view(64, 7, 7, 3)
    assert _cos(gw, wr.grad.permute(0, 2, 3, 1)) > 0.999


def test_stem_direct_conv_matches_patch_gemm(cuda):
    """Stem convolution straight from the uint8 image (no patch matrix): forward, BN statistics and weight
    gradient against the im2col + GEMM path and conv2d."""
    from kubedl_b200.ops import nn as K
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(0)
    N, S = 3, 224
    img = torch.randint(0, 256, (N, S, S, 3), dtype=torch.uint8, device=cuda)
    w = (torch.randn(64, 7, 7, 3, device=cuda) * 0.05).bfloat16()
    wp = K.stem_weight_padded(w)
    y, (partial, nparts) = K.stem_conv_fwd(img, wp, want_stats=True)
    xn = K.u8_normalize(img).float().permute(0, 3, 1, 2)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xn, wr, stride=2, padding=3)
    torch.testing.assert_close(y.float(), yr.permute(0, 2, 3, 1), atol=5e-2, rtol=3e-2)
    y2 = g.nt(K.stem_im2col(img), wp).view(N, 112, 112, 64)
    assert float((y.float() - y2.float()).abs().max()) < 2e-2
    # statistics of the stored bf16 values
    s = partial[:nparts].sum(0)
    yf = y.float().view(-1, 64)
    torch.testing.assert_close(s[0], yf.sum(0), atol=0.5, rtol=1e-3)
    torch.testing.assert_close(s[1], (yf * yf).sum(0), atol=0.5, rtol=1e-3)
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    gw = K.stem_weight_unpad(K.stem_conv_wgrad(img, dy)).view(64, 7, 7, 3)
    assert _cos(gw, wr.grad.permute(0, 2, 3, 1)) > 0.999
    err = float((gw.float() - wr.grad.permute(0, 2, 3, 1)).abs().max() / wr.grad.abs().max())
    assert err < 2e-2, err


@pytest.mark.parametrize("N", [3, 150])
def test_stem_space_to_depth_conv_matches_conv2d(cuda, N):
    """The stem as a 4x4 convolution on the 2x2 space-to-depth image, operand tiles read by TMA through a tensor
    map with overlapping windows (csrc/stem_s2d.cu): the packed image, forward, BN statistics and weight gradient
    against conv2d; N = 150 gives every CTA one or two tiles, including an image boundary inside a CTA's range."""
    from kubedl_b200.ops import nn as K
    torch.manual_seed(0)
    S = 224
    img = torch.randint(0, 256, (N, S, S, 3), dtype=torch.uint8, device=cuda)
    w = (torch.randn(64, 7, 7, 3, device=cuda) * 0.05).bfloat16()
    wp = K.stem_weight_padded(w)
    s2d = K.stem_s2d_pack(img)
    xn = K.u8_normalize(img)                                             # [N,224,224,3] bf16
    # reference packing: S[n, R+2, C+2, (dy*2+dx)*3+c] = xn[n, 2R+dy, 2C+dx, c], zero elsewhere
    ref = torch.zeros(N, 115, 115, 16, dtype=torch.bfloat16, device=cuda)
    ref[:, 2:114, 2:114, :12] = xn.view(N, 112, 2, 112, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(N, 112, 112, 12)
    assert torch.equal(s2d, ref)
    y, (partial, nparts) = K.stem_s2d_fwd(s2d, wp, want_stats=True)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xn.float().permute(0, 3, 1, 2), wr, stride=2, padding=3)
    torch.testing.assert_close(y.float(), yr.permute(0, 2, 3, 1), atol=5e-2, rtol=3e-2)
    s = partial[:nparts].sum(0)
    yf = y.float().view(-1, 64)
    torch.testing.assert_close(s[0], yf.sum(0), atol=0.5 * N, rtol=1e-3)
    torch.testing.assert_close(s[1], (yf * yf).sum(0), atol=0.5 * N, rtol=1e-3)
    dy = torch.randn_like(y)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    gwp = K.stem_s2d_wgrad(s2d, dy)
    assert float(gwp.view(64, 7, 24)[:, :, 21:].abs().max()) == 0.0     # padding columns are written as zeros
    gw = K.stem_weight_unpad(gwp).view(64, 7, 7, 3)
    assert _cos(gw, wr.grad.permute(0, 2, 3, 1)) > 0.999
    err = float((gw.float() - wr.grad.permute(0, 2, 3, 1)).abs().max() / wr.grad.abs().max())
    assert err < 2e-2, err
    assert torch.equal(gwp, K.stem_s2d_wgrad(s2d, dy))                   # fixed summation order
