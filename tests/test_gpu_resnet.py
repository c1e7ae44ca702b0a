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
    logits = net.forward(x, train=True)
    loss, dlogits = K.softmax_xent(logits, lab, V=1000)
    net.backward(dlogits)
    torch.cuda.synchronize()

    xr = x.float().permute(0, 3, 1, 2).contiguous()
    lr = tv(xr)
    loss_r = torch.nn.functional.cross_entropy(lr, lab)
    loss_r.backward()

    assert torch.isfinite(logits.float()).all()
    assert _cos(logits, lr) > 0.995, _cos(logits, lr)
    assert abs(float(loss[0]) - float(loss_r)) < 0.05 * max(1.0, float(loss_r))
    g = net.arena
    checks = {
        "fc.weight": (g.grad("fc.weight"), tv.fc.weight.grad),
        "layer4.2.conv3": (g.grad("layer4.2.conv3").permute(0, 3, 1, 2), tv.layer4[2].conv3.weight.grad),
        "layer3.0.downsample.0": (g.grad("layer3.0.downsample.0").permute(0, 3, 1, 2),
                                  tv.layer3[0].downsample[0].weight.grad),
        "layer2.1.conv2": (g.grad("layer2.1.conv2").permute(0, 3, 1, 2), tv.layer2[1].conv2.weight.grad),
        "layer1.0.conv1": (g.grad("layer1.0.conv1").permute(0, 3, 1, 2), tv.layer1[0].conv1.weight.grad),
        "layer3.2.bn2.weight": (g.grad("layer3.2.bn2.weight"), tv.layer3[2].bn2.weight.grad),
        "conv1": (g.grad("conv1").permute(0, 3, 1, 2), tv.conv1.weight.grad),
    }
    for name, (a, b) in checks.items():
        c = _cos(a, b)
        assert c > 0.97, f"{name}: cosine {c}"
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
