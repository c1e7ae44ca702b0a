import sys, torch, torchvision
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from test_gpu_resnet import _load_into_torchvision, _cos
from kubedl_b200.models.resnet import ResNet50
from kubedl_b200.ops import nn as K
cuda = torch.device("cuda", 0)
torch.manual_seed(0)
net = ResNet50(cuda, gemm=None, seed=3)
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
lr = tv(xr); loss_r = torch.nn.functional.cross_entropy(lr, lab); loss_r.backward()
print("loss", float(loss[0]), float(loss_r), "cos logits", _cos(logits, lr))
tvp = dict(tv.named_parameters())
for sp in net.arena.layout:
    g = net.arena.grad(sp.name).float()
    key = sp.name if sp.name.endswith((".weight", ".bias")) else sp.name + ".weight"
    r = tvp[key].grad
    if g.dim() == 4: g = g.permute(0, 3, 1, 2)
    print(f"{sp.name:32s} cos={_cos(g, r):+.4f} |ours|={g.norm().item():.3e} |ref|={r.norm().item():.3e} max={g.abs().max().item():.2e} nan={torch.isnan(g).any().item()}")
