"""Embedding dispatch/combine kernels (N4/N5) and the DIN model vs plain PyTorch fp32 autograd
(single GPU: world=1 exercises the same code path with all rows local)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_forward(table, mlp, user, cand, hist):
    D = 64
    u, c = table[user], table[cand]
    mask = hist >= 0
    h = table[hist.clamp(min=0)] * mask[..., None]
    s = (h * c[:, None, :]).sum(-1) / D ** 0.5
    s = s.masked_fill(~mask, float("-inf"))
    a = torch.softmax(s, 1)
    p = (a[..., None] * h).sum(1)
    x = torch.cat([u, c, p, p * c], 1)
    (w1, b1), (w2, b2), (w3, b3) = mlp
    h1 = torch.relu(x @ w1.t() + b1)
    h2 = torch.relu(h1 @ w2.t() + b2)
    return (h2 @ w3.t() + b3)[:, 0], x


def test_din_forward_backward_matches_autograd(cuda):
    from kubedl_b200.models.din import DINModel, synthetic_batch
    torch.manual_seed(0)
    V, L_, B = 50_000, 20, 512
    m = DINModel(vocab=V, hist_len=L_, lr=0.1, emb_lr=0.5, seed=1)
    user, cand, hist, label = synthetic_batch(B, L_, V, cuda, seed=3)
    table0 = m.shard.clone()
    a = m.arena
    mlp = [(a.w32(f"w{i}").clone().requires_grad_(True), a.w32(f"b{i}").clone().requires_grad_(True)) for i in (1, 2, 3)]
    table = table0.clone().requires_grad_(True)
    logit_ref, x_ref = _ref_forward(table, mlp, user, cand, hist)
    loss_ref = torch.nn.functional.binary_cross_entropy_with_logits(logit_ref, label)
    loss_ref.backward()

    logit = m.forward(user, cand, hist)
    x = m.saved[3]
    torch.testing.assert_close(x.float(), x_ref.detach(), atol=2e-3, rtol=2e-2)
    torch.testing.assert_close(logit, logit_ref.detach(), atol=3e-2, rtol=5e-2)
    loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, label)
    m.backward((torch.sigmoid(logit) - label) / B)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) < 2e-2
    # dense grads
    for i in (1, 2):
        g, r = a.grad(f"w{i}").float(), mlp[i - 1][0].grad
        cos = float((g.flatten() @ r.flatten()) / (g.norm() * r.norm() + 1e-12))
        assert cos > 0.99, (i, cos)
    g3 = a.grad("w3")[0].float()
    assert float((g3 @ mlp[2][0].grad[0]) / (g3.norm() * mlp[2][0].grad[0].norm() + 1e-12)) > 0.99
    # sparse update == -lr * dense table gradient, including duplicate ids (atomics coalesce)
    delta = (m.shard - table0)
    want = -0.5 * table.grad
    touched = want.abs().sum(1) > 0
    assert int(touched.sum()) > 100
    assert float(delta[~touched].abs().max()) == 0.0          # untouched rows are untouched
    d, w = delta[touched].flatten(), want[touched].flatten()
    cos = float((d @ w) / (d.norm() * w.norm() + 1e-12))
    assert cos > 0.98, cos
    assert abs(float(d.norm() / w.norm()) - 1.0) < 0.1


def test_din_training_reduces_loss(cuda):
    from kubedl_b200.models.din import DINModel, synthetic_batch
    V, L_, B = 20_000, 16, 1024
    m = DINModel(vocab=V, hist_len=L_, lr=0.05, emb_lr=20.0, seed=2)   # mean loss: row grads carry 1/B
    losses = []
    for it in range(80):
        user, cand, hist, label = synthetic_batch(B, L_, V, cuda, seed=it % 4)
        losses.append(float(m.step(user, cand, hist, label)))
    assert all(l == l for l in losses)
    assert sum(losses[-5:]) / 5 < sum(losses[:5]) / 5 - 0.02, (losses[:5], losses[-5:])


def test_din_fused_loss_and_bias_gradient_kernels(cuda):
    """kdl_din_bce_grad == BCE-with-logits + its gradient; kdl_relu_bwd_colsum == mask by (h > 0) + column sums."""
    from kubedl_b200.models.din import DINModel
    m = DINModel(vocab=4096, hist_len=8, seed=5)
    torch.manual_seed(7)
    B = 4096
    z = (torch.randn(B, 8, device=cuda) * 3).bfloat16()
    label = (torch.rand(B, device=cuda) < 0.4).float()
    loss, dz = m.loss_and_dz(z, label)
    logit = z[:, 0].float()
    ref_loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, label)
    ref_dz = ((torch.sigmoid(logit) - label) / B).bfloat16()
    assert abs(float(loss) - float(ref_loss)) < 1e-5 * max(1.0, float(ref_loss))
    torch.testing.assert_close(dz[:, 0].float(), ref_dz.float(), atol=1e-8, rtol=1e-2)
    assert float(dz[:, 1:].abs().max()) == 0.0
    for (M, N, relu) in [(4096, 512, True), (4096, 256, True), (4096, 8, False), (1000, 64, True), (77, 136, True)]:
        dh = torch.randn(M, N, device=cuda).bfloat16()
        h = torch.randn(M, N, device=cuda).bfloat16()
        dbias = torch.empty(N, dtype=torch.bfloat16, device=cuda)
        want = dh.float() * (h > 0) if relu else dh.float()
        got = m._relu_bwd_bias(dh.clone(), h if relu else None, dbias)
        assert torch.equal(got.float(), want.bfloat16().float())
        torch.testing.assert_close(dbias.float(), want.sum(0), atol=0.5, rtol=2e-2)
        dbias2 = torch.empty_like(dbias)
        m._relu_bwd_bias(dh.clone(), h if relu else None, dbias2)       # counters reset themselves; deterministic
        assert torch.equal(dbias, dbias2)


def test_din_step_matches_autograd_loss_and_bias_grads(cuda):
    """The fused training step (loss kernel -> backward_dz) produces the reference loss and bias gradients."""
    from kubedl_b200.models.din import DINModel, synthetic_batch
    V, L_, B = 50_000, 20, 512
    m = DINModel(vocab=V, hist_len=L_, lr=0.0, emb_lr=0.0, seed=1)
    user, cand, hist, label = synthetic_batch(B, L_, V, cuda, seed=3)
    a = m.arena
    mlp = [(a.w32(f"w{i}").clone().requires_grad_(True), a.w32(f"b{i}").clone().requires_grad_(True)) for i in (1, 2, 3)]
    table = m.shard.clone().requires_grad_(True)
    logit_ref, _ = _ref_forward(table, mlp, user, cand, hist)
    loss_ref = torch.nn.functional.binary_cross_entropy_with_logits(logit_ref, label)
    loss_ref.backward()
    import os
    os.environ["KUBEDL_DIN_GRAPH"] = "0"
    try:
        loss = m.step(user, cand, hist, label)
    finally:
        os.environ.pop("KUBEDL_DIN_GRAPH", None)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) < 2e-2
    for i in (1, 2):
        g, r = a.grad(f"b{i}").float(), mlp[i - 1][1].grad
        cos = float((g @ r) / (g.norm() * r.norm() + 1e-12))
        assert cos > 0.99, (i, cos)
    assert abs(float(a.grad("b3")[0]) - float(mlp[2][1].grad[0])) < 2e-2 * max(1.0, abs(float(mlp[2][1].grad[0])))
    assert float(a.grad("w3")[1:].float().abs().max()) == 0.0 and float(a.grad("b3")[1:].float().abs().max()) == 0.0


@pytest.mark.parametrize("dim,dtype", [(64, torch.bfloat16), (128, torch.float32), (36, torch.float32), (256, torch.bfloat16)])
def test_sharded_embedding_gather_and_scatter_add(cuda, dim, dtype):
    """Generic dispatch / combine (kdl_emb_gather / kdl_emb_scatter_add) vs index_select / index_add_."""
    from kubedl_b200.parallel.embedding import ShardedEmbedding
    torch.manual_seed(9)
    V, n = 5000, 3001
    emb = ShardedEmbedding(V, dim, seed=3)
    table0 = emb.shard[:V].clone()
    ids = torch.randint(0, V, (n,), device=cuda)
    ids[::17] = -1                                     # padding
    ids[5:40] = 7                                      # duplicates
    out = emb.lookup(ids.view(-1, 1), dtype=dtype)
    assert out.shape == (n, 1, dim)
    ref = table0[ids.clamp(min=0)] * (ids >= 0)[:, None]
    torch.testing.assert_close(out.view(n, dim).float(), ref.to(dtype).float(), atol=0, rtol=0)
    grad = torch.randn(n, dim, device=cuda).to(dtype)
    emb.apply_grad(ids, grad, lr=0.5)
    want = table0.clone()
    valid = ids >= 0
    want.index_add_(0, ids[valid], grad[valid].float(), alpha=-0.5)
    torch.testing.assert_close(emb.shard[:V], want, atol=2e-4, rtol=1e-4)
    sd = emb.state_dict()
    emb.shard.zero_()
    emb.load_state_dict(sd)
    torch.testing.assert_close(emb.shard[:V], want, atol=2e-4, rtol=1e-4)
