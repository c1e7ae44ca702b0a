"""GBDT kernels (N3) vs a plain PyTorch fp64 oracle (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(n, F, dev, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    X = torch.randn(n, F, device=dev, generator=g)
    w = torch.randn(F, device=dev, generator=g)
    logit = X @ w + 0.5 * X[:, 0] * X[:, 1] - 0.3 * (X[:, 2] > 0.5).float()
    y = (torch.rand(n, device=dev, generator=g) < torch.sigmoid(logit)).float()
    return X, y


def test_histogram_matches_index_add(cuda):
    from kubedl_b200.models import gbdt
    X, y = _data(50_000, 28, cuda)
    tr = gbdt.GBDTTrainer(X, y, max_depth=3)
    lib = tr.lib
    from kubedl_b200.ops import lib as L
    L.check(lib.kdl_gbdt_grad_logistic(tr.pred.data_ptr(), tr.y.data_ptr(), tr.gh.data_ptr(), tr.n, L.stream_ptr()))
    torch.testing.assert_close(tr.gh[:, 0], torch.sigmoid(tr.pred) - y)
    # fake node assignment with 4 nodes, build nodes 1 and 2
    tr.node = torch.randint(0, 4, (tr.n,), device=cuda, dtype=torch.int32)
    order = torch.sort(tr.node, stable=True)
    counts = torch.bincount(tr.node.long(), minlength=4)
    seg = torch.zeros(5, dtype=torch.int32, device=cuda)
    seg[1:] = counts.cumsum(0).int()
    build = torch.tensor([1, 2], dtype=torch.int32, device=cuda)
    hist = tr._build_hist(order.indices.int(), seg, build, 2).clone()
    ref = torch.zeros(2, 28, 256, 2, dtype=torch.float64, device=cuda)
    for k, nd in enumerate((1, 2)):
        m = tr.node == nd
        for f in range(28):
            ref[k, f].index_add_(0, tr.bins[m, f].long(), tr.gh[m].double())
    torch.testing.assert_close(hist.double(), ref, atol=1e-2, rtol=1e-3)
    # second call (node_done counters self-reset, level-parity buffer switch)
    hist2 = tr._build_hist(order.indices.int(), seg, build, 2)
    torch.testing.assert_close(hist2.double(), ref, atol=1e-2, rtol=1e-3)


def test_tree_matches_oracle_and_loss_decreases(cuda):
    from kubedl_b200.models import gbdt
    X, y = _data(200_000, 28, cuda, seed=1)
    tr = gbdt.GBDTTrainer(X, y, max_depth=5, learning_rate=0.3)
    from kubedl_b200.ops import lib as L
    L.check(tr.lib.kdl_gbdt_grad_logistic(tr.pred.data_ptr(), tr.y.data_ptr(), tr.gh.data_ptr(), tr.n, L.stream_ptr()))
    gh0 = tr.gh.clone()
    tree = tr.grow_tree()
    splits, leaf_ref, node_ref = gbdt.reference_tree_torch(tr.bins, gh0, 28, 5, 1.0, 0.3)
    agree = total = 0
    for d, (feat, thr, gain, valid) in enumerate(splits):
        s = tree.splits[d]
        same = (s[:, 0].long() == feat) & (s[:, 1].long() == thr)
        agree += int(same.sum()); total += same.numel()
        torch.testing.assert_close(s[:, 2].double(), gain.clamp(min=0), atol=0.05, rtol=2e-2)
    assert agree / total > 0.9, (agree, total)
    same_leaf = (tr.node.long() == node_ref).float().mean().item()
    assert same_leaf > 0.95, same_leaf
    if same_leaf == 1.0:
        torch.testing.assert_close(tree.leaf_value.double(), leaf_ref, atol=1e-3, rtol=1e-2)
    l0 = float(torch.nn.functional.binary_cross_entropy_with_logits(torch.zeros_like(y), y))
    losses = [float(tr.logloss())]
    for _ in range(4):
        tr.grow_tree()
        losses.append(float(tr.logloss()))
    assert losses[0] < l0 and all(b < a for a, b in zip(losses, losses[1:])), (l0, losses)
    # the stored trees reproduce the running prediction
    torch.testing.assert_close(tr.predict_bins(tr.bins), tr.pred, atol=1e-4, rtol=1e-4)
    # and so does the exported model file, from raw rows, on either device (what a Predict job does)
    import os
    import tempfile
    from kubedl_b200.models import gbdt as G
    with tempfile.TemporaryDirectory() as d:
        G.save_model(os.path.join(d, "m", "model.pt"), G.export_model(tr.cuts, tr.trees, learning_rate=tr.lr))
        model = G.load_model(os.path.join(d, "m", "model.pt"))
    assert len(model["trees"]) == len(tr.trees)
    torch.testing.assert_close(G.predict_margin(model, X), tr.pred, atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(G.predict_margin(model, X[:4096].cpu()), tr.pred[:4096].cpu(), atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("n,nodes", [(1000, 1), (100_003, 8), (1_000_000, 128), (257, 256)])
def test_partition_groups_rows_by_node(cuda, n, nodes):
    """Counting-sort row partition: a permutation of the rows, node-major, with matching segments."""
    import ctypes as C
    from kubedl_b200.models import gbdt  # noqa: F401  (registers the signatures)
    from kubedl_b200.ops import lib as L
    lib = L.lib()
    torch.manual_seed(n)
    node = torch.randint(0, nodes, (n,), device=cuda, dtype=torch.int32)
    ridx = torch.empty(n, dtype=torch.int32, device=cuda)
    seg = torch.empty(nodes + 1, dtype=torch.int32, device=cuda)
    counts = torch.empty(nodes, dtype=torch.int32, device=cuda)
    ws = torch.empty(lib.kdl_gbdt_partition_workspace_bytes(nodes), dtype=torch.uint8, device=cuda)
    L.check(lib.kdl_gbdt_partition(node.data_ptr(), n, nodes, ridx.data_ptr(), seg.data_ptr(), counts.data_ptr(),
                                   ws.data_ptr(), ws.numel(), L.stream_ptr()), "kdl_gbdt_partition")
    torch.cuda.synchronize()
    ref_counts = torch.bincount(node.long(), minlength=nodes)
    assert torch.equal(counts.long(), ref_counts)
    assert torch.equal(seg.long(), torch.cat([torch.zeros(1, dtype=torch.long, device=cuda), ref_counts.cumsum(0)]))
    assert torch.equal(torch.sort(ridx.long()).values, torch.arange(n, device=cuda))       # a permutation
    grouped = node[ridx.long()]
    assert bool((grouped[1:] >= grouped[:-1]).all())                                          # node-major
