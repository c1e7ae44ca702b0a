"""tcgen05 GEMM (kernel N6) vs torch.matmul in fp32 (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel_err(out, ref):
    return float((out.float() - ref).abs().max() / (ref.abs().max() + 1e-6))


@pytest.mark.parametrize("M,N,K", [
    (128, 128, 64), (256, 64, 64), (1000, 256, 64), (4096, 512, 2048), (16, 1000, 2048),
    (3136, 64, 256), (777, 192, 328), (12544, 2048, 512),
])
def test_gemm_nt(cuda, M, N, K):
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(0)
    a = torch.randn(M, K, device=cuda).bfloat16()
    b = torch.randn(N, K, device=cuda).bfloat16()
    out = g.nt(a, b)
    ref = a.float() @ b.float().t()
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)


def test_gemm_nt_epilogue(cuda):
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(1)
    M, N, K = 512, 1000, 256
    a = torch.randn(M, K, device=cuda).bfloat16()
    b = torch.randn(N, K, device=cuda).bfloat16()
    bias = torch.randn(N, device=cuda).bfloat16()
    add = torch.randn(M, N, device=cuda).bfloat16()
    out = g.nt(a, b, bias=bias, addend=add, relu=True)
    ref = torch.relu(a.float() @ b.float().t() + bias.float() + add.float())
    assert _rel_err(out, ref) < 1e-2


@pytest.mark.parametrize("M,N,K", [(256, 128, 64), (1024, 64, 256), (3000, 256, 1024), (640, 2048, 512)])
def test_gemm_nn(cuda, M, N, K):
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(2)
    a = torch.randn(M, K, device=cuda).bfloat16()
    b = torch.randn(K, N, device=cuda).bfloat16()
    add = torch.randn(M, N, device=cuda).bfloat16()
    out = g.nn(a, b)
    ref = a.float() @ b.float()
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)
    out2 = g.nn(a, b, addend=add)
    assert _rel_err(out2, ref + add.float()) < 1e-2


@pytest.mark.parametrize("K,M,N,split", [
    (4096, 64, 64, 0), (50000, 256, 64, 0), (1024, 128, 128, 1), (8192, 2048, 512, 0),
    (3000, 1000, 2048, 0), (100352, 64, 256, 0),
    # few output elements, many partial planes: the shared-element reduction, with plane counts that are not
    # multiples of its sub-lane count
    (20000, 64, 64, 23), (30000, 128, 64, 50), (12800, 192, 64, 19), (40000, 64, 128, 101),
])
def test_gemm_tn_splitk(cuda, K, M, N, split):
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(3)
    a = torch.randn(K, M, device=cuda).bfloat16()
    b = torch.randn(K, N, device=cuda).bfloat16()
    out = g.tn(a, b, split_k=split)
    ref = a.float().t() @ b.float()
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)
    out2 = g.tn(a, b, split_k=split)
    assert torch.equal(out, out2)  # deterministic split-K


@pytest.mark.parametrize("mode,M,N,K", [
    ("nt", 2048, 256, 256), ("nt", 2000, 1000, 512), ("nt", 1300, 320, 264), ("nt", 8192, 2048, 1024),
    ("nn", 2048, 512, 256), ("nn", 5000, 1024, 768), ("tn", 1024, 256, 512), ("tn", 1536, 768, 2048),
])
def test_gemm_2cta_path(cuda, mode, M, N, K):
    """Shapes that the dispatcher sends to the cta_group::2 kernel (N >= 256, M >= 1024, K >= 256),
    with ragged M / N / K so TMA clipping and zero fill are exercised on both CTAs of the pair."""
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(7)
    if mode == "nt":
        a = torch.randn(M, K, device=cuda).bfloat16(); b = torch.randn(N, K, device=cuda).bfloat16()
        bias = torch.randn(N, device=cuda).bfloat16(); add = torch.randn(M, N, device=cuda).bfloat16()
        out = g.nt(a, b, bias=bias, addend=add, relu=True)
        ref = torch.relu(a.float() @ b.float().t() + bias.float() + add.float())
    elif mode == "nn":
        a = torch.randn(M, K, device=cuda).bfloat16(); b = torch.randn(K, N, device=cuda).bfloat16()
        out = g.nn(a, b)
        ref = a.float() @ b.float()
    else:
        a = torch.randn(K, M, device=cuda).bfloat16(); b = torch.randn(K, N, device=cuda).bfloat16()
        out = g.tn(a, b, split_k=1)
        ref = a.float().t() @ b.float()
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)


@pytest.mark.parametrize("M,N,K", [(5000, 64, 64), (30000, 256, 64), (4096, 1024, 256), (1500, 2048, 512), (777, 192, 328)])
def test_gemm_fused_bn_stats_and_bn_from_stats(cuda, M, N, K):
    """Column statistics from the GEMM epilogue (1-CTA and 2-CTA kernels) == statistics of the stored
    bf16 output, and BN fed by them == BN that recomputes them."""
    from kubedl_b200.ops.gemm import Gemm
    from kubedl_b200.ops import nn as Kn
    g = Gemm(cuda)
    torch.manual_seed(11)
    a = torch.randn(M, K, device=cuda).bfloat16(); b = (torch.randn(N, K, device=cuda) * 0.2).bfloat16()
    y, (partial, nparts) = g.nt_stats(a, b)
    ref = a.float() @ b.float().t()
    assert _rel_err(y, ref) < 1e-2
    s = partial[:nparts].double().sum(0)
    yf = y.double()
    torch.testing.assert_close(s[0], yf.sum(0), atol=1e-2 * M ** 0.5, rtol=1e-3)
    torch.testing.assert_close(s[1], (yf * yf).sum(0), atol=1e-2 * M ** 0.5, rtol=1e-3)
    gamma, beta = torch.rand(N, device=cuda) + 0.5, torch.randn(N, device=cuda) * 0.1
    y1, sv1 = Kn.bn_fwd(y, gamma, beta, None, None, relu=True)
    y2, sv2 = Kn.bn_fwd(y, gamma, beta, None, None, relu=True, stats=(partial, nparts))
    torch.testing.assert_close(sv2, sv1, atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(y2.float(), y1.float(), atol=2e-2, rtol=2e-2)


def test_nn_compact_stride2_addend(cuda):
    """dgrad GEMM + the compact gradient of a 1x1/stride-2 sibling branch added at even pixels."""
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(3)
    for (n, H, W, K, N) in [(2, 8, 8, 64, 256), (3, 28, 28, 128, 512), (2, 56, 56, 64, 256)]:
        M = n * H * W
        a = torch.randn(M, K, device=cuda).bfloat16()
        b = (torch.randn(K, N, device=cuda) * 0.1).bfloat16()
        addc = torch.randn(n, H // 2, W // 2, N, device=cuda).bfloat16()
        out = g.nn(a, b, addend=addc.view(-1, N), addend_hw=(H, W))
        ref = (a.float() @ b.float()).view(n, H, W, N)
        ref[:, ::2, ::2, :] += addc.float()
        err = float((out.view(n, H, W, N).float() - ref).abs().max() / ref.abs().max())
        assert err < 1.5e-2, (n, H, W, K, N, err)


@pytest.mark.parametrize("M,N,K", [
    (256, 128, 64),        # one tile, one chunk pair
    (50000, 256, 64),      # 1-CTA kernel, many tiles per CTA (register + L2 prefetch across tiles), M tail
    (4099, 320, 128),      # partial N tile: the element-wise tail path
    (6000, 1024, 256),     # 2-CTA kernel, 4 chunks per warp and tile, M tail
    (40000, 512, 128),     # 1-CTA kernel, 4 column tiles
])
def test_nn_addend_prefetch_pipeline(cuda, M, N, K):
    """The addend registers are loaded one chunk (and one tile) ahead of where they are added: every
    chunk must still receive ITS rows and columns, on both kernels, including tails."""
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(21)
    a = torch.randn(M, K, device=cuda).bfloat16()
    b = (torch.randn(K, N, device=cuda) * 0.05).bfloat16()
    add = torch.randn(M, N, device=cuda).bfloat16()
    out = g.nn(a, b, addend=add)
    ref = a.float() @ b.float() + add.float()
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)
    # the addend dominates the product here, so a chunk that got another chunk's addend shows up element-wise
    assert float((out.float() - ref).abs().max()) < 0.1


@pytest.mark.parametrize("M,N,K", [(256, 64, 64), (50000, 256, 64), (6000, 1024, 256), (12544, 2048, 512)])
def test_nn_masked_addend_matches_masked_copy(cuda, M, N, K):
    """out = a @ b + mask * addend with the 1-bit ReLU mask BatchNorm saves (bit j of byte v <-> element 8v+j)
    == the same GEMM fed a separately masked copy of the addend."""
    from kubedl_b200.ops.gemm import Gemm
    g = Gemm(cuda)
    torch.manual_seed(22)
    a = torch.randn(M, K, device=cuda).bfloat16()
    b = (torch.randn(K, N, device=cuda) * 0.05).bfloat16()
    add = torch.randn(M, N, device=cuda).bfloat16()
    keep = torch.rand(M, N, device=cuda) < 0.6
    weights = (1 << torch.arange(8, device=cuda, dtype=torch.int32))
    mask = (keep.view(-1, 8).int() * weights).sum(1).to(torch.uint8)
    out = g.nn(a, b, addend=add, addend_mask=mask)
    ref = a.float() @ b.float() + add.float() * keep
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)
    assert float((out.float() - ref).abs().max()) < 0.1
    want = g.nn(a, b, addend=(add.float() * keep).bfloat16())
    assert torch.equal(out, want)
