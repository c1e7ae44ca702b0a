"""The layout algebra behind csrc/stem_s2d.cu, checked on the CPU: a 7x7 / stride-2 / pad-3 convolution on a 3-channel
image equals a 4x4 / stride-1 convolution on the zero-bordered 2x2 space-to-depth image with the filter re-indexed as
k2 = kh'*64 + kw'*16 + (dy*2+dx)*3 + c  <->  (kh, kw, c) = (2kh'+dy-1, 2kw'+dx-1, c).  (The kernels themselves are
tested on the GPU in tests/test_gpu_resnet.py; no counterpart in the reference.)"""
import torch


def k2_to_padded(k2: int) -> int:
    """Column kh*24 + kw*3 + c of the padded [64][168] filter for space-to-depth column k2, -1 where k2 is padding
    (the same function as s2d_k_to_padded in csrc/stem_s2d.cu)."""
    khp, kwp, ch = k2 >> 6, (k2 >> 4) & 3, k2 & 15
    if ch >= 12:
        return -1
    dy, dx, c = ch // 6, (ch // 3) & 1, ch % 3
    kh, kw = 2 * khp + dy - 1, 2 * kwp + dx - 1
    if kh < 0 or kw < 0:
        return -1
    return kh * 24 + kw * 3 + c


def test_mapping_is_a_bijection_onto_the_147_taps():
    cols = [k2_to_padded(k) for k in range(256)]
    used = [c for c in cols if c >= 0]
    assert len(used) == len(set(used)) == 147
    assert all((c % 24) < 21 and c < 168 for c in used)


def test_space_to_depth_convolution_equals_the_strided_one():
    torch.manual_seed(0)
    x = torch.randn(2, 224, 224, 3)
    w = torch.randn(5, 7, 7, 3)
    ref = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), stride=2, padding=3).permute(0, 2, 3, 1)
    S = torch.zeros(2, 115, 115, 16)
    S[:, 2:114, 2:114, :12] = x.view(2, 112, 2, 112, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(2, 112, 112, 12)
    wp = torch.zeros(5, 168)
    wp.view(5, 7, 24)[:, :, :21] = w.reshape(5, 7, 21)
    W2 = torch.zeros(5, 256)
    for k2 in range(256):
        k = k2_to_padded(k2)
        if k >= 0:
            W2[:, k2] = wp[:, k]
    y = torch.zeros(2, 112, 112, 5)
    for khp in range(4):
        # what one TMA box of the kernel holds: for every output pixel, the 4 pixels x 16 channels starting at it
        win = S[:, khp:khp + 112].unfold(2, 4, 1).permute(0, 1, 2, 4, 3).reshape(2, 112, 112, 64)
        y += win @ W2[:, khp * 64:(khp + 1) * 64].t()
    torch.testing.assert_close(y, ref, atol=2e-4, rtol=1e-4)
