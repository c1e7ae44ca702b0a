"""The allocation-free BatchNorm entry point, driven through the C ABI exactly as a foreign
caller would (ctypes, raw pointers, a raw stream handle).

Some users of libkubedl_b200.so are not allowed to allocate on the device: the whole GPU may
belong to another component's memory pool, or the call sits where cudaMalloc's device-wide
stall cannot be tolerated.  For them `kdl_bn_fwd_static_scratch` keeps its scratch in a fixed
array inside the library.  This test checks that it returns what the ordinary entry point
(`kdl_bn_fwd`, caller-owned workspace) returns, twice in a row — the array is shared between
calls and still holds the previous call's data the second time — and that a shape whose scratch
need exceeds the built-in limit is refused with a readable message instead of overrunning.

The file name sorts last on purpose: a `pytest -x` run reaches it only after everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind(lib):
    vp, i64, ci, cf = C.c_void_p, C.c_int64, C.c_int, C.c_float
    lib.kdl_bn_fwd_static_scratch.restype = ci
    lib.kdl_bn_fwd_static_scratch.argtypes = [vp, vp, vp, i64, ci, vp, vp, vp, vp, cf, cf, ci, vp, vp]
    lib.kdl_bn_static_scratch_limit.restype = C.c_size_t
    lib.kdl_bn_static_scratch_limit.argtypes = []
    lib.kdl_bn_fwd.restype = ci
    lib.kdl_bn_fwd.argtypes = [vp, vp, vp, i64, ci, vp, vp, vp, vp, cf, cf, ci, vp, vp, C.c_size_t, vp]
    lib.kdl_bn_workspace_bytes.restype = C.c_size_t
    lib.kdl_bn_workspace_bytes.argtypes = [i64, ci]
    lib.kdl_last_error.restype = C.c_char_p
    return lib


@pytest.mark.parametrize("M,C_,use_res", [(6000, 256, True), (20000, 64, False)])
def test_static_scratch_matches_workspace_entry_point(cuda, M, C_, use_res):
    from kubedl_b200.ops import build
    lib = _bind(C.CDLL(str(build.ensure_built())))
    assert lib.kdl_bn_static_scratch_limit() == 512 * 1024
    stream = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(0)
    gamma = torch.rand(C_, device=cuda) + 0.5
    beta = torch.randn(C_, device=cuda) * 0.1
    need = lib.kdl_bn_workspace_bytes(M, C_)
    assert 0 < need <= lib.kdl_bn_static_scratch_limit()
    ws = torch.zeros(need, dtype=torch.uint8, device=cuda)
    for call in range(2):  # two calls in a row, different data each time
        x = (torch.randn(M, C_, device=cuda) * (1.0 + call) + 0.25 * call).bfloat16()
        res = torch.randn(M, C_, device=cuda).bfloat16() if use_res else None
        rp = res.data_ptr() if use_res else None
        y_ref, y_new = torch.empty_like(x), torch.empty_like(x)
        save_ref = torch.empty(2, C_, device=cuda)
        save_new = torch.empty(2, C_, device=cuda)
        rm_ref, rv_ref = torch.zeros(C_, device=cuda), torch.ones(C_, device=cuda)
        rm_new, rv_new = torch.zeros(C_, device=cuda), torch.ones(C_, device=cuda)
        rc = lib.kdl_bn_fwd(x.data_ptr(), rp, y_ref.data_ptr(), M, C_, gamma.data_ptr(), beta.data_ptr(),
                            rm_ref.data_ptr(), rv_ref.data_ptr(), 1e-5, 0.1, 1, save_ref.data_ptr(),
                            ws.data_ptr(), ws.numel(), stream)
        assert rc == 0, lib.kdl_last_error()
        rc = lib.kdl_bn_fwd_static_scratch(x.data_ptr(), rp, y_new.data_ptr(), M, C_, gamma.data_ptr(),
                                           beta.data_ptr(), rm_new.data_ptr(), rv_new.data_ptr(), 1e-5,
                                           0.1, 1, save_new.data_ptr(), stream)
        assert rc == 0, lib.kdl_last_error()
        torch.cuda.synchronize()
        torch.testing.assert_close(y_new.float(), y_ref.float(), atol=3e-2, rtol=2e-2)
        torch.testing.assert_close(save_new, save_ref, atol=1e-3, rtol=1e-3)
        torch.testing.assert_close(rm_new, rm_ref, atol=1e-3, rtol=1e-3)
        torch.testing.assert_close(rv_new, rv_ref, atol=1e-3, rtol=1e-3)
        # and both agree with plain PyTorch in fp32
        xf = x.float()
        z = (xf - xf.mean(0)) * torch.rsqrt(xf.var(0, unbiased=False) + 1e-5) * gamma + beta
        if use_res:
            z = z + res.float()
        torch.testing.assert_close(y_new.float(), torch.relu(z), atol=3e-2, rtol=2e-2)


def test_static_scratch_rejects_oversized_shape(cuda):
    from kubedl_b200.ops import build
    lib = _bind(C.CDLL(str(build.ensure_built())))
    M, C_ = 400000, 1000  # C not a multiple of 64 -> one wide slice -> several MB of partials
    assert lib.kdl_bn_workspace_bytes(M, C_) > lib.kdl_bn_static_scratch_limit()
    x = torch.zeros(8, C_, device=cuda, dtype=torch.bfloat16)  # never touched: the call must bail out first
    g = torch.ones(C_, device=cuda)
    save = torch.empty(2, C_, device=cuda)
    rc = lib.kdl_bn_fwd_static_scratch(x.data_ptr(), None, x.data_ptr(), M, C_, g.data_ptr(), g.data_ptr(),
                                       None, None, 1e-5, 0.1, 1, save.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.kdl_last_error().decode()
    assert "scratch" in msg and "kdl_bn_fwd" in msg, msg
    torch.cuda.synchronize()
