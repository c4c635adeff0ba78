"""Allocation-free GroupNorm entry point (afb_groupnorm_tokens_static_ws, include/adaface_b200.h): the statistics
workspace is a `static __device__` array inside the library instead of a caller buffer.  Collected last on purpose.
Compared, through the C ABI, with the existing afb_groupnorm_tokens on the same
inputs, on two calls in a row that share the built-in scratch."""
import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu


def _both(lib, check, x, layout, gamma, beta, b, c, hw, groups, eps, act, pre_add=None):
    dev = x.device
    s = torch.cuda.current_stream().cuda_stream
    y_ref = torch.empty((b, hw, c), dtype=torch.bfloat16, device=dev)
    y_new = torch.full((b, hw, c), float("nan"), dtype=torch.bfloat16, device=dev)
    ws = torch.empty(int(lib.afb_groupnorm_ws_floats(b, hw, groups)), dtype=torch.float32, device=dev)
    add_ptr, add_ld = (None, 0) if pre_add is None else (pre_add.data_ptr(), pre_add.stride(0))
    check(lib.afb_groupnorm_tokens(x.data_ptr(), layout, gamma.data_ptr(), beta.data_ptr(), y_ref.data_ptr(),
                                   ws.data_ptr(), b, c, hw, groups, eps, act, add_ptr, add_ld, s), "afb_groupnorm_tokens")
    torch.cuda.synchronize()
    # everything the new entry point touches is allocated above: the call itself must not move the allocator
    before = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    check(lib.afb_groupnorm_tokens_static_ws(x.data_ptr(), layout, gamma.data_ptr(), beta.data_ptr(), y_new.data_ptr(),
                                             b, c, hw, groups, eps, act, add_ptr, add_ld, s),
          "afb_groupnorm_tokens_static_ws")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == before
    return y_ref, y_new


def test_groupnorm_static_ws_matches_existing_entry_point_twice(cuda_device):
    from adaface_b200 import _lib
    lib, d = _lib.load(), cuda_device
    g = torch.Generator().manual_seed(1234)
    cases = [
        # (batch, channels, h*w, layout, act, pre_add): two-kernel token-major path with chunk partials ...
        (4, 320, 64 * 64, 1, 0, False),
        # ... then a call with another shape, SiLU and the per-sample pre-add over the SAME scratch (stale contents)
        (3, 640, 32 * 32, 1, 1, True),
        # fp32 NCHW layout (the reference's calling convention) and the single-launch cluster path (scratch unused)
        (2, 320, 16 * 16, 0, 0, False),
        (2, 1280, 8 * 8, 1, 1, False),
    ]
    for rep in range(2):                                   # "two calls in a row" for every case, too
        for b, c, hw, layout, act, with_add in cases:
            xf = torch.randn((b, hw, c), generator=g) * 1.5 + 0.7
            gamma = (1 + 0.1 * torch.randn(c, generator=g)).to(d)
            beta = (0.1 * torch.randn(c, generator=g)).to(d)
            x = xf.transpose(1, 2).contiguous().to(d) if layout == 0 else xf.to(torch.bfloat16).to(d)
            add = torch.randn((b, c), generator=g).to(d) if with_add else None
            y_ref, y_new = _both(lib, _lib.check, x, layout, gamma, beta, b, c, hw, 32, 1e-6, act, add)
            assert torch.isfinite(y_new.float()).all()
            # same kernels, same geometry; the in-kernel group reductions use shared-memory atomics, so the bar is
            # the GroupNorm tests' own (test_kernels_gpu.py: 4e-3), not bit equality
            assert rel_l2(y_new.float(), y_ref.float()) < 4e-3, (rep, b, c, hw, layout, act)


def test_groupnorm_static_ws_rejects_sizes_beyond_the_built_in_scratch(cuda_device):
    from adaface_b200 import _lib
    lib, d = _lib.load(), cuda_device
    assert int(lib.afb_groupnorm_static_ws_max_floats()) == 256 * 1024
    b, c, hw, groups = 125, 64, 16, 32                     # 2 * 125 * 32 * 33 = 264000 > 262144 floats
    x = torch.zeros((b, hw, c), dtype=torch.bfloat16, device=d)
    y = torch.empty_like(x)
    gamma, beta = torch.ones(c, device=d), torch.zeros(c, device=d)
    rc = lib.afb_groupnorm_tokens_static_ws(x.data_ptr(), 1, gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), b, c, hw,
                                            groups, 1e-6, 0, None, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 2
    msg = lib.afb_last_error().decode()
    assert "built-in scratch" in msg and "afb_groupnorm_tokens" in msg
    with pytest.raises(_lib.AdafaceB200Error):
        _lib.check(rc, "afb_groupnorm_tokens_static_ws")
    b = 124                                                 # the documented limit itself is accepted
    x = torch.randn((b, hw, c), device=d).to(torch.bfloat16)
    y = torch.empty_like(x)
    _lib.check(lib.afb_groupnorm_tokens_static_ws(x.data_ptr(), 1, gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), b, c,
                                                  hw, groups, 1e-6, 0, None, 0, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all()
