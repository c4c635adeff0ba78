"""Kernel-level parity on the B200: every C-ABI entry point against the CPU oracle (oracle/port.py) or the
plain fp32 formula of the same op, on seeded inputs.  Run with `-m gpu`."""
import math

import pytest
import torch

from conftest import load_golden, rel_l2
from oracle import params as P
from oracle import port

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2          # north_star: per-block relative L2 <= 1e-2 in bf16


@pytest.fixture(scope="module")
def ops(cuda_device):
    from adaface_b200 import ops as o
    return o


def _r(key, shape, scale=1.0, seed=101):
    return P.randn(key, shape, seed, scale)


# ---------------------------------------------------------------------------------------------- K14
@pytest.mark.parametrize("n", [(8, 4, 64, 64), (3, 4, 8, 8), (1, 1, 1, 7)])
@pytest.mark.parametrize("sigma", [0.0, 0.3])
def test_cfg_ddim_step_bit_exact(ops, cuda_device, n, sigma):
    x, ec, eu, nz = (_r(f"ddim.{k}", n) for k in ("x", "ec", "eu", "nz"))
    a_t, a_prev, s1m = 0.0413, 0.0521, math.sqrt(1 - 0.0413)
    a_t, a_prev, s1m = (float(torch.tensor(v, dtype=torch.float32)) for v in (a_t, a_prev, s1m))
    ref_prev, ref_x0 = port.cfg_ddim_update(x, ec, eu, 7.5, a_t, a_prev, sigma, s1m, nz if sigma else None)
    d = cuda_device
    got_prev, got_x0 = ops.cfg_ddim_step(x.to(d), ec.to(d), eu.to(d), nz.to(d) if sigma else None, 7.5, a_t, a_prev,
                                         sigma, s1m, 1.0)
    assert torch.equal(got_x0.cpu(), ref_x0)
    assert torch.equal(got_prev.cpu(), ref_prev)


# ---------------------------------------------------------------------------------------------- K1 / K3
@pytest.mark.parametrize("c,hw", [(320, 16), (640, 8), (1280, 8), (320, 12), (2560, 8), (1920, 8), (960, 16)])
def test_groupnorm_tokens_both_layouts(ops, cuda_device, c, hw):
    b = 2
    x = _r("gn.x", (b, c, hw, hw)) * 1.5 + 0.7
    g, be = 1 + 0.1 * _r("gn.g", (c,)), 0.1 * _r("gn.b", (c,))
    ref = port.group_norm(x, g, be, 32, 1e-6).reshape(b, c, hw * hw).transpose(1, 2)
    d = cuda_device
    y0 = ops.groupnorm_tokens(x.to(d), g.to(d), be.to(d), 32, 1e-6)
    assert rel_l2(y0.float(), ref) < 4e-3                 # bf16 output rounding only
    xb = x.to(d).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ref1 = port.group_norm(xb.float().cpu(), g, be, 32, 1e-6).reshape(b, c, hw * hw).transpose(1, 2)
    y1 = ops.groupnorm_tokens(xb, g.to(d), be.to(d), 32, 1e-6)
    assert rel_l2(y1.float(), ref1) < 4e-3
    # fused SiLU (the ResBlocks' GroupNorm32 -> SiLU), both layouts
    silu = torch.nn.functional.silu
    assert rel_l2(ops.groupnorm_tokens(x.to(d), g.to(d), be.to(d), 32, 1e-5, 1).float(),
                  silu(port.group_norm(x, g, be, 32, 1e-5)).reshape(b, c, hw * hw).transpose(1, 2)) < 4e-3
    assert rel_l2(ops.groupnorm_tokens(xb, g.to(d), be.to(d), 32, 1e-5, 1).float(),
                  silu(port.group_norm(xb.float().cpu(), g, be, 32, 1e-5)).reshape(b, c, hw * hw).transpose(1, 2)) < 4e-3


def test_groupnorm_pre_add_and_add_bias_residual(ops, cuda_device):
    """The ResBlock folds: GroupNorm(x + pre_add[b, c]) with SiLU, and a + b + bias in one pass."""
    b, c, hw = 3, 640, 12
    d = cuda_device
    x = (_r("gnp.x", (b, c, hw, hw)) * 1.5 + 0.3).to(torch.bfloat16)
    g, be = 1 + 0.1 * _r("gnp.g", (c,)), 0.1 * _r("gnp.b", (c,))
    big = _r("gnp.add", (b, 3 * c)).to(d)
    add = big[:, c:2 * c]                                        # strided rows
    xb = x.to(d).contiguous(memory_format=torch.channels_last)
    ref = torch.nn.functional.silu(port.group_norm(x.float() + add.cpu()[:, :, None, None], g, be, 32, 1e-5))
    y = ops.groupnorm_tokens(xb, g.to(d), be.to(d), 32, 1e-5, 1, add)
    assert rel_l2(y.float(), ref.reshape(b, c, hw * hw).transpose(1, 2)) < 4e-3
    a2 = _r("abr.a", (b, hw * hw, c)).to(torch.bfloat16)
    b2 = _r("abr.b", (b, hw * hw, c)).to(torch.bfloat16)
    bias = _r("abr.bias", (c,))
    out = ops.add_bias_residual(a2.to(d), b2.to(d), bias.to(d))
    assert rel_l2(out.float(), a2.float() + b2.float() + bias) < 3e-3
    out = ops.add_bias_residual(a2.to(d), None, None)
    assert torch.equal(out.cpu(), a2)


@pytest.mark.parametrize("c", [320, 640, 1280, 2880])
def test_layernorm(ops, cuda_device, c):
    rows = 77
    x = (_r("ln.x", (rows, c)) * 2 + 0.3).to(torch.bfloat16)
    g, be = 1 + 0.1 * _r("ln.g", (c,)), 0.1 * _r("ln.b", (c,))
    d = cuda_device
    ref = port.layer_norm(x.float(), g, be)
    assert rel_l2(ops.layernorm(x.to(d), g.to(d), be.to(d), 1e-5).float(), ref) < 4e-3
    ref = port.layer_norm(x.float())
    assert rel_l2(ops.layernorm(x.to(d), None, None, 1e-5).float(), ref) < 4e-3


@pytest.mark.parametrize("c1,c2,hw", [(640, 320, 32), (320, 320, 64), (1280, 640, 32), (1280, 1280, 16)])
def test_groupnorm_virtual_concat(ops, cuda_device, c1, c2, hw):
    """GroupNorm32 + SiLU of th.cat([h, skip], 1) (decoder ResBlocks, openaimodel.py:1004-1010, 201-205) read from the
    two sources: identical to the kernel on the materialised concatenation, and within bf16 rounding of the oracle."""
    b, d = 2, cuda_device
    x1 = (_r("gnc.x1", (b, c1, hw, hw)) * 1.3 + 0.2).to(torch.bfloat16)
    x2 = (_r("gnc.x2", (b, c2, hw, hw)) * 0.7 - 0.4).to(torch.bfloat16)
    g, be = 1 + 0.1 * _r("gnc.g", (c1 + c2,)), 0.1 * _r("gnc.b", (c1 + c2,))
    cl = torch.channels_last
    y = ops.groupnorm_tokens_cat(x1.to(d).contiguous(memory_format=cl), x2.to(d).contiguous(memory_format=cl), g.to(d),
                                 be.to(d), 32, 1e-5, 1)
    xc = torch.cat([x1, x2], dim=1)
    y_mat = ops.groupnorm_tokens(xc.to(d).contiguous(memory_format=cl), g.to(d), be.to(d), 32, 1e-5, 1)
    ref = torch.nn.functional.silu(port.group_norm(xc.float(), g, be, 32, 1e-5)).reshape(b, c1 + c2, hw * hw).transpose(1, 2)
    assert rel_l2(y.float(), ref) < 4e-3
    assert rel_l2(y.float(), y_mat.float()) < 2e-3        # (different chunking / kernel variant: not bit-identical)


# ---------------------------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("m,n,k", [(128, 160, 64), (256, 320, 320), (512, 960, 320), (200, 640, 768),
                                   (1024, 1280, 5120), (64, 1280, 1280), (4096, 320, 1280), (154, 128, 768),
                                   # K >= 1280 with enough row blocks: the 320-column wide tiles (csrc/gemm.cu kWide),
                                   # ragged row tail
                                   (16384, 640, 1280), (9500, 1280, 1280), (5000, 960, 2560)])
def test_linear_store(ops, cuda_device, m, n, k):
    a = _r("g.a", (m, k)).to(torch.bfloat16)
    w = _r("g.w", (n, k), k ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("g.b", (n,))
    res = _r("g.r", (m, n)).to(torch.bfloat16)
    ref = a.float() @ w.float().t() + bias
    d = cuda_device
    out = ops.linear(a.to(d), w.to(d), bias.to(d), None, False)
    assert rel_l2(out.float(), ref) < 4e-3
    out = ops.linear(a.to(d), w.to(d), None, res.to(d), True)
    assert rel_l2(out, ref - bias + res.float()) < 1e-4          # fp32 out: only accumulation-order error
    out = ops.linear(a.to(d), w.to(d), bias.to(d), res.float().to(d), False)
    assert rel_l2(out.float(), ref + res.float()) < 4e-3


@pytest.mark.parametrize("b,cin,cout,h,w", [(2, 64, 160, 8, 8), (3, 320, 320, 16, 16), (2, 640, 320, 64, 64),
                                            (16, 1280, 1280, 8, 8), (1, 128, 128, 4, 32), (2, 192, 640, 32, 32),
                                            (1, 64, 32, 2, 128), (3, 64, 160, 8, 8), (5, 128, 320, 4, 4),
                                            # sub-box tiles (widths that are not a power of two: config 5's 96 / 48 / 24 / 12)
                                            (2, 128, 320, 96, 96), (3, 64, 160, 24, 24), (5, 64, 128, 12, 12), (2, 64, 160, 40, 20),
                                            (17, 1280, 1280, 12, 12)])
def test_conv3x3_implicit_gemm(ops, cuda_device, b, cin, cout, h, w):
    """3x3 / stride 1 / zero padding 1 convolution (ResBlock / Upsample convs, openaimodel.py:95-123,203-238) on the
    tcgen05 GEMM kernel against torch's fp32 conv2d of the same bf16-rounded operands."""
    x = _r("cv.x", (b, cin, h, w)).to(torch.bfloat16)
    wt = _r("cv.w", (cout, cin, 3, 3), (9 * cin) ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("cv.b", (cout,))
    res = _r("cv.r", (b, cout, h, w)).to(torch.bfloat16)
    ref = torch.nn.functional.conv2d(x.float(), wt.float(), bias, padding=1)
    d = cuda_device
    xt = x.to(d).permute(0, 2, 3, 1).reshape(b * h * w, cin).contiguous()
    wp = wt.to(d).permute(0, 2, 3, 1).reshape(cout, 9 * cin).contiguous()
    rt = res.to(d).permute(0, 2, 3, 1).reshape(b * h * w, cout).contiguous()
    to_nchw = lambda y: y.float().view(b, h, w, cout).permute(0, 3, 1, 2)
    assert rel_l2(to_nchw(ops.conv3x3(xt, wp, bias.to(d), None, h, w)), ref) < 4e-3
    assert rel_l2(to_nchw(ops.conv3x3(xt, wp, None, rt, h, w)), ref - bias[None, :, None, None] + res.float()) < 4e-3
    # channel slice of a wider NHWC buffer (pixel stride != Cin)
    wide = torch.cat([xt, xt.flip(1)], dim=1)
    assert rel_l2(to_nchw(ops.conv3x3(wide[:, :cin], wp, bias.to(d), None, h, w)), ref) < 4e-3


@pytest.mark.parametrize("b,cout,h,w", [(2, 320, 64, 64), (3, 320, 24, 40), (1, 64, 5, 7)])
def test_conv3x3_four_input_channels(ops, cuda_device, b, cout, h, w):
    """Conv2d(4, Cout, 3, padding=1) on the latent (openaimodel.py:575-579): im2col kernel + tcgen05 GEMM vs torch."""
    x = _r("c4.x", (b, 4, h, w)).to(torch.bfloat16)
    wt = _r("c4.w", (cout, 4, 3, 3), 36 ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("c4.b", (cout,))
    ref = torch.nn.functional.conv2d(x.float(), wt.float(), bias, padding=1)
    d = cuda_device
    xt = x.to(d).permute(0, 2, 3, 1).reshape(b * h * w, 4).contiguous()
    w64 = torch.zeros((cout, 64), dtype=torch.bfloat16, device=d)
    w64[:, :36] = wt.to(d).permute(0, 2, 3, 1).reshape(cout, 36)
    y = ops.conv3x3_c4(xt, w64, bias.to(d), h, w)
    assert rel_l2(y.float().view(b, h, w, cout).permute(0, 3, 1, 2), ref) < 4e-3


@pytest.mark.parametrize("b,c,cout,h,w", [(2, 320, 4, 64, 64), (3, 64, 8, 20, 24), (1, 128, 3, 9, 17), (2, 320, 4, 96, 96)])
def test_conv3x3_few_output_channels(ops, cuda_device, b, c, cout, h, w):
    """Conv2d(C, <= 8, 3, padding=1) -> fp32 NCHW (the U-Net's last convolution, openaimodel.py:716-720) vs torch."""
    x = _r("cf.x", (b, c, h, w)).to(torch.bfloat16)
    wt = _r("cf.w", (cout, c, 3, 3), (9 * c) ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("cf.b", (cout,))
    ref = torch.nn.functional.conv2d(x.float(), wt.float(), bias, padding=1)
    d = cuda_device
    xt = x.to(d).permute(0, 2, 3, 1).reshape(b * h * w, c).contiguous()
    w8 = torch.zeros((8, 9 * c), dtype=torch.bfloat16, device=d)
    w8[:cout] = wt.to(d).permute(0, 2, 3, 1).reshape(cout, 9 * c)
    y = ops.conv3x3_few_out(xt, w8, bias.to(d), cout, h, w)
    assert y.shape == (b, cout, h, w) and y.dtype == torch.float32
    assert rel_l2(y, ref) < 1e-4          # bf16 operands, fp32 accumulate and output


@pytest.mark.parametrize("b,cin,cout,h,w", [(2, 320, 320, 64, 64), (16, 640, 640, 32, 32), (16, 1280, 1280, 16, 16),
                                            (3, 64, 160, 16, 16), (1, 128, 128, 8, 64), (5, 64, 32, 8, 8),
                                            (2, 64, 160, 96, 96), (3, 128, 128, 48, 24)])      # sub-box output tiles
def test_conv3x3_stride2_implicit_gemm(ops, cuda_device, b, cin, cout, h, w):
    """3x3 / stride 2 / zero padding 1 convolution (Downsample.forward, openaimodel.py:138-165): the same implicit-GEMM
    kernel with a TMA element stride of 2 on the image axes, against torch's fp32 conv2d."""
    x = _r("cs.x", (b, cin, h, w)).to(torch.bfloat16)
    wt = _r("cs.w", (cout, cin, 3, 3), (9 * cin) ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("cs.b", (cout,))
    ref = torch.nn.functional.conv2d(x.float(), wt.float(), bias, stride=2, padding=1)
    d = cuda_device
    xt = x.to(d).permute(0, 2, 3, 1).reshape(b * h * w, cin).contiguous()
    wp = wt.to(d).permute(0, 2, 3, 1).reshape(cout, 9 * cin).contiguous()
    y = ops.conv3x3_s2(xt, wp, bias.to(d), h, w)
    assert y.shape == (b * (h // 2) * (w // 2), cout)
    assert rel_l2(y.float().view(b, h // 2, w // 2, cout).permute(0, 3, 1, 2), ref) < 4e-3


@pytest.mark.parametrize("b,cin,cout,h,w", [(2, 64, 160, 8, 8), (2, 128, 128, 16, 16), (16, 1280, 1280, 8, 8),
                                            (1, 640, 640, 32, 32), (3, 64, 320, 4, 32), (3, 128, 128, 8, 8),
                                            (5, 64, 160, 4, 4),
                                            (2, 64, 160, 48, 48), (3, 128, 128, 24, 24), (5, 64, 160, 12, 12)])   # sub-box tiles
def test_conv3x3_upsample_fold(ops, cuda_device, b, cin, cout, h, w):
    """Upsample.forward (openaimodel.py:95-123): nearest 2x + 3x3 conv as four folded 2x2 phase convolutions of the
    low-resolution image, against torch's fp32 interpolate + conv2d."""
    x = _r("up.x", (b, cin, h, w)).to(torch.bfloat16)
    wt = _r("up.w", (cout, cin, 3, 3), (9 * cin) ** -0.5)
    bias = 0.1 * _r("up.b", (cout,))
    ref = torch.nn.functional.conv2d(torch.nn.functional.interpolate(x.float(), scale_factor=2, mode="nearest"), wt, bias,
                                     padding=1)
    d = cuda_device
    wf = ops.fold_upsample_weight(wt)
    # the fold itself is exact in fp32: check it on the CPU against the reference formulation
    xf = x.float()
    xp = torch.nn.functional.pad(xf, (1, 1, 1, 1))
    chk = torch.zeros_like(ref)
    for py in (0, 1):
        for px in (0, 1):
            acc = 0
            for a in (0, 1):
                for c in (0, 1):
                    wt_ac = wf[(py * 2 + px) * cout:(py * 2 + px + 1) * cout, (a * 2 + c) * cin:(a * 2 + c + 1) * cin]
                    acc = acc + torch.einsum("bchw,oc->bohw", xp[:, :, py + a:py + a + h, px + c:px + c + w], wt_ac)
            chk[:, :, py::2, px::2] = acc + bias[None, :, None, None]
    assert rel_l2(chk, ref) < 1e-5
    xt = x.to(d).permute(0, 2, 3, 1).reshape(b * h * w, cin).contiguous()
    y = ops.conv3x3_up2(xt, wf.to(torch.bfloat16).to(d), bias.repeat(4).to(d), h, w)
    y = y.float().view(b, 2 * h, 2 * w, cout).permute(0, 3, 1, 2)
    assert rel_l2(y, ref) < 5e-3


def test_conv3x3_unsupported_geometry_raises(ops, cuda_device):
    from adaface_b200._lib import AdafaceB200Error
    d = cuda_device
    x = torch.zeros((9 * 9, 64), dtype=torch.bfloat16, device=d)
    wp = torch.zeros((32, 9 * 64), dtype=torch.bfloat16, device=d)
    with pytest.raises(AdafaceB200Error):
        ops.conv3x3(x, wp, None, None, 9, 9)             # widths must be a multiple of 4 (or divide 128)
    with pytest.raises(AdafaceB200Error):
        ops.conv3x3(torch.zeros((64, 48), dtype=torch.bfloat16, device=d),
                    torch.zeros((32, 9 * 48), dtype=torch.bfloat16, device=d), None, None, 8, 8)   # Cin % 64


@pytest.mark.parametrize("m,n,k1,k2", [(256, 320, 320, 320), (300, 640, 640, 320), (1024, 1280, 1280, 1280), (130, 160, 64, 40)])
def test_linear_virtual_concat_along_k(ops, cuda_device, m, n, k1, k2):
    """A = [a1 | a2] along K read from two tensors (the decoder's 1x1 skip convolution over th.cat([h, hs.pop()], 1),
    openaimodel.py:1004-1010, 277) against the GEMM on the materialised concatenation and the fp32 formula."""
    d = cuda_device
    a1 = _r("lc.a1", (m, k1)).to(torch.bfloat16).to(d)
    a2 = _r("lc.a2", (m, k2)).to(torch.bfloat16).to(d)
    w = _r("lc.w", (n, k1 + k2), (k1 + k2) ** -0.5).to(torch.bfloat16).to(d)
    bias = (0.1 * _r("lc.b", (n,))).to(d)
    res = _r("lc.r", (m, n)).to(torch.bfloat16).to(d)
    ref = torch.cat([a1, a2], 1).float() @ w.float().t() + bias + res.float()
    out = ops.linear_cat(a1, a2, w, bias, res)
    assert rel_l2(out.float(), ref) < 4e-3
    if (k1 + k2) % 8 == 0:
        mat = ops.linear(torch.cat([a1, a2], 1).contiguous(), w, bias, res, False)
        assert torch.equal(out, mat)          # same tiles, same accumulation order
    assert rel_l2(ops.linear_cat(a1, a2, w, None, None).float(), ref - bias - res.float()) < 4e-3


def test_linear_strided_views(ops, cuda_device):
    """A operand and residual as column slices of wider buffers (lda/ldr != K/N)."""
    m, n, k = 256, 320, 320
    big = _r("g.big", (m, 3 * k)).to(torch.bfloat16).to(cuda_device)
    w = _r("g.w2", (n, k), k ** -0.5).to(torch.bfloat16).to(cuda_device)
    a = big[:, k:2 * k]
    out = ops.linear(a, w, None, None, True)
    assert rel_l2(out, a.float() @ w.float().t()) < 1e-4


@pytest.mark.parametrize("m,c", [(256, 320), (128, 640), (77, 1280)])
def test_linear_geglu(ops, cuda_device, m, c):
    a = _r("gg.a", (m, c)).to(torch.bfloat16)
    w = _r("gg.w", (8 * c, c), c ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("gg.b", (8 * c,))
    h = a.float() @ w.float().t() + bias
    val, gate = h.chunk(2, dim=-1)
    ref = val * port.gelu_erf(gate)
    d = cuda_device
    wi, bi = ops.interleave_geglu(w.to(d), bias.to(d))
    out = ops.linear_geglu(a.to(d), wi, bi)
    assert out.shape == (m, 4 * c)
    assert rel_l2(out.float(), ref) < 5e-3


@pytest.mark.parametrize("m,c", [(256, 320), (200, 640), (77, 1280), (9500, 1280)])   # the last one: wide tiles
def test_linear_layernorm_fold(ops, cuda_device, m, c):
    """LayerNorm folded around the GEMMs (attention.py:269-283): the producer GEMM emits per-row sum / sum of squares
    of the bf16 values it stores, the consumer applies rstd * (acc - mean * colsum) + t.  Checked against
    torch.layer_norm + matmul in fp32 on the SAME bf16 activations."""
    d = cuda_device
    a0 = _r("lnf.a0", (m, c)).to(torch.bfloat16)
    w0 = _r("lnf.w0", (c, c), c ** -0.5).to(torch.bfloat16)
    b0 = 0.5 + 0.2 * _r("lnf.b0", (c,))                                  # a non-zero row mean on purpose
    res = _r("lnf.res", (m, c)).to(torch.bfloat16)
    x, stats = ops.linear_stats(a0.to(d), w0.to(d), b0.to(d), res.to(d))
    assert stats.shape == (m, ops.stats_parts(c), 2)
    xf = x.float().cpu()
    assert rel_l2(x, a0.float() @ w0.float().t() + b0 + res.float()) < 5e-3
    assert torch.allclose(stats[..., 0].sum(1).cpu(), xf.sum(-1), rtol=1e-4, atol=1e-2)
    assert torch.allclose(stats[..., 1].sum(1).cpu(), (xf * xf).sum(-1), rtol=1e-4, atol=1e-2)

    gamma, beta, eps = 1.0 + 0.3 * _r("lnf.g", (c,)), 0.2 * _r("lnf.be", (c,)), 1e-5
    ln = torch.nn.functional.layer_norm(xf, (c,), gamma, beta, eps)
    # plain consumer (to_q / to_k / to_v)
    w = _r("lnf.w", (3 * c, c), c ** -0.5)
    wg = (w * gamma[None, :]).to(torch.bfloat16)
    out = ops.linear_ln(x, stats, wg.to(d), wg.float().sum(1).to(d), (w @ beta).to(d), eps)
    assert rel_l2(out.float(), ln @ w.t()) < 6e-3
    # gated consumer (GEGLU projection)
    w2, bias2 = _r("lnf.w2", (8 * c, c), c ** -0.5), 0.1 * _r("lnf.b2", (8 * c,))
    wg2 = (w2 * gamma[None, :]).to(torch.bfloat16)
    wi, ti = ops.interleave_geglu(wg2.to(d), (w2 @ beta + bias2).to(d))
    out2 = ops.linear_ln_geglu(x, stats, wi, wi.float().sum(1).contiguous(), ti, eps)
    val, gate = (ln @ w2.t() + bias2).chunk(2, dim=-1)
    assert rel_l2(out2.float(), val * port.gelu_erf(gate)) < 8e-3


@pytest.mark.parametrize("b,c,hw", [(2, 320, 16), (3, 640, 8), (2, 1280, 8)])
def test_linear_nchw(ops, cuda_device, b, c, hw):
    n_tok = hw * hw
    a = _r("gn.a", (b * n_tok, c)).to(torch.bfloat16)
    w = _r("gn.w", (c, c), c ** -0.5).to(torch.bfloat16)
    bias = 0.1 * _r("gn.bias", (c,))
    x_in = _r("gn.xin", (b, c, hw, hw))
    ref = (a.float() @ w.float().t() + bias).reshape(b, n_tok, c).transpose(1, 2).reshape(b, c, hw, hw) + x_in
    d = cuda_device
    out = ops.linear_nchw(a.to(d), w.to(d), bias.to(d), x_in.to(d))
    assert rel_l2(out, ref) < 1e-4


# ---------------------------------------------------------------------------------------------- attention
def _attn_ref(q, k, v, heads, scale, mask=None):
    b, nq, c = q.shape
    d = c // heads
    sim = port.attention_scores(q, k, heads, scale)
    if mask is not None:
        sim = sim.masked_fill(~mask.bool()[:, None, None, :], -torch.finfo(sim.dtype).max)
    attn = sim.softmax(dim=-1)
    vh = v.reshape(b, -1, heads, d).permute(0, 2, 1, 3)
    return torch.matmul(attn, vh).permute(0, 2, 1, 3).reshape(b, nq, c)


@pytest.mark.parametrize("c,nq,nk", [(320, 256, 256), (320, 1024, 1024), (320, 256, 77), (640, 256, 256),
                                     (640, 64, 77), (1280, 64, 64), (1280, 256, 77), (320, 144, 144), (320, 4096, 4096)])
def test_attention_matches_oracle(ops, cuda_device, c, nq, nk):
    b, heads = 2, 8
    q = _r("at.q", (b, nq, c)).to(torch.bfloat16)
    k = _r("at.k", (b, nk, c)).to(torch.bfloat16)
    v = _r("at.v", (b, nk, c)).to(torch.bfloat16)
    scale = (c // heads) ** -0.5
    ref = _attn_ref(q.float(), k.float(), v.float(), heads, scale)
    d = cuda_device
    out = ops.attention(q.to(d), k.to(d), v.to(d), heads, scale, None)
    assert rel_l2(out.float(), ref) < BF16_TOL


def test_attention_fused_qkv_views_and_sharp_softmax(ops, cuda_device):
    """q/k/v as column slices of one [B,N,3C] buffer; large logits exercise the lazy-rescale path."""
    b, n, c, heads = 2, 512, 320, 8
    qkv = _r("at.qkv", (b, n, 3 * c), 3.0).to(torch.bfloat16)
    q, k, v = qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]
    scale = (c // heads) ** -0.5
    ref = _attn_ref(q.float(), k.float(), v.float(), heads, scale)
    qd = qkv.to(cuda_device)
    out = ops.attention(qd[..., :c], qd[..., c:2 * c], qd[..., 2 * c:], heads, scale, None)
    assert rel_l2(out.float(), ref) < BF16_TOL


@pytest.mark.parametrize("c,n", [(320, 256), (640, 256), (1280, 64)])
def test_attention_key_mask(ops, cuda_device, c, n):
    b, heads = 2, 8
    q = _r("am.q", (b, n, c)).to(torch.bfloat16)
    k = _r("am.k", (b, n, c)).to(torch.bfloat16)
    v = _r("am.v", (b, n, c)).to(torch.bfloat16)
    mask = (_r("am.m", (b, n)) > -0.3)
    mask[1, :] = False                                   # a fully masked sample: uniform average, like the reference
    scale = (c // heads) ** -0.5
    ref = _attn_ref(q.float(), k.float(), v.float(), heads, scale, mask)
    d = cuda_device
    out = ops.attention(q.to(d), k.to(d), v.to(d), heads, scale, mask.to(d))
    assert rel_l2(out.float(), ref) < BF16_TOL


@pytest.mark.parametrize("c,nq,nk,masked", [(320, 4096, 77, False), (320, 200, 1, False), (320, 129, 128, False),
                                            (640, 1024, 77, False), (640, 130, 65, True), (1280, 256, 77, False),
                                            (1280, 70, 128, True), (320, 300, 77, True)])
def test_attention_short_key_sequence(ops, cuda_device, c, nq, nk, masked):
    """kv_len <= 128 takes the one-query-tile configuration (several CTAs per SM, 128/256 TMEM columns each): ragged
    query lengths, 1..128 keys (the partial second key tile is masked inside the kernel), optional key mask."""
    b, heads = 3, 8
    q = _r("as.q", (b, nq, c)).to(torch.bfloat16)
    k = _r("as.k", (b, nk, c)).to(torch.bfloat16)
    v = _r("as.v", (b, nk, c)).to(torch.bfloat16)
    mask = None
    if masked:
        mask = (_r("as.m", (b, nk)) > -0.5)
        mask[:, 0] = True
    scale = (c // heads) ** -0.5
    ref = _attn_ref(q.float(), k.float(), v.float(), heads, scale, mask)
    d = cuda_device
    out = ops.attention(q.to(d), k.to(d), v.to(d), heads, scale, None if mask is None else mask.to(d))
    assert torch.isfinite(out).all()
    assert rel_l2(out.float(), ref) < BF16_TOL


# ---------------------------------------------------------------------------------------------- K12 / K13
def test_explicit_cross_attention_matches_fused(ops, cuda_device):
    """cross_scores + softmax_pv (the materialised path used by save_attn_vars / conv attention) against the oracle
    and against the fused tcgen05 kernel."""
    d = cuda_device
    for c, n in ((320, 256), (640, 100), (1280, 64)):
        b, t, heads = 2, 77, 8
        q = _r("xa.q", (b, n, c)).to(torch.bfloat16)
        k = _r("xa.k", (b, t, c)).to(torch.bfloat16)
        v = _r("xa.v", (b, t, c)).to(torch.bfloat16)
        scale = (c // heads) ** -0.5
        sim = ops.cross_scores(q.to(d), k.to(d), heads, scale)
        ref_sim = port.attention_scores(q.float(), k.float(), heads, scale)
        assert sim.shape == (b, heads, n, t) and rel_l2(sim, ref_sim) < 1e-5
        out, attn = ops.softmax_pv(sim, v.to(d), True)
        ref_attn = ref_sim.softmax(dim=-1)
        assert rel_l2(attn, ref_attn) < 1e-5
        vh = v.float().reshape(b, t, heads, c // heads).permute(0, 2, 1, 3)
        ref_out = torch.matmul(ref_attn, vh).permute(0, 2, 1, 3).reshape(b, n, c)
        assert rel_l2(out.float(), ref_out) < 4e-3
        fused = ops.attention(q.to(d), k.to(d), v.to(d), heads, scale, None)
        assert rel_l2(out.float(), fused.float()) < 6e-3
        out2, empty = ops.softmax_pv(sim, v.to(d), False)
        assert empty.numel() == 0 and torch.equal(out2, out)


@pytest.mark.parametrize("ks", [2, 3, 4])
def test_conv_attn_rows_vs_reference_golden(ops, cuda_device, ks):
    """replace_rows_by_conv_attn (ldm/util.py:701-879) against the output of the reference function itself."""
    from conftest import load_golden
    name = "conv_attn_ks4" if ks == 4 else "conv_attn"       # ks = 4: its own fixture (16 subject tokens, util.py:754)
    g = load_golden(name)
    heads, dh, hw, b, t, seed = [int(v) for v in g["meta"]]
    n = hw * hw
    q = P.randn(f"{name}.q", (b * heads, n, dh), seed)
    k = P.randn(f"{name}.k", (b * heads, t, dh), seed)
    sim = P.randn(f"{name}.sim", (b * heads, n, t), seed)
    inst = torch.tensor([0, 2], dtype=torch.int32)
    tok = torch.tensor([list(range(6, 6 + ks * ks))] * 2, dtype=torch.int32)
    d = cuda_device
    tokens = lambda x, l: x.reshape(b, heads, l, dh).permute(0, 2, 1, 3).reshape(b, l, heads * dh).to(torch.bfloat16).to(d)
    s = sim.reshape(b, heads, n, t).to(d).contiguous()
    ops.conv_attn_rows(s, tokens(q, n), tokens(k, t), inst.to(d), tok.to(d), hw, hw, ks, dh ** -0.5)
    ref = torch.as_tensor(g[f"out_ks{ks}"]).reshape(b, heads, n, t)
    s = s.cpu()
    replaced = torch.zeros(b, t, dtype=torch.bool)
    replaced[inst.long()[:, None], tok.long()] = True
    keep = ~replaced[:, None, None, :].expand_as(ref)
    assert torch.equal(s[keep], ref[keep])                      # untouched columns / samples are untouched
    assert rel_l2(s[~keep], ref[~keep]) < 6e-3                  # q, k rounded to bf16


@pytest.mark.parametrize("ca_idx,n", [(0, 256), (2, 64), (4, 64), (0, 4096)])
def test_ada_pool_and_head(ops, cuda_device, ca_idx, n):
    seed, b, K, r = 15, 2, 9, 12
    sd = P.init_ada_embedding(seed, K, r, 2)
    c = P.CA_INFEAT_DIMS[ca_idx]
    x = P.randn(f"adag.{ca_idx}.x", (b, n, c), seed).to(torch.bfloat16)
    q = P.randn(f"adag.{ca_idx}.q", (b, n, c), seed, 2.0).to(torch.bfloat16)
    to_k = P.draw(f"adag.{ca_idx}.to_k", (c, 768), seed)
    time_emb = P.randn("adag.time_emb", (b, 1280), seed)
    probe = P.randn(f"adag.{ca_idx}.probe", (768,), seed, 0.5)
    extra = P.randn(f"adag.{ca_idx}.extra", (b, 768), seed, 0.5)
    ref_out, ref_pool = port.ada_embedding(sd, ca_idx, x.float(), q.float(), to_k, time_emb, probe, extra, K, r)
    d = cuda_device
    g = lambda key: sd[key].to(d)
    pool_q = ops.ada_pool_prep(to_k.to(d), probe.to(d), extra.to(d), g(f"poolers.{ca_idx}.lora_to_fg_q.weight").squeeze(-1),
                               g(f"poolers.{ca_idx}.lora_to_bg_q.weight").squeeze(-1),
                               g(f"poolers.{ca_idx}.lora_to_k.weight").squeeze(-1))
    fg, bg, a_fg, a_bg = ops.ada_pool(x.to(d), q.to(d), pool_q, None, True)
    assert rel_l2(fg, ref_pool["fg_out"]) < 2e-3
    assert rel_l2(bg, ref_pool["bg_out"]) < 2e-3
    assert rel_l2(a_fg, ref_pool["attn_fg"]) < 2e-3
    assert rel_l2(a_bg, ref_pool["attn_bg"]) < 2e-3
    basis = torch.cat([sd["pre_vecs"], sd["basis_vecs"]], dim=1).to(d)
    out = ops.ada_head(fg, bg, time_emb.to(d), int(0.25 * c), g(f"layer_coeff_maps.{ca_idx}.weight"),
                       g(f"layer_coeff_maps.{ca_idx}.bias"), basis, sd["bias"][ca_idx].to(d))
    assert rel_l2(out, ref_out) < 5e-3
    # splice: the head kernel writes the K vectors as bf16 into the given rows of a [B, 77, 768] context
    ctx = torch.zeros((b + 1, 77, 768), dtype=torch.bfloat16, device=d)
    pos = torch.arange(5, 5 + K, dtype=torch.int32, device=d)
    out2 = ops.ada_head_splice(fg, bg, time_emb.to(d), int(0.25 * c), g(f"layer_coeff_maps.{ca_idx}.weight"),
                               g(f"layer_coeff_maps.{ca_idx}.bias"), basis, sd["bias"][ca_idx].to(d), ctx, pos)
    assert torch.equal(out2, out)
    assert torch.equal(ctx[:b, 5:5 + K], out.to(torch.bfloat16))
    assert float(ctx[:b, :5].abs().max()) == 0 and float(ctx[b].abs().max()) == 0 and float(ctx[:b, 5 + K:].abs().max()) == 0


@pytest.mark.parametrize("ca_idx,n,masked", [(0, 256, False), (4, 64, True), (0, 4096, False)])
def test_ada_head_pooled_matches_separate_launches(ops, cuda_device, ca_idx, n, masked):
    """afb_ada_head_pooled (the head merges the pooling partials in its prologue; opt-in through AFB_ADA_FUSE=1) against
    the default chain afb_ada_pool (with its combine launch) + afb_ada_head: same partials, same merge order in fp32, so
    the pooled features agree to rounding and the K vectors to 1e-5; the cached-bg variant feeds the given bg stream."""
    seed, b, K, r = 16, 2, 9, 12
    sd = P.init_ada_embedding(seed, K, r, 2)
    c = P.CA_INFEAT_DIMS[ca_idx]
    d = cuda_device
    x = P.randn(f"adaf.{ca_idx}.x", (b, n, c), seed).to(torch.bfloat16).to(d)
    q = P.randn(f"adaf.{ca_idx}.q", (b, n, c), seed, 2.0).to(torch.bfloat16).to(d)
    to_k = P.draw(f"adaf.{ca_idx}.to_k", (c, 768), seed).to(d)
    time_emb = P.randn("adaf.time_emb", (b, 1280), seed).to(d)
    probe = P.randn(f"adaf.{ca_idx}.probe", (768,), seed, 0.5).to(d)
    extra = P.randn(f"adaf.{ca_idx}.extra", (b, 768), seed, 0.5).to(d)
    g = lambda key: sd[key].to(d)
    pool_q = ops.ada_pool_prep(to_k, probe, extra, g(f"poolers.{ca_idx}.lora_to_fg_q.weight").squeeze(-1),
                               g(f"poolers.{ca_idx}.lora_to_bg_q.weight").squeeze(-1),
                               g(f"poolers.{ca_idx}.lora_to_k.weight").squeeze(-1))
    mask = None
    if masked:
        mask = (torch.arange(n, device=d)[None, :] % 3 != 1).expand(b, n).contiguous()
    td = int(0.25 * c)
    cw, cb = g(f"layer_coeff_maps.{ca_idx}.weight"), g(f"layer_coeff_maps.{ca_idx}.bias")
    basis = torch.cat([sd["pre_vecs"], sd["basis_vecs"]], dim=1).to(d)
    bias = sd["bias"][ca_idx].to(d)
    pos = torch.arange(5, 5 + K, dtype=torch.int32, device=d)
    fg, bg, _, _ = ops.ada_pool(x, q, pool_q, mask, False)
    for cached in (None, torch.flip(bg, dims=(0,)).contiguous()):
        ctx_a = torch.zeros((b + 1, 77, 768), dtype=torch.bfloat16, device=d)
        ctx_b = torch.zeros_like(ctx_a)
        want = ops.ada_head_splice(fg, bg if cached is None else cached, time_emb, td, cw, cb, basis, bias, ctx_a, pos)
        out, fg2, bg2 = ops.ada_pool_head_splice(x, q, pool_q, mask, cached, time_emb, td, cw, cb, basis, bias, ctx_b, pos)
        assert rel_l2(fg2, fg) < 1e-6 and rel_l2(bg2, bg) < 1e-6
        assert rel_l2(out, want) < 1e-5
        assert rel_l2(ctx_b.float(), ctx_a.float()) < 1e-3        # bf16 rows: at most a last-bit flip on a few elements
        assert float(ctx_b[b].abs().max()) == 0 and float(ctx_b[:b, :5].abs().max()) == 0


def test_ada_pool_prep_batched_equals_per_layer(ops, cuda_device):
    """afb_ada_pool_prep_batched (every layer / subject of a U-Net call in two launches) against the per-layer entry
    point: same arithmetic, so identical results; a second run() recomputes in place."""
    d, b = cuda_device, 3
    items = []
    for c in (320, 640, 1280, 320):
        lg = c // 8
        items.append(tuple(t.to(d) for t in (_r(f"pb.tok{c}", (c, 768), 768 ** -0.5), _r(f"pb.fg{c}", (768,)),
                                               _r(f"pb.bg{c}", (b, 768)), _r(f"pb.wf{c}", (lg, lg), lg ** -0.5),
                                               _r(f"pb.wb{c}", (lg, lg), lg ** -0.5), _r(f"pb.wk{c}", (lg, lg), lg ** -0.5))))
    batch = ops.AdaPrepBatch(items, b)
    batch.run()
    for it, pq in zip(items, batch.pool_q):
        ref = ops.ada_pool_prep(*it)
        assert pq.shape == ref.shape and torch.equal(pq, ref)
    for pq in batch.pool_q:
        pq.zero_()
    batch.run()
    assert all(torch.equal(pq, ops.ada_pool_prep(*it)) for it, pq in zip(items, batch.pool_q))


def test_casts_round_trip(ops, cuda_device):
    x = _r("cast.x", (3, 1001)).to(cuda_device)
    y = ops.to_bf16(x)
    assert torch.equal(y, x.to(torch.bfloat16))
    assert torch.equal(ops.to_f32(y), y.float())


def test_unsupported_arguments_raise(ops, cuda_device):
    from adaface_b200._lib import AdafaceB200Error
    d = cuda_device
    q = torch.zeros((1, 64, 8 * 24), dtype=torch.bfloat16, device=d)
    with pytest.raises(AdafaceB200Error):
        ops.attention(q, q, q, 8, 1.0, None)                     # head_dim 24 is not an SD-1.5 head size
    with pytest.raises(AdafaceB200Error):
        ops.linear(torch.zeros((8, 64), dtype=torch.bfloat16), torch.zeros((32, 64), dtype=torch.bfloat16), None, None, False)


# ---------------------------------------------------------------------------------------------- N2 static embedding
def test_static_layerwise_embedding_vs_reference_golden(ops, cuda_device):
    """afb_static_layerwise_embedding against the reference StaticLayerwiseEmbedding (golden) and, without pre_vecs /
    bias and at other sizes, against the port."""
    from adaface_b200.embedding import StaticLayerwiseEmbedding
    g = load_golden("static_emb")
    seed, K, r, L, _ = (int(v) for v in g["meta"])
    mod = StaticLayerwiseEmbedding(L, K, 768, r, init_vecs=torch.zeros(2, 768))
    mod.load_state_dict(P.init_static_embedding(seed, K, r, 2), strict=True)
    out = mod.to(cuda_device)()
    assert out.shape == (L, K, 768) and out.dtype == torch.float32
    assert rel_l2(out, torch.from_numpy(g["static"])) < 1e-5
    for (L2, K2, r2, D2, n_pre, has_bias) in [(16, 1, 6, 768, 0, False), (4, 3, 5, 1024, 1, True), (25, 16, 12, 768, 3, True)]:
        p = {"basis_rand_weights": _r("se.rw", (L2, K2, r2)), "basis_comm_weights": _r("se.cw", (1, K2, r2)),
             "basis_vecs": _r("se.bv", (K2, r2 - n_pre, D2))}
        if n_pre:
            p["pre_vecs"] = _r("se.pv", (K2, n_pre, D2))
        if has_bias:
            p["bias"] = _r("se.b", (L2, K2, D2), 0.02)
        ref = port.static_layerwise_embedding(p, D2)
        d = cuda_device
        got = ops.static_layerwise_embedding(p["basis_rand_weights"].to(d), p["basis_comm_weights"].to(d),
                                             p["pre_vecs"].to(d) if n_pre else None, p["basis_vecs"].to(d),
                                             p["bias"].to(d) if has_bias else None)
        assert rel_l2(got, ref) < 1e-5, (L2, K2, r2)


@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_splice_static_embedding_vs_reference_golden(ops, cuda_device, out_dtype):
    """first_token_pos + splice_static_embedding: bit-exact copy semantics against the reference's statements (golden)
    and the port, including a prompt without the placeholder, a repeated placeholder and one at the end of the prompt
    whose trailing vectors fall off the 77 positions."""
    from adaface_b200.embedding import splice_static_embedding
    from oracle.gen_golden import static_inputs
    g = load_golden("static_emb")
    seed, K, r, L, token = (int(v) for v in g["meta"])
    static = torch.from_numpy(g["static"])
    emb, tok = static_inputs("static_emb", seed, K=K)
    d = cuda_device
    out, tok_rep = splice_static_embedding(emb.to(d), tok.to(d), token, static.to(d), L, out_dtype)
    assert out.dtype == out_dtype and out.shape == (emb.shape[0] * L, 77, 768) and tok_rep.shape == (emb.shape[0] * L, 77)
    want = torch.from_numpy(g["spliced"])
    assert torch.equal(out[:, :, ::16].cpu(), want.to(out_dtype))
    assert torch.equal(out.cpu(), port.splice_static_embedding(emb, tok, token, static, L).to(out_dtype))
    assert torch.equal(ops.first_token_pos(tok.to(d), token).cpu(), torch.tensor([5, -1, 20], dtype=torch.int32))
    tok2 = tok.clone()
    tok2[1, 77 - 3] = token                                  # only 3 of the K vectors fit
    out2, _ = splice_static_embedding(emb.to(d), tok2.to(d), token, static.to(d), L, out_dtype)
    assert torch.equal(out2[L:2 * L, 74:77].cpu(), static[:, :3].to(out_dtype))
    assert torch.equal(out2[L:2 * L, :74].cpu(), emb[1, :74].expand(L, -1, -1).to(out_dtype))


def test_ada_pool_prep_converted_inputs_stay_alive(ops, cuda_device):
    """a bf16 to_k weight, an expanded 1-D bg query and non-contiguous lora weights all need a converted
    copy; the copies must outlive the launch (a freed temporary could be handed to the next conversion)."""
    d, b, c = cuda_device, 4, 640
    lg = c // 8
    to_k = _r("pp.tok", (c, 768), 768 ** -0.5).to(d)
    fg, bg1 = _r("pp.fg", (768,)).to(d), _r("pp.bg", (768,)).to(d)
    wf, wb, wk = (_r(f"pp.{n}", (lg, lg), lg ** -0.5).to(d) for n in ("wf", "wb", "wk"))
    to_k_bf16 = to_k.to(torch.bfloat16)
    ref = ops.ada_pool_prep(to_k_bf16.float(), fg, bg1[None].expand(b, -1).contiguous(), wf, wb, wk)
    for _ in range(3):          # repeated: the allocator has freed blocks of matching sizes to hand out
        out = ops.ada_pool_prep(to_k_bf16, fg.double(), bg1[None].expand(b, -1), wf.t().contiguous().t(), wb.double(), wk.half().float().t().t())
        ref2 = ops.ada_pool_prep(to_k_bf16.float(), fg, bg1[None].expand(b, -1).contiguous(), wf, wb, wk.half().float())
        assert torch.equal(out, ref2)
    assert rel_l2(ref, ops.ada_pool_prep(to_k, fg, bg1[None].expand(b, -1).contiguous(), wf, wb, wk)) < 5e-3
