"""Tensor-level wrappers of the C ABI (include/adaface_b200.h), registered as torch.library custom ops in
the `adaface_b200::` namespace.  PyTorch is plumbing here: device memory, streams, op registration.

Every op launches hand-written sm_100a kernels on torch's current CUDA stream; none has a CPU or eager
fallback - a non-CUDA tensor or a missing library raises.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib
from ._lib import AttnArgs, GemmArgs, check

Tensor = torch.Tensor


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: Optional[Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.AdafaceB200Error("adaface_b200 ops run on CUDA tensors only (no CPU fallback)")


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# Optional per-launch timing (bench.py's kernel breakdown / roofline): CUDA events on the launching stream
# around every C-ABI call, with the call's algorithmic FLOPs and bytes.  Off (None) on the normal path.
_TIMING: Optional[List[tuple]] = None
# which part of the U-Net a launch belongs to, for the breakdown only: "st" = the SpatialTransformer / AdaEmbedding /
# CFG-DDIM hot path, "resblock" = the kernels the U-Net shell borrows for ResBlocks, Upsample and the output head
_SCOPE = "st"


class scope:
    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        global _SCOPE
        self.prev, _SCOPE = _SCOPE, self.name

    def __exit__(self, *exc):
        global _SCOPE
        _SCOPE = self.prev


def start_timing() -> None:
    global _TIMING
    _TIMING = []


def stop_timing() -> List[Dict]:
    """Returns one record per C-ABI call since start_timing(): name, shape tag, ms, flops, bytes."""
    global _TIMING
    recs, _TIMING = _TIMING or [], None
    torch.cuda.synchronize()
    return [{"name": n, "tag": tag, "ms": e0.elapsed_time(e1), "flops": fl, "bytes": by, "scope": sc}
            for n, tag, fl, by, e0, e1, sc in recs]


def _run(name: str, fn, *args, flops: float = 0.0, nbytes: float = 0.0, tag: str = "") -> None:
    if _TIMING is None:
        check(fn(*args), name)
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(fn(*args), name)
    e1.record()
    _TIMING.append((name, tag, flops, nbytes, e0, e1, _SCOPE))


# ------------------------------------------------------------------------------------------- K14
@torch.library.custom_op("adaface_b200::cfg_ddim_step", mutates_args=())
def cfg_ddim_step(x: Tensor, e_cond: Tensor, e_uncond: Tensor, noise: Optional[Tensor], guidance_scale: float,
                  a_t: float, a_prev: float, sigma_t: float, sqrt_one_minus_at: float,
                  temperature: float) -> Tuple[Tensor, Tensor]:
    """ddim.py:252-296.  fp32; returns (x_prev, pred_x0)."""
    _need_cuda(x, e_cond, e_uncond, noise)
    for t in (x, e_cond, e_uncond, noise):
        if t is not None and (t.dtype != torch.float32 or t.shape != x.shape):
            raise ValueError("cfg_ddim_step: fp32 tensors of one shape expected")
    x, e_cond, e_uncond = x.contiguous(), e_cond.contiguous(), e_uncond.contiguous()
    noise = None if noise is None else noise.contiguous()
    x_prev, pred_x0 = torch.empty_like(x), torch.empty_like(x)
    _run("cfg_ddim_step", _lib.load().afb_cfg_ddim_step, _ptr(x), _ptr(e_cond), _ptr(e_uncond), _ptr(noise),
         _ptr(x_prev), _ptr(pred_x0), x.numel(), guidance_scale, a_t, a_prev, sigma_t, sqrt_one_minus_at, temperature,
         _stream(), nbytes=4.0 * x.numel() * (5 + (noise is not None)))
    return x_prev, pred_x0


@cfg_ddim_step.register_fake
def _(x, e_cond, e_uncond, noise, guidance_scale, a_t, a_prev, sigma_t, sqrt_one_minus_at, temperature):
    return torch.empty_like(x), torch.empty_like(x)


# ------------------------------------------------------------------------------------------- K1 / K3
@torch.library.custom_op("adaface_b200::groupnorm_tokens", mutates_args=())
def groupnorm_tokens(x: Tensor, gamma: Tensor, beta: Tensor, groups: int, eps: float, act: int = 0,
                     pre_add: Optional[Tensor] = None) -> Tensor:
    """GroupNorm over x [B,C,h,w] -> normalised tokens bf16 [B, h*w, C] (attention.py:325-327).
    fp32 NCHW-contiguous and bf16 channels-last inputs are consumed in place; anything else is rejected.
    act = 1 fuses SiLU after the affine (the U-Net ResBlocks' GroupNorm32 -> SiLU); pre_add (fp32 [B, C], rows may be
    strided) is added to x before the statistics (a ResBlock's conv bias + time-embedding add)."""
    _need_cuda(x, gamma, beta)
    b, c, h, w = x.shape
    hw = h * w
    if x.dtype == torch.float32 and x.is_contiguous():
        layout = 0
    elif x.dtype == torch.bfloat16 and x.permute(0, 2, 3, 1).is_contiguous():
        layout = 1
    else:
        raise ValueError("groupnorm_tokens: x must be fp32 NCHW-contiguous or bf16 channels-last")
    lib = _lib.load()
    ws = torch.empty(int(lib.afb_groupnorm_ws_floats(b, hw, groups)), dtype=torch.float32, device=x.device)
    y = torch.empty((b, hw, c), dtype=torch.bfloat16, device=x.device)
    add_ld = 0
    if pre_add is not None:
        if pre_add.dtype != torch.float32 or pre_add.shape != (b, c) or pre_add.stride(1) != 1:
            raise ValueError("groupnorm_tokens: pre_add must be fp32 [B, C] with contiguous rows")
        add_ld = pre_add.stride(0)
    _run("groupnorm_tokens", lib.afb_groupnorm_tokens, _ptr(x), layout, _ptr(gamma), _ptr(beta), _ptr(y), _ptr(ws), b, c,
         hw, groups, eps, act, _ptr(pre_add), add_ld, _stream(), nbytes=float(x.numel()) * (x.element_size() + 2), tag=f"b{b}n{hw}C{c}")
    return y


@groupnorm_tokens.register_fake
def _(x, gamma, beta, groups, eps, act=0, pre_add=None):
    b, c, h, w = x.shape
    return x.new_empty((b, h * w, c), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::groupnorm_tokens_cat", mutates_args=())
def groupnorm_tokens_cat(x1: Tensor, x2: Tensor, gamma: Tensor, beta: Tensor, groups: int, eps: float, act: int = 0) -> Tensor:
    """GroupNorm (+ SiLU) of th.cat([x1, x2], dim=1) without materialising the concatenation (decoder ResBlocks,
    openaimodel.py:1004-1010): x1 [B,C1,h,w], x2 [B,C2,h,w] bf16 channels-last -> bf16 tokens [B, h*w, C1+C2]."""
    _need_cuda(x1, x2, gamma, beta)
    b, c1, h, w = x1.shape
    c2 = x2.shape[1]
    if x2.shape != (b, c2, h, w) or x1.dtype != torch.bfloat16 or x2.dtype != torch.bfloat16 \
            or not x1.permute(0, 2, 3, 1).is_contiguous() or not x2.permute(0, 2, 3, 1).is_contiguous():
        raise ValueError("groupnorm_tokens_cat: two bf16 channels-last tensors of the same batch and size expected")
    hw, c = h * w, c1 + c2
    lib = _lib.load()
    ws = torch.empty(int(lib.afb_groupnorm_ws_floats(b, hw, groups)), dtype=torch.float32, device=x1.device)
    y = torch.empty((b, hw, c), dtype=torch.bfloat16, device=x1.device)
    _run("groupnorm_tokens", lib.afb_groupnorm_tokens_cat, _ptr(x1), c1, _ptr(x2), 1, _ptr(gamma), _ptr(beta), _ptr(y),
         _ptr(ws), b, c, hw, groups, eps, act, None, 0, _stream(), nbytes=float(b * hw * c) * 4, tag=f"b{b}n{hw}C{c1}+{c2}")
    return y


@groupnorm_tokens_cat.register_fake
def _(x1, x2, gamma, beta, groups, eps, act=0):
    b, c1, h, w = x1.shape
    return x1.new_empty((b, h * w, c1 + x2.shape[1]), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::layernorm", mutates_args=())
def layernorm(x: Tensor, gamma: Optional[Tensor], beta: Optional[Tensor], eps: float) -> Tensor:
    """LayerNorm over the last dim of a contiguous bf16 tensor (attention.py:267-269)."""
    _need_cuda(x, gamma, beta)
    if x.dtype != torch.bfloat16 or not x.is_contiguous():
        raise ValueError("layernorm: contiguous bf16 input expected")
    y = torch.empty_like(x)
    c = x.shape[-1]
    _run("layernorm", _lib.load().afb_layernorm_bf16, _ptr(x), _ptr(gamma), _ptr(beta), _ptr(y), x.numel() // c, c, eps,
         _stream(), nbytes=4.0 * x.numel(), tag=f"C{c}")
    return y


@layernorm.register_fake
def _(x, gamma, beta, eps):
    return torch.empty_like(x)


# ------------------------------------------------------------------------------------------- GEMMs
def _gemm(a: Tensor, w: Tensor, out: Tensor, bias: Optional[Tensor], residual: Optional[Tensor], epilogue: int,
          ldo: int, ldr: int, hw: int = 0, stats_out: Optional[Tensor] = None, ln_stats: Optional[Tensor] = None,
          ln_colsum: Optional[Tensor] = None, ln_eps: float = 0.0, conv_hw: Optional[Tuple[int, int]] = None,
          conv_fold: bool = False, a2: Optional[Tensor] = None, conv_stride: int = 1) -> None:
    _need_cuda(a, w, out, bias, residual, stats_out, ln_stats, ln_colsum)
    if ln_stats is not None:
        if ln_colsum is None or ln_stats.dtype != torch.float32 or ln_colsum.dtype != torch.float32 \
                or not ln_stats.is_contiguous() or ln_stats.dim() != 3 or ln_stats.shape[0] != a.shape[0] \
                or ln_stats.shape[2] != 2 or ln_colsum.numel() != w.shape[0]:
            raise ValueError("gemm: ln_stats must be fp32 [M, parts, 2] and ln_colsum fp32 [N]")
    if a.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or a.stride(-1) != 1 or w.stride(-1) != 1:
        raise ValueError("gemm: bf16 operands with contiguous K expected")
    if bias is not None and bias.dtype != torch.float32:
        raise ValueError("gemm: bias must be fp32")
    if a2 is not None and (a2.dtype != torch.bfloat16 or a2.stride(-1) != 1 or a2.shape[0] != a.shape[0]):
        raise ValueError("gemm: a2 must be bf16 [M, K2] with contiguous rows")
    args = GemmArgs(a=_ptr(a), w=_ptr(w), out=_ptr(out), bias=_ptr(bias), residual=_ptr(residual),
                    a2=_ptr(a2), k_split=a.shape[1] if a2 is not None else 0, lda2=a2.stride(0) if a2 is not None else 0,
                    m=a.shape[0], n=w.shape[0], k=w.shape[1] if (conv_hw or a2 is not None) else a.shape[1], lda=a.stride(0), ldw=w.stride(0),
                    ldo=ldo, ldr=ldr, conv_h=conv_hw[0] if conv_hw else 0, conv_w=conv_hw[1] if conv_hw else 0, conv_fold=1 if conv_fold else 0,
                    conv_stride=conv_stride, epilogue=epilogue, out_dtype=_lib.DT_F32 if out.dtype == torch.float32 else _lib.DT_BF16,
                    residual_dtype=_lib.DT_F32 if (residual is not None and residual.dtype == torch.float32)
                    else _lib.DT_BF16, hw=hw, row_stats_out=_ptr(stats_out), ln_stats=_ptr(ln_stats),
                    ln_colsum=_ptr(ln_colsum), ln_parts=0 if ln_stats is None else ln_stats.shape[1], ln_eps=ln_eps)
    if conv_stride == 2:
        args.m = a.shape[0] // 4              # output pixels; `a` holds the twice-as-large input image
    m, n, k = int(args.m), w.shape[0], w.shape[1] if (conv_hw or a2 is not None) else a.shape[1]
    _run(("conv3x3_up2_bf16" if conv_fold else "conv3x3_bf16") if conv_hw else "gemm_bf16", _lib.load().afb_gemm_bf16, C.byref(args), _stream(), flops=2.0 * m * n * k,
         nbytes=2.0 * (m * (k if a2 is not None else a.shape[1]) + n * k) + out.numel() * out.element_size(), tag=f"{m}x{n}x{k}e{epilogue}" + ("c" if a2 is not None else "") + ("s" if stats_out is not None else "") + ("l" if ln_stats is not None else "")
         + ("r" if residual is not None else ""))


@torch.library.custom_op("adaface_b200::conv3x3", mutates_args=())
def conv3x3(x: Tensor, w: Tensor, bias: Optional[Tensor], residual: Optional[Tensor], h: int, width: int) -> Tensor:
    """3x3 / stride 1 / zero-padding 1 convolution as an implicit GEMM on the tcgen05 GEMM kernel (ResBlock and
    Upsample convolutions, openaimodel.py:95-123,203-205,236-238).  x: bf16 NHWC pixels [B*h*width, Cin];
    w: bf16 [Cout, 9*Cin] = the channels-last memory of the Conv2d weight (k = (ky*3 + kx)*Cin + c); bias fp32
    [Cout]; residual bf16 [B*h*width, Cout].  Returns bf16 [B*h*width, Cout] (NHWC)."""
    if x.dim() != 2 or w.dim() != 2 or w.shape[1] != 9 * x.shape[1] or x.shape[0] % (h * width) != 0:
        raise ValueError("conv3x3: x [B*h*w, Cin] and w [Cout, 9*Cin] expected")
    out = torch.empty((x.shape[0], w.shape[0]), dtype=torch.bfloat16, device=x.device)
    if residual is not None and (residual.shape != out.shape or residual.stride(-1) != 1 or residual.dtype != torch.bfloat16):
        raise ValueError("conv3x3: residual must be bf16 [B*h*w, Cout] with contiguous rows")
    _gemm(x, w, out, bias, residual, _lib.EPI_STORE, out.stride(0), 0 if residual is None else residual.stride(0),
          conv_hw=(h, width))
    return out


@conv3x3.register_fake
def _(x, w, bias, residual, h, width):
    return x.new_empty((x.shape[0], w.shape[0]), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::conv3x3_s2", mutates_args=())
def conv3x3_s2(x: Tensor, w: Tensor, bias: Optional[Tensor], h: int, width: int) -> Tensor:
    """3x3 / stride 2 / zero-padding 1 convolution (Downsample.forward, openaimodel.py:138-165) on the same implicit-GEMM
    kernel: TMA reads tap (dy, dx) of output pixel (i, j) at input pixel (2i + dy, 2j + dx) with an element stride of 2.
    x: bf16 NHWC pixels [B*h*width, Cin] of the INPUT image (h, width even); returns bf16 [B*(h/2)*(width/2), Cout]."""
    if x.dim() != 2 or w.dim() != 2 or w.shape[1] != 9 * x.shape[1] or x.shape[0] % (h * width) != 0 or h % 2 or width % 2:
        raise ValueError("conv3x3_s2: x [B*h*w, Cin] with even h, w and w [Cout, 9*Cin] expected")
    out = torch.empty((x.shape[0] // 4, w.shape[0]), dtype=torch.bfloat16, device=x.device)
    _gemm(x, w, out, bias, None, _lib.EPI_STORE, out.stride(0), 0, conv_hw=(h // 2, width // 2), conv_stride=2)
    return out


@conv3x3_s2.register_fake
def _(x, w, bias, h, width):
    return x.new_empty((x.shape[0] // 4, w.shape[0]), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::conv3x3_c4", mutates_args=())
def conv3x3_c4(x: Tensor, w64: Tensor, bias: Optional[Tensor], h: int, width: int) -> Tensor:
    """Conv2d(4, Cout, 3, padding=1) on the latent (the U-Net's first convolution, openaimodel.py:575-579): an im2col
    into 64-wide rows, then the tcgen05 GEMM.  x: bf16 NHWC pixels [B*h*width, 4]; w64: bf16 [Cout, 64] with column
    (ky*3 + kx)*4 + c and zeros from 36 on; bias fp32 [Cout].  Returns bf16 [B*h*width, Cout] (NHWC)."""
    _need_cuda(x, w64, bias)
    if x.dim() != 2 or x.shape[1] != 4 or x.dtype != torch.bfloat16 or not x.is_contiguous() or w64.shape[1] != 64 \
            or x.shape[0] % (h * width) != 0:
        raise ValueError("conv3x3_c4: contiguous bf16 x [B*h*w, 4] and w64 [Cout, 64] expected")
    cols = torch.empty((x.shape[0], 64), dtype=torch.bfloat16, device=x.device)
    _run("im2col3x3_c4", _lib.load().afb_im2col3x3_c4, _ptr(x), _ptr(cols), x.shape[0] // (h * width), h, width, _stream(),
         nbytes=float(x.numel() * 2 * 9 + cols.numel() * 2), tag=f"b{x.shape[0] // (h * width)}h{h}w{width}")
    out = torch.empty((x.shape[0], w64.shape[0]), dtype=torch.bfloat16, device=x.device)
    _gemm(cols, w64, out, bias, None, _lib.EPI_STORE, out.stride(0), 0)
    return out


@conv3x3_c4.register_fake
def _(x, w64, bias, h, width):
    return x.new_empty((x.shape[0], w64.shape[0]), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::conv3x3_few_out", mutates_args=())
def conv3x3_few_out(x: Tensor, w8: Tensor, bias: Optional[Tensor], cout: int, h: int, width: int) -> Tensor:
    """Conv2d(C, cout <= 8, 3, padding=1) (the U-Net's last convolution, openaimodel.py:716-720).  x: bf16 NHWC pixels
    [B*h*width, C], C % 64 == 0; w8: bf16 [8, 9*C] with column (ky*3 + kx)*C + c, rows >= cout zero; bias fp32 [cout].
    Returns fp32 NCHW [B, cout, h, width]."""
    _need_cuda(x, w8, bias)
    if x.dim() != 2 or x.dtype != torch.bfloat16 or not x.is_contiguous() or w8.shape != (8, 9 * x.shape[1]) \
            or w8.dtype != torch.bfloat16 or not w8.is_contiguous() or x.shape[0] % (h * width) != 0:
        raise ValueError("conv3x3_few_out: contiguous bf16 x [B*h*w, C] and w8 [8, 9*C] expected")
    b = x.shape[0] // (h * width)
    out = torch.empty((b, cout, h, width), dtype=torch.float32, device=x.device)
    _run("conv3x3_few_out", _lib.load().afb_conv3x3_few_out, _ptr(x), _ptr(w8), _ptr(bias), _ptr(out), b, h, width, x.shape[1],
         cout, _stream(), flops=2.0 * x.shape[0] * cout * 9 * x.shape[1], nbytes=float(x.numel() * 2 + out.numel() * 4),
         tag=f"b{b}h{h}w{width}C{x.shape[1]}")
    return out


@conv3x3_few_out.register_fake
def _(x, w8, bias, cout, h, width):
    return x.new_empty((x.shape[0] // (h * width), cout, h, width), dtype=torch.float32)


def fold_upsample_weight(weight: Tensor) -> Tensor:
    """Conv2d weight [Cout, Cin, 3, 3] -> [4*Cout, 4*Cin] fp32 for `conv3x3_up2`: nearest-2x upsampling followed by the
    3x3 convolution equals, for output phase (py, px), a 2x2 convolution of the low-resolution image whose tap (a, c)
    (low-res offset (py-1+a, px-1+c)) is the sum of the 3x3 taps that land on that pixel:
    py = 0: a=0 <- ky {0}, a=1 <- ky {1, 2};  py = 1: a=0 <- ky {0, 1}, a=1 <- ky {2}  (same for px / kx)."""
    cout, cin = weight.shape[:2]
    w = weight.detach().float()
    sel = {0: ([0], [1, 2]), 1: ([0, 1], [2])}
    out = w.new_zeros((2, 2, cout, 2, 2, cin))
    for py in (0, 1):
        for px in (0, 1):
            for a in (0, 1):
                for c in (0, 1):
                    out[py, px, :, a, c, :] = w[:, :, sel[py][a], :][:, :, :, sel[px][c]].sum(dim=(2, 3))
    return out.reshape(4 * cout, 4 * cin)


@torch.library.custom_op("adaface_b200::conv3x3_up2", mutates_args=())
def conv3x3_up2(x: Tensor, w_folded: Tensor, bias4: Optional[Tensor], h: int, width: int) -> Tensor:
    """Upsample.forward (openaimodel.py:95-123): conv3x3(interpolate(x, 2x nearest)) + bias without materialising the
    upsampled image.  x: bf16 NHWC pixels [B*h*width, Cin] of the LOW-resolution image; w_folded bf16 [4*Cout, 4*Cin]
    (`fold_upsample_weight`); bias4 fp32 [4*Cout] (the bias repeated per phase).  Returns bf16 [B*2h*2width, Cout]."""
    if x.dim() != 2 or w_folded.dim() != 2 or w_folded.shape[1] != 4 * x.shape[1] or w_folded.shape[0] % 4 \
            or x.shape[0] % (h * width) != 0:
        raise ValueError("conv3x3_up2: x [B*h*w, Cin] and w_folded [4*Cout, 4*Cin] expected")
    cout = w_folded.shape[0] // 4
    out = torch.empty((4 * x.shape[0], cout), dtype=torch.bfloat16, device=x.device)
    _gemm(x, w_folded, out, bias4, None, _lib.EPI_STORE, cout, 0, conv_hw=(h, width), conv_fold=True)
    return out


@conv3x3_up2.register_fake
def _(x, w_folded, bias4, h, width):
    return x.new_empty((4 * x.shape[0], w_folded.shape[0] // 4), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::linear", mutates_args=())
def linear(a: Tensor, w: Tensor, bias: Optional[Tensor], residual: Optional[Tensor], out_fp32: bool) -> Tensor:
    """out[M,N] = a[M,K] @ w[N,K]^T + bias (+ residual[M,N]).  bf16 operands, fp32 accumulate."""
    if a.dim() != 2 or w.dim() != 2 or a.shape[1] != w.shape[1]:
        raise ValueError("linear: a [M,K] and w [N,K] expected")
    out = torch.empty((a.shape[0], w.shape[0]), dtype=torch.float32 if out_fp32 else torch.bfloat16, device=a.device)
    if residual is not None and (residual.shape != out.shape or residual.stride(-1) != 1):
        raise ValueError("linear: residual must be [M,N] with contiguous rows")
    _gemm(a, w, out, bias, residual, _lib.EPI_STORE, out.stride(0), 0 if residual is None else residual.stride(0))
    return out


@linear.register_fake
def _(a, w, bias, residual, out_fp32):
    return a.new_empty((a.shape[0], w.shape[0]), dtype=torch.float32 if out_fp32 else torch.bfloat16)


@torch.library.custom_op("adaface_b200::linear_cat", mutates_args=())
def linear_cat(a1: Tensor, a2: Tensor, w: Tensor, bias: Optional[Tensor], residual: Optional[Tensor]) -> Tensor:
    """out[M,N] = cat([a1, a2], 1) @ w[N, K1+K2]^T + bias (+ residual) without materialising the concatenation: the
    GEMM's K loop reads its first K1 columns from a1 and the rest from a2 (K1 % 64 == 0).  bf16 in / bf16 out."""
    if a1.dim() != 2 or a2.dim() != 2 or w.dim() != 2 or a1.shape[1] + a2.shape[1] != w.shape[1] or a1.shape[1] % 64:
        raise ValueError("linear_cat: a1 [M,K1], a2 [M,K2], w [N,K1+K2] with K1 %% 64 == 0 expected")
    out = torch.empty((a1.shape[0], w.shape[0]), dtype=torch.bfloat16, device=a1.device)
    if residual is not None and (residual.shape != out.shape or residual.stride(-1) != 1 or residual.dtype != torch.bfloat16):
        raise ValueError("linear_cat: residual must be bf16 [M,N] with contiguous rows")
    _gemm(a1, w, out, bias, residual, _lib.EPI_STORE, out.stride(0), 0 if residual is None else residual.stride(0), a2=a2)
    return out


@linear_cat.register_fake
def _(a1, a2, w, bias, residual):
    return a1.new_empty((a1.shape[0], w.shape[0]), dtype=torch.bfloat16)


def stats_parts(n: int) -> int:
    """Partials per row that `linear_stats` emits for n output columns (afb_gemm_stats_parts)."""
    return int(_lib.load().afb_gemm_stats_parts(int(n)))


@torch.library.custom_op("adaface_b200::linear_stats", mutates_args=())
def linear_stats(a: Tensor, w: Tensor, bias: Optional[Tensor], residual: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """`linear` with bf16 output that also returns the per-row (sum, sum of squares) partials of what it stored,
    fp32 [M, parts, 2]: the LayerNorm statistics of the output, for a consumer GEMM with the LayerNorm folded in."""
    out = torch.empty((a.shape[0], w.shape[0]), dtype=torch.bfloat16, device=a.device)
    if residual is not None and (residual.shape != out.shape or residual.stride(-1) != 1 or residual.dtype != torch.bfloat16):
        raise ValueError("linear_stats: residual must be bf16 [M,N] with contiguous rows")
    stats = torch.empty((a.shape[0], stats_parts(w.shape[0]), 2), dtype=torch.float32, device=a.device)
    _gemm(a, w, out, bias, residual, _lib.EPI_STORE, out.stride(0), 0 if residual is None else residual.stride(0),
          stats_out=stats)
    return out, stats


@linear_stats.register_fake
def _(a, w, bias, residual):
    n = w.shape[0]
    parts = stats_parts(n)
    return a.new_empty((a.shape[0], n), dtype=torch.bfloat16), a.new_empty((a.shape[0], parts, 2), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::linear_ln", mutates_args=())
def linear_ln(a: Tensor, stats: Tensor, w_gamma: Tensor, colsum: Tensor, bias: Tensor, eps: float) -> Tensor:
    """LayerNorm(a) @ W^T + b with the LayerNorm folded into the GEMM: w_gamma = W * gamma (bf16), colsum[n] =
    sum_k w_gamma[n,k], bias[n] = sum_k W[n,k] * beta[k] + b[n]; `stats` are a's row partials from `linear_stats`.
    out = rstd * (a @ w_gamma^T - mean * colsum) + bias (attention.py:269-283 + :157-159)."""
    out = torch.empty((a.shape[0], w_gamma.shape[0]), dtype=torch.bfloat16, device=a.device)
    _gemm(a, w_gamma, out, bias, None, _lib.EPI_STORE, out.stride(0), 0, ln_stats=stats, ln_colsum=colsum, ln_eps=eps)
    return out


@linear_ln.register_fake
def _(a, stats, w_gamma, colsum, bias, eps):
    return a.new_empty((a.shape[0], w_gamma.shape[0]), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::linear_ln_geglu", mutates_args=())
def linear_ln_geglu(a: Tensor, stats: Tensor, w_interleaved: Tensor, colsum: Tensor, bias: Tensor, eps: float) -> Tensor:
    """GEGLU(LayerNorm(a)) with the LayerNorm folded in (rows, colsum and bias interleaved by `interleave_geglu`)."""
    out = torch.empty((a.shape[0], w_interleaved.shape[0] // 2), dtype=torch.bfloat16, device=a.device)
    _gemm(a, w_interleaved, out, bias, None, _lib.EPI_GEGLU, out.stride(0), 0, ln_stats=stats, ln_colsum=colsum,
          ln_eps=eps)
    return out


@linear_ln_geglu.register_fake
def _(a, stats, w_interleaved, colsum, bias, eps):
    return a.new_empty((a.shape[0], w_interleaved.shape[0] // 2), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::linear_geglu", mutates_args=())
def linear_geglu(a: Tensor, w_interleaved: Tensor, bias_interleaved: Tensor) -> Tensor:
    """GEGLU (attention.py:32-39): w/bias rows pre-interleaved by `interleave_geglu`; returns bf16 [M, N/2]."""
    out = torch.empty((a.shape[0], w_interleaved.shape[0] // 2), dtype=torch.bfloat16, device=a.device)
    _gemm(a, w_interleaved, out, bias_interleaved, None, _lib.EPI_GEGLU, out.stride(0), 0)
    return out


@linear_geglu.register_fake
def _(a, w_interleaved, bias_interleaved):
    return a.new_empty((a.shape[0], w_interleaved.shape[0] // 2), dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::linear_nchw", mutates_args=())
def linear_nchw(a: Tensor, w: Tensor, bias: Tensor, residual: Tensor) -> Tensor:
    """proj_out + 'b (h w) c -> b c h w' + x_in (attention.py:335-341): a [B*hw, K] tokens, residual fp32
    NCHW-contiguous [B, N, h, w]; returns fp32 [B, N, h, w]."""
    if residual.dtype != torch.float32 or not residual.is_contiguous():
        raise ValueError("linear_nchw: fp32 NCHW-contiguous residual expected")
    out = torch.empty_like(residual)
    hw = residual.shape[2] * residual.shape[3]
    _gemm(a, w, out, bias, residual, _lib.EPI_NCHW, 0, 0, hw)
    return out


@linear_nchw.register_fake
def _(a, w, bias, residual):
    return torch.empty_like(residual)


def interleave_geglu(weight: Tensor, bias: Tensor) -> Tuple[Tensor, Tensor]:
    """Reorders GEGLU.proj rows [value(4C); gate(4C)] into 32-row groups (32 value rows, their 32 gate rows, ...)
    so that two adjacent 32-column accumulator chunks of the GEMM epilogue hold both halves of 32 outputs."""
    half = weight.shape[0] // 2
    if half % 32:
        raise ValueError("interleave_geglu: inner dim must be a multiple of 32")
    idx = torch.arange(half, device=weight.device).reshape(-1, 32)
    order = torch.cat([idx, idx + half], dim=1).reshape(-1)
    return weight[order].contiguous(), bias[order].contiguous()


# ------------------------------------------------------------------------------------------- K5 / K8
def _ld3(t: Tensor, name: str) -> int:
    if t.dim() != 3 or t.dtype != torch.bfloat16 or t.stride(2) != 1 or t.stride(0) != t.shape[1] * t.stride(1):
        raise ValueError(f"attention: {name} must be bf16 [B, L, C] with contiguous channels and a uniform row pitch")
    return t.stride(1)


@torch.library.custom_op("adaface_b200::attention", mutates_args=())
def attention(q: Tensor, k: Tensor, v: Tensor, heads: int, scale: float, key_mask: Optional[Tensor]) -> Tensor:
    """softmax(q k^T * scale [+ key mask]) v with heads concatenated along channels (attention.py:198-242).
    q [B,Nq,C], k/v [B,Nk,C] may be column slices of a fused projection buffer; key_mask uint8/bool [B,Nk]."""
    _need_cuda(q, k, v, key_mask)
    b, nq, c = q.shape
    nk = k.shape[1]
    if k.shape != v.shape or k.shape[0] != b or k.shape[2] != c or c % heads:
        raise ValueError("attention: shape mismatch")
    out = torch.empty((b, nq, c), dtype=torch.bfloat16, device=q.device)
    if key_mask is not None:
        key_mask = key_mask.reshape(b, nk).to(torch.uint8).contiguous()
    args = AttnArgs(q=_ptr(q), k=_ptr(k), v=_ptr(v), out=_ptr(out), key_mask=_ptr(key_mask), batch=b, heads=heads,
                    head_dim=c // heads, q_len=nq, kv_len=nk, ldq=_ld3(q, "q"), ldk=_ld3(k, "k"), ldv=_ld3(v, "v"),
                    ldo=c, scale=scale)
    _run("attention_fwd", _lib.load().afb_attention_fwd, C.byref(args), _stream(), flops=4.0 * b * nq * nk * c,
         nbytes=2.0 * b * c * (2 * nq + 2 * nk), tag=f"b{b}q{nq}k{nk}d{c // heads}")
    return out


@attention.register_fake
def _(q, k, v, heads, scale, key_mask):
    return q.new_empty(q.shape, dtype=torch.bfloat16)


# ------------------------------------------------------------------------------------------- K8b (explicit cross-attention)
@torch.library.custom_op("adaface_b200::cross_scores", mutates_args=())
def cross_scores(q: Tensor, k: Tensor, heads: int, scale: float) -> Tensor:
    """sim fp32 [B, heads, Nq, Nk] = scale * q k^T per head (attention.py:199), q/k bf16 [B, L, heads*d]."""
    _need_cuda(q, k)
    ldq, ldk = _ld3(q, "q"), _ld3(k, "k")
    b, nq, c = q.shape
    sim = torch.empty((b, heads, nq, k.shape[1]), dtype=torch.float32, device=q.device)
    _run("cross_scores", _lib.load().afb_cross_scores, _ptr(q), _ptr(k), _ptr(sim), b, heads, c // heads, nq, k.shape[1],
         ldq, ldk, scale, _stream(), flops=2.0 * b * nq * k.shape[1] * c, tag=f"b{b}q{nq}k{k.shape[1]}d{c // heads}")
    return sim


@cross_scores.register_fake
def _(q, k, heads, scale):
    return q.new_empty((q.shape[0], heads, q.shape[1], k.shape[1]), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::conv_attn_rows", mutates_args=("sim",))
def conv_attn_rows(sim: Tensor, q: Tensor, k: Tensor, inst_batch: Tensor, tok: Tensor, h: int, w: int, ks: int,
                   scale: float) -> None:
    """replace_rows_by_conv_attn (ldm/util.py:701-879) in place on sim [B, heads, h*w, Nk]: inst_batch int32 [I] are
    the samples that hold the subject, tok int32 [I, ks*ks] its first ks*ks token positions."""
    _need_cuda(sim, q, k, inst_batch, tok)
    if sim.dtype != torch.float32 or not sim.is_contiguous() or sim.dim() != 4:
        raise ValueError("conv_attn_rows: sim must be contiguous fp32 [B, heads, N, T]")
    if inst_batch.dtype != torch.int32 or tok.dtype != torch.int32 or not tok.is_contiguous() \
            or tok.shape != (inst_batch.numel(), ks * ks):
        raise ValueError("conv_attn_rows: inst_batch int32 [I] and tok int32 [I, ks*ks] expected")
    b, heads, n, t = sim.shape
    if n != h * w:
        raise ValueError("conv_attn_rows: sim rows do not match the feature map size")
    ldq, ldk = _ld3(q, "q"), _ld3(k, "k")
    ws = torch.empty((inst_batch.numel(), heads, n), dtype=torch.float32, device=sim.device)
    _run("conv_attn_rows", _lib.load().afb_conv_attn_rows, _ptr(sim), _ptr(q), _ptr(k), _ptr(inst_batch), _ptr(tok), _ptr(ws),
         b, heads, q.shape[2] // heads, h, w, t, ldq, ldk, inst_batch.numel(), ks, scale, _stream(),
         flops=2.0 * inst_batch.numel() * n * ks * ks * q.shape[2], tag=f"i{inst_batch.numel()}n{n}ks{ks}")


@torch.library.custom_op("adaface_b200::softmax_pv", mutates_args=())
def softmax_pv(sim: Tensor, v: Tensor, want_attn: bool) -> Tuple[Tensor, Tensor]:
    """(out bf16 [B, Nq, heads*d], attn fp32 [B, heads, Nq, Nk] or an empty tensor): softmax over the keys and
    attn . v (attention.py:238-242)."""
    _need_cuda(sim, v)
    if sim.dtype != torch.float32 or not sim.is_contiguous() or sim.dim() != 4:
        raise ValueError("softmax_pv: sim must be contiguous fp32 [B, heads, N, T]")
    ldv = _ld3(v, "v")
    b, heads, nq, t = sim.shape
    out = torch.empty((b, nq, v.shape[2]), dtype=torch.bfloat16, device=sim.device)
    attn = torch.empty_like(sim) if want_attn else torch.empty((0,), dtype=torch.float32, device=sim.device)
    _run("softmax_pv", _lib.load().afb_softmax_pv, _ptr(sim), _ptr(v), _ptr(attn) if want_attn else None, _ptr(out), b, heads,
         v.shape[2] // heads, nq, t, ldv, out.stride(1), _stream(), flops=2.0 * b * nq * t * v.shape[2],
         tag=f"b{b}q{nq}k{t}d{v.shape[2] // heads}")
    return out, attn


@softmax_pv.register_fake
def _(sim, v, want_attn):
    out = sim.new_empty((sim.shape[0], sim.shape[2], v.shape[2]), dtype=torch.bfloat16)
    return out, (torch.empty_like(sim) if want_attn else sim.new_empty((0,)))


# ------------------------------------------------------------------------------------------- K12 / K13
@torch.library.custom_op("adaface_b200::ada_pool_prep", mutates_args=())
def ada_pool_prep(to_k_weight: Tensor, fg_q_emb: Tensor, bg_q_emb: Tensor, lora_to_fg_q: Tensor,
                  lora_to_bg_q: Tensor, lora_to_k: Tensor) -> Tensor:
    """Step-invariant part of AttentionalPooler (embedding_manager.py:186-235) folded through lora_to_k (:227):
    -> pooling queries pool_q fp32 [B, 2, C]."""
    _need_cuda(to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k)
    c, ctx_dim = to_k_weight.shape
    b = bg_q_emb.shape[0]
    f = lambda t: t.to(torch.float32).contiguous()
    pool_q = torch.empty((b, 2, c), dtype=torch.float32, device=to_k_weight.device)
    lib = _lib.load()
    ws = torch.empty(int(lib.afb_ada_pool_prep_ws_floats(b, c)), dtype=torch.float32, device=to_k_weight.device)
    # the converted copies must outlive the launch: a temporary freed right after data_ptr() can be handed out again by
    # the caching allocator for the next conversion, and the kernel would read overwritten weights
    w, fg, bg, wf, wb, wk = map(f, (to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k))
    if fg.numel() != ctx_dim or bg.shape != (b, ctx_dim):
        raise ValueError("ada_pool_prep: fg_q_emb must hold ctx_dim values and bg_q_emb must be [B, ctx_dim]")
    _run("ada_pool_prep", lib.afb_ada_pool_prep, _ptr(w), ctx_dim, _ptr(fg), _ptr(bg), _ptr(wf), _ptr(wb), _ptr(wk),
         _ptr(pool_q), _ptr(ws), b, c, _stream(), nbytes=4.0 * c * ctx_dim)
    del w, fg, bg, wf, wb, wk      # stream-ordered free: the allocator reuses them only behind the launch
    return pool_q


@ada_pool_prep.register_fake
def _(to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k):
    return to_k_weight.new_empty((bg_q_emb.shape[0], 2, to_k_weight.shape[0]), dtype=torch.float32)


class AdaPrepBatch:
    """All (cross-attention layer, subject) pooling-query preparations of one U-Net call as ONE pair of launches
    (afb_ada_pool_prep_batched).  The descriptor table is uploaded once; `run()` recomputes every pool_q in place, so
    the call is CUDA-graph capturable and the outputs (`pool_q[i]`, fp32 [B, 2, C_i]) keep their addresses.
    items: (to_k_weight [C, ctx], fg_q_emb [ctx], bg_q_emb [B, ctx], lora_to_fg_q, lora_to_bg_q, lora_to_k) fp32."""

    def __init__(self, items, batch: int):
        lib = _lib.load()
        f = lambda t: t.detach().to(torch.float32).contiguous()
        self.keep, self.pool_q, table = [], [], []
        self.batch, self.ctx_dim, self.max_c = batch, int(items[0][0].shape[1]), 0
        dev = items[0][0].device
        for to_k_w, fg, bg, wf, wb, wk in items:
            _need_cuda(to_k_w, fg, bg, wf, wb, wk)
            c = int(to_k_w.shape[0])
            if to_k_w.shape[1] != self.ctx_dim or bg.reshape(batch, -1).shape[1] != self.ctx_dim:
                raise ValueError("AdaPrepBatch: every item must share ctx_dim and batch")
            ts = [f(to_k_w), f(fg.reshape(-1)), f(bg.reshape(batch, -1)), f(wf), f(wb), f(wk)]
            pq = torch.empty((batch, 2, c), dtype=torch.float32, device=dev)
            ws = torch.empty(int(lib.afb_ada_pool_prep_ws_floats(batch, c)), dtype=torch.float32, device=dev)
            self.keep.append(ts + [ws])
            self.pool_q.append(pq)
            self.max_c = max(self.max_c, c)
            table.append(_lib.AdaPrepItem(*[t.data_ptr() for t in ts], pq.data_ptr(), ws.data_ptr(), c, 0))
        raw = bytes(memoryview((_lib.AdaPrepItem * len(table))(*table)).cast("B"))
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.n = len(table)
        self.nbytes = 4.0 * sum(float(t[0].numel()) for t in self.keep)

    def run(self) -> None:
        _run("ada_pool_prep", _lib.load().afb_ada_pool_prep_batched, _ptr(self.table), self.n, self.max_c, self.ctx_dim,
             self.batch, _stream(), nbytes=self.nbytes, tag=f"batched{self.n}")


@torch.library.custom_op("adaface_b200::ada_pool", mutates_args=())
def ada_pool(x: Tensor, q: Tensor, pool_q: Tensor, key_mask: Optional[Tensor],
             want_attn: bool, drop_p: float = 0.0, seed: int = 0) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """AttentionalPooler main pass (embedding_manager.py:146-321): x, q bf16 [B,N,C] -> fg_out, bg_out fp32 [B,C]
    and, when want_attn, attn_fg/attn_bg fp32 [B*8, 1, N] (empty tensors otherwise).  drop_p > 0: train-mode dropout on
    the attention probabilities (:296) with the mask a counter-based hash of (seed, element) - see the header."""
    _need_cuda(x, q, pool_q, key_mask)
    if x.dtype != torch.bfloat16 or q.dtype != torch.bfloat16 or not x.is_contiguous() or not q.is_contiguous():
        raise ValueError("ada_pool: contiguous bf16 x and q expected")
    b, n, c = x.shape
    if pool_q.shape != (b, 2, c) or pool_q.dtype != torch.float32:
        raise ValueError("ada_pool: pool_q must be fp32 [B, 2, C] (from ada_pool_prep)")
    lib = _lib.load()
    dev = x.device
    ws = torch.empty(int(lib.afb_ada_pool_ws_floats(b, n, c)), dtype=torch.float32, device=dev)
    fg = torch.empty((b, c), dtype=torch.float32, device=dev)
    bg = torch.empty((b, c), dtype=torch.float32, device=dev)
    a_fg = torch.empty((b * 8, 1, n) if want_attn else (0,), dtype=torch.float32, device=dev)
    a_bg = torch.empty((b * 8, 1, n) if want_attn else (0,), dtype=torch.float32, device=dev)
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    pool_q = pool_q.contiguous()
    _run("ada_pool", lib.afb_ada_pool, _ptr(x), _ptr(q), _ptr(pool_q), _ptr(key_mask), _ptr(fg),
         _ptr(bg), _ptr(a_fg) if want_attn else None, _ptr(a_bg) if want_attn else None, _ptr(ws), b, n, c, float(drop_p),
         int(seed) & 0xFFFFFFFF, _stream(), nbytes=4.0 * b * n * c, tag=f"b{b}n{n}C{c}")
    return fg, bg, a_fg, a_bg


@ada_pool.register_fake
def _(x, q, pool_q, key_mask, want_attn, drop_p=0.0, seed=0):
    b, n, c = x.shape
    shape = (b * 8, 1, n) if want_attn else (0,)
    f = lambda s: x.new_empty(s, dtype=torch.float32)
    return f((b, c)), f((b, c)), f(shape), f(shape)


def _ada_head_impl(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias, ctx_out, token_pos):
    _need_cuda(infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, bias, ctx_out, token_pos)
    b, c = infeat_fg.shape
    if basis.dim() == 4:                      # zero-shot: per-instance basis vectors [B, K, r, D] (:860-868)
        if basis.shape[0] != b:
            raise ValueError("ada_head: per-instance basis must be [B, K, r, D]")
        k, r, d = basis.shape[1:]
        basis_stride = k * r * d
    else:
        k, r, d = basis.shape
        basis_stride = 0
    f = lambda t: None if t is None else t.to(torch.float32).contiguous()
    infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, bias = map(
        f, (infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, bias))
    if coeff_w.shape != (k * r, 2 * c + td):
        raise ValueError(f"ada_head: coefficient map {tuple(coeff_w.shape)} does not match ({k * r}, {2 * c + td})")
    stride = 0
    if ctx_out is not None:
        if token_pos is None or token_pos.dtype != torch.int32 or token_pos.numel() != k:
            raise ValueError("ada_head: token_pos must be int32 [K] when ctx_out is given")
        if ctx_out.dtype != torch.bfloat16 or ctx_out.dim() != 3 or ctx_out.shape[2] != d or ctx_out.stride(2) != 1 \
                or ctx_out.stride(1) != d or ctx_out.shape[0] < b:
            raise ValueError("ada_head: ctx_out must be bf16 [B, T, D] with contiguous rows")
        stride = ctx_out.stride(0)
    out = torch.empty((b, k, d), dtype=torch.float32, device=infeat_fg.device)
    _run("ada_head", _lib.load().afb_ada_head, _ptr(infeat_fg), _ptr(infeat_bg), _ptr(time_emb), time_emb.stride(0),
         _ptr(coeff_w), _ptr(coeff_b), _ptr(basis), _ptr(bias), _ptr(out), _ptr(ctx_out), _ptr(token_pos), stride,
         basis_stride, b, c, td, k, r, d, _stream(), nbytes=4.0 * (coeff_w.numel() + basis.numel()))
    return out


@torch.library.custom_op("adaface_b200::ada_pool_head_splice", mutates_args=("ctx_out",))
def ada_pool_head_splice(x: Tensor, q: Tensor, pool_q: Tensor, key_mask: Optional[Tensor], cached_bg: Optional[Tensor],
                         time_emb: Tensor, td: int, coeff_w: Tensor, coeff_b: Tensor, basis: Tensor, bias: Optional[Tensor],
                         ctx_out: Tensor, token_pos: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """`ada_pool` + `ada_head_splice` in two launches instead of three: the pooling kernel leaves its chunk partials in the
    workspace and the head merges them in its prologue (inference path of one cross-attention layer).
    Returns (out fp32 [B, K, D], fg_out, bg_out fp32 [B, C]); cached_bg replaces the pooled bg stream as the head's input."""
    _need_cuda(x, q, pool_q, key_mask, cached_bg, time_emb, coeff_w, coeff_b, basis, bias, ctx_out, token_pos)
    if x.dtype != torch.bfloat16 or q.dtype != torch.bfloat16 or not x.is_contiguous() or not q.is_contiguous():
        raise ValueError("ada_pool_head_splice: contiguous bf16 x and q expected")
    b, n, c = x.shape
    if pool_q.shape != (b, 2, c) or pool_q.dtype != torch.float32:
        raise ValueError("ada_pool_head_splice: pool_q must be fp32 [B, 2, C] (from ada_pool_prep)")
    if basis.dim() == 4:
        if basis.shape[0] != b:
            raise ValueError("ada_pool_head_splice: per-instance basis must be [B, K, r, D]")
        k, r, d = basis.shape[1:]
        basis_stride = k * r * d
    else:
        k, r, d = basis.shape
        basis_stride = 0
    f = lambda t: None if t is None else t.to(torch.float32).contiguous()
    cached_bg, time_emb, coeff_w, coeff_b, basis, bias = map(f, (cached_bg, time_emb, coeff_w, coeff_b, basis, bias))
    if coeff_w.shape != (k * r, 2 * c + td):
        raise ValueError(f"ada_pool_head_splice: coefficient map {tuple(coeff_w.shape)} does not match ({k * r}, {2 * c + td})")
    if token_pos.dtype != torch.int32 or token_pos.numel() != k or ctx_out.dtype != torch.bfloat16 or ctx_out.dim() != 3 \
            or ctx_out.shape[2] != d or ctx_out.stride(2) != 1 or ctx_out.stride(1) != d or ctx_out.shape[0] < b:
        raise ValueError("ada_pool_head_splice: ctx_out must be bf16 [B, T, D] with contiguous rows, token_pos int32 [K]")
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    pool_q = pool_q.contiguous()
    lib = _lib.load()
    dev = x.device
    ws = torch.empty(int(lib.afb_ada_pool_ws_floats(b, n, c)), dtype=torch.float32, device=dev)
    fg = torch.empty((b, c), dtype=torch.float32, device=dev)
    bg = torch.empty((b, c), dtype=torch.float32, device=dev)
    out = torch.empty((b, k, d), dtype=torch.float32, device=dev)
    _run("ada_pool", lib.afb_ada_pool, _ptr(x), _ptr(q), _ptr(pool_q), _ptr(key_mask), None, None, None, None, _ptr(ws), b, n, c,
         0.0, 0, _stream(), nbytes=4.0 * b * n * c, tag=f"b{b}n{n}C{c}")
    _run("ada_head", lib.afb_ada_head_pooled, _ptr(ws), n, _ptr(fg), _ptr(bg), _ptr(cached_bg), _ptr(time_emb), time_emb.stride(0),
         _ptr(coeff_w), _ptr(coeff_b), _ptr(basis), _ptr(bias), _ptr(out), _ptr(ctx_out), _ptr(token_pos), ctx_out.stride(0),
         basis_stride, b, c, td, k, r, d, _stream(), nbytes=4.0 * (coeff_w.numel() + basis.numel()))
    return out, fg, bg


@ada_pool_head_splice.register_fake
def _(x, q, pool_q, key_mask, cached_bg, time_emb, td, coeff_w, coeff_b, basis, bias, ctx_out, token_pos):
    b, _, c = x.shape
    f = lambda s_: x.new_empty(s_, dtype=torch.float32)
    return f((b, basis.shape[-3], basis.shape[-1])), f((b, c)), f((b, c))


@torch.library.custom_op("adaface_b200::ada_head", mutates_args=())
def ada_head(infeat_fg: Tensor, infeat_bg: Tensor, time_emb: Tensor, td: int, coeff_w: Tensor, coeff_b: Tensor,
             basis: Tensor, bias: Optional[Tensor]) -> Tensor:
    """AdaEmbedding head (embedding_manager.py:847-893): -> out fp32 [B, K, 768]."""
    return _ada_head_impl(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias, None, None)


@ada_head.register_fake
def _(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias):
    return infeat_fg.new_empty((infeat_fg.shape[0], basis.shape[-3], basis.shape[-1]), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::ada_head_splice", mutates_args=("ctx_out",))
def ada_head_splice(infeat_fg: Tensor, infeat_bg: Tensor, time_emb: Tensor, td: int, coeff_w: Tensor, coeff_b: Tensor,
                    basis: Tensor, bias: Optional[Tensor], ctx_out: Tensor, token_pos: Tensor) -> Tensor:
    """ada_head that also writes the K vectors, as bf16, into rows token_pos (int32 [K]) of ctx_out (bf16 [B', T, 768],
    B' >= B) - the splice of the subject-token rows into the layer's context."""
    return _ada_head_impl(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias, ctx_out, token_pos)


@ada_head_splice.register_fake
def _(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias, ctx_out, token_pos):
    return infeat_fg.new_empty((infeat_fg.shape[0], basis.shape[-3], basis.shape[-1]), dtype=torch.float32)


# ------------------------------------------------------------------------------------------- static embedding (N2)
def _f32c(t: Optional[Tensor], what: str) -> Optional[Tensor]:
    if t is None:
        return None
    if t.dtype != torch.float32:
        raise ValueError(f"{what}: fp32 tensor expected")
    return t.contiguous()


@torch.library.custom_op("adaface_b200::static_layerwise_embedding", mutates_args=())
def static_layerwise_embedding(basis_rand_weights: Tensor, basis_comm_weights: Tensor, pre_vecs: Optional[Tensor],
                               basis_vecs: Tensor, bias: Optional[Tensor]) -> Tensor:
    """StaticLayerwiseEmbedding.forward (embedding_manager.py:497-537): rand [L,K,r], comm [1,K,r], pre_vecs
    [K,N,D] or None, basis_vecs [K,r-N,D], bias [L,K,D] or None -> fp32 [L,K,D]."""
    _need_cuda(basis_rand_weights, basis_comm_weights, pre_vecs, basis_vecs, bias)
    rw, cw = _f32c(basis_rand_weights, "basis_rand_weights"), _f32c(basis_comm_weights, "basis_comm_weights")
    pv, bv, bs = _f32c(pre_vecs, "pre_vecs"), _f32c(basis_vecs, "basis_vecs"), _f32c(bias, "bias")
    L, K, r = rw.shape
    n_pre = 0 if pv is None else pv.shape[1]
    D = bv.shape[-1]
    if cw.numel() != K * r or bv.shape[0] != K or bv.shape[1] != r - n_pre or (pv is not None and pv.shape != (K, n_pre, D)) \
            or (bs is not None and bs.shape != (L, K, D)):
        raise ValueError("static_layerwise_embedding: inconsistent parameter shapes")
    out = torch.empty((L, K, D), device=rw.device, dtype=torch.float32)
    _run("static_layerwise_embedding", _lib.load().afb_static_layerwise_embedding, _ptr(rw), _ptr(cw), _ptr(pv), n_pre,
         _ptr(bv), _ptr(bs), _ptr(out), L, K, r, D, _stream(), nbytes=4.0 * (L * K * D * 2 + K * r * D))
    return out


@static_layerwise_embedding.register_fake
def _(basis_rand_weights, basis_comm_weights, pre_vecs, basis_vecs, bias):
    L, K, _ = basis_rand_weights.shape
    return basis_vecs.new_empty((L, K, basis_vecs.shape[-1]), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::first_token_pos", mutates_args=())
def first_token_pos(tokens: Tensor, token: int) -> Tensor:
    """tokens int64 [B,N] -> int32 [B]: first position of `token` in each prompt, -1 if absent (ldm/util.py:2114)."""
    _need_cuda(tokens)
    if tokens.dtype != torch.int64 or tokens.dim() != 2:
        raise ValueError("first_token_pos: int64 [B, N] token ids expected")
    tokens = tokens.contiguous()
    out = torch.empty((tokens.shape[0],), device=tokens.device, dtype=torch.int32)
    _run("first_token_pos", _lib.load().afb_first_token_pos, _ptr(tokens), int(token), _ptr(out), tokens.shape[0],
         tokens.shape[1], _stream(), nbytes=8.0 * tokens.numel())
    return out


@first_token_pos.register_fake
def _(tokens, token):
    return tokens.new_empty((tokens.shape[0],), dtype=torch.int32)


@torch.library.custom_op("adaface_b200::splice_static_embedding", mutates_args=())
def splice_static_embedding(embedded_text: Tensor, first_pos: Tensor, static_emb: Tensor, to_bf16: bool) -> Tensor:
    """The layer-repeated prompt with the subject rows spliced in (embedding_manager.py:1341-1352, 1510-1552):
    embedded_text fp32 [B,N,D], first_pos int32 [B], static_emb fp32 [L,K,D] -> [B*L, N, D] fp32 (or bf16)."""
    _need_cuda(embedded_text, first_pos, static_emb)
    et, se = _f32c(embedded_text, "embedded_text"), _f32c(static_emb, "static_emb")
    if first_pos.dtype != torch.int32 or et.dim() != 3 or first_pos.numel() != et.shape[0] or se.dim() != 3 \
            or se.shape[-1] != et.shape[-1]:
        raise ValueError("splice_static_embedding: embedded_text [B,N,D], first_pos int32 [B], static_emb [L,K,D] expected")
    B, N, D = et.shape
    L, K, _ = se.shape
    out = torch.empty((B * L, N, D), device=et.device, dtype=torch.bfloat16 if to_bf16 else torch.float32)
    first_pos = first_pos.contiguous()
    _run("splice_static_embedding", _lib.load().afb_splice_static_embedding, _ptr(et), _ptr(first_pos), _ptr(se),
         None if to_bf16 else _ptr(out), _ptr(out) if to_bf16 else None, B, N, L, K, D, _stream(),
         nbytes=float(B * L * N * D) * (4 + (2 if to_bf16 else 4)))
    return out


@splice_static_embedding.register_fake
def _(embedded_text, first_pos, static_emb, to_bf16):
    B, N, D = embedded_text.shape
    return embedded_text.new_empty((B * static_emb.shape[0], N, D), dtype=torch.bfloat16 if to_bf16 else torch.float32)


# ------------------------------------------------------------------------------------------- elementwise
@torch.library.custom_op("adaface_b200::add_bias_residual", mutates_args=())
def add_bias_residual(a: Tensor, b: Optional[Tensor], bias: Optional[Tensor]) -> Tensor:
    """a + b + bias over bf16 token-major [..., C] tensors (b, bias optional; bias fp32 [C])."""
    _need_cuda(a, b, bias)
    if a.dtype != torch.bfloat16 or not a.is_contiguous() or (b is not None and (b.dtype != a.dtype or b.shape != a.shape
                                                                                 or not b.is_contiguous())):
        raise ValueError("add_bias_residual: contiguous bf16 tensors of one shape expected")
    c = a.shape[-1]
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != c):
        raise ValueError("add_bias_residual: bias must be fp32 [C]")
    out = torch.empty_like(a)
    _run("add_bias_residual", _lib.load().afb_add_bias_residual, _ptr(a), _ptr(b), _ptr(bias), _ptr(out), a.numel() // c, c,
         _stream(), nbytes=2.0 * a.numel() * (2 + (b is not None)))
    return out


@add_bias_residual.register_fake
def _(a, b, bias):
    return torch.empty_like(a)


# ------------------------------------------------------------------------------------------- casts
@torch.library.custom_op("adaface_b200::to_bf16", mutates_args=())
def to_bf16(x: Tensor) -> Tensor:
    _need_cuda(x)
    if x.dtype != torch.float32:
        raise ValueError("to_bf16: fp32 input expected")
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _run("cast_f32_to_bf16", _lib.load().afb_cast_f32_to_bf16, _ptr(x), _ptr(y), x.numel(), _stream(),
         nbytes=6.0 * x.numel())
    return y


@to_bf16.register_fake
def _(x):
    return x.new_empty(x.shape, dtype=torch.bfloat16)


@torch.library.custom_op("adaface_b200::to_f32", mutates_args=())
def to_f32(x: Tensor) -> Tensor:
    _need_cuda(x)
    if x.dtype != torch.bfloat16:
        raise ValueError("to_f32: bf16 input expected")
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _run("cast_bf16_to_f32", _lib.load().afb_cast_bf16_to_f32, _ptr(x), _ptr(y), x.numel(), _stream(),
         nbytes=6.0 * x.numel())
    return y


@to_f32.register_fake
def _(x):
    return x.new_empty(x.shape, dtype=torch.float32)


# =========================================================================================== backward (inversion step)
# BASELINE.json configs[3]: the U-Net is frozen, the AdaEmbedding parameters are the trainable set.  The reference gets
# its gradients from torch autograd (ddpm.py:599 manual_backward); here every op on the path from the block output back
# to those parameters carries an analytic backward built from the kernels of csrc/backward.cu and from the forward GEMM
# (dA = dC W against the frozen weight).  Not built: the backward of self-attention (long key sequences), of GroupNorm and
# of the pooler towards the layer features x / q - ops that would need them raise instead of returning wrong gradients.
_WT_CACHE: Dict[Tuple, Tensor] = {}


def _transposed(w: Tensor) -> Tensor:
    """bf16 [N, K] weight -> contiguous [K, N] (the `w` operand of the dX GEMM), cached per weight tensor."""
    key = (w.data_ptr(), w._version, tuple(w.shape), tuple(w.stride()))
    t = _WT_CACHE.get(key)
    if t is None:
        if len(_WT_CACHE) > 512:
            _WT_CACHE.clear()
        t = w.detach().t().contiguous()
        _WT_CACHE[key] = t
    return t


def _frozen(*ws: Optional[Tensor]) -> None:
    for w in ws:
        if w is not None and w.requires_grad:
            raise NotImplementedError("adaface_b200: the inversion step trains the embedding parameters only - weight "
                                      "gradients of the U-Net GEMMs are not built; freeze the U-Net (requires_grad_(False))")


def _dx_gemm(dout: Tensor, w: Tensor) -> Tensor:
    """dA [M, K] = dOut [M, N] @ W [N, K]: the forward GEMM with the transposed weight."""
    return linear(dout.contiguous(), _transposed(w), None, None, False)


@torch.library.custom_op("adaface_b200::geglu_bwd", mutates_args=())
def geglu_bwd(h: Tensor, dout: Tensor) -> Tensor:
    """h bf16 [M, 2H] (interleaved value / gate column order of the forward GEMM), dout bf16 [M, H] -> dh bf16 [M, 2H]."""
    _need_cuda(h, dout)
    if h.dtype != torch.bfloat16 or dout.dtype != torch.bfloat16 or h.dim() != 2 or dout.shape != (h.shape[0], h.shape[1] // 2):
        raise ValueError("geglu_bwd: h bf16 [M, 2H] and dout bf16 [M, H] expected")
    h, dout = h.contiguous(), dout.contiguous()
    dh = torch.empty_like(h)
    _run("geglu_bwd", _lib.load().afb_geglu_bwd, _ptr(h), _ptr(dout), _ptr(dh), h.shape[0], dout.shape[1], _stream(),
         nbytes=2.0 * (2 * h.numel() + dout.numel()))
    return dh


@geglu_bwd.register_fake
def _(h, dout):
    return torch.empty_like(h)


@torch.library.custom_op("adaface_b200::layernorm_bwd", mutates_args=())
def layernorm_bwd(a: Tensor, g: Tensor, add: Optional[Tensor], eps: float) -> Tensor:
    """Backward of the affine-free LayerNorm over the last dim: a, g = d(a_hat) bf16 [M, C] (+ add: the gradient that
    reaches `a` along the residual stream) -> da bf16 [M, C]."""
    _need_cuda(a, g, add)
    if a.dtype != torch.bfloat16 or g.dtype != torch.bfloat16 or a.shape != g.shape or (add is not None and (
            add.dtype != torch.bfloat16 or add.shape != a.shape)):
        raise ValueError("layernorm_bwd: bf16 tensors of one shape expected")
    a, g = a.contiguous(), g.contiguous()
    add = None if add is None else add.contiguous()
    da = torch.empty_like(a)
    c = a.shape[-1]
    _run("layernorm_bwd", _lib.load().afb_layernorm_bwd, _ptr(a), _ptr(g), _ptr(add), _ptr(da), a.numel() // c, c, eps, _stream(),
         nbytes=2.0 * a.numel() * (3 + (add is not None)))
    return da


@layernorm_bwd.register_fake
def _(a, g, add, eps):
    return torch.empty_like(a)


@torch.library.custom_op("adaface_b200::attention_bwd_short", mutates_args=())
def attention_bwd_short(q: Tensor, k: Tensor, v: Tensor, dout: Tensor, heads: int, scale: float,
                        want_dq: bool) -> Tuple[Tensor, Tensor, Tensor]:
    """Backward of `attention` for a short key sequence (the 77-token context): -> (dq bf16 [B,Nq,C] or an empty tensor,
    dk fp32 [B,Nk,C], dv fp32 [B,Nk,C])."""
    _need_cuda(q, k, v, dout)
    b, nq, c = q.shape
    nk = k.shape[1]
    dout = dout.contiguous()
    if dout.dtype != torch.bfloat16 or dout.shape != q.shape:
        raise ValueError("attention_bwd_short: dout must be bf16 of q's shape")
    dq = torch.empty((b, nq, c) if want_dq else (0,), dtype=torch.bfloat16, device=q.device)
    dk = torch.zeros((b, nk, c), dtype=torch.float32, device=q.device)
    dv = torch.zeros((b, nk, c), dtype=torch.float32, device=q.device)
    args = AttnArgs(q=_ptr(q), k=_ptr(k), v=_ptr(v), out=None, key_mask=None, batch=b, heads=heads, head_dim=c // heads,
                    q_len=nq, kv_len=nk, ldq=_ld3(q, "q"), ldk=_ld3(k, "k"), ldv=_ld3(v, "v"), ldo=c, scale=scale)
    _run("attention_bwd", _lib.load().afb_attention_bwd_short, C.byref(args), _ptr(dout), _ptr(dq) if want_dq else None, _ptr(dk),
         _ptr(dv), c, c, _stream(), flops=10.0 * b * nq * nk * c, tag=f"b{b}q{nq}k{nk}d{c // heads}")
    return dq, dk, dv


@attention_bwd_short.register_fake
def _(q, k, v, dout, heads, scale, want_dq):
    f = lambda t: t.new_empty(t.shape, dtype=torch.float32)
    return (torch.empty_like(q) if want_dq else q.new_empty((0,))), f(k), f(v)


@torch.library.custom_op("adaface_b200::ada_head_bwd", mutates_args=())
def ada_head_bwd(infeat_fg: Tensor, infeat_bg: Tensor, time_emb: Tensor, td: int, coeff_w: Tensor, coeff_b: Tensor,
                 basis: Tensor, grad_out: Tensor, want_bias: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    """Backward of `ada_head` (shared basis [K, r, D]): -> (d_coeff_w, d_coeff_b, d_basis, d_bias [K, D] or empty,
    d_infeat_fg, d_infeat_bg), all fp32."""
    _need_cuda(infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, grad_out)
    b, c = infeat_fg.shape
    if basis.dim() != 3:
        raise NotImplementedError("ada_head_bwd: the zero-shot per-instance basis has no trainable tensor on this path")
    k, r, d = basis.shape
    f = lambda t: t.detach().to(torch.float32).contiguous()
    infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, grad_out = map(
        f, (infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis, grad_out))
    dev = infeat_fg.device
    z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
    d_cw, d_cb, d_basis = z(*coeff_w.shape), z(*coeff_b.shape), z(k, r, d)
    d_bias = z(k, d) if want_bias else z(0)
    ws = torch.empty((b, 2 * c + td), dtype=torch.float32, device=dev)
    d_fg, d_bg = torch.empty((b, c), dtype=torch.float32, device=dev), torch.empty((b, c), dtype=torch.float32, device=dev)
    _run("ada_head_bwd", _lib.load().afb_ada_head_bwd, _ptr(infeat_fg), _ptr(infeat_bg), _ptr(time_emb), time_emb.stride(0),
         _ptr(coeff_w), _ptr(coeff_b), _ptr(basis), _ptr(grad_out), _ptr(d_cw), _ptr(d_cb), _ptr(d_basis),
         _ptr(d_bias) if want_bias else None, _ptr(ws), _ptr(d_fg), _ptr(d_bg), b, c, td, k, r, d, _stream(),
         nbytes=8.0 * (coeff_w.numel() + basis.numel()))
    return d_cw, d_cb, d_basis, d_bias, d_fg, d_bg


@ada_head_bwd.register_fake
def _(infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, grad_out, want_bias):
    f = lambda t: t.new_empty(t.shape, dtype=torch.float32)
    k, _, d = basis.shape
    return f(coeff_w), f(coeff_b), f(basis), infeat_fg.new_empty((k, d) if want_bias else (0,), dtype=torch.float32), \
        f(infeat_fg), f(infeat_bg)


@torch.library.custom_op("adaface_b200::ada_pool_fwd_state", mutates_args=())
def ada_pool_fwd_state(x: Tensor, q: Tensor, pool_q: Tensor, key_mask: Optional[Tensor], drop_p: float = 0.0,
                       seed: int = 0) -> Tensor:
    """What both pooler backwards need from the forward, in one fp32 vector: the merged softmax statistics (max, sum) per
    (sample, head) [B * 16] followed by the pooled outputs fg [B, C] and bg [B, C].  One re-run of the pooling kernel per
    layer and backward (the forward op does not return its statistics), shared by `ada_pool_bwd` and
    `ada_pool_bwd_features`."""
    _need_cuda(x, q, pool_q, key_mask)
    b, n, c = x.shape
    lib = _lib.load()
    dev = x.device
    x, q = x.contiguous(), q.contiguous()
    pool_q = pool_q.detach().to(torch.float32).contiguous()
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    ws = torch.empty(int(lib.afb_ada_pool_ws_floats(b, n, c)), dtype=torch.float32, device=dev)
    state = torch.empty((b * 16 + 2 * b * c,), dtype=torch.float32, device=dev)
    fg, bg = state[b * 16:b * 16 + b * c], state[b * 16 + b * c:]
    _run("ada_pool", lib.afb_ada_pool, _ptr(x), _ptr(q), _ptr(pool_q), _ptr(key_mask), _ptr(fg), _ptr(bg), None, None, _ptr(ws),
         b, n, c, float(drop_p), int(seed) & 0xFFFFFFFF, _stream(), nbytes=4.0 * b * n * c, tag=f"b{b}n{n}C{c}")
    off = int(lib.afb_ada_pool_stats_offset(b, n, c))       # the forward's merged (max, sum) per (sample, head)
    state[:b * 16].copy_(ws[off:off + b * 16])
    return state


@ada_pool_fwd_state.register_fake
def _(x, q, pool_q, key_mask, drop_p=0.0, seed=0):
    return x.new_empty((x.shape[0] * 16 + 2 * x.shape[0] * x.shape[2],), dtype=torch.float32)


def _split_fwd_state(state: Tensor, b: int, c: int):
    if state.dtype != torch.float32 or state.numel() != b * 16 + 2 * b * c or not state.is_contiguous():
        raise ValueError("ada_pool backward: fwd_state must be the fp32 vector of ada_pool_fwd_state")
    return state[:b * 16], state[b * 16:b * 16 + b * c], state[b * 16 + b * c:]


@torch.library.custom_op("adaface_b200::ada_pool_bwd", mutates_args=())
def ada_pool_bwd(x: Tensor, q: Tensor, pool_q: Tensor, key_mask: Optional[Tensor], d_fg: Tensor, d_bg: Tensor,
                 drop_p: float = 0.0, seed: int = 0, fwd_state: Optional[Tensor] = None) -> Tensor:
    """Backward of `ada_pool` w.r.t. the folded pooling queries: needs the forward's softmax statistics and pooled outputs
    (`fwd_state` from `ada_pool_fwd_state`, else the forward is re-run here), then one streaming pass over x and q ->
    d_pool_q fp32 [B, 2, C]."""
    _need_cuda(x, q, pool_q, key_mask, d_fg, d_bg, fwd_state)
    b, n, c = x.shape
    lib = _lib.load()
    dev = x.device
    x, q = x.contiguous(), q.contiguous()
    pool_q = pool_q.detach().to(torch.float32).contiguous()
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    if fwd_state is None:
        fwd_state = ada_pool_fwd_state(x, q, pool_q, key_mask, drop_p, seed)
    stats, fg, bg = _split_fwd_state(fwd_state, b, c)
    d_fg, d_bg = d_fg.detach().to(torch.float32).contiguous(), d_bg.detach().to(torch.float32).contiguous()
    d_pool_q = torch.empty((b, 2, c), dtype=torch.float32, device=dev)
    _run("ada_pool_bwd", lib.afb_ada_pool_bwd, _ptr(x), _ptr(q), _ptr(pool_q), _ptr(key_mask), _ptr(stats), _ptr(fg), _ptr(bg),
         _ptr(d_fg), _ptr(d_bg), _ptr(d_pool_q), b, n, c, float(drop_p), int(seed) & 0xFFFFFFFF, _stream(),
         nbytes=4.0 * b * n * c, tag=f"b{b}n{n}C{c}")
    return d_pool_q


@ada_pool_bwd.register_fake
def _(x, q, pool_q, key_mask, d_fg, d_bg, drop_p=0.0, seed=0, fwd_state=None):
    return x.new_empty((x.shape[0], 2, x.shape[2]), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::ada_pool_prep_bwd", mutates_args=())
def ada_pool_prep_bwd(to_k_weight: Tensor, fg_q_emb: Tensor, bg_q_emb: Tensor, lora_to_fg_q: Tensor, lora_to_bg_q: Tensor,
                      lora_to_k: Tensor, d_pool_q: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """Backward of `ada_pool_prep` w.r.t. the three grouped-conv weights [C/8, C/8] (the to_k projection is frozen, the
    query embeddings are inputs): re-runs the projection for its workspace, then one small kernel."""
    _need_cuda(to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k, d_pool_q)
    c, ctx_dim = to_k_weight.shape
    b = bg_q_emb.shape[0]
    f = lambda t: t.detach().to(torch.float32).contiguous()
    w, fg, bg, wf, wb, wk, dq = map(f, (to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k, d_pool_q))
    lib = _lib.load()
    dev = w.device
    pool_q = torch.empty((b, 2, c), dtype=torch.float32, device=dev)
    ws = torch.empty(int(lib.afb_ada_pool_prep_ws_floats(b, c)), dtype=torch.float32, device=dev)
    _run("ada_pool_prep", lib.afb_ada_pool_prep, _ptr(w), ctx_dim, _ptr(fg), _ptr(bg), _ptr(wf), _ptr(wb), _ptr(wk), _ptr(pool_q),
         _ptr(ws), b, c, _stream(), nbytes=4.0 * c * ctx_dim)
    d_wf, d_wb, d_wk = torch.zeros_like(wf), torch.zeros_like(wb), torch.zeros_like(wk)
    _run("ada_pool_prep_bwd", lib.afb_ada_pool_prep_bwd, _ptr(ws), _ptr(wf), _ptr(wb), _ptr(wk), _ptr(dq), _ptr(d_wf), _ptr(d_wb),
         _ptr(d_wk), b, c, _stream(), nbytes=12.0 * wk.numel())
    return d_wf, d_wb, d_wk


@ada_pool_prep_bwd.register_fake
def _(to_k_weight, fg_q_emb, bg_q_emb, lora_to_fg_q, lora_to_bg_q, lora_to_k, d_pool_q):
    f = lambda t: t.new_empty(t.shape, dtype=torch.float32)
    return f(lora_to_fg_q), f(lora_to_bg_q), f(lora_to_k)


# ---- autograd formulas (torch.library.register_autograd): dX only against frozen weights -----------------------------
def _bf16c(g: Tensor) -> Tensor:
    return g.to(torch.bfloat16).contiguous()


def _linear_setup(ctx, inputs, output):
    a, w, bias, residual, out_fp32 = inputs
    _frozen(w, bias)
    ctx.save_for_backward(w)
    ctx.a_needs, ctx.res_needs = a.requires_grad, residual is not None and residual.requires_grad
    ctx.a_dtype, ctx.res_dtype = a.dtype, (None if residual is None else residual.dtype)


def _linear_backward(ctx, g):
    (w,) = ctx.saved_tensors
    gb = _bf16c(g)
    da = _dx_gemm(gb, w).to(ctx.a_dtype) if ctx.a_needs else None
    dres = g.to(ctx.res_dtype) if ctx.res_needs else None
    return da, None, None, dres, None


linear.register_autograd(_linear_backward, setup_context=_linear_setup)


def _linear_stats_setup(ctx, inputs, output):
    a, w, bias, residual = inputs
    _frozen(w, bias)
    ctx.save_for_backward(w)
    ctx.a_needs, ctx.res_needs = a.requires_grad, residual is not None and residual.requires_grad


def _linear_stats_backward(ctx, g, g_stats):
    # the row statistics are a side output consumed by the LayerNorm-folded GEMMs, whose backward differentiates the
    # LayerNorm itself (from the activations): no gradient flows through `stats`
    (w,) = ctx.saved_tensors
    gb = _bf16c(g)
    return (_dx_gemm(gb, w) if ctx.a_needs else None), None, None, (gb if ctx.res_needs else None)


linear_stats.register_autograd(_linear_stats_backward, setup_context=_linear_stats_setup)


def _linear_ln_setup(ctx, inputs, output):
    a, stats, w_gamma, colsum, bias, eps = inputs
    _frozen(w_gamma, colsum, bias)
    ctx.save_for_backward(a, w_gamma)
    ctx.eps = eps


def _linear_ln_backward(ctx, g):
    a, w_gamma = ctx.saved_tensors
    d_hat = _dx_gemm(_bf16c(g), w_gamma)                       # gradient w.r.t. the normalised activations
    return layernorm_bwd(a, d_hat, None, ctx.eps), None, None, None, None, None


linear_ln.register_autograd(_linear_ln_backward, setup_context=_linear_ln_setup)


def _linear_ln_geglu_setup(ctx, inputs, output):
    a, stats, w_interleaved, colsum, bias, eps = inputs
    _frozen(w_interleaved, colsum, bias)
    ctx.save_for_backward(a, stats, w_interleaved, colsum, bias)
    ctx.eps = eps


def _linear_ln_geglu_backward(ctx, g):
    a, stats, w_i, colsum, bias = ctx.saved_tensors
    # the forward never writes the 8C-wide projection: recompute it (same interleaved column order), then
    # d(a_hat) = dh W_interleaved and the LayerNorm backward
    h = linear_ln(a, stats, w_i, colsum, bias, ctx.eps)
    dh = geglu_bwd(h, _bf16c(g))
    d_hat = _dx_gemm(dh, w_i)
    return layernorm_bwd(a, d_hat, None, ctx.eps), None, None, None, None, None


linear_ln_geglu.register_autograd(_linear_ln_geglu_backward, setup_context=_linear_ln_geglu_setup)


def _attention_setup(ctx, inputs, output):
    q, k, v, heads, scale, key_mask = inputs
    if k.shape[1] > 128 or key_mask is not None:
        raise NotImplementedError("adaface_b200: only the backward of the short-key (77-token cross-) attention is built; "
                                  "self-attention / masked attention cannot take part in a differentiated graph yet")
    ctx.save_for_backward(q, k, v)
    ctx.heads, ctx.scale, ctx.q_needs = heads, scale, q.requires_grad


def _attention_backward(ctx, g):
    q, k, v = ctx.saved_tensors
    dq, dk, dv = attention_bwd_short(q, k, v, _bf16c(g), ctx.heads, ctx.scale, ctx.q_needs)
    return (dq if ctx.q_needs else None), dk.to(k.dtype), dv.to(v.dtype), None, None, None


attention.register_autograd(_attention_backward, setup_context=_attention_setup)


def _ada_head_setup(ctx, inputs, output):
    infeat_fg, infeat_bg, time_emb, td, coeff_w, coeff_b, basis, bias = inputs
    ctx.save_for_backward(infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis)
    ctx.td, ctx.has_bias = td, bias is not None


def _ada_head_backward(ctx, g):
    infeat_fg, infeat_bg, time_emb, coeff_w, coeff_b, basis = ctx.saved_tensors
    d_cw, d_cb, d_basis, d_bias, d_fg, d_bg = ada_head_bwd(infeat_fg, infeat_bg, time_emb, ctx.td, coeff_w, coeff_b, basis,
                                                           g.contiguous(), ctx.has_bias)
    # the time embedding carries no trainable tensor on this path: no gradient is produced for it
    return d_fg, d_bg, None, None, d_cw, d_cb, d_basis, (d_bias if ctx.has_bias else None)


ada_head.register_autograd(_ada_head_backward, setup_context=_ada_head_setup)


def _ada_pool_prep_setup(ctx, inputs, output):
    to_k_weight, fg_q_emb, bg_q_emb, wf, wb, wk = inputs
    _frozen(to_k_weight)
    ctx.save_for_backward(to_k_weight, fg_q_emb, bg_q_emb, wf, wb, wk)


def _ada_pool_prep_backward(ctx, g):
    to_k_weight, fg_q_emb, bg_q_emb, wf, wb, wk = ctx.saved_tensors
    d_wf, d_wb, d_wk = ada_pool_prep_bwd(to_k_weight, fg_q_emb, bg_q_emb, wf, wb, wk, g.contiguous())
    return None, None, None, d_wf.to(wf.dtype), d_wb.to(wb.dtype), d_wk.to(wk.dtype)


ada_pool_prep.register_autograd(_ada_pool_prep_backward, setup_context=_ada_pool_prep_setup)


# ---- second stage: self-attention, GroupNorm and the pooler's gradient towards the layer features --------------------
@torch.library.custom_op("adaface_b200::attention_lse", mutates_args=())
def attention_lse(q: Tensor, k: Tensor, v: Tensor, heads: int, scale: float,
                  key_mask: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """`attention` that also returns the log2-domain log-sum-exp of every query row, fp32 [B, heads, Nq] - what the
    backward needs to rebuild the probabilities.  Used instead of `attention` while gradients are recorded."""
    _need_cuda(q, k, v)
    b, nq, c = q.shape
    nk = k.shape[1]
    if k.shape != v.shape or k.shape[0] != b or k.shape[2] != c or c % heads:
        raise ValueError("attention_lse: shape mismatch")
    out = torch.empty((b, nq, c), dtype=torch.bfloat16, device=q.device)
    lse = torch.empty((b, heads, nq), dtype=torch.float32, device=q.device)
    if key_mask is not None:
        key_mask = key_mask.reshape(b, nk).to(torch.uint8).contiguous()
    args = AttnArgs(q=_ptr(q), k=_ptr(k), v=_ptr(v), out=_ptr(out), key_mask=_ptr(key_mask), batch=b, heads=heads,
                    head_dim=c // heads, q_len=nq, kv_len=nk, ldq=_ld3(q, "q"), ldk=_ld3(k, "k"), ldv=_ld3(v, "v"),
                    ldo=c, scale=scale, lse=_ptr(lse))
    _run("attention_fwd", _lib.load().afb_attention_fwd, C.byref(args), _stream(), flops=4.0 * b * nq * nk * c,
         nbytes=2.0 * b * c * (2 * nq + 2 * nk), tag=f"b{b}q{nq}k{nk}d{c // heads}")
    return out, lse


@attention_lse.register_fake
def _(q, k, v, heads, scale, key_mask=None):
    return q.new_empty(q.shape, dtype=torch.bfloat16), q.new_empty((q.shape[0], heads, q.shape[1]), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::attention_bwd", mutates_args=())
def attention_bwd(q: Tensor, k: Tensor, v: Tensor, out: Tensor, lse: Tensor, dout: Tensor, heads: int,
                  scale: float, key_mask: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor]:
    """Backward of `attention_lse` for any key length (self-attention): -> (dq, dk, dv) fp32 [B, L, C].  key_mask
    (uint8 / bool [B, Nk], 0 = masked) as in the forward: masked keys pass no gradient to q / k (attention.py:223-232)."""
    _need_cuda(q, k, v, out, lse, dout)
    b, nq, c = q.shape
    nk = k.shape[1]
    out, dout, lse = out.contiguous(), dout.contiguous(), lse.contiguous()
    if dout.dtype != torch.bfloat16 or dout.shape != q.shape or out.shape != q.shape:
        raise ValueError("attention_bwd: out / dout must be bf16 of q's shape")
    dev = q.device
    dq = torch.empty((b, nq, c), dtype=torch.float32, device=dev)
    dk = torch.empty((b, nk, c), dtype=torch.float32, device=dev)
    dv = torch.empty((b, nk, c), dtype=torch.float32, device=dev)
    ws = torch.empty((b * heads * nq,), dtype=torch.float32, device=dev)
    if key_mask is not None:
        key_mask = key_mask.reshape(b, nk).to(torch.uint8).contiguous()
    args = AttnArgs(q=_ptr(q), k=_ptr(k), v=_ptr(v), out=_ptr(out), key_mask=_ptr(key_mask), batch=b, heads=heads,
                    head_dim=c // heads, q_len=nq, kv_len=nk, ldq=_ld3(q, "q"), ldk=_ld3(k, "k"), ldv=_ld3(v, "v"),
                    ldo=c, scale=scale, lse=_ptr(lse))
    _run("attention_bwd", _lib.load().afb_attention_bwd, C.byref(args), _ptr(dout), _ptr(ws), _ptr(dq), _ptr(dk), _ptr(dv), c,
         _stream(), flops=10.0 * b * nq * nk * c, tag=f"b{b}q{nq}k{nk}d{c // heads}")
    return dq, dk, dv


@attention_bwd.register_fake
def _(q, k, v, out, lse, dout, heads, scale, key_mask=None):
    f = lambda t: t.new_empty(t.shape, dtype=torch.float32)
    return f(q), f(k), f(v)


@torch.library.custom_op("adaface_b200::self_attention_qkv", mutates_args=())
def self_attention_qkv(qkv: Tensor, heads: int, scale: float, key_mask: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Self-attention on the fused projection buffer qkv bf16 [B, N, 3C] (q | k | v column blocks) -> (out bf16 [B, N, C],
    lse).  Same kernels as `attention_lse` on the three column slices; the point is the backward: ONE gradient tensor of
    qkv's shape written in place by the attention backward (row pitch 3C), where slicing would have autograd zero-fill
    three qkv-sized buffers, copy the three slices in and add them up."""
    c = qkv.shape[2] // 3
    return attention_lse(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], heads, scale, key_mask)


@self_attention_qkv.register_fake
def _(qkv, heads, scale, key_mask=None):
    b, n, c3 = qkv.shape
    return qkv.new_empty((b, n, c3 // 3), dtype=torch.bfloat16), qkv.new_empty((b, heads, n), dtype=torch.float32)


@torch.library.custom_op("adaface_b200::self_attention_qkv_bwd", mutates_args=())
def self_attention_qkv_bwd(qkv: Tensor, out: Tensor, lse: Tensor, dout: Tensor, heads: int, scale: float,
                           key_mask: Optional[Tensor] = None) -> Tensor:
    """Backward of `self_attention_qkv`: d_qkv bf16 [B, N, 3C] (dq | dk | dv written by afb_attention_bwd with ldg = 3C)."""
    _need_cuda(qkv, out, lse, dout)
    b, n, c3 = qkv.shape
    c = c3 // 3
    out, dout, lse = out.contiguous(), dout.contiguous(), lse.contiguous()
    if dout.dtype != torch.bfloat16 or dout.shape != (b, n, c) or out.shape != (b, n, c) or qkv.stride(2) != 1 or qkv.stride(1) < c3:
        raise ValueError("self_attention_qkv_bwd: bf16 out / dout [B, N, C] and a row-contiguous qkv [B, N, 3C] expected")
    dev = qkv.device
    d = torch.empty((b, n, c3), dtype=torch.float32, device=dev)
    ws = torch.empty((b * heads * n,), dtype=torch.float32, device=dev)
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    ld = _ld3(qkv, "qkv")
    base = qkv.data_ptr()
    args = AttnArgs(q=base, k=base + 2 * c, v=base + 4 * c, out=_ptr(out), key_mask=_ptr(key_mask), batch=b, heads=heads,
                    head_dim=c // heads, q_len=n, kv_len=n, ldq=ld, ldk=ld, ldv=ld, ldo=c, scale=scale, lse=_ptr(lse))
    _run("attention_bwd", _lib.load().afb_attention_bwd, C.byref(args), _ptr(dout), _ptr(ws), d.data_ptr(), d.data_ptr() + 4 * c,
         d.data_ptr() + 8 * c, c3, _stream(), flops=10.0 * b * n * n * c, tag=f"b{b}q{n}k{n}d{c // heads}")
    return d.to(torch.bfloat16)


@self_attention_qkv_bwd.register_fake
def _(qkv, out, lse, dout, heads, scale, key_mask=None):
    return torch.empty_like(qkv, memory_format=torch.contiguous_format)


def _self_attention_qkv_setup(ctx, inputs, output):
    qkv, heads, scale = inputs[:3]
    key_mask = inputs[3] if len(inputs) > 3 else None
    out, lse = output
    ctx.save_for_backward(qkv, out, lse, key_mask)
    ctx.heads, ctx.scale = heads, scale


def _self_attention_qkv_backward(ctx, g, g_lse):
    qkv, out, lse, key_mask = ctx.saved_tensors
    return self_attention_qkv_bwd(qkv, out, lse, _bf16c(g), ctx.heads, ctx.scale, key_mask).to(qkv.dtype), None, None, None


self_attention_qkv.register_autograd(_self_attention_qkv_backward, setup_context=_self_attention_qkv_setup)


def self_attention_auto(qkv: Tensor, heads: int, scale: float, key_mask: Optional[Tensor]) -> Tensor:
    """Self-attention on the fused q | k | v buffer: `attention` on its column slices, or - while gradients are recorded
    through it - `self_attention_qkv`, whose backward hands autograd one packed gradient."""
    c = qkv.shape[2] // 3
    if torch.is_grad_enabled() and qkv.requires_grad and qkv.shape[1] > 128:
        return self_attention_qkv(qkv, heads, scale, key_mask)[0]
    return attention_auto(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], heads, scale, key_mask)


def _attention_lse_setup(ctx, inputs, output):
    q, k, v, heads, scale = inputs[:5]
    key_mask = inputs[5] if len(inputs) > 5 else None
    out, lse = output
    ctx.save_for_backward(q, k, v, out, lse, key_mask)
    ctx.heads, ctx.scale = heads, scale


def _attention_lse_backward(ctx, g, g_lse):
    q, k, v, out, lse, key_mask = ctx.saved_tensors
    dq, dk, dv = attention_bwd(q, k, v, out, lse, _bf16c(g), ctx.heads, ctx.scale, key_mask)
    return dq.to(q.dtype), dk.to(k.dtype), dv.to(v.dtype), None, None, None


attention_lse.register_autograd(_attention_lse_backward, setup_context=_attention_lse_setup)


def attention_auto(q: Tensor, k: Tensor, v: Tensor, heads: int, scale: float, key_mask: Optional[Tensor]) -> Tensor:
    """`attention`, or - while gradients are recorded through a long key sequence - `attention_lse`, whose backward
    covers self-attention.  (The short-key backward of `attention` needs no saved statistics.)"""
    if torch.is_grad_enabled() and k.shape[1] > 128 and (q.requires_grad or k.requires_grad or v.requires_grad):
        return attention_lse(q, k, v, heads, scale, key_mask)[0]
    return attention(q, k, v, heads, scale, key_mask)


@torch.library.custom_op("adaface_b200::groupnorm_tokens_bwd", mutates_args=())
def groupnorm_tokens_bwd(x: Tensor, grad_y: Tensor, gamma: Tensor, groups: int, eps: float, beta: Optional[Tensor] = None,
                         act: int = 0, pre_add: Optional[Tensor] = None) -> Tensor:
    """Backward of `groupnorm_tokens` for a bf16 channels-last x [B,C,h,w] and grad_y bf16 [B, h*w, C] -> dx bf16
    channels-last [B,C,h,w]; act = 1 with beta: through the fused SiLU, pre_add: what the forward added to x."""
    _need_cuda(x, grad_y, gamma, beta, pre_add)
    b, c, h, w = x.shape
    if x.dtype != torch.bfloat16 or not x.permute(0, 2, 3, 1).is_contiguous():
        raise NotImplementedError("groupnorm_tokens_bwd: only the bf16 channels-last layout is built")
    grad_y = _bf16c(grad_y)
    dx = torch.empty_like(x)                                  # preserves channels-last
    f = lambda t: None if t is None else t.detach().float().contiguous()
    gamma, beta = f(gamma), f(beta)
    add_ld = 0
    if pre_add is not None:
        if pre_add.dtype != torch.float32 or pre_add.shape != (b, c) or pre_add.stride(1) != 1:
            raise ValueError("groupnorm_tokens_bwd: pre_add must be fp32 [B, C] with contiguous rows")
        add_ld = pre_add.stride(0)
    ws = torch.empty((b * groups * 4,), dtype=torch.float32, device=x.device)
    _run("groupnorm_tokens_bwd", _lib.load().afb_groupnorm_tokens_bwd, _ptr(x), _ptr(grad_y), _ptr(gamma), _ptr(beta), _ptr(pre_add),
         add_ld, act, _ptr(dx), _ptr(ws), b, c, h * w, groups, eps, _stream(), nbytes=6.0 * x.numel(), tag=f"b{b}n{h * w}C{c}")
    return dx


@groupnorm_tokens_bwd.register_fake
def _(x, grad_y, gamma, groups, eps, beta=None, act=0, pre_add=None):
    return torch.empty_like(x)


def _gn_setup(ctx, inputs, output):
    x, gamma, beta, groups, eps, act, pre_add = inputs
    _frozen(gamma, beta, pre_add)
    ctx.save_for_backward(x, gamma, beta, pre_add)
    ctx.groups, ctx.eps, ctx.act = groups, eps, act


def _gn_backward(ctx, g):
    x, gamma, beta, pre_add = ctx.saved_tensors
    return groupnorm_tokens_bwd(x, g, gamma, ctx.groups, ctx.eps, beta, ctx.act, pre_add), None, None, None, None, None, None


groupnorm_tokens.register_autograd(_gn_backward, setup_context=_gn_setup)


# 3x3 convolution against a frozen weight: dX = the same implicit GEMM with the transposed, spatially flipped weight
# (w [Cout, 9*Cin], k = (ky*3 + kx)*Cin + c  ->  [Cin, 9*Cout], k' = ((2-ky)*3 + (2-kx))*Cout + o); cached per weight
_CONV_DX_W = {}


def conv3x3_dx_weight(w: Tensor) -> Tensor:
    key = (w.data_ptr(), w._version, tuple(w.shape))
    hit = _CONV_DX_W.get(key)
    if hit is not None and hit[0]() is w:          # the same tensor object, not a new one that reuses the address
        return hit[1]
    cout, k9 = w.shape
    cin = k9 // 9
    with torch.no_grad():
        c = w.detach().view(cout, 3, 3, cin).flip(1, 2).permute(3, 1, 2, 0).reshape(cin, 9 * cout).contiguous()
    if len(_CONV_DX_W) > 256:
        _CONV_DX_W.clear()
    _CONV_DX_W[key] = (weakref.ref(w), c)
    return c


def _conv3x3_setup(ctx, inputs, output):
    x, w, bias, residual, h, width = inputs
    _frozen(w, bias)
    if w.shape[0] % 64:
        raise NotImplementedError("conv3x3 backward: Cout must be a multiple of 64 (it is the K dimension of the dX convolution)")
    ctx.save_for_backward(w)
    ctx.hw = (h, width)
    ctx.x_needs, ctx.res_needs = x.requires_grad, residual is not None and residual.requires_grad
    ctx.x_dtype, ctx.res_dtype = x.dtype, (None if residual is None else residual.dtype)


def _conv3x3_backward(ctx, g):
    (w,) = ctx.saved_tensors
    gb = _bf16c(g)
    dx = conv3x3(gb, conv3x3_dx_weight(w), None, None, ctx.hw[0], ctx.hw[1]).to(ctx.x_dtype) if ctx.x_needs else None
    dres = g.to(ctx.res_dtype) if ctx.res_needs else None
    return dx, None, None, dres, None, None


conv3x3.register_autograd(_conv3x3_backward, setup_context=_conv3x3_setup)


def _abr_setup(ctx, inputs, output):
    a, b, bias = inputs
    _frozen(bias)
    ctx.b_needs = b is not None and b.requires_grad
    ctx.a_needs = a.requires_grad


def _abr_backward(ctx, g):
    return (g if ctx.a_needs else None), (g if ctx.b_needs else None), None


add_bias_residual.register_autograd(_abr_backward, setup_context=_abr_setup)


@torch.library.custom_op("adaface_b200::ada_pool_bwd_features", mutates_args=())
def ada_pool_bwd_features(x: Tensor, q: Tensor, pool_q: Tensor, key_mask: Optional[Tensor], d_fg: Tensor,
                          d_bg: Tensor, drop_p: float = 0.0, seed: int = 0,
                          fwd_state: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """Backward of `ada_pool` towards the layer features: -> (dx, dq) bf16 [B, N, C] (`fwd_state` as in `ada_pool_bwd`)."""
    _need_cuda(x, q, pool_q, key_mask, d_fg, d_bg, fwd_state)
    b, n, c = x.shape
    lib = _lib.load()
    dev = x.device
    x, q = x.contiguous(), q.contiguous()
    pool_q = pool_q.detach().to(torch.float32).contiguous()
    if key_mask is not None:
        key_mask = key_mask.reshape(b, n).to(torch.uint8).contiguous()
    if fwd_state is None:
        fwd_state = ada_pool_fwd_state(x, q, pool_q, key_mask, drop_p, seed)
    stats, fg, bg = _split_fwd_state(fwd_state, b, c)
    d_fg, d_bg = d_fg.detach().to(torch.float32).contiguous(), d_bg.detach().to(torch.float32).contiguous()
    dx, dq = torch.empty_like(x), torch.empty_like(q)
    _run("ada_pool_bwd_features", lib.afb_ada_pool_bwd_features, _ptr(x), _ptr(q), _ptr(pool_q), _ptr(key_mask), _ptr(stats),
         _ptr(fg), _ptr(bg), _ptr(d_fg), _ptr(d_bg), _ptr(dx), _ptr(dq), b, n, c, float(drop_p), int(seed) & 0xFFFFFFFF,
         _stream(), nbytes=8.0 * b * n * c, tag=f"b{b}n{n}C{c}")
    return dx, dq


@ada_pool_bwd_features.register_fake
def _(x, q, pool_q, key_mask, d_fg, d_bg, drop_p=0.0, seed=0, fwd_state=None):
    return torch.empty_like(x), torch.empty_like(q)


def _ada_pool_setup2(ctx, inputs, output):
    x, q, pool_q, key_mask, want_attn = inputs[:5]
    ctx.drop_p, ctx.seed = (inputs[5], inputs[6]) if len(inputs) > 5 else (0.0, 0)
    ctx.save_for_backward(x, q, pool_q, key_mask)
    ctx.feat_needs, ctx.q_needs = x.requires_grad or q.requires_grad, pool_q.requires_grad


def _ada_pool_backward2(ctx, g_fg, g_bg, g_afg, g_abg):
    x, q, pool_q, key_mask = ctx.saved_tensors
    z = lambda g: torch.zeros((x.shape[0], x.shape[2]), dtype=torch.float32, device=x.device) if g is None else g
    g_fg, g_bg = z(g_fg), z(g_bg)
    state = ada_pool_fwd_state(x, q, pool_q.detach(), key_mask, ctx.drop_p, ctx.seed) if (ctx.q_needs or ctx.feat_needs) else None
    d_pq = ada_pool_bwd(x, q, pool_q, key_mask, g_fg, g_bg, ctx.drop_p, ctx.seed, state) if ctx.q_needs else None
    dx = dq = None
    if ctx.feat_needs:
        dx, dq = ada_pool_bwd_features(x, q, pool_q, key_mask, g_fg, g_bg, ctx.drop_p, ctx.seed, state)
    return dx, dq, d_pq, None, None, None, None


ada_pool.register_autograd(_ada_pool_backward2, setup_context=_ada_pool_setup2)
