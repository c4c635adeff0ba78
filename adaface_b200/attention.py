"""Host-side mirror of the reference's transformer modules for the U-Net hot path
(reference: ldm/modules/attention.py - GEGLU :32, FeedForward :42, Normalize :71, CrossAttention :147,
BasicTransformerBlock :260, SpatialTransformer :287).

Same constructor arguments, forward signatures, externally visible attributes (`heads`, `scale`,
`infeat_size`, `use_conv_attn_kernel_size`, `save_attn_vars`, `cached_activations`, `is_training`,
`save_feat`/`feat`, `in_channels`) and state_dict keys as the reference, so an SD-1.5 checkpoint loads
unchanged and `openaimodel.py` can construct these classes in place of its own.

All arithmetic is done by the sm_100a kernels behind the C ABI (adaface_b200/ops.py): bf16 operands,
fp32 accumulation / statistics.  There is no torch-eager or CPU fallback; a missing CUDA library or a
CPU tensor raises.  PyTorch holds the parameters and the device buffers.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from . import ops

Tensor = torch.Tensor


def exists(val) -> bool:
    return val is not None


def default(val, d):
    return val if exists(val) else (d() if callable(d) else d)


def zero_module(module: nn.Module) -> nn.Module:
    """attention.py:62-68: the reference zero-initialises SpatialTransformer.proj_out."""
    for p in module.parameters():
        p.detach().zero_()
    return module


def Normalize(in_channels: int) -> nn.GroupNorm:
    """attention.py:71-72.  Only the parameters live here; the arithmetic is ops.groupnorm_tokens."""
    return nn.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6, affine=True)


class _PackedWeights:
    """Kernel-layout (bf16, fused / interleaved) copies of a module's fp32 parameters, rebuilt whenever a
    parameter is replaced, modified in place, moved or cast (tracked by data_ptr and tensor version)."""

    def _pack_key(self) -> Tuple:
        return tuple((p.data_ptr(), p._version, p.dtype) for p in self.parameters(recurse=True))

    def packed(self) -> Dict[str, Tensor]:
        key = self._pack_key()
        cache = self.__dict__.get("_pack_cache")
        if cache is None or cache[0] != key:
            with torch.no_grad():
                cache = (key, self._build_pack())
            self.__dict__["_pack_cache"] = cache
        return cache[1]

    def _build_pack(self) -> Dict[str, Tensor]:      # pragma: no cover - overridden
        raise NotImplementedError


def _bf16(t: Tensor) -> Tensor:
    return t.detach().to(torch.bfloat16).contiguous()


def _f32(t: Optional[Tensor]) -> Optional[Tensor]:
    return None if t is None else t.detach().to(torch.float32).contiguous()


def as_bf16_tokens(x: Tensor) -> Tensor:
    """[..., C] activations -> contiguous bf16 (the cast kernel when fp32 comes in)."""
    if not x.is_cuda:
        raise ops._lib.AdafaceB200Error("adaface_b200 modules run on CUDA tensors only (no CPU fallback)")
    if x.dtype == torch.bfloat16:
        return x.contiguous()
    if x.dtype != torch.float32:
        x = x.float()
    return ops.to_bf16(x)


def like_input(y_bf16: Tensor, ref: Tensor) -> Tensor:
    """Returns the bf16 kernel output in the caller's dtype (fp32 in the oracle configs)."""
    if ref.dtype == torch.bfloat16:
        return y_bf16
    y = ops.to_f32(y_bf16)
    return y if ref.dtype == torch.float32 else y.to(ref.dtype)


def key_mask_from_image_mask(mask: Optional[Tensor], b: int, n: int) -> Optional[Tensor]:
    """attention.py:223-232: mask [B, ...] -> key mask uint8 [B, N] (non-zero = keep)."""
    if mask is None:
        return None
    m = mask.reshape(mask.shape[0], -1)
    if m.shape != (b, n):
        raise ValueError(f"attention mask of shape {tuple(mask.shape)} does not flatten to [{b}, {n}]")
    return m.bool().to(torch.uint8).contiguous()


def resolve_context(context, layer_attn_components=None):
    """attention.py:184-193.  A callable context returns (context, placeholder2indices); a tuple context is
    (v_context, k_context).  Extension used by the AdaEmbedding wiring: a callable with the
    attribute `wants_layer_attn_components = True` is called with {'x', 'q', 'to_k', 'infeat_size'}."""
    placeholder2indices = None
    if callable(context):
        if getattr(context, "wants_layer_attn_components", False):
            context, placeholder2indices = context(layer_attn_components)
        else:
            context, placeholder2indices = context()
    if isinstance(context, (list, tuple)):
        v_context, k_context = context
    else:
        v_context = k_context = context
    return v_context, k_context, placeholder2indices


class GEGLU(nn.Module, _PackedWeights):
    """attention.py:32-39: x, gate = proj(x).chunk(2); x * gelu(gate) - one GEMM with a gated epilogue."""

    def __init__(self, dim_in: int, dim_out: int):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def _build_pack(self):
        w, b = ops.interleave_geglu(_bf16(self.proj.weight), _f32(self.proj.bias))
        return {"w": w, "b": b}

    def forward_tokens(self, x2d: Tensor) -> Tensor:
        p = self.packed()
        return ops.linear_geglu(x2d, p["w"], p["b"])

    def forward(self, x: Tensor) -> Tensor:
        y = self.forward_tokens(as_bf16_tokens(x).view(-1, x.shape[-1]))
        return like_input(y, x).view(*x.shape[:-1], y.shape[-1])


class FeedForward(nn.Module, _PackedWeights):
    """attention.py:42-59.  SD-1.5 uses glu=True; the plain Linear+GELU variant is not on the path."""

    def __init__(self, dim: int, dim_out: Optional[int] = None, mult: int = 4, glu: bool = False, dropout: float = 0.):
        super().__init__()
        if not glu:
            raise NotImplementedError("adaface_b200.FeedForward: only the gated (GEGLU) variant used by SD-1.5 is built")
        inner_dim = int(dim * mult)
        dim_out = default(dim_out, dim)
        self.net = nn.Sequential(GEGLU(dim, inner_dim), nn.Dropout(dropout), nn.Linear(inner_dim, dim_out))

    def _build_pack(self):
        return {"w2": _bf16(self.net[2].weight), "b2": _f32(self.net[2].bias)}

    def _check_dropout(self):
        if self.training and self.net[1].p > 0:
            raise NotImplementedError("adaface_b200: dropout > 0 in training mode is not supported on the kernel path")

    def forward_tokens(self, x2d: Tensor, residual: Optional[Tensor] = None) -> Tensor:
        """x2d bf16 [M, C] -> ff(x) (+ residual) bf16 [M, C]."""
        self._check_dropout()
        p = self.packed()
        g = self.net[0].forward_tokens(x2d)
        return ops.linear(g, p["w2"], p["b2"], residual, False)

    def forward(self, x: Tensor) -> Tensor:
        y = self.forward_tokens(as_bf16_tokens(x).view(-1, x.shape[-1]))
        return like_input(y, x).view(x.shape[:-1] + (y.shape[-1],))


class CrossAttention(nn.Module, _PackedWeights):
    """attention.py:147-257.  `forward(x, context=None, mask=None)`; context may be None (self-attention), a
    tensor, a (v_context, k_context) tuple or a callable returning (context, placeholder2indices)."""

    def __init__(self, query_dim: int, context_dim: Optional[int] = None, heads: int = 8, dim_head: int = 64,
                 dropout: float = 0.):
        super().__init__()
        inner_dim = dim_head * heads
        context_dim = default(context_dim, query_dim)
        self.scale = dim_head ** -0.5
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner_dim, bias=False)
        self.to_k = nn.Linear(context_dim, inner_dim, bias=False)
        self.to_v = nn.Linear(context_dim, inner_dim, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner_dim, query_dim), nn.Dropout(dropout))
        self.save_attn_vars = False
        self.cached_activations = None
        self.use_conv_attn_kernel_size = -1
        self.infeat_size = None
        self.is_training = True

    def _build_pack(self):
        wq, wk, wv = _bf16(self.to_q.weight), _bf16(self.to_k.weight), _bf16(self.to_v.weight)
        pack = {"wq": wq, "wk": wk, "wv": wv, "wkv": torch.cat([wk, wv], dim=0).contiguous(),
                "wo": _bf16(self.to_out[0].weight), "bo": _f32(self.to_out[0].bias)}
        if wq.shape[1] == wk.shape[1]:
            pack["wqkv"] = torch.cat([wq, wk, wv], dim=0).contiguous()
        return pack

    def _check_modes(self, placeholder2indices, context_provided: bool):
        if self.training and self.to_out[1].p > 0:
            raise NotImplementedError("adaface_b200: dropout > 0 in training mode is not supported on the kernel path")
        if self.save_attn_vars and not context_provided:
            # attention.py:245-255 on a self-attention layer would cache a [B, 8, N, N] matrix (8.6 GB at 64x64);
            # the reference only ever reads the cross-attention layers' activations (openaimodel.py:985-1000)
            raise NotImplementedError("adaface_b200: save_attn_vars is built for cross-attention layers only")

    def _conv_attn_subjects(self, placeholder2indices):
        """attention.py:208-216: [(inst_batch int32 [I], tok int32 [I, ks*ks])] per subject string, or []."""
        ks = int(self.use_conv_attn_kernel_size or -1)
        if placeholder2indices is None or ks <= 1:          # ks == 1 is a no-op in the reference (ldm/util.py:720)
            return []
        if ks > 4:
            raise ValueError("use_conv_attn_kernel_size must be 2, 3 or 4 (ldm/util.py:735-744)")
        out = []
        for subj_string in placeholder2indices:
            idx_b, idx_n = [torch.as_tensor(t) for t in placeholder2indices[subj_string]]
            n_inst = int(torch.unique(idx_b).numel())
            per = idx_b.numel() // n_inst
            if per < ks * ks:
                raise ValueError(f"conv attention with ks={ks} needs {ks * ks} subject tokens, the prompt has {per}")
            inst = idx_b.reshape(n_inst, per)[:, 0]
            tok = idx_n.reshape(n_inst, per)[:, :ks * ks]
            out.append((inst.to(torch.int32), tok.to(torch.int32).contiguous()))
        return out

    # ---- fused token-level paths used by BasicTransformerBlock (bf16 [B, N, C] in, bf16 [B*N, C] out) ----
    def self_attention_tokens(self, x: Tensor, key_mask: Optional[Tensor], residual: Optional[Tensor]) -> Tensor:
        """to_out(softmax(q k^T) v) (+ residual) with q, k, v from ONE fused [3C, C] projection."""
        self._check_modes(None, False)
        b, n, c = x.shape
        p = self.packed()
        inner = p["wq"].shape[0]
        qkv = ops.linear(x.view(b * n, c), p["wqkv"], None, None, False).view(b, n, 3 * inner)
        a = ops.self_attention_auto(qkv, self.heads, self.scale, key_mask)
        return ops.linear(a.view(b * n, inner), p["wo"], p["bo"], residual, False)

    def out_projection(self, a2d: Tensor, residual: Optional[Tensor], want_stats: bool = False):
        """to_out[0] (+ residual); with want_stats also the row partials a LayerNorm-folded consumer needs."""
        p = self.packed()
        if want_stats:
            return ops.linear_stats(a2d, p["wo"], p["bo"], residual)
        return ops.linear(a2d, p["wo"], p["bo"], residual, False)

    def attend_fused_qkv(self, qkv: Tensor, key_mask: Optional[Tensor]) -> Tensor:
        """qkv bf16 [B, N, 3*inner] (q | k | v column blocks) -> softmax(q k^T * scale) v as [B*N, inner]."""
        self._check_modes(None, False)
        b, n, w3 = qkv.shape
        inner = w3 // 3
        a = ops.self_attention_auto(qkv, self.heads, self.scale, key_mask)
        return a.view(b * n, inner)

    def project_query(self, x: Tensor) -> Tensor:
        b, n, c = x.shape
        return ops.linear(x.view(b * n, c), self.packed()["wq"], None, None, False).view(b, n, -1)

    def cross_attention_tokens(self, q: Tensor, v_context: Tensor, k_context: Tensor,
                               residual: Optional[Tensor], want_stats: bool = False, placeholder2indices=None):
        """q bf16 [B, N, C] (already projected); contexts [B, T, ctx_dim] -> to_out(attn) (+ residual)."""
        b, n, inner = q.shape
        p = self.packed()
        if v_context is k_context:
            ctx = as_bf16_tokens(k_context)
            t = ctx.shape[1]
            kv = ops.linear(ctx.view(b * t, ctx.shape[2]), p["wkv"], None, None, False).view(b, t, 2 * inner)
            k, v = kv[..., :inner], kv[..., inner:]
        else:
            kc, vc = as_bf16_tokens(k_context), as_bf16_tokens(v_context)
            k = ops.linear(kc.view(-1, kc.shape[2]), p["wk"], None, None, False).view(b, kc.shape[1], inner)
            v = ops.linear(vc.view(-1, vc.shape[2]), p["wv"], None, None, False).view(b, vc.shape[1], inner)
        subjects = self._conv_attn_subjects(placeholder2indices)
        if self.save_attn_vars or subjects:
            # optional modes that need the score matrix itself: explicit scores -> row replacement -> softmax . v
            sim = ops.cross_scores(q, k, self.heads, self.scale)                                # :199
            for inst, tok in subjects:                                                          # :208-216
                ops.conv_attn_rows(sim, q, k, inst.to(q.device), tok.to(q.device), self.infeat_size[0],
                                   self.infeat_size[1], int(self.use_conv_attn_kernel_size), self.scale)
            a, attn = ops.softmax_pv(sim, v, self.save_attn_vars)                               # :238-242
            if self.save_attn_vars:                                                             # :245-255
                d = inner // self.heads
                self.cached_activations = {
                    "q": q.view(b, n, self.heads, d).permute(0, 2, 1, 3).float() * math.sqrt(self.scale),
                    "attn": attn, "attnscore": sim}
        else:
            a = ops.attention_auto(q, k, v, self.heads, self.scale, None)
        return self.out_projection(a.view(b * n, inner), residual, want_stats)

    def forward(self, x: Tensor, context=None, mask: Optional[Tensor] = None) -> Tensor:
        b, n, c = x.shape
        xb = as_bf16_tokens(x)
        if not exists(context):
            out = self.self_attention_tokens(xb, key_mask_from_image_mask(mask, b, n), None)
        else:
            q = self.project_query(xb)
            comps = {"x": xb, "q": q, "to_k": self.to_k, "infeat_size": self.infeat_size}
            v_context, k_context, placeholder2indices = resolve_context(context, comps)
            self._check_modes(placeholder2indices, True)
            if exists(mask):
                # the reference would mask whole key columns of the 77-token context (NaN hazard noted at
                # attention.py:218-222); a key mask is only meaningful for self-attention
                raise ValueError("CrossAttention: mask is only supported for self-attention (context=None)")
            out = self.cross_attention_tokens(q, v_context, k_context, None, False, placeholder2indices)
        return like_input(out, x).view(b, n, -1)


class BasicTransformerBlock(nn.Module, _PackedWeights):
    """attention.py:260-285.  Gradient checkpointing is a no-op in the reference."""

    def __init__(self, dim: int, n_heads: int, d_head: int, dropout: float = 0., context_dim: Optional[int] = None,
                 gated_ff: bool = True, checkpoint: bool = True):
        super().__init__()
        self.attn1 = CrossAttention(query_dim=dim, heads=n_heads, dim_head=d_head, dropout=dropout)
        self.ff = FeedForward(dim, dropout=dropout, glu=gated_ff)
        self.attn2 = CrossAttention(query_dim=dim, context_dim=context_dim, heads=n_heads, dim_head=d_head,
                                    dropout=dropout)
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.norm3 = nn.LayerNorm(dim)
        self.checkpoint = checkpoint

    def _pack_key(self):
        ps = [p for n in (self.norm1, self.norm2, self.norm3) for p in n.parameters()]
        ps += [self.attn1.to_q.weight, self.attn1.to_k.weight, self.attn1.to_v.weight, self.attn2.to_q.weight]
        ps += list(self.ff.net[0].proj.parameters())
        return tuple((p.data_ptr(), p._version, p.dtype) for p in ps)

    @staticmethod
    def _fold(weight: Tensor, bias: Optional[Tensor], norm: nn.LayerNorm):
        """LayerNorm folded into the Linear that consumes it:  LN(x) W^T + b = rstd (x W'^T - mean s) + t with
        W' = W diag(gamma) (bf16, what the tensor cores multiply), s = row sums of that bf16 W', t = W beta + b."""
        w = weight.detach().float()
        wg = (w * norm.weight.detach().float()[None, :]).to(torch.bfloat16).contiguous()
        t = w @ norm.bias.detach().float()
        if bias is not None:
            t = t + bias.detach().float()
        return wg, t.contiguous()

    def _build_pack(self):
        out = {}
        for i, n in enumerate((self.norm1, self.norm2, self.norm3), 1):
            out[f"g{i}"], out[f"b{i}"] = _f32(n.weight), _f32(n.bias)
        a1 = self.attn1
        if a1.to_q.weight.shape[1] == a1.to_k.weight.shape[1]:
            wqkv = torch.cat([a1.to_q.weight, a1.to_k.weight, a1.to_v.weight], dim=0)
            out["qkv_w"], out["qkv_t"] = self._fold(wqkv, None, self.norm1)
            out["qkv_s"] = out["qkv_w"].float().sum(dim=1).contiguous()
        out["q2_w"], out["q2_t"] = self._fold(self.attn2.to_q.weight, None, self.norm2)
        out["q2_s"] = out["q2_w"].float().sum(dim=1).contiguous()
        proj = self.ff.net[0].proj
        wg, t = self._fold(proj.weight, proj.bias, self.norm3)
        out["ff_w"], out["ff_t"] = ops.interleave_geglu(wg, t)
        out["ff_s"] = out["ff_w"].float().sum(dim=1).contiguous()
        return out

    def forward_tokens(self, x: Tensor, context=None, key_mask: Optional[Tensor] = None,
                       x_stats: Optional[Tensor] = None) -> Tensor:
        """x bf16 [B, N, C] -> bf16 [B, N, C].  Residual adds ride in the GEMM epilogues; the three LayerNorms are
        folded into the GEMMs that consume them: the GEMM that PRODUCES the residual stream emits per-row sum /
        sum-of-squares partials in its epilogue (`x_stats` for the incoming x, from proj_in), and the consumer applies
        rstd * (acc - mean * colsum) + t.  No LayerNorm kernel, no normalised copy of the activations."""
        b, n, c = x.shape
        p = self.packed()
        x2d = x.view(b * n, c)
        a1, a2 = self.attn1, self.attn2
        # ---- x1 = attn1(norm1(x)) + x                                                             :277
        if x_stats is not None and "qkv_w" in p:
            qkv = ops.linear_ln(x2d, x_stats, p["qkv_w"], p["qkv_s"], p["qkv_t"], self.norm1.eps)
        else:
            n1 = ops.layernorm(x2d, p["g1"], p["b1"], self.norm1.eps)
            qkv = ops.linear(n1, a1.packed()["wqkv"], None, None, False)
        x1, st1 = a1.out_projection(a1.attend_fused_qkv(qkv.view(b, n, -1), key_mask), x2d, True)
        # ---- x2 = attn2(norm2(x1), context) + x1                                                  :280-281
        if exists(context):
            q = ops.linear_ln(x1, st1, p["q2_w"], p["q2_s"], p["q2_t"], self.norm2.eps).view(b, n, -1)
            comps = {"x": x1.view(b, n, c), "q": q, "to_k": a2.to_k, "infeat_size": a2.infeat_size}
            v_context, k_context, placeholder2indices = resolve_context(context, comps)
            a2._check_modes(placeholder2indices, True)
            x2, st2 = a2.cross_attention_tokens(q, v_context, k_context, x1, True, placeholder2indices)
        else:                                                                                   # attn2 degenerates to self-attention
            n2 = ops.layernorm(x1, p["g2"], p["b2"], self.norm2.eps)
            qkv2 = ops.linear(n2, a2.packed()["wqkv"], None, None, False)
            x2, st2 = a2.out_projection(a2.attend_fused_qkv(qkv2.view(b, n, -1), None), x1, True)
        # ---- ff(norm3(x2)) + x2                                                                   :283
        self.ff._check_dropout()
        g = ops.linear_ln_geglu(x2, st2, p["ff_w"], p["ff_s"], p["ff_t"], self.norm3.eps)
        fp = self.ff.packed()
        return ops.linear(g, fp["w2"], fp["b2"], x2, False).view(b, n, c)

    def forward(self, x: Tensor, context=None, mask: Optional[Tensor] = None) -> Tensor:
        b, n, _ = x.shape
        y = self.forward_tokens(as_bf16_tokens(x), context, key_mask_from_image_mask(mask, b, n))
        return like_input(y, x)

    _forward = forward


class SpatialTransformer(nn.Module, _PackedWeights):
    """attention.py:287-341.  Two activation formats are consumed in place:
      * fp32 NCHW-contiguous (the reference's own format): GroupNorm reads NCHW and emits tokens, proj_out's
        epilogue transposes back and adds x_in in fp32;
      * bf16 channels-last (what the bf16 U-Net shell of adaface_b200.unet runs in): tokens ARE the memory
        layout, so no transposition happens at either end.
    """

    def __init__(self, in_channels: int, n_heads: int, d_head: int, depth: int = 1, dropout: float = 0.,
                 context_dim: Optional[int] = None):
        super().__init__()
        self.in_channels = in_channels
        inner_dim = n_heads * d_head
        self.norm = Normalize(in_channels)
        self.proj_in = nn.Conv2d(in_channels, inner_dim, kernel_size=1, stride=1, padding=0)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner_dim, n_heads, d_head, dropout=dropout, context_dim=context_dim)
             for _ in range(depth)])
        self.proj_out = zero_module(nn.Conv2d(inner_dim, in_channels, kernel_size=1, stride=1, padding=0))
        self.save_feat = False

    def _pack_key(self):
        mods = (self.norm, self.proj_in, self.proj_out)
        return tuple((p.data_ptr(), p._version, p.dtype) for m in mods for p in m.parameters())

    def _build_pack(self):
        wi, wo = self.proj_in.weight, self.proj_out.weight
        return {"gn_g": _f32(self.norm.weight), "gn_b": _f32(self.norm.bias),
                "wi": _bf16(wi.reshape(wi.shape[0], wi.shape[1])), "bi": _f32(self.proj_in.bias),
                "wo": _bf16(wo.reshape(wo.shape[0], wo.shape[1])), "bo": _f32(self.proj_out.bias)}

    def forward(self, x: Tensor, context=None, mask: Optional[Tensor] = None) -> Tensor:
        if x.dim() != 4:
            raise ValueError("SpatialTransformer expects x of shape [B, C, H, W]")
        b, c, h, w = x.shape
        n = h * w
        p = self.packed()
        channels_last_bf16 = x.dtype == torch.bfloat16 and x.permute(0, 2, 3, 1).is_contiguous()
        if channels_last_bf16:
            x_in = x
        else:
            x_in = x.float().contiguous()                                                        # fp32 NCHW
        t = ops.groupnorm_tokens(x_in, p["gn_g"], p["gn_b"], self.norm.num_groups, self.norm.eps)   # :325
        t, t_stats = ops.linear_stats(t.view(b * n, c), p["wi"], p["bi"], None)                  # :326-327
        inner = t.shape[1]
        t = t.view(b, n, inner)
        key_mask = None
        if exists(mask):                                                                         # :332
            mask2 = F.interpolate(mask.float(), size=(h, w), mode="nearest")
            key_mask = key_mask_from_image_mask(mask2, b, n)
        for block in self.transformer_blocks:
            block.attn2.infeat_size = (h, w)                                                     # :330
            t = block.forward_tokens(t, context, key_mask, t_stats)
            t_stats = None                                   # deeper blocks (depth > 1) normalise their own input
        if self.save_feat:                                                                       # :337-338
            self.feat = t.view(b, h, w, inner).permute(0, 3, 1, 2)
        t2d = t.view(b * n, inner)
        if channels_last_bf16:
            res = x.permute(0, 2, 3, 1).reshape(b * n, c)
            y = ops.linear(t2d, p["wo"], p["bo"], res, False)                                    # :340-341
            return y.view(b, h, w, c).permute(0, 3, 1, 2)
        y = ops.linear_nchw(t2d, p["wo"], p["bo"], x_in)                                         # :335-341
        return y if x.dtype == torch.float32 else y.to(x.dtype)
