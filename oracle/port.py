"""CPU oracle ("port") for the AdaPrompt U-Net transformer hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
may import it, and there only as the checker / the timed CPU baseline.  The product path
(``adaface_b200``) never imports this module and fails loudly without its CUDA library.

What it is: a plain fp32 (optionally fp64) torch-on-CPU restatement, in functional form over a
``state_dict``-style mapping, of the reference algorithms on the path.  Every function cites the
reference file:line it follows (paths relative to the reference root, zyt334/AdaFace).

Pinning: the reference has no tests and no golden vectors for this path, so the
port is pinned against outputs of the *reference modules themselves*, imported unmodified in the
build container by ``oracle/gen_golden.py`` and committed as ``tests/golden/*.npz``;
``tests/test_oracle_golden.py`` replays them.  The reference is Python and cannot travel to the GPU
box, which is why the fixtures are committed together with their generating script.
"""
from __future__ import annotations

import math
from typing import Callable, Mapping, Optional, Sequence, Tuple, Union

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor
Params = Mapping[str, Tensor]

# U-Net block index -> cross-attention layer slot (openaimodel.py:876-877, embedding_manager.py:554-555)
LAYER_IDX2CA = {1: 0, 2: 1, 4: 2, 5: 3, 7: 4, 8: 5, 12: 6, 16: 7,
                17: 8, 18: 9, 19: 10, 20: 11, 21: 12, 22: 13, 23: 14, 24: 15}
CA_INFEAT_DIMS = [320, 320, 640, 640, 1280, 1280, 1280, 1280,
                  1280, 1280, 640, 640, 640, 320, 320, 320]      # embedding_manager.py:551-552


# --------------------------------------------------------------------------------------------
# Norms (attention.py:71-72 GroupNorm(32, C, eps=1e-6); attention.py:267-269 LayerNorm eps=1e-5)
# --------------------------------------------------------------------------------------------
def group_norm(x: Tensor, weight: Tensor, bias: Tensor, groups: int = 32, eps: float = 1e-6) -> Tensor:
    b, c = x.shape[:2]
    xg = x.reshape(b, groups, -1)
    mean = xg.mean(dim=2, keepdim=True)
    var = xg.var(dim=2, unbiased=False, keepdim=True)
    xn = ((xg - mean) * torch.rsqrt(var + eps)).reshape(x.shape)
    shape = (1, c) + (1,) * (x.dim() - 2)
    return xn * weight.reshape(shape) + bias.reshape(shape)


def layer_norm(x: Tensor, weight: Optional[Tensor] = None, bias: Optional[Tensor] = None,
               eps: float = 1e-5) -> Tensor:
    mean = x.mean(dim=-1, keepdim=True)
    var = x.var(dim=-1, unbiased=False, keepdim=True)
    y = (x - mean) * torch.rsqrt(var + eps)
    if weight is not None:
        y = y * weight
    if bias is not None:
        y = y + bias
    return y


def gelu_erf(x: Tensor) -> Tensor:
    """Exact (erf) GELU: F.gelu default, attention.py:39."""
    return 0.5 * x * (1.0 + torch.erf(x * (1.0 / math.sqrt(2.0))))


# --------------------------------------------------------------------------------------------
# CrossAttention.forward (attention.py:172-257)
# --------------------------------------------------------------------------------------------
def resolve_context(context):
    """attention.py:184-193: callable -> (context, placeholder2indices); tuple -> (v_ctx, k_ctx)."""
    placeholder2indices = None
    if callable(context):
        context, placeholder2indices = context()
    if isinstance(context, (list, tuple)):
        v_context, k_context = context
    else:
        v_context = k_context = context
    return v_context, k_context, placeholder2indices


def attention_scores(q: Tensor, k: Tensor, heads: int, scale: float) -> Tensor:
    """sim = q k^T * scale with heads folded into batch: attention.py:198-199. Returns [B,h,Nq,Nk]."""
    b, nq, c = q.shape
    d = c // heads
    qh = q.reshape(b, nq, heads, d).permute(0, 2, 1, 3)
    kh = k.reshape(b, k.shape[1], heads, d).permute(0, 2, 1, 3)
    return torch.matmul(qh, kh.transpose(-1, -2)) * scale


def cross_attention(x: Tensor, p: Params, prefix: str, context=None, mask: Optional[Tensor] = None,
                    heads: int = 8, return_vars: bool = False):
    """attention.py:172-257.  `mask` is the key mask [B,1,h,w] (self-attention only, :223-232).

    Conv-attention row replacement (attention.py:208-216) is handled by `replace_rows_by_conv_attn`
    when the caller passes `conv_attn=(placeholder2indices, infeat_size, ks)` through `context`'s
    callable; the default path (ks <= 0) never calls it.
    """
    wq, wk, wv = p[prefix + "to_q.weight"], p[prefix + "to_k.weight"], p[prefix + "to_v.weight"]
    wo, bo = p[prefix + "to_out.0.weight"], p[prefix + "to_out.0.bias"]
    q = x @ wq.t()                                             # :176
    if context is None:
        v_context = k_context = x                              # :178-182
    else:
        v_context, k_context, _ = resolve_context(context)
    k = k_context @ wk.t()                                     # :195
    v = v_context @ wv.t()                                     # :196
    b, nq, c = q.shape
    d = c // heads
    scale = d ** -0.5                                          # :153
    sim = attention_scores(q, k, heads, scale)                 # [B,h,Nq,Nk]
    if mask is not None:                                       # :223-232
        keep = mask.reshape(b, -1).bool()[:, None, None, :]
        sim = sim.masked_fill(~keep, -torch.finfo(sim.dtype).max)
    attn = sim.softmax(dim=-1)                                 # :238
    vh = v.reshape(b, v.shape[1], heads, d).permute(0, 2, 1, 3)
    out = torch.matmul(attn, vh)                               # :240
    out = out.permute(0, 2, 1, 3).reshape(b, nq, c)            # :242
    out = out @ wo.t() + bo                                    # :243
    if return_vars:                                            # :245-255
        qh = q.reshape(b, nq, heads, d).permute(0, 2, 1, 3)
        return out, {"q": qh * math.sqrt(scale), "attn": attn, "attnscore": sim}
    return out


def geglu_feedforward(x: Tensor, p: Params, prefix: str) -> Tensor:
    """FeedForward with GEGLU: attention.py:32-59 (Dropout p=0)."""
    h = x @ p[prefix + "net.0.proj.weight"].t() + p[prefix + "net.0.proj.bias"]
    a, g = h.chunk(2, dim=-1)                                  # :38
    h = a * gelu_erf(g)                                        # :39
    return h @ p[prefix + "net.2.weight"].t() + p[prefix + "net.2.bias"]


def basic_transformer_block(x: Tensor, p: Params, prefix: str, context=None,
                            mask: Optional[Tensor] = None, heads: int = 8) -> Tensor:
    """BasicTransformerBlock._forward: attention.py:275-285."""
    n1 = layer_norm(x, p[prefix + "norm1.weight"], p[prefix + "norm1.bias"])
    x1 = cross_attention(n1, p, prefix + "attn1.", None, mask, heads) + x          # :277
    n2 = layer_norm(x1, p[prefix + "norm2.weight"], p[prefix + "norm2.bias"])
    x2 = x1 + cross_attention(n2, p, prefix + "attn2.", context, None, heads)      # :280-281
    n3 = layer_norm(x2, p[prefix + "norm3.weight"], p[prefix + "norm3.bias"])
    return geglu_feedforward(n3, p, prefix + "ff.") + x2                           # :283


def spatial_transformer(x: Tensor, p: Params, context=None, mask: Optional[Tensor] = None,
                        heads: int = 8, prefix: str = "", depth: int = 1) -> Tensor:
    """SpatialTransformer.forward: attention.py:321-341."""
    b, c, h, w = x.shape
    x_in = x
    y = group_norm(x, p[prefix + "norm.weight"], p[prefix + "norm.bias"])          # :325
    wi = p[prefix + "proj_in.weight"].reshape(c, c)
    y = y.reshape(b, c, h * w).transpose(1, 2) @ wi.t() + p[prefix + "proj_in.bias"]   # :326-327
    mask2 = None
    if mask is not None:                                                           # :332
        mask2 = F.interpolate(mask, size=(h, w), mode="nearest")
    for i in range(depth):
        y = basic_transformer_block(y, p, f"{prefix}transformer_blocks.{i}.", context, mask2, heads)
    wo = p[prefix + "proj_out.weight"].reshape(c, c)
    y = y @ wo.t() + p[prefix + "proj_out.bias"]                                   # :335-340
    return y.transpose(1, 2).reshape(b, c, h, w) + x_in                            # :341


def spatial_transformer_ada(x: Tensor, p: Params, context, ada_subjects, ca_layer_idx: int, time_emb: Tensor,
                            heads: int = 8, prefix: str = "", io: Optional[dict] = None) -> Tensor:
    """SpatialTransformer.forward (attention.py:321-341, depth 1) with the per-layer AdaEmbedding call site composed in.

    The reference snapshot defines AdaEmbedding but has no live call site for it (SURVEY.md finding F1: the
    `get_layer_ada_conditioning` hook at embedding_manager.py:1754 was removed), so this composition is the BUILD'S
    RECONSTRUCTION of SURVEY.md 3.2, not a restatement of reference lines: after attn2's query projection
    (attention.py:176) and before its key / value projections (:195-196), every subject's AdaEmbedding
    (`ada_embedding`, embedding_manager.py:779-910) turns this layer's (x1, q) into K vectors that overwrite the subject's
    token rows of this layer's context for the first `n_instances` samples (the cond half of the CFG batch, F8).

    ada_subjects: sequence of dicts {params, positions [K], probes [16,768], extra_means [16,n,768], K, r, n_instances,
    cached_bg_from (optional index of the subject whose bg pooling is reused)}.  `context` as for spatial_transformer
    (tensor, tuple or callable); a shared (not split) context is assumed."""
    b, c, h, w = x.shape
    pre = prefix + "transformer_blocks.0."
    t = group_norm(x, p[prefix + "norm.weight"], p[prefix + "norm.bias"])                       # :325
    t = t.reshape(b, c, h * w).transpose(1, 2) @ p[prefix + "proj_in.weight"].reshape(c, c).t() + p[prefix + "proj_in.bias"]
    n1 = layer_norm(t, p[pre + "norm1.weight"], p[pre + "norm1.bias"])
    x1 = cross_attention(n1, p, pre + "attn1.", None, None, heads) + t                          # :277
    n2 = layer_norm(x1, p[pre + "norm2.weight"], p[pre + "norm2.bias"])
    q = n2 @ p[pre + "attn2.to_q.weight"].t()                                                   # :176
    v_context, k_context, _ = resolve_context(context)
    ctx = k_context.clone()
    bg_cache = {}
    for i, s in enumerate(ada_subjects):
        n = s["n_instances"]
        cached = bg_cache.get(s.get("cached_bg_from")) if s.get("cached_bg_from") is not None else None
        vecs, pooled = ada_embedding(s["params"], ca_layer_idx, x1[:n], q[:n], p[pre + "attn2.to_k.weight"], time_emb[:n],
                                     s["probes"][ca_layer_idx], s["extra_means"][ca_layer_idx][:n], s["K"], s["r"],
                                     infeat_size=(h, w), cached_pooler_bg_out=cached)
        bg_cache[i] = pooled["bg_out"]
        ctx[:n, list(s["positions"])] = vecs
        if io is not None:
            io.setdefault("ada_vecs", {})[(ca_layer_idx, i)] = vecs
    x2 = x1 + cross_attention(n2, p, pre + "attn2.", ctx, None, heads)                          # :280-281
    n3 = layer_norm(x2, p[pre + "norm3.weight"], p[pre + "norm3.bias"])
    x3 = geglu_feedforward(n3, p, pre + "ff.") + x2                                             # :283
    y = x3 @ p[prefix + "proj_out.weight"].reshape(c, c).t() + p[prefix + "proj_out.bias"]      # :335-340
    return y.transpose(1, 2).reshape(b, c, h, w) + x                                            # :341


# --------------------------------------------------------------------------------------------
# Conv-attention rows (ldm/util.py:701-879), optional mode (use_conv_attn_kernel_size > 1)
# --------------------------------------------------------------------------------------------
def replace_rows_by_conv_attn(sim: Tensor, q: Tensor, k: Tensor, subj_indices, infeat_size, ks: int,
                              heads: int, scale: float) -> Tensor:
    """Restatement of ldm/util.py:701-879 for conv_attn_mix_weight=1, shift_attn_maps_for_diff_embs=True.

    sim: [B*h, N, T]; q: [B*h, N, d]; k: [B*h, T, d]; subj_indices = (batch_idx[], token_idx[]) with
    M = ks*ks consecutive tokens per instance.
    """
    if ks <= 1:
        return sim
    H, W = infeat_size
    bh, n, d = q.shape
    b = bh // heads
    idx_b, idx_n = [torch.as_tensor(t) for t in subj_indices]
    m_all = ks * ks
    inst = idx_b.reshape(-1, idx_b.numel() // len(torch.unique(idx_b)))[:, 0]
    tok = idx_n.reshape(len(inst), -1)[:, :m_all]
    pad = {2: (0, 1, 0, 1), 3: (1, 1, 1, 1), 4: (1, 2, 1, 2)}[ks]
    out = sim.clone()
    qv = q.reshape(b, heads, n, d)
    kv = k.reshape(b, heads, k.shape[1], d)
    for ii, bi in enumerate(inst.tolist()):
        # q as an image [1, h*d, H, W]; subject keys as a grouped conv weight [h, d, ks, ks]
        q_img = qv[bi].permute(0, 2, 1).reshape(1, heads * d, H, W)
        kw = kv[bi][:, tok[ii]]                                    # [h, M, d]
        kw = kw.permute(0, 2, 1).reshape(heads, d, ks, ks)
        q_pad = F.pad(q_img, pad)
        conv = F.conv2d(q_pad, kw, groups=heads) * (scale / (ks ** 1.5))   # [1, h, H, W]
        for mi in range(m_all):
            dy, dx = mi // ks, mi % ks
            oy, ox = dy - pad[2], dx - pad[0]
            shifted = torch.zeros_like(conv)
            ys0, ys1 = max(0, oy), min(H, H + oy)
            xs0, xs1 = max(0, ox), min(W, W + ox)
            shifted[:, :, ys0:ys1, xs0:xs1] = conv[:, :, ys0 - oy:ys1 - oy, xs0 - ox:xs1 - ox]
            out.reshape(b, heads, n, -1)[bi, :, :, tok[ii, mi]] = shifted.reshape(heads, n)
    return out


# --------------------------------------------------------------------------------------------
# AttentionalPooler / AdaEmbedding (embedding_manager.py:100-321, 539-910)
# --------------------------------------------------------------------------------------------
def attentional_pooler(x: Tensor, q: Tensor, to_k_weight: Tensor, fg_q_emb: Tensor, bg_q_emb: Tensor,
                       lora_to_k: Tensor, lora_to_fg_q: Tensor, lora_to_bg_q: Tensor,
                       img_mask: Optional[Tensor] = None, infeat_size=None, n_heads: int = 8,
                       attn_keep_scale: Optional[Tensor] = None):
    """AttentionalPooler.forward: embedding_manager.py:146-321.  Eval mode by default; attn_keep_scale [B, h, 2, N]
    (0 or 1 / (1 - p)) stands for `self.attn_drop(attn)` in train mode (:296) with the mask given explicitly (the
    tests pass the mask the kernels regenerate from their seed, `pool_keep_scale`).

    x, q: [B,N,C]; to_k_weight: the layer's attn2.to_k weight [C,768]; fg_q_emb [768]; bg_q_emb [B,768];
    lora_*: grouped 1x1 Conv1d weights [C/8, C/8, 1] (8 groups, C/8 -> C/64 each; :124-126).
    """
    b, n, c = x.shape
    hs = n_heads ** -0.5
    k_ln = layer_norm(q)                                                   # :160 (no affine)
    v = (x + k_ln) * hs                                                    # :178
    k = v                                                                  # :181
    fg_q = (fg_q_emb[None] @ to_k_weight.t()).expand(b, -1)[:, None]       # :193-196  [B,1,C]
    bg_q = (bg_q_emb[:, None] @ to_k_weight.t())                           # :201-203  [B,1,C]
    fg_q_ln, bg_q_ln = layer_norm(fg_q), layer_norm(bg_q)                  # :208-209
    dg = c // n_heads                                                      # in-channels per group
    lg = lora_to_k.shape[0] // n_heads                                     # out-channels per group

    def grouped(t, w):                                                     # t [B,L,C] -> [B,L,C/8]
        tg = t.reshape(t.shape[0], t.shape[1], n_heads, dg)
        wg = w.reshape(n_heads, lg, dg)
        return torch.einsum("blgd,god->blgo", tg, wg).reshape(t.shape[0], t.shape[1], n_heads * lg)

    lora_fg_q = grouped(fg_q_ln, lora_to_fg_q)                             # :219
    lora_bg_q = grouped(bg_q_ln, lora_to_bg_q)                             # :222
    lora_k = grouped(k, lora_to_k) * hs                                    # :227
    lora_q = torch.cat([lora_fg_q, lora_bg_q], dim=1) * hs * hs            # :235   [B,2,C/8]
    lq = lora_q.reshape(b, 2, n_heads, lg).permute(0, 2, 1, 3)             # :238   [B,h,2,lg]
    lk = lora_k.reshape(b, n, n_heads, lg).permute(0, 2, 1, 3)             #        [B,h,N,lg]
    vh = v.reshape(b, n, n_heads, dg).permute(0, 2, 1, 3)                  #        [B,h,N,dg]
    sim = torch.matmul(lq, lk.transpose(-1, -2))                           # :247   [B,h,2,N]
    sim = 0.5 * sim + 0.5 * sim.mean(dim=1, keepdim=True)                  # :251-257
    if img_mask is not None:                                               # :259-269
        m = F.interpolate(img_mask, size=infeat_size, mode="nearest").reshape(b, 1, 1, -1).bool()
        sim = sim.masked_fill(~m, -torch.finfo(sim.dtype).max)
    attn = sim.reshape(b, n_heads, -1).softmax(dim=-1).reshape(sim.shape)  # :273-286 joint over 2N
    if attn_keep_scale is not None:                                        # :296 train-mode dropout
        attn = attn * attn_keep_scale
    attn_fg, attn_bg = attn[:, :, 0:1], attn[:, :, 1:2]
    fg_out = torch.matmul(attn_fg, vh).permute(0, 2, 1, 3).reshape(b, c)   # :307,316
    bg_out = torch.matmul(attn_bg, vh).permute(0, 2, 1, 3).reshape(b, c)   # :312,316
    return {"fg_out": fg_out, "bg_out": bg_out,
            "attn_fg": attn_fg.reshape(b * n_heads, 1, n), "attn_bg": attn_bg.reshape(b * n_heads, 1, n)}


def pool_keep_scale(seed: int, b: int, n_tokens: int, drop_p: float, n_heads: int = 8) -> Tensor:
    """The keep-scale [B, h, 2, N] of the pooling kernels' train-mode attention dropout (csrc/common.cuh
    pool_drop_scale): element (b, g, i, t) hashes (seed, ((b*8 + g)*2 + i)*N + t) with the murmur3 finaliser in 32-bit
    wrap-around arithmetic; kept (scale 1 / (1 - p)) when the top 24 bits / 2^24 >= p."""
    m32 = 0xFFFFFFFF
    idx = torch.arange(b * n_heads * 2 * n_tokens, dtype=torch.int64)
    h = ((idx & m32) * 0x9E3779B1 + (seed & m32)) & m32
    h = h ^ (h >> 16)
    h = (h * 0x85EBCA6B) & m32
    h = h ^ (h >> 13)
    h = (h * 0xC2B2AE35) & m32
    h = h ^ (h >> 16)
    u = (h >> 8).to(torch.float32) * (1.0 / 16777216.0)
    keep = (u >= torch.tensor(drop_p, dtype=torch.float32)).to(torch.float32) / (1.0 - drop_p)
    return keep.reshape(b, n_heads, 2, n_tokens)


def ada_embedding(p: Params, ca_layer_idx: int, x: Tensor, q: Tensor, to_k_weight: Tensor,
                  time_emb: Tensor, layer_subj_emb_probe: Tensor, layer_static_extra_emb_mean: Tensor,
                  K: int, r: int, img_mask: Optional[Tensor] = None, infeat_size=None,
                  out_emb_dim: int = 768, cached_pooler_bg_out: Optional[Tensor] = None):
    """AdaEmbedding.forward with fp32 arithmetic (autocast neutralised): embedding_manager.py:779-910.

    `p` holds the AdaEmbedding state_dict (keys poolers.{i}.lora_to_*.weight, layer_coeff_maps.{i}.*,
    pre_vecs, basis_vecs, bias).  H=2 configuration (fg and bg embeddings), see SURVEY.md F2.
    """
    i = ca_layer_idx
    pooled = attentional_pooler(x, q, to_k_weight, layer_subj_emb_probe, layer_static_extra_emb_mean,
                                p[f"poolers.{i}.lora_to_k.weight"], p[f"poolers.{i}.lora_to_fg_q.weight"],
                                p[f"poolers.{i}.lora_to_bg_q.weight"], img_mask, infeat_size)   # :815-819
    infeat_fg = pooled["fg_out"]
    infeat_bg = cached_pooler_bg_out if cached_pooler_bg_out is not None else pooled["bg_out"]  # :806-825
    infeat = torch.cat([infeat_fg, infeat_bg], dim=-1)                    # :831
    c = x.shape[-1]
    td = int(0.25 * c)                                                    # :612-619,684 (TIME_H = 1)
    time_feat = time_emb[:, :td]                                          # :849
    feat = torch.cat([layer_norm(infeat), layer_norm(time_feat)], dim=1)  # LNCat3 :51-58, :855
    coeffs = feat @ p[f"layer_coeff_maps.{i}.weight"].t() + p[f"layer_coeff_maps.{i}.bias"]
    coeffs = coeffs.reshape(-1, K, r)                                     # :858
    if "pre_vecs" in p and p["pre_vecs"] is not None:
        basis = torch.cat([p["pre_vecs"], p["basis_vecs"]], dim=1)        # :874  [K,r,768]
    else:
        basis = p["basis_vecs"]
    out = torch.einsum("bkr,krd->bkd", coeffs, basis)                     # :881
    out = layer_norm(out) / math.sqrt(out_emb_dim)                        # :882-884 (per-(layer,k) LN, no affine)
    if "bias" in p and p["bias"] is not None:
        out = out + p["bias"][i][None]                                    # :886-893
    return out, pooled


# --------------------------------------------------------------------------------------------
# StaticLayerwiseEmbedding + the splice of its rows into the layerwise prompt
# --------------------------------------------------------------------------------------------
def static_layerwise_embedding(p: Params, out_emb_dim: int = 768) -> Tensor:
    """StaticLayerwiseEmbedding.forward, not zero-shot: embedding_manager.py:497-537.

    p: basis_rand_weights [L,K,r], basis_comm_weights [1,K,r], basis_vecs [K,r-N,D], optional pre_vecs [K,N,D],
    bias [L,K,D] (absent = 0).  Returns [L,K,D]."""
    weights = p["basis_rand_weights"] + p["basis_comm_weights"]                          # :501
    basis = p["basis_vecs"]
    if "pre_vecs" in p and p["pre_vecs"] is not None:
        basis = torch.cat([p["pre_vecs"], basis], dim=1)                                 # :516-518
    L, K, _ = weights.shape
    unnorm = torch.einsum("lkr,krd->lkd", weights, basis)                                # :522
    ln = F.layer_norm(unnorm, (out_emb_dim,), None, None, 1e-5)                          # :529 (no affine)
    out = ln / math.sqrt(out_emb_dim)                                                    # :530
    bias = p.get("bias", None)
    return out if bias is None else out + bias                                           # :534


def splice_static_embedding(embedded_text: Tensor, tokenized_text: Tensor, placeholder_token: int,
                            subj_static_embedding: Tensor, num_layers: int = 16) -> Tensor:
    """EmbeddingManager.get_static_embedding, inference branch for one placeholder: embedding_manager.py:1341-1352
    (repeat over the layers, instance-major / layer-minor), :1360-1366 (first occurrence per instance),
    :1510-1552 (vector k of layer l goes to token position first + k of row b*L + l).

    embedded_text [B,N,D], tokenized_text [B,N] (int), subj_static_embedding [L,K,D].  Returns [B*L,N,D]."""
    B, N, D = embedded_text.shape
    L, K = num_layers, subj_static_embedding.shape[1]
    out = embedded_text.unsqueeze(1).repeat(1, L, 1, 1).reshape(B * L, N, D).clone()
    for b in range(B):
        hits = (tokenized_text[b] == placeholder_token).nonzero().flatten()
        if hits.numel() == 0:
            continue
        first = int(hits[0])
        for k in range(K):
            out[b * L:(b + 1) * L, first + k] = subj_static_embedding[:, k].to(out.dtype)
    return out


# --------------------------------------------------------------------------------------------
# DDIM schedule + CFG combine + update (ddim.py:28-68,135-296; diffusionmodules/util.py:21-77)
# --------------------------------------------------------------------------------------------
def make_beta_schedule_linear(n_timestep: int = 1000, linear_start: float = 0.00085,
                              linear_end: float = 0.012) -> np.ndarray:
    """diffusionmodules/util.py:21-25 ('linear'), values from v1-inference-ada.yaml:5-6."""
    return np.linspace(linear_start ** 0.5, linear_end ** 0.5, n_timestep, dtype=np.float64) ** 2


def make_ddim_schedule(num_steps: int = 50, num_ddpm: int = 1000, eta: float = 0.0,
                       linear_start: float = 0.00085, linear_end: float = 0.012):
    """ddim.py:28-68 + util.py:46-77 (uniform discretisation). Returns float32 arrays like the reference
    (alphas_cumprod is converted to float32 first: ddpm.py:244-265 registers it as float32)."""
    betas = make_beta_schedule_linear(num_ddpm, linear_start, linear_end)
    alphas_cumprod = np.cumprod(1.0 - betas, axis=0).astype(np.float32)
    c = num_ddpm // num_steps
    timesteps = np.asarray(list(range(0, num_ddpm, c))) + 1               # util.py:46-62
    alphas = alphas_cumprod[timesteps]
    alphas_prev = np.asarray([alphas_cumprod[0]] + alphas_cumprod[timesteps[:-1]].tolist(),
                             dtype=np.float32)                            # util.py:68
    # the reference mixes a float64 ndarray (alphas_prev) with a float32 tensor (alphas) here, i.e. the
    # expression is evaluated in float64 on float32-valued inputs and rounded when the scalar is used
    a64, ap64 = alphas.astype(np.float64), alphas_prev.astype(np.float64)
    sigmas = (eta * np.sqrt((1 - ap64) / (1 - a64) * (1 - a64 / ap64))).astype(np.float32)
    return {"timesteps": timesteps, "alphas": alphas, "alphas_prev": alphas_prev, "sigmas": sigmas,
            "sqrt_one_minus_alphas": np.sqrt(1.0 - alphas).astype(np.float32)}


def guidance_schedule(guidance_scale, total_steps: int):
    """Linear guidance annealing: ddim.py:169-180,215-218. Returns the scale used at loop step i."""
    if isinstance(guidance_scale, (list, tuple)):
        g_max, g_min = guidance_scale
    else:                                           # scalar: reference would crash; treat as (g, g)
        g_max = g_min = guidance_scale
    delta = (g_max - g_min) / (total_steps - 1)
    out, g = [], g_max
    for _ in range(total_steps):
        out.append(g)
        g = g - delta
    return out


def cfg_ddim_update(x: Tensor, e_cond: Tensor, e_uncond: Tensor, guidance_scale: float, a_t: float,
                    a_prev: float, sigma_t: float, sqrt_one_minus_at: float,
                    noise: Optional[Tensor] = None, temperature: float = 1.0):
    """ddim.py:252-296 with float32 scalars, same operation order."""
    f32 = lambda v: torch.tensor(v, dtype=x.dtype)
    e_t = e_uncond + guidance_scale * (e_cond - e_uncond)                 # :260
    pred_x0 = (x - f32(sqrt_one_minus_at) * e_t) / f32(a_t).sqrt()        # :279
    dir_xt = (1.0 - f32(a_prev) - f32(sigma_t) ** 2).sqrt() * e_t         # :283
    nz = 0.0 if noise is None else f32(sigma_t) * noise * temperature     # :286-287
    x_prev = f32(a_prev).sqrt() * pred_x0 + dir_xt + nz                   # :295
    return x_prev, pred_x0


def ddim_sample(apply_model: Callable[[Tensor, Tensor, Tensor], Tensor], x_T: Tensor, c_cond: Tensor,
                c_uncond: Tensor, num_steps: int = 50, guidance_scale=(10.0, 4.0), eta: float = 0.0,
                generator: Optional[torch.Generator] = None, return_trace: bool = False):
    """ddim_sampling + p_sample_ddim: ddim.py:135-296.  CFG batch order (cond, uncond): :238-252.

    `apply_model(x_in [2b,...], t_in [2b], ctx [2b*16,77,768]) -> eps [2b,...]`.
    """
    sch = make_ddim_schedule(num_steps, eta=eta)
    scales = guidance_schedule(guidance_scale, num_steps)
    img = x_T
    b = x_T.shape[0]
    trace = []
    for i, step in enumerate(np.flip(sch["timesteps"])):
        index = num_steps - i - 1                                          # :186
        ts = torch.full((b,), int(step), dtype=torch.long)
        x_in = torch.cat([img, img])                                       # :231
        t_in = torch.cat([ts, ts])
        ctx = torch.cat([c_cond, c_uncond])                                # :243/:249
        e_c, e_u = apply_model(x_in, t_in, ctx).chunk(2)                   # :252
        noise = None
        if eta > 0:
            noise = torch.randn(img.shape, generator=generator)
        img, pred_x0 = cfg_ddim_update(img, e_c, e_u, scales[i], float(sch["alphas"][index]),
                                       float(sch["alphas_prev"][index]), float(sch["sigmas"][index]),
                                       float(sch["sqrt_one_minus_alphas"][index]), noise)
        if return_trace:
            trace.append(img.clone())
    return (img, trace) if return_trace else img


# --------------------------------------------------------------------------------------------
# U-Net shell around the path (openaimodel.py:167-281 ResBlock, :95-165 up/down, :827-1052 forward).
# Outside the hot path (ResBlock convolutions stay on torch); restated here only so that the full
# DDIM step has a CPU checker on the GPU box.
# --------------------------------------------------------------------------------------------
def timestep_embedding(timesteps: Tensor, dim: int, max_period: int = 10000) -> Tensor:
    """diffusionmodules/util.py:154-174."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half)
    args = timesteps[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def _gn32(x, p, name):
    return group_norm(x.float(), p[name + ".weight"], p[name + ".bias"], 32, 1e-5)   # util.py:202-219


def res_block(x: Tensor, emb: Tensor, p: Params, prefix: str) -> Tensor:
    """ResBlock._forward without up/down, use_scale_shift_norm=False: openaimodel.py:259-279."""
    h = F.silu(_gn32(x, p, prefix + "in_layers.0"))
    h = F.conv2d(h, p[prefix + "in_layers.2.weight"], p[prefix + "in_layers.2.bias"], padding=1)
    e = F.silu(emb) @ p[prefix + "emb_layers.1.weight"].t() + p[prefix + "emb_layers.1.bias"]
    h = h + e[:, :, None, None]
    h = F.silu(_gn32(h, p, prefix + "out_layers.0"))
    h = F.conv2d(h, p[prefix + "out_layers.3.weight"], p[prefix + "out_layers.3.bias"], padding=1)
    if (prefix + "skip_connection.weight") in p:
        x = F.conv2d(x, p[prefix + "skip_connection.weight"], p[prefix + "skip_connection.bias"])
    return x + h


def upsample(x: Tensor, p: Params, prefix: str) -> Tensor:
    """Upsample.forward with use_conv=True: openaimodel.py:113-123."""
    x = F.interpolate(x, scale_factor=2, mode="nearest")
    return F.conv2d(x, p[prefix + "conv.weight"], p[prefix + "conv.bias"], padding=1)


def unet_forward(p: Params, x: Tensor, timesteps: Tensor, context: Tensor, model_channels: int = 320,
                 channel_mult: Sequence[int] = (1, 2, 4, 4), num_res_blocks: int = 2,
                 attention_resolutions: Sequence[int] = (4, 2, 1), heads: int = 8,
                 st_fn: Optional[Callable] = None, block_outputs: Optional[dict] = None,
                 ada_subjects=None, block_inputs: Optional[dict] = None) -> Tensor:
    """UNetModel.forward with use_layerwise_context=True, no conv-attn: openaimodel.py:827-1052.

    context: [16*B,77,768], instance-major / layer-minor (embedding_manager.py:1345-1349).
    `st_fn(x, params, prefix, layer_context)` defaults to the port's spatial_transformer.
    `ada_subjects` (see spatial_transformer_ada) composes the per-layer AdaEmbedding call site into every conditioned
    layer - the build's reconstruction of SURVEY.md 3.2 (F1), the workload bench.py times.  `block_inputs` /
    `block_outputs` collect every SpatialTransformer's input / output by U-Net block index; block_outputs["emb"] is the
    time embedding.
    """
    b = x.shape[0]
    ctx = context.reshape(b, 16, -1, context.shape[-1]).permute(1, 0, 2, 3)        # :866
    emb = timestep_embedding(timesteps, model_channels)                             # :846
    emb = emb @ p["time_embed.0.weight"].t() + p["time_embed.0.bias"]
    emb = F.silu(emb) @ p["time_embed.2.weight"].t() + p["time_embed.2.bias"]       # :847

    def run_st(h, prefix, layer_idx):
        lc = ctx[LAYER_IDX2CA[layer_idx]]
        layer_context = lambda: ((lc, lc), None)                                    # :894-896,920
        if block_inputs is not None:
            block_inputs[layer_idx] = h
        if st_fn is not None:
            out = st_fn(h, p, prefix, layer_context)
        elif ada_subjects:
            out = spatial_transformer_ada(h, p, layer_context, ada_subjects, LAYER_IDX2CA[layer_idx], emb, heads, prefix)
        else:
            out = spatial_transformer(h, p, layer_context, None, heads, prefix)
        if block_outputs is not None:
            block_outputs[layer_idx] = out
        return out

    if block_outputs is not None:
        block_outputs["emb"] = emb

    hs = []
    h = F.conv2d(x, p["input_blocks.0.0.weight"], p["input_blocks.0.0.bias"], padding=1)
    hs.append(h)
    layer_idx, blk, ds = 1, 1, 1
    for level, mult in enumerate(channel_mult):
        for _ in range(num_res_blocks):
            h = res_block(h, emb, p, f"input_blocks.{blk}.0.")
            if ds in attention_resolutions:
                h = run_st(h, f"input_blocks.{blk}.1.", layer_idx)
            hs.append(h)
            blk += 1
            layer_idx += 1
        if level != len(channel_mult) - 1:
            h = F.conv2d(h, p[f"input_blocks.{blk}.0.op.weight"], p[f"input_blocks.{blk}.0.op.bias"],
                         stride=2, padding=1)                                       # Downsample :138-164
            hs.append(h)
            blk += 1
            layer_idx += 1
            ds *= 2
    h = res_block(h, emb, p, "middle_block.0.")
    h = run_st(h, "middle_block.1.", layer_idx)
    h = res_block(h, emb, p, "middle_block.2.")
    layer_idx += 1
    blk = 0
    for level, mult in list(enumerate(channel_mult))[::-1]:
        for i in range(num_res_blocks + 1):
            h = torch.cat([h, hs.pop()], dim=1)                                     # :1018-1019
            h = res_block(h, emb, p, f"output_blocks.{blk}.0.")
            j = 1
            if ds in attention_resolutions:
                h = run_st(h, f"output_blocks.{blk}.1.", layer_idx)
                j = 2
            if level and i == num_res_blocks:
                h = F.interpolate(h, scale_factor=2, mode="nearest")                # Upsample :113-123
                h = F.conv2d(h, p[f"output_blocks.{blk}.{j}.conv.weight"],
                             p[f"output_blocks.{blk}.{j}.conv.bias"], padding=1)
                ds //= 2
            blk += 1
            layer_idx += 1
    h = F.silu(_gn32(h, p, "out.0"))
    return F.conv2d(h, p["out.2.weight"], p["out.2.bias"], padding=1)               # :693-697,1052


# --------------------------------------------------------------------------------------------
# Algorithmic FLOPs (SURVEY.md section 8d / BASELINE.md section 3)
# --------------------------------------------------------------------------------------------
def st_flops(n_tokens: int, c: int, t_ctx: int = 77, ctx_dim: int = 768) -> float:
    return 40.0 * n_tokens * c * c + 4.0 * n_tokens * n_tokens * c + 4.0 * n_tokens * t_ctx * c \
        + 4.0 * t_ctx * ctx_dim * c


def unet_st_flops(latent_hw: int = 64) -> float:
    n = latent_hw * latent_hw
    ladder = [(n, 320, 5), (n // 4, 640, 5), (n // 16, 1280, 5), (n // 64, 1280, 1)]
    return sum(cnt * st_flops(nn_, c) for nn_, c, cnt in ladder)
