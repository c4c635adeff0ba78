"""Host-side mirror of the per-layer adaptive subject embedding (reference:
ldm/modules/embedding_manager.py - LNCat3 :38-58, AttentionalPooler :100-321, AdaEmbedding :539-910).

Constructor arguments, `forward` signatures, return values and state_dict keys follow the reference.  The
arithmetic runs in two kernel groups (adaface_b200/csrc/ada.cu): K12, one streaming pass over the layer's
`x` and `q` with an online joint softmax over the (fg | bg) x token axis, and K13, the coefficient map +
basis combination + per-vector LayerNorm head.  Eval-mode semantics (the pooler's Dropout(0.1) is
train-only); fp32 statistics on bf16 activations, i.e. the fp32 restatement of the reference rather than
its forced CPU-bf16 autocast.

In the reference snapshot AdaEmbedding is defined but not wired into the U-Net; the wiring
offered here (`AdaPromptContext` in adaface_b200/unet.py) reconstructs the removed call site from the
class's own contract.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn.functional as F
from torch import nn

from . import ops
from .attention import _PackedWeights, _f32, as_bf16_tokens

Tensor = torch.Tensor

CA_INFEAT_DIMS = [320, 320, 640, 640, 1280, 1280, 1280, 1280, 1280, 1280, 640, 640, 640, 320, 320, 320]
LAYER_IDX2CA_LAYER_IDX = {1: 0, 2: 1, 4: 2, 5: 3, 7: 4, 8: 5, 12: 6, 16: 7,
                          17: 8, 18: 9, 19: 10, 20: 11, 21: 12, 22: 13, 23: 14, 24: 15}


class LNCat3(nn.Module):
    """embedding_manager.py:38-58.  Parameter-free (both LayerNorms are non-affine); its arithmetic is part
    of the K13 kernel, the module exists for structural parity."""

    def __init__(self, chan1: int, chan2: int, chan3: int = 0):
        super().__init__()
        self.chan1, self.chan2, self.chan3 = chan1, chan2, chan3


class AttentionalPooler(nn.Module, _PackedWeights):
    """embedding_manager.py:100-321."""

    def __init__(self, layer_idx: int, feat_dim: int, feat_reduction_ratio: int = 8, infeat_grad_scale: float = 0.5):
        super().__init__()
        self.n_heads = 8
        self.layer_inner_dim = feat_dim
        self.lora_dim = int(feat_dim / feat_reduction_ratio)
        if self.lora_dim * 8 != feat_dim:
            raise NotImplementedError("adaface_b200.AttentionalPooler: feat_reduction_ratio must be 8 (kernel layout)")
        conv = lambda: nn.Conv1d(feat_dim, self.lora_dim, kernel_size=1, groups=self.n_heads, bias=False)
        self.lora_to_k, self.lora_to_fg_q, self.lora_to_bg_q = conv(), conv(), conv()
        self.conv1d_extra_scale = self.n_heads ** -0.5
        self.layer_idx = layer_idx
        self.infeat_grad_scale = infeat_grad_scale
        self.is_fgbg_competitive = True
        self.attn_drop = nn.Dropout(0.1)
        self.out_drop = nn.Dropout(0.1)

    def _build_pack(self):
        return {k: _f32(getattr(self, k).weight.squeeze(-1)) for k in ("lora_to_k", "lora_to_fg_q", "lora_to_bg_q")}

    def _differentiated(self) -> bool:
        """True inside an inversion step: gradients are being recorded and this pooler's grouped-conv weights train."""
        return torch.is_grad_enabled() and any(getattr(self, k).weight.requires_grad
                                               for k in ("lora_to_k", "lora_to_fg_q", "lora_to_bg_q"))

    def prepare_queries(self, to_k: nn.Module, fg_q_emb: Tensor, bg_q_emb: Tensor, batch: int) -> Tensor:
        """fg / bg pooling queries [B, 2, C] folded through lora_to_k (embedding_manager.py:186-235, :227)."""
        if self._differentiated():
            # the parameters themselves (views, not the detached kernel-layout copies): autograd routes the gradients of
            # ops.ada_pool_prep (ops.ada_pool_prep_bwd) to them
            p = {k: getattr(self, k).weight.squeeze(-1) for k in ("lora_to_k", "lora_to_fg_q", "lora_to_bg_q")}
        else:
            p = self.packed()
        if bg_q_emb.dim() == 1:
            bg_q_emb = bg_q_emb[None].expand(batch, -1)
        return ops.ada_pool_prep(to_k.weight.detach(), fg_q_emb.reshape(-1), bg_q_emb.reshape(batch, -1),
                                 p["lora_to_fg_q"], p["lora_to_bg_q"], p["lora_to_k"])

    def forward(self, layer_attn_components: Dict, fg_q_emb: Tensor, bg_q_emb: Tensor, img_mask: Optional[Tensor] = None,
                debug: bool = False, want_attn: bool = True):
        if self.training:
            raise NotImplementedError("adaface_b200.AttentionalPooler: the kernel path implements eval mode "
                                      "(Dropout(0.1) on the attention is train-only); call .eval()")
        x, q = layer_attn_components["x"], layer_attn_components["q"]
        to_k = layer_attn_components["to_k"]
        infeat_size = layer_attn_components.get("infeat_size")
        b, n, c = x.shape
        # the pooling queries depend on the prompt and the weights only, not on the layer features: a caller may
        # have computed them ahead (prepare_queries, e.g. on a side stream) and passes them along
        pool_q = layer_attn_components.get("pool_q")
        if pool_q is None:
            pool_q = self.prepare_queries(to_k, fg_q_emb, bg_q_emb, b)
        key_mask = None
        if img_mask is not None:                                                                # :259-269
            m = F.interpolate(img_mask.float(), size=infeat_size, mode="nearest").reshape(b, -1)
            key_mask = m.bool().to(torch.uint8).contiguous()
        fg, bg, a_fg, a_bg = ops.ada_pool(as_bf16_tokens(x), as_bf16_tokens(q), pool_q, key_mask, want_attn)
        out = {"fg_out": fg, "bg_out": bg}
        if want_attn:
            out["attn_fg"], out["attn_bg"] = a_fg, a_bg
        return out


class AdaEmbedding(nn.Module, _PackedWeights):
    """embedding_manager.py:539-910.  Built for the attention-pooler variant in its two-stream (H = 2)
    configuration - the only one in which the reference's own forward is self-consistent -
    plus the bg-token variant that reuses a cached bg pooling (`use_cached_bg`)."""

    def __init__(self, num_layers: int = 16, num_vectors_per_subj_token: int = 1, fg_emb_count: int = 1,
                 bg_emb_count: int = 0, use_cached_bg: bool = False, out_emb_dim: int = 768, r: int = 12,
                 init_string=None, init_vecs: Optional[Tensor] = None, ca_infeat_dims=CA_INFEAT_DIMS,
                 layer_idx2ca_layer_idx=LAYER_IDX2CA_LAYER_IDX, has_bias: bool = True, use_attn_pooler: bool = True,
                 attn_pooler_feat_reduction_ratio: int = 8, token_string: str = "", placeholder_is_bg: bool = False,
                 do_zero_shot: bool = False, device_type: str = "cuda"):
        super().__init__()
        if not use_attn_pooler:
            raise NotImplementedError("adaface_b200.AdaEmbedding: only use_attn_pooler=True is built")
        if do_zero_shot and not has_bias:
            raise ValueError("Zero-shot AdaEmbedding must set has_bias=True")
        if num_layers != len(layer_idx2ca_layer_idx):
            raise ValueError(f"num_layers={num_layers} != len(layer_idx2ca_layer_idx)={len(layer_idx2ca_layer_idx)}")
        self.do_zero_shot = do_zero_shot
        self.token_string = token_string
        self.num_layers = num_layers
        self.out_emb_dim = out_emb_dim
        self.K = K = num_vectors_per_subj_token
        self.fg_emb_count, self.bg_emb_count = fg_emb_count, bg_emb_count
        if fg_emb_count + bg_emb_count > K:
            raise ValueError(f"fg_emb_count={fg_emb_count} + bg_emb_count={bg_emb_count} > num_vectors_per_subj_token={K}")
        self.placeholder_is_bg = placeholder_is_bg
        self.use_cached_bg = use_cached_bg
        if use_cached_bg:
            self.cached_bg_weight = nn.Parameter(torch.tensor(0.5), requires_grad=True)
        self.is_fg_only = fg_emb_count == K
        self.is_bg_only = bg_emb_count == K
        self.is_one_stream_only = self.is_fg_only or self.is_bg_only
        self.r = r
        self.use_attn_pooler = use_attn_pooler
        self.attn_pooler_feat_reduction_ratio = attn_pooler_feat_reduction_ratio
        self.emb_infeat_types = [0] * fg_emb_count + [1] * bg_emb_count + [2] * (K - fg_emb_count - bg_emb_count)
        self.device_type = device_type
        self.layer_idx2ca_layer_idx = dict(layer_idx2ca_layer_idx)
        self.ca_layer_idx2layer_idx = {v: k for k, v in self.layer_idx2ca_layer_idx.items()}
        self.H = H = 1 if self.is_one_stream_only else 2
        time_h = max([2 if t == 2 else 1 for t in self.emb_infeat_types] or [1])
        self.TD_frac = 0.25 * time_h

        if init_vecs is not None:
            if init_vecs.shape[1] != out_emb_dim or init_vecs.shape[0] > num_layers:
                raise ValueError(f"AdaEmbedding LoRA init vectors shape {tuple(init_vecs.shape)} must be "
                                 f"(<={num_layers}, {out_emb_dim})")
            self.N = n0 = init_vecs.shape[0]
            self.pre_vecs = None if do_zero_shot else nn.Parameter(init_vecs.unsqueeze(0).repeat(K, 1, 1))
        else:
            self.N = n0 = 0
            self.pre_vecs = None
        if not do_zero_shot:
            basis = F.normalize(torch.randn(K, r - n0, out_emb_dim), dim=-1) / 4.
            basis[:, -1] = 0
            self.basis_vecs = nn.Parameter(basis)
        else:
            self.basis_vecs = None

        self.ca_infeat_dims = list(ca_infeat_dims)
        self.poolers = nn.ModuleList([AttentionalPooler(i, self.ca_infeat_dims[i], attn_pooler_feat_reduction_ratio, 1)
                                      for i in range(num_layers)])
        self.TDs = [int(self.TD_frac * d) for d in self.ca_infeat_dims]
        self.layer_coeff_maps = nn.ModuleList([nn.Linear(d * H + td, r * K, bias=True)
                                               for d, td in zip(self.ca_infeat_dims, self.TDs)])
        self.layers_out_lns = nn.ModuleList([nn.ModuleList([nn.LayerNorm(out_emb_dim, elementwise_affine=False)
                                                            for _ in range(K)]) for _ in range(num_layers)])
        self.layer_lncat3s = nn.ModuleList([LNCat3(d * H, td) for d, td in zip(self.ca_infeat_dims, self.TDs)])
        self.reduce_fg_bg_cross_weights()
        self.has_bias = has_bias
        self.bias = nn.Parameter(torch.zeros(num_layers, K, out_emb_dim)) if (has_bias and not do_zero_shot) else None
        self.call_count = 0
        self.debug = False

    @torch.no_grad()
    def reduce_fg_bg_cross_weights(self, reduced_layer_idx: Optional[int] = None):
        """embedding_manager.py:719-773 (init-time): caps the cross-stream weights of the coefficient maps -
        bg->fg-embedding and fg->bg-embedding blocks - at 1 % of the matching in-stream mean magnitude."""
        if self.H == 1:
            return
        max_ratio = 0.01
        layers = range(self.num_layers) if reduced_layer_idx is None else [reduced_layer_idx]
        for li in layers:
            d = self.ca_infeat_dims[li]
            w = self.layer_coeff_maps[li].weight.data.view(self.K, self.r, -1)
            for k, kind in enumerate(self.emb_infeat_types):
                if kind == 2:
                    continue
                own, cross = (slice(0, d), slice(d, 2 * d)) if kind == 0 else (slice(d, 2 * d), slice(0, d))
                own_mean = w[k][:, own].abs().mean().item()
                cross_mean = w[k][:, cross].abs().mean().item()
                w[k][:, cross] *= min(1, max_ratio * own_mean / (cross_mean + 1e-6))

    def _pack_key(self):
        ps = [p for p in (self.pre_vecs, self.basis_vecs, self.bias) if p is not None]
        ps += [p for m in self.layer_coeff_maps for p in m.parameters()]
        return tuple((p.data_ptr(), p._version, p.dtype) for p in ps)

    def _build_pack(self):
        out = {}
        if self.basis_vecs is not None:
            basis = self.basis_vecs if self.pre_vecs is None else torch.cat([self.pre_vecs, self.basis_vecs], dim=1)
            out["basis"] = _f32(basis)                                                          # :874
        for i, m in enumerate(self.layer_coeff_maps):
            out[f"cw{i}"], out[f"cb{i}"] = _f32(m.weight), _f32(m.bias)
        if self.bias is not None:
            out["bias"] = _f32(self.bias)
        return out

    def forward(self, layer_idx: int, layer_attn_components: Dict, time_emb: Tensor, layer_subj_emb_probe: Tensor,
                layer_static_extra_emb_mean: Tensor, img_mask: Optional[Tensor] = None,
                cached_pooler_bg_out: Optional[Tensor] = None, zs_basis_vecs: Optional[Tensor] = None,
                splice_into: Optional[Tensor] = None, splice_positions: Optional[Tensor] = None):
        """Reference signature plus two optional arguments used by the U-Net wiring: `splice_into` (bf16 [B, T, 768]
        layer context) and `splice_positions` (int32 [K]) make the head kernel write the K vectors straight into the
        subject-token rows of the context, instead of a separate cast + index_put."""
        ca = self.layer_idx2ca_layer_idx[layer_idx]
        if self.H != 2:
            # the reference concatenates fg and bg features unconditionally (:831) into a Linear sized for one
            # stream when H == 1, and raises
            raise RuntimeError("AdaEmbedding.forward needs both fg and bg embeddings (fg_emb_count, bg_emb_count > 0)")
        if self.use_cached_bg and cached_pooler_bg_out is None:
            raise ValueError("AdaEmbedding(use_cached_bg=True) needs cached_pooler_bg_out")             # :789-792
        if self.do_zero_shot and zs_basis_vecs is None:
            raise ValueError("Zero-shot AdaEmbedding requires zs_basis_vecs")                       # :861
        pooled = self.poolers[ca](layer_attn_components, layer_subj_emb_probe, layer_static_extra_emb_mean,
                                  img_mask=img_mask, want_attn=getattr(self, "want_pooler_attn", True))   # :815-819
        infeat_fg = pooled.pop("fg_out")                                                        # :823-829
        infeat_bg = cached_pooler_bg_out if self.use_cached_bg else pooled["bg_out"]
        differentiated = torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (
            self.layer_coeff_maps[ca].weight, self.layer_coeff_maps[ca].bias, self.basis_vecs, self.pre_vecs, self.bias,
            infeat_fg, infeat_bg))
        if differentiated and not self.do_zero_shot:
            # inversion step: hand the parameters themselves to the op (ops.ada_head_bwd produces their gradients) and
            # write the K rows into a COPY of the context with a differentiable index_copy instead of the in-kernel splice
            m = self.layer_coeff_maps[ca]
            basis = self.basis_vecs if self.pre_vecs is None else torch.cat([self.pre_vecs, self.basis_vecs], dim=1)
            bias = None if self.bias is None else self.bias[ca]
            out = ops.ada_head(infeat_fg, infeat_bg, time_emb, self.TDs[ca], m.weight, m.bias, basis, bias)
            if splice_into is not None:
                ctx_new = splice_into.clone()
                n = out.shape[0]
                ctx_new[:n] = ctx_new[:n].index_copy(1, splice_positions.long(), out.to(ctx_new.dtype))
                pooled["ctx"] = ctx_new
            return out, pooled
        p = self.packed()
        if self.do_zero_shot:
            # per-instance basis [B, K, r, 768] from the subject basis generator, no bias; the coefficient scaler
            # (gen_gradient_scaler, :656,868) is the identity in the forward pass                   :860-868,886-890
            if zs_basis_vecs.dim() != 4 or tuple(zs_basis_vecs.shape[1:]) != (self.K, self.r, self.out_emb_dim):
                raise ValueError(f"zs_basis_vecs must be [B, {self.K}, {self.r}, {self.out_emb_dim}]")
            basis, bias = zs_basis_vecs[:infeat_fg.shape[0]], None
        else:
            basis, bias = p["basis"], (None if self.bias is None else p["bias"][ca])
        if splice_into is None:
            out = ops.ada_head(infeat_fg, infeat_bg, time_emb, self.TDs[ca], p[f"cw{ca}"], p[f"cb{ca}"], basis,
                               bias)                                                            # :847-893
        else:
            out = ops.ada_head_splice(infeat_fg, infeat_bg, time_emb, self.TDs[ca], p[f"cw{ca}"], p[f"cb{ca}"],
                                      basis, bias, splice_into, splice_positions)
        return out, pooled


class StaticLayerwiseEmbedding(nn.Module):
    """embedding_manager.py:360-537, the learned ("slow", not zero-shot) form: 16 x K static subject
    vectors as a low-rank combination of K sets of r basis vectors.  Constructor arguments, parameter names
    (`pre_vecs`, `basis_rand_weights`, `basis_vecs`, `basis_comm_weights`, `bias`) and their initialisation follow the
    reference, so its state_dict loads; `forward()` is one kernel (`afb_static_layerwise_embedding`).  The zero-shot form
    of the reference only reshapes the vectors a subject-basis generator hands it (:504-511) and is a view here."""

    def __init__(self, num_layers: int = 16, num_vectors_per_subj_token: int = 1, out_emb_dim: int = 768, r: int = 6,
                 init_noise_stds=(0.1, 0.04), init_string=None, init_vecs: Optional[Tensor] = None,
                 init_vec_weights: Optional[Tensor] = None, has_bias: bool = True, token_string: str = "",
                 do_zero_shot: bool = False, device_type: str = "cuda"):
        super().__init__()
        if r > min(num_layers, out_emb_dim):                                                       # :392-395
            raise ValueError(f"StaticLayerwiseEmbedding LoRA rank {r} must be less or equal than "
                             f"{min(num_layers, out_emb_dim)}")
        self.do_zero_shot, self.token_string, self.num_layers = do_zero_shot, token_string, num_layers
        self.K, self.out_emb_dim, self.r = num_vectors_per_subj_token, out_emb_dim, r
        self.has_bias, self.device_type = has_bias, device_type
        K = self.K
        if init_vecs is not None:
            if init_vecs.shape[1] != out_emb_dim or init_vecs.shape[0] > num_layers:               # :398-401
                raise ValueError(f"StaticLayerwiseEmbedding init vectors shape {tuple(init_vecs.shape)} must be "
                                 f"(<={num_layers}, {out_emb_dim})")
            N = self.N = init_vecs.shape[0]
            self.pre_vecs = None if do_zero_shot else nn.Parameter(init_vecs.unsqueeze(0).repeat(K, 1, 1).clone())
        else:
            N = self.N = 0
            self.pre_vecs = None
        self.basis_rand_weights = nn.Parameter(torch.randn(num_layers, K, r))                      # :417
        if not do_zero_shot:
            bv = F.normalize(torch.randn(K, r - N, out_emb_dim), dim=-1) / 4.0                      # :424-426
            bv[-1] = 0                                                                             # :428
            self.basis_vecs = nn.Parameter(bv)
        else:
            self.basis_vecs = None
        self.basis_comm_weights = nn.Parameter(torch.ones(1, K, r) / r)                            # :438
        if init_vecs is not None:                                                                  # :442-477
            with torch.no_grad():
                self.basis_comm_weights.fill_(1.0 / N)
                self.basis_comm_weights[:, :, N:] *= 0.4
                if init_vec_weights is not None:
                    assert len(init_vec_weights) == len(init_vecs)
                    self.basis_comm_weights[:, :, :N] = init_vec_weights.reshape(1, 1, -1)
                self.basis_rand_weights[:, :, :N] *= init_noise_stds[1]
                self.basis_rand_weights[:, :, N:] *= init_noise_stds[0]
        if has_bias and not do_zero_shot:
            self.bias = nn.Parameter(torch.zeros(num_layers, K, out_emb_dim))                      # :481-483
        else:
            self.bias = 0

    def forward(self, static_zs_embs: Optional[Tensor] = None) -> Tensor:
        """-> [16, K, 768] fp32 (zero-shot: `static_zs_embs` [B, 16, K, 768] -> [(B 16), K, 768], :504-511)."""
        if self.do_zero_shot:
            b, l, k, d = static_zs_embs.shape
            out = static_zs_embs.reshape(b * l, k, d)
            self.bias, self.has_bias = out, True
            return out
        bias = self.bias if isinstance(self.bias, torch.Tensor) else None
        return ops.static_layerwise_embedding(self.basis_rand_weights.detach(), self.basis_comm_weights.detach(),
                                              None if self.pre_vecs is None else self.pre_vecs.detach(),
                                              self.basis_vecs.detach(), None if bias is None else bias.detach())


def splice_static_embedding(embedded_text: Tensor, tokenized_text: Tensor, placeholder_token: int,
                            subj_static_embedding: Tensor, num_unet_ca_layers: int = 16,
                            out_dtype: torch.dtype = torch.float32):
    """The inference branch of EmbeddingManager.get_static_embedding for one placeholder string
    (embedding_manager.py:1341-1352, 1360-1366, 1510-1552): the prompt's token embeddings `[B, N, 768]` repeated over the
    16 cross-attention layers (instance-major / layer-minor) with the K static vectors of each layer written at the K
    positions that start at the FIRST occurrence of the placeholder in each prompt; prompts without it pass through.
    Returns (embedded_text [B*16, N, 768], tokenized_text [B*16, N]) like the reference.  Two launches, no host sync."""
    if subj_static_embedding.shape[0] != num_unet_ca_layers:
        raise ValueError("splice_static_embedding: one static embedding per cross-attention layer expected")
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("splice_static_embedding: fp32 or bf16 output")
    first = ops.first_token_pos(tokenized_text, int(placeholder_token))
    out = ops.splice_static_embedding(embedded_text, first, subj_static_embedding, out_dtype == torch.bfloat16)
    B, N = tokenized_text.shape
    tok = tokenized_text.unsqueeze(1).expand(B, num_unet_ca_layers, N).reshape(B * num_unet_ca_layers, N)
    return out, tok
