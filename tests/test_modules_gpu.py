"""Module-level parity on the B200: the host-side mirror of the reference's modules (adaface_b200.attention,
.embedding, .ddim, .unet), running the CUDA kernels through the C ABI, against (a) the golden outputs of the
UNMODIFIED reference modules (tests/golden/*.npz) and (b) the CPU oracle on the same seeded inputs.

Tolerances: north_star's per-block relative L2 <= 1e-2 in bf16; final 50-step DDIM latents cosine >= 0.995.
Run with `-m gpu`."""
import math

import numpy as np
import pytest
import torch

from conftest import cosine, load_golden, rel_l2
from oracle import params as P
from oracle import port

pytestmark = pytest.mark.gpu

BLOCK_TOL = 1e-2


@pytest.fixture(scope="module")
def afb(cuda_device):
    import adaface_b200
    return adaface_b200


def _st_module(afb, c, sd, dev):
    st = afb.SpatialTransformer(c, 8, c // 8, depth=1, context_dim=768)
    st.load_state_dict(sd, strict=True)
    return st.to(dev).eval()


def _st_case(name):
    g = load_golden(name)
    c, hw, b, seed, sub = [int(v) for v in g["meta"]]
    return g, c, hw, b, seed, sub, P.init_spatial_transformer(c, seed), P.randn(f"{name}.x", (b, c, hw, hw), seed), \
        P.randn(f"{name}.ctx", (b, 77, 768), seed)


# ------------------------------------------------------------------------------------------- SpatialTransformer
@pytest.mark.parametrize("name", ["st_small", "st_c640", "st_c1280", "st_config1"])
def test_spatial_transformer_vs_reference_golden(afb, cuda_device, name):
    """fp32 NCHW in / out, callable + tuple context - the reference's own calling convention (config 1 = st_config1)."""
    g, c, hw, b, seed, sub, sd, x, ctx = _st_case(name)
    st = _st_module(afb, c, sd, cuda_device)
    cd = ctx.to(cuda_device)
    with torch.no_grad():
        y = st(x.to(cuda_device), context=lambda: ((cd, cd), None))
        y_plain = st(x.to(cuda_device), context=cd)
    assert y.dtype == torch.float32 and y.shape == x.shape and y.is_contiguous()
    assert torch.equal(y, y_plain)
    assert rel_l2(y[:, :, ::sub, ::sub], g["out"]) < BLOCK_TOL
    # what the block ADDS to its input (the residual passes through in fp32) must match too
    ref_full = port.spatial_transformer(x, sd, lambda: ((ctx, ctx), None)) if hw <= 16 else None
    if ref_full is not None:
        assert rel_l2(y.cpu() - x, ref_full - x) < BLOCK_TOL


def test_spatial_transformer_mask_and_split_context(afb, cuda_device):
    name = "st_mask_vk"
    g, c, hw, b, seed, _, sd, x, ctx_v = _st_case(name)
    ctx_k = P.randn(f"{name}.ctx_k", ctx_v.shape, seed)
    mask = (P.randn(f"{name}.mask", (b, 1, 64, 64), seed) > -0.3).float()
    st = _st_module(afb, c, sd, cuda_device)
    d = cuda_device
    with torch.no_grad():
        y = st(x.to(d), context=lambda: ((ctx_v.to(d), ctx_k.to(d)), None), mask=mask.to(d))
    assert rel_l2(y, g["out"]) < BLOCK_TOL


@pytest.mark.parametrize("c,hw", [(320, 16), (640, 8), (1280, 8)])
def test_spatial_transformer_bf16_channels_last_path(afb, cuda_device, c, hw):
    """The format the bf16 U-Net shell runs in: tokens are the memory layout, output stays channels-last bf16."""
    seed, b = 21, 3
    sd = P.init_spatial_transformer(c, seed)
    x = P.randn("stcl.x", (b, c, hw, hw), seed)
    ctx = P.randn("stcl.ctx", (b, 77, 768), seed)
    st = _st_module(afb, c, sd, cuda_device)
    xb = x.to(cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        y = st(xb, context=ctx.to(cuda_device))
    assert y.dtype == torch.bfloat16 and y.shape == x.shape and y.is_contiguous(memory_format=torch.channels_last)
    ref = port.spatial_transformer(xb.float().cpu(), sd, ctx)
    assert rel_l2(y.float(), ref) < BLOCK_TOL


def test_spatial_transformer_weight_update_invalidates_packed_weights(afb, cuda_device):
    c, hw, b, seed = 320, 8, 1, 22
    sd = P.init_spatial_transformer(c, seed)
    st = _st_module(afb, c, sd, cuda_device)
    x = P.randn("stw.x", (b, c, hw, hw), seed).to(cuda_device)
    ctx = P.randn("stw.ctx", (b, 77, 768), seed).to(cuda_device)
    with torch.no_grad():
        y0 = st(x, context=ctx)
        sd2 = P.init_spatial_transformer(c, seed + 1)
        st.load_state_dict(sd2)
        y1 = st(x, context=ctx)
    ref1 = port.spatial_transformer(x.cpu(), sd2, ctx.cpu())
    assert rel_l2(y1, ref1) < BLOCK_TOL and rel_l2(y0, ref1) > 0.05


# ------------------------------------------------------------------------------------------- blocks
def test_cross_attention_standalone(afb, cuda_device):
    g = load_golden("attn_vars")
    c, n, b, seed = [int(v) for v in g["meta"]]
    name = "attn_vars"
    ca = afb.CrossAttention(c, context_dim=768, heads=8, dim_head=c // 8)
    sd = {k: P.draw(f"{name}.{k}", tuple(v.shape), seed) for k, v in ca.state_dict().items()}
    ca.load_state_dict(sd)
    ca = ca.to(cuda_device).eval()
    x = P.randn(f"{name}.x", (b, n, c), seed)
    ctx = P.randn(f"{name}.ctx", (b, 77, 768), seed)
    with torch.no_grad():
        out = ca(x.to(cuda_device), context=ctx.to(cuda_device))
    assert out.dtype == torch.float32
    assert rel_l2(out, g["out"]) < BLOCK_TOL


def test_basic_transformer_block_and_self_attention(afb, cuda_device):
    c, n, b, seed = 640, 64, 2, 23
    blk = afb.BasicTransformerBlock(c, 8, c // 8, context_dim=768)
    sd = {k: P.draw(f"btb.{k}", tuple(v.shape), seed) for k, v in blk.state_dict().items()}
    blk.load_state_dict(sd)
    blk = blk.to(cuda_device).eval()
    x = P.randn("btb.x", (b, n, c), seed)
    ctx = P.randn("btb.ctx", (b, 77, 768), seed)
    mask = (P.randn("btb.mask", (b, 1, 8, 8), seed) > -0.5).float()
    with torch.no_grad():
        y = blk(x.to(cuda_device), context=ctx.to(cuda_device), mask=mask.to(cuda_device))
    ref = port.basic_transformer_block(x, sd, "", ctx, mask)
    assert rel_l2(y, ref) < BLOCK_TOL
    # self-attention through the public CrossAttention.forward
    with torch.no_grad():
        a = blk.attn1(x.to(cuda_device), mask=mask.to(cuda_device))
    assert rel_l2(a, port.cross_attention(x, sd, "attn1.", None, mask)) < BLOCK_TOL


def test_unsupported_modes_raise(afb, cuda_device):
    ca = afb.CrossAttention(320, context_dim=768, heads=8, dim_head=40).to(cuda_device).eval()
    x = torch.zeros((1, 64, 320), device=cuda_device)
    ca.save_attn_vars = True
    with pytest.raises(NotImplementedError):                 # the capture exists for cross-attention layers only
        ca(x)
    ff = afb.FeedForward(320, glu=True)
    with pytest.raises(NotImplementedError):
        afb.FeedForward(320, glu=False)
    assert ff is not None


def test_cross_attention_save_attn_vars_vs_reference_golden(afb, cuda_device):
    """attention.py:245-255: cached q * d^-1/4, attn and attnscore as [B, 8, N, 77], against the reference module."""
    g = load_golden("attn_vars")
    c, n, b, seed = [int(v) for v in g["meta"]]
    name = "attn_vars"
    ca = afb.CrossAttention(c, context_dim=768, heads=8, dim_head=c // 8)
    sd = {k: P.draw(f"{name}.{k}", tuple(v.shape), seed) for k, v in ca.state_dict().items()}
    ca.load_state_dict(sd)
    ca = ca.to(cuda_device).eval()
    ca.save_attn_vars = True
    x = P.randn(f"{name}.x", (b, n, c), seed)
    ctx = P.randn(f"{name}.ctx", (b, 77, 768), seed)
    with torch.no_grad():
        out = ca(x.to(cuda_device), context=ctx.to(cuda_device))
    assert rel_l2(out, g["out"]) < BLOCK_TOL
    act = ca.cached_activations
    assert set(act) == {"q", "attn", "attnscore"}
    for k in ("q", "attn", "attnscore"):
        assert tuple(act[k].shape) == tuple(g[k].shape), k
        assert rel_l2(act[k], g[k]) < BLOCK_TOL, k


@pytest.mark.parametrize("ks", [2, 3, 4])
def test_cross_attention_conv_attn_rows(afb, cuda_device, ks):
    """attention.py:208-216 through CrossAttention.forward: a callable context that returns placeholder2indices turns
    on the convolutional-attention row replacement for the subject tokens (ks*ks of them) of the listed samples."""
    c, hw, b, seed = 320, 8, 3, 31
    n, heads = hw * hw, 8
    ca = afb.CrossAttention(c, context_dim=768, heads=heads, dim_head=c // heads)
    sd = {k: P.draw(f"cva.{k}", tuple(v.shape), seed) for k, v in ca.state_dict().items()}
    ca.load_state_dict(sd)
    ca = ca.to(cuda_device).eval()
    ca.use_conv_attn_kernel_size, ca.infeat_size = ks, (hw, hw)
    x = P.randn("cva.x", (b, n, c), seed)
    ctx = P.randn("cva.ctx", (b, 77, 768), seed)
    per = max(9, ks * ks)                                        # ks = 4 needs 16 subject tokens per instance
    idx_b = torch.tensor([0] * per + [2] * per)
    idx_n = torch.tensor(list(range(4, 4 + per)) * 2)
    p2i = {"z": (idx_b, idx_n)}
    cd = ctx.to(cuda_device)
    with torch.no_grad():
        out = ca(x.to(cuda_device), context=lambda: ((cd, cd), p2i))
        plain = ca(x.to(cuda_device), context=cd)
    # oracle: scores -> row replacement -> softmax -> . v -> to_out
    d = c // heads
    q, k, v = x @ sd["to_q.weight"].t(), ctx @ sd["to_k.weight"].t(), ctx @ sd["to_v.weight"].t()
    sim = port.attention_scores(q, k, heads, d ** -0.5)
    split = lambda t_: t_.reshape(b, t_.shape[1], heads, d).permute(0, 2, 1, 3).reshape(b * heads, t_.shape[1], d)
    sim = port.replace_rows_by_conv_attn(sim.reshape(b * heads, n, 77), split(q), split(k), (idx_b, idx_n), (hw, hw), ks,
                                         heads, d ** -0.5).reshape(b, heads, n, 77)
    o = torch.matmul(sim.softmax(dim=-1), v.reshape(b, 77, heads, d).permute(0, 2, 1, 3)).permute(0, 2, 1, 3).reshape(b, n, c)
    ref = o @ sd["to_out.0.weight"].t() + sd["to_out.0.bias"]
    assert rel_l2(out, ref) < BLOCK_TOL
    assert rel_l2(out[1], plain[1]) < 6e-3 and rel_l2(out[0], plain[0]) > 1e-2     # sample 1 has no subject (explicit vs fused path: bf16 P)


# ------------------------------------------------------------------------------------------- configs 3 and 5
def test_spatial_transformer_two_subjects_ada_wiring(afb, cuda_device):
    """BASELINE.json configs[2] structure ('a z hugging a y'): two subjects with their own AdaEmbedding weights, K = 9
    rows each at positions 2..10 and 13..21, recomputed from this layer's features and spliced into the context of
    the first n_instances samples (the cond half); the remaining samples keep the static context.  Oracle: the port's
    functions composed along the reconstructed call site."""
    from adaface_b200.unet import AdaPromptContext, AdaSubject
    c, hw, b, n_inst, seed, K, r, layer_idx, ca = 320, 16, 4, 2, 33, 9, 12, 1, 0
    n = hw * hw
    d = cuda_device
    sd = P.init_spatial_transformer(c, seed)
    st = _st_module(afb, c, sd, d)
    x = P.randn("ms.x", (b, c, hw, hw), seed)
    ctx = P.randn("ms.ctx", (b, 77, 768), seed)
    time_emb = P.randn("ms.time_emb", (b, 1280), seed)
    subjects, oracle_subjects = [], []
    for si, pos in enumerate((list(range(2, 11)), list(range(13, 22)))):
        asd = P.init_ada_embedding(seed + 10 * si, K, r, 2)
        ada = afb.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4, r=r, init_vecs=torch.zeros(2, 768))
        ada.load_state_dict(asd, strict=True)
        ada = ada.to(d).eval()
        probes = P.randn(f"ms.{si}.probes", (16, 768), seed, 0.5)
        extra = P.randn(f"ms.{si}.extra", (16, n_inst, 768), seed, 0.5)
        subjects.append(AdaSubject(ada, pos, probes.to(d), extra.to(d), n_instances=n_inst))
        oracle_subjects.append((asd, pos, probes[ca], extra[ca]))
    ada_ctx = AdaPromptContext(subjects)
    lc = ctx.to(d).to(torch.bfloat16)

    def with_ada(comps):
        new = ada_ctx.apply(layer_idx, comps, lc, time_emb.to(d))
        return (new, new), None
    with_ada.wants_layer_attn_components = True
    with torch.no_grad():
        y = st(x.to(d), context=with_ada)

    # ---- oracle
    pre = "transformer_blocks.0."
    t = port.group_norm(x, sd["norm.weight"], sd["norm.bias"])
    t = t.reshape(b, c, n).transpose(1, 2) @ sd["proj_in.weight"].reshape(c, c).t() + sd["proj_in.bias"]
    n1 = port.layer_norm(t, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"])
    x1 = port.cross_attention(n1, sd, pre + "attn1.") + t
    n2 = port.layer_norm(x1, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"])
    q = n2 @ sd[pre + "attn2.to_q.weight"].t()
    ctx2 = ctx.clone()
    for asd, pos, probe, extra in oracle_subjects:
        vecs, _ = port.ada_embedding(asd, ca, x1[:n_inst], q[:n_inst], sd[pre + "attn2.to_k.weight"], time_emb[:n_inst],
                                     probe, extra, K, r, infeat_size=(hw, hw))
        ctx2[:n_inst, pos] = vecs
    x2 = x1 + port.cross_attention(n2, sd, pre + "attn2.", ctx2)
    n3 = port.layer_norm(x2, sd[pre + "norm3.weight"], sd[pre + "norm3.bias"])
    x3 = port.geglu_feedforward(n3, sd, pre + "ff.") + x2
    ref = (x3 @ sd["proj_out.weight"].reshape(c, c).t() + sd["proj_out.bias"]).transpose(1, 2).reshape(b, c, hw, hw) + x
    assert rel_l2(y, ref) < BLOCK_TOL
    assert rel_l2(y.cpu() - x, ref - x) < 2 * BLOCK_TOL
    # the spliced rows really changed the result for the conditioned samples only
    with torch.no_grad():
        y_static = st(x.to(d), context=ctx.to(d))
    assert rel_l2(y[n_inst:], y_static[n_inst:]) < 1e-6 and rel_l2((y - y_static)[:n_inst].cpu(), torch.zeros(1)) > 0


def test_config5_shapes_96x96_latent(afb, cuda_device):
    """BASELINE.json configs[4]: 768x768 images = 96x96 latents, N = 9216 / 2304 tokens.  The eager score tensor is
    2.7 GB per sample per layer, so the full-size self-attention is checked (a) against the oracle on a slice of the
    query rows and (b) through its invariance to how the queries are chunked; the second-level block (48x48, C = 640)
    is compared in full."""
    from adaface_b200 import ops
    d = cuda_device
    b, n, c, heads = 2, 96 * 96, 320, 8
    qkv = P.randn("c5.qkv", (b, n, 3 * c), 41, 1.5).to(torch.bfloat16)
    qd = qkv.to(d)
    scale = (c // heads) ** -0.5
    full = ops.attention(qd[..., :c], qd[..., c:2 * c], qd[..., 2 * c:], heads, scale, None)
    rows = torch.arange(1000, 1000 + 384)
    qf, kf, vf = qkv[..., :c].float(), qkv[..., c:2 * c].float(), qkv[..., 2 * c:].float()
    sim = port.attention_scores(qf[:, rows], kf, heads, scale)
    ref = torch.matmul(sim.softmax(-1), vf.reshape(b, n, heads, c // heads).permute(0, 2, 1, 3)).permute(0, 2, 1, 3)
    assert rel_l2(full[:, rows].float(), ref.reshape(b, len(rows), c)) < 6e-3
    part = torch.cat([ops.attention(qd[:, lo:hi, :c].contiguous(), qd[..., c:2 * c], qd[..., 2 * c:], heads, scale, None)
                      for lo, hi in ((0, 512), (512, 5000), (5000, n))], dim=1)
    assert torch.equal(part, full)
    # second level of the 96x96 ladder: 48x48 tokens, C = 640, with the 16 identity tokens of the Arc2Face slot (4..19)
    c2, hw2, seed = 640, 48, 42
    sd = P.init_spatial_transformer(c2, seed)
    st = _st_module(afb, c2, sd, d)
    x = P.randn("c5.x", (1, c2, hw2, hw2), seed)
    ctx = P.randn("c5.ctx", (1, 77, 768), seed)
    ctx[:, 4:20] = P.randn("c5.id", (1, 16, 768), seed, 0.7)
    with torch.no_grad():
        y = st(x.to(d), context=ctx.to(d))
    assert rel_l2(y, port.spatial_transformer(x, sd, ctx)) < BLOCK_TOL


# ------------------------------------------------------------------------------------------- AdaEmbedding
def test_ada_embedding_module_vs_reference_golden(afb, cuda_device):
    """All 16 layers against the reference class called standalone (H=2: K=9, fg 5 / bg 4, r=12; eval; fp32
    arithmetic, i.e. the forced CPU autocast neutralised - SURVEY.md 8c)."""
    g = load_golden("ada")
    seed, b, K, r = [int(v) for v in g["meta"]]
    sd = P.init_ada_embedding(seed, K, r, 2)
    ada = afb.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4, r=r,
                           init_vecs=torch.zeros(2, 768))
    ada.load_state_dict(sd, strict=True)
    ada = ada.to(cuda_device).eval()
    layer_of = {v: k for k, v in port.LAYER_IDX2CA.items()}
    d = cuda_device
    worst = 0.0
    for ca_idx in range(16):
        c = P.CA_INFEAT_DIMS[ca_idx]
        hw = 16 if c == 320 else 8
        n = hw * hw
        x = P.randn(f"ada.{ca_idx}.x", (b, n, c), seed)
        q = P.randn(f"ada.{ca_idx}.q", (b, n, c), seed, 2.0)
        to_k = torch.nn.Linear(768, c, bias=False)
        to_k.weight.data.copy_(P.draw(f"ada.{ca_idx}.to_k", (c, 768), seed))
        time_emb = P.randn("ada.time_emb", (b, 1280), seed)
        probe = P.randn(f"ada.{ca_idx}.probe", (768,), seed, 0.5)
        extra = P.randn(f"ada.{ca_idx}.extra", (b, 768), seed, 0.5)
        comps = {"x": x.to(d), "q": q.to(d), "to_k": to_k.to(d), "infeat_size": (hw, hw)}
        with torch.no_grad():
            out, pooled = ada(layer_of[ca_idx], comps, time_emb.to(d), probe.to(d), extra.to(d))
        assert out.shape == (b, K, 768) and "fg_out" not in pooled
        e = rel_l2(out, g[f"out_{ca_idx}"])
        worst = max(worst, e)
        assert e < BLOCK_TOL, (ca_idx, e)
        assert rel_l2(pooled["bg_out"], g[f"bg_{ca_idx}"]) < BLOCK_TOL
        assert rel_l2(pooled["attn_fg"], g[f"attn_fg_{ca_idx}"]) < BLOCK_TOL
    # masked pooling on layer 0
    x = P.randn("ada.0.x", (b, 256, 320), seed)
    q = P.randn("ada.0.q", (b, 256, 320), seed, 2.0)
    to_k = torch.nn.Linear(768, 320, bias=False)
    to_k.weight.data.copy_(P.draw("ada.0.to_k", (320, 768), seed))
    mask = (P.randn("ada.mask", (b, 1, 64, 64), seed) > -0.3).float()
    with torch.no_grad():
        out, _ = ada(1, {"x": x.to(d), "q": q.to(d), "to_k": to_k.to(d), "infeat_size": (16, 16)},
                     P.randn("ada.time_emb", (b, 1280), seed).to(d), P.randn("ada.0.probe", (768,), seed, 0.5).to(d),
                     P.randn("ada.0.extra", (b, 768), seed, 0.5).to(d), img_mask=mask.to(d))
    assert rel_l2(out, g["out_0_masked"]) < BLOCK_TOL
    print(f"AdaEmbedding worst layer rel-L2 {worst:.2e}")


def test_ada_embedding_zero_shot_basis(afb, cuda_device):
    """embedding_manager.py:860-868,886-890: do_zero_shot takes per-instance basis vectors [B, K, r, 768] and no bias.
    Pinned through the golden-checked slow path: with every instance given the slow model's own basis the two must
    agree exactly, and instance i with its own basis must equal the slow model run with that basis."""
    seed, b, K, r, ca_idx = 35, 3, 9, 12, 2
    d = cuda_device
    sd = P.init_ada_embedding(seed, K, r, 2)
    slow = afb.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4, r=r, init_vecs=torch.zeros(2, 768))
    slow.load_state_dict(sd, strict=True)
    with torch.no_grad():
        slow.bias.zero_()
    slow = slow.to(d).eval()
    zs = afb.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4, r=r, do_zero_shot=True)
    zs.load_state_dict({k: v for k, v in sd.items() if k in zs.state_dict()}, strict=True)
    zs = zs.to(d).eval()
    assert zs.basis_vecs is None and zs.bias is None and zs.pre_vecs is None
    c = P.CA_INFEAT_DIMS[ca_idx]
    layer_idx = {v: k for k, v in port.LAYER_IDX2CA.items()}[ca_idx]
    to_k = torch.nn.Linear(768, c, bias=False).to(d)
    comps = {"x": P.randn("zs.x", (b, 64, c), seed).to(d), "q": P.randn("zs.q", (b, 64, c), seed, 2.0).to(d), "to_k": to_k,
             "infeat_size": (8, 8)}
    args = (P.randn("zs.t", (b, 1280), seed).to(d), P.randn("zs.probe", (768,), seed, 0.5).to(d),
            P.randn("zs.extra", (b, 768), seed, 0.5).to(d))
    shared = torch.cat([slow.pre_vecs, slow.basis_vecs], dim=1).detach()                       # [K, r, 768]
    with torch.no_grad():
        ref, _ = slow(layer_idx, comps, *args)
        out, _ = zs(layer_idx, comps, *args, zs_basis_vecs=shared[None].expand(b, -1, -1, -1).contiguous())
        assert torch.equal(out, ref)
        per_inst = shared[None] + 0.1 * P.randn("zs.basis", (b, K, r, 768), seed).to(d)
        out2, _ = zs(layer_idx, comps, *args, zs_basis_vecs=per_inst)
        for i in range(b):
            slow.pre_vecs.copy_(per_inst[i, :, :2])
            slow.basis_vecs.copy_(per_inst[i, :, 2:])
            ref_i, _ = slow(layer_idx, comps, *args)
            assert torch.equal(out2[i], ref_i[i])
        with pytest.raises(ValueError):
            zs(layer_idx, comps, *args)


# ------------------------------------------------------------------------------------------- DDIM
def _toy(x, t, c):
    return torch.tanh(x * 0.7 + c.mean(dim=(1, 2)).reshape(-1, 1, 1, 1)) * (1.0 + t.reshape(-1, 1, 1, 1) / 1000.0)


class _ToyModel:
    def __init__(self, dev):
        betas = port.make_beta_schedule_linear()
        ac = np.cumprod(1.0 - betas)
        f = lambda a: torch.tensor(a, dtype=torch.float32, device=dev)
        self.num_timesteps, self.device = 1000, dev
        self.betas, self.alphas_cumprod = f(betas), f(ac)
        self.alphas_cumprod_prev = f(np.append(1.0, ac[:-1]))
        self.calls = []

    def apply_model(self, x, t, c):
        self.calls.append(int(t[0]))
        return _toy(x, t, c)


def test_ddim_sampler_vs_reference_golden(afb, cuda_device):
    """p_sample_ddim and the full 50-step loop (guidance annealing, timestep order, (cond, uncond) batch order)
    on the reference's toy model; the toy model itself is plain torch on the GPU, the update is the K14 kernel."""
    g = load_golden("ddim")
    seed, b = [int(v) for v in g["meta"]]
    d = cuda_device
    x = P.randn("ddim.x", (b, 4, 8, 8), seed).to(d)
    c = P.randn("ddim.c", (b, 77, 768), seed).to(d)
    uc = P.randn("ddim.uc", (b, 77, 768), seed).to(d)
    model = _ToyModel(d)
    s = afb.DDIMSampler(model)
    s.make_schedule(50, ddim_eta=0.0)
    for index in (49, 17, 0):
        t = torch.full((b,), int(s.ddim_timesteps[index]), dtype=torch.long, device=d)
        x_prev, pred_x0 = s.p_sample_ddim(x, c, t, index=index, guidance_scale=7.5, unconditional_conditioning=uc)
        # the toy model's tanh runs in torch on the GPU (a few ulp from the CPU's), the update itself is bit-exact
        assert rel_l2(x_prev, g[f"x_prev_{index}"]) < 1e-5
        assert rel_l2(pred_x0, g[f"pred_x0_{index}"]) < 1e-5
    for tag, gs in (("10_4", [10.0, 4.0]), ("10_10", [10.0, 10.0])):
        model.calls.clear()
        samples, inter = s.sample(S=50, batch_size=b, shape=(4, 8, 8), conditioning=c, guidance_scale=gs,
                                  unconditional_conditioning=uc, eta=0.0, x_T=x)
        assert model.calls == list(g[f"t_seq_{tag}"])
        assert rel_l2(samples, g[f"samples_{tag}"]) < 2e-5, tag
        assert len(inter["x_inter"]) >= 2
    # no guidance: a single model call per step
    model.calls.clear()
    x_prev, _ = s.p_sample_ddim(x, c, torch.full((b,), 981, dtype=torch.long, device=d), index=49)
    e = _toy(x, torch.full((b,), 981, device=d), c)
    ref_prev, _ = port.cfg_ddim_update(x.cpu(), e.cpu(), e.cpu(), 1.0, *s.step_scalars(49)[:3], s.step_scalars(49)[3])
    assert rel_l2(x_prev, ref_prev) < 1e-5 and len(model.calls) == 1


# ------------------------------------------------------------------------------------------- full U-Net
def _load_unet(afb, seed, dev):
    from adaface_b200.unet import SD15_UNET_CONFIG, UNetModel
    with torch.device("meta"):
        unet = UNetModel(**SD15_UNET_CONFIG)
    sd = P.init_unet(seed)
    unet.load_state_dict(sd, assign=True, strict=True)
    return unet.to(dev).eval()


@pytest.mark.slow
def test_unet_step_vs_reference_golden(afb, cuda_device):
    """One full SD-1.5 U-Net call (B = 2: cond + uncond, 64x64 latent, layerwise context) against the reference
    UNetModel's eps and the output of each of its 16 SpatialTransformers (bf16 end to end, incl. the torch ResBlocks)."""
    g = load_golden("unet_step")
    seed, hw = [int(v) for v in g["meta"]]
    d = cuda_device
    unet = _load_unet(afb, seed, d)
    x_T = P.randn("unet_step.x_T", (1, 4, hw, hw), seed)
    c_cond = P.randn("unet_step.c_cond", (16, 77, 768), seed)
    c_uncond = P.randn("unet_step.c_uncond", (16, 77, 768), seed)
    outs = []
    hooks = [m.register_forward_hook(lambda mod, inp, out: outs.append(out)) for _, m in unet.spatial_transformers()]
    with torch.no_grad():
        eps = unet(torch.cat([x_T, x_T]).to(d), torch.full((2,), 981, dtype=torch.long, device=d),
                   context=torch.cat([c_cond, c_uncond]).to(d),
                   extra_info={"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1,
                               "placeholder2indices": None})
    for h in hooks:
        h.remove()
    assert len(outs) == 16 and eps.dtype == torch.float32
    errs = [rel_l2(o[:, ::8, ::4, ::4].float(), g[f"st_out_sub_{i}"]) for i, o in enumerate(outs)]
    e_eps = rel_l2(eps, g["eps"])
    print("per-ST rel-L2 through the U-Net:", " ".join(f"{e:.1e}" for e in errs), f"| eps {e_eps:.2e}")
    # errors accumulate across 25 bf16 blocks (ResBlocks included); the per-block bar (1e-2) is checked with
    # exact inputs above - here the chain must stay within 3e-2
    assert max(errs) < 3e-2 and e_eps < 3e-2
    assert cosine(eps.cpu(), g["eps"]) > 0.999


def test_unet_config5_geometry_96x96(afb, cuda_device):
    """BASELINE.json configs[4] geometry (768x768 -> 96x96 latent, N = 9216 / 2304 / 576 / 144): the full U-Net call runs
    through the fallback geometry of the shell (no 128-pixel boxes at width 96: cuDNN convolutions, interpolate +
    conv Upsample) with the virtual concatenations and every path kernel at the large sizes.  No CPU oracle at this size
    (the eager score tensors need > 20 GB), so the check is batch independence: a two-sample call must
    equal the two one-sample calls, and the output must be finite and non-trivial."""
    d = cuda_device
    unet = _load_unet(afb, 17, d)
    x = P.randn("cfg5.x", (2, 4, 96, 96), 23).to(d)
    ctx = P.randn("cfg5.ctx", (32, 77, 768), 23).to(d)
    info = {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1, "placeholder2indices": None}
    t = torch.full((2,), 501, dtype=torch.long, device=d)
    with torch.no_grad():
        both = unet(x, t, context=ctx, extra_info=info)
        one = [unet(x[i:i + 1], t[i:i + 1], context=ctx[16 * i:16 * (i + 1)], extra_info=info) for i in range(2)]
    assert both.shape == (2, 4, 96, 96) and bool(torch.isfinite(both).all())
    assert float(both.std()) > 1e-3
    # (cuDNN may pick different algorithms per batch size: bf16-level agreement, not bit equality)
    assert rel_l2(both, torch.cat(one)) < 2e-2


@pytest.mark.slow
@pytest.mark.parametrize("name", ["ddim10_small", "ddim50"])
def test_ddim_full_loop_vs_reference_golden(afb, cuda_device, name):
    """north_star: final DDIM latents cosine >= 0.995 against the reference's fp32 path (n_samples = 1 + CFG (10, 4),
    eta = 0): ddim50 = 50 steps at 64x64 (config 2 scaled to one sample), ddim10_small = 10 steps at 16x16."""
    from adaface_b200.unet import DiffusionModel
    g = load_golden(name)
    seed, hw, steps = [int(v) for v in g["meta"]]
    d = cuda_device
    unet = _load_unet(afb, seed, d)
    model = DiffusionModel(unet, use_cuda_graph=True)
    sampler = afb.DDIMSampler(model)
    x_T = P.randn(f"{name}.x_T", (1, 4, hw, hw), seed).to(d)
    c_cond = P.randn(f"{name}.c_cond", (16, 77, 768), seed).to(d)
    c_uncond = P.randn(f"{name}.c_uncond", (16, 77, 768), seed).to(d)
    info = {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1,
            "placeholder2indices": None}
    samples, _ = sampler.sample(S=steps, batch_size=1, shape=(4, hw, hw), conditioning=(c_cond, ["a z"], info),
                                guidance_scale=[10.0, 4.0], unconditional_conditioning=(c_uncond, [""], info),
                                eta=0.0, x_T=x_T)
    cos = cosine(samples.cpu(), g["samples"])
    print(f"{name}: final latent cosine {cos:.5f}, rel-L2 {rel_l2(samples, g['samples']):.3e}")
    assert cos >= 0.995


# ---------------------------------------------------------------------------------------------- N1: ResBlock / Upsample
def test_resblock_and_upsample_vs_reference_golden(afb, cuda_device):
    """SURVEY.md 8f N1 (openaimodel.py:95-123, 167-281): the U-Net shell's ResBlock (fused GroupNorm+SiLU with the
    bias / time-embedding fold, skip-GEMM tail) - also fed with the decoder's virtual concatenation - and the
    upsample-folded Upsample against outputs of the reference's own modules."""
    from adaface_b200 import unet as U
    from oracle.gen_golden import RESBLOCK_CASES, RESBLOCK_CH_STRIDE, resblock_inputs
    g = load_golden("resblock")
    seed, d, cl = 21, cuda_device, torch.channels_last
    for tag, cin, cout, hw in RESBLOCK_CASES:
        rb = U.ResBlock(cin, 1280, 0.0, out_channels=cout)
        rb.load_state_dict(P.init_resblock(cin, cout, seed), strict=True)
        rb = rb.to(d).eval()
        x, emb = resblock_inputs(tag, cin, hw, seed)
        xb = x.to(d).to(torch.bfloat16).contiguous(memory_format=cl)
        with torch.no_grad():
            if tag.startswith("cat"):
                half = cin // 2
                pair = U.CatPair(xb[:, :half].contiguous(memory_format=cl), xb[:, half:].contiguous(memory_format=cl))
                assert pair.virtual()                            # concatenations stay virtual (both GroupNorm kernels read two sources)
                y = rb(pair, emb.to(d))
            else:
                y = rb(xb, emb.to(d))
        assert y.shape == (2, cout, hw, hw)
        assert rel_l2(y.float()[:, ::RESBLOCK_CH_STRIDE], g[tag]) < 1e-2, tag
    up = U.Upsample(640, True)
    up.load_state_dict(P.init_upsample(640, seed), strict=True)
    up = up.to(d).eval()
    x = P.randn("rb.up.x", (2, 640, 16, 16), seed).to(d).to(torch.bfloat16).contiguous(memory_format=cl)
    with torch.no_grad():
        y = up(x)
    assert y.shape == (2, 640, 32, 32)
    assert rel_l2(y.float()[:, ::RESBLOCK_CH_STRIDE], g["up"]) < 1e-2


def test_reference_checkpoint_to_layerwise_prompt(cuda_device):
    """SURVEY.md 8f N3 + N2 end to end: the reference-format checkpoint fixture (live reference modules, pickled) ->
    this package's StaticLayerwiseEmbedding on the GPU == what the reference modules computed (golden), then spliced into
    a batch of prompts == the oracle's splice, bit for bit."""
    import os
    from adaface_b200.checkpoint import load_embedding_checkpoint
    from adaface_b200.embedding import splice_static_embedding
    from conftest import load_golden
    fixture = os.path.join(os.path.dirname(__file__), "golden", "embeddings_gs-toy.pt")
    if not os.path.exists(fixture):
        pytest.skip("checkpoint fixture not generated")
    g = load_golden("embeddings_ckpt")
    ck = load_embedding_checkpoint(fixture)
    static = ck.static_embeddings(cuda_device)
    for km in ("z", "y"):
        assert static[km].shape == (16, ck.token2num_vectors[km], 768)
        assert rel_l2(static[km], torch.from_numpy(g[f"static_{km}"])) < 1e-5
    emb = P.randn("ckpt.prompt", (4, 77, 768), 5, 0.05)
    tok = torch.full((4, 77), 320, dtype=torch.long)
    tok[0, 3] = ck.string_to_token["z"]
    tok[2, 70] = ck.string_to_token["z"]
    tok[2, 75] = ck.string_to_token["z"]
    tok[3, 76] = ck.string_to_token["z"]                  # last position: one of the four vectors fits
    out, tok16 = splice_static_embedding(emb.to(cuda_device), tok.to(cuda_device), ck.string_to_token["z"], static["z"], 16,
                                         torch.bfloat16)
    # (the oracle's splice writes all K rows and would run off the prompt for sample 3: restated per vector here)
    want = emb.unsqueeze(1).repeat(1, 16, 1, 1).reshape(64, 77, 768).clone()
    for b, first in ((0, 3), (2, 70), (3, 76)):
        for k in range(4):
            if first + k < 77:
                want[b * 16:(b + 1) * 16, first + k] = static["z"].cpu()[:, k]
    assert torch.equal(out.cpu(), want.to(torch.bfloat16))
    assert tok16.shape == (64, 77) and torch.equal(tok16[16:32].cpu(), tok[1].expand(16, -1))
