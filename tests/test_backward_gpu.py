"""Backward of the inversion step (BASELINE.json configs[3]: backward through cross-attention, gradients only on the
AdaEmbedding parameters; SURVEY.md 8d config 4: gradients rel-L2 <= 2e-2 against fp32 autograd through the reference
modules).  The reference has no backward code of its own - torch autograd differentiates its modules - so the oracle
here is fp32 torch autograd through oracle/port.py (the restatement pinned to the reference's forward outputs) on the
same seeded inputs.  Kernel level first, then one SpatialTransformer with the AdaEmbedding call site wired in.

Built so far: everything between the block output and the AdaEmbedding parameters of that layer.  Not built (and raising
instead of returning wrong gradients): self-attention / GroupNorm backward and the pooler's gradient towards the layer
features, i.e. the chain that carries a layer's gradient further down into earlier layers of the U-Net."""
import math

import pytest
import torch

from conftest import rel_l2
from oracle import params as P
from oracle import port

pytestmark = pytest.mark.gpu

GRAD_TOL = 2e-2


@pytest.fixture(scope="module")
def ops(cuda_device):
    from adaface_b200 import ops as o
    return o


def _r(key, shape, scale=1.0, seed=61):
    return P.randn(key, shape, seed, scale)


@pytest.mark.parametrize("m,n,k", [(300, 320, 320), (1000, 640, 2560), (77 * 3, 1280, 768)])
def test_linear_dx_and_residual_grad(ops, cuda_device, m, n, k):
    d = cuda_device
    a, w, bias, res = _r("lb.a", (m, k)), _r("lb.w", (n, k), k ** -0.5), _r("lb.b", (n,), 0.1), _r("lb.r", (m, n))
    go = _r("lb.go", (m, n))
    ab, wb, rb = a.to(torch.bfloat16), w.to(torch.bfloat16), res.to(torch.bfloat16)
    a_ref, r_ref = ab.float().requires_grad_(True), rb.float().requires_grad_(True)
    (a_ref @ wb.float().t() + bias + r_ref).backward(go.to(torch.bfloat16).float())
    a_d, r_d = ab.to(d).requires_grad_(True), rb.to(d).requires_grad_(True)
    out = ops.linear(a_d, wb.to(d), bias.to(d), r_d, False)
    out.backward(go.to(torch.bfloat16).to(d))
    assert rel_l2(a_d.grad.float(), a_ref.grad) < 6e-3 and rel_l2(r_d.grad.float(), r_ref.grad) < 1e-6
    with pytest.raises(NotImplementedError):                  # weight gradients are not built: the U-Net is frozen
        ops.linear(a_d, wb.to(d).requires_grad_(True), None, None, False)


@pytest.mark.parametrize("m,c", [(500, 320), (260, 1280)])
def test_layernorm_folded_geglu_backward(ops, cuda_device, m, c):
    """d/da of GEGLU(LayerNorm(a)) through linear_ln_geglu: recomputed projection, geglu_bwd, dX GEMM, layernorm_bwd."""
    from adaface_b200.attention import BasicTransformerBlock
    d = cuda_device
    a = (_r("gb.a", (m, c)) * 1.3 + 0.2).to(torch.bfloat16)
    gamma, beta = 1 + 0.1 * _r("gb.g", (c,)), 0.1 * _r("gb.be", (c,))
    w, b = _r("gb.w", (8 * c, c), c ** -0.5), _r("gb.b", (8 * c,), 0.1)
    go = _r("gb.go", (m, 4 * c)).to(torch.bfloat16)
    a_ref = a.float().requires_grad_(True)
    h = port.layer_norm(a_ref, gamma, beta) @ w.t() + b
    av, gv = h.chunk(2, dim=-1)
    (av * port.gelu_erf(gv)).backward(go.float())
    # the product's fold: W' = W diag(gamma) (bf16), colsum, t = W beta + b, rows interleaved for the gated epilogue
    norm = torch.nn.LayerNorm(c)
    with torch.no_grad():
        norm.weight.copy_(gamma); norm.bias.copy_(beta)
    wg, t = BasicTransformerBlock._fold(w, b, norm)
    wi, ti = ops.interleave_geglu(wg, t)
    colsum = wi.float().sum(dim=1).contiguous()
    # row statistics as the producer GEMM would emit them: (sum, sum of squares) partials of the bf16 row
    parts = ops.stats_parts(c)
    stats = torch.zeros((m, parts, 2))
    stats[:, 0, 0], stats[:, 0, 1] = a.float().sum(1), (a.float() ** 2).sum(1)
    a_d = a.to(d).requires_grad_(True)
    out = ops.linear_ln_geglu(a_d, stats.to(d), wi.to(d), colsum.to(d), ti.to(d), 1e-5)
    assert rel_l2(out.float(), (av * port.gelu_erf(gv)).detach()) < 1e-2
    out.backward(go.to(d))
    assert rel_l2(a_d.grad.float(), a_ref.grad) < GRAD_TOL


@pytest.mark.parametrize("c,nq,nk", [(320, 700, 77), (640, 256, 77), (1280, 64, 77), (320, 100, 128)])
def test_cross_attention_backward(ops, cuda_device, c, nq, nk):
    d = cuda_device
    b, heads = 2, 8
    q, k, v = (_r(f"ab.{n}", (b, l, c), 0.8).to(torch.bfloat16) for n, l in (("q", nq), ("k", nk), ("v", nk)))
    go = _r("ab.go", (b, nq, c)).to(torch.bfloat16)
    scale = (c // heads) ** -0.5
    qr, kr, vr = (t.float().requires_grad_(True) for t in (q, k, v))
    sim = port.attention_scores(qr, kr, heads, scale)
    o = torch.matmul(sim.softmax(-1), vr.reshape(b, nk, heads, c // heads).permute(0, 2, 1, 3)).permute(0, 2, 1, 3).reshape(b, nq, c)
    o.backward(go.float())
    qd, kd, vd = (t.to(d).requires_grad_(True) for t in (q, k, v))
    out = ops.attention(qd, kd, vd, heads, scale, None)
    out.backward(go.to(d))
    assert rel_l2(qd.grad.float(), qr.grad) < GRAD_TOL
    assert rel_l2(kd.grad.float(), kr.grad) < GRAD_TOL and rel_l2(vd.grad.float(), vr.grad) < GRAD_TOL
    # the fused K|V projection buffer: k and v as column slices, gradient of the buffer assembled by autograd
    kv = torch.cat([k, v], dim=-1).to(d).requires_grad_(True)
    ops.attention(q.to(d), kv[..., :c], kv[..., c:], heads, scale, None).backward(go.to(d))
    assert rel_l2(kv.grad.float(), torch.cat([kr.grad, vr.grad], dim=-1)) < GRAD_TOL
    with pytest.raises(NotImplementedError):                  # self-attention backward is not built
        long_k = _r("ab.lk", (b, 256, c)).to(torch.bfloat16).to(d).requires_grad_(True)
        ops.attention(q.to(d), long_k, long_k, heads, scale, None)


def _ada_module(afb, asd, dev, K, r):
    ada = afb.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4, r=r, init_vecs=torch.zeros(2, 768))
    ada.load_state_dict(asd, strict=True)
    ada = ada.to(dev).eval()
    ada.want_pooler_attn = False
    return ada


def _layer_keys(ca):
    return ([f"poolers.{ca}.lora_to_k.weight", f"poolers.{ca}.lora_to_fg_q.weight", f"poolers.{ca}.lora_to_bg_q.weight",
             f"layer_coeff_maps.{ca}.weight", f"layer_coeff_maps.{ca}.bias", "pre_vecs", "basis_vecs"])


@pytest.mark.parametrize("ca,n", [(0, 400), (3, 256), (6, 64)])
def test_ada_embedding_parameter_gradients(cuda_device, ca, n):
    """AdaEmbedding.forward (pooler + head) at one layer: gradients of every tensor it trains - the three grouped-conv
    pooler weights, the coefficient map, the basis vectors and the bias - against fp32 autograd through the port."""
    import adaface_b200 as afb
    d = cuda_device
    K, r, b, seed = 9, 12, 3, 71
    c = port.CA_INFEAT_DIMS[ca]
    layer_idx = {v: k for k, v in port.LAYER_IDX2CA.items()}[ca]
    asd = P.init_ada_embedding(seed, K, r, 2)
    x, q = (_r(f"ag.{t}", (b, n, c), 1.2).to(torch.bfloat16) for t in ("x", "q"))
    to_k_w = _r("ag.tok", (c, 768), 768 ** -0.5)
    time_emb, probe, extra = _r("ag.te", (b, 1280)), _r("ag.pr", (768,), 0.5), _r("ag.ex", (b, 768), 0.5)
    go = _r("ag.go", (b, K, 768))
    # ---- oracle: fp32 autograd through the port
    ref_p = {k: v.clone().requires_grad_(True) for k, v in asd.items()}
    out_ref, _ = port.ada_embedding(ref_p, ca, x.float(), q.float(), to_k_w, time_emb, probe, extra, K, r, infeat_size=None)
    out_ref.backward(go)
    # ---- product
    ada = _ada_module(afb, asd, d, K, r)
    to_k = torch.nn.Linear(768, c, bias=False).to(d)
    with torch.no_grad():
        to_k.weight.copy_(to_k_w)
    to_k.requires_grad_(False)
    comps = {"x": x.to(d), "q": q.to(d), "to_k": to_k, "infeat_size": None}
    out, _ = ada(layer_idx, comps, time_emb.to(d), probe.to(d), extra.to(d))
    assert rel_l2(out, out_ref.detach()) < 5e-3
    out.backward(go.to(d))
    got = dict(ada.named_parameters())
    errs = {}
    for key in _layer_keys(ca):
        errs[key] = rel_l2(got[key].grad, ref_p[key].grad)
    errs["bias"] = rel_l2(got["bias"].grad[ca], ref_p["bias"].grad[ca])
    print(f"ca {ca} (C={c}, N={n}) gradient rel-L2:", " ".join(f"{k.split('.')[-2] if '.' in k else k}={v:.1e}" for k, v in errs.items()))
    assert max(errs.values()) < GRAD_TOL, errs
    # nothing else received a gradient: only this layer's tensors are on the path
    other = [k for k, p in got.items() if p.grad is not None and k not in _layer_keys(ca) + ["bias"]]
    assert all(float(got[k].grad.abs().max()) == 0.0 for k in other)


@pytest.mark.parametrize("c,hw,ca", [(320, 16, 0), (640, 8, 3)])
def test_inversion_step_through_one_spatial_transformer(cuda_device, c, hw, ca):
    """The block-level inversion step: frozen SpatialTransformer, two subjects' AdaEmbeddings wired in, loss = MSE of the
    block output against a target; the gradient runs back through proj_out, the GEGLU feed-forward (LayerNorm fold),
    attn2's output projection, the 77-key cross-attention and its fused K|V projection into the spliced context rows, the
    AdaEmbedding head and the pooler.  Every trained tensor of both subjects against fp32 autograd through the port's
    composed block."""
    import adaface_b200 as afb
    from adaface_b200.unet import AdaPromptContext, AdaSubject
    d = cuda_device
    b, n_inst, seed, K, r = 4, 2, 83, 9, 12
    layer_idx = {v: k for k, v in port.LAYER_IDX2CA.items()}[ca]
    sd = P.init_spatial_transformer(c, seed)
    x = _r("inv.x", (b, c, hw, hw)).to(torch.bfloat16)
    ctx = _r("inv.ctx", (b, 77, 768)).to(torch.bfloat16)
    target = _r("inv.tgt", (b, c, hw, hw))
    time_emb = _r("inv.te", (b, 1280))
    subjects = P.ada_subject_inputs("inv", 2, n_inst, seed, K, r)
    # ---- oracle
    ref_subjects = []
    for s in subjects:
        ref_subjects.append(dict(s, params={k: v.clone().requires_grad_(True) for k, v in s["params"].items()}))
    y_ref = port.spatial_transformer_ada(x.float(), sd, ctx.float(), ref_subjects, ca, time_emb)
    loss_ref = ((y_ref - target) ** 2).mean()
    loss_ref.backward()
    # ---- product
    st = afb.SpatialTransformer(c, 8, c // 8, depth=1, context_dim=768)
    st.load_state_dict(sd, strict=True)
    st = st.to(d).eval().requires_grad_(False)
    adas = [_ada_module(afb, s["params"], d, K, r) for s in subjects]
    ada_ctx = AdaPromptContext([AdaSubject(a, s["positions"], s["probes"].to(d), s["extra_means"].to(d), n_instances=n_inst)
                                for a, s in zip(adas, subjects)])
    lc = ctx.to(d)

    def with_ada(comps):
        new = ada_ctx.apply(layer_idx, comps, lc, time_emb.to(d))
        return (new, new), None
    with_ada.wants_layer_attn_components = True
    xd = x.to(d).contiguous(memory_format=torch.channels_last)
    y = st(xd, context=with_ada)
    assert rel_l2(y.float(), y_ref.detach()) < 1e-2
    loss = ((y.float() - target.to(d)) ** 2).mean()
    loss.backward()
    assert abs(float(loss) - float(loss_ref)) < 2e-2 * float(loss_ref)
    worst = {}
    for si, (ada, rs) in enumerate(zip(adas, ref_subjects)):
        got = dict(ada.named_parameters())
        for key in _layer_keys(ca):
            worst[f"s{si}.{key}"] = rel_l2(got[key].grad, rs["params"][key].grad)
        worst[f"s{si}.bias"] = rel_l2(got["bias"].grad[ca], rs["params"]["bias"].grad[ca])
    print(f"inversion step, C={c}: gradient rel-L2 per trained tensor:", " ".join(f"{k}={v:.1e}" for k, v in worst.items()))
    assert max(worst.values()) < GRAD_TOL, worst
    assert not lc.requires_grad and torch.equal(lc.cpu(), ctx)          # the caller's context was not written to
