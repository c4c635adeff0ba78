"""Generates tests/golden/*.npz by running the UNMODIFIED reference modules (imported from
/root/reference, CPU, fp32) on seeded inputs.  Build-container only; the fixtures it writes are
committed so that the GPU box (no reference there) can pin both the port and the CUDA path.

TEST INFRASTRUCTURE ONLY (see oracle/port.py header).

    python -m oracle.gen_golden                 # all fast cases
    python -m oracle.gen_golden unet_step ddim50    # the slow ones (minutes)

Inputs and weights are NOT stored: they are re-drawn from oracle/params.py by name and seed.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch

from . import params as P
from .port import LAYER_IDX2CA
from . import port
from . import ref_harness as RH

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _save(name, **arrays):
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    out = {}
    for k, v in arrays.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        out[k] = np.asarray(v)
    np.savez_compressed(path, **out)
    print(f"wrote {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


def _ref_st(ref, c, sd):
    st = ref.attention.SpatialTransformer(c, 8, c // 8, depth=1, context_dim=768)
    missing, unexpected = st.load_state_dict(sd, strict=True)
    return st.eval()


def st_inputs(case, c, hw, b, seed):
    x = P.randn(f"{case}.x", (b, c, hw, hw), seed)
    ctx = P.randn(f"{case}.ctx", (b, 77, 768), seed)
    return x, ctx


def case_st(ref, name, c, hw, b=2, seed=11, sub=1):
    sd = P.init_spatial_transformer(c, seed)
    st = _ref_st(ref, c, sd)
    x, ctx = st_inputs(name, c, hw, b, seed)
    with torch.no_grad():
        y = st(x, context=lambda: ((ctx, ctx), None))            # callable + tuple context
        y_plain = st(x, context=ctx)                             # a plain tensor also works
    assert torch.equal(y, y_plain)
    _save(name, out=y[:, :, ::sub, ::sub], out_norm=y.norm(), delta_norm=(y - x).norm(),
          meta=np.array([c, hw, b, seed, sub]))


def case_st_mask_vk(ref, name="st_mask_vk", c=320, hw=16, b=2, seed=12):
    sd = P.init_spatial_transformer(c, seed)
    st = _ref_st(ref, c, sd)
    x, ctx_v = st_inputs(name, c, hw, b, seed)
    ctx_k = P.randn(f"{name}.ctx_k", (b, 77, 768), seed)
    mask = (P.randn(f"{name}.mask", (b, 1, 64, 64), seed) > -0.3).float()
    with torch.no_grad():
        y = st(x, context=lambda: ((ctx_v, ctx_k), None), mask=mask)
    _save(name, out=y, meta=np.array([c, hw, b, seed, 1]))


def case_attn_vars(ref, name="attn_vars", c=320, n=64, b=2, seed=13):
    """CrossAttention with save_attn_vars=True: attention.py:245-255."""
    ca = ref.attention.CrossAttention(c, context_dim=768, heads=8, dim_head=c // 8).eval()
    sd = {k: P.draw(f"{name}.{k}", tuple(v.shape), seed) for k, v in ca.state_dict().items()}
    ca.load_state_dict(sd)
    ca.save_attn_vars = True
    x = P.randn(f"{name}.x", (b, n, c), seed)
    ctx = P.randn(f"{name}.ctx", (b, 77, 768), seed)
    with torch.no_grad():
        out = ca(x, context=ctx)
    act = ca.cached_activations
    _save(name, out=out, q=act["q"], attn=act["attn"], attnscore=act["attnscore"],
          meta=np.array([c, n, b, seed]))


def case_conv_attn(ref, name="conv_attn", seed=14):
    """replace_rows_by_conv_attn: ldm/util.py:701-879, ks in {2,3}."""
    heads, d, hw, b, t = 8, 40, 8, 3, 77
    n = hw * hw
    q = P.randn(f"{name}.q", (b * heads, n, d), seed)
    k = P.randn(f"{name}.k", (b * heads, t, d), seed)
    sim = P.randn(f"{name}.sim", (b * heads, n, t), seed)
    idx_b = torch.tensor([0] * 9 + [2] * 9)
    idx_n = torch.tensor(list(range(6, 15)) * 2)
    outs = {}
    for ks in (2, 3):
        outs[f"out_ks{ks}"] = ref.util.replace_rows_by_conv_attn(sim, q, k, (idx_b, idx_n), (hw, hw), ks,
                                                                heads, d ** -0.5)
    _save(name, meta=np.array([heads, d, hw, b, t, seed]), **outs)


def case_conv_attn_ks4(ref, name="conv_attn_ks4", seed=24):
    """replace_rows_by_conv_attn with ks = 4 (the asymmetric 1,2,1,2 padding branch, ldm/util.py:754): 16 subject tokens
    per instance at positions 6..21."""
    heads, d, hw, b, t = 8, 40, 8, 3, 77
    n = hw * hw
    q = P.randn(f"{name}.q", (b * heads, n, d), seed)
    k = P.randn(f"{name}.k", (b * heads, t, d), seed)
    sim = P.randn(f"{name}.sim", (b * heads, n, t), seed)
    idx_b = torch.tensor([0] * 16 + [2] * 16)
    idx_n = torch.tensor(list(range(6, 22)) * 2)
    out = ref.util.replace_rows_by_conv_attn(sim, q, k, (idx_b, idx_n), (hw, hw), 4, heads, d ** -0.5)
    _save(name, meta=np.array([heads, d, hw, b, t, seed]), out_ks4=out)


def build_ref_ada(ref, seed, K=9, r=12):
    sd = P.init_ada_embedding(seed, K, r, 2)
    ada = ref.embedding_manager.AdaEmbedding(num_vectors_per_subj_token=K, fg_emb_count=5, bg_emb_count=4,
                                             r=r, init_vecs=torch.zeros(2, 768), device_type="cpu")
    res = ada.load_state_dict(sd, strict=True)
    return ada.eval(), sd


def ada_inputs(name, ca_idx, n, b, seed):
    c = P.CA_INFEAT_DIMS[ca_idx]
    x = P.randn(f"{name}.{ca_idx}.x", (b, n, c), seed)
    q = P.randn(f"{name}.{ca_idx}.q", (b, n, c), seed, 2.0)
    to_k = P.draw(f"{name}.{ca_idx}.to_k", (c, 768), seed)
    time_emb = P.randn(f"{name}.time_emb", (b, 1280), seed)
    probe = P.randn(f"{name}.{ca_idx}.probe", (768,), seed, 0.5)
    extra = P.randn(f"{name}.{ca_idx}.extra", (b, 768), seed, 0.5)
    return x, q, to_k, time_emb, probe, extra


def case_ada(ref, name="ada", seed=15, b=2):
    """AdaEmbedding standalone with the section-3.2 contract, H=2 (K=9, fg 5 / bg 4, r=12), eval mode,
    fp32 (autocast neutralised) plus the as-written (CPU bf16 autocast) output for layer 0."""
    ada, _ = build_ref_ada(ref, seed)
    layer_of = {v: k for k, v in LAYER_IDX2CA.items()}
    arrays = {}
    for ca_idx in range(16):
        c = P.CA_INFEAT_DIMS[ca_idx]
        hw = 16 if c == 320 else 8
        n = hw * hw
        x, q, to_k, time_emb, probe, extra = ada_inputs(name, ca_idx, n, b, seed)
        lin = torch.nn.Linear(768, c, bias=False)
        lin.weight.data.copy_(to_k)
        comps = {"x": x, "q": q, "to_k": lin, "infeat_size": (hw, hw)}
        with torch.no_grad(), ref.fp32_autocast():
            out, pooled = ada(layer_of[ca_idx], comps, time_emb, probe, extra)
        arrays[f"out_{ca_idx}"] = out
        arrays[f"bg_{ca_idx}"] = pooled["bg_out"]
        arrays[f"attn_fg_{ca_idx}"] = pooled["attn_fg"]
        if ca_idx == 0:
            with torch.no_grad():
                out_bf16, _ = ada(layer_of[ca_idx], comps, time_emb, probe, extra)
            arrays["out_0_as_written_autocast"] = out_bf16.float()
    # masked variant on layer 0
    x, q, to_k, time_emb, probe, extra = ada_inputs(name, 0, 256, b, seed)
    lin = torch.nn.Linear(768, 320, bias=False)
    lin.weight.data.copy_(to_k)
    mask = (P.randn(f"{name}.mask", (b, 1, 64, 64), seed) > -0.3).float()
    with torch.no_grad(), ref.fp32_autocast():
        out, pooled = ada(1, {"x": x, "q": q, "to_k": lin, "infeat_size": (16, 16)}, time_emb, probe, extra,
                          img_mask=mask)
    arrays["out_0_masked"] = out
    _save(name, meta=np.array([seed, b, 9, 12]), **arrays)


def case_ddim(ref, name="ddim", seed=16):
    """Schedule (ddim.py:28-68), guidance annealing (:169-218), p_sample_ddim (:222-296) on a toy model."""
    b = 3

    def toy(x, t, c):
        # a deterministic stand-in for the U-Net; cond/uncond halves differ through c
        return torch.tanh(x * 0.7 + c.mean(dim=(1, 2)).reshape(-1, 1, 1, 1)) * (1.0 + t.reshape(-1, 1, 1, 1) / 1000.0)

    model = ref.ModelStandIn(apply_fn=toy)
    sampler = ref.CPUDDIMSampler(model)
    sampler.make_schedule(ddim_num_steps=50, ddim_eta=0.0, verbose=False)
    arrays = {"timesteps": sampler.ddim_timesteps, "alphas": np.asarray(sampler.ddim_alphas),
              "alphas_prev": np.asarray(sampler.ddim_alphas_prev), "sigmas": np.asarray(sampler.ddim_sigmas),
              "sqrt_one_minus_alphas": np.asarray(sampler.ddim_sqrt_one_minus_alphas)}
    x = P.randn(f"{name}.x", (b, 4, 8, 8), seed)
    c = P.randn(f"{name}.c", (b, 77, 768), seed)
    uc = P.randn(f"{name}.uc", (b, 77, 768), seed)
    with torch.no_grad():
        for index in (49, 17, 0):
            t = torch.full((b,), int(sampler.ddim_timesteps[index]), dtype=torch.long)
            x_prev, pred_x0 = sampler.p_sample_ddim(x, c, t, index=index, guidance_scale=7.5,
                                                   unconditional_conditioning=uc)
            arrays[f"x_prev_{index}"] = x_prev
            arrays[f"pred_x0_{index}"] = pred_x0
        for tag, gs in (("10_4", [10.0, 4.0]), ("10_10", [10.0, 10.0])):
            model.calls.clear()
            samples, _ = sampler.sample(S=50, batch_size=b, shape=(4, 8, 8), conditioning=c, verbose=False,
                                        guidance_scale=gs, unconditional_conditioning=uc, eta=0.0, x_T=x)
            arrays[f"samples_{tag}"] = samples
            arrays[f"t_seq_{tag}"] = np.array([int(tt[0]) for _, tt in model.calls])
        # eta > 0 consumes RNG (ddim.py:286); pin with the global generator
        torch.manual_seed(1234)
        samples, _ = sampler.sample(S=50, batch_size=b, shape=(4, 8, 8), conditioning=c, verbose=False,
                                    guidance_scale=[10.0, 4.0], unconditional_conditioning=uc, eta=0.5, x_T=x)
        arrays["samples_eta05_seed1234"] = samples
        arrays["sigmas_eta05"] = np.asarray(sampler.ddim_sigmas)
    _save(name, meta=np.array([seed, b]), **arrays)


def build_ref_unet(ref, seed):
    t0 = time.time()
    with torch.device("meta"):
        unet = ref.openaimodel.UNetModel(**RH.SD15_UNET_CFG)
    ref_keys = {k: tuple(v.shape) for k, v in unet.state_dict().items()}
    spec = dict(P.unet_spec())
    assert ref_keys == spec, "oracle/params.py:unet_spec does not match the reference UNetModel state_dict"
    sd = P.init_unet(seed)
    unet.load_state_dict(sd, assign=True, strict=True)
    print(f"reference U-Net built + loaded in {time.time() - t0:.1f}s "
          f"({sum(v.numel() for v in sd.values()) / 1e6:.1f} M params)")
    return unet.eval(), sd


def unet_inputs(name, n_samples, hw, seed):
    x_T = P.randn(f"{name}.x_T", (n_samples, 4, hw, hw), seed)
    c_cond = P.randn(f"{name}.c_cond", (n_samples * 16, 77, 768), seed)
    c_uncond = P.randn(f"{name}.c_uncond", (16, 77, 768), seed).repeat(n_samples, 1, 1)
    return x_T, c_cond, c_uncond


def case_unet_step(ref, name="unet_step", seed=17, hw=64):
    """One full SD-1.5 U-Net call (B=2: one cond + one uncond) with per-SpatialTransformer outputs."""
    unet, _ = build_ref_unet(ref, seed)
    x_T, c_cond, c_uncond = unet_inputs(name, 1, hw, seed)
    x_in = torch.cat([x_T, x_T])
    t_in = torch.full((2,), 981, dtype=torch.long)
    ctx = torch.cat([c_cond, c_uncond])
    arrays = {}
    hooks = []
    idx = [0]

    def hook(mod, inp, out):
        arrays[f"st_out_norm_{idx[0]}"] = out.norm()
        arrays[f"st_in_norm_{idx[0]}"] = inp[0].norm()
        arrays[f"st_out_sub_{idx[0]}"] = out[:, ::8, ::4, ::4]
        idx[0] += 1

    for m in unet.modules():
        if isinstance(m, ref.attention.SpatialTransformer):
            hooks.append(m.register_forward_hook(hook))
    t0 = time.time()
    with torch.no_grad():
        eps = unet(x_in, t_in, context=ctx, context_in=None, extra_info=RH.default_extra_info())
    print(f"reference U-Net forward B=2 {hw}x{hw}: {time.time() - t0:.1f}s")
    assert idx[0] == 16
    _save(name, eps=eps, meta=np.array([seed, hw]), **arrays)


def case_ddim50(ref, name="ddim50", seed=18, hw=64, steps=50):
    """`steps`-step DDIM, n_samples=1 + CFG (10,4), eta=0, through the reference DDIMSampler and U-Net."""
    unet, _ = build_ref_unet(ref, seed)
    model = ref.ModelStandIn(unet=unet)
    sampler = ref.CPUDDIMSampler(model)
    x_T, c_cond, c_uncond = unet_inputs(name, 1, hw, seed)
    # p_sample_ddim sums the two prompt lists (ddim.py:245)
    cond = (c_cond, ["a z"], RH.default_extra_info())
    uncond = (c_uncond, [""], RH.default_extra_info())
    trace = {}

    def img_callback(pred_x0, i):
        if i in (0, 4, 9, 24, 49):
            trace[f"pred_x0_step{i}"] = pred_x0.clone()

    t0 = time.time()
    with torch.no_grad():
        samples, inter = sampler.sample(S=steps, batch_size=1, shape=(4, hw, hw), conditioning=cond, verbose=False,
                                        guidance_scale=[10.0, 4.0], unconditional_conditioning=uncond, eta=0.0,
                                        x_T=x_T, img_callback=img_callback)
    print(f"reference {steps}-step DDIM (B=2, {hw}x{hw}): {time.time() - t0:.1f}s")
    _save(name, samples=samples, meta=np.array([seed, hw, steps]), **trace)


def case_unet_step_b16(ref, name="unet_step_b16", seed=17, hw=64, n_samples=8):
    """The U-Net call of BASELINE.json configs[1] at its full batch: n_samples = 8 + the CFG pair = 16 samples at 64x64
    through the UNMODIFIED reference U-Net (~9 GB of fp32 scores per 64x64 self-attention layer).  eps is stored
    sub-sampled (every other pixel) plus in full for the first cond / uncond pair; per-SpatialTransformer outputs
    sub-sampled like `unet_step`.  Same weights as `unet_step` (seed 17)."""
    unet, _ = build_ref_unet(ref, seed)
    x_T, c_cond, c_uncond = unet_inputs(name, n_samples, hw, seed)
    x_in = torch.cat([x_T, x_T])
    t_in = torch.full((2 * n_samples,), 981, dtype=torch.long)
    ctx = torch.cat([c_cond, c_uncond])
    arrays, idx = {}, [0]

    def hook(mod, inp, out):
        arrays[f"st_out_sub_{idx[0]}"] = out[::4, ::8, ::4, ::4]       # samples 0, 4 (cond) and 8, 12 (uncond)
        idx[0] += 1

    for m in unet.modules():
        if isinstance(m, ref.attention.SpatialTransformer):
            m.register_forward_hook(hook)
    t0 = time.time()
    with torch.no_grad():
        eps = unet(x_in, t_in, context=ctx, context_in=None, extra_info=RH.default_extra_info())
    print(f"reference U-Net forward B={2 * n_samples} {hw}x{hw}: {time.time() - t0:.1f}s")
    assert idx[0] == 16
    _save(name, eps_sub=eps[:, :, ::2, ::2], eps_pair0=eps[[0, n_samples]], meta=np.array([seed, hw, n_samples]), **arrays)


def case_ddim50_ada(name, n_subjects, seed=18, hw=64, steps=50):
    """`steps`-step DDIM with the AdaEmbedding call site composed into every conditioned layer (one subject: the
    benchmarked workload of bench.py at n_samples = 1; two subjects: BASELINE.json configs[2] 'a z hugging a y').
    The reference has no live call site for AdaEmbedding, so this fixture comes from oracle/port.py -
    `unet_forward(ada_subjects=...)`, whose parts are pinned to the reference by `unet_step`, `ddim50` and `ada` - and not
    from the reference modules.  n_samples = 1 + CFG (10, 4), eta = 0; U-Net weights as `ddim50` (seed 18)."""
    sd = P.init_unet(seed)
    x_T, c_cond, c_uncond = unet_inputs(name, 1, hw, seed)
    subjects = P.ada_subject_inputs(name, n_subjects, 1, seed)
    trace = {}
    step = [0]

    def apply_model(x_in, t_in, ctx):
        out = port.unet_forward(sd, x_in, t_in, ctx, ada_subjects=subjects)
        step[0] += 1
        return out

    t0 = time.time()
    with torch.no_grad():
        samples, tr = port.ddim_sample(apply_model, x_T, c_cond, c_uncond, steps, (10.0, 4.0), return_trace=True)
    for i in (0, 9, 24, steps - 1):
        if i < steps:
            trace[f"x_step{i}"] = tr[i]
    print(f"port {steps}-step DDIM with {n_subjects} AdaEmbedding subject(s) (B=2, {hw}x{hw}): {time.time() - t0:.1f}s")
    _save(name, samples=samples, meta=np.array([seed, hw, steps, n_subjects]), **trace)


FAST = ["st_small", "st_c640", "st_c1280", "st_mask_vk", "st_config1", "attn_vars", "conv_attn", "ada", "ddim"]
SLOW = ["unet_step", "ddim10_small", "ddim50"]


RESBLOCK_CH_STRIDE = 4
RESBLOCK_CASES = (("same", 320, 320, 16), ("skip", 320, 640, 16), ("cat", 640, 320, 32), ("cat_small", 2560, 1280, 8))


def resblock_inputs(tag, cin, hw, seed, b=2):
    return P.randn(f"rb.{tag}.x", (b, cin, hw, hw), seed), P.randn(f"rb.{tag}.emb", (b, 1280), seed)


def case_resblock(ref, name="resblock", seed=21):
    """The reference's ResBlock (openaimodel.py:167-281) and Upsample (:95-123) - SURVEY.md 8f N1 - on seeded inputs."""
    out = {}
    for tag, cin, cout, hw in RESBLOCK_CASES:
        rb = ref.openaimodel.ResBlock(cin, 1280, 0.0, out_channels=cout, dims=2, use_checkpoint=False, use_scale_shift_norm=False)
        rb.load_state_dict(P.init_resblock(cin, cout, seed), strict=True)
        x, emb = resblock_inputs(tag, cin, hw, seed)
        with torch.no_grad():
            out[tag] = rb.eval()(x, emb)[:, ::RESBLOCK_CH_STRIDE]      # every 4th channel keeps the fixture small
    up = ref.openaimodel.Upsample(640, True, dims=2)
    up.load_state_dict(P.init_upsample(640, seed), strict=True)
    with torch.no_grad():
        out["up"] = up.eval()(P.randn("rb.up.x", (2, 640, 16, 16), seed))[:, ::RESBLOCK_CH_STRIDE]
    _save(name, **out)


STATIC_TOKEN = 49408        # first id past the CLIP vocabulary: what add_tokens gives the placeholder


def static_inputs(name, seed, b=3, n=77, K=9):
    """Token embeddings and token ids of b prompts: prompt 0 holds the placeholder (followed by its K-1 filler
    slots) at position 5 and once more at 40 (only the first counts, embedding_manager.py:1363-1366), prompt 1 has
    none, prompt 2 holds it at position 20."""
    emb = P.randn(f"{name}.embedded_text", (b, n, 768), seed, 0.05)
    tok = (P.randn(f"{name}.tokens", (b, n), seed).abs() * 1000).long() % 49000 + 1
    tok[0, 5] = STATIC_TOKEN
    tok[0, 40] = STATIC_TOKEN
    tok[2, 20] = STATIC_TOKEN
    return emb, tok


def case_static(ref, name="static_emb", seed=22, K=9, r=12, L=16):
    """SURVEY 8f N2.  StaticLayerwiseEmbedding.forward run on the unmodified reference class
    (embedding_manager.py:497-537), and the splice of its rows into the layer-repeated prompt by the reference's own
    statements of EmbeddingManager.get_static_embedding (the method itself needs a CLIP tokenizer/embedder, absent
    here): repeat :1350, where :1360, extract_first_index_in_each_instance :1366 (the reference's function,
    ldm/util.py:2114), per-k assignment :1510-1552 with REAL_OCCURS_IN_BATCH repeats :1546."""
    sd = P.init_static_embedding(seed, K, r, 2)
    em = ref.embedding_manager
    mod = em.StaticLayerwiseEmbedding(num_layers=L, num_vectors_per_subj_token=K, out_emb_dim=768, r=r,
                                      init_vecs=torch.zeros(2, 768), has_bias=True, device_type="cpu")
    mod.load_state_dict(sd, strict=True)
    with torch.no_grad():
        static = mod.eval()()
    emb, tok = static_inputs(name, seed, K=K)
    B, N = tok.shape
    embedded_text = emb.unsqueeze(1).repeat(1, L, 1, 1).view(B * L, N, -1).clone()
    tokenized_text = tok.unsqueeze(1).repeat(1, L, 1).view(B * L, N)
    placeholder_indices = torch.where(tokenized_text == STATIC_TOKEN)
    first = ref.util.extract_first_index_in_each_instance(placeholder_indices)
    real_occurs = first[0].numel() // L
    for k in range(K):
        embedded_text[(first[0], first[1] + k)] = static[:, k].repeat(real_occurs, 1)
    _save(name, meta=np.array([seed, K, r, L, STATIC_TOKEN]), static=static, spliced=embedded_text[:, :, ::16])


CKPT_FIXTURE = "embeddings_gs-toy.pt"


def case_embeddings_ckpt(ref, name="embeddings_ckpt", seed=23, L=16):
    """SURVEY 8f N3.  A checkpoint in the reference's own on-disk format: the dict of EmbeddingManager.save
    (embedding_manager.py:1820-1834) with LIVE reference StaticLayerwiseEmbedding modules (a subject "z", K=4, r=6, two
    init vectors; a background "y", K=2, r=4, none), written by torch.save, i.e. a pickle naming
    ldm.modules.embedding_manager.StaticLayerwiseEmbedding.  EmbeddingManager itself cannot be built here (it wants a
    CLIP embedder), so the dict is assembled by hand with the reference's container types (:986-989, :1215-1216).
    The .npz holds what the reference modules compute, for the loader's parity test."""
    from collections import OrderedDict
    em = ref.embedding_manager
    specs = {"z": (4, 6, 2), "y": (2, 4, 0)}
    embedders = torch.nn.ParameterDict()
    outs = {}
    for i, (km, (K, r, n_init)) in enumerate(specs.items()):
        mod = em.StaticLayerwiseEmbedding(num_layers=L, num_vectors_per_subj_token=K, out_emb_dim=768, r=r,
                                          init_vecs=torch.zeros(n_init, 768) if n_init else None, has_bias=True,
                                          token_string=km, device_type="cpu")
        sd = {k: P.draw(f"ckpt.{km}.{k}", tuple(v.shape), seed + i, "normal", 0.2) for k, v in mod.state_dict().items()}
        mod.load_state_dict(sd, strict=True)
        embedders[km] = mod
        with torch.no_grad():
            outs[f"static_{km}"] = mod.eval()()
    ckpt = {"string_to_token": OrderedDict(z=torch.tensor(49408), y=torch.tensor(49409)),
            "string_to_static_embedder": embedders,
            "string_to_subj_basis_generator_dict": torch.nn.ModuleDict(),
            "token2num_vectors": {"z": 4, "y": 2},
            "emb_global_scale_scores": torch.nn.Parameter(torch.tensor([0.3, -0.2])),
            "use_conv_attn_kernel_size": 3,
            "placeholder_strings": ["z", "y"], "subject_strings": ["z"], "background_strings": ["y"],
            "ca_q_bns": torch.nn.ModuleDict({"1": torch.nn.BatchNorm2d(320, affine=False)}),
            "ca_outfeat_lns": torch.nn.ModuleDict({"1": torch.nn.LayerNorm(320, elementwise_affine=False)}),
            "do_zero_shot": False}
    path = os.path.join(GOLDEN_DIR, CKPT_FIXTURE)
    torch.save(ckpt, path)
    print(f"wrote {path} ({os.path.getsize(path) // 1024} KiB)")
    _save(name, meta=np.array([seed, L]), **outs)


def main(argv):
    ref = RH.load()
    torch.set_num_threads(os.cpu_count() or 1)
    todo = argv or FAST
    for name in todo:
        t0 = time.time()
        if name == "st_small":
            case_st(ref, "st_small", 320, 16)
        elif name == "st_c640":
            case_st(ref, "st_c640", 640, 8)
        elif name == "st_c1280":
            case_st(ref, "st_c1280", 1280, 8)
        elif name == "st_config1":
            case_st(ref, "st_config1", 320, 64, sub=4)          # BASELINE.json configs[0]
        elif name == "st_mask_vk":
            case_st_mask_vk(ref)
        elif name == "attn_vars":
            case_attn_vars(ref)
        elif name == "conv_attn":
            case_conv_attn(ref)
        elif name == "conv_attn_ks4":
            case_conv_attn_ks4(ref)
        elif name == "ada":
            case_ada(ref)
        elif name == "ddim":
            case_ddim(ref)
        elif name == "unet_step":
            case_unet_step(ref)
        elif name == "ddim50":
            case_ddim50(ref)
        elif name == "resblock":
            case_resblock(ref)
        elif name == "static_emb":
            case_static(ref)
        elif name == "embeddings_ckpt":
            case_embeddings_ckpt(ref)
        elif name == "unet_step_b16":
            case_unet_step_b16(ref)
        elif name == "ddim50_ada1":
            case_ddim50_ada("ddim50_ada1", 1)
        elif name == "ddim50_ada2":
            case_ddim50_ada("ddim50_ada2", 2)
        elif name == "ddim10_ada2_small":
            case_ddim50_ada("ddim10_ada2_small", 2, seed=19, hw=16, steps=10)
        elif name == "ddim10_small":
            case_ddim50(ref, "ddim10_small", seed=19, hw=16, steps=10)
        else:
            raise SystemExit(f"unknown case {name}")
        print(f"[{name}] {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main(sys.argv[1:])
