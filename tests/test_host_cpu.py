"""CPU-side checks of the host layer: the C-ABI library loads and exports every symbol the header declares,
module state-dict keys equal the reference's, the DDIM schedule equals the reference's golden values, the
product path fails loudly without CUDA, and the sample sharding works at world_size 2 over gloo."""
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden
from oracle import params as P
from oracle import port


# ------------------------------------------------------------------------------------------- C ABI
def _header_functions():
    text = open(os.path.join(ROOT, "include", "adaface_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(afb_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_loads_and_exports_every_declared_symbol():
    import ctypes
    from adaface_b200 import _lib, build
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    declared = _header_functions()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/adaface_b200.h but not exported"
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared, "ctypes signature table and header disagree"
    lib.afb_abi_version.restype = ctypes.c_int
    assert lib.afb_abi_version() >= 1


def test_sass_contains_blackwell_tensor_and_tma_instructions():
    """The GEMM / attention kernels must be tcgen05 + TMA code (UTC*MMA, UTMALDG in SASS), not mma.sync."""
    import shutil
    import subprocess
    from adaface_b200 import build
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not available")
    sass = subprocess.run(["cuobjdump", "-sass", build.build()], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass or "UTCMMA" in sass
    assert "UTMALDG" in sass
    assert "LDTM" in sass
    # warp-level mma.sync is allowed off the default forward path only: the self-attention backward of the inversion step
    # (csrc/backward.cu attn_bwd_mma3_kernel / attn_bwd_mma2_kernel / attn_bwd_short_mma_kernel), the U-Net's last convolution, 320 -> 4 channels, whose 4 output columns cannot
    # feed a 128-row tcgen05 tile (csrc/conv_small.cu), and the explicit (materialised-score) cross-attention of the
    # save_attn_vars / conv-attention modes (csrc/attn_explicit.cu); every GEMM / attention kernel of the default path
    # must be tcgen05
    legacy = set()
    for block in sass.split("Function : ")[1:]:
        name, _, body = block.partition("\n")
        if "HMMA.16816" in body:
            legacy.add(name.strip())
    # (+ the AdaEmbedding pooling kernel: [tokens x C] x [C x 16] and [C x tokens] x [tokens x 8] contractions, far too
    #  skinny for a 128 x N tcgen05 tile)
    allowed = ("attn_bwd_mma2_kernel", "attn_bwd_mma3_kernel", "attn_bwd_short_mma_kernel", "conv3x3_few_out_kernel", "cross_scores_kernel", "softmax_pv_kernel", "ada_pool_mma_kernel")
    assert legacy and all(any(a in n for a in allowed) for n in legacy), legacy


# ------------------------------------------------------------------------------------------- state dicts
def test_spatial_transformer_state_dict_matches_reference_keys():
    from adaface_b200 import SpatialTransformer
    for c in (320, 640, 1280):
        st = SpatialTransformer(c, 8, c // 8, depth=1, context_dim=768)
        have = {k: tuple(v.shape) for k, v in st.state_dict().items()}
        assert have == dict(P.spatial_transformer_spec(c))
        # the reference zero-initialises proj_out (attention.py:313-317)
        assert float(st.proj_out.weight.abs().max()) == 0.0 and float(st.proj_out.bias.abs().max()) == 0.0
        attn2 = st.transformer_blocks[0].attn2
        for attr in ("heads", "scale", "save_attn_vars", "cached_activations", "use_conv_attn_kernel_size",
                     "infeat_size", "is_training"):
            assert hasattr(attn2, attr)
        assert attn2.scale == (c // 8) ** -0.5 and attn2.heads == 8


def test_unet_state_dict_matches_reference_keys():
    from adaface_b200.unet import SD15_UNET_CONFIG, UNetModel
    with torch.device("meta"):
        unet = UNetModel(**SD15_UNET_CONFIG)
    have = {k: tuple(v.shape) for k, v in unet.state_dict().items()}
    want = dict(P.unet_spec())
    assert set(have) == set(want)
    assert have == want
    layers = [i for i, _ in unet.spatial_transformers()]
    assert layers == sorted(port.LAYER_IDX2CA)                 # openaimodel.py:876-877
    assert sum(v.numel() for v in unet.state_dict().values()) == 859_520_964


def test_ada_embedding_state_dict_and_cross_weight_reduction():
    from adaface_b200 import AdaEmbedding
    torch.manual_seed(0)
    ada = AdaEmbedding(num_vectors_per_subj_token=9, fg_emb_count=5, bg_emb_count=4, r=12,
                       init_vecs=torch.zeros(2, 768), device_type="cpu")
    have = {k: tuple(v.shape) for k, v in ada.state_dict().items()}
    assert have == dict(P.ada_embedding_spec(9, 12, 2))
    assert ada.H == 2 and ada.TDs == [int(0.25 * c) for c in P.CA_INFEAT_DIMS]
    # embedding_manager.py:719-773: cross-stream blocks are at most 1 % of the in-stream mean magnitude
    w = ada.layer_coeff_maps[0].weight.view(9, 12, -1)
    d = 320
    assert float(w[0][:, d:2 * d].abs().mean()) <= 0.0101 * float(w[0][:, :d].abs().mean())
    assert float(w[8][:, :d].abs().mean()) <= 0.0101 * float(w[8][:, d:2 * d].abs().mean())
    one = AdaEmbedding(num_vectors_per_subj_token=1, fg_emb_count=1, bg_emb_count=0, device_type="cpu")
    with pytest.raises(RuntimeError):                          # H == 1 is not self-consistent
        one(1, {}, None, None, None)


# ------------------------------------------------------------------------------------------- DDIM host logic
class _Model:
    def __init__(self):
        betas = port.make_beta_schedule_linear()
        ac = np.cumprod(1.0 - betas)
        self.num_timesteps = 1000
        self.betas = torch.tensor(betas, dtype=torch.float32)
        self.alphas_cumprod = torch.tensor(ac, dtype=torch.float32)
        self.alphas_cumprod_prev = torch.tensor(np.append(1.0, ac[:-1]), dtype=torch.float32)
        self.device = torch.device("cpu")

    def apply_model(self, x, t, c):
        return torch.tanh(x)


def test_ddim_schedule_matches_reference_golden():
    from adaface_b200 import DDIMSampler
    from adaface_b200.ddim import guidance_schedule
    g = load_golden("ddim")
    s = DDIMSampler(_Model())
    s.make_schedule(50, ddim_eta=0.0)
    assert np.array_equal(s.ddim_timesteps, g["timesteps"])
    assert list(s.ddim_timesteps[:3]) == [1, 21, 41] and s.ddim_timesteps[-1] == 981        # ddim.py:30-36
    for mine, key in ((s.ddim_alphas, "alphas"), (s.ddim_alphas_prev, "alphas_prev"), (s.ddim_sigmas, "sigmas"),
                      (s.ddim_sqrt_one_minus_alphas, "sqrt_one_minus_alphas")):
        np.testing.assert_allclose(mine, g[key], rtol=2e-6, atol=0, err_msg=key)
    s.make_schedule(50, ddim_eta=0.5)
    np.testing.assert_allclose(s.ddim_sigmas, g["sigmas_eta05"], rtol=2e-6)
    sc = guidance_schedule([10.0, 4.0], 50)
    assert sc == port.guidance_schedule((10.0, 4.0), 50)
    assert sc[0] == 10.0 and abs(sc[-1] - 4.0) < 1e-9
    assert guidance_schedule(7.5, 3) == [7.5, 7.5, 7.5]


def test_product_path_fails_loudly_on_cpu():
    from adaface_b200 import CrossAttention, DDIMSampler, SpatialTransformer
    from adaface_b200._lib import AdafaceB200Error
    st = SpatialTransformer(320, 8, 40, depth=1, context_dim=768).eval()
    with pytest.raises(AdafaceB200Error):
        st(torch.zeros(1, 320, 8, 8), context=torch.zeros(1, 77, 768))
    with pytest.raises(AdafaceB200Error):
        CrossAttention(320, 768, 8, 40)(torch.zeros(1, 64, 320), context=torch.zeros(1, 77, 768))
    s = DDIMSampler(_Model())
    s.make_schedule(50)
    x = torch.zeros(1, 4, 8, 8)
    with pytest.raises(AdafaceB200Error):
        s.p_sample_ddim(x, torch.zeros(1, 77, 768), torch.full((1,), 981), index=49, guidance_scale=2.0,
                        unconditional_conditioning=torch.zeros(1, 77, 768))


def test_product_package_does_not_import_the_oracle():
    import subprocess
    code = ("import sys; import adaface_b200, adaface_b200.unet, adaface_b200.sampling; "
            "bad = [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')]; "
            "assert not bad, bad")
    subprocess.run([sys.executable, "-c", code], check=True, cwd=ROOT)
    for fn in os.listdir(os.path.join(ROOT, "adaface_b200")):
        if fn.endswith(".py"):
            src = open(os.path.join(ROOT, "adaface_b200", fn)).read()
            assert "import oracle" not in src and "from oracle" not in src, fn


def test_geglu_interleave_is_a_row_permutation():
    from adaface_b200.ops import interleave_geglu
    w = torch.arange(64 * 3, dtype=torch.float32).reshape(64, 3)
    b = torch.arange(64, dtype=torch.float32)
    wi, bi = interleave_geglu(w, b)
    w = torch.arange(128 * 3, dtype=torch.float32).reshape(128, 3)
    b = torch.arange(128, dtype=torch.float32)
    wi, bi = interleave_geglu(w, b)
    # rows 0..31 are value rows 0..31, rows 32..63 their gate rows (64..95), then value 32..63, gate 96..127
    assert bi.tolist() == list(range(0, 32)) + list(range(64, 96)) + list(range(32, 64)) + list(range(96, 128))
    assert torch.equal(wi[:, 0] / 3, bi)


# ------------------------------------------------------------------------------------------- sharding
def test_shard_bounds_and_context_slicing():
    from adaface_b200.sampling import shard_bounds, shard_layerwise_context
    assert shard_bounds(8, 2) == [(0, 4), (4, 8)]
    assert shard_bounds(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert shard_bounds(1, 4) == [(0, 1), (1, 1), (1, 1), (1, 1)]
    ctx = torch.arange(6 * 16).reshape(6 * 16, 1, 1)
    part = shard_layerwise_context(ctx, 2, 4, 6)
    assert part.shape[0] == 32 and int(part[0]) == 32 and int(part[-1]) == 63


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port_no, n, out_dir):
    import torch.distributed as dist
    from adaface_b200.sampling import sample_sharded
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(42)
    x_T = torch.randn((n, 4, 8, 8), generator=g)                       # same tensor on every rank
    c = torch.randn((n * 16, 77, 8), generator=g)
    uc = torch.randn((16, 77, 8), generator=g).repeat(n, 1, 1)
    seen = []

    def run_local(x, cc, cu):
        seen.append((x.shape[0], cc.shape[0], cu.shape[0]))
        return x * 2.0 + cc.reshape(x.shape[0], 16, -1).mean(dim=(1, 2)).reshape(-1, 1, 1, 1)

    out = sample_sharded(run_local, x_T, c, uc)
    ref = x_T * 2.0 + c.reshape(n, 16, -1).mean(dim=(1, 2)).reshape(-1, 1, 1, 1)
    assert torch.allclose(out, ref), "gathered result differs from the single-process result"
    assert seen[0][1] == 16 * seen[0][0] and seen[0][2] == 16 * seen[0][0]
    torch.save({"n_local": seen[0][0]}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [8, 5])
def test_sample_sharded_world_size_2_gloo(tmp_path, n):
    import torch.multiprocessing as mp
    port_no = _free_port()
    mp.spawn(_worker, args=(2, port_no, n, str(tmp_path)), nprocs=2, join=True)
    counts = [torch.load(os.path.join(tmp_path, f"r{r}.pt"))["n_local"] for r in range(2)]
    assert sum(counts) == n and abs(counts[0] - counts[1]) <= 1


def _grad_worker(rank, world, port_no, out_dir):
    import torch.distributed as dist
    from adaface_b200.sampling import FlatGradAllReduce
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    torch.manual_seed(0)                                                # same parameters on every rank
    params = [torch.nn.Parameter(torch.randn(9, 12, 768)), torch.nn.Parameter(torch.randn(108, 720)),
              torch.nn.Parameter(torch.randn(5)), torch.nn.Parameter(torch.randn(3), requires_grad=False)]
    g = torch.Generator().manual_seed(100 + rank)                       # different gradients per rank
    grads = [torch.randn(p.shape, generator=g) for p in params[:3]]
    params[0].grad, params[1].grad = grads[0].clone(), grads[1].clone()
    if rank == 0:
        params[2].grad = grads[2].clone()                               # unused on rank 1: contributes zeros
    reducer = FlatGradAllReduce(params)
    flat = reducer()
    assert flat.numel() == sum(p.numel() for p in params[:3]) and params[3].grad is None
    first = [p.grad.clone() for p in params[:3]]
    # attached mode: the gradients live inside the flat buffer, a call is only the all-reduce
    params[0].grad, params[1].grad, params[2].grad = grads[0].clone(), grads[1].clone(), None
    if rank == 0:
        params[2].grad = grads[2].clone()
    reducer.attach()
    assert all(reducer._is_view(p, off) for p, off in zip(reducer.params, reducer.offsets))
    reducer()
    assert all(torch.equal(p.grad, f) for p, f in zip(params[:3], first)), "attached mode differs from copy mode"
    torch.save({"grads": first, "own": grads}, os.path.join(out_dir, f"g{rank}.pt"))
    dist.destroy_process_group()


def test_inversion_gradient_allreduce_world_size_2_gloo(tmp_path):
    """SURVEY.md 8e, inversion mode: one flat-buffer all-reduce averages the embedding gradients over the ranks."""
    import torch.multiprocessing as mp
    mp.spawn(_grad_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"g{r}.pt")) for r in range(2))
    for i in range(3):
        want = (r0["own"][i] + (r1["own"][i] if i < 2 else 0)) / 2
        assert torch.allclose(r0["grads"][i], want, atol=1e-6) and torch.equal(r0["grads"][i], r1["grads"][i])


def test_conv3x3_dx_weight_is_the_transposed_flipped_filter():
    """ops.conv3x3's backward runs the forward kernel on `conv3x3_dx_weight(w)`: check on CPU, with torch's own conv2d
    and autograd, that this matrix ([Cin, 9*Cout], taps flipped) is the filter whose convolution is dX; the result is cached
    per weight TENSOR (a new tensor at a reused address must not hit)."""
    import torch.nn.functional as F
    from adaface_b200 import ops
    cout, cin = 64, 128
    w = torch.randn(cout, 9 * cin).to(torch.bfloat16)                       # [Cout, (ky, kx, c)] as _Bf16Conv._w_native packs it
    wt = ops.conv3x3_dx_weight(w)
    assert wt.shape == (cin, 9 * cout) and ops.conv3x3_dx_weight(w) is wt
    W = w.float().view(cout, 3, 3, cin).permute(0, 3, 1, 2)
    x = torch.randn(2, cin, 6, 5, requires_grad=True)
    y = F.conv2d(x, W, padding=1)
    g = torch.randn_like(y)
    y.backward(g)
    dx = F.conv2d(g, wt.float().view(cin, 3, 3, cout).permute(0, 3, 1, 2), padding=1)
    assert torch.allclose(dx, x.grad, atol=1e-3, rtol=1e-4)
    w2 = w.clone()
    assert ops.conv3x3_dx_weight(w2) is not wt


def test_captured_grad_step_fails_loudly_without_cuda_or_attached_gradients():
    """sampling.CapturedGradStep (the inversion step as one CUDA graph) has no CPU counterpart and needs every .grad to be
    a view of the all-reduce's flat buffer: both are errors, not silent fallbacks."""
    from adaface_b200.sampling import CapturedGradStep, FlatGradAllReduce
    p = torch.nn.Parameter(torch.randn(4, 3))
    reducer = FlatGradAllReduce([p])
    with pytest.raises(RuntimeError, match="CUDA"):
        CapturedGradStep(lambda x: (p * x).sum(), (torch.ones(4, 3),), reducer.attach())
    if torch.cuda.is_available():
        q = torch.nn.Parameter(torch.randn(4, 3, device="cuda"))
        with pytest.raises(ValueError, match="attach"):
            CapturedGradStep(lambda x: (q * x).sum(), (torch.ones(4, 3, device="cuda"),), FlatGradAllReduce([q]))


def test_reference_embedding_checkpoint_loads_without_ldm():
    """a checkpoint written by torch.save with LIVE reference modules (pickled class path
    ldm.modules.embedding_manager.StaticLayerwiseEmbedding; fixture made by oracle/gen_golden.py from the imported
    reference) loads where no `ldm` package exists, into this package's StaticLayerwiseEmbedding, with the bookkeeping of
    EmbeddingManager.load (renaming, subject/background strings, token2num_vectors, global scales)."""
    import subprocess
    import sys
    fixture = os.path.join(os.path.dirname(__file__), "golden", "embeddings_gs-toy.pt")
    if not os.path.exists(fixture):
        pytest.skip("checkpoint fixture not generated")
    code = f"""
import sys, torch
from adaface_b200.checkpoint import load_embedding_checkpoint, OpaqueReferenceModule
from adaface_b200.embedding import StaticLayerwiseEmbedding
ck = load_embedding_checkpoint({fixture!r} + ":z-zz")
assert not any(m == "ldm" or m.startswith("ldm.") for m in sys.modules)
assert ck.subject_strings == ["zz"] and ck.background_strings == ["y"] and ck.placeholder_strings == ["zz", "y"]
assert ck.token2num_vectors == {{"zz": 4, "y": 2}} and ck.string_to_token == {{"zz": 49408, "y": 49409}}
assert ck.use_conv_attn_kernel_size == 3 and ck.do_zero_shot is False
assert abs(ck.emb_global_scale_scores["zz"] - 0.3) < 1e-6 and abs(ck.emb_global_scales()["y"] - (torch.sigmoid(torch.tensor(-0.2)).item() + 0.5)) < 1e-6
z, y = ck.string_to_static_embedder["zz"], ck.string_to_static_embedder["y"]
assert type(z) is StaticLayerwiseEmbedding and (z.K, z.r, z.N, z.num_layers) == (4, 6, 2, 16) and (y.K, y.r, y.N) == (2, 4, 0)
assert sorted(z.state_dict()) == ["basis_comm_weights", "basis_rand_weights", "basis_vecs", "bias", "pre_vecs"]
assert sorted(y.state_dict()) == ["basis_comm_weights", "basis_rand_weights", "basis_vecs", "bias"]
assert set(ck.opaque) == {{"string_to_subj_basis_generator_dict", "ca_q_bns", "ca_outfeat_lns"}}
try:
    load_embedding_checkpoint([{fixture!r}, {fixture!r}])
except ValueError as e:
    assert "Duplicate key" in str(e)
else:
    raise AssertionError("duplicate subject placeholder must raise like the reference")
print("ok")
"""
    env = dict(os.environ, PYTHONPATH=os.path.dirname(os.path.dirname(__file__)))
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert res.returncode == 0 and "ok" in res.stdout, res.stderr[-2000:]
    # the arithmetic of the loaded parameters, checked on the CPU with the oracle port against the reference's outputs
    from adaface_b200.checkpoint import load_embedding_checkpoint
    from conftest import load_golden, rel_l2
    from oracle import port
    g = load_golden("embeddings_ckpt")
    ck = load_embedding_checkpoint(fixture)
    for km in ("z", "y"):
        sd = {k: v.detach() for k, v in ck.string_to_static_embedder[km].state_dict().items()}
        assert rel_l2(port.static_layerwise_embedding(sd), g[f"static_{km}"]) < 1e-5


def test_built_library_stays_current_when_the_checkout_moves(tmp_path):
    """The GPU box runs a COPY of the repository under another path with the prebuilt .so inside: the staleness check
    must not depend on where the checkout lives (a rebuild there, by every rank of a torchrun job at once, once produced
    a half-linked library), and concurrent build() callers must serialise on the file lock."""
    import shutil
    import subprocess
    import sys
    from adaface_b200 import build as B
    if not os.path.exists(B.LIB_PATH):
        pytest.skip("library not built")
    assert B.is_current()
    root = tmp_path / "elsewhere"
    shutil.copytree(os.path.join(ROOT, "adaface_b200"), root / "adaface_b200", ignore=shutil.ignore_patterns("__pycache__"))
    shutil.copytree(os.path.join(ROOT, "include"), root / "include")
    code = ("import sys; from adaface_b200 import build as B; assert B.PKG_DIR.startswith(sys.argv[1]), B.PKG_DIR; "
            "assert B.is_current(); print(B.build())")
    procs = [subprocess.Popen([sys.executable, "-c", code, str(root)], cwd=root, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                              text=True, env=dict(os.environ, PYTHONPATH=str(root))) for _ in range(3)]
    for p in procs:
        out, err = p.communicate(timeout=600)
        assert p.returncode == 0, err[-1500:]
        assert out.strip().endswith("libadaface_b200.so")


def test_embedding_checkpoint_reader_refuses_globals_outside_its_allow_list(tmp_path):
    """the reader runs a pickle with weights_only=False; anything that is not a reference class, torch,
    numpy or a plain container must be refused instead of imported."""
    import pickle
    from adaface_b200.checkpoint import load_embedding_checkpoint

    class Evil:
        def __reduce__(self):
            return (os.system, ("echo pwned > /dev/null",))

    path = tmp_path / "embeddings_gs-evil.pt"
    torch.save({"string_to_static_embedder": Evil()}, str(path))
    with pytest.raises(pickle.UnpicklingError, match="allow-list"):
        load_embedding_checkpoint(str(path))


def test_loader_refuses_a_library_built_against_another_header(monkeypatch):
    """a prebuilt .so whose argument structs may differ from include/adaface_b200.h must not be loaded."""
    from adaface_b200 import _lib, build
    _lib.load()                                     # current library: loads
    monkeypatch.setattr(_lib, "_LIB", None)
    monkeypatch.setattr(build, "is_current", lambda: True)
    monkeypatch.setattr(build, "header_digest", lambda: "0" * 64)
    with pytest.raises(_lib.AdafaceB200Error, match="another revision"):
        _lib.load()
