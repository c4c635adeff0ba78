"""Deterministic random-init parameters and inputs for the oracle, the goldens and the GPU parity tests.

TEST INFRASTRUCTURE ONLY (see oracle/port.py header).

The reference constructors cannot run on the GPU box (the reference is not there), so weights are
drawn here, tensor by tensor, from a CPU ``torch.Generator`` seeded by (seed, crc32(key)) and loaded
into BOTH the imported reference modules (oracle/gen_golden.py, build container only) and the
product modules / the port.  CPU ``randn`` with a fixed seed is reproducible across machines for a
given torch version, so fixture inputs never need to be stored - only reference outputs are.

Zero-initialised modules of the reference (SpatialTransformer.proj_out attention.py:313-317,
UNetModel.out[2] and ResBlock.out_layers[3] openaimodel.py:233-235,693-697) are randomised like the
rest, otherwise a random-init block is the identity.
"""
from __future__ import annotations

import zlib
from typing import Dict, Iterable, Tuple

import torch

from .port import CA_INFEAT_DIMS


def _gen(seed: int, key: str) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((seed * 1000003 + zlib.crc32(key.encode())) & 0x7FFFFFFF)
    return g


def draw(key: str, shape: Tuple[int, ...], seed: int, kind: str = "auto", gain: float = 1.0) -> torch.Tensor:
    """One tensor.  Matrices/conv kernels ~ N(0, gain/sqrt(fan_in)); norm scales ~ 1 + 0.1 N; biases ~ 0.05 N."""
    g = _gen(seed, key)
    shape = tuple(shape)
    if kind == "auto":
        if len(shape) >= 2:
            kind = "matrix"
        elif key.endswith("weight"):
            kind = "scale"
        else:
            kind = "bias"
    if kind == "matrix":
        fan_in = 1
        for s in shape[1:]:
            fan_in *= s
        return torch.randn(shape, generator=g) * (gain / fan_in ** 0.5)
    if kind == "scale":
        return 1.0 + 0.1 * torch.randn(shape, generator=g)
    if kind == "bias":
        return 0.05 * torch.randn(shape, generator=g)
    if kind == "normal":
        return torch.randn(shape, generator=g) * gain
    raise ValueError(kind)


def spatial_transformer_spec(c: int, context_dim: int = 768, prefix: str = "") -> Iterable[Tuple[str, Tuple[int, ...]]]:
    """state_dict keys/shapes of SpatialTransformer(c, 8, c//8, depth=1, context_dim): attention.py:295-319."""
    t = prefix + "transformer_blocks.0."
    yield prefix + "norm.weight", (c,)
    yield prefix + "norm.bias", (c,)
    yield prefix + "proj_in.weight", (c, c, 1, 1)
    yield prefix + "proj_in.bias", (c,)
    for a, cd in (("attn1.", c), ("attn2.", context_dim)):
        yield t + a + "to_q.weight", (c, c)
        yield t + a + "to_k.weight", (c, cd)
        yield t + a + "to_v.weight", (c, cd)
        yield t + a + "to_out.0.weight", (c, c)
        yield t + a + "to_out.0.bias", (c,)
    yield t + "ff.net.0.proj.weight", (8 * c, c)
    yield t + "ff.net.0.proj.bias", (8 * c,)
    yield t + "ff.net.2.weight", (c, 4 * c)
    yield t + "ff.net.2.bias", (c,)
    for n in ("norm1", "norm2", "norm3"):
        yield t + n + ".weight", (c,)
        yield t + n + ".bias", (c,)
    yield prefix + "proj_out.weight", (c, c, 1, 1)
    yield prefix + "proj_out.bias", (c,)


def init_spatial_transformer(c: int, seed: int, context_dim: int = 768, prefix: str = "") -> Dict[str, torch.Tensor]:
    return {k: draw(k, s, seed) for k, s in spatial_transformer_spec(c, context_dim, prefix)}


def unet_spec(model_channels: int = 320, channel_mult=(1, 2, 4, 4), num_res_blocks: int = 2,
              attention_resolutions=(4, 2, 1), context_dim: int = 768, in_ch: int = 4, out_ch: int = 4):
    """state_dict keys/shapes of the SD-1.5 UNetModel (openaimodel.py:447-697, v1-inference-ada.yaml:35-51)."""
    mc = model_channels
    ted = 4 * mc
    yield "time_embed.0.weight", (ted, mc)
    yield "time_embed.0.bias", (ted,)
    yield "time_embed.2.weight", (ted, ted)
    yield "time_embed.2.bias", (ted,)

    def res(prefix, cin, cout):
        yield prefix + "in_layers.0.weight", (cin,)
        yield prefix + "in_layers.0.bias", (cin,)
        yield prefix + "in_layers.2.weight", (cout, cin, 3, 3)
        yield prefix + "in_layers.2.bias", (cout,)
        yield prefix + "emb_layers.1.weight", (cout, ted)
        yield prefix + "emb_layers.1.bias", (cout,)
        yield prefix + "out_layers.0.weight", (cout,)
        yield prefix + "out_layers.0.bias", (cout,)
        yield prefix + "out_layers.3.weight", (cout, cout, 3, 3)
        yield prefix + "out_layers.3.bias", (cout,)
        if cin != cout:
            yield prefix + "skip_connection.weight", (cout, cin, 1, 1)
            yield prefix + "skip_connection.bias", (cout,)

    yield "input_blocks.0.0.weight", (mc, in_ch, 3, 3)
    yield "input_blocks.0.0.bias", (mc,)
    chans = [mc]
    ch, ds, blk = mc, 1, 1
    for level, mult in enumerate(channel_mult):
        for _ in range(num_res_blocks):
            yield from res(f"input_blocks.{blk}.0.", ch, mult * mc)
            ch = mult * mc
            if ds in attention_resolutions:
                yield from spatial_transformer_spec(ch, context_dim, f"input_blocks.{blk}.1.")
            chans.append(ch)
            blk += 1
        if level != len(channel_mult) - 1:
            yield f"input_blocks.{blk}.0.op.weight", (ch, ch, 3, 3)
            yield f"input_blocks.{blk}.0.op.bias", (ch,)
            chans.append(ch)
            blk += 1
            ds *= 2
    yield from res("middle_block.0.", ch, ch)
    yield from spatial_transformer_spec(ch, context_dim, "middle_block.1.")
    yield from res("middle_block.2.", ch, ch)
    blk = 0
    for level, mult in list(enumerate(channel_mult))[::-1]:
        for i in range(num_res_blocks + 1):
            ich = chans.pop()
            yield from res(f"output_blocks.{blk}.0.", ch + ich, mult * mc)
            ch = mult * mc
            j = 1
            if ds in attention_resolutions:
                yield from spatial_transformer_spec(ch, context_dim, f"output_blocks.{blk}.1.")
                j = 2
            if level and i == num_res_blocks:
                yield f"output_blocks.{blk}.{j}.conv.weight", (ch, ch, 3, 3)
                yield f"output_blocks.{blk}.{j}.conv.bias", (ch,)
                ds //= 2
            blk += 1
    yield "out.0.weight", (ch,)
    yield "out.0.bias", (ch,)
    yield "out.2.weight", (out_ch, ch, 3, 3)
    yield "out.2.bias", (out_ch,)


def resblock_spec(cin: int, cout: int, ted: int = 1280, prefix: str = ""):
    """state_dict keys/shapes of one ResBlock(cin, ted, dropout, out_channels=cout): openaimodel.py:183-257."""
    yield prefix + "in_layers.0.weight", (cin,)
    yield prefix + "in_layers.0.bias", (cin,)
    yield prefix + "in_layers.2.weight", (cout, cin, 3, 3)
    yield prefix + "in_layers.2.bias", (cout,)
    yield prefix + "emb_layers.1.weight", (cout, ted)
    yield prefix + "emb_layers.1.bias", (cout,)
    yield prefix + "out_layers.0.weight", (cout,)
    yield prefix + "out_layers.0.bias", (cout,)
    yield prefix + "out_layers.3.weight", (cout, cout, 3, 3)
    yield prefix + "out_layers.3.bias", (cout,)
    if cin != cout:
        yield prefix + "skip_connection.weight", (cout, cin, 1, 1)
        yield prefix + "skip_connection.bias", (cout,)


def init_resblock(cin: int, cout: int, seed: int, ted: int = 1280) -> Dict[str, torch.Tensor]:
    return {k: draw(f"rb{cin}x{cout}." + k, s, seed) for k, s in resblock_spec(cin, cout, ted)}


def init_upsample(ch: int, seed: int) -> Dict[str, torch.Tensor]:
    """Upsample(ch, use_conv=True): openaimodel.py:95-123."""
    return {"conv.weight": draw(f"up{ch}.conv.weight", (ch, ch, 3, 3), seed), "conv.bias": draw(f"up{ch}.conv.bias", (ch,), seed)}


def init_unet(seed: int, **cfg) -> Dict[str, torch.Tensor]:
    return {k: draw(k, s, seed) for k, s in unet_spec(**cfg)}


def ada_embedding_spec(K: int = 9, r: int = 12, n_init: int = 2, out_dim: int = 768):
    """state_dict keys/shapes of AdaEmbedding(num_vectors_per_subj_token=K, fg/bg = 5/4, r, init_vecs [n_init,768])
    in its H=2 configuration: embedding_manager.py:546-716."""
    yield "pre_vecs", (K, n_init, out_dim)
    yield "basis_vecs", (K, r - n_init, out_dim)
    for i, c in enumerate(CA_INFEAT_DIMS):
        lg = c // 8
        yield f"poolers.{i}.lora_to_k.weight", (lg, lg, 1)
        yield f"poolers.{i}.lora_to_fg_q.weight", (lg, lg, 1)
        yield f"poolers.{i}.lora_to_bg_q.weight", (lg, lg, 1)
    for i, c in enumerate(CA_INFEAT_DIMS):
        td = int(0.25 * c)
        yield f"layer_coeff_maps.{i}.weight", (r * K, 2 * c + td)
        yield f"layer_coeff_maps.{i}.bias", (r * K,)
    yield "bias", (len(CA_INFEAT_DIMS), K, out_dim)


def init_ada_embedding(seed: int, K: int = 9, r: int = 12, n_init: int = 2) -> Dict[str, torch.Tensor]:
    out = {}
    for k, s in ada_embedding_spec(K, r, n_init):
        if k in ("pre_vecs", "basis_vecs"):
            out[k] = draw(k, s, seed, "normal", 0.25)
        elif k == "bias":
            out[k] = draw(k, s, seed, "normal", 0.02)
        else:
            out[k] = draw(k, s, seed)
    return out


def static_embedding_spec(K: int = 9, r: int = 12, n_init: int = 2, num_layers: int = 16, out_dim: int = 768):
    """state_dict keys/shapes of StaticLayerwiseEmbedding(num_layers, K, out_dim, r, init_vecs [n_init, 768],
    has_bias=True): embedding_manager.py:374-490."""
    if n_init > 0:
        yield "pre_vecs", (K, n_init, out_dim)
    yield "basis_rand_weights", (num_layers, K, r)
    yield "basis_vecs", (K, r - n_init, out_dim)
    yield "basis_comm_weights", (1, K, r)
    yield "bias", (num_layers, K, out_dim)


def init_static_embedding(seed: int, K: int = 9, r: int = 12, n_init: int = 2) -> Dict[str, torch.Tensor]:
    out = {}
    for k, s in static_embedding_spec(K, r, n_init):
        if k in ("pre_vecs", "basis_vecs"):
            out[k] = draw("static." + k, s, seed, "normal", 0.25)
        elif k == "bias":
            out[k] = draw("static." + k, s, seed, "normal", 0.02)
        elif k == "basis_comm_weights":
            out[k] = draw("static." + k, s, seed, "normal", 0.2) + 1.0 / r
        else:
            out[k] = draw("static." + k, s, seed, "normal", 0.1)
    return out


def randn(key: str, shape, seed: int, scale: float = 1.0) -> torch.Tensor:
    """A named synthetic input tensor."""
    return draw(key, tuple(shape), seed, "normal", scale)


def ada_subject_inputs(name: str, n_subjects: int, n_instances: int, seed: int, K: int = 9, r: int = 12):
    """Seeded per-subject inputs of the AdaEmbedding wiring (BASELINE.json configs[1] with one subject,
    configs[2] 'a z hugging a y' with two): an AdaEmbedding state_dict (K = 9 vectors, fg 5 / bg 4, rank r), the K token
    positions the subject occupies in the 77-token prompt (2..10, then 13..21), the per-layer fg probe [16, 768] and the
    per-layer bg query [16, n_instances, 768].  Shared by oracle/gen_golden.py, the GPU tests and bench.py's checker."""
    out = []
    for si in range(n_subjects):
        first = 2 + 11 * si
        out.append({"params": init_ada_embedding(seed + 10 * si, K, r, 2), "positions": list(range(first, first + K)),
                    "probes": randn(f"{name}.{si}.probes", (16, 768), seed, 0.5),
                    "extra_means": randn(f"{name}.{si}.extra", (16, n_instances, 768), seed, 0.5),
                    "K": K, "r": r, "n_instances": n_instances})
    return out
