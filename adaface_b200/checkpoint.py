"""Reader of the reference's subject-embedding checkpoints `embeddings_gs-<step>.pt`.

The reference writes them with `EmbeddingManager.save` (ldm/modules/embedding_manager.py:1820-1834): a `torch.save`
pickle of a dict whose values are LIVE `nn.Module`s - `string_to_static_embedder` holds one
`ldm.modules.embedding_manager.StaticLayerwiseEmbedding` per placeholder string - so `torch.load` can only rebuild them
where the `ldm` package is importable.  This reader needs no `ldm`: every `ldm.*` class named in the pickle is
rebuilt as an opaque `nn.Module` shell that just carries the pickled attributes, parameters and sub-modules, and the
static embedders are then converted to this package's `StaticLayerwiseEmbedding` (same constructor arguments and
state_dict keys), which computes on the GPU.  What the inference branch of `EmbeddingManager.load` (:1840-2052) does
with the dict is mirrored: the optional `path:from-to,from-to` placeholder renaming (:1857-1866), subject / background
string bookkeeping (:2007-2016), `token2num_vectors` (:2025-2026), the per-placeholder `emb_global_scale_scores`
(:2003-2004) and `use_conv_attn_kernel_size` (:1874-1875).

Not converted (kept as opaque shells under `EmbeddingCheckpoint.opaque`): the zero-shot subject-basis generators
(`string_to_subj_basis_generator_dict`: a CLIP-based image-to-embedding network, outside SURVEY.md section 8) and the
`ca_q_bns` / `ca_outfeat_lns` normalisers (training-loss statistics only).
"""
from __future__ import annotations

import pickle
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Union

import torch
from torch import nn

from .embedding import StaticLayerwiseEmbedding

REFERENCE_PREFIXES = ("ldm.",)
_SHELLS: Dict[tuple, type] = {}


class OpaqueReferenceModule(nn.Module):
    """Stand-in for a pickled instance of a reference class: carries its attributes / parameters / sub-modules, so
    `state_dict()` works, but has no forward."""
    reference_class = ""

    def forward(self, *args, **kwargs):
        raise NotImplementedError(f"{self.reference_class} was loaded as an opaque shell (no `ldm` package here); "
                                  "only its state is available")


def _shell(module: str, name: str) -> type:
    key = (module, name)
    if key not in _SHELLS:
        _SHELLS[key] = type(name, (OpaqueReferenceModule,), {"reference_class": f"{module}.{name}", "__module__": __name__})
    return _SHELLS[key]


# What an `embeddings_gs-*.pt` legitimately names beside the reference's own classes: torch tensors / parameters /
# storages and their rebuild helpers, torch.nn modules (the pickled embedders are live nn.Modules), plain containers and
# numpy scalars / arrays.  Anything else is refused: a pickle from a third party must not be able to import and call
# arbitrary globals (os.system, builtins.eval, ...) just because this reader runs with weights_only=False.
_ALLOWED_MODULE_PREFIXES = ("torch", "collections", "numpy", "_codecs")
_ALLOWED_BUILTINS = {"set", "frozenset", "list", "dict", "tuple", "int", "float", "bool", "str", "bytes", "bytearray",
                     "complex", "slice", "range", "object"}
_DENIED = {("torch", "load"), ("torch.serialization", "load"), ("torch.hub", "load"), ("numpy", "load")}


class _Unpickler(pickle.Unpickler):
    def find_class(self, module: str, name: str):
        if module.startswith(REFERENCE_PREFIXES):
            return _shell(module, name)
        root = module.split(".", 1)[0]
        allowed = (root in _ALLOWED_MODULE_PREFIXES and (module, name) not in _DENIED) or \
            (module in ("builtins", "__builtin__") and name in _ALLOWED_BUILTINS)
        if not allowed:
            raise pickle.UnpicklingError(f"embedding checkpoint names {module}.{name}, which is not on the allow-list of "
                                         "this reader (torch / torch.nn / collections / numpy / plain containers)")
        return super().find_class(module, name)


class _PickleModule:
    """The `pickle_module` handed to torch.load: the stdlib pickle with the class lookup redirected."""
    __name__ = "adaface_b200_checkpoint_pickle"
    Unpickler = _Unpickler
    load = staticmethod(lambda f, **kw: _Unpickler(f, **kw).load())
    loads = staticmethod(pickle.loads)
    dump, dumps, Pickler = staticmethod(pickle.dump), staticmethod(pickle.dumps), pickle.Pickler
    HIGHEST_PROTOCOL, DEFAULT_PROTOCOL = pickle.HIGHEST_PROTOCOL, pickle.DEFAULT_PROTOCOL
    UnpicklingError, PicklingError = pickle.UnpicklingError, pickle.PicklingError


def _to_static_embedder(obj: Any, do_zero_shot_default: bool) -> StaticLayerwiseEmbedding:
    """An opaque StaticLayerwiseEmbedding shell (or anything with its attributes) -> this package's module."""
    if isinstance(obj, StaticLayerwiseEmbedding):
        return obj
    if isinstance(obj, torch.Tensor):
        raise ValueError("checkpoint stores a plain embedding tensor for this placeholder (non-layerwise textual "
                         "inversion); only StaticLayerwiseEmbedding embedders are supported")
    attrs = obj.__dict__
    do_zero_shot = attrs.get("do_zero_shot", do_zero_shot_default)       # older checkpoints lack it (:2031-2033)
    if do_zero_shot:
        raise ValueError("zero-shot StaticLayerwiseEmbedding holds no parameters; its vectors come from a subject-basis "
                         "generator, which is outside this package")
    sd = obj.state_dict()
    L, K, r = sd["basis_rand_weights"].shape
    D = sd["basis_vecs"].shape[-1]
    n_pre = sd["pre_vecs"].shape[1] if "pre_vecs" in sd else 0
    mod = StaticLayerwiseEmbedding(num_layers=L, num_vectors_per_subj_token=K, out_emb_dim=D, r=r,
                                   init_vecs=torch.zeros(n_pre, D) if n_pre else None, has_bias="bias" in sd,
                                   token_string=attrs.get("token_string", ""), do_zero_shot=False)
    mod.load_state_dict(sd, strict=True)
    return mod.eval()


@dataclass
class EmbeddingCheckpoint:
    string_to_token: Dict[str, Any] = field(default_factory=dict)
    string_to_static_embedder: Dict[str, StaticLayerwiseEmbedding] = field(default_factory=dict)
    token2num_vectors: Dict[str, int] = field(default_factory=dict)
    emb_global_scale_scores: Dict[str, float] = field(default_factory=dict)
    use_conv_attn_kernel_size: Optional[int] = None
    subject_strings: List[str] = field(default_factory=list)
    background_strings: List[str] = field(default_factory=list)
    do_zero_shot: bool = False
    opaque: Dict[str, Any] = field(default_factory=dict)

    @property
    def placeholder_strings(self) -> List[str]:
        return self.subject_strings + self.background_strings              # :2044

    def emb_global_scales(self) -> Dict[str, float]:
        """sigmoid(score) + 0.5 per placeholder (embedding_manager.py:1744)."""
        return {k: float(torch.sigmoid(torch.tensor(v)) + 0.5) for k, v in self.emb_global_scale_scores.items()}

    def static_embeddings(self, device) -> Dict[str, torch.Tensor]:
        """placeholder -> [16, K, 768] fp32 on `device` (one kernel launch each)."""
        return {k: m.to(device)() for k, m in self.string_to_static_embedder.items()}


def read_raw(path: str) -> dict:
    """torch.load of one checkpoint file with every `ldm.*` class replaced by an opaque shell."""
    return torch.load(path, map_location="cpu", weights_only=False, pickle_module=_PickleModule)


def load_embedding_checkpoint(ckpt_paths: Union[str, Sequence[str]], do_zero_shot: bool = False) -> EmbeddingCheckpoint:
    """EmbeddingManager.load (embedding_manager.py:1840-2052) for sampling.  `ckpt_paths`: one path or several; a path
    may carry `:from-to,from-to` to rename placeholders (:1857-1866)."""
    if isinstance(ckpt_paths, str):
        ckpt_paths = [ckpt_paths]
    out = EmbeddingCheckpoint(do_zero_shot=do_zero_shot)
    for spec in ckpt_paths:
        parts = spec.split(":")
        mapper = dict(m.split("-") for m in parts[1].split(",")) if len(parts) == 2 else {}
        ckpt = read_raw(parts[0])
        ckpt_bg = ckpt.get("background_strings", [])
        if ckpt.get("use_conv_attn_kernel_size", None) is not None:
            out.use_conv_attn_kernel_size = ckpt["use_conv_attn_kernel_size"]
        for key in ("string_to_subj_basis_generator_dict", "ca_q_bns", "ca_outfeat_lns"):
            if key in ckpt:
                out.opaque[key] = ckpt[key]
        scores = ckpt.get("emb_global_scale_scores", None)
        for idx, km in enumerate(ckpt["placeholder_strings"]):
            km2 = mapper.get(km, km)
            if km2 in out.string_to_token:
                if km2 in out.background_strings:
                    continue                                               # :1996-1999: duplicate background, ignored
                raise ValueError(f"Duplicate key {km}->{km2} in {parts[0]}")
            if scores is not None:
                out.emb_global_scale_scores[km2] = float(scores.detach()[idx])
            if km in ckpt_bg:
                if km2 not in out.background_strings:
                    out.background_strings.append(km2)
            elif km2 not in out.subject_strings:
                out.subject_strings.append(km2)
            # the token id is the tokenizer's (:2018-2020); the stored one is kept for tokenizers extended the same way
            tok = ckpt["string_to_token"].get(km)
            out.string_to_token[km2] = int(tok) if isinstance(tok, torch.Tensor) else tok
            out.string_to_static_embedder[km2] = _to_static_embedder(ckpt["string_to_static_embedder"][km], do_zero_shot)
            if km in ckpt.get("token2num_vectors", {}):
                out.token2num_vectors[km2] = ckpt["token2num_vectors"][km]
    return out
