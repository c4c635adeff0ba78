"""Imports the UNMODIFIED reference modules from /root/reference on CPU, with the harness shims
SURVEY.md section 8(c) lists.  Build-container only: /root/reference does not exist on the GPU box, so
nothing in tests/, smoke() or bench.py imports this file.  Used by oracle/gen_golden.py.

TEST INFRASTRUCTURE ONLY (see oracle/port.py header).
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ADAFACE_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "ldm"))


def _install_omegaconf_stub():
    # openaimodel.py:480 imports omegaconf.listconfig.ListConfig only for an isinstance-style check.
    if "omegaconf" in sys.modules:
        return
    oc = types.ModuleType("omegaconf")
    lc = types.ModuleType("omegaconf.listconfig")

    class ListConfig(list):
        pass

    lc.ListConfig = ListConfig
    oc.listconfig = lc
    oc.ListConfig = ListConfig
    sys.modules["omegaconf"] = oc
    sys.modules["omegaconf.listconfig"] = lc


def load():
    """Returns a namespace with the reference classes on the path."""
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    _install_omegaconf_stub()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import torch
    from ldm.modules import attention as ref_attention
    from ldm.modules import embedding_manager as ref_em
    from ldm.modules.diffusionmodules import openaimodel as ref_unet
    from ldm.models.diffusion import ddim as ref_ddim
    from ldm import util as ref_util

    class CPUDDIMSampler(ref_ddim.DDIMSampler):
        # ddim.py:22-26 moves every buffer to cuda unconditionally; keep it where it is.
        def register_buffer(self, name, attr):
            setattr(self, name, attr)

    class ModelStandIn:
        """What DDIMSampler needs from LatentDiffusion (ddpm.py:244-265, :2292, :5495-5505)."""

        def __init__(self, unet=None, apply_fn=None):
            from ldm.modules.diffusionmodules.util import make_beta_schedule
            import numpy as np
            betas = make_beta_schedule("linear", 1000, linear_start=0.00085, linear_end=0.012)
            alphas_cumprod = np.cumprod(1.0 - betas, axis=0)
            alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
            f32 = lambda a: torch.tensor(a, dtype=torch.float32)
            self.num_timesteps = 1000
            self.betas = f32(betas)
            self.alphas_cumprod = f32(alphas_cumprod)
            self.alphas_cumprod_prev = f32(alphas_cumprod_prev)
            self.device = torch.device("cpu")
            self.unet = unet
            self.apply_fn = apply_fn
            self.calls = []

        def apply_model(self, x, t, c):
            self.calls.append((tuple(x.shape), t.clone()))
            if self.apply_fn is not None:
                return self.apply_fn(x, t, c)
            ctx, ctx_in, extra_info = c
            return self.unet(x, t, context=ctx, context_in=ctx_in, extra_info=extra_info)

    @contextlib.contextmanager
    def fp32_autocast():
        """AdaEmbedding.forward hard-codes torch.autocast(enabled=True) (embedding_manager.py:797), which is
        bf16 on CPU.  For the fp32 oracle replace it by a null context for the duration of the call."""
        saved = torch.autocast

        class _Null:
            def __init__(self, *a, **k):
                pass

            def __enter__(self):
                return self

            def __exit__(self, *a):
                return False

        torch.autocast = _Null
        try:
            yield
        finally:
            torch.autocast = saved

    ns = types.SimpleNamespace(
        torch=torch, attention=ref_attention, embedding_manager=ref_em, openaimodel=ref_unet,
        ddim=ref_ddim, util=ref_util, CPUDDIMSampler=CPUDDIMSampler, ModelStandIn=ModelStandIn,
        fp32_autocast=fp32_autocast)
    return ns


SD15_UNET_CFG = dict(image_size=32, in_channels=4, out_channels=4, model_channels=320,
                     attention_resolutions=[4, 2, 1], num_res_blocks=2, channel_mult=[1, 2, 4, 4],
                     num_heads=8, use_spatial_transformer=True, transformer_depth=1, context_dim=768,
                     use_checkpoint=True, legacy=False)          # v1-inference-ada.yaml:35-51


def default_extra_info():
    # the U-Net only runs with a dict extra_info, layerwise context and an integer kernel size.
    return {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1,
            "placeholder2indices": None}
