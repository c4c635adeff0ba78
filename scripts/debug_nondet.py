"""Developer check: is a U-Net call reproducible?  Runs the same call twice eagerly and reports, per module, the
relative difference between the two runs (0 = bit-identical)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import params as P
from adaface_b200.unet import SD15_UNET_CONFIG, UNetModel
from adaface_b200.attention import SpatialTransformer

hw = int(sys.argv[1]) if len(sys.argv) > 1 else 16
d = torch.device("cuda:0")
with torch.device("meta"):
    unet = UNetModel(**SD15_UNET_CONFIG)
unet.load_state_dict(P.init_unet(17), assign=True, strict=True)
unet = unet.to(d).eval()
x = P.randn("gk.x", (2, 4, hw, hw), 5).to(d)
t = torch.full((2,), 500, dtype=torch.long, device=d)
ctx = P.randn("gk.ctx", (32, 77, 768), 5).to(d)
info = {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1, "placeholder2indices": None}
runs = []
for rep in range(3):
    outs = []
    hooks = []
    for name, m in unet.named_modules():
        if isinstance(m, SpatialTransformer) or type(m).__name__ in ("ResBlock", "Upsample", "Downsample"):
            hooks.append(m.register_forward_hook(lambda mod, inp, out, name=name: outs.append((name, out.float().clone()))))
    with torch.no_grad():
        eps = unet(x, t, context=ctx, extra_info=info)
    for h in hooks:
        h.remove()
    outs.append(("eps", eps.float().clone()))
    runs.append(outs)
for (n0, a), (_, b), (_, c) in zip(*runs):
    r1 = float((a - b).norm() / a.norm()); r2 = float((b - c).norm() / a.norm())
    if r1 > 0 or r2 > 0:
        print(f"{n0:40s} {tuple(a.shape)} run0-vs-1 {r1:.2e}  run1-vs-2 {r2:.2e}")
print("done")
