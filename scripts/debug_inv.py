"""Developer check of the inversion step at full size: finiteness of the loss and of every gradient, step by step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from adaface_b200.unet import AdaPromptContext, AdaSubject, DiffusionModel
hw = int(sys.argv[1]) if len(sys.argv) > 1 else 64
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0")
unet, adas = bench.build_modules(1)
unet, ada = unet.to(dev).requires_grad_(False), adas[0].to(dev)
x_all, c_all, _, probes, extra = bench.synthetic_inputs(b, 1, hw)
x_start, ctx = x_all.to(dev), c_all.to(dev)
ada_ctx = AdaPromptContext([AdaSubject(ada, bench.subject_positions(0), probes[0].to(dev), extra[0].to(dev), n_instances=b)])
info = {"use_layerwise_context": True, "is_training": True, "use_conv_attn_kernel_size": -1, "placeholder2indices": None, "ada_prompt": ada_ctx}
model = DiffusionModel(unet)
gen = torch.Generator(device=dev).manual_seed(1234)
for it in range(2):
    t = torch.randint(0, 1000, (b,), device=dev, generator=gen)
    noise = torch.randn(x_start.shape, device=dev, generator=gen)
    x_noisy = model.q_sample(x_start, t, noise)
    eps = unet(x_noisy, t, context=ctx, extra_info=info)
    loss = ((eps.float() - noise) ** 2).mean()
    print("iter", it, "t", t.tolist(), "eps finite", bool(torch.isfinite(eps).all()), "loss", float(loss.detach()))
    ada.zero_grad()
    loss.backward()
    bad = [(k, int((~torch.isfinite(p.grad)).sum())) for k, p in ada.named_parameters() if p.grad is not None and not bool(torch.isfinite(p.grad).all())]
    print("  non-finite grads:", bad[:8], "of", len(bad))
    gn = sum(float(p.grad.float().pow(2).sum()) for p in ada.parameters() if p.grad is not None and bool(torch.isfinite(p.grad).all())) ** 0.5
    print("  grad norm (finite ones)", gn)
