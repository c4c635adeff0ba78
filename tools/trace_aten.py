"""Developer probe: which ATen ops (with input shapes) still launch kernels inside one eager U-Net call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from adaface_b200.unet import SD15_UNET_CONFIG, UNetModel, init_random_

dev = torch.device("cuda:0")
with torch.device(dev):
    unet = UNetModel(**SD15_UNET_CONFIG)
init_random_(unet, 0).eval()
x = torch.randn(16, 4, 64, 64, device=dev)
ctx = torch.randn(256, 77, 768, device=dev)
t = torch.full((16,), 501, dtype=torch.long, device=dev)
info = {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1, "placeholder2indices": None}
with torch.no_grad():
    for _ in range(2):
        unet(x, t, context=ctx, extra_info=info)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=True) as prof:
        unet(x, t, context=ctx, extra_info=info)
        torch.cuda.synchronize()
rows = []
for e in prof.key_averages(group_by_input_shape=True, group_by_stack_n=4):
    if e.key.startswith("aten::") and e.self_device_time_total > 0:
        rows.append((e.self_device_time_total, e.count, e.key, str(e.input_shapes)[:90], " <- ".join(s.split("/")[-1] for s in e.stack[:3])[:150]))
for r in sorted(rows, reverse=True)[:25]:
    print(f"{r[0]:9.1f} us  x{r[1]:<3d} {r[2]:28s} {r[3]}  {r[4]}")
