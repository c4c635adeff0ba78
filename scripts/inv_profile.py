#!/usr/bin/env python
"""Where the device time of one eager inversion step goes, kernel by kernel (torch.profiler, CUDA activities): this repo's
kernels next to the torch / cuDNN kernels of the U-Net shell (ResBlocks, resampling, loss).  Usage: python scripts/inv_profile.py [out.txt]"""
import sys

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402


def main(out):
    from adaface_b200.sampling import FlatGradAllReduce
    from adaface_b200.unet import AdaPromptContext, AdaSubject, DiffusionModel
    dev = torch.device("cuda", 0)
    torch.backends.cudnn.benchmark = True
    b = bench.N_SAMPLES
    unet, adas = bench.build_modules(1)
    unet, ada = unet.to(dev).requires_grad_(False), adas[0].to(dev)
    x_all, c_all, _, probes, extra = bench.synthetic_inputs(b, 1, bench.LATENT)
    x_start, ctx = x_all[:b].to(dev), c_all[:16 * b].to(dev)
    ada_ctx = AdaPromptContext([AdaSubject(ada, bench.subject_positions(0), probes[0].to(dev), extra[0][:, :b].contiguous().to(dev),
                                           n_instances=b)])
    info = {"use_layerwise_context": True, "is_training": True, "use_conv_attn_kernel_size": -1, "placeholder2indices": None,
            "ada_prompt": ada_ctx}
    model = DiffusionModel(unet, use_cuda_graph=False)
    params = [p for p in ada.parameters() if p.requires_grad]
    reducer = FlatGradAllReduce(params).attach()
    gen = torch.Generator(device=dev).manual_seed(1234)

    def step():
        t = torch.randint(0, 1000, (b,), device=dev, generator=gen)
        noise = torch.randn(x_start.shape, device=dev, generator=gen)
        eps = unet(model.q_sample(x_start, t, noise), t, context=ctx, extra_info=info)
        reducer.flat.zero_()
        ((eps.float() - noise) ** 2).mean().backward()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    table = prof.key_averages().table(sort_by="cuda_time_total", row_limit=60, max_name_column_width=90)
    open(out, "w").write(table)
    print(table[:6000])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/inv_profile.txt")
