#!/usr/bin/env python
"""bench.py - DDIM steps/s of the AdaPrompt-conditioned SD-1.5 U-Net step on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # N > 1: launched by torchrun, one rank per GPU
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm's CPU path (oracle port)

One "step" = one DDIM step of BASELINE.json configs[1]: the full SD-1.5 U-Net at 64x64 latents (512x512) on
n_samples = 8 plus their classifier-free-guidance pair (batch 16, (cond, uncond) order), layerwise context
[16*B, 77, 768], the 16 AdaEmbedding layers recomputed from the layer features (K = 9 subject vectors, fg 5 / bg 4),
followed by the CFG combine + DDIM update.  Random-init weights of that architecture (zero-initialised modules
re-randomised), synthetic inputs.  Multi-GPU shards samples across ranks (weak scaling: 8 samples per
GPU), no collective inside the step; `value` = DDIM steps of an 8-sample batch completed per second over all ranks.

Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_SAMPLES = 8
LATENT = 64
DDIM_STEPS = 50
GUIDANCE = (10.0, 4.0)          # scripts/stable_txt2img.py --scale default: annealed 10 -> 4
SUBJ_K, SUBJ_FG, SUBJ_BG, SUBJ_R = 9, 5, 4, 12
SUBJ_POS = list(range(2, 2 + SUBJ_K))


def st_flops(n, c, t=77, d=768):
    """Algorithmic FLOPs of one SpatialTransformer per sample."""
    return 40.0 * n * c * c + 4.0 * n * n * c + 4.0 * n * t * c + 4.0 * t * d * c


def unet_st_flops(hw=None):
    hw = LATENT if hw is None else hw
    n = hw * hw
    return sum(k * st_flops(m, c) for m, c, k in ((n, 320, 5), (n // 4, 640, 5), (n // 16, 1280, 5), (n // 64, 1280, 1)))


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            m = json.load(open(path))
            p.update({k: float(m[k]) for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in m})
            p["source"] = "measured"
        except Exception:
            pass
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ================================================================================================ reference arm
def cpu_unet_step_time(steps, warmup, threads=None):
    """The reference algorithm's CPU path for this workload: oracle/port.py (fp32 torch-on-CPU restatement of the
    reference modules, pinned to the reference's outputs) - one U-Net call at 64x64 for ONE sample and its CFG pair
    (batch 2) plus the CFG/DDIM update.  8 such samples make one step of the benchmarked workload."""
    import torch
    from oracle import params as P
    from oracle import port
    if threads:
        torch.set_num_threads(threads)
    sd = P.init_unet(17)
    sch = port.make_ddim_schedule(DDIM_STEPS)
    x = P.randn("bench.x_T", (1, 4, LATENT, LATENT), 42)
    cc = P.randn("bench.c", (16, 77, 768), 3)
    cu = P.randn("bench.uc", (16, 77, 768), 3)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            idx = DDIM_STEPS - 1 - (i % DDIM_STEPS)
            t0 = time.perf_counter()
            ts = torch.full((2,), int(sch["timesteps"][idx]), dtype=torch.long)
            e_c, e_u = port.unet_forward(sd, torch.cat([x, x]), ts, torch.cat([cc, cu])).chunk(2)
            x, _ = port.cfg_ddim_update(x, e_c, e_u, 10.0, float(sch["alphas"][idx]), float(sch["alphas_prev"][idx]),
                                        0.0, float(sch["sqrt_one_minus_alphas"][idx]))
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return sum(times) / len(times), torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = args.steps if args.steps is not None else 3
    warmup = args.warmup if args.warmup is not None else 1
    t_pair, threads = cpu_unet_step_time(steps, warmup)
    ms = t_pair * N_SAMPLES * 1e3
    val = 1e3 / ms
    sample = (f"oracle/port.py fp32 on {threads} host threads: U-Net + CFG/DDIM update for 1 sample + its CFG pair "
              f"(batch 2, 64x64) = 1/8 of a step, x8")
    line = {"impl": "reference", "metric": "ddim_steps_per_s", "value": val, "unit": "steps/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(1),
            "cpu_baseline": {"value": val, "unit": "steps/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=_RESULT_OUT, flush=True)


def workload_config(world):
    px = 8 * LATENT
    which = "BASELINE.json configs[1]" if (LATENT, N_SAMPLES) == (64, 8) else \
        "geometry and batch of BASELINE.json configs[4] (the ArcFace identity tokens are synthetic context rows like all others)"
    return {"workload": f"SD-1.5 U-Net DDIM step {px}x{px} ({LATENT}x{LATENT} latent), n_samples={N_SAMPLES} + CFG pair "
                        f"(batch {2 * N_SAMPLES}) per GPU, guidance (10,4) annealed, 16 AdaEmbedding layers (K=9, fg5/bg4, "
                        f"r=12), layerwise context [16B,77,768], random-init weights ({which})",
            "n_samples_per_gpu": N_SAMPLES, "batch_incl_cfg_per_gpu": 2 * N_SAMPLES, "latent": [4, LATENT, LATENT],
            "ddim_steps_per_image": DDIM_STEPS, "parallelism": f"samples sharded over {world} GPU(s), no collective "
            "inside the step", "resblocks": "Upsample = nearest-2x folded into four 2x2 phase convolutions on this repo's implicit-GEMM tcgen05 kernel; "
                         "ResBlock 3x3 / stride-2 / 4-channel convolutions and concat on torch/cuDNN bf16 channels-last unless "
                         "AFB_CONV=1 (outside the hot path, included in the step)",
            "l2": "no explicit flush: each step streams 1.7 GB of bf16 weights plus > 2 GB of activations, far above "
                  "the 126 MB L2"}


# ================================================================================================ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    steps = args.steps if args.steps is not None else 20
    warmup = max(args.warmup if args.warmup is not None else 3, 3)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.backends.cudnn.benchmark = True       # the torch ResBlock convolutions pick their fastest cuDNN algorithm
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from adaface_b200 import AdaEmbedding, DDIMSampler, _lib, ops
    from adaface_b200.unet import (SD15_UNET_CONFIG, AdaPromptContext, AdaSubject, DiffusionModel, UNetModel,
                                   init_random_)
    torch.manual_seed(0)
    with torch.device(dev):
        unet = UNetModel(**SD15_UNET_CONFIG)
        ada = AdaEmbedding(num_vectors_per_subj_token=SUBJ_K, fg_emb_count=SUBJ_FG, bg_emb_count=SUBJ_BG, r=SUBJ_R,
                           init_vecs=torch.randn(2, 768))
    init_random_(unet, 0).eval()
    init_random_(ada, 1).eval()
    ada.want_pooler_attn = False
    model = DiffusionModel(unet, use_cuda_graph=not args.no_graph)
    sampler = DDIMSampler(model)
    sampler.match_reference_rng = False
    sampler.make_schedule(DDIM_STEPS, ddim_eta=0.0)
    scales = __import__("adaface_b200.ddim", fromlist=["guidance_schedule"]).guidance_schedule(list(GUIDANCE), DDIM_STEPS)

    # every rank draws the same global tensors and keeps its own 8-sample slice (weak scaling)
    g = torch.Generator(device="cpu").manual_seed(42)
    b = N_SAMPLES
    x_T_h = torch.randn((b, 4, LATENT, LATENT), generator=g).pin_memory()
    g.manual_seed(3 + rank)
    c_h = torch.randn((b * 16, 77, 768), generator=g).pin_memory()
    uc_h = torch.randn((16, 77, 768), generator=g).repeat(b, 1, 1).pin_memory()
    probes = torch.randn((16, 768), generator=g).to(dev) * 0.5
    extra = torch.randn((16, b, 768), generator=g).to(dev) * 0.5
    ada_ctx = AdaPromptContext([AdaSubject(ada, SUBJ_POS, probes, extra, n_instances=b)])
    info = {"use_layerwise_context": True, "is_training": False, "use_conv_attn_kernel_size": -1,
            "placeholder2indices": None, "ada_prompt": ada_ctx}
    x_d, c_d, uc_d = x_T_h.to(dev), c_h.to(dev), uc_h.to(dev)
    cond, uncond = (c_d, ["a z"] * b, info), (uc_d, [""] * b, info)
    x_out_h = torch.empty_like(x_T_h).pin_memory()

    def ddim_step(x, i):
        idx = DDIM_STEPS - 1 - (i % DDIM_STEPS)
        ts = torch.full((b,), int(sampler.ddim_timesteps[idx]), device=dev, dtype=torch.long)
        x_prev, _ = sampler.p_sample_ddim(x, cond, ts, index=idx, guidance_scale=scales[i % DDIM_STEPS],
                                          unconditional_conditioning=uncond)
        return x_prev

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.no_grad():
        # ---- warm-up (builds packed weights, captures the U-Net graph)
        n_launch0 = _lib.launch_count()
        x = x_d
        for i in range(warmup):
            x = ddim_step(x, i)
        torch.cuda.synchronize()

        # launches of OUR kernels per step: count one eager step through the C ABI (graph replays do not re-enter it)
        model.use_cuda_graph = False
        n0 = _lib.launch_count()
        ddim_step(x_d, 0)
        torch.cuda.synchronize()
        launches_per_step = _lib.launch_count() - n0
        model.use_cuda_graph = not args.no_graph

        if args.profile_step:
            # for `ncu --profile-from-start off`: exactly one eager (non-graph) step between cudaProfilerStart/Stop
            model.use_cuda_graph = False
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            ddim_step(x_d, 0)
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            model.use_cuda_graph = not args.no_graph

        # ---- timed region: device-resident inputs
        clocks = ClockSampler(local)
        barrier()
        clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t_wall = time.perf_counter()
        e0.record()
        x = x_d
        for i in range(steps):
            x = ddim_step(x, i)
        e1.record()
        barrier()
        t_wall = time.perf_counter() - t_wall
        clk = clocks.stop()
        ms_dev = max_over_ranks(e0.elapsed_time(e1) / steps)
        ms_wall = max_over_ranks(t_wall * 1e3 / steps)
        finite = bool(torch.isfinite(x).all())

        # ---- e2e: same step through the public API with HOST buffers (pinned), H2D of the step's inputs and D2H of
        # its result inside the timed region
        # The context of step i+1 does not depend on step i: its H2D copy is issued on a copy stream into a staging
        # buffer while step i computes (double buffering), and moved into the step's (graph-captured) input buffers
        # by a device copy when the step starts.  x_t and the read-back of x_prev stay on the compute stream.
        copy_stream = torch.cuda.Stream()
        stage = [(torch.empty_like(c_d), torch.empty_like(uc_d)) for _ in range(2)]
        staged = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def prefetch(i):
            k = i & 1
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[k])          # the step that last read this staging buffer is done with it
                stage[k][0].copy_(c_h, non_blocking=True)
                stage[k][1].copy_(uc_h, non_blocking=True)
                staged[k].record(copy_stream)

        def e2e_step(i):
            k = i & 1
            main = torch.cuda.current_stream()
            x_d.copy_(x_T_h, non_blocking=True)
            main.wait_event(staged[k])
            c_d.copy_(stage[k][0], non_blocking=True)
            uc_d.copy_(stage[k][1], non_blocking=True)
            consumed[k].record(main)
            prefetch(i + 1)                                   # overlaps this step's compute
            xp = ddim_step(x_d, i)
            x_out_h.copy_(xp, non_blocking=True)
            main.synchronize()

        for k in range(2):
            consumed[k].record(torch.cuda.current_stream())
        prefetch(0)
        for i in range(2):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(2, 2 + steps):
            e2e_step(i)
        barrier()
        ms_e2e = max_over_ranks((time.perf_counter() - t0) * 1e3 / steps)
        h2d = x_T_h.numel() * 4 + c_h.numel() * 4 + uc_h.numel() * 4
        d2h = x_out_h.numel() * 4

        # ---- kernel breakdown: one eager step with CUDA events around every C-ABI launch.  A device-side sleep lets
        # the host run ahead so that the events bracket back-to-back kernels, not launch gaps.
        model.use_cuda_graph = False
        x_d.copy_(x_T_h)
        torch.cuda.synchronize()
        torch.cuda._sleep(int(0.06 * 1.9e9))
        ops.start_timing()
        ddim_step(x_d, 0)
        recs = ops.stop_timing()
        model.use_cuda_graph = not args.no_graph

    if args.dump_kernels and rank == 0:
        by_tag = {}
        for r in recs:
            e = by_tag.setdefault((r["name"], r["tag"]), {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "n": 0})
            e["ms"] += r["ms"]; e["flops"] += r["flops"]; e["bytes"] += r["bytes"]; e["n"] += 1
        with open(args.dump_kernels, "w") as fh:
            fh.write("kernel,shape,launches,total_ms,avg_us,tflops,gbs\n")
            for (n, tag), e in sorted(by_tag.items(), key=lambda kv: -kv[1]["ms"]):
                fh.write(f"{n},{tag},{e['n']},{e['ms']:.4f},{1e3 * e['ms'] / e['n']:.1f},"
                         f"{e['flops'] / (e['ms'] * 1e-3) / 1e12:.1f},{e['bytes'] / (e['ms'] * 1e-3) / 1e9:.0f}\n")
    groups = {}
    for r in recs:
        gk = groups.setdefault(r["name"], {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "launches": 0})
        gk["ms"] += r["ms"]; gk["flops"] += r["flops"]; gk["bytes"] += r["bytes"]; gk["launches"] += 1
    # ResBlock / Upsample convolutions on this repo's implicit-GEMM kernel are outside the
    # SpatialTransformer path: reported in `kernels` and `resblock_conv`, not in `st_path`
    NON_ST = ("conv3x3_bf16", "conv3x3_up2_bf16")
    is_st = lambda r: r.get("scope", "st") == "st" and r["name"] not in NON_ST
    st_ms = sum(r["ms"] for r in recs if is_st(r))
    prep_ms = sum(r["ms"] for r in recs if is_st(r) and r["name"] == "ada_pool_prep")
    shell_ms = sum(r["ms"] for r in recs if not is_st(r))          # this repo's kernels inside ResBlocks / Upsample / head
    conv_ms = sum(v["ms"] for k, v in groups.items() if k in NON_ST)
    conv_flops = sum(v["flops"] for k, v in groups.items() if k in NON_ST)
    pk = peaks()
    st_alg_flops = unet_st_flops() * 2 * b
    # dominant kernel = the single launch configuration (kernel, shape) with the most device time in a step among
    # the tensor-bound ones; all its launches are identical, so "average launch duration" is well defined
    by_shape = {}
    for r in recs:
        if r["flops"] > 0 and is_st(r):
            e = by_shape.setdefault((r["name"], r["tag"]), {"ms": 0.0, "flops": 0.0, "n": 0})
            e["ms"] += r["ms"]; e["flops"] += r["flops"]; e["n"] += 1
    (top, top_tag), tg = max(by_shape.items(), key=lambda kv: kv[1]["ms"])
    achieved = tg["flops"] / (tg["ms"] * 1e-3) / 1e12
    traffic = None
    try:    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_dram_traffic.json"))).get(f"{top}:{top_tag}")
    except Exception:
        pass
    roofline = {"kernel": f"{top}[{top_tag}]", "bound": "tensor", "achieved": achieved, "peak": pk["bf16_tflops_sustained"],
                "unit": "TFLOP/s", "frac": achieved / pk["bf16_tflops_sustained"], "traffic": traffic,
                "traffic_unit": "bytes per launch (ncu dram read + write)" if traffic is not None else None,
                "peak_source": f"{pk['source']} (MEASURED_PEAKS.json bf16_tflops_sustained: kernel timed inside a long step)",
                "launches_per_step": tg["n"], "avg_launch_us": 1e3 * tg["ms"] / tg["n"], "ms_per_step": tg["ms"],
                "flops_per_launch": tg["flops"] / tg["n"],
                "note": "algorithmic FLOPs (unpadded head dim / context) of one launch / its average CUDA-event time "
                        "in one instrumented eager step; per-group figures for every kernel are under 'kernels'"}
    breakdown = {k: {"ms": round(v["ms"], 4), "launches": v["launches"],
                     "tflops": (v["flops"] / (v["ms"] * 1e-3) / 1e12) if v["flops"] else None,
                     "gbs": (v["bytes"] / (v["ms"] * 1e-3) / 1e9) if not v["flops"] else None} for k, v in groups.items()}

    if rank == 0:
        value = world * 1e3 / ms_dev
        line = {"metric": "ddim_steps_per_s", "value": value, "unit": "steps/s", "n_gpus": world, "steps": steps,
                "warmup": warmup, "ms_per_step": ms_dev, "ms_per_step_wall": ms_wall, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": workload_config(world), "clocks": clk,
                "e2e": {"value": world * 1e3 / ms_e2e, "unit": "steps/s", "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "api": "DDIMSampler.p_sample_ddim on pinned host x_T / context, result read back; every step copies its "
                               "x_t and its whole context from pinned host memory (the context copy of step i+1 is issued on a "
                               "copy stream while step i computes) and reads x_prev back before the next step starts"},
                "gpu_launches": launches_per_step * steps, "gpu_launches_per_step": launches_per_step,
                "cuda_graph": not args.no_graph, "finite": finite, "roofline": roofline,
                "st_path": {"ms_per_step": st_ms, "algorithmic_tflop_per_step": st_alg_flops / 1e12,
                            "tflops": st_alg_flops / (st_ms * 1e-3) / 1e12,
                            "frac_of_bf16_peak": st_alg_flops / (st_ms * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
                            "ms_per_step_critical_path": st_ms - prep_ms,
                            "tflops_critical_path": st_alg_flops / ((st_ms - prep_ms) * 1e-3) / 1e12,
                            "note": "SpatialTransformer + AdaEmbedding + CFG/DDIM kernels only (sum of CUDA-event times "
                                    "of their C-ABI launches in one eager step); the *_critical_path figures leave out "
                                    "ada_pool_prep, which runs on a side stream concurrently with the first ResBlocks"},
                "shell_kernels_ms_per_step": shell_ms,
                "non_st_ms_per_step": ms_dev - st_ms,
                "resblock_conv": {"ms_per_step": conv_ms, "tflops": (conv_flops / (conv_ms * 1e-3) / 1e12) if conv_ms else None,
                                  "note": "convolutions that ran as implicit GEMMs on this repo's tcgen05 GEMM kernel (Upsample folds by "
                                          "default, every 3x3 ResBlock convolution with AFB_CONV=1); part of non_st_ms_per_step"},
                "kernels": breakdown}
        if world == 1 and not args.no_cpu_baseline:
            n_cpu = 5 if LATENT <= 64 else 1                     # ~10-15 s of CPU work on the box's host cores
            t_pair, threads = cpu_unet_step_time(n_cpu, 1)
            line["cpu_baseline"] = {"value": 1.0 / (t_pair * N_SAMPLES), "unit": "steps/s", "cores": threads,
                                    "kind": "port", "sample": f"oracle/port.py fp32 on {threads} host threads: U-Net call + "
                                    f"CFG/DDIM update for 1 sample + its CFG pair (batch 2, {LATENT}x{LATENT}), mean of "
                                    f"{n_cpu} calls after 1 warm-up = {t_pair:.2f} s; a step is {N_SAMPLES} of those "
                                    f"(x{N_SAMPLES} extrapolation)"}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), file=_RESULT_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_RESULT_OUT = sys.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-step", action="store_true",
                    help="bracket one eager step with cudaProfilerStart/Stop (ncu --profile-from-start off)")
    ap.add_argument("--dump-kernels", default=None, help="write the per-shape kernel breakdown of one step as CSV")
    ap.add_argument("--workload", default="config2", choices=["config2", "config5"],
                    help="config2 (default, the benchmarked one): 64x64 latents, 8 samples + CFG pair; config5: 96x96 "
                         "latents, 16 samples + CFG pair (BASELINE.json configs[4] geometry; an extra data point)")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON result: libraries that chat on file descriptor 1 (NCCL prints
    # "NCCL version ..." there when NCCL_DEBUG=VERSION is set on the box) are sent to stderr instead
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.workload == "config5":
        global N_SAMPLES, LATENT
        N_SAMPLES, LATENT = 16, 96
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
