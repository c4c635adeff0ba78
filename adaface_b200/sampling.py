"""Multi-GPU sampling: shard the sample / prompt axis across ranks, one process per GPU.

The reference samples on a single GPU (scripts/stable_txt2img.py:238,342); each sample's DDIM trajectory
depends only on its own x_T and context, so the path shards with NO collective inside the step loop
(SURVEY.md 8e).  The only exchange is one all-gather of the final latents (NCCL over NVLink on the GPU box,
gloo in the CPU tests).  A sample's CFG pair (cond, uncond) always stays on one rank.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

Tensor = torch.Tensor


def shard_bounds(n_items: int, world_size: int) -> List[Tuple[int, int]]:
    """Contiguous, balanced [lo, hi) ranges; the first n_items % world_size ranks hold one extra item."""
    if n_items < 0 or world_size <= 0:
        raise ValueError("shard_bounds: n_items >= 0 and world_size > 0 expected")
    base, extra = divmod(n_items, world_size)
    out, lo = [], 0
    for r in range(world_size):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def shard_layerwise_context(context: Tensor, lo: int, hi: int, n_items: int) -> Tensor:
    """context [16*n, T, D] is instance-major / layer-minor (embedding_manager.py:1345-1349) or [n, T, D]."""
    if context.shape[0] % n_items:
        raise ValueError(f"context batch {context.shape[0]} is not a multiple of the sample count {n_items}")
    per = context.shape[0] // n_items
    return context[lo * per:hi * per]


def all_gather_samples(local: Tensor, bounds: Sequence[Tuple[int, int]], group=None) -> Tensor:
    """Gathers per-rank [n_r, ...] latents into [n, ...] on every rank (ranks may hold unequal counts)."""
    world = len(bounds)
    if world == 1 or not (dist.is_available() and dist.is_initialized()):
        return local
    counts = [hi - lo for lo, hi in bounds]
    width = max(counts)
    item_shape = tuple(local.shape[1:])
    padded = local.new_zeros((width,) + item_shape)
    padded[:local.shape[0]] = local
    bufs = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(bufs, padded.contiguous(), group=group)
    return torch.cat([b[:c] for b, c in zip(bufs, counts)], dim=0)


def sample_sharded(run_local: Callable[[Tensor, Tensor, Tensor], Tensor], x_T: Tensor, c_cond: Tensor,
                   c_uncond: Tensor, rank: Optional[int] = None, world_size: Optional[int] = None,
                   gather: bool = True, group=None) -> Tensor:
    """Runs `run_local(x_T_shard, c_cond_shard, c_uncond_shard) -> latents_shard` on this rank's slice of the
    samples and all-gathers the result.  x_T [n, 4, h, w] must be the SAME tensor on every rank (drawn from one
    seed), so that the union over ranks equals the single-GPU run sample for sample."""
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    n = x_T.shape[0]
    bounds = shard_bounds(n, world_size)
    lo, hi = bounds[rank]
    if hi > lo:
        out = run_local(x_T[lo:hi], shard_layerwise_context(c_cond, lo, hi, n),
                        shard_layerwise_context(c_uncond, lo, hi, n))
    else:
        out = x_T[:0].clone()
    return all_gather_samples(out, bounds, group) if gather else out


class FlatGradAllReduce:
    """The one collective of inversion mode (BASELINE.json configs[3]): the U-Net is frozen and
    replicated, only the AdaEmbedding / StaticLayerwiseEmbedding parameters (3.8 M floats = 15 MB) carry gradients,
    so a data-parallel optimizer step needs ONE all-reduce: every gradient is copied into a persistent flat fp32
    buffer, summed over the ranks (NCCL over NVLink; gloo in the CPU tests), divided by the world size and copied
    back - the arithmetic of the reference's Lightning-DDP wrapper (main.py:862) on its trainable set.  The gradients
    come from the backward kernels registered on the custom ops (csrc/backward.cu).
    Parameters whose `.grad` is None on a rank contribute zeros (DDP's find_unused_parameters semantics)."""

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatGradAllReduce: no trainable parameters")
        dev = self.params[0].device
        if any(p.device != dev for p in self.params):
            raise ValueError("FlatGradAllReduce: parameters on different devices")
        self.group = group
        self.offsets, n = [], 0
        for p in self.params:
            self.offsets.append(n)
            n += p.numel()
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)

    def _is_view(self, p, off) -> bool:
        return (p.grad is not None and p.grad.dtype == torch.float32 and p.grad.is_contiguous()
                and p.grad.data_ptr() == self.flat.data_ptr() + 4 * off)

    @torch.no_grad()
    def attach(self) -> "FlatGradAllReduce":
        """Makes every fp32 parameter's `.grad` a view of the flat buffer (keeping what it held), so that backward
        accumulates straight into it and a call is just the all-reduce: no gather / scatter copies (83 tensors for
        an AdaEmbedding).  `zero_grad(set_to_none=True)` detaches the views; the call then falls back to copies."""
        for p, off in zip(self.params, self.offsets):
            if p.dtype != torch.float32 or self._is_view(p, off):
                continue
            view = self.flat[off:off + p.numel()].view_as(p)
            if p.grad is not None:
                view.copy_(p.grad)
            else:
                view.zero_()
            p.grad = view
        return self

    @torch.no_grad()
    def __call__(self) -> Tensor:
        """Averages the gradients over the ranks in place; returns the flat buffer (a view of what was reduced)."""
        views = [self._is_view(p, off) for p, off in zip(self.params, self.offsets)]
        for p, off, v in zip(self.params, self.offsets, views):
            if v:
                continue
            seg = self.flat[off:off + p.numel()]
            if p.grad is not None:
                seg.copy_(p.grad.reshape(-1))
            else:
                seg.zero_()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
            self.flat.div_(dist.get_world_size(self.group))
        for p, off, v in zip(self.params, self.offsets, views):
            if v:
                continue
            g = self.flat[off:off + p.numel()].view_as(p)
            if p.grad is None:
                p.grad = g.clone().to(p.dtype)
            else:
                p.grad.copy_(g)
        return self.flat


class CapturedGradStep:
    """Forward + loss + backward of one inversion step (ddpm.py:519-637: training_step -> guided_denoise ->
    calc_recon_loss -> manual_backward) as ONE CUDA graph.  The step is ~750 launches of this repo's kernels plus the
    torch ops of the U-Net shell, and issued eagerly it is bound by the host (the kernels of one batch-8 step add up to
    about 2/3 of its wall time); replayed as a graph the device runs them back to back.

    `loss_fn(*inputs) -> scalar` must be shape-static and free of host synchronisation; the parameters' `.grad` must be
    views of `reducer.flat` (`FlatGradAllReduce.attach()`), which the graph zeroes first, so that a replay leaves this
    step's gradients in the flat buffer, ready for the all-reduce and the optimizer (both stay outside the graph: NCCL
    and the optimizer's step counter).  Train-mode AttentionalPooler dropout takes its seed on the host per call, so a
    captured step would replay ONE mask: modules must be in eval mode (checked by the caller that owns them).  No
    autograd graph from an earlier eager step over the same parameters may be alive at construction (its AccumulateGrad
    nodes are bound to the stream they ran on and would pull that stream into the capture).
    Not a CPU path: capture needs a CUDA device."""

    def __init__(self, loss_fn: Callable[..., Tensor], example_inputs: Sequence[Tensor], reducer: FlatGradAllReduce,
                 warmup: int = 3):
        if not example_inputs or not all(t.is_cuda for t in example_inputs):
            raise RuntimeError("CapturedGradStep: CUDA example inputs expected (CUDA graphs have no CPU counterpart)")
        for p_, off in zip(reducer.params, reducer.offsets):
            if not reducer._is_view(p_, off):
                raise ValueError("CapturedGradStep: call FlatGradAllReduce.attach() first - every .grad must live in the flat buffer")
        self.reducer = reducer
        self.inputs = [t.detach().clone() for t in example_inputs]
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):                         # warm-up off the capture stream (allocator, lazy inits)
            for _ in range(max(warmup, 1)):
                reducer.flat.zero_()
                loss_fn(*self.inputs).backward()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            reducer.flat.zero_()
            loss = loss_fn(*self.inputs)
            loss.backward()
        self.loss = loss.detach()                             # drops the captured autograd graph and its AccumulateGrad nodes
        del loss
        for p_, off in zip(reducer.params, reducer.offsets):
            if not reducer._is_view(p_, off):
                raise RuntimeError("CapturedGradStep: backward replaced a .grad tensor during capture")

    def __call__(self, *inputs: Tensor) -> Tensor:
        """Copies the inputs into the graph's buffers and replays; returns the (graph-owned) loss tensor."""
        if len(inputs) != len(self.inputs):
            raise ValueError(f"CapturedGradStep: {len(self.inputs)} inputs expected")
        for dst, src in zip(self.inputs, inputs):
            if dst.shape != src.shape or dst.dtype != src.dtype:
                raise ValueError("CapturedGradStep: input shape / dtype differs from the captured one")
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.loss
