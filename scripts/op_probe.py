#!/usr/bin/env python
"""Device time of single ops without host starvation: each case is captured as a CUDA graph of N back-to-back calls and
the replay is timed with CUDA events.  Usage: python scripts/op_probe.py gn|ada|attn_bwd"""
import sys

import torch

sys.path.insert(0, ".")
from adaface_b200 import ops  # noqa: E402


def timed(fn, n=20, reps=5):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (n * reps)


def gn_cases():
    d = "cuda"
    for b, hw, c, c2, act in [(16, 4096, 320, 0, 1), (16, 1024, 640, 0, 1), (16, 1024, 320, 0, 1), (16, 256, 1280, 0, 1), (16, 64, 1280, 0, 1),
                              (16, 64, 2560, 0, 1), (16, 256, 640, 0, 1), (16, 1024, 640, 640, 1), (16, 256, 1280, 1280, 1),
                              (16, 1024, 640, 320, 1), (16, 256, 1920, 0, 1), (16, 1024, 1280, 0, 0), (16, 256, 1280, 0, 0)]:
        h = int(hw ** 0.5)
        ga, be = torch.ones(c + c2, device=d), torch.zeros(c + c2, device=d)
        x = torch.randn(b, c, h, h, device=d).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        if c2:
            x2 = torch.randn(b, c2, h, h, device=d).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            us = timed(lambda: ops.groupnorm_tokens_cat(x, x2, ga, be, 32, 1e-5, act))
        else:
            us = timed(lambda: ops.groupnorm_tokens(x, ga, be, 32, 1e-5, act))
        mb = b * hw * (c + c2) * 4 / 1e6
        print(f"groupnorm b{b} n{hw} C{c}{'+%d' % c2 if c2 else ''} act{act}: {us:7.2f} us  {mb / us * 1e3 / 1e3:6.2f} TB/s (read + write once)")


def ada_cases():
    """The AdaEmbedding chain of one cross-attention layer as the U-Net runs it (8 cond samples): pooling (main + combine
    launches) and the head with the context splice; queries prepared once outside (they are on a side stream in the step)."""
    d = "cuda"
    K, r, b = 9, 12, 8
    for n, c, td in [(4096, 320, 320), (1024, 640, 320), (256, 1280, 320), (64, 1280, 320)]:
        x = torch.randn(b, n, c, device=d).to(torch.bfloat16)
        q = torch.randn(b, n, c, device=d).to(torch.bfloat16)
        pool_q = torch.randn(b, 2, c, device=d)
        te = torch.randn(b, 1280, device=d)
        f = 2 * c + td
        cw, cb = torch.randn(K * r, f, device=d) * f ** -0.5, torch.zeros(K * r, device=d)
        basis, bias = torch.randn(K, r, 768, device=d), torch.zeros(K, 768, device=d)
        ctx = torch.zeros(2 * b, 77, 768, device=d, dtype=torch.bfloat16)
        pos = torch.arange(2, 2 + K, device=d, dtype=torch.int32)

        def pool():
            return ops.ada_pool(x, q, pool_q, None, False)

        def chain():
            fg, bg, _, _ = ops.ada_pool(x, q, pool_q, None, False)
            return ops.ada_head_splice(fg, bg, te, td, cw, cb, basis, bias, ctx, pos)

        t_pool, t_chain = timed(pool), timed(chain)
        print(f"ada chain b{b} n{n} C{c}: pool (main + combine) {t_pool:6.2f} us, pool + head {t_chain:6.2f} us, head alone {t_chain - t_pool:6.2f} us")


def attn_bwd_cases():
    """Self-attention backward of the inversion step (batch 8, SD-1.5 head sizes); AFB_ATTN_BWD=mma2 / simt select the
    earlier versions (read once per process)."""
    import os
    d = "cuda"
    for b, n, c, heads in [(8, 4096, 320, 8), (8, 1024, 640, 8), (8, 256, 1280, 8), (8, 64, 1280, 8)]:
        dh = c // heads
        qkv = (torch.randn(b, n, 3 * c, device=d) * 0.5).to(torch.bfloat16)
        q, k, v = qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]
        out, lse = ops.attention_lse(q, k, v, heads, dh ** -0.5)
        dout = torch.randn(b, n, c, device=d).to(torch.bfloat16)
        us = timed(lambda: ops.attention_bwd(q, k, v, out, lse, dout, heads, dh ** -0.5), n=3, reps=3)
        fl = 10.0 * b * n * n * c
        print(f"attention_bwd[{os.environ.get('AFB_ATTN_BWD', 'mma3')}] b{b} n{n} d{dh}: {us:9.1f} us  {fl / us / 1e6:7.1f} TFLOP/s")
        kc = (torch.randn(b, 77, 2 * c, device=d) * 0.5).to(torch.bfloat16)
        us = timed(lambda: ops.attention_bwd_short(q, kc[..., :c], kc[..., c:], dout, heads, dh ** -0.5, True), n=3, reps=3)
        fl = 10.0 * b * n * 77 * c
        print(f"attention_bwd_short[{os.environ.get('AFB_ATTN_BWD', 'mma3')}] b{b} n{n} k77 d{dh}: {us:9.1f} us  {fl / us / 1e6:7.1f} TFLOP/s")


if __name__ == "__main__":
    with torch.no_grad():
        {"gn": gn_cases, "ada": ada_cases, "attn_bwd": attn_bwd_cases}[sys.argv[1]]()
