"""Inversion-mode collective on the GPUs: FlatGradAllReduce over NCCL, one process per GPU.
torchrun --nproc-per-node N scripts/nccl_grad_allreduce.py  ->  one JSON line from rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adaface_b200.embedding import AdaEmbedding            # noqa: E402
from adaface_b200.sampling import FlatGradAllReduce        # noqa: E402


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    torch.manual_seed(0)
    ada = AdaEmbedding(num_vectors_per_subj_token=9, fg_emb_count=5, bg_emb_count=4, r=12,
                       init_vecs=torch.zeros(2, 768)).cuda()
    params = [p for p in ada.parameters() if p.requires_grad]
    for p in params:
        p.grad = torch.full_like(p, float(rank + 1))
    reducer = FlatGradAllReduce(params)
    reducer()
    want = sum(range(1, world + 1)) / world
    ok = all(bool(torch.all(p.grad == want)) for p in params)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        reducer()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # attached mode: gradients are views of the flat buffer, a call is the all-reduce and the division only
    for p in params:
        p.grad = torch.full_like(p, float(rank + 1))
    reducer.attach()
    reducer()
    ok = ok and all(bool(torch.all(p.grad == want)) for p in params)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        reducer()
    e1.record()
    torch.cuda.synchronize()
    ms2 = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda")
    dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"check": "flat_grad_allreduce", "world": world, "ok": ok, "params": len(params),
                          "floats": reducer.flat.numel(), "ms_per_call_max_over_ranks": round(float(ms), 4),
                          "ms_per_call_attached": round(float(ms2), 4)}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
