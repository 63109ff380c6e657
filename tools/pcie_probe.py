"""Pinned H2D, D2H and simultaneous H2D+D2H copy bandwidth with ALL ranks copying at once: the ceiling of
bench.py's e2e figure at N GPUs.  Plain page-locked cudaMemcpyAsync traffic, one copy per direction per repetition.

    python tools/pcie_probe.py                                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py

Prints one JSON line (rank 0): per-GPU and aggregate GB/s for each pattern, and the e2e ceiling in samples/s
(a round-trip sample costs 4 bytes each way)."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 30
h1 = torch.empty(n, dtype=torch.uint8).pin_memory()
h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda")
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def h2d():
    with torch.cuda.stream(s1):
        d1.copy_(h1, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


def both():
    h2d()
    d2h()


def t(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item())


res = {k: n / t(f) / 1e9 for k, f in (("h2d", h2d), ("d2h", d2h), ("both_per_direction", both))}
if rank == 0:
    print(json.dumps({"n_gpus": world, "GBps_per_gpu": res,
                      "aggregate_GBps": {k: v * world * (2 if k.startswith("both") else 1) for k, v in res.items()},
                      "e2e_ceiling_Msamples_per_s_aggregate": res["both_per_direction"] / 4 * world * 1e3,
                      "how": "1 GiB page-locked buffers per rank, plain cudaMemcpyAsync, all ranks at once, max over ranks"}))
if world > 1:
    dist.destroy_process_group()
