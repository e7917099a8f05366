"""Timing of the distributed screened-Poisson solve (multi.DistributedPoisson) under torchrun, against the
single-GPU solve of the same image on rank 0.  torchrun --nproc-per-node N tools/gpu_dist_poisson.py [W H]..."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import gradient_based_path_tracing_b200 as g
from gradient_based_path_tracing_b200 import _capi, multi

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
sc = g.Scene.named("cbox_gdpt_64")
ctx = sc.context(local)
lib, check = _capi.lib(), _capi.check
sizes = [(1920, 1080), (3840, 2160)]
if len(sys.argv) > 2:
    sizes = [(int(sys.argv[i]), int(sys.argv[i + 1])) for i in range(1, len(sys.argv) - 1, 2)]
for w, h in sizes:
    n = w * h * 3
    gen = torch.Generator(device=dev); gen.manual_seed(1)
    full = [torch.randn(n, dtype=torch.float64, device=dev, generator=gen) for _ in range(5)]     # same on every rank
    bands = multi.partition_rows(h, world)
    y0, y1 = bands[rank]
    mine = [torch.zeros_like(f) for f in full]
    for m, f in zip(mine, full):
        m[y0 * w * 3:y1 * w * 3] = f[y0 * w * 3:y1 * w * 3]
    cx, cy, out = (torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(3))
    dp = multi.DistributedPoisson(w, h, bands, dist, torch, multi.CudaStages(ctx, _capi), device=dev)
    times = []
    for it in range(6):
        work = [m.clone() for m in mine]
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        dp.solve(*work, cx, cy, out, alpha=0.04)
        dp.gather_rows([out], dst=0)
        torch.cuda.synchronize(); dist.barrier()
        times.append((time.perf_counter() - t0) * 1e3)
    if rank == 0:
        p = lambda t: C.c_void_p(t.data_ptr())
        fx, fy, fo = (torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(3))
        st = []
        for it in range(4):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            check(lib.gdpt_merge_device(ctx, w, h, p(full[1]), p(full[3]), p(full[2]), p(full[4]), p(fx), p(fy)))
            check(lib.gdpt_poisson_device(ctx, w, h, p(full[0]), p(fx), p(fy), 0.04, p(fo)))
            check(lib.gdpt_synchronize(ctx))
            st.append((time.perf_counter() - t0) * 1e3)
        rel = float(torch.linalg.norm(out - fo) / torch.linalg.norm(fo))
        print("poisson %dx%d: distributed on %d GPUs %.2f ms (min of %s), single GPU %.2f ms, rel-L2 difference %.2e"
              % (w, h, world, min(times[1:]), ["%.1f" % t for t in times], min(st[1:]), rel), flush=True)
dist.barrier()
dist.destroy_process_group()
