"""Concurrent host<->device copy bandwidth of all ranks (torchrun): tells whether the e2e path of a multi-GPU run is
bounded by the box's PCIe / host-memory fabric rather than by the kernels."""
import os, time
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
    del os.environ["NCCL_DEBUG"]
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 256 << 20
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.zero_()
h2 = torch.empty(n // 4, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("h2d", "h2d+d2h"):
    for _ in range(2):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(10):
        with torch.cuda.stream(s1):
            d.copy_(h, non_blocking=True)
        if mode != "h2d":
            with torch.cuda.stream(s2):
                h2.copy_(d[: n // 4], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    bw = torch.tensor([10 * n / dt / 1e9], device="cuda")
    if world > 1:
        lo = bw.clone(); dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(bw, op=dist.ReduceOp.SUM)
    else:
        lo = bw
    if rank == 0:
        print(f"{mode}: {world} rank(s), H2D aggregate {bw.item():.1f} GB/s, slowest rank {lo.item():.1f} GB/s", flush=True)
