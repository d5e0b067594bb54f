"""Dev probe (torchrun): cost of the 28 MB gradient exchange pieces."""
import os, datetime
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=120), device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
P = 500000
parts = [torch.randn(P, k, device=dev) for k in (3, 3, 1, 3, 4)]
flat = torch.zeros(P * 14, device=dev)
work = torch.randn(64 << 20, device=dev)

def timeit(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)

def cat(): torch.cat([p.flatten() for p in parts], out=flat)
def ar(): dist.all_reduce(flat)
def both(): cat(); ar()
def compute(): work.mul_(1.0001)          # ~0.08 ms stand-in kernel
def compute_ar(): compute(); ar()
r = dict(cat=timeit(cat), allreduce=timeit(ar), both=timeit(both), compute=timeit(compute), compute_ar=timeit(compute_ar))
if rank == 0: print(world, {k: round(v, 4) for k, v in r.items()})
dist.destroy_process_group()
