"""Dev probe: per-kernel device time of one C3 step (surfel raster + fused SSIM + loss glue) via torch.profiler."""
import sys, types
sys.path.insert(0, ".")
import torch
from torch.profiler import profile, ProfilerActivity
import bench

captured = {}
orig_event = torch.cuda.Event
args = types.SimpleNamespace(steps=3)
# run the leg once to build everything, capturing its step() through a tiny hook: re-create the closure instead
import inspect
src = inspect.getsource(bench.surfel_leg)
src = src.replace("    steps, warm = max(3, min(args.steps, 10)), 3\n", "    captured['step'] = step\n    return {}\n    steps, warm = 3, 3\n")
ns = dict(bench.__dict__); ns["captured"] = captured
exec(src, ns)
ns["surfel_leg"](args, 0)
step = captured["step"]
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
agg = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        a = agg.setdefault(e.name[:90], [0, 0.0]); a[0] += 1; a[1] += e.device_time
tot = sum(v[1] for v in agg.values())
print(f"total device time per step {tot/3/1e3:.3f} ms")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:40]:
    print(f"{v[1]/3:9.1f} us  x{v[0]//3:<3d} {k}")
