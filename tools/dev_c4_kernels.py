"""Dev probe: per-kernel device time and GPU-idle share of one C4 mapping step (bench.mapping_leg) via torch.profiler."""
import sys, types, inspect
sys.path.insert(0, ".")
import torch
from torch.profiler import profile, ProfilerActivity
import bench

captured = {}
src = inspect.getsource(bench.mapping_leg)
src = src.replace("    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)\n", "    captured['step'] = step\n    return {}\n", 1)
ns = dict(bench.__dict__); ns["captured"] = captured
exec(src, ns)
ns["mapping_leg"](types.SimpleNamespace(steps=3), 0, 1, 0)
step = captured["step"]
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start; t1 = max(e.time_range.end for e in ev)
busy = 0; cur = t0
for e in ev:
    s, en = e.time_range.start, e.time_range.end
    busy += max(0, en - max(s, cur)); cur = max(cur, en)
print(f"span per step {(t1-t0)/3/1e3:.3f} ms, busy {busy/3/1e3:.3f} ms, launches per step {len(ev)//3}")
agg = {}
for e in ev:
    a = agg.setdefault(e.name[:80], [0, 0.0]); a[0] += 1; a[1] += e.device_time
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:28]:
    print(f"{v[1]/3:9.1f} us  x{v[0]//3:<3d} {k}")

print("by count:")
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:22]:
    print(f"x{v[0]//3:<4d} {v[1]/3:9.1f} us  {k}")
