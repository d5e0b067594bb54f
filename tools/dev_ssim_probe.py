import sys, torch
sys.path.insert(0, ".")
from pings_b200 import _lib as L
from pings_b200.fused_ssim import fused_ssim
for (H, W) in [(1080, 1920), (370, 1226)]:
    img = torch.rand(1, 3, H, W, device="cuda", requires_grad=True)
    tgt = torch.rand(1, 3, H, W, device="cuda")
    for _ in range(3):
        img.grad = None; fused_ssim(img, tgt).backward()
    torch.cuda.synchronize()
    L.timing_enable(True)
    for _ in range(5):
        img.grad = None; fused_ssim(img, tgt).backward()
    ph = {}
    for n, ms in L.timing_drain(): ph.setdefault(n, []).append(ms)
    L.timing_enable(False)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        img.grad = None; fused_ssim(img, tgt).backward()
    b.record(); torch.cuda.synchronize()
    print(H, W, {k: round(sum(v) / len(v), 4) for k, v in ph.items()}, "module fwd+bwd ms", round(a.elapsed_time(b) / 20, 4))
