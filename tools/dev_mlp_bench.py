"""Dev probe: decoder MLP forward throughput, SIMT vs tcgen05."""
import sys, types
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from pings_b200 import decoder as D
from test_mlp_tc_gpu import _decoder

def bench(dec, x, impl, n=20):
    for _ in range(3):
        D.mlp_forward_impl(dec, x, impl)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        D.mlp_forward_impl(dec, x, impl)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

for (inf, hid, lv, out, N) in [(8, 64, 1, 1, 1 << 20), (32, 64, 2, 5, 1 << 20), (32, 128, 2, 32, 1 << 20), (32, 64, 2, 5, 200000)]:
    dec = _decoder(inf, hid, lv, out, False)
    x = torch.randn(N, inf + 3, device="cuda")
    flop = 2.0 * N * ((inf + 3) * hid + (lv - 1) * hid * hid + hid * out)
    byts = N * (inf + 3 + out) * 4
    for name, impl in (("simt", D.MLP_SIMT), ("tcgen05", D.MLP_TCGEN05)):
        ms = bench(dec, x, impl)
        print(f"in={inf+3} hid={hid} lv={lv} out={out} N={N} {name:8s}: {ms:.3f} ms  {flop/ms/1e9:.2f} TFLOP/s  {byts/ms/1e6:.0f} GB/s")
