"""tcgen05/TMEM decoder MLP (pg_mlp_tc.cuh) against a plain torch fp32 reference of the same op
(DEC = model/decoder.py:62-98) and against the SIMT kernel.  Tolerance: 3xTF32 split accumulation is
fp32-level, so |y - ref| <= 2e-5 * max|ref| + 1e-6 (the SIMT kernel meets the same bar)."""
import types

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _decoder(in_feat, hidden, levels, out_dim, leaky, bias=True, seed=0):
    from pings_b200.decoder import Decoder
    cfg = types.SimpleNamespace(mlp_leaky_relu=leaky, infer_bs=4096, mlp_bias_on=bias, main_loss_type="bce",
                                logistic_gaussian_ratio=0.55, sigma_sigmoid_m=0.1, device="cuda")
    torch.manual_seed(seed)
    dec = Decoder(cfg, in_feat, hidden, levels, out_dim)
    with torch.no_grad():
        for lin in dec._linears():
            lin.weight.mul_(1.7)
            if lin.bias is not None:
                lin.bias.uniform_(-0.3, 0.3)
    return dec


def _torch_ref(dec, x):
    h = x
    hidden = []
    for lin in dec.layers:
        h = F.linear(h, lin.weight, lin.bias)
        h = F.leaky_relu(h, 0.01) if dec.use_leaky_relu else F.relu(h)
        hidden.append(h)
    return F.linear(h, dec.lout.weight, dec.lout.bias), hidden


CASES = [  # in_feat (+3 pos), hidden, levels, out, leaky, bias, N
    (8, 64, 1, 1, False, True, 5000),      # C1 SDF decoder
    (32, 64, 2, 5, False, True, 20000),    # spawn-sized heads
    (32, 128, 2, 32, True, True, 9001),    # two weight pieces per hidden layer
    (16, 40, 1, 3, True, False, 129),      # hidden not a multiple of 16, no bias, ragged tile
    (5, 24, 3, 64, False, True, 1),        # single row, widest supported head
]


@pytest.mark.parametrize("case", CASES)
def test_tcgen05_mlp_matches_fp32_reference(case):
    from pings_b200 import decoder as D
    in_feat, hidden, levels, out_dim, leaky, bias, N = case
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        dec = _decoder(in_feat, hidden, levels, out_dim, leaky, bias)
        g = torch.Generator(device="cuda").manual_seed(N)
        x = torch.randn(N, in_feat + 3, device="cuda", generator=g) * 2.0
        with torch.no_grad():
            ref, ref_h = _torch_ref(dec, x)
        y_tc, h_tc = D.mlp_forward_impl(dec, x, D.MLP_TCGEN05, save_hidden=True)
        y_si, h_si = D.mlp_forward_impl(dec, x, D.MLP_SIMT, save_hidden=True)
        torch.cuda.synchronize()
        tol = 2e-5 * float(ref.abs().max()) + 1e-6
        assert float((y_tc - ref).abs().max()) <= tol, ("tcgen05 vs torch", float((y_tc - ref).abs().max()), tol)
        assert float((y_si - ref).abs().max()) <= tol
        for l in range(levels):
            tl = 2e-5 * float(ref_h[l].abs().max()) + 1e-6
            assert float((h_tc[l] - ref_h[l]).abs().max()) <= tl
            assert float((h_si[l] - ref_h[l]).abs().max()) <= tl
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def test_auto_dispatch_and_autograd_use_tensor_path_for_large_batches():
    """mlp_batch-sized calls go through the tensor-core forward and still back-propagate (SIMT backward)."""
    from pings_b200 import _lib as L
    dec = _decoder(32, 64, 2, 4, False)
    x = torch.randn(8192, 35, device="cuda", requires_grad=True)
    L.timing_enable(True)
    y = dec.mlp_batch(x)
    names = [n for n, _ in L.timing_drain()]
    L.timing_enable(False)
    assert "mlp_forward_tc" in names
    y.square().sum().backward()
    xr = x.detach().clone().requires_grad_(True)
    ref, _ = _torch_ref(dec, xr)
    ref.square().sum().backward()
    assert float((x.grad - xr.grad).abs().max()) <= 1e-4 * float(xr.grad.abs().max()) + 1e-6
