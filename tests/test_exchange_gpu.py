"""pg_multimem_allreduce_f32 (the gradient exchange as one kernel over NVSwitch multicast) against
ncclAllReduce on the same random data: the sums agree to fp32 reassociation, replicas hold identical bits, two calls in
a row work (the signal pads return to zero).  No counterpart in the reference (single-GPU).
Runs tools/dev_multimem.py under torchrun on min(2, visible GPUs) ranks."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_multimem_allreduce_matches_nccl():
    n = min(2, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29613", os.path.join(ROOT, "tools", "dev_multimem.py"), "--quick"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    if p.returncode != 0 and "no multicast support" in (p.stdout + p.stderr):
        pytest.skip("GPUs without NVSwitch multicast: the exchange falls back to ncclAllReduce (FlatGradAllReduce)")
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("{")][-1]
    d = json.loads(line)
    assert d["world"] == n
    for k in ("c4_4MB", "c2_28MB"):
        assert d[k]["replicas_identical"] and d[k]["max_rel_err_vs_nccl"] < 1e-6
