"""Multi-GPU probe / test driver of the multimem gradient exchange (run under torchrun, one rank per GPU):
correctness against ncclAllReduce on random data (replicas identical, sum within fp32 reassociation) and device time
of both on the C2 (28 MB) and C4 (4.2 MB) exchange sizes.  Prints one JSON line on rank 0.
usage: python -m torch.distributed.run --nproc-per-node N tools/dev_multimem.py [--quick]"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pings_b200.parallel import MultimemGradAllReduce   # noqa: E402


def main():
    quick = "--quick" in sys.argv
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    for name, numel in (("c4_4MB", 1060353), ("c2_28MB", 7000000)):
        ex = MultimemGradAllReduce(numel, dev)            # collective; runs the self-check
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        x = torch.randn(numel, device=dev, generator=g)
        ref = x.clone()
        dist.all_reduce(ref)
        for rep in range(2):                              # twice in a row: the signal pads must return to zero
            ex.flat.copy_(x)
            ex.all_reduce()
            torch.cuda.synchronize()
            err = float((ex.flat - ref).abs().max() / ref.abs().max())
            assert err < 1e-6, (name, rep, err)
        # replicas identical: compare a checksum of the bits over ranks
        bits = ex.flat.view(torch.int32).to(torch.int64).sum().view(1)
        lo, hi = bits.clone(), bits.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert int(lo) == int(hi), "replicas differ"
        res = {"max_rel_err_vs_nccl": err, "replicas_identical": True}
        if not quick:
            def timed(fn, n=30):
                for _ in range(5):
                    fn()
                dist.barrier()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(n):
                    fn()
                b.record()
                torch.cuda.synchronize()
                t = torch.tensor([a.elapsed_time(b) / n], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return round(float(t), 4)
            y = x.clone()
            res["nccl_ms"] = timed(lambda: dist.all_reduce(y))
            for ctas in (8, 16, 32, 64):
                ex.max_ctas = ctas
                res[f"multimem_ms_{ctas}ctas"] = timed(ex.all_reduce)
            res["bytes"] = numel * 4
        out[name] = res
        del ex
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
