"""Data-parallel plumbing for the hot path (one process per GPU, torch.distributed).

The reference is single-GPU; this is the new design of DESIGN.md §6:
views and SDF query rows are independent units and are sharded with no data-path collective;
training adds exactly one exchange step per iteration, the sum of the parameter gradients
(Gaussian parameters / neural-point features / decoder weights), done as ONE all-reduce over a
flat fp32 buffer (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Iterable, Sequence

import torch
import torch.distributed as dist


def world() -> tuple[int, int]:
    """(rank, world_size); (0, 1) when no process group is initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_views(views: Sequence, rank: int, world_size: int) -> list:
    """Round-robin view sharding: rank r renders views[r::world] (a batch of V views costs
    ceil(V / world) renders per rank; cameras of a trajectory are interleaved so that the ranks'
    workloads have the same depth complexity)."""
    return list(views[rank::world_size])


def shard_rows(n: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous split of n query rows: [lo, hi) of this rank (sizes differ by at most 1)."""
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class FlatGradAllReduce:
    """Sum (or average) the gradients of `params` over all ranks with a single collective.

    Gradients are packed into one persistent flat buffer (so NCCL sees one large message: the
    cost on NVSwitch is launch latency, not link count), all-reduced, and written back into
    `p.grad`.  Parameters whose grad is None on this rank (e.g. neural points not visible in this
    rank's views) contribute zeros.  With `side_stream=True` on CUDA the collective runs on its own
    stream and `wait()` joins it, so that it overlaps independent backward work (the SDF branch).
    """

    def __init__(self, params: Iterable[torch.Tensor], average: bool = True, group=None, side_stream: bool = False,
                 multimem: bool = False):
        """multimem=True (CUDA, world > 1; a collective call): the flat buffer is allocated in symmetric memory and the
        exchange is `MultimemGradAllReduce`'s kernel (in-switch reduction over NVSwitch) instead of ncclAllReduce; when
        the system has no multicast support on any rank the buffer silently stays a plain tensor reduced by NCCL
        (`self.backend` says which)."""
        self.params = [p for p in params]
        if not self.params:
            raise ValueError("FlatGradAllReduce: no parameters")
        dev, dt = self.params[0].device, self.params[0].dtype
        for p in self.params:
            if p.device != dev or p.dtype != dt:
                raise ValueError("FlatGradAllReduce: all parameters must share one device and dtype "
                                 f"(got {p.device}/{p.dtype} next to {dev}/{dt})")
        self.sizes = [p.numel() for p in self.params]
        self._mm = None
        if multimem and dev.type == "cuda" and dt == torch.float32 and world()[1] > 1:
            try:
                self._mm = MultimemGradAllReduce(sum(self.sizes), dev, group)
            except Exception:   # noqa: BLE001
                self._mm = None
            ok = torch.tensor([1 if self._mm is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok) == 0:
                self._mm = None
        self.backend = "multimem" if self._mm is not None else "collective"
        self.flat = self._mm.flat if self._mm is not None else torch.zeros(sum(self.sizes), dtype=dt, device=dev)
        self.views = list(self.flat.split(self.sizes))
        self.average = average
        self.group = group
        # high priority: the exchange's few CTAs get SM slots ahead of the compute kernels they overlap
        self.stream = torch.cuda.Stream(device=dev, priority=-1) if (side_stream and dev.type == "cuda") else None
        self._work = None

    def attach(self) -> None:
        """Make every `p.grad` a view of the flat buffer and zero it: autograd then accumulates straight into
        the buffer (AccumulateGrad adds in place into an existing .grad), so `launch()` / `wait()` need no
        pack / scatter copies (2 copies per parameter per step otherwise).  Call once per step INSTEAD of
        setting the grads to None; a caller that replaces `p.grad` afterwards simply falls back to the copies.

        Not combinable with side_stream=True: the collective would then reduce the very buffer that any backward
        work still running after launch() accumulates into (a data race), so that combination raises."""
        if self.stream is not None:
            raise RuntimeError("FlatGradAllReduce.attach() cannot be used with side_stream=True: gradients would "
                               "accumulate into the buffer while it is being reduced")
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v.view(p.shape)

    def _attached(self, p: torch.Tensor, v: torch.Tensor) -> bool:
        g = p.grad
        return g is not None and g.data_ptr() == v.data_ptr() and g.numel() == v.numel() and g.is_contiguous()

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * self.flat.element_size()

    def pack(self) -> None:
        for p, v in zip(self.params, self.views):
            if self._attached(p, v):
                continue
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad.reshape(-1))

    def launch(self) -> None:
        """pack + start the all-reduce (asynchronous when a side stream is used)."""
        rank, ws = world()
        self.pack()
        if ws == 1:
            return
        if self.stream is not None:
            self.stream.wait_stream(torch.cuda.current_stream(self.flat.device))
            with torch.cuda.stream(self.stream):
                self._reduce(ws)
        else:
            self._reduce(ws)

    def _reduce(self, ws: int) -> None:
        if self._mm is not None:
            self._mm.all_reduce(scale=(1.0 / ws) if self.average else 1.0)   # the average is folded into the kernel
            return
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        if self.average:
            self.flat.div_(ws)

    def wait(self) -> None:
        """Join the collective and scatter the reduced gradients back into p.grad."""
        if self.stream is not None:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.stream)
        for p, v in zip(self.params, self.views):
            if self._attached(p, v):
                continue
            if p.grad is None:
                p.grad = v.reshape(p.shape).clone()
            else:
                p.grad.copy_(v.reshape(p.shape))

    def __call__(self) -> None:
        self.launch()
        self.wait()


class MultimemGradAllReduce:
    """The gradient exchange as ONE kernel over NVSwitch multicast instead of ncclAllReduce.

    `flat` is a flat fp32 buffer in symmetric memory (torch.distributed._symmetric_memory does the plumbing: the
    allocation, the handle exchange, the multicast mapping and the signal pads); producers write their gradients
    straight into it (`param_grads_out=` of the rasterizer backward, FlatGradAllReduce.attach-style views) and
    `all_reduce()` launches `pg_multimem_allreduce_f32` on the current stream: barrier, in-switch reduction of this
    rank's slice (multimem.ld_reduce), broadcast (multimem.st), barrier.  Replicas receive identical bits.

    Raises when the GPUs have no multicast support (no NVSwitch): the caller then uses FlatGradAllReduce (NCCL).
    The constructor is a collective (rendezvous + a self-check against a known sum)."""

    def __init__(self, numel: int, device, group=None, max_ctas: int = 0):
        import torch.distributed._symmetric_memory as symm
        from . import _lib as L
        self._L = L
        group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.numel = int(numel)
        padded = (self.numel + 3) // 4 * 4
        self.device = torch.device(device)
        self._store = symm.empty(padded, dtype=torch.float32, device=self.device)
        self._hdl = symm.rendezvous(self._store, group)
        mc = int(self._hdl.multicast_ptr)
        if mc == 0:
            raise RuntimeError("MultimemGradAllReduce: no multicast support on this system (NVSwitch required)")
        # the handle describes the allocation block: this tensor may sit at an offset inside it
        self._mc = mc + (self._store.data_ptr() - int(self._hdl.buffer_ptrs[self.rank]))
        self._pads = int(self._hdl.signal_pad_ptrs_dev)
        self._pad_bytes = int(type(self._hdl).signal_pad_size)
        self.max_ctas = int(max_ctas)
        self.flat = self._store[:self.numel]
        self._self_check()

    def all_reduce(self, scale: float = 1.0) -> None:
        """In place on `flat`, on torch's current stream of the buffer's device."""
        L = self._L
        with L.device_guard(self.device):
            L.check(L.lib().pg_multimem_allreduce_f32(self._mc, self._pads, self._pad_bytes, self._store.numel(), self.rank,
                                                      self.world, float(scale), self.max_ctas, L.raw_stream(self.device)),
                    "multimem all-reduce")

    def _self_check(self) -> None:
        n = self._store.numel()
        idx = torch.arange(n, device=self.device, dtype=torch.float32)
        self._store.copy_((idx % 251.0) * float(self.rank + 1))
        self.all_reduce()
        want = (idx % 251.0) * float(self.world * (self.world + 1) // 2)
        ok = bool(torch.equal(self._store, want))
        self._store.zero_()
        torch.cuda.synchronize(self.device)
        if not ok:
            raise RuntimeError("MultimemGradAllReduce: self-check failed (multicast mapping does not address this buffer)")


class FlatAdamW:
    """The replicated-parameter update of a data-parallel step as ONE fused kernel (SURVEY.md §8e names it the serial
    part of the C4 step; the reference builds a per-tensor AdamW in utils/tools.py:142 setup_optimizer).

    The parameters are re-pointed at consecutive slices of one flat buffer (`p.data` becomes a view) that is itself
    a leaf parameter whose `.grad` is the flat exchange buffer of a FlatGradAllReduce (attach() mode), so
    `torch.optim.AdamW(fused=True)` sees a single tensor: one launch per step instead of one multi-tensor launch
    over ~30 tensors plus their bookkeeping."""

    def __init__(self, exchange: "FlatGradAllReduce", lr: float = 1e-3, weight_decay: float = 0.0, betas=(0.9, 0.999),
                 eps: float = 1e-15):
        self.exchange = exchange
        flat = torch.empty_like(exchange.flat)
        off = 0
        for p, n in zip(exchange.params, exchange.sizes):
            flat[off:off + n].copy_(p.data.reshape(-1))
            p.data = flat[off:off + n].view(p.shape)
            off += n
        self.flat = torch.nn.Parameter(flat)
        self.flat.grad = exchange.flat
        fused = flat.is_cuda
        self.opt = torch.optim.AdamW([self.flat], lr=lr, weight_decay=weight_decay, betas=betas, eps=eps, fused=fused)

    def step(self) -> None:
        self.flat.grad = self.exchange.flat
        self.opt.step()


def replicas_in_sync(tensors: Iterable[torch.Tensor], group=None) -> bool:
    """Debug check of DESIGN.md §6: replicated parameters must be bit-identical on every rank after a
    step.  Compares an int64 checksum of the raw bits (all-gather of one scalar per rank)."""
    rank, ws = world()
    acc = None
    for t in tensors:
        bits = t.detach().contiguous().view(torch.uint8).to(torch.int64)
        w = torch.arange(1, bits.numel() + 1, dtype=torch.int64, device=bits.device)
        s = (bits.flatten() * (w % 65521)).sum()
        acc = s if acc is None else acc * 31 + s
    if ws == 1:
        return True
    gathered = [torch.zeros_like(acc) for _ in range(ws)]
    dist.all_gather(gathered, acc, group=group)
    return all(bool(g == gathered[0]) for g in gathered)
