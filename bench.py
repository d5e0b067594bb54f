#!/usr/bin/env python
"""bench.py — headline benchmark of the pings_b200 hot path (contract in the task statement).

Workload at every N (BASELINE.json configs[1]): 3DGS rasterizer forward +
backward with camera-pose gradients, 500,000 synthetic Gaussians, one 1226x370 KITTI-shaped
view per step per GPU.  metric = raster fwd+bwd Mpix/s (whole job).  A "step" = one view:
preprocess -> depth sort -> binning -> blend -> blend backward -> per-Gaussian backward
(+ NCCL all-reduce of the Gaussian-parameter gradients when N > 1, the training exchange step).

  value : inputs resident in HBM, CUDA events around each step (L2 flushed between steps).
  e2e   : the same step through the reference-facing API (GaussianRasterizer autograd module)
          with HOST (pinned) Gaussian parameters: H2D copies, forward, L1+depth loss, backward,
          D2H of loss + pose gradient inside the timed region.
  roofline : dominant kernel (by device time, measured with CUDA events on the launching stream
          via the library's phase timer), algorithmic bytes of SURVEY.md §8d / DESIGN.md.
  cpu_baseline : oracle/ CPU restatement timed on this box's host cores on a bounded sample.
  extra.sdf : the second half of the metric (SDF queries/s, config C1) when that path is built.

`--impl reference` times the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libref_r3.so, built
from /root/reference's own sources) on the same scene; the reference has no CPU rasterizer, so that
is its only implementation of this path.  If that library is absent the
arm falls back to the CPU restatement.
"""
from __future__ import annotations

import argparse
from types import SimpleNamespace
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

C2 = dict(P=500_000, W=1226, H=370, seed=0)


# ----------------------------------------------------------------------------------------------
def peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=float(d["hbm_gbs"]), bf16_tflops=float(d.get("bf16_tflops", 2250.0)),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(smax) if smax else None,
                    samples=len(sm), reasons=sorted(reasons))


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        torch.cuda.set_device(local)
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=180))
    else:
        torch.cuda.set_device(0)
    return rank, world, local


def barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


# ----------------------------------------------------------------------------------------------
def make_scene(rank: int):
    """C2 scene; every rank sees the same Gaussians from its own seeded pose (view sharding)."""
    from pings_b200 import synth
    cam0 = synth.kitti_camera(C2["seed"], C2["W"], C2["H"])
    sc = synth.gaussians_in_view(cam0, C2["P"], C2["seed"])
    cam = cam0
    if rank > 0:  # other ranks: nearby poses 0.1 m apart along the optical axis, so that every rank has the
        # same amount of work within ~1 % (weak scaling then shows the exchange cost, not view-to-view variation)
        T = cam0.T_wc.clone()
        T[:3, 3] += T[:3, 2] * (0.1 * rank)
        cam = synth.Camera(cam0.W, cam0.H, cam0.fx, cam0.fy, cam0.cx, cam0.cy, T)
    ct = synth.cotangents(cam, C2["seed"])
    g = torch.Generator().manual_seed(77)
    target = torch.rand(3, C2["H"], C2["W"], generator=g)
    return cam, sc, ct, target


def algorithmic_bytes(P: int, R: int, W: int, H: int, Q: int = -1, surfel: bool = False) -> dict:
    """Per-launch algorithmic HBM bytes of every kernel of a step.  Keys are the
    phase names the library's phase timer reports (one per kernel for the band binning)."""
    WH = W * H
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    NC = min(4096, max(148, (4 << 20) // T))   # bin_max_chunks (pg_raster_common.cuh)
    nL = ((gx + 31) // 32) * gy                # (row, band) lists
    nch = min(4096, max(256, ((1 << 20) // max(nL, 1)) & ~7))   # band_row_chunks
    out = {
        "preprocess_fwd": (160 if surfel else 104) * P,      # 44 B in + 60 B out per Gaussian (surfel: + normal, zx, zy)
        # bucket sort (band path): histogram reads the keys (4), scatter reads them again and writes (key, id) pairs
        # (4 + 8), the local sort reads the pairs, gathers the rects and writes ids + rects (8 + 8 + 12); per Gaussian,
        # counting every Gaussian as visible.  (The 4-pass LSD path moves 112 B per Gaussian.)
        "depth_sort": (44 if os.environ.get("PG_SORT_IMPL", "b")[0] != "r" else 112) * P,
        # blend kernels: HBM sees the id list once, each record once (re-reads by other tiles hit L2) and the
        # images; the per-instance record / gradient traffic (l2_bytes below) is L2 and shared-memory work
        "render_fwd": 4 * R + (68 if surfel else 48) * P + (44 if surfel else 28) * WH,
        "render_bwd": 4 * R + (68 if surfel else 48) * P + 2 * (64 if surfel else 48) * P + (68 if surfel else 52) * WH,
        "preprocess_bwd": (215 if surfel else 190) * P,
    }
    if Q >= 0:   # band binning, Q band records, S segments of 128 records
        S = Q // 128 + nL
        out.update({
            "band_rows_kernel<count>": 8 * P + 4 * nL * nch,                 # depth-ordered rects in, count matrix out
            "band_list_scan_kernel": 2 * 4 * nL * nch,
            "band_rows_kernel<write>": 8 * P + 4 * P + 4 * nL * nch + 4 * Q,  # rect + id in, one 4-B record per (Gaussian, row, band)
            "band_expand_kernel<count>": 4 * Q + 8 * Q + 4 * 32 * S,          # record + its rect (gather) in, per-tile segment counts out
            "band_tile_scan_kernel": 2 * 4 * 32 * S + 8 * T,
            "band_expand_kernel<write>": 4 * Q + 8 * Q + 4 * 32 * S + 4 * R,  # + one 4-B id per instance out
        })
        out["binning"] = sum(v for k, v in out.items() if k.startswith("band_"))
    else:
        out["binning"] = 2 * 12 * P + 4 * R + 16 * T * NC + 8 * T
    return out


def l2_bytes(P: int, R: int) -> dict:
    """Per-launch L2-level traffic of the blend kernels: 4-B id + 48-B record per (tile, Gaussian) instance,
    plus the 48-B gradient record accumulated with atomics per instance in the backward."""
    return {"render_fwd": 52 * R, "render_bwd": 52 * R + 48 * R}


BAND_LISTS: dict = {}   # {"n": (row, band) lists of the benchmarked tile grid}


def roofline_tables(phase_ms: dict, ab: dict, pk: dict, l2: dict | None = None):
    """-> (per-kernel table, per-phase table with the band kernels folded into `binning`, name of the longest kernel)."""
    def row(ms, b):
        gbs = b / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        return dict(ms=round(ms, 4), alg_bytes=int(b), gbs=round(gbs, 1), frac=round(gbs / pk["hbm_gbs"], 4))
    per_kernel = {k: row(ms, ab.get(k, 0)) for k, ms in phase_ms.items()}
    folded: dict = {}
    for k, ms in phase_ms.items():
        kk = "binning" if k.startswith("band_") else k
        folded[kk] = folded.get(kk, 0.0) + ms
    per_phase = {k: row(ms, ab.get(k, 0)) for k, ms in folded.items()}
    for k, b in (l2 or {}).items():
        if k in per_phase:
            per_phase[k]["l2_bytes"] = b
            per_phase[k]["l2_gbs"] = round(b / (per_phase[k]["ms"] * 1e-3) / 1e9, 1) if per_phase[k]["ms"] > 0 else 0.0
    dom = max(phase_ms, key=phase_ms.get)
    return per_kernel, per_phase, dom


# Issue-slot view of the blend kernels (they are not HBM-bound: DRAM < 1 % in ncu): warp instructions executed per
# launch on the C2 scene from the committed ncu --set full capture (profiles/r02_ncu_full_c2.csv,
# smsp__inst_executed.sum; a property of the scene, not of the clock) over the kernel's LIVE duration, against the
# B200 issue peak of 148 SMs x 4 schedulers x 1 warp-instruction / clock at the SM clock sampled during the run.
ISSUE_PEAK_GINST = 148 * 4 * 1.965      # G warp-instructions/s: 148 SMs x 4 schedulers at the 1.965 GHz maximum SM clock
ISSUE_INST: dict = {"render_fwd": 76318981, "render_bwd": 92405504, "preprocess_fwd": 8841762,
                    "preprocess_bwd": 14840160, "depth_sort": 4850044}   # depth_sort = the four bucket-sort kernels


def issue_roofline(phase_ms: dict, sm_mhz: float, pairs: int) -> dict:
    """Per kernel: G warp-instructions/s achieved vs the issue peak; for the blend kernels also pixel-Gaussian
    pairs/s (pairs = sum over pixels of the list position of the last contributor, i.e. what the walk visits) and
    thread-instructions per pair."""
    peak = 148 * 4 * (sm_mhz or 1965.0) * 1e-3
    out = {"peak_ginst_s": round(peak, 1), "source": "profiles/r02_ncu_full_c2.csv (smsp__inst_executed.sum), live durations"}
    for k, inst in ISSUE_INST.items():
        ms = phase_ms.get(k, 0.0)
        if ms <= 0:
            continue
        g = inst / (ms * 1e-3) / 1e9
        row = {"warp_inst_per_launch": inst, "ginst_s": round(g, 1), "frac": round(g / peak, 4)}
        if k.startswith("render_") and pairs > 0:
            row["pairs_per_launch"] = int(pairs)
            row["gpairs_s"] = round(pairs / (ms * 1e-3) / 1e9, 2)
            row["thread_inst_per_pair"] = round(inst * 32 / pairs, 1)
        out[k] = row
    return out


def launches_per_step(T: int, P: int) -> int:
    """Kernel launches of one fwd+bwd (radix pass = hist + 3 scan + scatter)."""
    scan = 3
    rp = 1 + scan + 1
    band = BAND_LISTS.get("n", 1 << 30) <= 1024
    if band and os.environ.get("PG_SORT_IMPL", "b")[0] != "r":
        pre = 1 + 4                      # preprocess + bucket sort: histogram, scan, scatter, local sort
    elif (P + 4095) // 4096 <= 296:      # one cooperative launch: 4 LSD passes + rect gather + weight scan
        pre = 1 + 1
    else:
        pre = 1 + 4 * rp + 1 + scan      # preprocess, depth argsort, rect gather, weight scan
    if BAND_LISTS.get("n", 1 << 30) <= 1024:
        binning = 6                      # band binning: rows x2, list scan, expand x2, tile scan
    elif T <= 12288:                     # one-level counting sort: count walk, matrix scan, ranges, scatter walk
        binning = 1 + scan + 1 + 1
    else:                                # radix fallback: emit, 2 tile passes, ranges
        binning = 1 + 2 * rp + 1
    return pre + binning + 1 + 1 + 1     # + blend fwd, blend bwd, per-Gaussian bwd


def run_ours(args, rank, world, local):
    from pings_b200 import _lib as L, _raster_c as RC
    from pings_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda", local)
    cam, sc_cpu, ct_cpu, target_cpu = make_scene(rank)
    P, W, H = C2["P"], C2["W"], C2["H"]
    sc = {k: v.to(dev) for k, v in sc_cpu.items()}
    ct = {k: v.to(dev) for k, v in ct_cpu.items()}
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    bg = torch.ones(3, device=dev)
    e = torch.empty(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    # two flat gradient buffers: the exchange of step k (side stream) overlaps the compute of step k+1
    # exchange buffers (double-buffered: the all-reduce of step k overlaps step k+1).  Default at N > 1: symmetric
    # memory + our multimem kernel (in-switch reduction); PG_EXCHANGE=nccl keeps ncclAllReduce (also the fallback when
    # the GPUs have no multicast support).
    mm = None
    if world > 1 and os.environ.get("PG_EXCHANGE", "multimem") != "nccl":
        try:
            from pings_b200.parallel import MultimemGradAllReduce
            mm = [MultimemGradAllReduce(P * 14, dev, max_ctas=int(os.environ.get("PG_EXCHANGE_CTAS", "0"))) for _ in range(2)]
        except Exception as ex_:   # noqa: BLE001  (no NVSwitch multicast: say so and use NCCL)
            print(f"[bench] multimem exchange unavailable ({type(ex_).__name__}: {ex_}); using NCCL", file=sys.stderr)
            mm = None
        ok_ = torch.tensor([1 if mm is not None else 0], device=dev)
        dist.all_reduce(ok_, op=dist.ReduceOp.MIN)
        if int(ok_) == 0:
            mm = None
    grads = [m.flat for m in mm] if mm is not None else [torch.zeros(P * 14, device=dev) for _ in range(2)]
    grads_flat = grads[0]

    def exchange_grads(k: int):
        if mm is not None:
            mm[k].all_reduce()
        else:
            dist.all_reduce(grads[k])
    side = torch.cuda.Stream(device=dev, priority=int(os.environ.get("PG_EXCHANGE_PRIO", "-1"))) if world > 1 else None
    step_no = [0]
    exch_ev: list = []

    last_binning: list = []

    def step_resident(exchange: bool = True):
        flat = grads[step_no[0] & 1]
        step_no[0] += 1
        R, color, radii, geom, binning, img, depth, opac, n_touched = RC.rasterize_gaussians(
            bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full,
            proj, cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, False)
        out = RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"],
                                              sc["rotations"], 1.0, e, view, full, proj, cam.tanfovx, cam.tanfovy,
                                              ct["color"], ct["depth"], e, 0, center, geom, R, binning, img, False,
                                              param_grads_out=flat)
        last_binning[:] = [binning, geom, img, RC.last_binning_capacity()]
        if world > 1 and exchange:
            # training exchange step: sum the Gaussian-parameter gradients over the view shards (the backward wrote
            # them straight into the flat buffer).  The all-reduce runs on a side stream and overlaps the next
            # view's forward/backward (views of a batch accumulate; the optimizer waits for all of them): at most
            # one exchange is in flight - this step first joins the previous one, inside its own timed region.
            cur = torch.cuda.current_stream(dev)
            cur.wait_stream(side)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ea.record(side)
                exchange_grads((step_no[0] - 1) & 1)
                eb.record(side)
                exch_ev.append((ea, eb))
        return R

    # ---- warm-up + timed region (device-resident) ----
    for _ in range(args.warmup):
        R = step_resident()
    if side is not None:
        torch.cuda.current_stream(dev).wait_stream(side)
    barrier(world)
    exch_ev.clear()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        R = step_resident()
        b.record()
        evs.append((a, b))
    if side is not None:   # the last exchange has nothing left to hide behind: its exposed part is timed as a tail
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        torch.cuda.current_stream(dev).wait_stream(side)
        b.record()
        evs.append((a, b))
    barrier(world)
    clocks = sampler.stop() if rank == 0 else None
    exchange_ms = (sum(x.elapsed_time(y) for x, y in exch_ev) / max(len(exch_ev), 1)) if exch_ev else 0.0
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * W * H / (ms_per_step * 1e-3) / 1e6

    # ---- e2e through the public API with host buffers ----
    # the host copy of the Gaussian parameters is ONE flat pinned buffer with the five tensors as views (the layout
    # pings_b200.parallel.FlatAdamW keeps), so a step's upload is a single 28 MB H2D copy
    keys = list(sc_cpu)
    sizes = [sc_cpu[k].numel() for k in keys]
    pinned_flat = torch.empty(sum(sizes), dtype=torch.float32).pin_memory()
    pinned, off = {}, 0
    for k, n_ in zip(keys, sizes):
        pinned[k] = pinned_flat[off:off + n_].view(sc_cpu[k].shape)
        pinned[k].copy_(sc_cpu[k])
        off += n_
    target = target_cpu.to(dev)
    target_depth = torch.rand(1, H, W, generator=torch.Generator().manual_seed(78)).mul_(60.0).add_(1.0).to(dev)
    from pings_b200.losses import photo_depth_loss
    h2d_bytes = sum(v.numel() * v.element_size() for v in pinned.values())
    host_out = [torch.empty(7, dtype=torch.float32).pin_memory() for _ in range(2)]
    out_done = [torch.cuda.Event() for _ in range(2)]
    losses: list = []
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, full, proj, 0, center, False, False)
    rast = GaussianRasterizer(rs)

    # Double-buffered input staging: the H2D copy of step k+1 runs on a copy stream while step k computes
    # (every copy is inside the timed region; only the first one is exposed).
    copy_stream = torch.cuda.Stream(device=dev)
    slots_flat = [torch.empty_like(pinned_flat, device=dev) for _ in range(2)]
    slots = []
    for s_ in range(2):
        d_, off = {}, 0
        for k, n_ in zip(keys, sizes):
            d_[k] = slots_flat[s_][off:off + n_].view(sc_cpu[k].shape)
            off += n_
        slots.append(d_)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    h2d_ev: list = []

    def upload(i: int):
        slot = i & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])
            ta, tb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ta.record(copy_stream)
            slots_flat[slot].copy_(pinned_flat, non_blocking=True)
            tb.record(copy_stream)
            h2d_ev.append((ta, tb))
            ready[slot].record(copy_stream)

    def step_e2e(i: int, prefetch: bool, more: bool = True):
        cur = torch.cuda.current_stream()
        if prefetch:
            if more:
                upload(i + 1)                  # overlaps this step's kernels
        else:
            upload(i)                          # serial variant: copy, then compute
        slot = i & 1
        cur.wait_event(ready[slot])
        leaves = {k: slots[slot][k].detach().requires_grad_(True) for k in pinned}
        theta = torch.zeros(1, 3, device=dev, requires_grad=True)
        rho = torch.zeros(1, 3, device=dev, requires_grad=True)
        m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
        if world > 1:
            rast.param_grads_out = grads[i & 1]
        color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                       colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                       rotations=leaves["rotations"], theta=theta, rho=rho)
        # the mapping loss of utils/mapper.py:1239,1252-1267 (photometric L1 + masked depth L1), fused op
        loss, _, _ = photo_depth_loss(color, target, depth, target_depth, opacity, 0.01, 0.5, 80.0, 0.5)
        loss.backward()
        consumed[slot].record(cur)
        if world > 1:
            # rast.param_grads_out: the backward kernel wrote the parameter gradients straight into this step's flat
            # exchange buffer (no pack copy); side-stream all-reduce as in the resident loop
            cur.wait_stream(side)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                exchange_grads(i & 1)
        res = torch.cat([loss.detach().view(1), rho.grad.flatten(), theta.grad.flatten()])
        host_out[slot].copy_(res, non_blocking=True)
        out_done[slot].record(cur)
        if prefetch:
            # pipelined loop: the host reads step i-1's result while step i runs (every step's result
            # still reaches the host inside the timed region; run_e2e collects the last one)
            if i > 0:
                while not out_done[slot ^ 1].query():
                    pass
                losses.append(float(host_out[slot ^ 1][0]))
        else:
            cur.synchronize()
            losses.append(float(host_out[slot][0]))

    def run_e2e(prefetch: bool) -> float:
        """K steps back to back, wall clock from the first upload to the last result on the host; the
        L2 flush between steps is inside the timed region (about 0.04 ms per step)."""
        for ev in consumed:
            ev.record()
        for i in range(args.warmup):
            if prefetch and i == 0:
                upload(0)
            step_e2e(i, prefetch, more=i + 1 < args.warmup)
        torch.cuda.synchronize()
        losses.clear()
        barrier(world)
        first = args.warmup
        for ev in consumed:
            ev.record()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if prefetch:
            upload(first)
        for i in range(first, first + args.steps):
            flush.fill_(1)
            step_e2e(i, prefetch, more=i + 1 < first + args.steps)
        if prefetch:
            last = (first + args.steps - 1) & 1
            out_done[last].synchronize()
            losses.append(float(host_out[last][0]))
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        assert len(losses) >= args.steps and all(x == x for x in losses[-args.steps:]), "e2e results missing"
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / args.steps

    e2e_serial_ms = run_e2e(False)
    h2d_ev.clear()
    e2e_ms_per_step = run_e2e(True)
    torch.cuda.synchronize()
    h2d_ms = sorted(x.elapsed_time(y) for x, y in h2d_ev[-args.steps:])
    h2d_ms = h2d_ms[len(h2d_ms) // 2] if h2d_ms else 0.0       # median device time of one step's upload
    t_ = torch.tensor([h2d_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_, op=dist.ReduceOp.MAX)
    h2d_ms = float(t_.item())
    if os.environ.get("PG_E2E_PROFILE") and rank == 0:   # dev only: where the pipelined step's host and device time go
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            run_e2e(True)
        print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=40, max_name_column_width=60), file=sys.stderr)
        print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=25, max_name_column_width=60), file=sys.stderr)
    e2e_value = world * W * H / (e2e_ms_per_step * 1e-3) / 1e6
    e2e_serial_value = world * W * H / (e2e_serial_ms * 1e-3) / 1e6

    # ---- per-phase attribution (separate pass; events on the launching stream) ----
    phases: dict[str, list[float]] = {}
    if rank == 0:
        L.timing_enable(True)
        for _ in range(3):
            flush.fill_(1)
            step_resident(exchange=False)   # rank-0-only pass: no collective here
            for name, ms in L.timing_drain():
                phases.setdefault(name, []).append(ms)
        L.timing_enable(False)
    barrier(world)
    if rank != 0:
        return None
    phase_ms = {k: sum(v) / len(v) for k, v in phases.items()}
    Q = int(L.lib().pg_raster_band_records(W, H, int(last_binning[3]), last_binning[0].data_ptr(), None)) if last_binning else -1
    BAND_LISTS["n"] = ((W + 15) // 16 + 31) // 32 * ((H + 15) // 16)
    ab = algorithmic_bytes(P, R, W, H, Q)
    pk = peaks()
    per_kernel, per_phase, dom = roofline_tables(phase_ms, ab, pk, l2_bytes(P, R))
    pairs = 0
    if last_binning:
        dv = RC.debug_views(0, P, W, H, int(R), last_binning[1], last_binning[0], last_binning[2])
        pairs = int(dv["n_contrib"].to(torch.int64).sum().item())
    issue = issue_roofline(phase_ms, float(clocks.get("sm_mhz") or 0.0), pairs) if P == C2["P"] else {}
    achieved = per_kernel[dom]["gbs"]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    line = {
        "metric": "raster fwd+bwd Mpix/s (3DGS, pose grads)", "value": round(value, 2), "unit": "Mpix/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2: 3DGS rasterizer fwd+bwd with pose grads, 500k synthetic Gaussians, one 1226x370 view "
                               "per GPU per step (BASELINE.json configs[1])",
                   "P": P, "W": W, "H": H, "num_rendered": int(R), "band_records": Q, "views_per_step": world,
                   "l2": "flushed between steps (256 MiB write)",
                   "exchange": ((("multimem all-reduce (our kernel: in-switch ld_reduce + multicast store over NVSwitch, "
                                  "symmetric memory)") if mm is not None else "NCCL all-reduce")
                                + " of Gaussian-parameter grads (28 MB) on a side stream, overlapping the next "
                                "view's compute; the last one is timed as an exposed tail") if world > 1 else "none (N=1)",
                   "exchange_ms": round(exchange_ms, 4)},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 2), "unit": "Mpix/s", "ms_per_step": round(e2e_ms_per_step, 4),
                "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": 28,
                "h2d_ms": round(h2d_ms, 4), "h2d_gbs": round(h2d_bytes / (h2d_ms * 1e-3) / 1e9, 1) if h2d_ms > 0 else None,
                "staging": "pipelined loop: H2D of step k+1 on a copy stream overlaps step k, the host reads step k's "
                           "result (loss + pose gradient, pinned D2H) while step k+1 runs; L2 flush between steps "
                           "inside the timed region; serial_value = copy, compute, read back to back",
                "serial_value": round(e2e_serial_value, 2), "serial_ms_per_step": round(e2e_serial_ms, 4)},
        "gpu_launches": launches_per_step(T, P) * args.steps,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": pk["hbm_gbs"],
                     "unit": "GB/s", "frac": round(achieved / pk["hbm_gbs"], 4), "traffic": TRAFFIC.get(dom),
                     "peak_source": pk["source"], "kernels": per_kernel, "phases": per_phase,
                     "issue": issue,
                     "note": "kernel = the single longest kernel of the step (per-kernel CUDA-event pairs on the launching "
                             "stream, separate pass); `phases` folds the six band-binning kernels into `binning`; the "
                             "blend kernels are issue-bound (see `issue`), HBM fractions are reported for all"},
    }
    return line


# per-launch dram bytes (dram__bytes_read.sum + dram__bytes_write.sum) from the committed ncu --set full capture of
# the C2 step (profiles/r02_ncu_full_c2.csv) and of the C1 SDF query (profiles/r02_ncu_full_sdf.csv); the capture
# runs the step without the L2 flush, so inputs produced by the previous kernel are L2 hits
TRAFFIC: dict = {"preprocess_fwd": 28.8e6, "depth_sort": 12.7e6, "band_rows_kernel<count>": 3.42e6,
                 "band_list_scan_kernel": 1.19e6, "band_rows_kernel<write>": 6.31e6, "band_expand_kernel<count>": 15.1e6,
                 "band_tile_scan_kernel": 2.80e6, "band_expand_kernel<write>": 51.2e6, "render_fwd": 3.39e6,
                 "render_bwd": 13.05e6, "preprocess_bwd": 50.7e6, "sdf_query": 11.7e6}


def cpu_baseline_line(args) -> dict | None:
    try:
        from oracle import cpu_raster
    except ImportError:
        return None
    return cpu_raster.bench_sample(C2, seconds_budget=20.0)


def run_reference(args, rank, world, local):
    """The unmodified reference CUDA rasterizer on the same scene (rank 0 only)."""
    if rank != 0:
        return None
    from oracle import ref_api
    P, W, H = C2["P"], C2["W"], C2["H"]
    if not ref_api.available("ref_r3"):
        cb = cpu_baseline_line(args)
        if cb is None:
            return {"impl": "reference", "unavailable": "oracle/_ref/libref_r3.so and oracle CPU port both missing"}
        return {"impl": "reference", "metric": "raster fwd+bwd Mpix/s (3DGS, pose grads)", "value": cb["value"],
                "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "higher_is_better": True, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    dev = torch.device("cuda", local)
    cam, sc_cpu, ct_cpu, _ = make_scene(0)
    sc = {k: v.to(dev) for k, v in sc_cpu.items()}
    ct = {k: v.to(dev) for k, v in ct_cpu.items()}
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    bg = torch.ones(3, device=dev)
    ref = ref_api.RefRaster3()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step():
        r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, campos=center, want_views=False)
        g = ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, r["radii"], ct["color"], ct["depth"],
                         campos=center)
        tau = g["tau"].sum(0)  # R3 diff_gaussian_rasterization/__init__.py:152
        return r["num_rendered"], tau

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    ms = 0.0
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        R, tau = step()
        torch.cuda.synchronize()
        ms += (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    ms_per_step = ms / args.steps
    value = W * H / (ms_per_step * 1e-3) / 1e6
    return {"impl": "reference", "metric": "raster fwd+bwd Mpix/s (3DGS, pose grads)", "value": round(value, 2),
            "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C2: 3DGS rasterizer fwd+bwd with pose grads, 500k synthetic Gaussians, one 1226x370 view",
                       "P": P, "W": W, "H": H, "num_rendered": int(R),
                       "note": "reference = unmodified submodules/diff-gaussian-rasterization-w-pose CUDA sources "
                               "compiled for sm_100a (oracle/_ref/libref_r3.so) on this GPU; the reference has no CPU "
                               "rasterizer. Host wall-clock around fwd+bwd (+ its torch.sum of dL_dtau), device synced."},
            "clocks": clocks,
            "cpu_baseline": {"value": round(value, 2), "unit": "Mpix/s", "cores": 0, "kind": "reference",
                             "sample": "full C2 view per step, reference CUDA kernels on the same B200 (no CPU "
                                       "implementation of this path exists in the reference)"},
            "e2e": {"value": round(value, 2), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}



# ----------------------------------------------------------------------------------------------
# Second half of the metric: SDF queries/s (config C1: one synthetic 64x1800 scan, PIN-SLAM settings)
def sdf_leg(args, local, world=1, rank=0, keep=None) -> dict:
    from pings_b200 import _lib as L, synth_neural as SN
    dev = torch.device("cuda", local)
    npm, dec, q = SN.c1_map_and_queries(device=dev)
    if rank > 0:  # query sharding: every rank holds the replicated map and takes its own scan-shaped batch
        g = torch.Generator().manual_seed(1000 + rank)
        q = (q.cpu() + 0.05 * torch.randn(q.shape, generator=g)).to(dev)
    N = q.shape[0]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    steps, warm = max(args.steps, 10), max(args.warmup, 3)
    for _ in range(warm):
        npm.sdf_fused(q, dec)
    barrier(world)
    evs = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        npm.sdf_fused(q, dec)
        b.record()
        evs.append((a, b))
    barrier(world)
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    if keep is not None and rank == 0:
        keep.append((q.clone(), npm.sdf_fused(q, dec)[0].clone()))
    # e2e: host queries -> H2D -> fused query -> D2H of the sdf values
    qh = q.cpu().pin_memory()
    out_h = torch.empty(N, dtype=torch.float32).pin_memory()
    def e2e():
        qd = qh.to(dev, non_blocking=True)
        sdf, _, cnt = npm.sdf_fused(qd, dec)
        out_h.copy_(sdf, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    for _ in range(warm):
        e2e()
    t_e2e = 0.0
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e2e()
        t_e2e += time.perf_counter() - t0
    e2e_ms = t_e2e * 1e3 / steps
    # training-shaped batch (bs = 10,000, run_pin_slam.yaml:24): query_feature + decoder + backward
    qb = q[:10000].clone()
    r = torch.randn(10000, device=dev)
    def train_step():
        npm.local_geo_features.grad = None
        geo, _, w, cnt, _ = npm.query_feature(qb, accumulate_stability=False)
        (dec.sdf(geo) * r).sum().backward()
    for _ in range(warm):
        train_step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        train_step()
    b.record()
    torch.cuda.synchronize()
    train_ms = a.elapsed_time(b) / steps
    t = torch.tensor([ms, e2e_ms, train_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, train_ms = [float(x) for x in t]
    # map-sized variant: ~5e5 neural points, queries spread over the whole map (slot-table probes miss L2)
    big = {}
    try:
        npm_b, dec_b, q_b = SN.large_map_and_queries(device=dev)
        if rank > 0:
            q_b = q_b[torch.randperm(q_b.shape[0], device=dev)]
        for _ in range(warm):
            npm_b.sdf_fused(q_b, dec_b)
        torch.cuda.synchronize()
        evb = []
        for _ in range(steps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            npm_b.sdf_fused(q_b, dec_b)
            b.record()
            evb.append((a, b))
        torch.cuda.synchronize()
        ms_b = sum(a.elapsed_time(b) for a, b in evb) / steps
        tb = torch.tensor([ms_b], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        ms_b = float(tb.item())
        Nb = q_b.shape[0]
        # per query: 81 probes x one 32-B sector each (what the hardware moves for a 4-B slot) + neighbour rows + output
        sect = Nb * (81 * 32 + 332 + 12)
        big = {"value": round(world * Nb / (ms_b * 1e-3), 1), "unit": "queries/s", "ms_per_step": round(ms_b, 4),
               "neural_points": int(npm_b.count()), "queries_per_step_per_gpu": Nb,
               "roofline": {"bound": "hbm", "alg_bytes_per_query": 600 + 332 + 12,
                            "achieved": round(Nb * (600 + 332 + 12) / (ms_b * 1e-3) / 1e9, 1),
                            "sector_bytes_per_query": 81 * 32 + 332 + 12,
                            "sector_gbs": round(sect / (ms_b * 1e-3) / 1e9, 1), "peak": peaks()["hbm_gbs"],
                            "frac": round(Nb * (600 + 332 + 12) / (ms_b * 1e-3) / 1e9 / peaks()["hbm_gbs"], 4),
                            "frac_sectors": round(sect / (ms_b * 1e-3) / 1e9 / peaks()["hbm_gbs"], 4)}}
        del npm_b
    except Exception as ex:   # noqa: BLE001
        big = f"failed: {ex!r}"
    pk = peaks()
    alg = N * (600 + 332 + 4 + 8)  # kNN 600 B + gather/IDW 332 B (int32 slot mirror), + outputs
    gbs = alg / (ms * 1e-3) / 1e9
    return {"sdf": {"metric": "SDF queries/s (fused kNN+IDW+decoder, C1)", "value": round(world * N / (ms * 1e-3), 1),
                    "unit": "queries/s", "ms_per_step": round(ms, 4), "queries_per_step_per_gpu": N,
                    "neural_points": int(npm.count()),
                    "slot_table": "int32 mirror [1e8] (400 MB) of the reference's int64 table",
                    "large_map": big,
                    "e2e": {"value": round(world * N / (e2e_ms * 1e-3), 1), "unit": "queries/s",
                            "h2d_bytes_per_step": N * 12, "d2h_bytes_per_step": N * 4},
                    "train_fwd_bwd_bs10000": {"value": round(world * 10000 / (train_ms * 1e-3), 1), "unit": "queries/s",
                                              "ms_per_step": round(train_ms, 4)},
                    "roofline": {"bound": "hbm", "kernel": "sdf_query", "achieved": round(gbs, 1), "peak": pk["hbm_gbs"],
                                 "unit": "GB/s", "frac": round(gbs / pk["hbm_gbs"], 4), "alg_bytes_per_query": 944,
                                 "traffic": TRAFFIC.get("sdf_query"),
                                 # the map is L2-resident at C1 and the kernel is issue-bound: warp instructions per query
                                 # from profiles/r02_ncu_full_sdf.csv (half-warp kernel) over the live duration
                                 "issue": ({"warp_inst_per_query": 666, "ginst_s": round(666 * N / (ms * 1e-3) / 1e9, 1),
                                            "peak_ginst_s": round(ISSUE_PEAK_GINST, 1),
                                            "frac": round(666 * N / (ms * 1e-3) / 1e9 / ISSUE_PEAK_GINST, 4)}
                                           if os.environ.get("PG_SDF_IMPL") is None else None)}}}


# ----------------------------------------------------------------------------------------------
# Config C3 (BASELINE.json configs[2]): surfel rasterizer fwd+bwd (RGB + depth + normal) with fused-SSIM loss,
# 1M surfels at 1920x1080, through the public autograd modules.  Reported under extra.surfel_ssim.
C3 = dict(P=1_000_000, W=1920, H=1080, seed=1)


def surfel_leg(args, local) -> dict:
    from pings_b200 import _lib as L, synth
    from pings_b200.diff_gaussian_surfel_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from pings_b200.fused_ssim import fused_ssim
    dev = torch.device("cuda", local)
    P, W, H = C3["P"], C3["W"], C3["H"]
    cam = synth.hd_camera(C3["seed"], W, H)
    sc = {k: v.to(dev) for k, v in synth.surfels_in_view(cam, P, C3["seed"]).items()}
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    g = torch.Generator().manual_seed(78)
    target = torch.rand(3, H, W, generator=g).to(dev)
    target_d = (torch.rand(1, H, W, generator=g) * 30 + 2).to(dev)
    target_n = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0).to(dev)
    cfg = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device=dev)
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0, view, full, proj,
                                       torch.tensor([0.0, 0.0, H - 1.0, W - 1.0], device=dev),
                                       torch.tensor([cam.cx / W, cam.cy / H], device=dev), 0, center, False, False, cfg)
    rast = GaussianRasterizer(rs)
    leaves = {k: sc[k].clone().requires_grad_(True) for k in sc}
    theta = torch.zeros(1, 3, device=dev, requires_grad=True)
    rho = torch.zeros(1, 3, device=dev, requires_grad=True)
    m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    info = {}

    def step():
        for t in list(leaves.values()) + [theta, rho, m2d]:
            t.grad = None
        color, normal, depth, opac, radii, contrib = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                          colors_precomp=leaves["colors_precomp"],
                                                          scales=leaves["scales"], rotations=leaves["rotations"],
                                                          theta=theta, rho=rho)
        l1 = (color - target).abs().mean()
        loss = 0.8 * l1 + 0.2 * (1.0 - fused_ssim(color.unsqueeze(0), target.unsqueeze(0)))   # config.py:274
        loss = loss + 0.02 * (depth - target_d).abs().mean() + 0.01 * (1 - (normal * target_n).sum(0)).mean() \
            + 0.01 * opac.mean()
        loss.backward()
        info["vis"] = radii
        return loss

    steps, warm = max(3, min(args.steps, 10)), 3
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    evs = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    # per-phase device time of the library kernels inside the step (the rest is the SSIM kernels + torch loss glue)
    c3_phases: dict[str, list[float]] = {}
    L.timing_enable(True)
    for _ in range(3):
        flush.fill_(1)
        step()
        for name, pms in L.timing_drain():
            c3_phases.setdefault(name, []).append(pms)
    L.timing_enable(False)
    c3_phase_ms = {k: round(sum(v) / len(v), 4) for k, v in c3_phases.items()}
    # standalone fused SSIM fwd+bwd on [1,3,1080,1920] (24 + 28 B/element algorithmic)
    img = torch.rand(1, 3, H, W, device=dev, requires_grad=True)
    tgt = target.unsqueeze(0)
    for _ in range(3):
        img.grad = None
        fused_ssim(img, tgt).backward()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        img.grad = None
        fused_ssim(img, tgt).backward()
    b.record()
    torch.cuda.synchronize()
    ssim_ms = a.elapsed_time(b) / 20
    pk = peaks()
    ssim_gbs = 3 * H * W * (24 + 28) / (ssim_ms * 1e-3) / 1e9
    return {"surfel_ssim": {"metric": "surfel raster fwd+bwd Mpix/s (RGB+depth+normal, fused-SSIM loss, C3)",
                            "value": round(W * H / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s", "ms_per_step": round(ms, 4),
                            "P": P, "W": W, "H": H, "visible": int((info["vis"] > 0).sum()),
                            "phases_ms": c3_phase_ms,
                            "note": "through GaussianRasterizer + fused_ssim autograd modules, loss and torch glue included",
                            "ssim_fwd_bwd": {"ms": round(ssim_ms, 4), "gbs": round(ssim_gbs, 1),
                                             "frac": round(ssim_gbs / pk["hbm_gbs"], 4), "alg_bytes_per_elem": 52}}}


# ----------------------------------------------------------------------------------------------
# Config C4 (BASELINE.json configs[3]; SURVEY.md §8d recipe, utils/mapper.py:995-1146): one PINGS mapping step on a
# synthetic KITTI-shaped LiDAR + RGB frame.  Map = the mirror's update() (pinned to the reference's on a golden,
# tests/test_maintenance_cpu.py) of ONE C1-style scan coloured by a procedural texture at voxel_size_m 0.25
# (config/run_kitti_gs.yaml:21-30: feature_dim 32, color_feature_dim 16, weighted_first False, gs_mlp_hidden_dim 128,
# view_concat, colour residual, 8 surfels per neural point).  A step = the 8-view batch (cameras along a 4 m straight
# path, each 1226x370): per view render() + RGB L1 / fused-SSIM / depth loss + backward; one SDF batch of 8192 rows
# ([N,6,35] -> 64 -> 1, sdf = sum_k w_k sdf_k); ONE flat gradient all-reduce; one fused AdamW step on the flat
# parameter buffer.  The 8 views are SHARDED views[rank::world] (strong scaling), the SDF rows bs/world per rank;
# replicated parameters are checked bit-identical across ranks after the last step.
C4 = dict(Fg=32, Fc=16, hidden=128, level=1, K=8, W=1226, H=370, res=0.25, sdf_bs=8192, views=8, seed=2, lr=1e-3)


class _ViewCam(SimpleNamespace):
    def full_patch(self, down_rate=0):
        s = 2 ** down_rate
        return torch.tensor([0, 0, int(self.image_height / s) - 1, int(self.image_width / s) - 1], dtype=torch.float32,
                            device=self.device)


def _c4_view(i: int, dev):
    """Camera i of the batch: at (0.5 i, 0, 0) m looking along +x (LiDAR frame: x forward, z up)."""
    import math
    from pings_b200 import synth
    W, H = C4["W"], C4["H"]
    T = torch.eye(4, dtype=torch.float64)
    T[:3, 0] = torch.tensor([0.0, -1.0, 0.0], dtype=torch.float64)   # camera x (right)  = -y world
    T[:3, 1] = torch.tensor([0.0, 0.0, -1.0], dtype=torch.float64)   # camera y (down)   = -z world
    T[:3, 2] = torch.tensor([1.0, 0.0, 0.0], dtype=torch.float64)    # camera z (forward) = +x world
    T[:3, 3] = torch.tensor([4.0 * i / (C4["views"] - 1), 0.0, 0.0], dtype=torch.float64)
    cam = synth.Camera(W, H, 718.856, 718.856, 607.19, 185.22, T)
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    vc = _ViewCam(device=dev, uid=i, image_width=W, image_height=H, FoVx=2 * math.atan(cam.tanfovx),
                  FoVy=2 * math.atan(cam.tanfovy), zfar=100.0, fx=cam.fx, fy=cam.fy,
                  prcppoint=torch.tensor([cam.cx / W, cam.cy / H], device=dev), projection_matrix=proj,
                  world_view_transform=view, full_proj_transform=full, camera_center=center,
                  cam_rot_delta=torch.zeros(3, device=dev, requires_grad=True),
                  cam_trans_delta=torch.zeros(3, device=dev, requires_grad=True),
                  exposure_mat=torch.eye(3, device=dev).requires_grad_(True),
                  exposure_offset=torch.zeros(1, 3, device=dev).requires_grad_(True),
                  exposure_a=torch.zeros(1, device=dev), exposure_b=torch.zeros(1, device=dev))
    g = torch.Generator().manual_seed(900 + i)
    target = torch.rand(3, H, W, generator=g).to(dev)
    target_d = (torch.rand(1, H, W, generator=g) * 40 + 3).to(dev)
    return vc, target, target_d


def mapping_leg(args, local, world, rank) -> dict:
    from pings_b200 import synth_neural as SN
    from pings_b200.decoder import Decoder
    from pings_b200.fused_ssim import fused_ssim
    from pings_b200.gaussian_renderer import render
    from pings_b200.losses import photo_depth_loss
    from pings_b200.neural_points import NeuralPoints, default_config
    from pings_b200.parallel import FlatAdamW, FlatGradAllReduce, replicas_in_sync, shard_rows, shard_views
    dev = torch.device("cuda", local)
    Fg, Fc, Hd, Lv, K, W, H = (C4[k] for k in ("Fg", "Fc", "hidden", "level", "K", "W", "H"))
    # ---- map: one coloured scan through update() (identical on every rank) ----
    g = torch.Generator().manual_seed(C4["seed"])
    pts = SN.lidar_scan(C4["seed"])
    cols = (0.5 + 0.5 * torch.sin(pts * torch.tensor([0.7, 1.1, 2.3]))).clamp(0, 1)    # procedural texture
    cfg = default_config(device=dev, voxel_size_m=C4["res"], search_alpha=0.8, feature_dim=Fg, color_feature_dim=Fc,
                         color_on=True, weighted_first=False, buffer_size=int(2e7), track_on=False,
                         temporal_local_map_off=True)
    npm = NeuralPoints(cfg)
    npm.travel_dist = torch.zeros(1, device=dev)
    npm.update(pts.to(dev), cols.to(dev), None, torch.zeros(3, device=dev), None, 0)
    n = npm.count()
    npm.geo_features = (0.1 * torch.randn(n + 1, Fg, generator=g)).to(dev)
    npm.color_features = (0.1 * torch.randn(n + 1, Fc, generator=g)).to(dev)
    npm.reset_local_map(torch.zeros(3, device=dev), None, 0)
    torch.manual_seed(C4["seed"])                  # identical decoder weights on every rank
    decs = {"gauss_xyz": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_rot": Decoder(cfg, Fg, Hd, Lv, 4, K, 0),
            "gauss_scale": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_alpha": Decoder(cfg, Fg, Hd, Lv, 1, K, 0),
            "gauss_color": Decoder(cfg, Fc, Hd, Lv, 3, K, 3)}
    sdf_dec = Decoder(cfg, Fg, 64, 1, 1)
    for d in list(decs.values()) + [sdf_dec]:
        d.to(dev)
    data = {"position": npm.local_neural_points, "orientation": npm.local_point_orientations,
            "color": npm.local_point_colors, "geo_feature": npm.local_geo_features,
            "color_feature": npm.local_color_features, "resolution": C4["res"],
            "free_mask": npm.local_free_gs_mask, "valid_mask": npm.local_valid_gs_mask}
    my_views = [_c4_view(i, dev) for i in shard_views(list(range(C4["views"])), rank, world)]
    q_all = (pts + 0.05 * torch.randn(pts.shape, generator=g))[torch.randperm(pts.shape[0], generator=g)[:C4["sdf_bs"]]]
    lo, hi = shard_rows(C4["sdf_bs"], rank, world)
    qb = q_all[lo:hi].contiguous().to(dev)
    sdf_target = torch.zeros(hi - lo, device=dev)
    bg = torch.zeros(3, device=dev)
    params = [npm.local_geo_features, npm.local_color_features]
    for d in list(decs.values()) + [sdf_dec]:
        params += [p for p in d.parameters()]
    exchange = FlatGradAllReduce(params, average=False, multimem=os.environ.get("PG_EXCHANGE", "multimem") != "nccl")
    opt = FlatAdamW(exchange, lr=C4["lr"])
    info = {}
    nv = float(C4["views"])

    def step():
        exchange.attach()   # grads are views of the flat exchange buffer, zeroed: no pack / scatter copies
        total = None
        for vc, target, target_d in my_views:
            for p in (vc.cam_rot_delta, vc.cam_trans_delta, vc.exposure_mat, vc.exposure_offset):
                p.grad = None
            res = render(vc, None, data, decs, None, bg, view_concat_on=True, learn_color_residual=True,
                         gs_type="gaussian_surfel", d2n_on=False)
            if res is None:
                raise RuntimeError("render() returned None for a C4 view")
            rgb = res["render"]
            l1d, _, _ = photo_depth_loss(rgb, target, res["surf_depth"], target_d, res["rend_alpha"], 0.05, 0.5, 80.0, 1e-3)
            loss = (0.8 * l1d + 0.2 * (1.0 - fused_ssim(rgb.unsqueeze(0), target.unsqueeze(0)))
                    + 0.01 * res["rend_alpha"].mean()) / nv            # sum over the 8 views = mean loss of the batch
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
            info["gaussians"] = int(res["gaussian_xyz"].shape[0])
        geo, _, w, cnt, _ = npm.query_feature(qb, accumulate_stability=False)
        sdf = (sdf_dec.sdf(geo) * w).sum(1).squeeze(1)                # weighted_first False: utils/mapper.py:2276-2279
        l_sdf = 0.5 * (sdf - sdf_target).abs().sum() / C4["sdf_bs"]
        l_sdf.backward()
        exchange()          # ONE all-reduce (sum over ranks = the loss of the global batch)
        opt.step()          # ONE fused AdamW launch on the flat parameter buffer
        return total + l_sdf.detach()

    if not _all_ok(len(my_views) > 0 or world > C4["views"], world, dev):
        raise RuntimeError("C4 setup failed on a rank")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    steps, warm = max(3, min(args.steps, 10)), 3
    for _ in range(warm):
        step()
    if os.environ.get("PG_C4_PROFILE") and rank == 0:      # dev only: host-side cost of the step by function
        import cProfile
        import pstats
        pr = cProfile.Profile()
        torch.cuda.synchronize()
        pr.enable()
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        pr.disable()
        st_ = pstats.Stats(pr, stream=sys.stderr)
        st_.sort_stats("cumulative").print_stats(45)
        st_.sort_stats("tottime").print_stats(30)
    barrier(world)
    evs = []
    loss = None
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        loss = step()
        b.record()
        evs.append((a, b))
    barrier(world)
    in_sync = replicas_in_sync([opt.flat])
    lt = loss.clone()
    if world > 1:
        dist.all_reduce(lt)            # every rank holds its share of the batch loss
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return {"mapping_step": {"metric": "PINGS mapping views/s: 8-view batch (render() + loss + backward) + SDF batch + "
                                       "gradient all-reduce + fused AdamW (C4)",
                             "value": round(C4["views"] / (ms * 1e-3), 2), "unit": "views/s", "ms_per_step": round(ms, 4),
                             "views_per_step": C4["views"], "views_per_rank": len(my_views), "scaling": "strong",
                             "mpix_per_s": round(C4["views"] * W * H / (ms * 1e-3) / 1e6, 2),
                             "neural_points": int(n), "local_points": int(npm.local_count()), "gaussians_per_point": K,
                             "gaussians_rendered_last_view": info.get("gaussians"),
                             "decoder": f"{Fg}+0/{Fc}+3 -> {Hd} x{Lv} -> K*out (5 heads); SDF [N,6,{Fg + 3}] -> 64 -> 1",
                             "sdf_rows_per_rank": hi - lo, "exchange_bytes": exchange.nbytes if world > 1 else 0, "exchange_backend": exchange.backend if world > 1 else "none",
                             "loss_after_steps": round(float(lt), 6), "steps_run": steps + warm,
                             "replicas_in_sync": bool(in_sync),
                             "note": "map built by update() from one coloured C1-style scan at voxel 0.25 m; host-driven "
                                     "step timed with CUDA events, L2 flushed between steps"}}

# ----------------------------------------------------------------------------------------------
# Config C5 (BASELINE.json configs[4]; reference caller inspect_pings.py:361 render_with_poses): 5M surfels, 64 novel
# 1080p views, forward only, views[rank::world], Gaussians replicated - "replicas only", no collective on the path.
C5 = dict(P=5_000_000, W=1920, H=1080, views=64, seed=3)


def _surfel_statics(cam, dev, W, H):
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    return dict(view=view, full=full, proj=proj, center=center, tanx=cam.tanfovx, tany=cam.tanfovy,
                prcp=torch.tensor([cam.cx / W, cam.cy / H], device=dev),
                bbox=torch.tensor([0.0, 0.0, H - 1.0, W - 1.0], device=dev))


def large_map_leg(args, local, world, rank) -> dict:
    from pings_b200 import _lib as L, _raster_c as RC, synth
    dev = torch.device("cuda", local)
    P, W, H, NV = C5["P"], C5["W"], C5["H"], C5["views"]
    sc = {k: v.to(dev) for k, v in synth.large_map(P, C5["seed"]).items()}
    cams = synth.loop_cameras(NV, W, H)
    mine = list(range(NV))[rank::world]
    st = {i: _surfel_statics(cams[i], dev, W, H) for i in mine}
    cfg = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device=dev)
    bg = torch.zeros(3, device=dev)
    e = torch.empty(0)

    def render_view(i):
        c = st[i]
        return RC.rasterize_gaussians_surfel(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                                             sc["rotations"], 1.0, e, c["view"], c["full"], c["proj"], c["prcp"], c["bbox"],
                                             c["tanx"], c["tany"], H, W, e, 0, c["center"], False, False, cfg)

    with torch.no_grad():
        for i in mine[:2]:
            render_view(i)
        barrier(world)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        Rs = []
        for i in mine:      # inputs (280 MB of Gaussians + 150 MB of per-Gaussian state) exceed the 126 MB L2: no flush needed
            Rs.append(render_view(i)[0])
        b.record()
        barrier(world)
        ms = a.elapsed_time(b)
        # e2e: the rendered RGB of every view reaches pinned host memory (24.9 MB / view), D2H on a copy stream
        # overlapping the next view; camera matrices come from the host (tiny H2D)
        host = [torch.empty(3, H, W, dtype=torch.float32).pin_memory() for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        done = [torch.cuda.Event() for _ in range(2)]
        cams_host = {i: [m.pin_memory() for m in cams[i].matrices()] for i in mine}
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k, i in enumerate(mine):
            c = st[i]
            for dst, src in zip((c["view"], c["full"], c["proj"], c["center"]), cams_host[i]):
                dst.copy_(src, non_blocking=True)
            out = render_view(i)
            ev = torch.cuda.Event()
            ev.record()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev)
                if k >= 2:
                    done[k & 1].synchronize()
                host[k & 1].copy_(out[1], non_blocking=True)
                out[1].record_stream(copy_stream)
                done[k & 1].record(copy_stream)
        copy_stream.synchronize()
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    res = {"metric": "large-map render Mpix/s (5M surfels, 64 novel 1080p views, forward only, C5)",
           "value": round(NV * W * H / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s", "ms_per_view": round(ms / len(mine), 4),
           "views": NV, "views_per_rank": len(mine), "P": P, "W": W, "H": H, "scaling": "strong (64 views split over N)",
           "num_rendered_mean": int(sum(Rs) / max(len(Rs), 1)),
           "e2e": {"value": round(NV * W * H / (e2e_ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
                   "h2d_bytes_per_view": 4 * (16 * 3 + 3), "d2h_bytes_per_view": 3 * H * W * 4},
           "l2": "inputs larger than L2 (430 MB of Gaussian data per view)", "collective": "none (replicas only)"}
    if rank == 0:
        phases: dict = {}
        L.timing_enable(True)
        with torch.no_grad():
            for i in mine[:3]:
                out = render_view(i)
                R = out[0]
                for name, pms in L.timing_drain():
                    phases.setdefault(name, []).append(pms)
        L.timing_enable(False)
        phase_ms = {k: sum(v) / len(v) for k, v in phases.items()}
        Q = int(L.lib().pg_raster_band_records(W, H, int(RC.last_binning_capacity()), out[8].data_ptr(), None))
        ab = algorithmic_bytes(P, int(R), W, H, Q, surfel=True)
        per_kernel, per_phase, dom = roofline_tables(phase_ms, ab, peaks())
        res["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["gbs"], "peak": peaks()["hbm_gbs"],
                           "unit": "GB/s", "frac": per_kernel[dom]["frac"], "kernels": per_kernel, "phases": per_phase,
                           "band_records": Q}
    return {"large_map_render": res}


# ----------------------------------------------------------------------------------------------
# A second 3DGS scene with PINGS-like occupancy (the C2 recipe is occlusion-degenerate, every tile list
# saturates after 2 % of its entries): 300k spawned-Gaussian sized, low-opacity splats, same camera and kernels.
def raster3_scene_leg(args, local, scene: str = "pings_like") -> dict:
    from pings_b200 import _lib as L, _raster_c as RC, synth
    dev = torch.device("cuda", local)
    W, H = C2["W"], C2["H"]
    cam = synth.kitti_camera(C2["seed"], W, H)
    sc_cpu = synth.pings_like_gaussians(cam)
    P = sc_cpu["means3D"].shape[0]
    sc = {k: v.to(dev) for k, v in sc_cpu.items()}
    ct = {k: v.to(dev) for k, v in synth.cotangents(cam, C2["seed"]).items()}
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    bg = torch.ones(3, device=dev)
    e = torch.empty(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    keep = {}

    def step():
        R, color, radii, geom, binning, img, depth, opac, n_touched = RC.rasterize_gaussians(
            bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full,
            proj, cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, False)
        RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0,
                                        e, view, full, proj, cam.tanfovx, cam.tanfovy, ct["color"], ct["depth"], e, 0,
                                        center, geom, R, binning, img, False)
        keep.update(R=R, geom=geom, binning=binning, img=img, radii=radii)
        return R

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    steps = max(5, min(args.steps, 20))
    evs = []
    for _ in range(steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        R = step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    dv = RC.debug_views(0, P, W, H, R, keep["geom"], keep["binning"], keep["img"])
    T = ((W + 15) // 16) * ((H + 15) // 16)
    ll = (dv["ranges"][:, 1] - dv["ranges"][:, 0]).float()
    nc = dv["n_contrib"].float().view(H, W)
    pad = torch.zeros(((H + 15) // 16) * 16, ((W + 15) // 16) * 16, device=dev)
    pad[:H, :W] = nc
    tile_max = pad.view((H + 15) // 16, 16, (W + 15) // 16, 16).amax(dim=(1, 3)).flatten()
    walked = float((tile_max / ll.clamp_min(1)).mean())
    phases: dict = {}
    L.timing_enable(True)
    for _ in range(3):
        flush.fill_(1)
        step()
        for name, pms in L.timing_drain():
            phases.setdefault(name, []).append(pms)
    L.timing_enable(False)
    phase_ms = {k: round(sum(v) / len(v), 4) for k, v in phases.items()}
    return {"raster_" + scene: {
        "metric": "raster fwd+bwd Mpix/s (3DGS, pose grads), PINGS-like occupancy", "value": round(W * H / (ms * 1e-3) / 1e6, 2),
        "unit": "Mpix/s", "ms_per_step": round(ms, 4), "P": P, "W": W, "H": H, "num_rendered": int(R),
        "instances_per_gaussian": round(R / max(int((keep["radii"] > 0).sum()), 1), 2),
        "mean_list_len": round(float(ll.mean()), 1), "fraction_of_list_walked": round(walked, 3),
        "kernels_ms": phase_ms, "l2": "flushed between steps"}}


def sdf_cpu_leg(budget_s: float = 15.0, check_against=None) -> dict:
    """The REFERENCE's own PyTorch CPU SDF path on this box's host cores (BASELINE.json north_star: "the reference's
    PyTorch CPU SDF/decoder path ... is the reported baseline"): model/neural_gaussians.py NeuralPoints.update ->
    query_feature(accumulate_stability=False) + model/decoder.py Decoder.sdf in infer_bs batches under no_grad
    (utils/mapper.py:2273-2289), imported from /root/reference or, on the GPU box, from the copy oracle/build_ref.py
    staged under oracle/_ref/py/.  Falls back to the restatement oracle/neural_cpu.py (kind "port") only when neither
    exists.  check_against: optional (queries, sdf) from the GPU path, compared on the first batch."""
    from pings_b200 import synth_neural as SN
    torch.set_num_threads(os.cpu_count() or 1)
    from oracle import ref_python as rp
    if not rp.available():
        return _sdf_cpu_leg_port(budget_s)
    cfg = rp.load_config("run_pin_slam.yaml")
    NP, Dec = rp.neural_points_cls(), rp.decoder_cls()
    pts = SN.lidar_scan(42, 1800)
    g = torch.Generator().manual_seed(43)
    npm = NP(cfg)
    npm.travel_dist = torch.zeros(1)
    npm.update(pts, None, None, torch.zeros(3), None, 0)
    n = npm.neural_points.shape[0]
    npm.geo_features = 0.1 * torch.randn(n + 1, cfg.feature_dim, generator=g)   # same draw as SN.c1_map_and_queries
    npm.reset_local_map(torch.zeros(3), None, 0)
    torch.manual_seed(42)
    dec = Dec(cfg, cfg.feature_dim, 64, 1, 1)
    q = (pts + 0.05 * torch.randn(pts.shape, generator=g)).contiguous()
    bs = int(cfg.infer_bs)

    def run():
        out = []
        with torch.no_grad():
            for h in range(0, q.shape[0], bs):
                geo, _, w, cnt, _ = npm.query_feature(q[h:h + bs], accumulate_stability=False)
                out.append(dec.sdf(geo))
        return torch.cat(out)

    sdf = run()
    times = []
    t_end = time.perf_counter() + budget_s
    while len(times) < 3 or (time.perf_counter() < t_end and len(times) < 20):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    best, med = min(times), statistics.median(times)
    res = {"value": round(q.shape[0] / med, 1), "best": round(q.shape[0] / best, 1), "unit": "queries/s",
           "cores": torch.get_num_threads(), "kind": "reference",
           "source": "staged copy oracle/_ref/py" if rp.staged_only() else rp.REF,
           "neural_points": int(n),
           "sample": f"full C1 scan ({q.shape[0]} queries, infer_bs {bs}, no_grad), reference NeuralPoints.query_feature + "
                     f"Decoder.sdf, median of {len(times)} runs"}
    if check_against is not None:
        qg, sg = check_against
        if qg.shape == q.shape and bool(torch.equal(qg.cpu(), q)):
            res["max_abs_diff_vs_gpu"] = float((sg.cpu() - sdf).abs().max())
    return res


def _sdf_cpu_leg_port(budget_s: float = 15.0) -> dict:
    """Fallback: the restatement oracle/neural_cpu.py (only when no copy of the reference Python is available)."""
    from oracle import neural_cpu as O
    from pings_b200 import synth_neural as SN
    npm, dec, q = SN.c1_map_and_queries(device="cpu")
    S = SN.oracle_state(npm)
    Ws = [l.weight.detach() for l in dec._linears()]
    Bs = [l.bias.detach() for l in dec._linears()]
    def run():
        with torch.no_grad():
            for h in range(0, q.shape[0], 80000):  # infer_bs batches (config.py:774)
                O.sdf_query(S, q[h:h + 80000], 6, True, Ws, Bs, dec.sdf_scale, time_filtering=False)
    run()
    times = []
    t_end = time.perf_counter() + budget_s
    while len(times) < 5 or (time.perf_counter() < t_end and len(times) < 20):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    best, med = min(times), statistics.median(times)
    return {"value": round(q.shape[0] / med, 1), "best": round(q.shape[0] / best, 1), "unit": "queries/s",
            "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"full C1 scan ({q.shape[0]} queries, infer_bs 80000, no_grad), median of {len(times)} runs"}


# ----------------------------------------------------------------------------------------------
# Reference-arm comparators (bench.py --impl reference): the UNMODIFIED reference CUDA code built into oracle/_ref on
# the other configurations - R3 on the PINGS-like scene, RS on C3 (fwd+bwd) and C5 (fwd), fused-ssim on the two
# shapes of FS/tests/test.py:59 and BASELINE.md rows B2/B3.  Same scenes, same L2 flush, device synced wall clock.
def _timed(fn, steps, flush):
    fn()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(steps):
        if flush is not None:
            flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ms += (time.perf_counter() - t0) * 1e3
    return ms / steps


def reference_extras(args, local) -> dict:
    from oracle import ref_api
    from pings_b200 import synth
    dev = torch.device("cuda", local)
    out: dict = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    steps = max(3, min(args.steps, 5))
    if ref_api.available("ref_r3"):
        try:
            W, H = C2["W"], C2["H"]
            cam = synth.kitti_camera(C2["seed"], W, H)
            sc = {k: v.to(dev) for k, v in synth.pings_like_gaussians(cam).items()}
            ct = {k: v.to(dev) for k, v in synth.cotangents(cam, C2["seed"]).items()}
            view, full, proj, center = [m.to(dev) for m in cam.matrices()]
            bg = torch.ones(3, device=dev)
            ref = ref_api.RefRaster3()
            info = {}
            def step():
                r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, campos=center, want_views=False)
                g = ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, r["radii"], ct["color"], ct["depth"],
                                 campos=center)
                g["tau"].sum(0)
                info["R"] = r["num_rendered"]
            ms = _timed(step, steps, flush)
            out["raster_pings_like"] = {"value": round(W * H / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
                                        "ms_per_step": round(ms, 4), "num_rendered": int(info["R"]),
                                        "impl": "reference R3 CUDA (oracle/_ref/libref_r3.so)"}
        except Exception as ex:
            out["raster_pings_like"] = f"failed: {ex!r}"
    if ref_api.available("ref_rs"):
        try:
            P, W, H = C3["P"], C3["W"], C3["H"]
            cam = synth.hd_camera(C3["seed"], W, H)
            sc = {k: v.to(dev) for k, v in synth.surfels_in_view(cam, P, C3["seed"]).items()}
            ct = {k: v.to(dev) for k, v in synth.cotangents(cam, C3["seed"], normal=True).items()}
            s = _surfel_statics(cam, dev, W, H)
            cfg = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device=dev)
            bg = torch.zeros(3, device=dev)
            ref = ref_api.RefRasterS()
            mats = (s["view"], s["full"], s["proj"])
            info = {}
            def step3():
                r = ref.forward(sc, mats, W, H, s["tanx"], s["tany"], bg, s["prcp"], s["bbox"], cfg, campos=s["center"],
                                want_views=False)
                g = ref.backward(sc, mats, s["tanx"], s["tany"], bg, s["prcp"], s["bbox"], cfg, r["radii"], ct["color"],
                                 ct["normal"], ct["depth"], ct["opac"], campos=s["center"])
                g["tau"].sum(0)
                info["R"] = r["num_rendered"]
            ms = _timed(step3, steps, flush)
            out["surfel_raster_c3"] = {"value": round(W * H / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
                                       "ms_per_step": round(ms, 4), "num_rendered": int(info["R"]),
                                       "note": "rasterizer fwd+bwd only (no loss glue, no SSIM)",
                                       "impl": "reference RS CUDA (oracle/_ref/libref_rs.so)"}
            del sc, ct
        except Exception as ex:
            out["surfel_raster_c3"] = f"failed: {ex!r}"
        try:
            P, W, H = C5["P"], C5["W"], C5["H"]
            sc = {k: v.to(dev) for k, v in synth.large_map(P, C5["seed"]).items()}
            cams = synth.loop_cameras(C5["views"], W, H)
            cfg = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device=dev)
            bg = torch.zeros(3, device=dev)
            ref = ref_api.RefRasterS()
            sub = list(range(0, C5["views"], 8))          # 8 of the 64 views (bounded sample of the same workload)
            sts = [_surfel_statics(cams[i], dev, W, H) for i in sub]
            def step5():
                for s5 in sts:
                    ref.forward(sc, (s5["view"], s5["full"], s5["proj"]), W, H, s5["tanx"], s5["tany"], bg, s5["prcp"],
                                s5["bbox"], cfg, campos=s5["center"], want_views=False)
            ms = _timed(step5, 2, None) / len(sub)
            out["large_map_render"] = {"value": round(W * H / (ms * 1e-3) / 1e6, 2), "unit": "Mpix/s",
                                       "ms_per_view": round(ms, 4), "views_timed": len(sub), "P": P,
                                       "impl": "reference RS CUDA forward (oracle/_ref/libref_rs.so), one GPU"}
            del sc
        except Exception as ex:
            out["large_map_render"] = f"failed: {ex!r}"
    mod = ref_api.ref_ssim_module()
    if mod is not None:
        try:
            res = {}
            for shape in ((1, 3, 1080, 1920), (5, 5, 1080, 1920)):
                g = torch.Generator().manual_seed(9)
                a = torch.rand(*shape, generator=g).to(dev)
                b = torch.rand(*shape, generator=g).to(dev)
                n = a.numel()
                dmap = torch.full_like(a, 1.0 / n)
                def ssim_step():
                    m, d1, d2, d3 = mod.fusedssim(0.01 ** 2, 0.03 ** 2, a, b, True)
                    mod.fusedssim_backward(0.01 ** 2, 0.03 ** 2, a, b, dmap, d1, d2, d3)
                ms = _timed(ssim_step, 10, None)
                res["x".join(map(str, shape))] = {"ms_fwd_bwd": round(ms, 4), "gelem_per_s": round(n / (ms * 1e-3) / 1e9, 2)}
            out["fused_ssim"] = {"impl": "reference fused-ssim CUDA (oracle/_ref/fused_ssim_cuda_ref.so, --use_fast_math)", **res}
        except Exception as ex:
            out["fused_ssim"] = f"failed: {ex!r}"
    return out


# ---- decoder MLP leg (extra.mlp_batch): the one dense contraction of the path ----------------------
def mlp_leg(args, local) -> dict:
    """Decoder.mlp_batch on a spawn_gaussians-sized batch (GR:609-714): 2^20 rows through a 35 -> 64 -> 64 -> 5
    head, tcgen05/TMEM kernel (3xTF32) with the fp32 SIMT kernel beside it."""
    import types
    from pings_b200 import decoder as D
    dev = torch.device("cuda", local)
    cfg = types.SimpleNamespace(mlp_leaky_relu=False, infer_bs=4096, mlp_bias_on=True, main_loss_type="bce",
                                logistic_gaussian_ratio=0.55, sigma_sigmoid_m=0.1, device=dev)
    torch.manual_seed(0)
    dec = D.Decoder(cfg, 32, 64, 2, 5)
    N = 1 << 20
    x = torch.randn(N, 35, device=dev)
    flop = 2.0 * N * (35 * 64 + 64 * 64 + 64 * 5)
    byts = N * (35 + 5) * 4
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(impl):
        for _ in range(max(3, args.warmup)):
            D.mlp_forward_impl(dec, x, impl)
        ms = 0.0
        for _ in range(args.steps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            D.mlp_forward_impl(dec, x, impl)
            b.record()
            b.synchronize()
            ms += a.elapsed_time(b)
        return ms / args.steps

    ms_tc, ms_simt = timed(D.MLP_TCGEN05), timed(D.MLP_SIMT)
    pk = peaks()
    tf32_peak = pk.get("bf16_tflops", 2250.0) / 2.0
    return {"mlp_batch": {
        "metric": "decoder MLP rows/s (mlp_batch, 35-64-64-5, fp32-accurate)", "value": round(N / (ms_tc * 1e-3), 1),
        "unit": "rows/s", "ms_per_step": round(ms_tc, 4), "rows": N, "simt_fp32_ms": round(ms_simt, 4),
        "roofline": {"bound": "tensor", "kernel": "mlp_tc_forward_kernel", "achieved": round(flop / ms_tc / 1e9, 2),
                     "peak": round(tf32_peak, 1), "unit": "TFLOP/s", "frac": round(flop / ms_tc / 1e9 / tf32_peak, 4),
                     "note": "algorithmic fp32 FLOPs; the kernel issues 3 TF32 MMAs per algorithmic product (3xTF32 "
                             "split) and is bound by its register epilogue (bias, activation, hi/lo split, restage), "
                             "peak = measured bf16 dense / 2",
                     "hbm_gbs": round(byts / ms_tc / 1e6, 1)}}}


def _all_ok(ok: bool, world: int, dev) -> bool:
    """Collective-safe go/no-go: every rank learns whether ALL ranks are fine before entering a leg's collectives."""
    if world == 1:
        return ok
    t = torch.tensor([1 if ok else 0], device=dev, dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t.item()))


def run_leg(name, fn, world, dev):
    """Run an optional leg on every rank.  A rank that fails records the error; the ranks then agree (MIN
    all-reduce) so that nobody is left waiting inside the leg's next collective, and the leg is reported as failed."""
    err = None
    try:
        res = fn()
    except Exception as ex:   # noqa: BLE001 - an optional leg must not take the headline down
        res, err = None, repr(ex)
    try:
        ok = _all_ok(err is None, world, dev)
    except Exception as ex:   # the process group itself is gone
        ok, err = False, err or repr(ex)
    if not ok:
        return {name: f"failed: {err or 'another rank failed'}"}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--legs", default="all", help="comma list of optional legs: sdf,mapping,large_map,surfel,pings_like,mlp")
    args = ap.parse_args()
    want = (lambda k: True) if args.legs == "all" else (lambda k: k in args.legs.split(","))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":  # rank 0 alone runs it; no process group is formed
        if int(os.environ.get("RANK", "0")) == 0:
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            world = int(os.environ.get("WORLD_SIZE", "1"))
            ref_line = run_reference(args, 0, world, local)
            ex = ref_line.setdefault("extra", {})
            ex["note"] = ("the reference is single-GPU (SURVEY.md 0.3): every number of this arm is ONE GPU, whatever "
                          "n_gpus the job was launched with")
            if args.legs != "none":
                try:
                    ex.update(reference_extras(args, local))
                except Exception as e:   # noqa: BLE001
                    ex["comparators"] = f"failed: {e!r}"
                try:  # the reference's own CPU SDF path, the reported baseline of the second metric half
                    ex["sdf"] = sdf_cpu_leg()
                except Exception as e:   # noqa: BLE001
                    ex["sdf"] = f"failed: {e!r}"
            print(json.dumps(ref_line), flush=True)
        return
    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    line = run_ours(args, rank, world, local)
    extra: dict = {}
    sdf_check: list = []
    if want("sdf"):
        extra.update(run_leg("sdf", lambda: sdf_leg(args, local, world, rank, keep=sdf_check), world, dev))
    if want("mapping"):
        extra.update(run_leg("mapping_step", lambda: mapping_leg(args, local, world, rank), world, dev))
    if want("large_map"):
        extra.update(run_leg("large_map_render", lambda: large_map_leg(args, local, world, rank), world, dev))
    if rank == 0:
        line.setdefault("extra", {}).update(extra)
        if world == 1:
            for key, name, fn in (("surfel", "surfel_ssim", lambda: surfel_leg(args, local)),
                                  ("pings_like", "raster_pings_like", lambda: raster3_scene_leg(args, local)),
                                  ("mlp", "mlp_batch", lambda: mlp_leg(args, local))):
                if want(key):
                    line["extra"].update(run_leg(name, fn, 1, dev))
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline_line(args)
            if cb is not None:
                line["cpu_baseline"] = cb
            if isinstance(line["extra"].get("sdf"), dict):
                try:
                    line["extra"]["sdf"]["cpu_baseline"] = sdf_cpu_leg(check_against=sdf_check[0] if sdf_check else None)
                except Exception as e:   # noqa: BLE001
                    line["extra"]["sdf"]["cpu_baseline"] = f"failed: {e!r}"
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
