#!/usr/bin/env python
"""bench.py -- density fields/s through the scattering path (I(q) + peaks) on B200.

Contract: ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line (rank 0).
A step = one pass of the hot path over one batch of synthetic fields per GPU (default: 256
gyroid/BCC 128^3 fp64 fields, BASELINE.json config 2).  ``value`` = device-resident whole-job
fields/s; ``e2e`` = the same through the host-buffer C-ABI call (pinned host -> device copies
inside the timed region); ``roofline`` = the dominant hand-written kernel against the measured
HBM peak; ``cpu_baseline`` = the numpy port of the reference on this host's cores.
``--impl reference`` times that CPU port alone.

The default line also carries the rest of BASELINE.json's metric as sub-records (each leg a few seconds):
``"voronoi"`` (256^3 / 4096 seeds / uint8 mask, config 3: voxels/s, roofline, e2e, cpu_baseline, parity),
``"sweep"`` (config 4 at its stated 256 fields of 256^3 per GPU: scatter + Voronoi per field) and, with
``--gpus N > 1``, ``"slab"`` (config 5: one 1024^3 field through the real NCCL path in both modes, stage split,
all-to-all GB/s, parity of the two modes) and ``"voronoi_sharded"`` (one 256^3 and one 512^3 grid by voxel slab,
strong scaling).  ``--workload voronoi|slab|sweep`` prints one of those records alone; ``--sub none`` drops them.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

PEAK_KW = dict(width=0.5, step=2000, cutoff=20, min_height=1)


def measured_traffic(key):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture, or None."""
    try:
        return json.loads((ROOT / "profiles" / "traffic.json").read_text()).get(key, {}).get("bytes")
    except Exception:
        return None


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU baseline (oracle port)
def _cpu_one_field(args):
    idx, n, cells = args
    from oracle import scatter_np
    from simubox_b200 import synth

    mat = synth.field_numpy(idx, (n, n, n), cells=cells)
    t0 = time.perf_counter()
    qI, _ = scatter_np.scattering_peak(mat, synth.box_for(cells), cutoff=30, threshold=1e-12)
    pk = scatter_np.show_peak(qI, **PEAK_KW)
    return time.perf_counter() - t0, qI, pk["peaks_index"]


def cpu_scatter_baseline(n, cells, fields, cores):
    """numpy port of Scatter.sacttering_peak + show_peak (oracle/scatter_np.py), one field per worker."""
    import multiprocessing as mp

    jobs = [(i, n, cells) for i in range(fields)]
    t0 = time.perf_counter()
    if cores > 1:
        with mp.get_context("spawn").Pool(cores) as pool:
            pool.map(_cpu_one_field, jobs[:cores])  # warm: imports, page-in
            t0 = time.perf_counter()
            res = pool.map(_cpu_one_field, jobs)
    else:
        res = [_cpu_one_field(j) for j in jobs]
    # throughput of the scattering work itself (field generation excluded): the per-field times were
    # measured while all workers ran concurrently, so memory-bandwidth contention is included
    busy = sum(r[0] for r in res) / max(cores, 1)
    return fields / busy, res


def _cpu_voronoi_chunk(args):
    pts, vox, box = args
    from oracle import voronoi_np

    return voronoi_np.label_voxels(pts, 0, vox, box=box, chunk=64)


def cpu_voronoi_baseline(pts, shape, box, voxels, cores):
    import multiprocessing as mp

    rng = np.random.default_rng(5)
    vox = np.stack([rng.integers(0, n, voxels) for n in shape], axis=1)
    parts = np.array_split(vox, max(cores, 1))
    t0 = time.perf_counter()
    if cores > 1:
        with mp.get_context("spawn").Pool(cores) as pool:
            pool.map(_cpu_voronoi_chunk, [(pts, p[:8], box) for p in parts])
            t0 = time.perf_counter()
            labs = pool.map(_cpu_voronoi_chunk, [(pts, p, box) for p in parts])
    else:
        labs = [_cpu_voronoi_chunk((pts, vox, box))]
    wall = time.perf_counter() - t0
    return voxels / wall, vox, np.concatenate(labs)


# ----------------------------------------------------------------------------- distributed plumbing
_CTX = None


def dist_setup(gpus):
    """(rank, world, local) -- the process group is created once per process."""
    global _CTX
    import torch

    if _CTX is None:
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        local = int(os.environ.get("LOCAL_RANK", "0"))
        if world > 1:
            import torch.distributed as dist

            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        _CTX = (rank, world, local)
    return _CTX


def barrier(world):
    import torch

    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world, device):
    import torch

    if world == 1:
        return float(x)
    import torch.distributed as dist

    t = torch.tensor([float(x)], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_steps(fn, steps, warmup, world, device):
    """W warm-up steps, then exactly K steps between barrier+synchronize, CUDA events on the
    launching (current) stream; returns max-over-ranks milliseconds per step."""
    import torch

    for _ in range(warmup):
        fn()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    barrier(world)
    return max_over_ranks(ms, world, device)


# ----------------------------------------------------------------------------- scatter workload
def run_scatter(args):
    import torch

    from simubox_b200 import Scatter, _lib, synth
    from simubox_b200.scatter import get_plan

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n, cells, B = args.grid, args.cells, args.fields
    L = synth.box_for(cells)
    shape = (n, n, n)
    # synthetic fields, generated on the device in slices (rank r gets fields r*B .. r*B+B-1)
    fields = torch.empty((B, *shape), dtype=torch.float64, device=dev)
    for b0 in range(0, B, 16):
        nb = min(16, B - b0)
        fields[b0:b0 + nb] = synth.fields_torch(rank * B + b0, nb, shape, cells=cells, device=dev)
    torch.cuda.synchronize()
    path = {"auto": 0, "dense": 1, "gather": 2, "pruned": 3, "fused": 4}[args.path]
    kw = dict(cutoff=30, threshold=1e-12, peaks=True, width=PEAK_KW["width"], step=PEAK_KW["step"],
              peak_cutoff=PEAK_KW["cutoff"], min_height=PEAK_KW["min_height"], max_batch=args.max_batch, device=local, path=path)
    plan = get_plan(shape, L, 30, 1e-12, max_batch=min(args.max_batch, B), device=local, path=path)
    last = {}

    def step_device():
        last["out"] = Scatter.sacttering_peak_batch(fields, L, **kw)

    # --- device-resident throughput (inputs already in HBM; 4.3 GB per pass >> 126 MB L2, no flush needed)
    clk = ClockSampler(local).__enter__()  # sampled across the device-timed, profiled and end-to-end regions
    step_device()
    t_wait = time.time()
    while not clk.rows and time.time() - t_wait < 3.0:
        time.sleep(0.05)
    plan.set_profiling(False)
    ms = timed_steps(step_device, args.steps, args.warmup, world, dev)
    # --- same steps again with per-stage CUDA events (transform vs hand-written kernel split): one profile per step,
    # warm-up step dropped, MEDIAN over the steps (a sum over all launches once let a single perturbed step put the
    # kernel time above the step time)
    plan.set_profiling(True)
    plan.profile(reset=True)
    per_step = []
    for i in range(args.steps + 1):
        step_device()
        pr = plan.profile(reset=True)  # synchronises on the step's last event
        if i:
            per_step.append(pr)
    barrier(world)
    plan.set_profiling(False)
    med = lambda key: float(np.median([pr[key] for pr in per_step]))
    own_per_step = med("own_launches")
    transform_ms_per_step, fused_ms_per_step = med("transform_ms"), med("fused_ms")
    fused_ms_per_launch = fused_ms_per_step / max(own_per_step, 1)
    prof = {"transform_ms": transform_ms_per_step, "fused_ms": fused_ms_per_step, "own_launches": own_per_step}

    # --- end to end through the host-buffer C-ABI call (pinned host memory).  The rank is bound to its GPU's NUMA-local
    # CPUs BEFORE the pinned allocation, so that first touch places the staging buffer on that node.
    from simubox_b200.distributed import bind_to_gpu_numa

    affinity0 = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa(local)
    host = torch.empty((B, *shape), dtype=torch.float64, pin_memory=True)
    host.copy_(fields)
    torch.cuda.synchronize()

    kw_host = dict(kw, max_batch=args.host_batch)

    def step_host():
        last["host"] = Scatter.sacttering_peak_batch(host, L, **kw_host)

    for _ in range(max(1, min(args.warmup, 2))):
        step_host()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        step_host()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.e2e_steps, world, dev)
    os.sched_setaffinity(0, affinity0)  # the CPU-baseline leg below uses every core this process was given
    clk.__exit__()
    clocks = clk.summary()
    ns = plan.n_shells
    h2d = B * n ** 3 * 8
    d2h = B * (ns * 8 + 4 + 256 * 4)
    assert np.array_equal(last["host"]["I"], last["out"]["I"].cpu().numpy()) or plan.path == 1, "host and device paths disagree"

    del host, fields
    last_out = {k: (v.cpu() if hasattr(v, "cpu") else v) for k, v in last["out"].items()}
    last.clear()
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    last["out"] = last_out
    peak, peak_src = measured_peak_gbs()
    spec_bytes = 16 * n * n * (n // 2 + 1)
    pipeline_bytes = 8 * n ** 3 + 2 * spec_bytes
    pathname = _lib.PATH_NAMES[plan.path]
    if plan.path == 4:  # fused pruned DFT: ONE kernel reads each field exactly once and finishes it
        kern, unit_bytes = (plan.kernel or "pruned_fused_kernel"), 8 * n ** 3
        launches_per_step = own_per_step
        fused_ms_per_launch = prof["transform_ms"] / max(prof["own_launches"], 1)
        if fused_ms_per_launch * launches_per_step > ms:
            # the kernel IS the step here (one launch per step); event pairs of the profiled pass cannot undercut the
            # un-profiled step they repeat, so a larger median means the profiled pass was perturbed: use the step time
            fused_ms_per_launch = ms / max(launches_per_step, 1)
    elif plan.path == 3:  # pruned DFT: pruned_planes_kernel reads each field exactly once (first event pair)
        kern, unit_bytes = "pruned_planes_kernel", 8 * n ** 3
        launches_per_step = own_per_step / 2
        fused_ms_per_launch = prof["transform_ms"] / max(prof["own_launches"] / 2, 1)
    else:
        kern, unit_bytes = ("shell_histogram_kernel" if plan.path == 1 else "shell_gather_kernel"), spec_bytes
        launches_per_step = own_per_step
    fields_per_launch = B / max(launches_per_step, 1)
    achieved = unit_bytes * fields_per_launch / (fused_ms_per_launch * 1e-3) / 1e9 if fused_ms_per_launch > 0 else None
    roof = {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": (achieved / peak) if achieved else None,
            "traffic": measured_traffic(f"{kern}|{n}^3|{int(round(fields_per_launch))} fields/launch"), "peak_source": peak_src,
            "algorithmic_bytes_per_field": unit_bytes, "fields_per_launch": fields_per_launch,
            "kernel_ms_per_launch": fused_ms_per_launch,
            "step_split_ms": ({kern: transform_ms_per_step} if plan.path == 4 else
                              {"pruned_planes_kernel": transform_ms_per_step, "pruned_finish_kernel": fused_ms_per_step} if plan.path == 3
                              else {"transform(cuFFT)": transform_ms_per_step, "own_kernels": fused_ms_per_step}),
            "pipeline": {"bytes_per_field": pipeline_bytes, "achieved": pipeline_bytes * B / (ms * 1e-3) / 1e9,
                         "frac": pipeline_bytes * B / (ms * 1e-3) / 1e9 / peak}}

    # --- CPU baseline: numpy port of the reference on a bounded sample of the same workload; also a parity check
    cores = len(os.sched_getaffinity(0))
    sample = args.cpu_fields if args.cpu_fields > 0 else max(cores, 8)
    cpu_rate, res = cpu_scatter_baseline(n, cells, sample, cores)
    gI = last["out"]["I"][:sample].cpu().numpy()
    gp, gn = last["out"]["peak_idx"][:sample].cpu().numpy(), last["out"]["n_peaks"][:sample].cpu().numpy()
    parity = True
    for b, (_, qI, pk) in enumerate(res):
        tol = 1e-9 * np.abs(qI[:-1, 1]) + 1e-12 * np.abs(qI[:, 1]).max()
        # the CPU sample is generated with libm, the GPU batch with CUDA sin/cos (1e-15 apart): 1e-8 check here,
        # the bit-for-bit-input parity tests live in tests/
        parity &= bool(np.array_equal(qI[:, 0], last["out"]["q"]))
        parity &= bool(np.all(np.abs(gI[b, :-1] - qI[:-1, 1]) <= 10 * tol + 1e-8 * np.abs(qI[:-1, 1])))
        parity &= bool(np.array_equal(gp[b, :gn[b]], pk))
    out = {
        "metric": "density fields/sec (I(q)+peaks)", "value": B * world / (ms * 1e-3), "unit": "fields/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{B} synthetic 3-component gyroid/BCC {n}^3 fp64 fields per GPU, scatter I(q) cutoff=30 + show_peak",
                   "grid": [n, n, n], "fields_per_gpu": B, "path": pathname, "fields_per_launch": min(args.max_batch, B), "host_chunk": args.host_batch,
                   "l2": "inputs (4.3 GB per pass) exceed the 126 MB L2; no flush needed", "parallelism": f"fields sharded, {world} rank(s), no collective"},
        "e2e": {"value": B * world / (e2e_ms * 1e-3), "unit": "fields/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms, "h2d_GBps": h2d / (e2e_ms * 1e-3) / 1e9, "numa_binding_rank0": numa, "api": "Scatter.sacttering_peak_batch(pinned host tensor) -> sbx_scatter_execute_host"},
        "gpu_launches": int(round(own_per_step * args.steps)),
        "roofline": roof,
        "cpu_baseline": {"value": cpu_rate, "unit": "fields/s", "cores": cores, "kind": "port", "txyz": "vectorised",
                         "sample": f"{sample} fields of the same generator through oracle/scatter_np.scattering_peak + show_peak, one field per worker process",
                         "note": "per-field rate of a bounded sample; the port builds txyz vectorised, ~3.6x faster per core than the reference's itertools.product (SURVEY.md 6), so GPU/CPU ratios are conservative"},
        "parity_vs_cpu_sample": parity,
        "clocks": clocks,
    }
    return out


# ----------------------------------------------------------------------------- voronoi workload
def run_voronoi(args):
    import torch

    from simubox_b200 import VoronoiCell, synth

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n = args.vgrid
    shape, box = (n, n, n), (float(n),) * 3
    per_axis = 16
    grids = args.vgrids  # independent grids per GPU per step (different seed sets, same mask)
    seeds = [torch.from_numpy(synth.seeds_lattice(shape, per_axis, 4.0, 1234 + rank * grids + i)).to(dev) for i in range(grids)]
    mask = (synth.fields_torch(1, 1, shape, cells=4, device=dev)[0] > 0.5).to(torch.uint8).contiguous()
    flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)  # 1 GiB >> 126 MB L2; also keeps the GPU busy while the host queues the step
    last = {}

    def step():
        for s in seeds:
            last["lab"] = VoronoiCell.label_grid(s, 0, shape=shape, box=box, mask=mask, device=local, return_tensor=True)

    def timed_labelling(steps, warmup):
        """Every step: flush L2 (untimed), then the labelling between two CUDA events on the launching stream.
        The flush runs ~0.3 ms, so the host has queued the whole step before the GPU reaches the first event."""
        for _ in range(warmup):
            flush.zero_()
            step()
        barrier(world)
        pairs = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step()
            e1.record()
            pairs.append((e0, e1))
        torch.cuda.synchronize()
        ms_ = sum(a.elapsed_time(b) for a, b in pairs) / steps
        barrier(world)
        return max_over_ranks(ms_, world, dev)

    with ClockSampler(local) as clk:
        ms = timed_labelling(args.steps, args.warmup)
    clocks = clk.summary()
    lab, st = VoronoiCell.label_grid(seeds[0], 0, shape=shape, box=box, mask=mask, device=local, return_tensor=True, stats=True)
    live = int(mask.sum().item())

    # end to end: seeds + mask on the host (pinned), labels back to the host
    h_seeds = [s.cpu().pin_memory() for s in seeds]
    h_mask = mask.cpu().pin_memory()
    h_lab = torch.empty(shape, dtype=torch.int32, pin_memory=True)

    def step_host():
        for s in h_seeds:
            d = VoronoiCell.label_grid(s.to(dev, non_blocking=True), 0, shape=shape, box=box,
                                       mask=h_mask.to(dev, non_blocking=True), device=local, return_tensor=True)
            h_lab.copy_(d, non_blocking=True)
        torch.cuda.synchronize()

    step_host()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        step_host()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.e2e_steps, world, dev)
    del flush
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = measured_peak_gbs()
    S = per_axis ** 3
    vox = n ** 3
    unit_bytes = 4 * vox + S * 32 + vox  # int32 labels + seed table + uint8 mask
    achieved = unit_bytes * grids / (ms * 1e-3) / 1e9
    cores = len(os.sched_getaffinity(0))
    cpu_rate, cvox, clab = cpu_voronoi_baseline(seeds[0].cpu().numpy(), shape, box, args.cpu_voxels, cores)
    full = VoronoiCell.label_grid(seeds[0], 0, shape=shape, box=box, device=local)
    parity = bool(np.array_equal(full[cvox[:, 0], cvox[:, 1], cvox[:, 2]], clab))
    out = {
        "metric": "Voronoi voxels/sec", "value": vox * grids * world / (ms * 1e-3), "unit": "voxels/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"periodic Voronoi labelling of {n}^3 grid, {S} jittered-lattice seeds, uint8 mask; {grids} grid(s) per GPU per step",
                   "l2": "1 GiB buffer written before every step, outside the CUDA-event pair that times the labelling", "parallelism": f"grids sharded, {world} rank(s), no collective"},
        "e2e": {"value": vox * grids * world / (e2e_ms * 1e-3), "unit": "voxels/s", "h2d_bytes_per_step": grids * (S * 24 + vox),
                "d2h_bytes_per_step": grids * vox * 4, "ms_per_step": e2e_ms, "api": "VoronoiCell.label_grid(host seeds, host mask) -> host labels"},
        "gpu_launches": 3 * grids * args.steps,
        "roofline": {"bound": "hbm", "kernel": "coarse_kernel+mid_kernel+label_kernel", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic(f"coarse_kernel+mid_kernel+label_kernel|{n}^3|{S} seeds"),
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_grid": unit_bytes, "evaluations_per_live_voxel": st["evaluations"] / max(live, 1),
                     "overflow_tiles": st["overflow_tiles"], "path": st.get("path"),
                     "note": "issue/latency bound, not HBM bound (DESIGN.md section 4): three launches, fp64 argmin over ~3.3 candidates per voxel"},
        "cpu_baseline": {"value": cpu_rate, "unit": "voxels/s", "cores": cores, "kind": "port",
                         "sample": f"{args.cpu_voxels} random voxels of the same grid through oracle/voronoi_np.label_voxels (27 images x {S} seeds per voxel)"},
        "parity_vs_cpu_sample": parity, "clocks": clocks,
    }
    return out


# ----------------------------------------------------------------------------- one giant field, slab decomposed
def _slab_mode(args, mode, local_slab, ctx, env=None):
    """One mode of the slab-decomposed scatter ('pruned' or 'fft') on this rank's planes: ms per step (CUDA events, max
    over ranks), stage split averaged over >= 10 profiled steps each started at a barrier, the rows."""
    import torch

    from simubox_b200 import distributed as D, synth

    rank, world, local = ctx
    dev = torch.device("cuda", local)
    n = args.grid
    env = dict(env or {}, SBX_SLAB_PATH=mode)
    os.environ.update(env)  # read by the library at plan creation
    slab = D.SlabScatter((n, n, n), synth.box_for(args.cells), cutoff=30, device=local)
    for k in env:
        os.environ.pop(k, None)
    out = torch.empty(slab.n_shells, dtype=torch.float64, device=dev)

    def step():
        slab.execute(local_slab, out=out)

    with ClockSampler(local) as clk:
        step()
        ms = timed_steps(step, args.steps, args.warmup, world, dev)
    slab.set_profiling(True)
    stages = []
    for _ in range(max(10, args.steps)):
        barrier(world)  # every rank enters the profiled step together: no rank skew inside the collective's stage
        step()
        torch.cuda.synchronize()
        stages.append(slab.profile())
    slab.set_profiling(False)
    prof = {k: float(np.mean([st[k] for st in stages])) for k in stages[0]}
    prof = {k: max_over_ranks(v, world, dev) for k, v in prof.items()}
    I = out.cpu().numpy()
    rec = {"path": "pruned_dft" if slab.path == 4 else "slab_fft", "exchange": slab.exchange, "ms_per_step": ms, "stage_ms": prof, "kernel": slab.kernel,
           "n_shells": int(slab.n_shells), "clocks": clk.summary()}
    q = slab.q.copy()
    slab.close()
    torch.cuda.empty_cache()
    return rec, I, q


def run_slab(args, modes=("pruned", "fft")):
    """BASELINE config 5: ONE --grid^3 fp64 field, N0/world planes per rank.  Both decompositions behind
    sbx_scatter_slab_execute are run through the real NCCL communicator: 'pruned' (each rank reduces its planes to the
    mode sums, one small all-reduce) and 'fft' (cuFFT 2D -> NCCL all-to-all -> cuFFT 1D -> fused histogram ->
    all-reduce, what north_star describes).  Strong scaling: the field is fixed, ranks split it."""
    import torch

    from simubox_b200 import distributed as D, synth

    ctx = dist_setup(args.gpus)
    rank, world, local = ctx
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n = args.grid
    lo, hi = D.slab_planes(n, rank, world)
    # this rank's planes of the synthetic field, generated on the device plane block by plane block
    local_slab = torch.empty((hi - lo, n, n), dtype=torch.float64, device=dev)
    for x0 in range(lo, hi, 32):
        x1 = min(hi, x0 + 32)
        local_slab[x0 - lo:x1 - lo] = synth.field_planes_torch(3, (n, n, n), x0, x1, cells=args.cells, device=dev)
    # "fft" runs the default exchange (fused pack + push into the peers' buffers over NVLink when CUDA IPC maps them,
    # else NCCL send/recv).  The other variants are the exchange-bound form of the mode, where every column of the half
    # spectrum travels (what a cutoff reaching the Nyquist column needs): fused / NCCL chunked / NCCL unchunked.
    variants = {"pruned": ("pruned", {}), "fft": ("fft", {}), "fft_allcols": ("fft", {"SBX_SLAB_ALLCOLS": "1"}),
                "fft_allcols_nccl": ("fft", {"SBX_SLAB_ALLCOLS": "1", "SBX_SLAB_EXCHANGE": "nccl"}),
                "fft_allcols_nccl_unchunked": ("fft", {"SBX_SLAB_ALLCOLS": "1", "SBX_SLAB_EXCHANGE": "nccl", "SBX_SLAB_CHUNKS": "1"})}
    names = list(modes) + ([v for v in variants if v.startswith("fft_allcols")] if "fft" in modes and world > 1 else [])
    if getattr(args, "slab_variants", ""):
        names = [v for v in args.slab_variants.split(",") if v in variants]
    modes = tuple(names)
    recs, rows = {}, {}
    for name in names:
        recs[name], rows[name], q = _slab_mode(args, variants[name][0], local_slab, ctx, variants[name][1])
    del local_slab
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = measured_peak_gbs()
    spec = 16 * n * n * (n // 2 + 1)
    out = {"metric": "density fields/sec (I(q), one slab-decomposed field)", "unit": "fields/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"one {n}^3 fp64 field, {world} slab(s) of {hi - lo} planes, cutoff=30", "grid": [n, n, n],
                      "l2": f"{8 * n ** 3 / world / 1e9:.1f} GB per rank per pass exceed the 126 MB L2"},
           "peak": peak, "peak_source": peak_src}
    for mode, rec in recs.items():
        pruned = rec["path"] == "pruned_dft"
        # pruned DFT: the field is read once.  slab FFT: field + write/read of the half spectrum through the 2D stage,
        # the exchange buffers and the 1D stage, then the scan
        bytes_field = 8 * n ** 3 if pruned else 8 * n ** 3 + 6 * spec
        rec["fields_per_s"] = 1.0 / (rec["ms_per_step"] * 1e-3)
        rec["roofline"] = {"bound": "hbm", "kernel": (rec["kernel"] or "pruned_blocked_kernel") if pruned else "cuFFT 2D + slab_pack_kernel + cuFFT 1D + shell_histogram_kernel",
                           "achieved": bytes_field / world / (rec["ms_per_step"] * 1e-3) / 1e9, "unit": "GB/s",
                           "frac": bytes_field / world / (rec["ms_per_step"] * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_field": bytes_field}
        rec["parallelism"] = ("all-reduce of the mode sums (no all-to-all)" if pruned else
                              ("pack kernel stores into the peers' buffers over NVLink (CUDA IPC) + 4-byte barrier all-reduce + all-reduce of the shell sums"
                               if rec.get("exchange") == "p2p_fused" else "NCCL all-to-all (send/recv group) + all-reduce"))
        if not pruned and world > 1 and rec["stage_ms"].get("all_to_all_ms"):
            allc = mode.startswith("fft_allcols")
            rec["columns_exchanged"] = "all (N2/2+1)" if allc else "only c with kv2[c] <= largest listed k^2"
            if rec.get("exchange") == "p2p_fused":
                rec["stage_note"] = "fused exchange: the transfer is inside stage A (pack kernel stores into the peers); 'all_to_all_ms' is the 4-byte arrival barrier"
            elif allc:
                sent = spec / world * (world - 1) / world  # bytes each GPU sends (and receives)
                rec["all_to_all_GBps_per_gpu"] = sent / (rec["stage_ms"]["all_to_all_ms"] * 1e-3) / 1e9
            # stage_ms comes from serialised (profiled) steps; ms_per_step from the pipelined ones
            serial = sum(rec["stage_ms"].values())
            rec["serial_stage_sum_ms"] = serial
            if rec.get("exchange") != "p2p_fused":
                rec["all_to_all_hidden_frac"] = max(0.0, min(1.0, (serial - rec["ms_per_step"]) / rec["stage_ms"]["all_to_all_ms"]))
    if "pruned" in rows and "fft" in rows:
        a, b = rows["pruned"], rows["fft"]
        full = len(a) - 1
        tol = 1e-9 * np.abs(a[:full]) + 1e-12 * np.abs(a[1:]).max()
        out["slab_parity"] = bool(np.all(np.abs(a[:full] - b[:full]) <= tol))
        out["slab_parity_worst_rel"] = float(np.max(np.abs(a[:full] - b[:full]) / np.maximum(np.abs(a[:full]), 1e-300)))
        for extra in ("fft_allcols", "fft_allcols_nccl", "fft_allcols_nccl_unchunked"):
            if extra in rows:
                out["slab_parity"] = bool(out["slab_parity"] and np.all(np.abs(a[:full] - rows[extra][:full]) <= tol))
    first = recs[modes[0]]
    out.update(value=first["fields_per_s"], ms_per_step=first["ms_per_step"], modes=recs, dc_row=float(rows[modes[0]][0]),
               gpu_launches=(3 if first["path"] == "pruned_dft" else 2) * args.steps, clocks=first["clocks"],
               roofline=dict(first["roofline"], peak=peak, peak_source=peak_src, traffic=measured_traffic(f"{first['roofline']['kernel']}|{n}^3|{world} slab(s)")))
    return out


# ----------------------------------------------------------------------------- one Voronoi grid across GPUs (voxel slabs)
def run_voronoi_sharded(args, grids=((256, 16), (512, 32))):
    """ONE grid labelled by voxel slab, seeds replicated, no collective; strong scaling against the
    single-launch labelling of the whole grid on one GPU.  Each rank checks its slab against its planes of a whole-grid
    labelling (bit for bit)."""
    import torch

    from simubox_b200 import VoronoiCell, distributed as D, synth

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    flush = torch.empty(1 << 29, dtype=torch.uint8, device=dev)
    recs = []
    for n, per_axis in grids:
        shape, box = (n, n, n), (float(n),) * 3
        pts = torch.from_numpy(synth.seeds_lattice(shape, per_axis, 4.0, 1234)).to(dev)
        mask = (synth.fields_torch(1, 1, shape, cells=max(1, n // 64), device=dev)[0] > 0.5).to(torch.uint8).contiguous()
        lo, hi = D.shard_range(n, rank, world)
        mslab = mask[lo:hi].contiguous()

        def timed(fn):
            for _ in range(3):
                flush.zero_(); fn()
            barrier(world)
            pairs = []
            for _ in range(args.steps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record()
                pairs.append((e0, e1))
            torch.cuda.synchronize()
            t = float(np.median([x.elapsed_time(y) for x, y in pairs]))
            barrier(world)
            return max_over_ranks(t, world, dev)

        last = {}
        ms = timed(lambda: last.__setitem__("slab", VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mslab, device=local, return_tensor=True, planes=(lo, hi))))
        whole_ms = timed(lambda: last.__setitem__("whole", VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, device=local, return_tensor=True)))
        ok = torch.equal(last["slab"], last["whole"][lo:hi])
        ok_all = max_over_ranks(0.0 if ok else 1.0, world, dev) == 0.0
        recs.append({"grid": [n, n, n], "seeds": per_axis ** 3, "planes_per_rank": hi - lo, "ms_per_grid": ms, "voxels_per_s": n ** 3 / (ms * 1e-3),
                     "one_gpu_whole_grid_ms": whole_ms, "speedup_vs_one_gpu": whole_ms / ms, "slabs_equal_whole_grid": bool(ok_all)})
        del mask, mslab, last
        torch.cuda.empty_cache()
    del flush
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    return {"metric": "Voronoi voxels/sec, one grid sharded by voxel slab", "unit": "voxels/s", "n_gpus": world, "scaling": "strong",
            "parallelism": f"planes [r N0/P, (r+1) N0/P) on rank r of {world}, seeds replicated, no collective (sbx_voronoi_plan_execute_range)",
            "l2": "512 MiB buffer written before every timed labelling", "steps": args.steps, "grids": recs}


# ----------------------------------------------------------------------------- phase-diagram sweep (config 4)
def run_sweep(args):
    """BASELINE config 4: a sweep of 256^3 fields sharded over the GPUs; per field scatter I(q)+peaks AND periodic
    Voronoi labelling of its masked grid with 4096 domain-centre seeds.  --sweep-fields = fields per GPU (weak scaling;
    256 per GPU = the 2048 fields of the config on 8 GPUs)."""
    import torch

    from simubox_b200 import Scatter, VoronoiCell, synth

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n, cells, B = args.vgrid, 4, args.sweep_fields
    shape, L = (n, n, n), synth.box_for(cells)
    fields = torch.empty((B, *shape), dtype=torch.float64, device=dev)
    masks = torch.empty((B, *shape), dtype=torch.uint8, device=dev)
    for b in range(B):
        fields[b] = synth.fields_torch(rank * B + b, 1, shape, cells=cells, device=dev)[0]
        masks[b] = (fields[b] > 0.5).to(torch.uint8)
    # every field has its own seed set (default_rng(1234 + i))
    seeds = torch.from_numpy(np.stack([synth.seeds_lattice(shape, 16, 4.0, 1234 + rank * B + b) for b in range(B)])).to(dev)
    kw = dict(cutoff=30, threshold=1e-12, peaks=True, width=PEAK_KW["width"], step=PEAK_KW["step"], peak_cutoff=PEAK_KW["cutoff"],
              min_height=PEAK_KW["min_height"], max_batch=min(32, B), device=local)
    box = (float(n),) * 3
    last = {}
    checks = torch.zeros(B, dtype=torch.int64, device=dev)

    # the labellings are latency-bound launches of partial waves: several in flight on their own streams (one plan
    # workspace each, plan_slot) overlap; the scatter batch runs on the main stream meanwhile
    n_streams = max(1, args.sweep_streams)
    side = [torch.cuda.Stream(device=dev) for _ in range(n_streams)] if n_streams > 1 else []

    def step():
        main = torch.cuda.current_stream(dev)
        for s_ in side:
            s_.wait_stream(main)
        last["scatter"] = Scatter.sacttering_peak_batch(fields, L, **kw)
        for b in range(B):
            if side:
                with torch.cuda.stream(side[b % n_streams]):
                    lab = VoronoiCell.label_grid(seeds[b], 0, shape=shape, box=box, mask=masks[b], device=local, return_tensor=True,
                                                 plan_slot=b % n_streams)
            else:
                lab = VoronoiCell.label_grid(seeds[b], 0, shape=shape, box=box, mask=masks[b], device=local, return_tensor=True)
            if b == B - 1:
                last["lab"] = lab
        for s_ in side:
            main.wait_stream(s_)

    steps = max(2, min(args.steps, 5))
    with ClockSampler(local) as clk:
        ms = timed_steps(step, steps, 3, world, dev)
    checks[0] = last["lab"].to(torch.int64).sum()  # nothing large leaves the device
    del fields, masks
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = measured_peak_gbs()
    bytes_field = 8 * n ** 3 + (4 * n ** 3 + 32 * 4096 + n ** 3)  # field read once + labels, seeds, mask
    return {
        "metric": "sweep fields/sec (I(q)+peaks and Voronoi labelling per field)", "value": B * world / (ms * 1e-3), "unit": "fields/s",
        "n_gpus": world, "steps": steps, "warmup": 3, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{B} synthetic {n}^3 fields per GPU ({B * world} in all), own seed set per field: scatter cutoff=30 + show_peak, and periodic Voronoi labels (4096 seeds, uint8 mask) per field",
                   "l2": f"{B * 8 * n ** 3 / 1e9:.1f} GB of fields per pass exceed the 126 MB L2", "parallelism": f"fields sharded, {world} rank(s), no collective",
                   "streams": f"scatter batch on the main stream, labellings round-robin on {n_streams} side stream(s)"},
        "roofline": {"bound": "hbm", "kernel": "pruned_ring_kernel + coarse/mid/label kernels", "achieved": bytes_field * B / (ms * 1e-3) / 1e9,
                     "peak": peak, "unit": "GB/s", "frac": bytes_field * B / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_field": bytes_field},
        "label_checksum_last_field": int(checks[0].item()),
        "gpu_launches": steps * (-(-B // min(32, B)) + 3 * B), "clocks": clk.summary(),
    }


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's own CPU path (numpy port: oracle/) on this host's cores; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = len(os.sched_getaffinity(0))
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    if args.workload == "voronoi":
        from simubox_b200 import synth

        n = args.vgrid
        pts = synth.seeds_lattice((n, n, n), 16, 4.0, 1234)
        rates = [cpu_voronoi_baseline(pts, (n, n, n), (float(n),) * 3, args.cpu_voxels, cores)[0] for _ in range(warm + steps)][warm:]
        rate, unit, metric = float(np.mean(rates)), "voxels/s", "Voronoi voxels/sec"
        sample = f"{args.cpu_voxels} random voxels of the {n}^3 / 4096-seed periodic grid per step"
        workload = f"periodic Voronoi labelling of {n}^3 grid, 4096 jittered-lattice seeds"
        per_step = args.cpu_voxels
    else:
        sample_n = args.cpu_fields if args.cpu_fields > 0 else max(cores, 8)
        rates = [cpu_scatter_baseline(args.grid, args.cells, sample_n, cores)[0] for _ in range(warm + steps)][warm:]
        rate, unit, metric = float(np.mean(rates)), "fields/s", "density fields/sec (I(q)+peaks)"
        sample = f"{sample_n} fields of the same generator per step, one field per worker process"
        workload = f"{args.fields} synthetic 3-component gyroid/BCC {args.grid}^3 fp64 fields per GPU, scatter I(q) cutoff=30 + show_peak"
        per_step = sample_n
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": rate, "unit": unit, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": per_step / rate * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": {"workload": workload, "bounded_sample": sample},
        "cpu_baseline": {"value": rate, "unit": unit, "cores": cores, "kind": "port", "txyz": "vectorised", "sample": sample,
                         "note": "same generator and per-unit work as the GPU arm's config, timed on a bounded sample per step (per-unit rate); numpy port of the reference with a vectorised txyz (~3.6x faster per core than the reference itself)"},
        "e2e": {"value": rate, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="scatter", choices=["scatter", "voronoi", "slab", "sweep"])
    ap.add_argument("--path", default="auto", choices=["auto", "dense", "gather", "pruned", "fused"])
    ap.add_argument("--grid", type=int, default=128)
    ap.add_argument("--cells", type=int, default=2)
    ap.add_argument("--fields", type=int, default=256)
    ap.add_argument("--max-batch", type=int, default=256, help="fields per kernel launch (device-resident path)")
    ap.add_argument("--host-batch", type=int, default=16, help="fields per H2D chunk of the host-streaming path")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-fields", type=int, default=0)
    ap.add_argument("--vgrid", type=int, default=256)
    ap.add_argument("--vgrids", type=int, default=1)
    ap.add_argument("--cpu-voxels", type=int, default=4096)
    ap.add_argument("--sweep-fields", type=int, default=256, help="256^3 fields per GPU of the sweep record (config 4: 2048 over 8 GPUs)")
    ap.add_argument("--sweep-streams", type=int, default=4, help="side streams the sweep's labellings are issued on (1 = all on the main stream)")
    ap.add_argument("--slab-grid", type=int, default=1024, help="grid of the one-field slab record (config 5)")
    ap.add_argument("--slab-variants", default="", help="--workload slab: comma list of pruned,fft,fft_allcols,fft_allcols_nccl,fft_allcols_nccl_unchunked")
    ap.add_argument("--sub", default="all", help="sub-records of the default line: all | none | comma list of voronoi,sweep,slab,voronoi_sharded")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    from simubox_b200 import build

    build.build()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload == "voronoi":
        rec = run_voronoi(args)
    elif args.workload == "slab":
        rec = run_slab(args)
    elif args.workload == "sweep":
        rec = run_sweep(args)
    else:
        rec = run_scatter(args)
        want = {"voronoi", "sweep", "slab", "voronoi_sharded"} if args.sub == "all" else set(filter(None, args.sub.split(","))) - {"none"}
        subs = {}
        sub_args = argparse.Namespace(**vars(args))
        sub_args.steps, sub_args.e2e_steps = min(args.steps, 10), min(args.e2e_steps, 2)

        def leg(name, fn):
            # a failing leg must not take the headline down with it: the error is recorded in its place
            t0 = time.time()
            try:
                r = fn()
            except Exception as e:  # noqa: BLE001
                r = {"error": f"{type(e).__name__}: {e}"[:300]}
            if isinstance(r, dict):
                r["leg_seconds"] = round(time.time() - t0, 1)
            subs[name] = r

        if "voronoi" in want:
            leg("voronoi", lambda: run_voronoi(sub_args))
        if "voronoi_sharded" in want and world > 1:
            leg("voronoi_sharded", lambda: run_voronoi_sharded(sub_args))
        if "slab" in want and world > 1:
            slab_args = argparse.Namespace(**vars(sub_args))
            slab_args.grid, slab_args.cells = args.slab_grid, 8
            leg("slab", lambda: run_slab(slab_args))
        if "sweep" in want:
            leg("sweep", lambda: run_sweep(sub_args))
        if rec is not None:
            rec.update({k: v for k, v in subs.items() if v is not None})
    if rec is not None:
        print(json.dumps(rec))
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
