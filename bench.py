#!/usr/bin/env python
"""bench.py -- density fields/s through the scattering path (I(q) + peaks) on B200.

Contract: ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line (rank 0).
A step = one pass of the hot path over one batch of synthetic fields per GPU (default: 256
gyroid/BCC 128^3 fp64 fields, BASELINE.json config 2).  ``value`` = device-resident whole-job
fields/s; ``e2e`` = the same through the host-buffer C-ABI call (pinned host -> device copies
inside the timed region); ``roofline`` = the dominant hand-written kernel against the measured
HBM peak; ``cpu_baseline`` = the numpy port of the reference on this host's cores.
``--impl reference`` times that CPU port alone.  ``--workload voronoi`` reports voxels/s of the
256^3 / 4096-seed periodic masked labelling (BASELINE.json config 3) instead.
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
def dist_setup(gpus):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


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
    # --- same steps again with per-stage CUDA events (transform vs hand-written kernel split)
    plan.set_profiling(True)
    plan.profile(reset=True)
    timed_steps(step_device, args.steps, 1, world, dev)
    prof = plan.profile(reset=True)
    plan.set_profiling(False)
    own_per_step = prof["own_launches"] / (args.steps + 1)
    fused_ms_per_launch = prof["fused_ms"] / max(prof["own_launches"], 1)
    transform_ms_per_step = prof["transform_ms"] / (args.steps + 1)
    fused_ms_per_step = prof["fused_ms"] / (args.steps + 1)

    # --- end to end through the host-buffer C-ABI call (pinned host memory)
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
    clk.__exit__()
    clocks = clk.summary()
    ns = plan.n_shells
    h2d = B * n ** 3 * 8
    d2h = B * (ns * 8 + 4 + 256 * 4)
    assert np.array_equal(last["host"]["I"], last["out"]["I"].cpu().numpy()) or plan.path == 1, "host and device paths disagree"

    if rank != 0:
        return
    peak, peak_src = measured_peak_gbs()
    spec_bytes = 16 * n * n * (n // 2 + 1)
    pipeline_bytes = 8 * n ** 3 + 2 * spec_bytes
    pathname = _lib.PATH_NAMES[plan.path]
    if plan.path == 4:  # fused pruned DFT: ONE kernel reads each field exactly once and finishes it
        kern, unit_bytes = (plan.kernel or "pruned_fused_kernel"), 8 * n ** 3
        launches_per_step = own_per_step
        fused_ms_per_launch = prof["transform_ms"] / max(prof["own_launches"], 1)
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
                "ms_per_step": e2e_ms, "h2d_GBps": h2d / (e2e_ms * 1e-3) / 1e9, "api": "Scatter.sacttering_peak_batch(pinned host tensor) -> sbx_scatter_execute_host"},
        "gpu_launches": int(round(own_per_step * args.steps)),
        "roofline": roof,
        "cpu_baseline": {"value": cpu_rate, "unit": "fields/s", "cores": cores, "kind": "port",
                         "sample": f"{sample} fields of the same generator through oracle/scatter_np.scattering_peak + show_peak, one field per worker process"},
        "parity_vs_cpu_sample": parity,
        "clocks": clocks,
    }
    print(json.dumps(out))


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
    if rank != 0:
        return
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
                     "overflow_tiles": st["overflow_tiles"]},
        "cpu_baseline": {"value": cpu_rate, "unit": "voxels/s", "cores": cores, "kind": "port",
                         "sample": f"{args.cpu_voxels} random voxels of the same grid through oracle/voronoi_np.label_voxels (27 images x {S} seeds per voxel)"},
        "parity_vs_cpu_sample": parity, "clocks": clocks,
    }
    print(json.dumps(out))


# ----------------------------------------------------------------------------- one giant field, slab decomposed
def run_slab(args):
    """BASELINE config 5: ONE --grid^3 fp64 field, N0/world planes per rank, cuFFT 2D -> NCCL all-to-all ->
    cuFFT 1D -> fused histogram -> all-reduce.  Strong scaling: the field is fixed, ranks split it."""
    import torch

    from simubox_b200 import distributed as D, synth

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n = args.grid
    L = synth.box_for(args.cells)
    slab = D.SlabScatter((n, n, n), L, cutoff=30, device=local)
    lo, hi = slab.planes
    # this rank's planes of the synthetic field, generated on the device plane block by plane block
    local_slab = torch.empty((hi - lo, n, n), dtype=torch.float64, device=dev)
    for x0 in range(lo, hi, 32):
        x1 = min(hi, x0 + 32)
        local_slab[x0 - lo:x1 - lo] = synth.field_planes_torch(3, (n, n, n), x0, x1, cells=args.cells, device=dev)
    out = torch.empty(slab.n_shells, dtype=torch.float64, device=dev)

    def step():
        slab.execute(local_slab, out=out)

    with ClockSampler(local) as clk:
        step()
        ms = timed_steps(step, args.steps, args.warmup, world, dev)
        slab.set_profiling(True)
        step()
        torch.cuda.synchronize()
        prof = slab.profile()
        slab.set_profiling(False)
    I = out.cpu().numpy()
    if rank != 0:
        slab.close()
        return
    peak, peak_src = measured_peak_gbs()
    spec = 16 * n * n * (n // 2 + 1)
    pruned = slab.path == 4
    # pruned DFT: the field is read once.  slab FFT: field + write/read of the half spectrum through the 2D stage,
    # the exchange buffers and the 1D stage, then the scan
    bytes_field = 8 * n ** 3 if pruned else 8 * n ** 3 + 6 * spec
    kern = (slab.kernel or "pruned_blocked_kernel") if pruned else "cuFFT 2D + slab_pack_kernel + cuFFT 1D + shell_histogram_kernel"
    comm = "all-reduce of the mode sums (no all-to-all)" if pruned else "NCCL all-to-all (send/recv group) + all-reduce"
    print(json.dumps({
        "metric": "density fields/sec (I(q), one slab-decomposed field)", "value": 1.0 / (ms * 1e-3), "unit": "fields/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"one {n}^3 fp64 field, {world} slab(s) of {hi - lo} planes, cutoff=30", "grid": [n, n, n],
                   "l2": f"{8 * n ** 3 / world / 1e9:.1f} GB per rank per pass exceed the 126 MB L2",
                   "path": "pruned_dft_blocked" if pruned else "slab_fft", "parallelism": f"slab decomposition over {world} rank(s): {comm}"},
        "stage_ms": prof, "n_shells": int(slab.n_shells), "dc_row": float(I[0]),
        "roofline": {"bound": "hbm", "kernel": kern,
                     "achieved": bytes_field / world / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": bytes_field / world / (ms * 1e-3) / 1e9 / peak,
                     "traffic": measured_traffic(f"{kern}|{n}^3|{world} slab(s)"), "peak_source": peak_src,
                     "algorithmic_bytes_per_field": bytes_field},
        "gpu_launches": (3 if pruned else 2) * args.steps, "clocks": clk.summary(),
    }))
    slab.close()


# ----------------------------------------------------------------------------- phase-diagram sweep (config 4)
def run_sweep(args):
    """BASELINE config 4: a sweep of 256^3 fields sharded over the GPUs; per field scatter I(q)+peaks AND periodic
    Voronoi labelling of its masked grid with 4096 domain-centre seeds.  --fields = fields per GPU (weak scaling)."""
    import torch

    from simubox_b200 import Scatter, VoronoiCell, synth

    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n, cells, B = args.vgrid, 4, args.fields
    shape, L = (n, n, n), synth.box_for(cells)
    fields = torch.empty((B, *shape), dtype=torch.float64, device=dev)
    masks = torch.empty((B, *shape), dtype=torch.uint8, device=dev)
    for b in range(B):
        fields[b] = synth.fields_torch(rank * B + b, 1, shape, cells=cells, device=dev)[0]
        masks[b] = (fields[b] > 0.5).to(torch.uint8)
    seeds = [torch.from_numpy(synth.seeds_lattice(shape, 16, 4.0, 77 + rank * B + b)).to(dev) for b in range(min(B, 16))]
    kw = dict(cutoff=30, threshold=1e-12, peaks=True, width=PEAK_KW["width"], step=PEAK_KW["step"], peak_cutoff=PEAK_KW["cutoff"],
              min_height=PEAK_KW["min_height"], max_batch=min(32, B), device=local)
    box = (float(n),) * 3
    last = {}

    def step():
        last["scatter"] = Scatter.sacttering_peak_batch(fields, L, **kw)
        for b in range(B):
            last["lab"] = VoronoiCell.label_grid(seeds[b % len(seeds)], 0, shape=shape, box=box, mask=masks[b], device=local, return_tensor=True)

    with ClockSampler(local) as clk:
        ms = timed_steps(step, args.steps, args.warmup, world, dev)
    if rank != 0:
        return
    peak, peak_src = measured_peak_gbs()
    bytes_field = 8 * n ** 3 + (4 * n ** 3 + 32 * 4096 + n ** 3)  # field read once + labels, seeds, mask
    print(json.dumps({
        "metric": "sweep fields/sec (I(q)+peaks and Voronoi labelling per field)", "value": B * world / (ms * 1e-3), "unit": "fields/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{B} synthetic {n}^3 fields per GPU: scatter cutoff=30 + show_peak, and periodic Voronoi labels (4096 seeds, uint8 mask) per field",
                   "l2": f"{B * 8 * n ** 3 / 1e9:.1f} GB of fields per pass exceed the 126 MB L2", "parallelism": f"fields sharded, {world} rank(s), no collective"},
        "roofline": {"bound": "hbm", "kernel": "pruned_ring_kernel + coarse/mid/label kernels", "achieved": bytes_field * B / (ms * 1e-3) / 1e9,
                     "peak": peak, "unit": "GB/s", "frac": bytes_field * B / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_field": bytes_field},
        "gpu_launches": args.steps * (-(-B // min(32, B)) + 3 * B), "clocks": clk.summary(),
    }))


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
        "cpu_baseline": {"value": rate, "unit": unit, "cores": cores, "kind": "port", "sample": sample},
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
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    from simubox_b200 import build

    build.build()
    if args.workload == "voronoi":
        run_voronoi(args)
    elif args.workload == "slab":
        run_slab(args)
    elif args.workload == "sweep":
        run_sweep(args)
    else:
        run_scatter(args)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
