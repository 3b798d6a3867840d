"""Concurrent host -> device copy bandwidth per GPU with 1 ... N ranks copying at once (run under torchrun):
pinned 1 GiB per rank, CUDA events on the copy stream, barrier before every timed burst.  Variants: default CPU
affinity / bound to the GPU's NUMA-local CPUs before the pinned allocation; one / two copy streams.
Prints one JSON line on rank 0.  Names the saturated link for DESIGN.md section 5."""
import json, os, sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from simubox_b200.distributed import bind_to_gpu_numa, gpu_numa_cpus

local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
GiB = 1 << 30
out = {"world": world, "variants": {}}
cpus, node = gpu_numa_cpus(local)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def gather(x):
    if world == 1:
        return [x]
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    return [float(p.item()) for p in parts]


def measure(streams):
    host = torch.empty(GiB, dtype=torch.uint8, pin_memory=True)
    host.fill_(1)
    d = torch.empty(GiB, dtype=torch.uint8, device=dev)
    ss = [torch.cuda.Stream() for _ in range(streams)]
    part = GiB // streams
    def burst():
        for i, s in enumerate(ss):
            with torch.cuda.stream(s):
                d[i * part:(i + 1) * part].copy_(host[i * part:(i + 1) * part], non_blocking=True)
    burst(); torch.cuda.synchronize()
    best = 0.0
    for _ in range(5):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in ss:
            s.wait_event(e0)
        for _ in range(4):
            burst()
        for s in ss:
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, 4 * GiB / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    del host, d
    return best


for name, bind, streams in (("default_1stream", False, 1), ("default_2streams", False, 2), ("numa_bound_1stream", True, 1), ("numa_bound_2streams", True, 2)):
    info = bind_to_gpu_numa(local) if bind else {}
    g = gather(measure(streams))
    out["variants"][name] = {"per_gpu_GBps": [round(x, 2) for x in g], "sum_GBps": round(sum(g), 1), "min_GBps": round(min(g), 2)}
    if bind:
        out["variants"][name]["binding_rank0"] = info
nodes = gather(float(node))
out["gpu_numa_nodes"] = [int(x) for x in nodes]
out["local_cpus_rank0"] = len(cpus)
if int(os.environ.get("RANK", "0")) == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
