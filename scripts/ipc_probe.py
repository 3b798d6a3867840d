"""Does CUDA IPC work between the ranks' processes on this box?  (torchrun, 2 ranks.)  Each rank cudaMalloc's a buffer,
exports its IPC handle, opens the peer's, writes a pattern into the PEER's buffer with a device-to-device copy, and
after a barrier checks what the peer wrote into its own."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
from cuda.bindings import runtime as rt

local = int(os.environ["LOCAL_RANK"]); rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 1 << 20
err, mine = rt.cudaMalloc(n * 4); assert err == rt.cudaError_t.cudaSuccess, err
err, h = rt.cudaIpcGetMemHandle(mine); assert err == rt.cudaError_t.cudaSuccess, err
blob = bytes(h.reserved)
blobs = [None] * world
dist.all_gather_object(blobs, blob)
peer = (rank + 1) % world
ph = rt.cudaIpcMemHandle_t(); ph.reserved = blobs[peer]
err, pptr = rt.cudaIpcOpenMemHandle(ph, rt.cudaIpcMemLazyEnablePeerAccess)
print(f"rank {rank}: open peer handle -> {err}", flush=True)
assert err == rt.cudaError_t.cudaSuccess, err
src = torch.full((n,), rank + 100, dtype=torch.int32, device="cuda")
err, = rt.cudaMemcpy(pptr, src.data_ptr(), n * 4, rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice); assert err == rt.cudaError_t.cudaSuccess, err
torch.cuda.synchronize(); dist.barrier()
out = torch.empty(n, dtype=torch.int32, device="cuda")
rt.cudaMemcpy(out.data_ptr(), mine, n * 4, rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice)
writer = (rank - 1) % world
ok = bool((out == writer + 100).all().item())
print(f"rank {rank}: peer wrote {int(out[0])}, expected {writer + 100}: {'IPC OK' if ok else 'IPC BROKEN'}", flush=True)
# bandwidth of a peer copy (copy engine over NVLink)
big = 256 << 20
err, a = rt.cudaMalloc(big); err, hb = rt.cudaIpcGetMemHandle(a)
bl = [None] * world; dist.all_gather_object(bl, bytes(hb.reserved))
pb = rt.cudaIpcMemHandle_t(); pb.reserved = bl[peer]
err, pa = rt.cudaIpcOpenMemHandle(pb, rt.cudaIpcMemLazyEnablePeerAccess)
s = torch.empty(big, dtype=torch.uint8, device="cuda")
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
st = torch.cuda.current_stream().cuda_stream
e0.record()
for _ in range(8):
    rt.cudaMemcpyAsync(pa, s.data_ptr(), big, rt.cudaMemcpyKind.cudaMemcpyDeviceToDevice, st)
e1.record(); torch.cuda.synchronize()
print(f"rank {rank}: peer copy {8 * big / (e0.elapsed_time(e1) * 1e-3) / 1e9:.0f} GB/s", flush=True)
dist.barrier()
rt.cudaIpcCloseMemHandle(pptr); rt.cudaIpcCloseMemHandle(pa)
dist.destroy_process_group()
