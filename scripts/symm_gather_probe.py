"""Probe: gather a per-rank device buffer into rank 0 with copy-engine peer writes through torch symmetric memory,
next to dist.gather (NCCL kernels).  torchrun --nproc-per-node N scripts/symm_gather_probe.py"""
import os
import sys
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.dup2(2, 1)
dist.init_process_group("nccl", device_id=dev)
n = 2 * 192 * 249216
src = torch.full((n,), rank + 1, dtype=torch.int32, device=dev)
import torch.distributed._symmetric_memory as symm_mem
buf = symm_mem.empty(world * n, dtype=torch.int32, device=dev)
hdl = symm_mem.rendezvous(buf, dist.group.WORLD.group_name)
root = hdl.get_buffer(0, (world * n,), torch.int32)
s = torch.cuda.Stream()
dist.barrier(); torch.cuda.synchronize()
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        root[rank * n:(rank + 1) * n].copy_(src, non_blocking=True)
        hdl.barrier(channel=0)
        e1.record(s)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ok = all(int(buf[r * n]) == r + 1 and int(buf[(r + 1) * n - 1]) == r + 1 for r in range(world))
        sys.stderr.write("symm peer-copy gather: %.3f ms  ok=%s\n" % (float(t), ok))
gl = [torch.empty_like(src) for _ in range(world)] if rank == 0 else None
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dist.gather(src, gl, dst=0); e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        sys.stderr.write("nccl gather: %.3f ms\n" % float(t))
dist.destroy_process_group()
