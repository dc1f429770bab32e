#!/usr/bin/env python
"""Host-link ceiling of a multi-GPU box: every rank copies the end-to-end leg's per-step volumes (56 MB up, 149 MB
down, pinned host memory, two streams) with no kernels in between, all ranks at once.  Compare with bench.py's e2e.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_ceiling.py -- [--bind-numa]
--bind-numa: bind the rank to the CPUs of its GPU's NUMA node BEFORE the pinned allocation (first-touch placement)."""
import os
import sys

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
node = -1
try:
    bus = torch.cuda.get_device_properties(lr).pci_bus_id
    dom = torch.cuda.get_device_properties(lr).pci_domain_id
    dev = torch.cuda.get_device_properties(lr).pci_device_id
    path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
    node = int(open(path).read())
except Exception as ex:  # noqa
    print(f"rank {rank}: numa node unknown ({ex})", file=sys.stderr)
if "--bind-numa" in sys.argv and node >= 0:
    cpus = []
    for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
        a, _, b = part.partition("-")
        cpus += list(range(int(a), int(b or a) + 1))
    allowed = sorted(set(cpus) & os.sched_getaffinity(0))
    if allowed:
        os.sched_setaffinity(0, allowed)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
UP, DOWN = 55987200, 149299200
h_up = torch.empty(UP, dtype=torch.uint8).pin_memory(); h_up.fill_(1)
h_dn = torch.empty(DOWN, dtype=torch.uint8).pin_memory(); h_dn.fill_(1)
d_up = torch.empty(UP, dtype=torch.uint8, device="cuda"); d_dn = torch.ones(DOWN, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def run(iters, up=True, down=True):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s_up.wait_event(e0); s_dn.wait_event(e0)
    for _ in range(iters):
        if up:
            with torch.cuda.stream(s_up):
                d_up.copy_(h_up, non_blocking=True)
        if down:
            with torch.cuda.stream(s_dn):
                h_dn.copy_(d_dn, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s_up); torch.cuda.current_stream().wait_stream(s_dn)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


run(3)
for name, up, down in (("up+down", True, True), ("up only", True, False), ("down only", False, True)):
    ms = run(20, up, down)
    if rank == 0:
        gb = ((UP if up else 0) + (DOWN if down else 0)) / ms / 1e6
        print(f"world {world} numa={'--numa' in sys.argv} {name}: {ms:.2f} ms per step-volume, {gb:.1f} GB/s per rank, "
              f"{gb * world:.1f} GB/s total; e2e ceiling from copies alone = {world * 24 / ms * 1000:.0f} frames/s")
print(f"rank {rank}: gpu numa node {node}, cpus {sorted(os.sched_getaffinity(0))[:4]}..({len(os.sched_getaffinity(0))})", file=sys.stderr)
if world > 1:
    dist.destroy_process_group()
