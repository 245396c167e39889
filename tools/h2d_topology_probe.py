#!/usr/bin/env python
"""Host -> device copy bandwidth per GPU, alone and together, and where the box puts the GPUs (PCIe switches, NUMA).

Run under torchrun with one rank per GPU (`python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1
tools/h2d_topology_probe.py`). It explains the `e2e` figure of `bench.py` at N > 1: the end-to-end path is bound by
the pinned-memory H2D copies, and those share PCIe uplinks and host DRAM between ranks. Nothing here is on the
product path; rank 0 prints JSON lines.
"""
import ctypes
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist

NBYTES = 1 << 30
REPS = 6


def gpu_numa(local):
    out = {}
    try:
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        out["bdf"] = bdf
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            out["numa_node"] = int(f.read().strip())
    except Exception as e:  # the probe reports what it could not read
        out["error"] = repr(e)
    return out


def set_mempolicy_bind(node):
    """set_mempolicy(MPOL_BIND, {node}) for this thread: later first-touch allocations land on `node`."""
    libc = ctypes.CDLL(None, use_errno=True)
    mask = ctypes.c_ulong(1 << node)
    rc = libc.syscall(238, 2, ctypes.byref(mask), ctypes.c_ulong(64))
    return rc, ctypes.get_errno()


def measure(host, dev):
    s = torch.cuda.current_stream()
    dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(REPS):
        dev.copy_(host, non_blocking=True)
    e1.record(s)
    torch.cuda.synchronize()
    return NBYTES * REPS / (e0.elapsed_time(e1) * 1e-3) / 1e9


def phase(name, active, host, dev, rank, world):
    """Ranks in `active` copy at the same time; returns the per-rank GB/s on every rank."""
    dist.barrier()
    torch.cuda.synchronize()
    v = measure(host, dev) if rank in active else 0.0
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    g = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(g, t)
    vals = [round(float(x.item()), 2) for x in g]
    if rank == 0:
        print(json.dumps({"phase": name, "active": sorted(active), "gbs": vals, "sum": round(sum(vals), 1)}), flush=True)
    return vals


def main():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    info = gpu_numa(local)
    info["affinity"] = len(os.sched_getaffinity(0))
    infos = [None] * world
    dist.all_gather_object(infos, info)
    if rank == 0:
        print(json.dumps({"gpus": infos}), flush=True)
        nodes = {}
        base = "/sys/devices/system/node"
        for d in sorted(os.listdir(base)) if os.path.isdir(base) else []:
            if d.startswith("node") and d[4:].isdigit():
                with open(f"{base}/{d}/cpulist") as f:
                    nodes[d] = f.read().strip()
        print(json.dumps({"numa_nodes": nodes, "cpu_count": os.cpu_count(),
                          "affinity": sorted(os.sched_getaffinity(0))[:4] + ["..."] + sorted(os.sched_getaffinity(0))[-2:]}),
              flush=True)
        try:
            print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=60).stdout, flush=True)
        except Exception as e:
            print("topo failed", e, flush=True)

    dev = torch.empty(NBYTES, dtype=torch.uint8, device="cuda")
    host = torch.empty(NBYTES, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    for r in range(world):
        phase(f"solo{r}", {r}, host, dev, rank, world)
    phase("all", set(range(world)), host, dev, rank, world)
    if world >= 2:
        for pair in ((0, 1), (0, 2), (0, 4), (2, 3), (4, 5), (6, 7)):
            if max(pair) < world:
                phase(f"pair{pair[0]}{pair[1]}", set(pair), host, dev, rank, world)
        if world >= 4:
            phase("even", set(range(0, world, 2)), host, dev, rank, world)
            phase("low_half", set(range(world // 2)), host, dev, rank, world)

    # Same copies out of memory bound to the GPU's own NUMA node (if the box tells us which and lets us).
    node = info.get("numa_node", -1)
    res = {"rank": rank, "node": node}
    if node is not None and node >= 0:
        res["set_mempolicy"] = set_mempolicy_bind(node)
    ress = [None] * world
    dist.all_gather_object(ress, res)
    if rank == 0:
        print(json.dumps({"mempolicy": ress}), flush=True)
    del host
    host2 = torch.empty(NBYTES, dtype=torch.uint8).pin_memory()
    host2.fill_(2)
    phase("all_after_mempolicy", set(range(world)), host2, dev, rank, world)
    # cudaHostRegister of first-touched anonymous memory (pages placed by this thread's policy, then pinned)
    plain = torch.empty(NBYTES, dtype=torch.uint8)
    plain.fill_(3)
    rc = torch.cuda.cudart().cudaHostRegister(plain.data_ptr(), NBYTES, 0)
    if rank == 0:
        print(json.dumps({"cudaHostRegister": int(rc)}), flush=True)
    phase("all_registered", set(range(world)), plain, dev, rank, world)
    phase("solo0_registered", {0}, plain, dev, rank, world)
    torch.cuda.cudart().cudaHostUnregister(plain.data_ptr())
    dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
