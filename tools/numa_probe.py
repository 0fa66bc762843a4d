#!/usr/bin/env python
"""Host-side placement probe for the multi-GPU `e2e` leg: where the container's CPUs and each GPU sit (NUMA), and what
N ranks copying 39 MB of page-locked host memory to their GPUs at once reach in aggregate, with the pinned buffer
allocated (a) wherever the rank happens to run, (b) after the rank has moved to the CPUs NVML reports as local to its
GPU.  usage: torchrun --nproc-per-node N tools/numa_probe.py"""
import glob
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist


def local_cpus(index):
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(index)
    n_words = (os.cpu_count() + 63) // 64
    mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
    cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
    bus = pynvml.nvmlDeviceGetPciInfo(h).busId
    bus = bus.decode() if isinstance(bus, bytes) else bus
    node = None
    for cand in (bus.lower(), bus.lower()[4:]):
        p = f"/sys/bus/pci/devices/{cand}/numa_node"
        if os.path.exists(p):
            node = int(open(p).read())
            break
    return cpus, node


def h2d_rate(buf_host, buf_dev, reps=30):
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        buf_dev.copy_(buf_host, non_blocking=True)
    torch.cuda.synchronize()
    dist.barrier()
    return reps * buf_host.numel() * buf_host.element_size() / (time.perf_counter() - t0) / 1e9


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {"rank": rank}
    allowed = os.sched_getaffinity(0)
    cpus, node = local_cpus(local)
    out.update(allowed=len(allowed), gpu_numa_node=node, gpu_local_cpus=len(cpus), usable_local=len(cpus & allowed))
    if rank == 0:
        nodes = {os.path.basename(p): open(p + "/cpulist").read().strip()
                 for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*"))}
        out["numa_nodes"] = nodes
        out["allowed_cpus"] = sorted(allowed)[:4] + ["...", max(allowed)]
        out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[-2500:]
    n = 39_321_600 // 4
    dev = torch.empty(n, dtype=torch.float32, device="cuda")
    a = torch.empty(n, dtype=torch.float32).pin_memory()
    a.fill_(1.0)
    out["as_launched_gbs"] = h2d_rate(a, dev)
    if cpus & allowed:
        os.sched_setaffinity(0, cpus & allowed)
        b = torch.empty(n, dtype=torch.float32)
        b.fill_(1.0)                       # first touch on the local node
        b = b.pin_memory()
        out["gpu_local_gbs"] = h2d_rate(b, dev)
        out["old_buffer_after_move_gbs"] = h2d_rate(a, dev)
    else:
        dist.barrier(); dist.barrier(); dist.barrier(); dist.barrier()
    gathered = [None] * world
    dist.all_gather_object(gathered, out)
    if rank == 0:
        for g in gathered:
            print(json.dumps(g))
        for k in ("as_launched_gbs", "gpu_local_gbs", "old_buffer_after_move_gbs"):
            print(k, "aggregate", round(sum(g.get(k, 0) for g in gathered), 1))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
