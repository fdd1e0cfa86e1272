"""Aggregate pinned H2D bandwidth with all ranks copying at once (torchrun), with and without binding each rank to
the CPUs local to its GPU before the pinned buffer is allocated.  Prints topology facts from rank 0."""
import os, subprocess, sys, time
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception as e:
        return repr(e)


def local_cpus():
    bus = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
    q = sh("nvidia-smi -i %d --query-gpu=pci.bus_id --format=csv,noheader" % local).lower()
    q = q[4:] if q.startswith("0000") and len(q) > 12 else q          # 00000000:1B:00.0 -> 0000:1b:00.0
    path = "/sys/bus/pci/devices/%s" % q
    node = sh("cat %s/numa_node" % path)
    cpul = sh("cat %s/local_cpulist" % path)
    return q, node, cpul


def parse_list(s):
    out = []
    for part in s.split(","):
        if "-" in part:
            a, b = part.split("-"); out += list(range(int(a), int(b) + 1))
        elif part.strip():
            out.append(int(part))
    return out


def measure(tag):
    nbytes = 123 * 1000 * 1000
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(40):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = 40 * nbytes / dt / 1e9
    t = torch.tensor([gbs], device=dev)
    if world > 1:
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        vals = [float(x.item()) for x in g]
    else:
        vals = [gbs]
    if rank == 0:
        print("%s: per-rank GB/s %s  aggregate %.1f" % (tag, ["%.1f" % v for v in vals], sum(vals)), flush=True)


bus, node, cpul = local_cpus()
allowed = sorted(os.sched_getaffinity(0))
print("rank %d gpu %s numa_node %s local_cpulist %s allowed_cpus %d..%d (%d)" % (rank, bus, node, cpul, allowed[0], allowed[-1], len(allowed)), flush=True)
if rank == 0:
    print(sh("nvidia-smi topo -m | head -14"))
    print(sh("ls /sys/devices/system/node/ | tr '\\n' ' '; grep -i 'mems_allowed_list\\|cpus_allowed_list' /proc/self/status"))
    print(sh("cat /sys/devices/system/node/node*/cpulist"))
measure("default")
want = [c for c in parse_list(cpul) if c in allowed] if cpul and cpul[0].isdigit() else []
if want:
    os.sched_setaffinity(0, want)
    measure("bound to GPU-local CPUs")
else:
    if rank == 0:
        print("no GPU-local CPUs inside the allowed set")
if world > 1:
    dist.destroy_process_group()
