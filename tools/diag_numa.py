"""8-GPU box diagnostics: topology, NUMA layout, and the aggregate pinned host->device bandwidth with one process per GPU
(what bounds the host-fed leg of bench.py at N GPUs).  Run under torchrun."""
import os, sys, time, glob, subprocess
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if rank == 0:
    print(subprocess.run("nvidia-smi topo -m | cut -c1-160; lscpu | grep -iE 'numa|socket|model name|^cpu\\(s\\)'; cat /sys/fs/cgroup/cpuset.cpus.effective /sys/fs/cgroup/cpuset.mems.effective 2>/dev/null; grep -i allowed /proc/self/status", shell=True, capture_output=True, text=True).stdout)
    for d in sorted(glob.glob("/sys/bus/pci/devices/*/numa_node")):
        try:
            cls = open(d.replace("numa_node", "class")).read().strip()
            if cls.startswith("0x0302") or cls.startswith("0x0300"):
                print(d.split("/")[-2], "numa", open(d).read().strip())
        except Exception:
            pass
    sys.stdout.flush()
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def bw(solo):
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    if solo is not None and rank != solo:
        if world > 1: dist.barrier()
        return None
    t0 = time.perf_counter()
    for _ in range(8): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    if world > 1: dist.barrier()
    return 8 * n / t / 1e9
all_bw = bw(None)
res = torch.tensor([all_bw], device="cuda")
if world > 1:
    g = [torch.zeros_like(res) for _ in range(world)]
    dist.all_gather(g, res)
else:
    g = [res]
solo = bw(0)
if rank == 0:
    print("H2D GB/s, all ranks at once:", [round(float(x), 1) for x in g], "sum", round(sum(float(x) for x in g), 1))
    print("H2D GB/s, rank 0 alone:", round(solo, 1))
    print("affinity", sorted(os.sched_getaffinity(0)))
if world > 1: dist.destroy_process_group()
