"""Host -> device bandwidth of the ranks of one box, one at a time and all together (why `e2e` of bench.py does not scale
with N): every rank copies a pinned 64 MB buffer to its GPU 20 times; CUDA-event time.  Also prints the PCIe / NUMA topology.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 tools/h2d_scaling.py"""
import os
import subprocess

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dev = torch.device("cuda", torch.cuda.current_device())
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
MB = 64
host = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
host.fill_(rank)
gpu = torch.empty(MB << 20, dtype=torch.uint8, device=dev)
back = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()


def timed(direction):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        (gpu.copy_(host, non_blocking=True) if direction == "h2d" else back.copy_(gpu, non_blocking=True))
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        (gpu.copy_(host, non_blocking=True) if direction == "h2d" else back.copy_(gpu, non_blocking=True))
    e1.record()
    torch.cuda.synchronize()
    return 20 * MB / 1024 / (e0.elapsed_time(e1) * 1e-3)       # GiB/s


def barrier():
    if world > 1:
        dist.barrier()


res = {}
for direction in ("h2d", "d2h"):
    alone = 0.0
    for r in range(world):                 # one rank at a time
        barrier()
        if r == rank:
            alone = timed(direction)
    barrier()
    together = timed(direction)            # all ranks at once
    res[direction] = (alone, together)
if world > 1:
    out = [None] * world
    dist.all_gather_object(out, (rank, res, sorted(os.sched_getaffinity(0))[:4], len(os.sched_getaffinity(0))))
else:
    out = [(rank, res, sorted(os.sched_getaffinity(0))[:4], len(os.sched_getaffinity(0)))]
if rank == 0:
    for r, d, aff, naff in out:
        print(f"rank {r}: h2d alone {d['h2d'][0]:6.1f} GiB/s, all ranks at once {d['h2d'][1]:6.1f};  d2h alone {d['d2h'][0]:6.1f}, "
              f"at once {d['d2h'][1]:6.1f}   (cpu affinity: {naff} cpus, first {aff})")
    print(f"aggregate h2d with all {world} ranks at once: {sum(d['h2d'][1] for _, d, _, _ in out):.1f} GiB/s; "
          f"d2h {sum(d['d2h'][1] for _, d, _, _ in out):.1f} GiB/s")
    for cmd in (["nvidia-smi", "topo", "-m"], ["lscpu"]):
        try:
            t = subprocess.run(cmd, capture_output=True, text=True, timeout=20).stdout
            if cmd[0] == "lscpu":
                t = "\n".join(l for l in t.splitlines() if any(k in l for k in ("NUMA", "Model name", "Socket", "CPU(s):", "Thread")))
            print(t)
        except Exception as ex:
            print(cmd, ex)
if world > 1:
    dist.destroy_process_group()
