#!/usr/bin/env python3
"""Raw host->device copy rates with all ranks copying at once (what bounds bench.py's e2e number):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29544 tools/h2d_probe.py

Every rank copies 1 GiB of pinned host memory to its GPU with cudaMemcpyAsync, 8 times, (a) alone, one rank
after the other, (b) all ranks together.  Prints one JSON line: per-GPU GB/s alone and together, the aggregate,
the CPU affinity / NUMA node nvidia-smi reports for each GPU, and the host's core and memory figures."""
import json
import os
import subprocess
import sys
import time


def main():
    import torch
    import torch.distributed as dist
    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    n = 1 << 30
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.fill_(7)
    dev = torch.empty(n, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def copy_rate(reps=8):
        dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        return n * reps / (time.perf_counter() - t0) / 1e9

    alone = 0.0
    for r in range(world):  # one rank at a time
        barrier()
        if r == rank:
            alone = copy_rate()
    barrier()
    t0 = time.perf_counter()
    together = copy_rate()
    barrier()
    vals = torch.tensor([alone, together], dtype=torch.float64, device="cuda")
    parts = [torch.empty_like(vals) for _ in range(world)]
    if world > 1:
        dist.all_gather(parts, vals)
    else:
        parts = [vals]
    if rank == 0:
        topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout
        aff = [l.split("\t") for l in topo.splitlines() if l.startswith("GPU")]
        mem_gb = int([l for l in open("/proc/meminfo") if l.startswith("MemTotal")][0].split()[1]) / 1e6
        print(json.dumps({"n_gpus": world, "bytes_per_copy": n,
                          "per_gpu_alone_gbs": [round(p[0].item(), 2) for p in parts],
                          "per_gpu_together_gbs": [round(p[1].item(), 2) for p in parts],
                          "aggregate_together_gbs": round(sum(p[1].item() for p in parts), 2),
                          "gpu_cpu_affinity": [[c.strip() for c in a[-4:-1]] for a in aff][:world],
                          "host_cores": os.cpu_count(), "host_mem_gb": round(mem_gb, 1)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
