#!/usr/bin/env python
"""Host->device upload probe for a multi-GPU box (VERDICT r01 'e2e weak scaling collapses at N >= 4').

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/h2d_probe.py > gpurun_out/h2d_probe.jsonl

Rank 0 prints the box's topology (nvidia-smi topo -m, NUMA nodes, CPU counts). Then every rank measures its pinned
host->device copy rate while the first k ranks copy CONCURRENTLY (k = 1, 2, 4, .. N), once with pinned memory allocated
where the launcher left the process ("unbound") and once after rve_ctx_bind_host_to_device ("bound": CPUs + preferred
NUMA node of the GPU's PCIe root). One JSON line per (placement, k). The last lines compare the two whole-array upload
paths of csrc/sync.cu on 10 M affine rows (copy engine + expand kernels vs one zero-copy kernel) on rank 0 alone and on
all ranks together.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ravengine_b200 import Context  # noqa: E402


def sh(cmd: str) -> str:
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=30).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return f"<{e}>"


def node_of_address(ptr: int) -> int:
    """NUMA node of the page holding `ptr` (move_pages query), -1 when unknown."""
    import ctypes as C

    libc = C.CDLL(None, use_errno=True)
    pages = (C.c_void_p * 1)(C.c_void_p(ptr & ~4095))
    status = (C.c_int * 1)(-1)
    rc = libc.syscall(279, 0, 1, pages, None, status, 0)  # SYS_move_pages on x86_64
    return int(status[0]) if rc == 0 else -1


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        print(json.dumps({"topo": sh("nvidia-smi topo -m"), "numa_nodes": sh("ls -d /sys/devices/system/node/node* | wc -l"),
                          "lscpu": sh("lscpu | grep -i -E 'model name|socket|numa|^CPU\\(s\\)'"),
                          "node_mem": sh("grep -H MemTotal /sys/devices/system/node/node*/meminfo"),
                          "affinity_at_start": sorted(os.sched_getaffinity(0))[:4] + ["..", len(os.sched_getaffinity(0))]}), flush=True)
    stream = torch.cuda.Stream()
    ctx = Context(local, stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    nbytes = 512 << 20
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    ks = [k for k in (1, 2, 4, 8) if k <= world]

    def concurrent_rates(host, tag, info):
        for k in ks:
            barrier()
            rate = 0.0
            if rank < k:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(stream):
                    dev.copy_(host, non_blocking=True)
                    e0.record(stream)
                    for _ in range(4):
                        dev.copy_(host, non_blocking=True)
                    e1.record(stream)
                e1.synchronize()
                rate = 4 * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
            rates = [rate]
            if world > 1:
                t = torch.zeros(world, device="cuda")
                dist.all_gather_into_tensor(t, torch.tensor([rate], device="cuda"))
                rates = [float(x) for x in t.tolist()]
            nodes = [info]
            if world > 1:
                nodes = [None] * world
                dist.all_gather_object(nodes, info)
            if rank == 0:
                print(json.dumps({"probe": "concurrent_h2d", "placement": tag, "active_ranks": k, "gbs_by_rank": rates[:k],
                                  "gbs_total": sum(rates[:k]), "per_rank_info": nodes[:k] if k == ks[-1] else None}), flush=True)

    host0 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    host0.fill_(1)
    concurrent_rates(host0, "unbound", {"host_page_node": node_of_address(host0.data_ptr()), "cpu": os.sched_getaffinity(0).__len__()})
    aff = ctx.bind_host_to_device()
    host1 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    host1.fill_(1)
    aff["host_page_node"] = node_of_address(host1.data_ptr())
    concurrent_rates(host1, "bound", aff)
    del host0, host1

    # the two whole-array upload paths on 10 M affine rows
    from ravengine_b200.vram_vector import SYNC_FULL, SYNC_FULL_PULL, BufferedVRAMVector

    n = 10_000_000
    for affine, eb in ((True, 48), (False, 64)):
        v = BufferedVRAMVector(ctx, eb, n, affine_rows=affine)
        v.host(np.float32)[:] = 1.0
        for mode, name in ((SYNC_FULL, "copy"), (SYNC_FULL_PULL, "pull")):
            for together in ([False, True] if world > 1 else [False]):
                barrier()
                rate = 0.0
                if together or rank == 0:
                    v.encode_sync(mode)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    for _ in range(3):
                        v.encode_sync(mode)
                    e1.record(stream)
                    e1.synchronize()
                    rate = 3 * n * eb / (e0.elapsed_time(e1) * 1e-3) / 1e9
                rates = [rate]
                if world > 1:
                    t = torch.zeros(world, device="cuda")
                    dist.all_gather_into_tensor(t, torch.tensor([rate], device="cuda"))
                    rates = [float(x) for x in t.tolist()]
                if rank == 0:
                    print(json.dumps({"probe": "vector_full_upload", "host_row_bytes": eb, "path": name, "all_ranks": together,
                                      "ms_per_upload_rank0": n * eb / rates[0] / 1e6 if rates[0] else None, "gbs_by_rank": rates}), flush=True)
        v.destroy()
    barrier()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
