#!/usr/bin/env python
"""Second host->device probe (one process, all GPUs of the box visible): where is the ceiling that h2d_probe.py found
(GPUs 0-3 share ~115 GB/s, all eight ~239 GB/s, on a VM that exposes a single NUMA node)?

  python tools/h2d_probe2.py > gpurun_out/h2d_probe2.jsonl

1. per-GPU rates for subsets of GPUs copying concurrently, each from its own pinned buffer;
2. a pool of pinned candidate buffers: for every candidate, GPUs 0-3 (then 4-7) all copy from THAT buffer concurrently -
   if guest memory is spread over two physical sockets, candidates differ and ranks could pick near buffers;
3. the same for a buffer backed by transparent huge pages (mmap + MADV_HUGEPAGE + cudaHostRegister): IOMMU pressure.
"""
from __future__ import annotations

import ctypes as C
import json
import mmap
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

NB = 256 << 20


def rates(gpus, hosts, devs, streams, reps=4):
    """every GPU in `gpus` copies hosts[g] -> devs[g] `reps` times, all concurrently; returns GB/s per GPU"""
    ev = {}
    for g in gpus:
        with torch.cuda.device(g), torch.cuda.stream(streams[g]):
            devs[g].copy_(hosts[g], non_blocking=True)  # warm
    for g in gpus:
        torch.cuda.synchronize(g)
    for g in gpus:
        with torch.cuda.device(g), torch.cuda.stream(streams[g]):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(streams[g])
            for _ in range(reps):
                devs[g].copy_(hosts[g], non_blocking=True)
            e1.record(streams[g])
            ev[g] = (e0, e1)
    out = []
    for g in gpus:
        ev[g][1].synchronize()
        out.append(reps * NB / (ev[g][0].elapsed_time(ev[g][1]) * 1e-3) / 1e9)
    return out


def main():
    n = torch.cuda.device_count()
    streams = {g: torch.cuda.Stream(device=g) for g in range(n)}
    devs = {g: torch.empty(NB, dtype=torch.uint8, device=f"cuda:{g}") for g in range(n)}
    own = {g: torch.empty(NB, dtype=torch.uint8).pin_memory() for g in range(n)}
    for h in own.values():
        h.fill_(1)
    subsets = [s for s in ([0], [4], [0, 1], [4, 5], [0, 4], [0, 1, 2, 3], [4, 5, 6, 7], [0, 1, 4, 5], [0, 2, 4, 6], list(range(8))) if max(s) < n]
    for s in subsets:
        r = rates(s, own, devs, streams)
        print(json.dumps({"probe": "subset", "gpus": s, "gbs": [round(x, 1) for x in r], "total": round(sum(r), 1)}), flush=True)
    # candidate pool
    pool = []
    try:
        for _ in range(12):
            t = torch.empty(NB, dtype=torch.uint8).pin_memory()
            t.fill_(2)
            pool.append(t)
    except RuntimeError as e:
        print(json.dumps({"probe": "pool", "note": f"stopped at {len(pool)} buffers: {e}"}), flush=True)
    groups = [g for g in ([0, 1, 2, 3], [4, 5, 6, 7]) if max(g) < n]
    for i, b in enumerate(pool):
        line = {"probe": "candidate", "buffer": i}
        for grp in groups:
            r = rates(grp, {g: b for g in grp}, devs, streams)
            line[f"gpus{grp[0]}-{grp[-1]}_total"] = round(sum(r), 1)
        print(json.dumps(line), flush=True)
    # transparent huge pages
    try:
        mm = mmap.mmap(-1, NB + (2 << 20), flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
        addr = C.addressof(C.c_char.from_buffer(mm))
        aligned = (addr + (2 << 20) - 1) & ~((2 << 20) - 1)
        libc = C.CDLL(None, use_errno=True)
        rc = libc.madvise(C.c_void_p(aligned), C.c_size_t(NB), 14)  # MADV_HUGEPAGE
        C.memset(C.c_void_p(aligned), 3, NB)
        rt = torch.cuda.cudart()
        err = rt.cudaHostRegister(aligned, NB, 1)  # portable
        thp = open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip()
        import numpy as np

        arr = np.ctypeslib.as_array((C.c_uint8 * NB).from_address(aligned))
        t = torch.from_numpy(arr)
        line = {"probe": "hugepage", "madvise_rc": rc, "register": str(err), "thp": thp,
                "anon_huge_kb": [l for l in open("/proc/self/smaps_rollup") if "AnonHuge" in l][0].split()[1]}
        for grp in groups + [[0]]:
            r = rates(grp, {g: t for g in grp}, devs, streams)
            line[f"gpus{grp[0]}-{grp[-1]}_total"] = round(sum(r), 1)
        print(json.dumps(line), flush=True)
        rt.cudaHostUnregister(aligned)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"probe": "hugepage", "error": repr(e)}), flush=True)


if __name__ == "__main__":
    main()
