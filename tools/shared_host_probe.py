#!/usr/bin/env python
"""Feasibility probe for cross-rank upload forwarding: can rank B pin (cudaHostRegister) and DMA from host memory that
rank A owns? Rank 0 creates a memfd, rank 1 maps it through /proc/<pid>/fd/<fd>, both register it, rank 1 copies from it.
  torchrun --nproc-per-node 2 tools/shared_host_probe.py"""
import ctypes as C
import json
import mmap
import os

import numpy as np
import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 256 << 20
    info = [None]
    if rank == 0:
        fd = os.memfd_create("rvs_probe")
        os.ftruncate(fd, nbytes)
        info = [(os.getpid(), fd)]
    dist.broadcast_object_list(info, src=0)
    pid, rfd = info[0]
    out = {"rank": rank}
    try:
        myfd = fd if rank == 0 else os.open(f"/proc/{pid}/fd/{rfd}", os.O_RDWR)
        mm = mmap.mmap(myfd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        addr = C.addressof(C.c_char.from_buffer(mm))
        arr = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(addr))
        if rank == 0:
            arr[:] = 7
        dist.barrier()
        err = torch.cuda.cudart().cudaHostRegister(addr, nbytes, 1 | 2)  # portable | mapped
        out["register"] = str(err)
        t = torch.from_numpy(arr)
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        dev.copy_(t, non_blocking=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(4):
            dev.copy_(t, non_blocking=True)
        e1.record()
        e1.synchronize()
        out["gbs"] = 4 * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
        out["sum_ok"] = bool(int(dev[:1000].sum().item()) == 7000)
        torch.cuda.cudart().cudaHostUnregister(addr)
    except Exception as e:  # noqa: BLE001
        out["error"] = repr(e)
    res = [None] * world
    dist.all_gather_object(res, out)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
