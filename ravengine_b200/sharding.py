"""Instance-range sharding of the cull across ranks (partition A). One process per GPU; torch.distributed
is the plumbing. Every rank culls its slice of each command's entity list into rank-local compact slices; the only
exchange step is the per-slot visible counts (all-gather, a few dozen bytes) followed by an exclusive scan over ranks,
which gives each rank's offset inside the global per-(draw, LOD) slice and the global instanceCount. The compacted ids
are gathered only when a single consumer needs them.

Slot layout (identical to the single-GPU one, shaders/defaultcull.csh:305-312, src/RenderEngine_Draw.cpp:1334-1362):
per material, command c with n_c entities and L_c LODs owns indirect slots [sum L_<c, +L_c) and the culling-buffer
range starting at sum_{c'<c} L_c' * n_c', LOD l at + n_c * l.

This module holds host logic only (integer plans + collectives); the scan itself is `rve_shard_offsets` on the device.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


def split_range(n: int, world: int) -> list[tuple[int, int]]:
    """Contiguous, balanced [lo, hi) ranges: rank r owns floor(r*n/world) .. floor((r+1)*n/world)."""
    return [(r * n // world, (r + 1) * n // world) for r in range(world)]


@dataclass(frozen=True)
class SlotPlan:
    """Slots of ONE material. counts[c] = entities of command c (on this rank, or globally), lods[c] = its LOD count."""
    counts: tuple[int, ...]
    lods: tuple[int, ...]

    @property
    def num_slots(self) -> int:
        return sum(self.lods)

    def slot_starts(self) -> list[int]:
        """Start of every slot's slice in the material's culling buffer, in slot order."""
        out, off = [], 0
        for n, l in zip(self.counts, self.lods):
            out += [off + n * k for k in range(l)]
            off += n * l
        return out

    @property
    def culling_slots(self) -> int:
        return sum(self.counts) * sum(self.lods)  # numEntities * numLODs (:1288)


def shard_plans(global_counts: list[tuple[int, ...]], lods: list[tuple[int, ...]], world: int) -> list[list[SlotPlan]]:
    """plans[rank][material] when every command's entity list is split by split_range."""
    plans = []
    for r in range(world):
        per_mat = []
        for counts, l in zip(global_counts, lods):
            per_mat.append(SlotPlan(tuple(split_range(n, world)[r][1] - split_range(n, world)[r][0] for n in counts), tuple(l)))
        plans.append(per_mat)
    return plans


def exchange_counts(local_counts, world: int):
    """all-gather of this rank's u32[slots] visible counts -> [world, slots] (same device / dtype)."""
    import torch
    import torch.distributed as dist

    out = torch.empty((world, local_counts.numel()), dtype=local_counts.dtype, device=local_counts.device)
    if world == 1:
        out[0] = local_counts
        return out
    try:
        dist.all_gather_into_tensor(out.view(-1), local_counts.contiguous())
    except (RuntimeError, NotImplementedError):  # older gloo builds
        parts = [torch.empty_like(local_counts) for _ in range(world)]
        dist.all_gather(parts, local_counts.contiguous())
        out = torch.stack(parts)
    return out


def gather_ids(local_ids, local_plan: list[SlotPlan], global_plan: list[SlotPlan], counts_all, offsets_all, rank: int, world: int,
               id_base_of_rank):
    """All-gather of the compacted id slices into the global culling-buffer layout (only when one consumer needs them).

    local_ids[m]: this rank's culling buffer of material m (torch int32/uint32-as-int32 tensor);
    counts_all / offsets_all: [world, total_slots] from exchange_counts / the exclusive scan (slots of all materials
    concatenated in material order); id_base_of_rank(r): value added to rank r's local ids to make them global.
    Returns one global culling buffer per material (filled with -1 where nothing was written)."""
    import torch
    import torch.distributed as dist

    dev = local_ids[0].device
    counts_host = counts_all.cpu().numpy().astype(np.int64)
    offs_host = offsets_all.cpu().numpy().astype(np.int64)
    # pack this rank's visible ids slot by slot, pad to the largest rank, all-gather
    packed, s = [], 0
    for m, plan in enumerate(local_plan):
        for start in plan.slot_starts():
            c = int(counts_host[rank, s])
            packed.append(local_ids[m][start:start + c])
            s += 1
    mine = torch.cat(packed) if packed else torch.empty(0, dtype=torch.int32, device=dev)
    per_rank = counts_host.sum(axis=1)
    cap = int(per_rank.max()) if per_rank.size else 0
    buf = torch.zeros(max(cap, 1), dtype=mine.dtype, device=dev)
    buf[: mine.numel()] = mine
    if world > 1:
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf)
    else:
        parts = [buf]
    out = [torch.full((max(p.culling_slots, 1),), -1, dtype=mine.dtype, device=dev) for p in global_plan]
    for r in range(world):
        pos, s = 0, 0
        base = id_base_of_rank(r)
        for m, plan in enumerate(global_plan):
            for gstart in plan.slot_starts():
                c = int(counts_host[r, s])
                if c:
                    dst = gstart + int(offs_host[r, s])
                    out[m][dst:dst + c] = parts[r][pos:pos + c] + base
                pos += c
                s += 1
    return out


class ShardedCull:
    """GPU-side driver of one rank. After `world.cull_static(...)` has run on this rank's shard, `exchange()` enqueues
    ONE kernel on the cull's stream that copies the per-slot instance counts out of the indirect buffers (the next
    frame re-initialises them), then - on a side stream, so the next frame's pyramid build and cull are not held up by
    collective latency - the NCCL all-gather of those counts and the device exclusive scan over ranks. Buffers are
    double-buffered; `finish()` joins the side stream (call it before reading offsets / totals, or at frame end).
    Nothing synchronises with the host."""

    def __init__(self, ctx, world_obj, materials: list[int], lods_per_material: list[int], world_size: int, torch_stream,
                 rank: int = 0, exchange: str = "auto"):
        """exchange: "p2p" = one kernel over NVLink peer memory (csrc/p2p.cu: peer stores + flags + scan, no NCCL call),
        "nccl" = all-gather + scan kernel, "auto" = p2p when every rank could map its peers' mailboxes."""
        import torch

        from .api import Context

        self.ctx, self.world_obj, self.mats, self.lods, self.ws, self.stream = ctx, world_obj, materials, lods_per_material, world_size, torch_stream
        self.slots = sum(lods_per_material)
        dev = torch.device("cuda", ctx.device)
        import os

        self.inline = os.environ.get("RVE_EXCHANGE_INLINE") == "1"  # diagnostics: exchange on the cull's own stream
        self.side = torch_stream if self.inline else torch.cuda.Stream(device=dev)
        self.ctx2 = Context(ctx.device, self.side.cuda_stream)  # same device memory, enqueues on the side stream
        self.t_local, self.t_all, self.t_off, self.t_tot = [], [], [], []
        self.b_local, self.b_all, self.b_off, self.b_tot, self.done = [], [], [], [], [None, None]
        with torch.cuda.stream(torch_stream):
            for _ in range(2):
                self.t_local.append(torch.zeros(self.slots, dtype=torch.int32, device=dev))
                self.t_all.append(torch.zeros((world_size, self.slots), dtype=torch.int32, device=dev))
                self.t_off.append(torch.zeros((world_size, self.slots), dtype=torch.int32, device=dev))
                self.t_tot.append(torch.zeros(self.slots, dtype=torch.int32, device=dev))
        torch_stream.synchronize()
        for i in range(2):
            self.b_local.append(ctx.wrap_tensor(self.t_local[i]))
            self.b_all.append(self.ctx2.wrap_tensor(self.t_all[i]))
            self.b_off.append(self.ctx2.wrap_tensor(self.t_off[i]))
            self.b_tot.append(self.ctx2.wrap_tensor(self.t_tot[i]))
        self.k = 0
        self._indirect = None
        self._indirect2 = None
        self.pending = None
        # fused: the exchange kernel gathers the counts itself (one launch less on the cull's stream); RVE_EXCHANGE_FUSED=0
        # restores the separate gather launch
        self.fused = os.environ.get("RVE_EXCHANGE_FUSED", "1") == "1" and not self.inline
        self.rank, self.mode = rank, "nccl"
        if world_size > 1 and exchange in ("auto", "p2p"):
            self._setup_p2p(strict=exchange == "p2p")

    def _setup_p2p(self, strict: bool) -> None:
        """Mailboxes: one cudaMalloc block per rank, zero-filled, exchanged as CUDA IPC handles through the process group."""
        import torch
        import torch.distributed as dist

        nbytes = self.ctx2.mailbox_bytes(self.ws, self.slots)
        ok, err = 1, ""
        try:
            self.mailbox = self.ctx2.create_buffer(nbytes)
            self.mailbox.zero()
            self.status = self.ctx2.create_buffer(16)
            self.status.zero()
            self.ctx2.synchronize()
            handle = self.ctx2.ipc_export(self.mailbox)
        except Exception as e:  # noqa: BLE001 - reported below, every rank must take the same branch
            ok, err, handle = 0, str(e), b""
        handles = [None] * self.ws
        dist.all_gather_object(handles, (ok, handle))
        peers = []
        if all(h[0] for h in handles):
            try:
                for q, (_, h) in enumerate(handles):
                    peers.append(self.mailbox if q == self.rank else self.ctx2.ipc_import(h, nbytes))
            except Exception as e:  # noqa: BLE001
                ok, err = 0, str(e)
        else:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=torch.device("cuda", self.ctx.device))
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            self.peers, self.mode = peers, "p2p"
        elif strict:
            raise RuntimeError(f"peer-memory count exchange unavailable: {err}")

    def _indirect_buffers(self):
        bufs = [self.world_obj.material_outputs(m)[1] for m in self.mats]
        return bufs

    def exchange(self, refresh_buffers: bool = False) -> int:
        """Returns the index (0/1) of the buffer set this frame's counts / offsets / totals land in."""
        import torch
        import torch.distributed as dist

        b = self.k & 1
        self.k += 1
        if self._indirect is None or refresh_buffers:  # handles change only when the shim reallocates (entity counts changed)
            self._indirect = self._indirect_buffers()
            self._indirect2 = None
        if self.mode == "p2p" and self.fused:
            # no launch on the cull's stream at all: the exchange kernel (side stream) reads the counts out of the
            # indirect buffers itself; begin_cull() keeps the next frame's re-initialisation behind it
            if self._indirect2 is None:
                self._indirect2 = [self.ctx2.wrap_pointer(x.ptr, x.nbytes, keepalive=x) for x in self._indirect]
            ev = torch.cuda.Event()
            ev.record(self.stream)
            self.side.wait_event(ev)
            self.ctx2.exchange_counts_p2p_indirect(self._indirect2, [0] * len(self.mats), self.lods, self.peers, self.rank, self.ws,
                                                   self.k, self.b_off[b], self.b_tot[b], self.status)
            self.done[b] = torch.cuda.Event()
            self.done[b].record(self.side)
            self.pending = self.done[b]
            return b
        if self.done[b] is not None:
            self.stream.wait_event(self.done[b])  # the exchange that last used this buffer set has finished
        self.ctx.gather_instance_counts_multi(self._indirect, [0] * len(self.mats), self.lods, self.b_local[b])
        ev = torch.cuda.Event()
        ev.record(self.stream)
        self.side.wait_event(ev)
        if self.mode == "p2p":
            if not hasattr(self, "b_local2"):
                self.b_local2 = [self.ctx2.wrap_tensor(t) for t in self.t_local]
            self.ctx2.exchange_counts_p2p(self.b_local2[b], self.slots, self.peers, self.rank, self.ws, self.k, self.b_off[b],
                                          self.b_tot[b], self.status)
        else:
            with torch.cuda.stream(self.side):
                if self.ws > 1:
                    dist.all_gather_into_tensor(self.t_all[b].view(-1), self.t_local[b])
                else:
                    self.t_all[b][0].copy_(self.t_local[b])
            self.ctx2.shard_offsets(self.b_all[b], self.ws, self.slots, self.b_off[b], self.b_tot[b])
        self.done[b] = torch.cuda.Event()
        self.done[b].record(self.side)
        return b

    def begin_cull(self) -> None:
        """Call between the pyramid build and the cull of the next frame: the cull re-initialises the indirect commands,
        which the (fused-gather) exchange of the previous frame may still be reading on the side stream."""
        if self.pending is not None:
            self.stream.wait_event(self.pending)
            self.pending = None

    def finish(self) -> None:
        self.stream.wait_stream(self.side)

    def timed_out(self) -> bool:
        """True when a peer failed to arrive within the exchange kernel's timeout (p2p mode; synchronises)."""
        if self.mode != "p2p":
            return False
        import numpy as np

        return bool(self.status.download(np.uint32)[0])

    @property
    def launch_count(self) -> int:
        return self.ctx2.launch_count
