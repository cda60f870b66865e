"""Instance-range sharding of the cull across ranks (partition A). One process per GPU; torch.distributed
is the plumbing. Every rank culls its slice of each command's entity list into rank-local compact slices; the only
exchange step is the per-slot visible counts (all-gather, a few dozen bytes) followed by an exclusive scan over ranks,
which gives each rank's offset inside the global per-(draw, LOD) slice and the global instanceCount. The compacted ids
are gathered only when a single consumer needs them - by a kernel that writes them straight into the consumer's global
culling buffers over NVLink peer memory (csrc/p2p.cu, rve_gather_ids_p2p), chained behind the count exchange.
Partition B (independent camera views) and the 2-D split (view range x instance range) are planned here too.

Slot layout (identical to the single-GPU one, shaders/defaultcull.csh:305-312, src/RenderEngine_Draw.cpp:1334-1362):
per material, command c with n_c entities and L_c LODs owns indirect slots [sum L_<c, +L_c) and the culling-buffer
range starting at sum_{c'<c} L_c' * n_c', LOD l at + n_c * l.

This module holds host logic only (integer plans + process-group plumbing); the scan and the gather run on the device.
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


@dataclass(frozen=True)
class GatherPlan:
    """Where every slot's ids travel in the device-side gather (rve_gather_ids_p2p): slot s reads this rank's slice at
    src_start[s] of local culling buffer slot_buffer[s] (= material index) and lands in the consumer's GLOBAL buffer of
    the same material at dst_start[s] + offsets[rank][s]."""
    slot_buffer: tuple[int, ...]
    src_start: tuple[int, ...]
    dst_start: tuple[int, ...]
    global_slots: tuple[int, ...]  # id slots of each material's global culling buffer (numEntities * numLODs)


def gather_plan(local_plan: list[SlotPlan], global_plan: list[SlotPlan]) -> GatherPlan:
    """Slot layout on both sides of the gather. The global side is the single-GPU layout of the whole scene
    (shaders/defaultcull.csh:305-312: cullingBufferOffset + numObjects * lod with GLOBAL numObjects)."""
    buf, src, dst = [], [], []
    for m, (lp, gp) in enumerate(zip(local_plan, global_plan)):
        assert lp.lods == gp.lods, "local and global plans must describe the same commands"
        ls, gs = lp.slot_starts(), gp.slot_starts()
        buf += [m] * len(ls)
        src += ls
        dst += gs
    return GatherPlan(tuple(buf), tuple(src), tuple(dst), tuple(p.culling_slots for p in global_plan))


def split_views(num_views: int, world: int) -> list[tuple[int, int]]:
    """Partition B: rank r owns the contiguous view range [r*V/R, (r+1)*V/R) - views are independent, no collective."""
    return split_range(num_views, world)


def assign_views(num_views: int, world: int, policy: str = "strided") -> list[list[int]]:
    """Which views each rank culls. "range": the contiguous ranges of split_views. "strided" (default): rank r takes views
    r, r + R, r + 2R, ... - neighbouring views of a camera grid see similar amounts of the scene, so dealing them round-robin
    balances the visible work (a contiguous split of BASELINE C5's 32 x 16 grid left one of two ranks with 1.9x the work)."""
    if policy == "range":
        return [list(range(lo, hi)) for lo, hi in split_views(num_views, world)]
    assert policy == "strided", policy
    return [list(range(r, num_views, world)) for r in range(world)]


@dataclass(frozen=True)
class Grid2D:
    """View range x instance range (C5: 512 views over 1 B instances). world = view_groups * instance_groups; rank
    r = vg * instance_groups + ig owns views split_views(V, view_groups)[vg] of instances split_range(n, instance_groups)[ig].
    The per-slot count exchange (and the optional id gather) runs among the instance_groups ranks that share a view group
    (`peers`), once per owned view; ranks of different view groups never talk."""
    world: int
    view_groups: int

    def __post_init__(self):
        assert self.view_groups >= 1 and self.world % self.view_groups == 0, "view groups must divide the world size"

    @property
    def instance_groups(self) -> int:
        return self.world // self.view_groups

    def coords(self, rank: int) -> tuple[int, int]:
        return rank // self.instance_groups, rank % self.instance_groups

    def peers(self, rank: int) -> list[int]:
        vg, _ = self.coords(rank)
        return [vg * self.instance_groups + i for i in range(self.instance_groups)]

    def owned(self, rank: int, num_views: int, num_instances: int) -> tuple[tuple[int, int], tuple[int, int]]:
        vg, ig = self.coords(rank)
        return split_views(num_views, self.view_groups)[vg], split_range(num_instances, self.instance_groups)[ig]


class ShardedViews:
    """Partition B on one rank: cull this rank's views of a batch with the view-batched kernel (rve_cull_views through
    the shim, <= 128 views per call; the reference runs one cull per view, src/RenderEngine_Draw.cpp:1769-1866,
    1870-1987). No data-path collective. After `cull_chunk(i)` each material's buffers hold the chunk's views as consecutive
    copies of the single-view layout (view v of the chunk at v * numEntities*numLODs ids / v * numLODs commands), exactly
    what rvs_cull_static_views produces for those views on one GPU."""

    MAX_VIEWS_PER_CALL = 128

    def __init__(self, world_obj, views: list, pyramids: list, rank: int, world_size: int, policy: str = "strided"):
        assert len(views) == len(pyramids)
        self.world_obj, self.rank, self.ws = world_obj, rank, world_size
        self.owned = assign_views(len(views), world_size, policy)[rank]  # global view indices, in the order they are culled
        self.views, self.pyramids = [views[i] for i in self.owned], [pyramids[i] for i in self.owned]
        k = self.MAX_VIEWS_PER_CALL
        self.chunks = [(i, min(i + k, len(self.views))) for i in range(0, len(self.views), k)]

    def cull_chunk(self, i: int) -> list[int]:
        """Enqueues the cull of chunk i; returns the GLOBAL indices of the views it covers (copy v of the output = the v-th)."""
        a, b = self.chunks[i]
        self.world_obj.cull_static_views(self.views[a:b], self.pyramids[a:b])
        return self.owned[a:b]

    def cull_all(self) -> None:
        for i in range(len(self.chunks)):
            self.cull_chunk(i)


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
        # Frame slots: with two frames in flight every slot has its own world (its own indirect / culling buffers), so the
        # exchange of frame f reads slot f & 1 while the cull of frame f + 1 re-initialises the other one - no wait between
        # them (add_frame_slot). Slot 0 is the world passed to the constructor.
        self.targets = [(world_obj, materials)]
        self._indirect = [None]
        self._indirect2 = [None]
        self._pending = [None]
        # fused: the exchange kernel gathers the counts itself (one launch less on the cull's stream); RVE_EXCHANGE_FUSED=0
        # restores the separate gather launch
        self.fused = os.environ.get("RVE_EXCHANGE_FUSED", "1") == "1" and not self.inline
        self.rank, self.mode = rank, "nccl"
        if world_size > 1 and exchange in ("auto", "p2p"):
            self._setup_p2p(strict=exchange == "p2p")

    def add_frame_slot(self, world_obj, materials: list[int]) -> int:
        """A second (third ..) set of output buffers: a world with the same materials / commands as slot 0's. Returns its slot."""
        self.targets.append((world_obj, materials))
        self._indirect.append(None)
        self._indirect2.append(None)
        self._pending.append(None)
        if getattr(self, "_local_ids2", None) is not None:
            self._local_ids2.append(None)
        return len(self.targets) - 1

    @property
    def pending(self):
        return self._pending[0]

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

    def setup_gather(self, local_plan: list[SlotPlan], global_plan: list[SlotPlan], consumers: list[int], id_base: int) -> None:
        """Device-side gather for `consumers` (ranks that need the whole compacted id lists). Consumers allocate one
        global culling buffer per material (global numEntities * numLODs ids, the single-GPU layout) and publish it to
        their peers as CUDA IPC handles; requires the peer-memory exchange (mode "p2p"), or world size 1."""
        import torch.distributed as dist

        assert self.ws == 1 or self.mode == "p2p", "the device-side gather runs over the peer-memory mailboxes"
        self.gplan = gather_plan(local_plan, global_plan)
        assert len(self.gplan.slot_buffer) == self.slots
        self.consumers, self.id_base = sorted(consumers), id_base
        self.global_ids = None
        if self.rank in self.consumers:
            self.global_ids = [self.ctx2.create_buffer(max(4 * n, 16)) for n in self.gplan.global_slots]
            self.ctx2.synchronize()
        nb = len(self.gplan.global_slots)
        self._dst = [[None] * nb for _ in range(self.ws)]
        if self.ws == 1:
            self._dst[0] = self.global_ids
            self._gather_boxes = getattr(self, "peers", None) or [self._solo_mailbox()]
        else:
            mine = [self.ctx2.ipc_export(b) for b in self.global_ids] if self.global_ids else []
            allh = [None] * self.ws
            dist.all_gather_object(allh, mine)
            for q in self.consumers:
                self._dst[q] = self.global_ids if q == self.rank else [self.ctx2.ipc_import(h, max(4 * n, 16))
                                                                        for h, n in zip(allh[q], self.gplan.global_slots)]
            self._gather_boxes = self.peers
        self._local_ids2 = [None] * len(self.targets)

    def _solo_mailbox(self):
        self.mailbox = self.ctx2.create_buffer(self.ctx2.mailbox_bytes(1, self.slots))
        self.mailbox.zero()
        self.status = self.ctx2.create_buffer(16)
        self.status.zero()
        self.peers = [self.mailbox]
        return self.mailbox

    def gather(self, b: int, slot: int = 0) -> None:
        """Enqueue the id gather of the exchange that just filled buffer set `b` (side stream, behind the exchange kernel)."""
        import torch

        if self._local_ids2[slot] is None:  # this rank's culling buffers, seen from the side-stream context
            w, mats = self.targets[slot]
            bufs = [w.material_outputs(m)[0] for m in mats]
            self._local_ids2[slot] = [self.ctx2.wrap_pointer(x.ptr, x.nbytes, keepalive=x) for x in bufs]
        mask = sum(1 << q for q in self.consumers)
        self.ctx2.gather_ids_p2p(self._local_ids2[slot], self.gplan, self._dst, mask, self.id_base, self.b_off[b], self.b_tot[b],
                                 self._gather_boxes, self.rank, self.ws, self.k, self.status)
        self.done[b] = torch.cuda.Event()
        self.done[b].record(self.side)
        self._pending[slot] = self.done[b]  # the next cull of this slot rewrites the local slices: it must wait for this gather too

    def _indirect_buffers(self, slot: int = 0):
        w, mats = self.targets[slot]
        return [w.material_outputs(m)[1] for m in mats]

    def exchange(self, refresh_buffers: bool = False, slot: int = 0) -> int:
        """Returns the index (0/1) of the buffer set this frame's counts / offsets / totals land in. `slot`: the frame slot
        (world) whose cull has just been enqueued."""
        import torch
        import torch.distributed as dist

        b = self.k & 1
        self.k += 1
        if self._indirect[slot] is None or refresh_buffers:  # handles change only when the shim reallocates (entity counts changed)
            self._indirect[slot] = self._indirect_buffers(slot)
            self._indirect2[slot] = None
        if self.mode == "p2p" and self.fused:
            # no launch on the cull's stream at all: the exchange kernel (side stream) reads the counts out of the
            # indirect buffers itself; begin_cull() keeps the slot's next re-initialisation behind it
            if self._indirect2[slot] is None:
                self._indirect2[slot] = [self.ctx2.wrap_pointer(x.ptr, x.nbytes, keepalive=x) for x in self._indirect[slot]]
            ev = torch.cuda.Event()
            ev.record(self.stream)
            self.side.wait_event(ev)
            self.ctx2.exchange_counts_p2p_indirect(self._indirect2[slot], [0] * len(self.mats), self.lods, self.peers, self.rank, self.ws,
                                                   self.k, self.b_off[b], self.b_tot[b], self.status)
            self.done[b] = torch.cuda.Event()
            self.done[b].record(self.side)
            self._pending[slot] = self.done[b]
            return b
        if self.done[b] is not None:
            self.stream.wait_event(self.done[b])  # the exchange that last used this buffer set has finished
        self.ctx.gather_instance_counts_multi(self._indirect[slot], [0] * len(self.mats), self.lods, self.b_local[b])
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

    def begin_cull(self, slot: int = 0) -> None:
        """Call before the cull of a frame (between its pyramid build and the cull when they are separate calls): the cull
        re-initialises the slot's indirect commands, which the (fused-gather) exchange of the slot's previous frame may still
        be reading on the side stream. With two frame slots that exchange is two frames old: the wait is over before it starts."""
        if self._pending[slot] is not None:
            self.stream.wait_event(self._pending[slot])
            self._pending[slot] = None

    def finish(self) -> None:
        self.stream.wait_stream(self.side)

    def close(self) -> None:
        """Joins the side stream and releases the side-stream context (peer mappings are unmapped with it)."""
        self.finish()
        self.ctx2.synchronize()
        self.ctx2.close()

    def timed_out(self) -> bool:
        """True when a peer failed to arrive within the exchange kernel's timeout (p2p mode; synchronises)."""
        if self.mode != "p2p":
            return False
        import numpy as np

        return bool(self.status.download(np.uint32)[0])

    @property
    def launch_count(self) -> int:
        return self.ctx2.launch_count
