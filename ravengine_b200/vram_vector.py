"""BufferedVRAMVector mirror (include/RavEngine/BufferedVRAMVector.hpp:33-209): a host-shared array with
per-element modification flags and a device-private copy, synchronised by EncodeSync
(src/BufferedVRAMVector.cpp:27-144). Marshalling only: the scan, the copies and the scatter kernel
live in librve_b200.so (csrc/sync.cu)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .api import Buffer, Context, _check

SYNC_RANGES, SYNC_PACKED, SYNC_PULL = _abi.SYNC_RANGES, _abi.SYNC_PACKED, _abi.SYNC_PULL
SYNC_FULL, SYNC_FULL_PULL = _abi.SYNC_FULL, _abi.SYNC_FULL_PULL


class BufferedVRAMVector:
    def __init__(self, ctx: Context, elem_bytes: int, count: int = 0, affine_rows: bool = False, shared_host: bool = False):
        """affine_rows: host rows are 48-B mat4x3 (columns c0..c3, xyz each); the private buffer keeps 64-B mat4 rows.
        shared_host: the host block is a memfd mapping another rank can map and pin (share() / Context.host_import)."""
        self.ctx, self.elem_bytes, self.affine_rows = ctx, elem_bytes, affine_rows
        h = C.c_void_p()
        flags = (_abi.VECTOR_AFFINE_ROWS if affine_rows else 0) | (_abi.VECTOR_SHARED_HOST if shared_host else 0)
        _check(ctx.lib, ctx.lib.rvs_vector_create_ex(ctx.h, elem_bytes, count, flags, C.byref(h)))
        self.h = h

    def share(self) -> tuple[int, int, int]:
        """(pid, fd, mapped bytes) of a shared_host vector's host block"""
        pid, fd, nb = C.c_int32(), C.c_int32(), C.c_uint64()
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_share(self.h, C.byref(pid), C.byref(fd), C.byref(nb)))
        return pid.value, fd.value, nb.value

    def encode_sync_rows(self, mode: int, row_begin: int, row_end: int) -> _abi.SyncStats:
        """whole-row upload (SYNC_FULL / SYNC_FULL_PULL) of rows [row_begin, row_end) only"""
        st = _abi.SyncStats()
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_encode_sync_rows(self.h, mode, row_begin, row_end, C.byref(st)))
        return st

    def __len__(self) -> int:
        return self.ctx.lib.rvs_vector_size(self.h)

    def resize(self, count: int) -> None:
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_resize(self.h, count))

    def host(self, dtype=np.uint8) -> np.ndarray:
        """GetHostBuffer(): a view of the pinned host block (invalid after the next growth)."""
        n = len(self) * self.elem_bytes
        if n == 0:
            return np.empty(0, dtype)
        raw = (C.c_uint8 * n).from_address(self.ctx.lib.rvs_vector_host(self.h))
        return np.frombuffer(raw, dtype=dtype)

    def tracking(self) -> np.ndarray:
        n = len(self)
        if n == 0:
            return np.empty(0, np.uint8)
        raw = (C.c_uint8 * n).from_address(self.ctx.lib.rvs_vector_tracking(self.h))
        return np.frombuffer(raw, dtype=np.uint8)

    def set_value_at(self, i: int, value: np.ndarray) -> None:
        value = np.ascontiguousarray(value)
        assert value.nbytes == self.elem_bytes
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_set(self.h, i, value.ctypes.data))

    def mark(self, begin: int, end: int) -> None:
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_mark(self.h, begin, end))

    def encode_sync(self, mode: int = SYNC_PACKED, direct_min_rows: int = 256) -> _abi.SyncStats:
        st = _abi.SyncStats()
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_encode_sync(self.h, mode, direct_min_rows, C.byref(st)))
        return st

    def pulled_rows(self) -> int:
        return int(self.ctx.lib.rvs_vector_pulled_rows(self.h))

    def last_ranges(self) -> np.ndarray:
        n = self.ctx.lib.rvs_vector_last_ranges(self.h, None, 0)
        out = np.zeros((n, 2), np.uint32)
        if n:
            self.ctx.lib.rvs_vector_last_ranges(self.h, out.ctypes.data_as(C.POINTER(C.c_uint32)), n)
        return out

    @property
    def private_buffer(self) -> Buffer:
        """GetPrivateBuffer(): ask again after every encode_sync (growth replaces it, like the reference)."""
        h = self.ctx.lib.rvs_vector_private_buffer(self.h)
        return Buffer(self.ctx, h, self.ctx.lib.rve_buffer_size(self.ctx.h, h))

    def destroy(self) -> None:
        if self.h:
            self.ctx.lib.rvs_vector_destroy(self.h)
            self.h = None


class SparseSet:
    """BufferedVRAMSparseSet<entity_id_t, entity_id_t> mirror (include/RavEngine/SparseSet.hpp:6-104,
    include/RavEngine/BufferedVRAMVector.hpp:129-170): a command's entity list. Marshalling only (csrc/sync.cu)."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        h = C.c_void_p()
        _check(ctx.lib, ctx.lib.rvs_sparse_set_create(ctx.h, C.byref(h)))
        self.h = h

    def __len__(self) -> int:
        return self.ctx.lib.rvs_sparse_set_dense_size(self.h)

    def emplace(self, sparse_index: int, value: int | None = None) -> None:
        _check(self.ctx.lib, self.ctx.lib.rvs_sparse_set_emplace(self.h, sparse_index, sparse_index if value is None else value))

    def erase(self, sparse_index: int) -> None:
        _check(self.ctx.lib, self.ctx.lib.rvs_sparse_set_erase(self.h, sparse_index))

    def emplace_many(self, sparse_indices: np.ndarray, values: np.ndarray | None = None) -> None:
        idx = np.ascontiguousarray(sparse_indices, dtype=np.uint32)
        val = idx if values is None else np.ascontiguousarray(values, dtype=np.uint32)
        _check(self.ctx.lib, self.ctx.lib.rvs_sparse_set_emplace_many(self.h, idx.ctypes.data, val.ctypes.data, len(idx)))

    def erase_many(self, sparse_indices: np.ndarray) -> None:
        idx = np.ascontiguousarray(sparse_indices, dtype=np.uint32)
        _check(self.ctx.lib, self.ctx.lib.rvs_sparse_set_erase_many(self.h, idx.ctypes.data, len(idx)))

    def has(self, sparse_index: int) -> bool:
        return bool(self.ctx.lib.rvs_sparse_set_has(self.h, sparse_index))

    def dense(self) -> np.ndarray:
        n = len(self)
        if n == 0:
            return np.empty(0, np.uint32)
        return np.ctypeslib.as_array((C.c_uint32 * n).from_address(self.ctx.lib.rvs_sparse_set_dense(self.h))).copy()

    def reverse_map(self) -> np.ndarray:
        n = len(self)
        if n == 0:
            return np.empty(0, np.uint32)
        return np.ctypeslib.as_array((C.c_uint32 * n).from_address(self.ctx.lib.rvs_sparse_set_reverse_map(self.h))).copy()

    def encode_sync(self, mode: int = SYNC_PACKED, direct_min_rows: int = 256) -> _abi.SyncStats:
        st = _abi.SyncStats()
        _check(self.ctx.lib, self.ctx.lib.rvs_vector_encode_sync(self.ctx.lib.rvs_sparse_set_vector(self.h), mode, direct_min_rows, C.byref(st)))
        return st

    @property
    def private_buffer(self) -> Buffer:
        h = self.ctx.lib.rvs_vector_private_buffer(self.ctx.lib.rvs_sparse_set_vector(self.h))
        return Buffer(self.ctx, h, self.ctx.lib.rve_buffer_size(self.ctx.h, h))

    def destroy(self) -> None:
        if self.h:
            self.ctx.lib.rvs_sparse_set_destroy(self.h)
            self.h = None


def encode_sync_batch(vectors, direct_min_rows: int = 256):
    """rvs_vectors_encode_sync_batch: EncodeSync (packed) of several BufferedVRAMVector / SparseSet objects of one context - one
    staging copy and one scatter launch for all of those whose changes are small. Returns one SyncStats per vector."""
    if not vectors:
        return []
    lib = vectors[0].ctx.lib
    hs = [lib.rvs_sparse_set_vector(v.h) if isinstance(v, SparseSet) else v.h for v in vectors]
    arr = (C.c_void_p * len(hs))(*hs)
    stats = (_abi.SyncStats * len(hs))()
    _check(lib, lib.rvs_vectors_encode_sync_batch(arr, len(hs), direct_min_rows, stats))
    return list(stats)
