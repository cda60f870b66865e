"""rve_sort_entity_ids_static (csrc/sort_ids.cu): the bitmap sort with its scratch in a file-scope device array, for callers
that cannot call cudaMalloc. Through the C-ABI, against the allocating entry point rve_sort_entity_ids (and np.sort), bit for
bit as every id test here, on calls in a row that share the one array. Collected last on purpose
(the name)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from ravengine_b200.api import RveError

pytestmark = pytest.mark.gpu

MAX_IDS = int(re.search(r"#define\s+RVE_SORT_STATIC_MAX_IDS\s+(\d+)u", (Path(__file__).resolve().parents[1] / "include" / "rve_b200.h").read_text()).group(1))


def _both(ctx, ids, id_limit):
    """(static result, allocating result) of one list through the two entry points, raw C-ABI calls"""
    count = len(ids)
    src = ctx.buffer_from_numpy(ids)
    out_static, out_alloc = ctx.create_buffer(4 * count), ctx.create_buffer(4 * count)
    out_static.upload(np.full(count, 0xFFFFFFFF, np.uint32))
    rc = ctx.lib.rve_sort_entity_ids_static(ctx.h, src.handle, count, id_limit, out_static.handle)
    assert rc == 0, ctx.lib.rve_last_error_string().decode()
    rc = ctx.lib.rve_sort_entity_ids(ctx.h, src.handle, count, id_limit, out_alloc.handle)
    assert rc == 0, ctx.lib.rve_last_error_string().decode()
    ctx.synchronize()
    a, b = out_static.download(np.uint32)[:count].copy(), out_alloc.download(np.uint32)[:count].copy()
    for x in (src, out_static, out_alloc):
        x.destroy()
    return a, b


def test_static_sort_equals_the_allocating_sort_on_calls_in_a_row(ctx):
    """A large id space first, then smaller ones with other ids: whatever the earlier call left in the shared array must not
    show (bitmap and counter are cleared per call); the maximum id space; ids past id_limit are dropped the same way."""
    rng = np.random.default_rng(2026)
    cases = [(200_000, 5_000_000), (50_000, 3_000_000), (33, 64), (1, 1), (300_000, MAX_IDS), (1000, 40_000)]
    for count, id_limit in cases:
        ids = rng.choice(id_limit, size=count, replace=False).astype(np.uint32)
        got, ref = _both(ctx, ids, id_limit)
        assert np.array_equal(got, ref), (count, id_limit)
        assert np.array_equal(got, np.sort(ids)), (count, id_limit)
    # ids >= id_limit are dropped: the tail of the output stays as it was, in both entry points alike
    ids = np.concatenate([rng.choice(10_000, size=500, replace=False), 10_000 + rng.choice(10_000, size=12, replace=False)]).astype(np.uint32)
    rng.shuffle(ids)
    count = len(ids)
    src, out = ctx.buffer_from_numpy(ids), ctx.create_buffer(4 * count)
    for fn in (ctx.sort_entity_ids_static, ctx.sort_entity_ids):
        out.upload(np.full(count, 0xFFFFFFFF, np.uint32))
        fn(src, count, 10_000, out)
        ctx.synchronize()
        got = out.download(np.uint32)[:count]
        assert np.array_equal(got[:500], np.sort(ids[ids < 10_000])) and np.all(got[500:] == 0xFFFFFFFF)
    src.destroy()
    out.destroy()


def test_static_sort_allocates_no_device_memory_and_rejects_larger_id_spaces(ctx):
    import torch

    rng = np.random.default_rng(7)
    count, id_limit = 100_000, 2_000_000
    ids = rng.choice(id_limit, size=count, replace=False).astype(np.uint32)
    src, out = ctx.buffer_from_numpy(ids), ctx.create_buffer(4 * count)
    torch.cuda.mem_get_info(ctx.device)  # torch's own CUDA set-up happens here, not inside the measured window
    ctx.sort_entity_ids_static(src, count, id_limit, out)  # resolves the array's address on this device, loads the kernels
    ctx.synchronize()
    free_before = torch.cuda.mem_get_info(ctx.device)[0]
    for _ in range(2):
        ctx.sort_entity_ids_static(src, count, id_limit, out)
    ctx.synchronize()
    assert torch.cuda.mem_get_info(ctx.device)[0] == free_before
    assert np.array_equal(out.download(np.uint32)[:count], np.sort(ids))
    # past the stated maximum: a clear error, nothing enqueued, the output untouched
    marker = np.full(count, 0xABCD1234, np.uint32)
    out.upload(marker)
    launches = ctx.launch_count
    with pytest.raises(RveError, match="RVE_SORT_STATIC_MAX_IDS"):
        ctx.sort_entity_ids_static(src, count, MAX_IDS + 1, out)
    rc = ctx.lib.rve_sort_entity_ids_static(ctx.h, src.handle, count, C.c_uint32(MAX_IDS + 1), out.handle)
    assert rc == -6  # RVE_ERR_UNSUPPORTED
    ctx.synchronize()
    assert ctx.launch_count == launches
    assert np.array_equal(out.download(np.uint32)[:count], marker)
    src.destroy()
    out.destroy()
