"""Dirty-range upload: the range scan against the restated reference loop
(src/BufferedVRAMVector.cpp:106-139), and - on the GPU - both sync modes leaving the private buffer
byte-identical to the host buffer."""
import numpy as np
import pytest

from oracle import host_tables


def ranges_of(flags):
    return host_tables.sync_ranges(np.array(flags, np.uint8)).tolist()


def test_range_scan_known_answers():
    # a single dirty element drags maxGapSize-1 clean ones along: closed at the index of the 5th clean flag
    assert ranges_of([1, 0, 0, 0, 0, 0, 0, 0]) == [[0, 5]]
    # a second dirty element inside the gap window extends the range ...
    assert ranges_of([1, 0, 0, 0, 0, 1]) == [[0, 6]]
    # ... one flag later it opens a new one
    assert ranges_of([1, 0, 0, 0, 0, 0, 1]) == [[0, 5], [6, 7]]
    assert ranges_of([0, 0, 0]) == []
    assert ranges_of([]) == []
    assert ranges_of([0, 0, 1]) == [[2, 3]]
    assert ranges_of([1] * 9) == [[0, 9]]
    # range still open at the end of the array
    assert ranges_of([0, 1, 0, 0]) == [[1, 4]]


def test_range_scan_clears_flags():
    t = np.array([0, 1, 1, 0, 0, 0, 0, 0, 0, 1], np.uint8)
    r = host_tables.sync_ranges(t)
    assert r.tolist() == [[1, 7], [9, 10]]
    assert not t.any()


def test_library_range_derivations_equal_the_restated_reference_loop():
    """CPU only (the two functions are plain host code of librve_b200.so): the scan the library runs over a flag array and the
    shortcut from the list of flagged indices against the restated reference loop (oracle/host_tables.py::sync_ranges), over
    random densities, clustered touches at every distance around maxGapSize, flags at both ends, empty and full arrays."""
    import ctypes as C

    from ravengine_b200 import _abi

    lib = _abi.load_library()
    rng = np.random.default_rng(5)
    cases = [np.zeros(0, np.uint8), np.zeros(17, np.uint8), np.ones(9, np.uint8), np.array([0, 0, 1], np.uint8), np.array([1] + [0] * 7, np.uint8)]
    for n in (1, 7, 64, 1000, 20_001):
        for density in (0.0005, 0.01, 0.2, 0.7):
            f = (rng.random(n) < density).astype(np.uint8)
            cases.append(f)
        f = np.zeros(n, np.uint8)
        base = rng.integers(0, max(n - 12, 1), size=max(n // 50, 1))
        for gap in range(1, 9):
            f[np.minimum(base + gap * (rng.random(len(base)) < 0.5), n - 1)] = 1
        f[[0, n - 1]] = 1
        cases.append(f)
    for f in cases:
        n = len(f)
        expect = host_tables.sync_ranges(f.copy())
        out = np.zeros((n // 2 + 2, 2), np.uint32)
        fl = np.ascontiguousarray(f)
        got = lib.rvs_sync_ranges_from_flags(fl.ctypes.data_as(C.POINTER(C.c_uint8)) if n else None, n, out.ctypes.data_as(C.POINTER(C.c_uint32)), len(out))
        assert got == len(expect) and np.array_equal(out[:got], expect)
        idx = np.flatnonzero(f).astype(np.uint32)
        out2 = np.zeros_like(out)
        got2 = lib.rvs_sync_ranges_from_indices(idx.ctypes.data_as(C.POINTER(C.c_uint32)) if len(idx) else None, len(idx), n,
                                                out2.ctypes.data_as(C.POINTER(C.c_uint32)), len(out2))
        assert got2 == len(expect) and np.array_equal(out2[:got2], expect), (n, idx[:10])


@pytest.mark.gpu
@pytest.mark.parametrize("elem_bytes,dtype", [(64, np.float32), (4, np.uint32), (2, np.uint16), (12, np.float32)])
@pytest.mark.parametrize("density", [0.0, 0.001, 0.05, 0.6, 1.0])
def test_sync_modes_match_host(ctx, elem_bytes, dtype, density):
    from ravengine_b200.vram_vector import SYNC_PACKED, SYNC_PULL, SYNC_RANGES, BufferedVRAMVector

    rng = np.random.default_rng(int(density * 1000) + elem_bytes)
    n = 50_003
    per = elem_bytes // np.dtype(dtype).itemsize
    for mode in (SYNC_RANGES, SYNC_PACKED, SYNC_PULL):
        v = BufferedVRAMVector(ctx, elem_bytes, n)
        h = v.host(dtype).reshape(n, per)
        h[:] = rng.integers(0, 60000, size=(n, per)).astype(dtype)
        st = v.encode_sync(mode, 64)  # Resize marked everything: one range
        assert mode == SYNC_PULL or (st.ranges == 1 and st.rows_in_ranges == n)
        pulled = n
        assert np.array_equal(v.private_buffer.download(dtype).reshape(-1, per)[:n], h)
        for frame in range(3):
            dirty = rng.random(n) < density
            if frame == 1:
                dirty[1000:3000] = True  # one long run: goes by direct copy in packed mode
            idx = np.nonzero(dirty)[0]
            h[idx] = rng.integers(0, 60000, size=(len(idx), per)).astype(dtype)
            v.tracking()[idx] = 1
            expect = host_tables.sync_ranges(dirty.astype(np.uint8))
            st = v.encode_sync(mode, 64)
            if mode == SYNC_PULL:  # no host scan: the kernel consumed (and cleared) the flags it was handed
                pulled += len(idx)
                assert np.array_equal(v.private_buffer.download(dtype).reshape(-1, per)[:n], h)
                assert v.pulled_rows() == pulled
                assert not v.tracking().any()
                continue
            assert np.array_equal(v.last_ranges(), expect)
            assert st.ranges == len(expect)
            assert not v.tracking().any()
            if mode == SYNC_RANGES:
                assert st.packed_rows == 0 and st.direct_copies == len(expect)
            else:
                lens = expect[:, 1] - expect[:, 0] if len(expect) else np.zeros(0, np.uint32)
                assert st.direct_copies == int((lens >= 64).sum())
                short = expect[lens < 64]
                assert st.packed_rows == sum(int(dirty[b:e].sum()) for b, e in short)
            assert np.array_equal(v.private_buffer.download(dtype).reshape(-1, per)[:n], h)
        v.destroy()


@pytest.mark.gpu
def test_growth_keeps_contents_and_marks_new_rows(ctx):
    from ravengine_b200.vram_vector import SYNC_PACKED, BufferedVRAMVector

    v = BufferedVRAMVector(ctx, 64, 1000)
    a = v.host(np.float32).reshape(-1, 16)
    a[:] = np.arange(16000, dtype=np.float32).reshape(-1, 16)
    v.encode_sync(SYNC_PACKED)
    before = a.copy()  # the view dies with the old pinned block at the next growth
    v.set_value_at(7, np.full(16, -1.0, np.float32))
    before[7] = -1.0
    v.resize(5000)  # capacity 1024 -> 8192: new pinned block, new private buffer, old rows carried over
    b = v.host(np.float32).reshape(-1, 16)
    assert np.array_equal(b[:1000], before)
    b[1000:] = 3.5
    st = v.encode_sync(SYNC_PACKED)
    assert st.ranges == 2  # row 7 (+ gap tail) and the 4000 new rows
    got = v.private_buffer.download(np.float32).reshape(-1, 16)[:5000]
    assert np.array_equal(got, b)
    assert got[7, 0] == -1.0 and got[4999, 15] == 3.5
    # the same through the GPU-pull mode, across another growth
    from ravengine_b200.vram_vector import SYNC_PULL

    keep = b.copy()
    v.resize(20000)
    c = v.host(np.float32).reshape(-1, 16)
    c[5000:] = -7.25
    c[11] = 42.0
    v.tracking()[11] = 1
    v.encode_sync(SYNC_PULL)
    got = v.private_buffer.download(np.float32).reshape(-1, 16)[:20000]
    keep[11] = 42.0
    assert np.array_equal(got[:5000], keep) and np.all(got[5000:] == -7.25)
    assert v.pulled_rows() == 15001
    v.destroy()


def _affine48(m64: np.ndarray) -> np.ndarray:
    """(n, 16) column-major mat4 rows -> (n, 12) mat4x3 rows (columns c0..c3, xyz each)."""
    return np.ascontiguousarray(m64.reshape(-1, 4, 4)[:, :, :3].reshape(-1, 12))


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 31, 50_003, 1_200_001])
def test_affine_rows_are_byte_identical_to_mat4_rows(ctx, n):
    """RVS_VECTOR_AFFINE_ROWS: 48-B host rows, 64-B private rows. Every sync mode must leave the private buffer
    byte-identical to a plain 64-B vector that holds the same (affine) matrices."""
    from ravengine_b200 import scenes
    from ravengine_b200.vram_vector import (SYNC_FULL, SYNC_FULL_PULL, SYNC_PACKED, SYNC_PULL, SYNC_RANGES, BufferedVRAMVector)

    rng = np.random.default_rng(n)
    want = scenes.random_transforms(rng, n, 500.0, 50.0).reshape(n, 16)
    assert np.all(want[:, 3::4] == np.array([0, 0, 0, 1], np.float32))  # the generator's matrices are affine
    for mode in (SYNC_RANGES, SYNC_PACKED, SYNC_PULL, SYNC_FULL, SYNC_FULL_PULL):
        v = BufferedVRAMVector(ctx, 48, n, affine_rows=True)
        ref = BufferedVRAMVector(ctx, 64, n)
        h = v.host(np.float32).reshape(n, 12)
        cur = want.copy()
        h[:] = _affine48(cur)
        ref.host(np.float32).reshape(n, 16)[:] = cur
        v.encode_sync(mode, 64)
        ref.encode_sync(SYNC_FULL)
        got = v.private_buffer.download(np.float32).reshape(-1, 16)[:n]
        assert got.tobytes() == ref.private_buffer.download(np.float32).reshape(-1, 16)[:n].tobytes() == cur.tobytes()
        for frame in range(2):
            dirty = rng.random(n) < (0.02 if frame == 0 else 0.7)
            if n > 5000:
                dirty[1000:3000] = True  # a long run: direct range in packed mode
            idx = np.nonzero(dirty)[0]
            cur[idx] = scenes.random_transforms(rng, len(idx), 100.0, 10.0).reshape(-1, 16) if len(idx) else cur[idx]
            h[idx] = _affine48(cur[idx]) if len(idx) else h[idx]
            if mode in (SYNC_FULL, SYNC_FULL_PULL):
                v.encode_sync(mode)  # flags are not consulted
            else:
                v.tracking()[idx] = 1
                st = v.encode_sync(mode, 64)
                if mode != SYNC_PULL:
                    assert np.array_equal(v.last_ranges(), host_tables.sync_ranges(dirty.astype(np.uint8)))
                assert not v.tracking().any()
            got = v.private_buffer.download(np.float32).reshape(-1, 16)[:n]
            assert got.tobytes() == cur.tobytes()
        v.destroy()
        ref.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("elem_bytes,dtype", [(64, np.float32), (4, np.uint32), (2, np.uint16)])
def test_full_modes_and_clean_vectors(ctx, elem_bytes, dtype):
    """RVS_SYNC_FULL / _FULL_PULL upload every row without reading flags; a vector nobody wrote to syncs without work."""
    from ravengine_b200.vram_vector import SYNC_FULL, SYNC_FULL_PULL, SYNC_PACKED, SYNC_PULL, BufferedVRAMVector

    rng = np.random.default_rng(elem_bytes)
    n = 70_001
    per = elem_bytes // np.dtype(dtype).itemsize
    for mode in (SYNC_FULL, SYNC_FULL_PULL):
        v = BufferedVRAMVector(ctx, elem_bytes, n)
        h = v.host(dtype).reshape(n, per)
        for _ in range(2):
            h[:] = rng.integers(0, 60000, size=(n, per)).astype(dtype)
            st = v.encode_sync(mode)
            assert st.direct_bytes == n * elem_bytes
            assert np.array_equal(v.private_buffer.download(dtype).reshape(-1, per)[:n], h)
        launches = ctx.launch_count
        for m in (SYNC_PACKED, SYNC_PULL):  # clean: no scan, no launch, no copy
            st = v.encode_sync(m)
            assert st.ranges == 0 and st.packed_bytes == 0 and ctx.launch_count == launches
        v.set_value_at(5, h[7].copy())
        h2 = h.copy()
        h2[5] = h[7]
        v.encode_sync(SYNC_PACKED)
        assert np.array_equal(v.private_buffer.download(dtype).reshape(-1, per)[:n], h2)
        v.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("affine", [True, False])
def test_forwarded_rows_land_in_the_owner_s_private_buffer(ctx, affine):
    """Upload forwarding (rve_shim.h): the owner uploads rows [0, k) of a shared-host vector itself, a helper context (here:
    a second stream of the same device, the host block mapped a second time through /proc/self/fd) reads rows [k, n) from
    the shared block and writes them into the owner's private buffer, ordered by the two flags of the protocol. The
    private buffer must equal a plain full upload, frame after frame."""
    import torch

    from ravengine_b200 import Context, scenes
    from ravengine_b200.vram_vector import SYNC_FULL, BufferedVRAMVector

    n, k = 300_007, 180_000
    eb, per = (48, 12) if affine else (64, 16)
    rng = np.random.default_rng(5)
    v = BufferedVRAMVector(ctx, eb, n, affine_rows=affine, shared_host=True)
    s2 = torch.cuda.Stream()
    helper = Context(ctx.device, s2.cuda_stream)
    pid, fd, nbytes = v.share()
    host_ptr = helper.host_import(pid, fd, nbytes)
    flags_owner, flags_helper = ctx.create_buffer(64), ctx.create_buffer(64)  # [0] "free" lives with the helper, [1] "landed" with the owner
    status = [ctx.create_buffer(16), helper.create_buffer(16)]
    h = v.host(np.float32).reshape(n, per)
    m64 = scenes.random_transforms(rng, n, 100.0, 10.0).reshape(n, 16)
    h[:] = _affine48(m64) if affine else m64
    v.encode_sync(SYNC_FULL)  # creates the private buffer
    ctx.synchronize()
    dst_h = helper.wrap_pointer(v.private_buffer.ptr, v.private_buffer.nbytes)
    f_helper_view = helper.wrap_pointer(flags_helper.ptr, 64)
    f_owner_view = helper.wrap_pointer(flags_owner.ptr, 64)
    for epoch in range(1, 4):
        m64 = scenes.random_transforms(rng, n, 100.0, 10.0).reshape(n, 16)
        h[:] = _affine48(m64) if affine else m64
        # helper first (as a faster rank would be): it must wait for the owner's "free" signal
        helper.flag_wait(f_helper_view, 0, epoch, status[1])
        helper.forward_rows(host_ptr, eb, affine, k, n - k, dst_h)
        helper.flag_signal(f_owner_view, 1, epoch)
        # owner: "my private buffer is free for this frame", own rows, wait for the forwarded rows
        ctx.flag_signal(flags_helper, 0, epoch)
        v.encode_sync_rows(SYNC_FULL, 0, k)
        ctx.flag_wait(flags_owner, 1, epoch, status[0])
        ctx.synchronize()
        helper.synchronize()
        got = v.private_buffer.download(np.float32).reshape(-1, 16)[:n]
        assert got.tobytes() == m64.tobytes()
        assert status[0].download(np.uint32)[0] == 0 and status[1].download(np.uint32)[0] == 0
    helper.host_release(host_ptr, nbytes)
    helper.close()
    v.destroy()


@pytest.mark.gpu
def test_batched_sync_of_many_vectors_matches_their_host_buffers(ctx):
    """rvs_vectors_encode_sync_batch: seven structures of different row sizes change a little per frame; those with small
    changes share ONE staging copy and ONE scatter launch, the others (2-byte rows, affine rows, a long dirty run, a vector that
    has never been synced) take their own EncodeSync - every private buffer must equal its host buffer after every frame, clean
    vectors cost nothing, and the launch count shows the sharing."""
    from ravengine_b200.vram_vector import BufferedVRAMVector, SparseSet, encode_sync_batch

    rng = np.random.default_rng(31)
    n = 20_000
    specs = [(4, np.uint32, False), (16, np.float32, False), (64, np.float32, False), (12, np.float32, False), (2, np.uint16, False),
             (48, np.float32, True)]
    vecs, hosts = [], []
    for eb, dt, affine in specs:
        v = BufferedVRAMVector(ctx, eb, n, affine_rows=affine)
        h = v.host(dt).reshape(n, eb // np.dtype(dt).itemsize)
        h[:] = rng.integers(0, 1000, size=h.shape).astype(dt)
        if affine:
            h[:] = rng.random(h.shape, dtype=np.float32)
        vecs.append(v)
        hosts.append(h)
    s = SparseSet(ctx)
    s.emplace_many(np.arange(5000, dtype=np.uint32))
    everything = vecs + [s]

    def check():
        for v, h, (eb, dt, affine) in zip(vecs, hosts, specs):
            got = v.private_buffer.download(dt)
            if affine:  # 48-B host rows land as 64-B mat4 rows: (0, 0, 0, 1) appended per column group - compare through the expansion rule
                g = got.reshape(-1, 16)[:n]
                exp = np.zeros((n, 16), np.float32)
                exp.reshape(n, 4, 4)[:, :, :3] = h.reshape(n, 4, 3)
                exp[:, 15] = 1.0
                assert np.array_equal(g, exp)
            else:
                assert np.array_equal(got.reshape(-1, h.shape[1])[:n], h)
        assert np.array_equal(s.private_buffer.download(np.uint32)[: len(s.dense())], s.dense())

    encode_sync_batch(everything)  # first sync: everything is marked, nothing qualifies as small
    ctx.synchronize()
    check()
    l0 = ctx.launch_count
    sts = encode_sync_batch(everything)  # nothing changed
    assert ctx.launch_count == l0 and all(st.ranges == 0 for st in sts)
    for frame in range(4):
        for v, h, (eb, dt, affine) in zip(vecs, hosts, specs):
            idx = rng.choice(n, size=int(rng.integers(1, 200)), replace=False)
            if frame == 2 and eb == 16:
                idx = np.arange(0, n)  # every row: more than the batch takes, this vector goes its own way (one direct copy)
            h[idx] = (rng.random((len(idx), h.shape[1])) * 100).astype(dt)
            v.tracking()[idx] = 1
        gone = rng.choice(s.dense(), size=40, replace=False)
        s.erase_many(gone)
        s.emplace_many((10_000 + frame * 100 + np.arange(40)).astype(np.uint32))
        l0 = ctx.launch_count
        sts = encode_sync_batch(everything)
        ctx.synchronize()
        shared = 5 if not (frame == 2) else 4  # 4-, 16-, 64-, 12-byte vectors + the set's list
        assert sum(1 for st in sts if st.packed_rows) >= shared
        # one launch for the shared block, one scatter each for the 2-byte and the affine vector (+ the long run's vector in frame 2: a copy, no launch)
        assert ctx.launch_count - l0 == 3, (frame, ctx.launch_count - l0)
        check()
    for v in vecs:
        v.destroy()
    s.destroy()


@pytest.mark.gpu
def test_ranges_from_the_touched_list_equal_the_reference_scan(ctx):
    """When every flag of a frame was set through SetValueAt / mark / Resize the sync derives its ranges from the list of touched
    indices instead of scanning the flags: the ranges must be the reference loop's (src/BufferedVRAMVector.cpp:106-139) for
    isolated touches, touches at every distance around maxGapSize, marks, growth, a shrink below touched rows, duplicates and
    the end of the array; a frame that took the raw tracking pointer falls back to the scan, and the frame after it is exact again."""
    from ravengine_b200.vram_vector import SYNC_PACKED, SYNC_RANGES, BufferedVRAMVector

    rng = np.random.default_rng(77)
    n = 30_000
    v = BufferedVRAMVector(ctx, 16, n)
    h = v.host(np.uint32).reshape(n, 4)
    h[:] = rng.integers(0, 1 << 30, size=(n, 4), dtype=np.uint32)
    v.encode_sync(SYNC_RANGES, 64)
    for frame in range(12):
        size = len(v)
        flags = np.zeros(size, np.uint8)
        if frame == 5:  # the raw pointer: not recorded, the sync scans
            idx = rng.choice(size, size=300, replace=False)
            v.tracking()[idx] = 1
            flags[idx] = 1
            h = v.host(np.uint32).reshape(size, 4)
            h[idx] = rng.integers(0, 1 << 30, size=(len(idx), 4), dtype=np.uint32)
        else:
            base = rng.choice(size - 40, size=150, replace=False)
            idx = np.unique(np.concatenate([base, base[:60] + rng.integers(1, 9, size=60), [0, size - 1, size - 3]]))  # gaps of 1..8 around maxGapSize 5
            for i in idx.tolist() + idx[:10].tolist():  # duplicates too
                v.set_value_at(int(i), rng.integers(0, 1 << 30, size=4, dtype=np.uint32))
            flags[idx] = 1
            if frame % 3 == 1:
                b = int(rng.integers(0, size - 500))
                v.mark(b, b + 37)
                h = v.host(np.uint32).reshape(size, 4)
                h[b:b + 37] = rng.integers(0, 1 << 30, size=(37, 4), dtype=np.uint32)
                flags[b:b + 37] = 1
            if frame % 4 == 2:  # shrink below touched rows, then grow past the old size: new rows start out modified
                v.resize(size - 200)
                v.resize(size + 100)
                flags = np.concatenate([flags[: size - 200], np.ones(300, np.uint8)])
        expect = host_tables.sync_ranges(flags.copy())
        st = v.encode_sync(SYNC_PACKED if frame % 2 else SYNC_RANGES, 64)
        assert np.array_equal(v.last_ranges(), expect), frame
        assert st.ranges == len(expect)
        m = len(v)
        assert np.array_equal(v.private_buffer.download(np.uint32).reshape(-1, 4)[:m], v.host(np.uint32).reshape(m, 4)), frame
        assert not v.tracking().any()
        v.encode_sync(SYNC_RANGES, 64)  # (the tracking() call above switched recording off for this empty frame: scan, nothing to do)
    v.destroy()
