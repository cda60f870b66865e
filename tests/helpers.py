"""Shared test helpers: run the same cull inputs through the oracle and through the C-ABI."""
from __future__ import annotations

import numpy as np

import oracle
from oracle import host_tables
from ravengine_b200 import CullingUBO


def make_ubo(view_proj_cm, cam_pos, num_cubos, layers=0xFFFFFFFF, single=False, shadow=False) -> CullingUBO:
    u = CullingUBO()
    u.viewProj[:] = [float(x) for x in view_proj_cm]
    u.camPos[:] = [float(x) for x in cam_pos]
    u.numCubos = num_cubos
    u.cameraRenderLayers = layers
    u.singleInstanceModeAndShadowMode = (1 if single else 0) | (2 if shadow else 0)
    return u


def oracle_cull(tables, ubo, model, layers, flags, pyr, dim, levels, groups=None, nthreads=1, taskflow=False):
    """Runs the oracle in place on copies of the table's output arrays; returns the bindless list."""
    bl = [None if b is None else b.copy() for b in tables.bindless]
    groups = (tables.total_entities + 63) // 64 if groups is None else groups
    oracle.cull(ubo, tables.cubos, model, layers, flags, pyr, dim, levels, bl, groups, nthreads, taskflow)
    return bl


def device_cull(ctx, tables, ubo, model, layers, flags, depth=None, pyr_width=1024, groups=None, pyramid_texels=None, deferred=False):
    """Same inputs through rve_cull. Returns the downloaded bindless list (same indexing as the oracle's)."""
    dev = [ctx.buffer_from_numpy(b) for b in tables.bindless]
    cubos = tables.cubos.copy()
    for f in ("idOutputBufferBindlessHandle", "entityIDInputBufferBindlessHandle", "indirectOutputBufferBindlessHandle",
              "lodDistanceBufferBindlessHandle"):
        # oracle handle -> device handle; values that name no oracle buffer (a dirty table tail) are passed through as they are
        cubos[f] = np.array([dev[int(h)].bindless if int(h) < len(dev) else int(h) for h in tables.cubos[f]], dtype=np.uint32)
    d_cubos = ctx.buffer_from_numpy(cubos.view(np.uint8))
    d_model, d_layers, d_flags = ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)
    pyr = ctx.create_pyramid(pyr_width)
    if depth is not None:
        d_depth = ctx.buffer_from_numpy(depth)
        ctx.build_pyramid(pyr, d_depth, depth.shape[1], depth.shape[0], deferred=deferred)
    elif pyramid_texels is not None:
        pyr.texels.upload(pyramid_texels)
    groups = (tables.total_entities + 63) // 64 if groups is None else groups
    ctx.cull(ubo, d_cubos, d_model, d_layers, d_flags, pyr, groups)
    ctx.synchronize()
    out = [b.download(np.uint8).view(h.dtype)[: h.size].reshape(h.shape) for b, h in zip(dev, tables.bindless)]
    pyr_host = pyr.texels.download(np.float32)
    for b in dev + [d_cubos, d_model, d_layers, d_flags]:
        b.destroy()
    return out, pyr_host


def compare_cull_outputs(tables, materials, ref, got):
    """Counts identical, each (draw, LOD) slice identical after sorting, every other byte untouched."""
    ci = 0
    for m in materials:
        n = len(m.commands)
        cb0 = tables.cubos[ci]
        h_cull, h_ind = int(cb0["idOutputBufferBindlessHandle"]), int(cb0["indirectOutputBufferBindlessHandle"])
        np.testing.assert_array_equal(ref[h_ind], got[h_ind], err_msg="indirect commands differ")
        rs = host_tables.sorted_slices(ref[h_ind], ref[h_cull], tables.cubos, range(ci, ci + n))
        gs = host_tables.sorted_slices(got[h_ind], got[h_cull], tables.cubos, range(ci, ci + n))
        for a, b in zip(rs, gs):
            np.testing.assert_array_equal(a, b, err_msg="visible id slice differs")
        # untouched slots keep the 0xFFFFFFFF fill on both sides
        assert int((ref[h_cull] != 0xFFFFFFFF).sum()) == int((got[h_cull] != 0xFFFFFFFF).sum())
        ci += n
    # inputs are never written
    for i, (a, b) in enumerate(zip(ref, got)):
        if a is not None and a.dtype == np.float32:
            np.testing.assert_array_equal(a, b)


def footmin_chain(texels, dim, levels):
    """The derived chain rve_cull samples (include/rve_b200.h, rve_depth_pyramid_footmin), from a texel chain: per level the
    min over the clamped 2x2 footprint {i-1,i} x {j-1,j}, an (n+1)^2 image, levels back to back."""
    import numpy as np

    out, off = [], 0
    for lv in range(levels):
        n = dim >> lv
        p = np.pad(texels[off:off + n * n].reshape(n, n), 1, mode="edge")
        out.append(np.minimum(np.minimum(p[:-1, :-1], p[:-1, 1:]), np.minimum(p[1:, :-1], p[1:, 1:])).ravel())
        off += n * n
    return np.concatenate(out)
