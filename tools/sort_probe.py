#!/usr/bin/env python
"""Per-kernel times of the entity-list sort (rve_sort_entity_ids) on a random list and on a churned contiguous range (what the
engine's lists look like), for a launch list:

  python tools/sort_probe.py && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv \
      --log-file gpurun_out/sort_launches.csv python tools/sort_probe.py
"""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from ravengine_b200 import Context
st = torch.cuda.Stream()
ctx = Context(0, st.cuda_stream)
rng = np.random.default_rng(4)
n_ids, id_space = 2_500_000, 10_000_000
rand = rng.choice(id_space, size=n_ids, replace=False).astype(np.uint32)
# churned contiguous range like C2's lists
from ravengine_b200.vram_vector import SparseSet
ss = SparseSet(ctx); base = np.arange(2_500_000, 5_000_000, dtype=np.uint32); ss.emplace_many(base)
gone = rng.choice(base, size=250_000, replace=False); ss.erase_many(gone); ss.emplace_many(rng.permutation(gone)); churn = ss.dense().copy()
for name, ids in (("random", rand), ("churned_range", churn)):
    src = ctx.buffer_from_numpy(ids); dst = ctx.create_buffer(4 * n_ids)
    for _ in range(3):
        ctx.sort_entity_ids(src, n_ids, id_space, dst)
    ctx.synchronize()
    assert np.array_equal(dst.download(np.uint32)[:n_ids], np.sort(ids)), name
print("ok")
