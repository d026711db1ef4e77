"""rt_trace_primary_noalloc (include/rt_b200.h): the allocation-free variant of rt_trace_primary_ex -- same kernel, the device
staging area is the library's file-scope __device__ array instead of a cudaMalloc'ed buffer.  No counterpart in the reference;
added for callers that cannot allocate.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
from conftest import bits

pytestmark = pytest.mark.gpu


def _scene(rt):
    hs = rt.synthetic_scene(3000, W=256, H=144, seed=77)
    sp, pl, mt, _ = hs.arrays()
    return rt.Scene(sp, pl, mt), rt.pixel_center_rays(hs.camera(0))


def test_noalloc_trace_equals_the_allocating_entry_point_twice_in_a_row(rt):
    """through the C-ABI, two calls in a row with different rays (the shared static array keeps the first call's contents):
    ids equal and t BIT-equal to rt_trace_primary / rt_trace_primary_ex -- the bar of the other id / t tests -- in exact and
    in production arithmetic; and the free device memory does not move across the calls"""
    import torch
    sc, rays = _scene(rt)
    assert len(rays) <= rt.lib().rt_noalloc_max_rays() == 262144
    rng = np.random.default_rng(3)
    rays2 = rays[rng.permutation(len(rays))[: len(rays) // 3]].copy()   # a shorter batch: stale results stay behind it
    rays2[:, 3:] += rng.normal(0, 0.02, (len(rays2), 3)).astype(np.float32)
    want = {(id(r), ex): sc.trace_primary(r, exact=ex) for r in (rays, rays2) for ex in (True, False)}
    sc.trace_primary_noalloc(rays[:16])                                  # set-up (symbol address, module load) before measuring
    out = tuple(np.empty(len(rays), d) for d in (np.int32, np.int32, np.float32))
    free0 = torch.cuda.mem_get_info(0)[0]
    for ex in (True, False):
        for r in (rays, rays2, rays):                                    # long, short, long again
            o = tuple(a[: len(r)] for a in out)
            ty, ix, t = sc.trace_primary_noalloc(r, exact=ex, out=o)
            wy, wi, wt = want[(id(r), ex)]
            assert (ty == wy).all() and (ix == wi).all() and (bits(t) == bits(wt)).all()
            assert (ty >= 0).mean() > 0.2
    assert torch.cuda.mem_get_info(0)[0] == free0                        # nothing was allocated (or freed) on the device
    sc.close()


def test_noalloc_trace_rejects_batches_beyond_the_static_scratch(rt):
    sc, rays = _scene(rt)
    n = rt.lib().rt_noalloc_max_rays() + 1
    big = np.resize(rays, (n, 6))
    with pytest.raises(rt.RtError) as e:
        sc.trace_primary_noalloc(big)
    assert e.value.code == 1 and "RT_NOALLOC_MAX_RAYS" in str(e.value) and "262144" in str(e.value)
    ty, ix, t = sc.trace_primary_noalloc(big[: n - 1])                   # exactly the maximum is accepted
    wy, wi, wt = sc.trace_primary(big[: n - 1])
    assert (ty == wy).all() and (ix == wi).all() and (bits(t) == bits(wt)).all()
    assert sc.trace_primary_noalloc(np.zeros((0, 6), np.float32))[0].shape == (0,)
    sc.close()
