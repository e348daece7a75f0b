"""GPU test (-m gpu) of rt_trace_rays_closest_static, the entry point that allocates nothing (static device scratch;
no counterpart in the reference, see include/rt_api.h).  Collected last on purpose (file name).
Its results must equal rt_trace_rays_closest bit for bit (the bar of every closest-hit test here), on two
calls in a row that share the scratch — the second smaller than the first, so stale contents would show.
"""
import numpy as np
import pytest

from conftest import fixture_scene, random_rays
from raytracer_b200 import librt

pytestmark = pytest.mark.gpu


def _same(a, b):
    return np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)) and np.array_equal(a[1], b[1])


def test_static_scratch_closest_hit_matches_allocating_entry_point():
    ctx = librt.Context(device=0)
    try:
        ctx.build_scene(fixture_scene("cornell_pbr"))
        cap = int(ctx.L.rt_static_max_rays())
        assert cap == 65536
        kw = dict(lo=(-4, -4, -4), hi=(4, 4, 4), target_lo=(-2, -2, -2), target_hi=(2, 2, 2))
        rays1 = random_rays(cap, 11, **kw)       # the full scratch
        rays2 = random_rays(20001, 12, **kw)     # then a shorter, ragged batch over the same memory
        ref1, ref2 = ctx.trace_rays_closest(rays1), ctx.trace_rays_closest(rays2)
        assert (ref1[1] != 0xFFFFFFFF).sum() > 1000 and (ref2[1] != 0xFFFFFFFF).sum() > 1000
        got1 = ctx.trace_rays_closest_static(rays1)
        got2 = ctx.trace_rays_closest_static(rays2)
        assert _same(got1, ref1), "first call differs from rt_trace_rays_closest"
        assert _same(got2, ref2), "second call (shared scratch) differs from rt_trace_rays_closest"
        assert _same(ctx.trace_rays_closest_static(rays1[:1]), (ref1[0][:1], ref1[1][:1]))
        tuv0, ids0 = ctx.trace_rays_closest_static(np.zeros((0, 8), np.float32))
        assert len(ids0) == 0 and len(tuv0) == 0
        # beyond the maximum: a clear error, nothing traced
        with pytest.raises(librt.RtError) as e:
            ctx.trace_rays_closest_static(random_rays(cap + 1, 13, **kw))
        assert "exceeds the static scratch" in str(e.value)
        # and the scratch still works afterwards
        assert _same(ctx.trace_rays_closest_static(rays2), ref2)
        # an edit that is not on the device yet: bringing it there allocates, so the static entry point refuses ...
        scene = fixture_scene("cornell_pbr")
        ctx.update_object_transform(0, np.asarray(scene.objects[0].transform.model_matrix, np.float32))
        with pytest.raises(librt.RtError) as e:
            ctx.trace_rays_closest_static(rays2)
        assert "does not allocate" in str(e.value)
        ctx.scene_info()  # ... until some other call has made the scene resident
        assert _same(ctx.trace_rays_closest_static(rays2), ctx.trace_rays_closest(rays2))
    finally:
        ctx.close()
