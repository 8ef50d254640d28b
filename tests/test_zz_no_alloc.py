"""vspg_debug_majorant_segments_noalloc (include/vspg.h): the majorant-segment hook staged in the library's fixed
file-scope device array instead of cudaMalloc'd blocks.  No counterpart in the reference;
checked against the allocating entry point, bit for bit, on two calls in a row -- the second one
smaller, with fewer segment slots, over a scratch array that still holds the first call's data.

Collected last (pytest --collect-only)."""
import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A
from pyvspg import api

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def own_ctx():
    """A context of its own: the test passes alone in a fresh process and after the other tests' session context."""
    c = api.Context(0)
    yield c
    c.close()


def _setup(ctx, name):
    cfg = H.small(name)
    ctx.set_scene(cfg.scene)
    ctx.set_guiding(cfg.guiding)
    ctx.set_integrator(cfg.ip)
    return cfg


def _same(a, b):
    (seg_a, cnt_a), (seg_b, cnt_b) = a, b
    assert np.array_equal(cnt_a, cnt_b)
    assert np.array_equal(seg_a.view(np.uint32), seg_b.view(np.uint32))     # unfilled slots included: both read zero


def test_noalloc_majorant_segments_equal_the_allocating_entry_point_twice_in_a_row(own_ctx):
    ctx = own_ctx
    cfg = _setup(ctx, "C2")
    rng = np.random.default_rng(21)
    rays1 = H.random_rays(rng, 4096, cfg.scene.eye)
    want1 = ctx.debug_majorant_segments(0, rays1, 64)
    got1 = ctx.debug_majorant_segments_noalloc(0, rays1, 64)
    _same(got1, want1)
    assert want1[1].max() > 8 and (want1[1] == 0).any()
    # second call, straight after: other rays, fewer of them, another slot count -- stale segments of call 1 must not show
    rays2 = H.random_rays(rng, 1000, cfg.scene.eye)[24:]          # (skip the axis-aligned ones: mostly short walks and misses)
    want2 = ctx.debug_majorant_segments(0, rays2, 24)
    got2 = ctx.debug_majorant_segments_noalloc(0, rays2, 24)
    _same(got2, want2)
    assert (got2[0][want2[1] == 0] == 0).all()                    # rays that miss the medium: all-zero rows
    # and the first inputs again reproduce the first outputs
    _same(ctx.debug_majorant_segments_noalloc(0, rays1, 64), want1)


def test_noalloc_on_another_medium_type_and_context(own_ctx):
    c2 = api.Context(0)        # a second context on the same device shares the scratch array (sequential use is fine)
    try:
        cfg = _setup(c2, "C5")     # rgbgrid: sigma_maj differs per channel
        rays = H.random_rays(np.random.default_rng(22), 2048, cfg.scene.eye)
        _same(c2.debug_majorant_segments_noalloc(0, rays, 48), c2.debug_majorant_segments(0, rays, 48))
    finally:
        c2.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(own_ctx):
    ctx = own_ctx
    cfg = _setup(ctx, "C2")
    rays = H.random_rays(np.random.default_rng(23), A.NOALLOC_MAX_RAYS + 1, cfg.scene.eye)
    with pytest.raises(api.VspgError) as e:
        ctx.debug_majorant_segments_noalloc(0, rays, 4)
    assert e.value.code == A.ERR_INVALID and "exceeds the fixed scratch" in str(e.value)
    with pytest.raises(api.VspgError) as e:
        ctx.debug_majorant_segments_noalloc(0, rays[:4097], 64)          # 4097 x 64 > 262144 segments
    assert e.value.code == A.ERR_INVALID and "exceeds the fixed scratch" in str(e.value)
    # the limits themselves are accepted, and n = 0 is a no-op
    seg, cnt = ctx.debug_majorant_segments_noalloc(0, rays[:A.NOALLOC_MAX_RAYS], A.NOALLOC_MAX_SEGMENTS // A.NOALLOC_MAX_RAYS)
    assert cnt.shape == (A.NOALLOC_MAX_RAYS,) and cnt.max() > 0
    seg, cnt = ctx.debug_majorant_segments_noalloc(0, rays[:0], 8)
    assert cnt.shape == (0,)
