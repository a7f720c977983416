"""om_render_read_view_noalloc: the allocation-free variant of om_render_read_view (fixed file-scope device scratch, no
counterpart in the reference).  Collected last on purpose.
Called through the C ABI and held to the existing entry point's bytes - RGB24 is integer output, so the
bar is equality - on calls in a row that would expose stale scratch (the reductions of modes 1 and 3 live there)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _handle(api, W, H, spp, seed):
    fz = api.FrozenHittableList.generate("random_scene", 1)
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, W / H, 0.1, 10.0)
    r = api.Renderer(fz, cam, api.render_params(W, H, spp, seed=seed))
    r.launch(0)
    r.sync()
    return r


def test_noalloc_view_equals_read_view_on_calls_in_a_row(api):
    big = _handle(api, 160, 96, 6, 5)     # more samples, other depths: its reductions must not leak into the next call
    small = _handle(api, 96, 64, 2, 9)
    try:
        want = {(id(r), m): r.read_view(m) for r in (big, small) for m in range(7)}
        # two calls in a row per mode, then the handles and the reduced modes interleaved
        order = [(r, m) for r in (big, small) for m in range(7) for _ in range(2)]
        order += [(big, 3), (small, 3), (big, 1), (small, 1), (small, 3), (big, 3), (small, 0), (big, 6)]
        buf = {id(big): np.zeros((96, 160, 3), np.uint8), id(small): np.zeros((64, 96, 3), np.uint8)}
        for r, m in order:
            got = r.read_view_noalloc(m, out=buf[id(r)])   # the caller's buffer is reused too
            assert (got == want[(id(r), m)]).all(), "mode %d differs from om_render_read_view" % m
        assert want[(id(big), 1)].max() > 0 and want[(id(small), 3)].max() > 0  # the reduced modes show something
    finally:
        big.close()
        small.close()


def test_noalloc_view_rejects_what_does_not_fit(api, A):
    r = _handle(api, 2048, 1025, 1, 3)    # 2 099 200 pixels > OM_VIEW_NOALLOC_MAX_PIXELS
    try:
        assert r.npx > A.VIEW_NOALLOC_MAX_PIXELS
        with pytest.raises(api.OmError) as e:
            r.read_view_noalloc(0)
        assert e.value.status == A.OM_ERR_INVALID and "OM_VIEW_NOALLOC_MAX_PIXELS" in str(e.value)
        assert r.read_view(0).shape == (1025, 2048, 3)   # the allocating entry point still serves it
    finally:
        r.close()
    s = _handle(api, 16, 16, 1, 3)
    try:
        with pytest.raises(api.OmError) as e:
            s.read_view_noalloc(7)
        assert e.value.status == A.OM_ERR_INVALID
    finally:
        s.close()
