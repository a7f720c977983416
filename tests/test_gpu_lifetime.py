"""Scene lifetime (include/ottomarcher.h, "Lifetime"): a render handle pins its scene; freeze / clear / destroy of the scene are refused
with OM_ERR_INVALID while a handle lives and succeed afterwards, and the handle keeps rendering the same bits throughout."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_scene_cannot_change_under_a_live_handle(api, A):
    from raytracingoneweekend_b200 import _lib
    L = _lib.load()
    fz = api.FrozenHittableList.generate("random_scene", 1)
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, 1.5, 0.1, 10.0)
    p = api.render_params(48, 32, 2, seed=4)
    r = api.Renderer(fz, cam, p)
    r.launch(0); r.sync()
    first = r.read_pixels().copy()
    r2 = api.Renderer(fz, cam, p)
    for name, args in (("om_scene_freeze", (fz.h, None, 0)), ("om_scene_clear", (fz.h,)), ("om_scene_destroy", (fz.h,))):
        assert getattr(L, name)(*args) == A.OM_ERR_INVALID, name
        assert b"render handle" in L.om_last_error()
    r2.close()
    assert L.om_scene_freeze(fz.h, None, 0) == A.OM_ERR_INVALID          # one handle is still alive
    r.reset(); r.launch(0); r.sync()                                        # ... and still renders from an intact scene
    again = r.read_pixels()
    assert (again["sum"].view(np.uint32) == first["sum"].view(np.uint32)).all() and (again["bloom"] == first["bloom"]).all()
    r.close()
    assert L.om_scene_freeze(fz.h, None, 0) == A.OM_OK                    # no handles left: the scene may be re-frozen
    r3 = api.Renderer(fz, cam, p)
    r3.launch(0); r3.sync()
    assert (r3.read_pixels()["sum"].view(np.uint32) == first["sum"].view(np.uint32)).all()
    r3.close()
