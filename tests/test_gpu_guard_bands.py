"""The guard bands of tests/conftest.py must notice a stray store (otherwise their silence over the other GPU tests means nothing)."""
import pytest

pytestmark = pytest.mark.gpu


def test_guard_bands_notice_a_store_outside_the_array(_device_guard_bands):
    import torch
    g = _device_guard_bands
    assert g is not None, "guard bands are switched off (FIMEXCU_TEST_NO_GUARDS)"
    t = torch.empty((3, 5), dtype=torch.float32, device="cuda")
    assert t.shape == (3, 5) and t.is_contiguous() and t.data_ptr() % 512 == 0
    like = torch.empty_like(t)
    assert like.shape == t.shape and like.data_ptr() != t.data_ptr()
    assert torch.empty(0, dtype=torch.int16, device="cuda").numel() == 0
    t.fill_(1.0)
    g.check_all()
    assert g.errors == []
    # one element past the end, one before the start
    u = torch.empty(7, dtype=torch.float32, device="cuda")
    torch.as_strided(u, (8,), (1,))[7] = 2.0
    g.check_all()
    assert len(g.errors) == 1 and "after" in g.errors[0] and "+0" in g.errors[0], g.errors
    g.errors.clear()
    v = torch.empty(7, dtype=torch.int16, device="cuda")
    raw = g.blocks[-1][0]
    raw[4095] = 0
    g.check_all()
    assert len(g.errors) == 1 and "before" in g.errors[0] and "-1 " in g.errors[0], g.errors
    g.errors.clear()
    del v


def test_wrapper_outputs_are_guarded(_device_guard_bands):
    """the arrays the Python mirrors hand to the library come from the guarded allocator"""
    import numpy as np
    import torch
    import fimex_b200
    g = _device_guard_bands
    rng = np.random.default_rng(5)
    px, py = rng.uniform(0, 30, 77 * 13), rng.uniform(0, 20, 77 * 13)
    ci = fimex_b200.CachedInterpolation("x", "y", 1, px, py, 31, 21, 77, 13)
    before = len(g.blocks)
    fld = rng.normal(size=(3, 21, 31)).astype(np.float32)
    out, _ = ci.interpolateValues(torch.from_numpy(ci.reduceInput(fld, 31, 21)).cuda())
    torch.cuda.synchronize()
    assert len(g.blocks) > before and any(b[0].data_ptr() + 4096 == out.data_ptr() for b in g.blocks)
