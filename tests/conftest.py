import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.oracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own code compiled into oracle/_ref (None -> tests that need it skip)."""
    import oracle_lib
    lib = oracle_lib.ref()
    if lib is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return lib


# ------------------------------------------------------------------------------------------------------------------------
# Guard bands around every device array a GPU test (or the Python wrappers under test) allocates with torch.empty /
# torch.empty_like: 4 KB of a known byte pattern before and after the array, checked when the test ends.  A kernel that
# stores outside its output -- a rim tile, a padded pitch, a level chunk one too long -- fails the test that ran it even
# when the stray store lands in memory nobody looks at.  (compute-sanitizer cannot be used on the GPU pool.)
# FIMEXCU_TEST_NO_GUARDS=1 switches the bands off.
GUARD_BYTES = 4096
GUARD_PATTERN = 0xA5
GUARD_HELD_LIMIT = 16 << 30  # bytes of guarded arrays kept alive for the end-of-test check; older ones are checked and let go


class _GuardBands:
    def __init__(self, torch):
        self.torch = torch
        self.orig_empty = torch.empty
        self.orig_empty_like = torch.empty_like
        self.blocks = []  # (raw uint8 tensor, payload bytes, description)
        self.held = 0
        self.errors = []

    def _check(self, raw, n, what):
        lo, hi = raw[:GUARD_BYTES], raw[GUARD_BYTES + n:]
        for side, band in (("before", lo), ("after", hi)):
            bad = (band != GUARD_PATTERN).nonzero()
            if bad.numel():
                first = int(bad[0]) if side == "after" else int(bad[-1]) - GUARD_BYTES
                self.errors.append("%d byte(s) written %s %s (nearest at offset %+d bytes from that end)" % (bad.numel(), side, what, first))

    def check_all(self):
        for raw, n, what in self.blocks:
            self._check(raw, n, what)
        self.blocks, self.held = [], 0

    def empty(self, *size, **kw):
        torch = self.torch
        extra = set(kw) - {"dtype", "device"}
        dev = kw.get("device")
        if extra or dev is None or torch.device(dev).type != "cuda":
            return self.orig_empty(*size, **kw)
        shape = tuple(size[0]) if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)) else tuple(int(s) for s in size)
        dtype = kw.get("dtype") or torch.get_default_dtype()
        n = 1
        for s in shape:
            n *= int(s)
        n *= self.orig_empty((), dtype=dtype).element_size()
        raw = self.orig_empty(n + 2 * GUARD_BYTES, dtype=torch.uint8, device=dev)
        raw[:GUARD_BYTES] = GUARD_PATTERN
        raw[GUARD_BYTES + n:] = GUARD_PATTERN
        if self.held + n > GUARD_HELD_LIMIT:
            self.check_all()
        self.blocks.append((raw, n, "a %s array of shape %s" % (str(dtype).replace("torch.", ""), shape)))
        self.held += n
        return raw[GUARD_BYTES:GUARD_BYTES + n].view(dtype).view(shape)

    def empty_like(self, t, **kw):
        if kw or not t.is_cuda:
            return self.orig_empty_like(t, **kw)
        return self.empty(tuple(t.shape), dtype=t.dtype, device=t.device)


@pytest.fixture(autouse=True)
def _device_guard_bands(request):
    if request.node.get_closest_marker("gpu") is None or os.environ.get("FIMEXCU_TEST_NO_GUARDS") == "1":
        yield None
        return
    import torch
    if not torch.cuda.is_available():
        yield None
        return
    g = _GuardBands(torch)
    torch.empty, torch.empty_like = g.empty, g.empty_like
    try:
        yield g
        torch.cuda.synchronize()
        g.check_all()
    finally:
        torch.empty, torch.empty_like = g.orig_empty, g.orig_empty_like
    assert not g.errors, "stores outside a device array: " + "; ".join(g.errors[:5])
