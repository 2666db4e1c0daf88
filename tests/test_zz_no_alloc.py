"""hs_dump_compact_noalloc — the allocation-free variant of hs_dump_compact (static device scratch,
no counterpart in the reference; include/hideseek/c_api.h) — against the existing entry point,
through the C ABI, on two calls in a row.  The records are integer / copied float state, so the
bar is the one tests/compact.py uses for exact comparisons: byte-identical.  The second call
dumps fewer, different worlds into the same scratch, so anything the first call left behind
(contact rows of a busier world, for one) would show.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from compact import compact_dt, hs_dump_compact, hs_set_debug

pytestmark = pytest.mark.gpu


def _noalloc(sim, w0, count):
    lib = sim._clib
    lib.hs_dump_compact_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64]
    out = np.full(count, 0, dtype=compact_dt)
    out.view(np.uint8)[:] = 0xA5                     # every byte must come from the call
    rc = lib.hs_dump_compact_noalloc(sim._h, w0, count, out.ctypes.data, out.nbytes)
    return rc, out


def test_noalloc_dump_equals_allocating_dump_twice():
    import gpu_hideseek as gh
    N, A = 200, 4
    sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=3)
    hs_set_debug(sim, 1)
    sim.init()
    act = sim.action_tensor().to_torch()
    gen = torch.Generator(device="cpu").manual_seed(17)
    for _ in range(30):
        a = torch.zeros(act.shape, dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 11, (N * A, 3), generator=gen, dtype=torch.int32)
        a[:, 3] = (torch.rand(N * A, generator=gen) < 0.3).int()
        act.copy_(a.to(act.device))
        sim.step()
    lib = sim._clib
    cap = lib.hs_dump_compact_noalloc_max()
    assert cap == 64
    assert lib.hs_compact_dump_bytes() == compact_dt.itemsize

    # call 1: a full scratch; call 2, right after it: fewer and other worlds, ordered by falling
    # contact count against call 1 so that stale rows would be visible
    for w0, count in ((7, cap), (130, 10)):
        want = hs_dump_compact(sim, w0, count)
        rc, got = _noalloc(sim, w0, count)
        assert rc == 0, lib.hs_last_error()
        assert got.tobytes() == want.tobytes(), (w0, count)
        assert int(got["num_contacts"].sum()) > 0 and int(got["num_bodies"].min()) > 0
    # and once more the first range: the array is shared, not consumed
    rc, again = _noalloc(sim, 7, cap)
    assert rc == 0 and again.tobytes() == hs_dump_compact(sim, 7, cap).tobytes()

    # beyond the static maximum: a clear error, nothing written
    rc, untouched = _noalloc(sim, 0, cap + 1)
    assert rc != 0
    msg = lib.hs_last_error().decode()
    assert "at most 64" in msg, msg
    assert (untouched.view(np.uint8) == 0xA5).all()
    # bad ranges are rejected like the allocating entry point rejects them
    assert _noalloc(sim, N - 3, 10)[0] != 0
    assert sim.overflow_count() == 0
