"""GPU (-m gpu), collected LAST on purpose: the allocation-free entry point
cuorb_is_in_frustum_noalloc (include/cuorb.h; no counterpart in the reference)
against the existing cuorb_is_in_frustum, through the C ABI, on two calls in a row.

Bar: the same as every other test of this path - bit-exact on every output array and on the returned count (the two
entry points launch the same kernel; only where the staging area lives differs: a fixed `static __device__` array
instead of a cudaMalloc'ed workspace).  The second call is smaller than the first and uses other inputs, so anything
the shared array kept from the first call (stale points beyond n, the previous in-view count) would show up.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUTS = (("in_view", np.uint8), ("proj_x", np.float32), ("proj_y", np.float32), ("proj_xr", np.float32),
        ("level", np.int32), ("view_cos", np.float32), ("depth", np.float32))


@pytest.fixture(scope="module")
def lib():
    from orbslam3_cpp_fork_b200 import _lib
    L = _lib.load()
    if L.cuorb_device_count() < 1:
        pytest.fail("no CUDA device: libcuorb has no CPU fallback")
    return L


def _call(fn, cam, P, N, min_d, max_d, limit):
    from orbslam3_cpp_fork_b200 import _lib
    n = len(min_d)
    out = {k: np.full(n, 99, dt) for k, dt in OUTS}          # poisoned: every entry must be written by the call
    rc = fn(0, C.byref(cam), n, _lib.ptr(P), _lib.ptr(N), _lib.ptr(min_d), _lib.ptr(max_d), C.c_float(limit),
            *[_lib.ptr(out[k]) for k, _ in OUTS])
    return rc, out


def _case(n, seed):
    from orbslam3_cpp_fork_b200 import _lib
    from tests.matching_cases import frustum_case
    cam, P, N, min_d, max_d = frustum_case(n, seed)
    ccam = _lib.FrustumCamera.from_buffer_copy(bytes(cam))   # same POD layout (include/cuorb.h cuorb_frustum_camera)
    return ccam, np.ascontiguousarray(P, np.float32), np.ascontiguousarray(N, np.float32), min_d, max_d


def test_noalloc_equals_existing_entry_point_two_calls_in_a_row(lib):
    from orbslam3_cpp_fork_b200 import _lib
    # expected results from the existing (allocating) entry point, computed first
    cases = [(_case(20000, 11), 0.5), (_case(5000, 12), 0.8)]
    want = [_call(lib.cuorb_is_in_frustum, *c, lim) for c, lim in cases]
    for rc, _ in want:
        _lib.check(rc)
    assert want[0][0] > 2000 and want[1][0] > 100            # the cases really have points in view
    # two calls of the new entry point, back to back, the second one smaller and on other inputs
    got = [_call(lib.cuorb_is_in_frustum_noalloc, *c, lim) for c, lim in cases]
    for i, ((rc_w, w), (rc_g, g)) in enumerate(zip(want, got)):
        _lib.check(rc_g)
        assert rc_g == rc_w, f"call {i}: in-view count {rc_g} != {rc_w}"
        for k, _ in OUTS:
            assert g[k].tobytes() == w[k].tobytes(), f"call {i}: {k} differs"     # bit-exact, floats included
    # and once more the first case: a larger call after a smaller one
    rc, g = _call(lib.cuorb_is_in_frustum_noalloc, *cases[0][0], cases[0][1])
    assert rc == want[0][0] and all(g[k].tobytes() == want[0][1][k].tobytes() for k, _ in OUTS)


def test_noalloc_limits_and_empty(lib):
    from orbslam3_cpp_fork_b200 import _lib
    cam, P, N, min_d, max_d = _case(64, 3)
    header = open(os.path.join(ROOT, "include", "cuorb.h")).read()
    nmax = int(re.search(r"#define\s+CUORB_FRUSTUM_NOALLOC_MAX\s+(\d+)", header).group(1))
    # exactly the maximum is accepted and equals the existing entry point
    camM, PM, NM, mnM, mxM = _case(nmax, 5)
    rc_w, w = _call(lib.cuorb_is_in_frustum, camM, PM, NM, mnM, mxM, 0.5)
    rc_g, g = _call(lib.cuorb_is_in_frustum_noalloc, camM, PM, NM, mnM, mxM, 0.5)
    assert _lib.check(rc_g) == _lib.check(rc_w)
    assert all(g[k].tobytes() == w[k].tobytes() for k, _ in OUTS)
    # one more than the maximum is rejected with a clear error and writes nothing
    n = nmax + 1
    big = [np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32), np.ones(n, np.float32), np.ones(n, np.float32)]
    rc, out = _call(lib.cuorb_is_in_frustum_noalloc, cam, *big, 0.5)
    assert rc == _lib.E_CAPACITY
    msg = _lib.last_error()
    assert "CUORB_FRUSTUM_NOALLOC_MAX" in msg and str(nmax) in msg, msg
    assert (out["in_view"] == 99).all()
    # n == 0 is a no-op
    assert lib.cuorb_is_in_frustum_noalloc(0, C.byref(cam), 0, None, None, None, None, C.c_float(0.5), *([None] * 7)) == 0
    # the array survives an error: a normal call still gives the right answer
    rc_w, w = _call(lib.cuorb_is_in_frustum, cam, P, N, min_d, max_d, 0.5)
    rc_g, g = _call(lib.cuorb_is_in_frustum_noalloc, cam, P, N, min_d, max_d, 0.5)
    assert rc_g == rc_w and all(g[k].tobytes() == w[k].tobytes() for k, _ in OUTS)


def test_noalloc_body_calls_no_allocator():
    """The entry point's own source (and the helper it uses) names no allocating call."""
    src = open(os.path.join(ROOT, "orbslam3_cpp_fork_b200", "csrc", "frameglue.cu")).read()
    a = src.index("unsigned char* frustum_scratch_ptr(int device)")
    body = src[a:]
    assert "int cuorb_is_in_frustum_noalloc(" in body
    code = re.sub(r"//[^\n]*", "", body)                      # comments may name what is NOT called
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "t_scratch", "new ", "malloc("):
        assert banned not in code, banned
    assert "cudaGetSymbolAddress" in code and "g_frustum_scratch" in code
