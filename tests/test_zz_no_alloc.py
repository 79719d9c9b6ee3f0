"""b200rt_intersect_closest_noalloc (include/b200rt.h; csrc/no_alloc.cu): the host-buffer closest-hit call whose device-side
staging is a file-scope static __device__ array instead of cudaMalloc'ed buffers.  No counterpart in the reference — added
for callers that cannot allocate.  Collected last on purpose;
it must pass alone in a fresh process and after every other test.

Through the C ABI, against the existing entry point b200rt_intersect_closest on the same rays: bit for bit (the tolerance of
every integer / traversal test in this repo), on two calls in a row — the second with other rays and a smaller n, so that
the array still holds the first call's records — for a BVH and a kd-tree; sizes beyond the stated maximum are rejected."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _same(a, b):
    assert np.array_equal(a["prim"], b["prim"])
    for f in ("t", "b0", "b1", "b2"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


def _noalloc(gpu, acc, rays):
    from b200rt import api
    L = gpu.lib()
    L.b200rt_intersect_closest_noalloc.restype = C.c_int
    L.b200rt_intersect_closest_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
    hits = np.zeros(len(rays), api.HIT_DTYPE)
    hits.view(np.uint8)[:] = 0x7f            # garbage the call must overwrite
    rc = L.b200rt_intersect_closest_noalloc(acc._h, rays.ctypes.data, len(rays), hits.ctypes.data, None)
    return rc, hits


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row(gpu, kind):
    from b200rt import scenes
    v = scenes.soup(4000, seed=9)
    acc = gpu.MakeAccelerator(kind, v)
    box = [-12, -12, -12, 12, 12, 12]
    first = scenes.random_rays(box, 60000, seed=3)
    second = scenes.random_rays(box, 20011, seed=4)
    for rays in (first, second):            # two calls in a row: the second finds the first call's records in the array
        rc, got = _noalloc(gpu, acc, rays)
        assert rc == 0, gpu.lib().b200rt_last_error()
        want = acc.Intersect(rays)          # the existing entry point (cudaMalloc'ed staging), same host rays
        _same(got, want)
        assert (got["prim"] != gpu.MISS).any() and (got["prim"] == gpu.MISS).any()
    # and once more with the first batch: nothing of the second call leaks into it
    rc, again = _noalloc(gpu, acc, first)
    assert rc == 0
    _same(again, acc.Intersect(first))


def test_sizes_beyond_the_static_scratch_are_rejected(gpu):
    from b200rt import scenes
    acc = gpu.MakeAccelerator("bvh", scenes.soup(200, seed=2))
    rays = scenes.random_rays([-12, -12, -12, 12, 12, 12], 65537, seed=1)
    rc, hits = _noalloc(gpu, acc, rays)
    assert rc == 1                          # B200RT_EINVAL
    msg = gpu.lib().b200rt_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "65536" in msg and "exceed" in msg
    assert (hits.view(np.uint8) == 0x7f).all()       # nothing was written
    rc, hits = _noalloc(gpu, acc, rays[:65536])      # the maximum itself is fine
    assert rc == 0
    _same(hits, acc.Intersect(rays[:65536]))
