"""b200rt_intersect_closest_async / _any_async / b200rt_intersect_wait (include/b200rt.h): enqueue + wait over HOST buffers.
Against the synchronous calls on the same rays, bit for bit, with several queues in flight at once and more chunks than the
pipeline has staging slots."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pinned(a):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_enqueue_then_wait_equals_the_blocking_calls(gpu, kind):
    import torch
    from b200rt import api, scenes
    v = scenes.soup(6000, seed=5)
    acc = gpu.MakeAccelerator(kind, v)
    box = [-12, -12, -12, 12, 12, 12]
    # 1.6 M rays = 7+ chunks of the 256 Ki-ray pipeline stage (4 staging slots), then two smaller queues behind it
    sizes = [1_600_000, 300_001, 70_000]
    rays = [_pinned(scenes.random_rays(box, n, seed=11 + i).astype(np.float32).reshape(-1, 8)) for i, n in enumerate(sizes)]
    hits = [torch.empty((n, 5), dtype=torch.int32).pin_memory() for n in sizes]
    occ = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes]
    for h in hits:
        h.fill_(0x7f7f7f7f)
    for o in occ:
        o.fill_(0x7f)
    for r, h, o in zip(rays, hits, occ):                      # everything enqueued, nothing waited for
        acc.intersect_p_ptr_async(r.data_ptr(), len(r), o.data_ptr())
        acc.intersect_ptr_async(r.data_ptr(), len(r), h.data_ptr())
    acc.wait()
    for r, h, o in zip(rays, hits, occ):
        want = acc.Intersect(r.numpy())
        got = h.numpy().view(api.HIT_DTYPE).reshape(-1)
        assert np.array_equal(got["prim"], want["prim"])
        for f in ("t", "b0", "b1", "b2"):
            assert np.array_equal(got[f].view(np.uint32), want[f].view(np.uint32)), f
        assert np.array_equal(o.numpy(), acc.IntersectP(r.numpy()).astype(np.uint8))
        assert (got["prim"] != gpu.MISS).any() and (got["prim"] == gpu.MISS).any()
    # a second round on the same buffers (the ring's slots have all been used), a blocking call in between
    acc.intersect_ptr_async(rays[1].data_ptr(), len(rays[1]), hits[1].data_ptr())
    mid = acc.Intersect(rays[2].numpy())
    acc.wait()
    assert np.array_equal(hits[1].numpy().view(api.HIT_DTYPE).reshape(-1)["prim"], acc.Intersect(rays[1].numpy())["prim"])
    assert np.array_equal(mid["prim"], hits[2].numpy().view(api.HIT_DTYPE).reshape(-1)["prim"])


def test_device_buffers_are_refused_by_the_enqueue_calls(gpu):
    import torch
    from b200rt import scenes
    acc = gpu.MakeAccelerator("bvh", scenes.soup(200, seed=2))
    r = torch.zeros((64, 8), dtype=torch.float32, device="cuda")
    h = torch.zeros((64, 5), dtype=torch.int32, device="cuda")
    with pytest.raises(Exception):
        acc.intersect_ptr_async(r.data_ptr(), 64, h.data_ptr())
    acc.wait()
