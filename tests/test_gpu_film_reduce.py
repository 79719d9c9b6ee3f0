"""b200rt_film_reduce (include/b200rt_wavefront.h): the ranks' films summed into rank 0's over CUDA IPC peer memory, the exchange
step of a render started as one process per rank without a communicator (the C++ `Integrator "b200wavefront"`).  Two real
processes share the box's GPU here; with one GPU per rank the same mapping goes over NVLink."""
import multiprocessing as mp
import os
import numpy as np
import pytest


def _rank_main(rank, world, rdv, seed, on_device, q):
    try:
        import torch
        from b200rt import film as bf
        rng = np.random.default_rng(seed + rank)
        film = rng.random((64, 96, 4), dtype=np.float32)
        heat = rng.integers(0, 1000, (6, 64, 96), dtype=np.uint32)
        if on_device:
            f = torch.from_numpy(film).cuda()
            h = torch.from_numpy(heat.view(np.int32)).cuda()
            out = bf.film_reduce_ipc(f, rank, world, rdv, heat=h)
            res = (f.cpu().numpy(), h.cpu().numpy().view(np.uint32))
        else:
            out = bf.film_reduce_ipc(film, rank, world, rdv, heat=heat)
            res = (film, heat)
        q.put((rank, out is not None, res[0], res[1], None))
    except Exception as e:   # noqa: BLE001
        q.put((rank, False, None, None, repr(e)))


@pytest.mark.gpu
@pytest.mark.parametrize("on_device", [False, True])
def test_three_ranks_sum_in_rank_order(tmp_path, on_device):
    world, seed = 3, 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_rank_main, args=(r, world, str(tmp_path), seed, on_device, q)) for r in (2, 1, 0)]
    for p in ps:
        p.start()
    got = {}
    for _ in ps:
        r, is_root, f, h, err = q.get(timeout=240)
        assert err is None, err
        got[r] = (is_root, f, h)
    for p in ps:
        p.join(60)
    films = [np.random.default_rng(seed + r).random((64, 96, 4), dtype=np.float32) for r in range(world)]
    heats = []
    for r in range(world):
        g = np.random.default_rng(seed + r)
        g.random((64, 96, 4), dtype=np.float32)
        heats.append(g.integers(0, 1000, (6, 64, 96), dtype=np.uint32))
    assert got[0][0] and not got[1][0] and not got[2][0]
    assert np.array_equal(got[0][1], (films[0] + films[1]) + films[2])        # rank order, bit for bit
    assert np.array_equal(got[0][2], heats[0] + heats[1] + heats[2])
    assert np.array_equal(got[1][1], films[1]) and np.array_equal(got[2][2], heats[2])   # peers keep their own
    assert os.listdir(tmp_path) == []


def test_bad_arguments_fail_before_any_device_work(tmp_path):
    from b200rt import film as bf, api
    f = np.zeros((4, 4, 4), np.float32)
    assert bf.film_reduce_ipc(f, 0, 1, "") is f                      # world 1: nothing to do
    with pytest.raises(api.B200RTError, match="rendezvous"):
        bf.film_reduce_ipc(f, 0, 2, "")
    with pytest.raises(api.B200RTError, match="not a directory"):
        bf.film_reduce_ipc(f, 0, 2, str(tmp_path / "missing"))
    with pytest.raises(api.B200RTError, match="rank"):
        bf.film_reduce_ipc(f, 2, 2, str(tmp_path))
