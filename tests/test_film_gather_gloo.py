"""CPU, world_size 2 over gloo: the N>1 path's host logic — tile ownership (tile t -> rank t % world),
the pixel lists, and the film gather to rank 0."""
import os
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))


def _worker(rank, world, port, xres, yres, tile, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200rt.film import FilmGather, tile_pixel_indices
    g = FilmGather(xres, yres, tile, rank, world, "cpu")
    film = torch.zeros((yres, xres, 4))
    mine = torch.from_numpy(tile_pixel_indices(xres, yres, tile, rank, world))
    # this rank "renders" its own tiles: value = f(pixel)
    film.view(-1, 4)[mine] = torch.stack([mine.float(), mine.float() * 2, mine.float() + 7, torch.ones(len(mine))], 1)
    full = g.gather(film)
    if rank == 0:
        np.save(out, full.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_reassembles_frame(tmp_path, world):
    xres, yres, tile = 70, 45, 16  # ragged edge tiles
    out = str(tmp_path / "film.npy")
    port = 29500 + os.getpid() % 2000 + world
    mp.spawn(_worker, args=(world, port, xres, yres, tile, out), nprocs=world, join=True)
    full = np.load(out)
    pix = np.arange(xres * yres, dtype=np.float32)
    want = np.stack([pix, pix * 2, pix + 7, np.ones_like(pix)], 1).reshape(yres, xres, 4)
    assert np.array_equal(full, want)


def test_tiles_partition_the_image():
    from b200rt.film import tile_pixel_indices
    for world in (1, 2, 4, 8):
        allpix = np.concatenate([tile_pixel_indices(700, 700, 16, r, world) for r in range(world)])
        assert len(allpix) == 700 * 700 and len(np.unique(allpix)) == 700 * 700


def test_render_tiles_matches_host_lists(built_lib):
    """b200rt_render_tiles (device-side tiling) agrees with the host pixel lists."""
    from b200rt.film import tile_pixel_indices
    from b200rt import scenes
    cam = scenes.look_at_camera((0, 0, 0), (1, 0, 0), (0, 1, 0), 40, 700, 700)
    for world in (1, 2, 8):
        for rank in range(world):
            w = built_lib.WavefrontPathIntegrator(cam, rank=rank, world=world)
            ntiles, npix = w.tiles()
            assert npix == len(tile_pixel_indices(700, 700, 16, rank, world))
