"""Film reassembly across ranks.

The reference merges FilmTiles under a mutex inside one process (Film::MergeFilmTile,
src/core/film.cpp:118-133).  Here each rank owns the tiles t with t % world == rank of the
reference's 16x16 tile grid (src/core/integrator.cpp:237-239) and the frame is reassembled on rank 0
(see FilmGather)."""
import numpy as np
import torch
import torch.distributed as dist


def tile_pixel_indices(xres, yres, tile, rank, world):
    """Flat pixel indices (y*xres + x) of the tiles owned by `rank`, in (tile, row, column) order."""
    ntx, nty = (xres + tile - 1) // tile, (yres + tile - 1) // tile
    out = []
    for t in range(rank, ntx * nty, world):
        tx, ty = t % ntx, t // ntx
        xs = np.arange(tx * tile, min(xres, (tx + 1) * tile))
        ys = np.arange(ty * tile, min(yres, (ty + 1) * tile))
        out.append((ys[:, None] * xres + xs[None, :]).ravel())
    return np.concatenate(out) if out else np.zeros(0, np.int64)


class FilmGather:
    """Reassembles the frame on rank 0.

    Each rank's film holds its own tiles and — for samples that sit exactly on a pixel boundary (FilmTile::AddSample's
    footprint then covers the neighbouring pixel too, film.h:136-141; Halton's first samples do) — contributions to the
    one-pixel halo around them, which may belong to another rank's tile.  Every other pixel is zero, so the frame is the
    SUM of the ranks' films: one reduce to rank 0 (NCCL over NVLink / NVSwitch on GPUs, gloo in the CPU tests).  A pixel
    that received no halo contribution is `own + 0 + ... + 0`, i.e. bit-identical to the single-rank film."""

    def __init__(self, xres, yres, tile, rank, world, device):
        self.rank, self.world, self.device = rank, world, device
        self.xres, self.yres = xres, yres
        self.bytes_per_gather = xres * yres * 16 * (world - 1)

    def gather(self, film):
        """film: (yres, xres, 4) tensor holding this rank's tiles (+ halo contributions).  Returns the full frame on
        rank 0 (the same tensor, completed in place), None elsewhere."""
        if self.world == 1:
            return film
        dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)
        return film if self.rank == 0 else None


def film_reduce_ipc(film, rank, world, rendezvous, heat=None, timeout_s=0.0):
    """The same sum without a communicator: b200rt_film_reduce (CUDA IPC peer memory, handles through files in `rendezvous`),
    the call the C++ `Integrator "b200wavefront"` makes when `pbrt` is started once per rank.  film: CUDA or CPU float32
    tensor / array of (yres, xres, 4); heat: optional uint32 planes.  Completed in place on rank 0."""
    import ctypes as C
    from . import api

    def ptr(t):
        return (t.data_ptr(), t.numel()) if hasattr(t, "data_ptr") else (t.ctypes.data, t.size)
    fp, fn = ptr(film)
    hp, hn = ptr(heat) if heat is not None else (None, 0)
    L = api.lib()
    L.b200rt_film_reduce.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_double]
    api._check(L.b200rt_film_reduce(str(rendezvous).encode(), rank, world, fp, fn, hp, hn, float(timeout_s)))
    return film if rank == 0 else None
