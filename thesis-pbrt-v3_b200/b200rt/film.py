"""Film reassembly across ranks.

The reference merges FilmTiles under a mutex inside one process (Film::MergeFilmTile,
src/core/film.cpp:118-133).  Here each rank owns the tiles t with t % world == rank of the
reference's 16x16 tile grid (src/core/integrator.cpp:237-239); with the box filter no pixel is
shared between tiles, so reassembly is a pure gather of each rank's tile pixels to rank 0
(torch.distributed.gather -> NCCL send/recv over NVLink on GPUs, gloo on CPU)."""
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
    """Precomputes the per-rank pixel lists once; gather() moves only this rank's pixels."""

    def __init__(self, xres, yres, tile, rank, world, device):
        self.rank, self.world, self.device = rank, world, device
        self.xres, self.yres = xres, yres
        self.idx = [torch.from_numpy(tile_pixel_indices(xres, yres, tile, r, world)).to(device) for r in range(world)]
        self.maxlen = max(len(i) for i in self.idx)
        self.send = torch.zeros((self.maxlen, 4), dtype=torch.float32, device=device)
        self.recv = [torch.zeros((self.maxlen, 4), dtype=torch.float32, device=device) for _ in range(world)] \
            if rank == 0 else None
        self.bytes_per_gather = self.maxlen * 16 * (world - 1)

    def gather(self, film):
        """film: (yres, xres, 4) tensor holding this rank's tiles.  Returns the full frame on rank 0
        (the same tensor, completed in place), None elsewhere."""
        flat = film.view(-1, 4)
        if self.world == 1:
            return film
        mine = self.idx[self.rank]
        self.send[: len(mine)] = flat[mine]
        dist.gather(self.send, self.recv, dst=0)
        if self.rank != 0:
            return None
        for r in range(1, self.world):
            flat[self.idx[r]] = self.recv[r][: len(self.idx[r])]
        return film
