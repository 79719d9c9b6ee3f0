"""Ray-set files: little-endian header {magic u32, version u32, n u64, flags u32, pad u32}
followed by n x 8 float32 (ox, oy, oz, tMax, dx, dy, dz, time) — the b200rt_ray record.  Used to hand
identical ray sets (camera rays R0, wavefront queues R1..Rk, random rays Rrand) to the GPU path and to
the CPU checkers."""
import struct
import numpy as np

MAGIC = 0x59523242  # "B2RY"
FLAG_ANY_HIT = 1    # the set is meant for IntersectP (shadow rays)
_HDR = struct.Struct("<IIQII")


def save(path, rays, any_hit=False):
    r = np.ascontiguousarray(rays, "<f4").reshape(-1, 8)
    with open(path, "wb") as f:
        f.write(_HDR.pack(MAGIC, 1, len(r), FLAG_ANY_HIT if any_hit else 0, 0))
        f.write(r.tobytes())


def load(path):
    """Returns (rays (n,8) float32, flags)."""
    with open(path, "rb") as f:
        hdr = f.read(_HDR.size)
        if len(hdr) != _HDR.size:
            raise ValueError("%s: truncated header" % path)
        magic, ver, n, flags, _ = _HDR.unpack(hdr)
        if magic != MAGIC or ver != 1:
            raise ValueError("%s: not a b200rt ray file (magic %#x version %d)" % (path, magic, ver))
        data = f.read()
    if len(data) != n * 32:
        raise ValueError("%s: expected %d rays (%d bytes), file holds %d bytes" % (path, n, n * 32, len(data)))
    return np.frombuffer(data, "<f4").reshape(-1, 8).copy(), flags
