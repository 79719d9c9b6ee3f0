"""ctypes binding for oracle/liboracle.so (the CPU restatement in oracle/oracle.c).
TEST INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")

HIT = np.dtype([("prim", "<u4"), ("t", "<f4"), ("b0", "<f4"), ("b1", "<f4"), ("b2", "<f4")])
COUNTERS = np.dtype([("rays", "<u8"), ("nodes_fetched", "<u8"), ("nodes_passed", "<u8"), ("leaf_visits", "<u8"),
                     ("prim_tests", "<u8"), ("index_reads", "<u8")])
MISS = 0xFFFFFFFF
KIND = {"bvh": 0, "kdtree": 1, "rbsp": 2, "rbspkd": 3, "bsp": 4, "bspkd": 5}

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        common = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                  C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_intersect_closest.argtypes = common
        L.orc_intersect_any.argtypes = common
        L.orc_degenerate_flags.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_triangle.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.orc_triangle.restype = C.c_int
        L.orc_sphere.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_sphere.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def degenerate_flags(tri_verts, tri_uvs=None):
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    uv = None if tri_uvs is None else np.ascontiguousarray(tri_uvs, np.float32).reshape(-1, 6)
    out = np.zeros(len(v), np.uint8)
    lib().orc_degenerate_flags(_p(v), _p(uv), len(v), _p(out))
    return out


class OracleAccel:
    """Reference-format tree (dict from refbind.RefAccel.export() or b200rt builders) + triangles
    in original primitive order; runs the restated Intersect / IntersectP loops."""

    def __init__(self, tree, tri_verts, tri_uvs=None, prim_kind=None, spheres=None):
        """prim_kind: optional (n_prims,) uint8, 1 = sphere; spheres: (n_spheres, 16) float32 = WorldToObject
        rows 0..2 (3x4), radius, 3 pad — the k-th sphere primitive uses row k."""
        self.prim_kind = None if prim_kind is None else np.ascontiguousarray(prim_kind, np.uint8)
        self.spheres = None if spheres is None else np.ascontiguousarray(spheres, np.float32).reshape(-1, 16)
        self.kind = KIND[tree["kind"]]
        self.nodes = np.ascontiguousarray(tree["nodes"], np.uint32)
        self.indices = np.ascontiguousarray(tree["indices"], np.uint32)
        d = tree.get("directions")
        self.directions = np.ascontiguousarray(d if d is not None and len(d) else np.zeros((3, 3)), np.float32)
        self.M = len(self.directions)
        self.bounds = np.ascontiguousarray(tree["bounds"], np.float32)
        self.verts = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
        self.degen = degenerate_flags(self.verts, tri_uvs)

    def _call(self, fn, rays, out, counters, per_ray, n_threads):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        cnt = np.zeros(1, COUNTERS) if counters else None
        pr = np.zeros(len(rays), COUNTERS) if per_ray else None
        fn(self.kind, _p(self.nodes), _p(self.indices), _p(self.directions), self.M, _p(self.bounds), _p(self.verts),
           _p(self.degen), len(self.verts), _p(rays), len(rays), _p(out), _p(cnt), _p(pr), n_threads,
           _p(self.prim_kind), _p(self.spheres))
        res = [out]
        if counters:
            res.append(cnt[0])
        if per_ray:
            res.append(pr)
        return res[0] if len(res) == 1 else tuple(res)

    def intersect(self, rays, counters=False, per_ray=False, n_threads=1):
        n = np.asarray(rays).size // 8
        return self._call(lib().orc_intersect_closest, rays, np.zeros(n, HIT), counters, per_ray, n_threads)

    def intersect_p(self, rays, counters=False, per_ray=False, n_threads=1):
        n = np.asarray(rays).size // 8
        return self._call(lib().orc_intersect_any, rays, np.zeros(n, np.uint8), counters, per_ray, n_threads)


def triangle(ray, verts, degenerate=False):
    """Single Triangle::Intersect; returns (hit, t, b0, b1, b2)."""
    r = np.ascontiguousarray(ray, np.float32).reshape(8)
    v = np.ascontiguousarray(verts, np.float32).reshape(9)
    o = [C.c_float() for _ in range(4)]
    hit = lib().orc_triangle(_p(r), _p(v), int(degenerate), *[C.byref(x) for x in o])
    return (bool(hit),) + tuple(x.value for x in o)


def sphere_params(centers_radii):
    """(n,4) (cx,cy,cz,radius) -> (n,16) sphere parameter rows for a pure-Translate ObjectToWorld
    (scenes/killeroo-simple:20-24): WorldToObject = Translate(-c)."""
    cr = np.asarray(centers_radii, np.float32).reshape(-1, 4)
    out = np.zeros((len(cr), 16), np.float32)
    out[:, 0] = out[:, 5] = out[:, 10] = 1
    out[:, 3], out[:, 7], out[:, 11] = -cr[:, 0], -cr[:, 1], -cr[:, 2]
    out[:, 12] = cr[:, 3]
    return out


def sphere(ray, params):
    """Single Sphere::Intersect; returns (hit, t)."""
    r = np.ascontiguousarray(ray, np.float32).reshape(8)
    sp = np.ascontiguousarray(params, np.float32).reshape(16)
    t = C.c_float()
    hit = lib().orc_sphere(_p(r), _p(sp), C.byref(t))
    return bool(hit), t.value
