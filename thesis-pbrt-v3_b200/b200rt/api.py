"""ctypes binding of include/b200rt.h + the Accelerator mirror (see package docstring)."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
MISS = 0xFFFFFFFF
RAY_DTYPE = np.dtype([("o", "<f4", 3), ("tMax", "<f4"), ("d", "<f4", 3), ("time", "<f4")])
HIT_DTYPE = np.dtype([("prim", "<u4"), ("t", "<f4"), ("b0", "<f4"), ("b1", "<f4"), ("b2", "<f4")])
RAY_STATS_DTYPE = np.dtype([("prim_tests", "<u4"), ("interior_visits", "<u4"), ("leaf_visits", "<u4"), ("nodes_fetched", "<u4")])
COUNTERS_FIELDS = ["rays", "nodes_fetched", "nodes_passed", "leaf_visits", "prim_tests", "index_reads"]
KINDS = {"bvh": 0, "kdtree": 1, "rbsp": 2, "rbspkd": 3, "bsp": 4, "bspkd": 5}
ERRORS = {1: "EINVAL", 2: "ENODEVICE", 3: "ECUDA", 4: "EUNSUPPORTED", 5: "ENOMEM"}


class B200RTError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200rt %s: %s" % (ERRORS.get(code, code), msg))
        self.code = code


class SceneDesc(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("nodes", C.c_void_p), ("node_count", C.c_uint32),
                ("prim_indices", C.c_void_p), ("n_indices", C.c_uint32), ("directions", C.c_void_p),
                ("n_directions", C.c_uint32), ("bounds", C.c_float * 6), ("tri_verts", C.c_void_p),
                ("tri_uvs", C.c_void_p), ("n_prims", C.c_uint32), ("prim_kind", C.c_void_p),
                ("sphere_params", C.c_void_p), ("n_spheres", C.c_uint32),
                ("tri_normals", C.c_void_p), ("prim_flags", C.c_void_p)]


class CountersStruct(C.Structure):
    _fields_ = [(f, C.c_uint64) for f in COUNTERS_FIELDS]


def lib_path():
    # B200RT_LIB: another build of the same library, for same-box A/B measurements only (scripts/ab_shade_inline.sh)
    return os.environ.get("B200RT_LIB") or os.path.join(_HERE, "libb200rt.so")


_lib = None


def lib():
    """Loads libb200rt.so.  Fails loudly when the CUDA extension has not been built — there is no
    fallback implementation."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise B200RTError(2, "libb200rt.so is not built (run `python thesis-pbrt-v3_b200/build.py`); "
                                 "b200rt has no CPU fallback")
        L = C.CDLL(p)
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
        L.b200rt_init.argtypes = [i32]
        L.b200rt_last_error.restype = C.c_char_p
        L.b200rt_version.restype = i32
        L.b200rt_launch_count.restype = u64
        L.b200rt_scene_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(vp)]
        L.b200rt_scene_destroy.argtypes = [vp]
        L.b200rt_scene_destroy.restype = None
        L.b200rt_scene_bounds.argtypes = [vp, C.POINTER(C.c_float * 6)]
        L.b200rt_intersect_closest.argtypes = [vp, vp, u64, vp, vp]
        L.b200rt_intersect_any.argtypes = [vp, vp, u64, vp, vp]
        L.b200rt_intersect_closest_async.argtypes = [vp, vp, u64, vp, vp]
        L.b200rt_intersect_any_async.argtypes = [vp, vp, u64, vp, vp]
        L.b200rt_intersect_wait.argtypes = [vp]
        L.b200rt_intersect_closest_counted.argtypes = [vp, vp, u64, vp, C.POINTER(CountersStruct), vp]
        L.b200rt_intersect_any_counted.argtypes = [vp, vp, u64, vp, C.POINTER(CountersStruct), vp]
        L.b200rt_intersect_closest_stats.argtypes = [vp, vp, u64, vp, vp, vp]
        L.b200rt_intersect_any_stats.argtypes = [vp, vp, u64, vp, vp, vp]
        L.b200rt_build_bvh.argtypes = [vp, u32, i32, i32, i32, C.POINTER(vp), C.POINTER(u32), C.POINTER(vp)]
        L.b200rt_build_kdtree.argtypes = [vp, u32, i32, i32, C.c_float, i32, i32, C.POINTER(vp), C.POINTER(u32),
                                          C.POINTER(vp), C.POINTER(u32), C.POINTER(C.c_float * 6)]
        L.b200rt_free.argtypes = [vp]
        L.b200rt_free.restype = None
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise B200RTError(rc, lib().b200rt_last_error().decode())


def init(device=0):
    _check(lib().b200rt_init(device))


def launch_count():
    return int(lib().b200rt_launch_count())


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _take(ptr, n, width, dtype=np.uint32):
    if n == 0:
        return np.zeros((0, width), dtype)
    arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint32)), (n, width)).copy()
    lib().b200rt_free(ptr)
    return arr.view(dtype) if dtype is not np.uint32 else arr


def make_rays(o, d, tmax=np.inf, time=0.0):
    """Packs origins/directions into the 32-byte ray records of b200rt_ray."""
    o = np.asarray(o, np.float32).reshape(-1, 3)
    r = np.zeros((len(o), 8), np.float32)
    r[:, 0:3] = o
    r[:, 3] = tmax
    r[:, 4:7] = np.asarray(d, np.float32).reshape(-1, 3)
    r[:, 7] = time
    return r


def build_bvh(tri_verts, maxnodeprims=4, intersectcost=8, traversalcost=1):
    """Host sweep-SAH BVH builder (csrc/build_bvh.cpp), params of CreateBVHAccelerator
    (reference src/accelerators/bvh.cpp:529-535).  Returns a reference-format tree dict."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    nodes, order, nc = C.c_void_p(), C.c_void_p(), C.c_uint32()
    _check(lib().b200rt_build_bvh(_p(v), len(v), maxnodeprims, intersectcost, traversalcost, C.byref(nodes),
                                  C.byref(nc), C.byref(order)))
    n = _take(nodes, nc.value, 8)
    o = _take(order, len(v), 1).reshape(-1)
    bounds = n[0, :6].view(np.float32).copy() if len(n) else np.zeros(6, np.float32)
    return dict(kind="bvh", nodes=n, indices=o, directions=np.zeros((0, 3), np.float32), bounds=bounds,
                n_prims=len(v))


def build_kdtree(tri_verts, intersectcost=80, traversalcost=1, emptybonus=0.0, maxprims=1, maxdepth=-1):
    """Host SAH kd-tree builder (csrc/build_kd.cpp), params of CreateKdTreeAccelerator
    (reference src/accelerators/kdtreeaccel.cpp:523-545)."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    nodes, idx, nc, ni = C.c_void_p(), C.c_void_p(), C.c_uint32(), C.c_uint32()
    b = (C.c_float * 6)()
    _check(lib().b200rt_build_kdtree(_p(v), len(v), intersectcost, traversalcost, emptybonus, maxprims, maxdepth,
                                     C.byref(nodes), C.byref(nc), C.byref(idx), C.byref(ni), C.byref(b)))
    n = _take(nodes, nc.value, 2)
    if ni.value:
        i = _take(idx, ni.value, 1).reshape(-1)
    else:
        lib().b200rt_free(idx)
        i = np.zeros(0, np.uint32)
    dirs = np.eye(3, dtype=np.float32)
    return dict(kind="kdtree", nodes=n, indices=i, directions=dirs, bounds=np.array(list(b), np.float32),
                n_prims=len(v))


def build_bvh_device(tri_verts, maxnodeprims=4, intersectcost=8, traversalcost=1):
    """build_bvh on the DEVICE (b200rt_build_bvh_device, csrc/bvh_build.cuh): nodes and primNumMapping byte-identical to
    build_bvh's."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    nodes, order, nc = C.c_void_p(), C.c_void_p(), C.c_uint32()
    st = BuildStats()
    L = lib()
    L.b200rt_build_bvh_device.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    _check(L.b200rt_build_bvh_device(_p(v), len(v), int(maxnodeprims), int(intersectcost), int(traversalcost),
                                     C.byref(nodes), C.byref(nc), C.byref(order), C.byref(st)))
    n = _take(nodes, nc.value, 8)
    o = _take(order, len(v), 1).reshape(-1)
    bounds = n[0, :6].view(np.float32).copy() if len(n) else np.zeros(6, np.float32)
    return dict(kind="bvh", nodes=n, indices=o, directions=np.zeros((0, 3), np.float32), bounds=bounds, n_prims=len(v),
                build=dict(levels=st.levels, candidates=st.candidates, kernel_launches=st.kernel_launches, build_ms=st.build_ms))


def build_kdtree_device(tri_verts, intersectcost=80, traversalcost=1, emptybonus=0.0, maxprims=1, maxdepth=-1):
    """build_kdtree on the DEVICE (b200rt_build_kdtree_device: the level-synchronous builder of csrc/rbsp_build.cuh in its
    kd mode).  Node array identical to build_kdtree's; the order of the indices inside a leaf may differ."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    nodes, idx, nc, ni = C.c_void_p(), C.c_void_p(), C.c_uint32(), C.c_uint32()
    b = (C.c_float * 6)()
    st = BuildStats()
    L = lib()
    L.b200rt_build_kdtree_device.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    _check(L.b200rt_build_kdtree_device(_p(v), len(v), int(intersectcost), int(traversalcost), float(emptybonus), int(maxprims),
                                        int(maxdepth), C.byref(nodes), C.byref(nc), C.byref(idx), C.byref(ni), C.byref(b), C.byref(st)))
    n = _take(nodes, nc.value, 2)
    if ni.value:
        i = _take(idx, ni.value, 1).reshape(-1)
    else:
        L.b200rt_free(idx)
        i = np.zeros(0, np.uint32)
    return dict(kind="kdtree", nodes=n, indices=i, directions=np.eye(3, dtype=np.float32), bounds=np.array(list(b), np.float32),
                n_prims=len(v), build=dict(levels=st.levels, candidates=st.candidates, kernel_launches=st.kernel_launches,
                                           build_ms=st.build_ms))


class BuildStats(C.Structure):
    _fields_ = [("levels", C.c_uint32), ("candidates", C.c_uint64), ("kernel_launches", C.c_uint64), ("build_ms", C.c_float)]


def build_rbsp(tri_verts, nbDirections=3, intersectcost=80, traversalcost=5, emptybonus=0.0, maxprims=1, maxdepth=-1,
               kdtraversalcost=None):
    """RBSP tree built ON THE DEVICE (csrc/build_rbsp.cu, level-synchronous: every candidate plane of a tree level cut and
    measured in parallel), params of CreateRBSPTreeAccelerator (reference src/accelerators/rbsp.cpp:549-570).  The node
    array equals the reference builder's byte for byte.  tree["build"] holds levels / candidates / launches / build_ms.
    kdtraversalcost not None selects RBSPKd::buildTree (rbspKd.cpp:187-470, CreateRBSPKdTreeAccelerator's default 1)."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
    nodes, idx, nc, ni = C.c_void_p(), C.c_void_p(), C.c_uint32(), C.c_uint32()
    b = (C.c_float * 6)()
    dirs = np.zeros((16, 3), np.float32)
    st = BuildStats()
    L = lib()
    L.b200rt_build_rbsp.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.b200rt_build_rbspkd.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    if kdtraversalcost is None:
        _check(L.b200rt_build_rbsp(_p(v), len(v), int(nbDirections), int(intersectcost), int(traversalcost), float(emptybonus),
                                   int(maxprims), int(maxdepth), C.byref(nodes), C.byref(nc), C.byref(idx), C.byref(ni),
                                   C.byref(b), _p(dirs), C.byref(st)))
    else:
        _check(L.b200rt_build_rbspkd(_p(v), len(v), int(nbDirections), int(intersectcost), int(traversalcost), int(kdtraversalcost),
                                     float(emptybonus), int(maxprims), int(maxdepth), C.byref(nodes), C.byref(nc), C.byref(idx),
                                     C.byref(ni), C.byref(b), _p(dirs), C.byref(st)))
    n = _take(nodes, nc.value, 2)
    if ni.value:
        i = _take(idx, ni.value, 1).reshape(-1)
    else:
        L.b200rt_free(idx)
        i = np.zeros(0, np.uint32)
    return dict(kind="rbsp" if kdtraversalcost is None else "rbspkd", nodes=n, indices=i, directions=dirs[:int(nbDirections)].copy(), bounds=np.array(list(b), np.float32),
                n_prims=len(v), build=dict(levels=st.levels, candidates=st.candidates, kernel_launches=st.kernel_launches,
                                           build_ms=st.build_ms))


class B200Accel:
    """Aggregate over a device-resident tree.  Mirrors pbrt's Aggregate (src/core/primitive.h:49-56,
    188-198) for ray BATCHES: WorldBound(), Intersect(rays) -> hits, IntersectP(rays) -> occluded."""

    def __init__(self, tree, tri_verts, tri_uvs=None, prim_kind=None, spheres=None, tri_normals=None, prim_flags=None):
        """prim_kind: optional (n_prims,) uint8 with 1 = full sphere; spheres: (n_spheres,16) float32 rows
        (WorldToObject 3x4, radius, pad) — see b200rt_scene_desc in include/b200rt.h.
        tri_normals: optional (n_prims, 9) world-space vertex normals (TriangleMesh::n); prim_flags: optional (n_prims,)
        uint8, bit 0 = reverseOrientation ^ transformSwapsHandedness, bit 1 = the primitive has vertex normals (set for
        every primitive when tri_normals is given and prim_flags is not)."""
        self.tree = tree
        v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
        pk = None if prim_kind is None else np.ascontiguousarray(prim_kind, np.uint8)
        sp = None if spheres is None else np.ascontiguousarray(spheres, np.float32).reshape(-1, 16)
        uv = None if tri_uvs is None else np.ascontiguousarray(tri_uvs, np.float32).reshape(-1, 6)
        tn = None if tri_normals is None else np.ascontiguousarray(tri_normals, np.float32).reshape(-1, 9)
        if tn is not None and prim_flags is None:
            prim_flags = np.full(len(v), 2, np.uint8)
        pf = None if prim_flags is None else np.ascontiguousarray(prim_flags, np.uint8)
        nodes = np.ascontiguousarray(tree["nodes"], np.uint32)
        idx = np.ascontiguousarray(tree["indices"], np.uint32)
        dirs = tree.get("directions")
        dirs = None if dirs is None or len(dirs) == 0 else np.ascontiguousarray(dirs, np.float32)
        d = SceneDesc()
        d.kind = KINDS[tree["kind"]]
        d.nodes, d.node_count = _p(nodes), len(nodes)
        d.prim_indices, d.n_indices = _p(idx), len(idx)
        d.directions, d.n_directions = _p(dirs), 0 if dirs is None else len(dirs)
        d.bounds = (C.c_float * 6)(*[float(x) for x in np.asarray(tree["bounds"], np.float32)])
        d.tri_verts, d.tri_uvs, d.n_prims = _p(v), _p(uv), len(v)
        d.prim_kind, d.sphere_params, d.n_spheres = _p(pk), _p(sp), 0 if sp is None else len(sp)
        d.tri_normals, d.prim_flags = _p(tn), _p(pf)
        h = C.c_void_p()
        _check(lib().b200rt_scene_create(C.byref(d), C.byref(h)))
        self._h = h
        self.kind = tree["kind"]
        self.n_prims = len(v)

    # -- Aggregate interface ------------------------------------------------------------------
    def WorldBound(self):
        b = (C.c_float * 6)()
        _check(lib().b200rt_scene_bounds(self._h, C.byref(b)))
        return np.array(list(b), np.float32)

    def Intersect(self, rays, counters=False, stream=None):
        """rays: (n,8) float32 host array.  Returns hits (HIT_DTYPE); with counters=True also a dict."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), HIT_DTYPE)
        if counters:
            c = CountersStruct()
            _check(lib().b200rt_intersect_closest_counted(self._h, _p(rays), len(rays), _p(hits), C.byref(c), stream))
            return hits, {f: int(getattr(c, f)) for f in COUNTERS_FIELDS}
        _check(lib().b200rt_intersect_closest(self._h, _p(rays), len(rays), _p(hits), stream))
        return hits

    def IntersectStats(self, rays, any_hit=False, stream=None):
        """Instrumented Intersect / IntersectP: (hits or occluded, per-ray RAY_STATS_DTYPE) — the reference's
        Ray::stats after each call (src/core/geometry.h:1078-1173)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        st = np.zeros(len(rays), RAY_STATS_DTYPE)
        if any_hit:
            out = np.zeros(len(rays), np.uint8)
            _check(lib().b200rt_intersect_any_stats(self._h, _p(rays), len(rays), _p(out), _p(st), stream))
        else:
            out = np.zeros(len(rays), HIT_DTYPE)
            _check(lib().b200rt_intersect_closest_stats(self._h, _p(rays), len(rays), _p(out), _p(st), stream))
        return out, st

    def IntersectP(self, rays, counters=False, stream=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint8)
        if counters:
            c = CountersStruct()
            _check(lib().b200rt_intersect_any_counted(self._h, _p(rays), len(rays), _p(occ), C.byref(c), stream))
            return occ, {f: int(getattr(c, f)) for f in COUNTERS_FIELDS}
        _check(lib().b200rt_intersect_any(self._h, _p(rays), len(rays), _p(occ), stream))
        return occ

    # -- raw-pointer variants (device buffers; used by bench.py and the wavefront driver) -----
    def intersect_ptr(self, rays_ptr, n, hits_ptr, stream=None):
        _check(lib().b200rt_intersect_closest(self._h, rays_ptr, n, hits_ptr, stream))

    def intersect_p_ptr(self, rays_ptr, n, occ_ptr, stream=None):
        _check(lib().b200rt_intersect_any(self._h, rays_ptr, n, occ_ptr, stream))

    # -- enqueue + wait (HOST buffers that stay alive and untouched until wait() returns): several queues in flight at once,
    #    e.g. the shadow queue and the next closest-hit queue of one bounce
    def intersect_ptr_async(self, rays_ptr, n, hits_ptr, stream=None):
        _check(lib().b200rt_intersect_closest_async(self._h, rays_ptr, n, hits_ptr, stream))

    def intersect_p_ptr_async(self, rays_ptr, n, occ_ptr, stream=None):
        _check(lib().b200rt_intersect_any_async(self._h, rays_ptr, n, occ_ptr, stream))

    def wait(self):
        _check(lib().b200rt_intersect_wait(self._h))

    def close(self):
        if getattr(self, "_h", None):
            lib().b200rt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_BVH_DEFAULTS = dict(maxnodeprims=4, intersectcost=8, traversalcost=1)
_KD_DEFAULTS = dict(intersectcost=80, traversalcost=1, emptybonus=0.0, maxprims=1, maxdepth=-1)
_RBSP_DEFAULTS = dict(nbDirections=3, intersectcost=80, traversalcost=5, emptybonus=0.0, maxprims=1, maxdepth=-1)
_IGNORED = {"splitalpha", "alphatype", "axisselectiontype", "axisselectionamount", "nbDirections"}


def MakeAccelerator(name, tri_verts, paramSet=None, tri_uvs=None, tree=None, tri_normals=None, prim_flags=None, builder="device"):
    """Mirror of MakeAccelerator(name, prims, paramSet) (reference src/core/api.cpp:790-1015) for
    the accelerators on the hot path.  `paramSet` uses the .pbrt parameter names.  "bvh", "kdtree", "rbsp" and "rbspkd" are built ON THE
    DEVICE by the level-synchronous builders (build_bvh_device / build_kdtree_device / build_rbsp; `builder="host"` selects the
    serial host builders for bvh / kdtree — same nodes); the general BSP families take a host-built tree (`tree=`,
    reference node format).  Unknown names return None after a warning, like the reference (api.cpp:1011-1012)."""
    ps = dict(paramSet or {})
    if tree is not None:
        return B200Accel(tree, tri_verts, tri_uvs, tri_normals=tri_normals, prim_flags=prim_flags)
    if name == "bvh":
        args = {k: ps.pop(k, v) for k, v in _BVH_DEFAULTS.items()}
        tree = (build_bvh_device if builder == "device" else build_bvh)(tri_verts, **args)
    elif name == "kdtree":
        args = {k: ps.pop(k, v) for k, v in _KD_DEFAULTS.items()}
        tree = (build_kdtree_device if builder == "device" else build_kdtree)(tri_verts, **args)
    elif name == "rbsp":
        args = {k: ps.pop(k, v) for k, v in _RBSP_DEFAULTS.items()}
        tree = build_rbsp(tri_verts, **args)
    elif name == "rbspkd":
        args = {k: ps.pop(k, v) for k, v in _RBSP_DEFAULTS.items()}
        tree = build_rbsp(tri_verts, kdtraversalcost=ps.pop("kdtraversalcost", 1), **args)
    else:
        import warnings
        warnings.warn('Accelerator "%s" unknown.' % name)
        return None
    for k in ps:
        if k not in _IGNORED:
            import warnings
            warnings.warn('Parameter "%s" not used' % k)  # api.cpp:1013 paramSet.ReportUnused()
    return B200Accel(tree, tri_verts, tri_uvs, tri_normals=tri_normals, prim_flags=prim_flags)
