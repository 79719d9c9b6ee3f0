"""ctypes binding for oracle/_ref/libpbrtref.so — the UNMODIFIED reference hot path (see
oracle/ref/ref_api.h).  TEST INFRASTRUCTURE ONLY: imported by tests/, by the golden-vector generator
(tests/golden/make_golden.py) and by bench.py's cpu_baseline / --impl reference legs.  Never by the
product path."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libpbrtref.so")

REF_HIT = np.dtype([("prim", "<i4"), ("t", "<f4"), ("uv", "<f4", 2), ("p", "<f4", 3)])
REF_CNT = np.dtype([("prim_tests", "<u4"), ("interior_visits", "<u4"), ("leaf_visits", "<u4")])


class TreeInfo(C.Structure):
    _fields_ = [("node_count", C.c_uint32), ("node_stride", C.c_uint32), ("n_indices", C.c_uint32),
                ("n_directions", C.c_uint32), ("n_prims", C.c_uint32), ("bounds", C.c_float * 6)]


def available():
    return os.path.exists(LIB_PATH)


_lib = None


def bind(path):
    """Declares the ref_api.h prototypes on the shared library at `path` (libpbrtref.so, or the host-plugin
    test library tests/hostplugin builds on top of the same objects)."""
    if True:
        L = C.CDLL(path)
        L.ref_scene_create.restype = C.c_void_p
        L.ref_scene_create.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        if hasattr(L, "ref_scene_create2"):
            L.ref_scene_create2.restype = C.c_void_p
            L.ref_scene_create2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        if hasattr(L, "ref_trace_enable"):
            L.ref_trace_enable.argtypes = [C.c_int]
            L.ref_trace_records.restype = C.c_uint64
            L.ref_trace_copy.argtypes = [C.c_void_p]
            L.ref_camera_matrices.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_scene_destroy.argtypes = [C.c_void_p]
        L.ref_scene_num_prims.argtypes = [C.c_void_p]
        L.ref_loopsubdiv.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_accel_create.restype = C.c_void_p
        L.ref_accel_create.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_accel_destroy.argtypes = [C.c_void_p]
        L.ref_accel_info.argtypes = [C.c_void_p, C.POINTER(TreeInfo)]
        L.ref_accel_export.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_intersect_p.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_prim_intersect.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]
        L.ref_prim_intersect_p.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]
        if hasattr(L, "ref_render"):
            L.ref_render.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_char_p, C.c_int, C.c_char_p, C.c_void_p,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    return L


def lib():
    global _lib
    if _lib is None:
        _lib = bind(LIB_PATH)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RefScene:
    """vector<shared_ptr<Primitive>>: one GeometricPrimitive per triangle, then the spheres."""

    def __init__(self, P, indices, uv=None, spheres=None, L=None, normals=None, reverse_orientation=False):
        """normals: optional (n_verts, 3) per-vertex normals (TriangleMesh::n); reverse_orientation: the shape's
        `ReverseOrientation` flag."""
        self.L = L or lib()
        self.P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        self.idx = np.ascontiguousarray(indices, np.int32).reshape(-1, 3)
        self.uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
        self.spheres = None if spheres is None else np.ascontiguousarray(spheres, np.float32).reshape(-1, 4)
        ns = 0 if self.spheres is None else len(self.spheres)
        self.N = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        if self.N is not None or reverse_orientation:
            self.h = self.L.ref_scene_create2(_p(self.P), len(self.P), _p(self.idx), len(self.idx), _p(self.uv), _p(self.N),
                                              int(bool(reverse_orientation)), _p(self.spheres), ns)
        else:
            self.h = self.L.ref_scene_create(_p(self.P), len(self.P), _p(self.idx), len(self.idx), _p(self.uv),
                                            _p(self.spheres), ns)
        self.n_prims = self.L.ref_scene_num_prims(self.h)

    def prim_intersect(self, prim, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), REF_HIT)
        self.L.ref_prim_intersect(self.h, int(prim), _p(rays), len(rays), _p(hits))
        return hits

    def prim_intersect_p(self, prim, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint8)
        self.L.ref_prim_intersect_p(self.h, int(prim), _p(rays), len(rays), _p(occ))
        return occ

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_scene_destroy(self.h)
            self.h = None


def loopsubdiv(P, indices, n_levels, with_normals=False):
    P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
    idx = np.ascontiguousarray(indices, np.int32).reshape(-1)
    oP, oI = C.c_void_p(), C.c_void_p()
    nv, nt = C.c_int(), C.c_int()
    rc = lib().ref_loopsubdiv(_p(P), len(P), _p(idx), len(idx), n_levels, C.byref(oP), C.byref(nv), C.byref(oI),
                              C.byref(nt))
    if rc != 0:
        raise RuntimeError("ref_loopsubdiv failed")
    outP = np.ctypeslib.as_array(C.cast(oP, C.POINTER(C.c_float)), (nv.value, 3)).copy()
    outI = np.ctypeslib.as_array(C.cast(oI, C.POINTER(C.c_int)), (nt.value, 3)).copy()
    lib().ref_free(oP)
    lib().ref_free(oI)
    if with_normals:   # the limit-surface normals the reference's mesh carries (TriangleMesh::n)
        lib().ref_loopsubdiv_normals.argtypes = [C.c_void_p]
        N = np.zeros((nv.value, 3), np.float32)
        if lib().ref_loopsubdiv_normals(_p(N)) != nv.value:
            raise RuntimeError("the subdivided mesh has no normals")
        return outP, outI, N
    return outP, outI


class RefAccel:
    """The reference's own accelerator built by its Create*Accelerator from a ParamSet."""

    def __init__(self, scene, kind, int_params=None, float_params=None):
        self.scene = scene
        self.kind = kind
        ip = int_params or {}
        fp = float_params or {}
        inames = (C.c_char_p * (len(ip) + 1))(*[k.encode() for k in ip], None)
        ivals = (C.c_int * max(1, len(ip)))(*ip.values())
        fnames = (C.c_char_p * (len(fp) + 1))(*[k.encode() for k in fp], None)
        fvals = (C.c_float * max(1, len(fp)))(*fp.values())
        self.L = scene.L
        self.h = self.L.ref_accel_create(scene.h, kind.encode(), inames, ivals, fnames, fvals)
        if not self.h:
            raise ValueError("unknown accelerator kind %r" % kind)

    def export(self):
        ti = TreeInfo()
        self.L.ref_accel_info(self.h, C.byref(ti))
        nodes = np.zeros((ti.node_count, ti.node_stride // 4), np.uint32)
        idx = np.zeros(ti.n_indices, np.uint32)
        dirs = np.zeros((ti.n_directions, 3), np.float32)
        self.L.ref_accel_export(self.h, _p(nodes), _p(idx), _p(dirs))
        # the general-BSP builder families share two node formats (BSP.h:122-184 / BSPKd.h:11-174)
        kind = {"bsppaper": "bsp", "bspcluster": "bsp", "bsppaperkd": "bspkd", "bspclusterfastkd": "bspkd"}.get(self.kind, self.kind)
        return dict(kind=kind, nodes=nodes, indices=idx, directions=dirs,
                    bounds=np.array(list(ti.bounds), np.float32), n_prims=ti.n_prims)

    def intersect(self, rays, counters=False, n_threads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), REF_HIT)
        cnt = np.zeros(len(rays), REF_CNT) if counters else None
        self.L.ref_intersect(self.h, _p(rays), len(rays), _p(hits), _p(cnt), n_threads)
        return (hits, cnt) if counters else hits

    def intersect_p(self, rays, counters=False, n_threads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint8)
        cnt = np.zeros(len(rays), REF_CNT) if counters else None
        self.L.ref_intersect_p(self.h, _p(rays), len(rays), _p(occ), _p(cnt), n_threads)
        return (occ, cnt) if counters else occ

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_accel_destroy(self.h)
            self.h = None


def trace_render(scene, *args, **kw):
    """render(..., n_threads=1) with the event trace on: returns (render's result, records (n, 16) uint32) — every
    sampler call and every ray of the render in execution order (oracle/ref/ref_render.cpp, tests/hostwf.parse_trace)."""
    kw["n_threads"] = 1
    scene.L.ref_trace_enable(1)
    try:
        out = render(scene, *args, **kw)
        n = scene.L.ref_trace_records()
        rec = np.zeros((n, 16), np.uint32)
        scene.L.ref_trace_copy(rec.ctypes.data)
    finally:
        scene.L.ref_trace_enable(0)
    return out, rec


def camera_matrices(eye, look, up, fov, xres, yres, camera_matrix=None):
    """(RasterToCamera, CameraToWorld) 4x4 of the PerspectiveCamera ref_render builds."""
    cam = np.array(list(eye) + list(look) + list(up) + [fov], np.float32)
    r2c, c2w = np.zeros(16, np.float32), np.zeros(16, np.float32)
    m = None if camera_matrix is None else np.ascontiguousarray(camera_matrix, np.float32).reshape(16)
    lib().ref_camera_matrices(_p(cam), xres, yres, _p(m), _p(r2c), _p(c2w))
    return r2c.reshape(4, 4), c2w.reshape(4, 4)


def render(scene, accel_kind, eye, look, up, fov, xres, yres, spp, maxdepth, light_pos, light_I, kd,
           sampler="random", n_threads=0, tmp_prefix=None, heat=False, prim_mat=None, emitter_L=None, timing=False,
           distant=None, cropwindow=None, infinite=None, camera_matrix=None, textures=None, extra_lights=None,
           lightsamplestrategy="uniform", texture_filter=None):
    """The reference's OWN render (SamplerIntegrator::Render + PathIntegrator::Li + Film, oracle/ref/ref_render.cpp)
    of `scene` with a matte material and one point light.  Returns (rgb (yres,xres,3), n_closest, n_shadow)."""
    import tempfile
    cam = np.array(list(eye) + list(look) + list(up) + [fov], np.float32)
    lp, li, k = (np.array(a, np.float32) for a in (light_pos, light_I, kd))
    cw_, ch_ = xres, yres
    if cropwindow is not None:   # Film "float cropwindow" [x0 x1 y0 y1]: the outputs hold Film::croppedPixelBounds only
        x0, x1, y0, y1 = (float(np.float32(c)) for c in cropwindow)
        f32 = np.float32
        cw_ = int(np.ceil(f32(xres) * f32(x1))) - int(np.ceil(f32(xres) * f32(x0)))
        ch_ = int(np.ceil(f32(yres) * f32(y1))) - int(np.ceil(f32(yres) * f32(y0)))
    out = np.zeros((ch_, cw_, 3), np.float32)
    nc, ns, secs = C.c_uint64(), C.c_uint64(), C.c_double()
    hm = np.zeros((6, ch_, cw_), np.uint32) if heat else None
    pk = None if prim_mat is None else np.ascontiguousarray(prim_mat, np.float32)
    if pk is not None:   # 8-float records, or the device's 20-float records (adds Material "uber")
        assert pk.ndim == 2 and pk.shape[1] in (8, 20), pk.shape
        if hasattr(scene.L, "ref_render_set_mat_stride"):
            scene.L.ref_render_set_mat_stride(int(pk.shape[1]))
        elif pk.shape[1] != 8:
            raise RuntimeError("libpbrtref.so predates ref_render_set_mat_stride: rebuild oracle/ref")
    el = None if emitter_L is None else np.ascontiguousarray(emitter_L, np.float32)
    if hasattr(scene.L, "ref_render_set_distant"):   # distant=(from, to): LightSource "distant" with L = light_I
        scene.L.ref_render_set_distant.argtypes = [C.c_void_p]
        ft = None if distant is None else np.array(list(distant[0]) + list(distant[1]), np.float32)
        scene.L.ref_render_set_distant(_p(ft))
    elif distant is not None:
        raise RuntimeError("libpbrtref.so predates ref_render_set_distant: rebuild oracle/ref")
    # infinite=(L rgb, LightToWorld 4x4 row-major): LightSource "infinite" with a constant map; camera_matrix = the
    # world-to-camera CTM of a `Transform [...]` before `Camera` (eye/look/up are then ignored, fov still applies)
    if hasattr(scene.L, "ref_render_set_infinite"):
        scene.L.ref_render_set_infinite.argtypes = [C.c_void_p, C.c_void_p]
        scene.L.ref_render_set_camera_matrix.argtypes = [C.c_void_p]
        il = None if infinite is None else np.ascontiguousarray(infinite[0], np.float32)
        im = None if infinite is None else np.ascontiguousarray(infinite[1], np.float32).reshape(16)
        scene.L.ref_render_set_infinite(_p(il), _p(im))
        scene.L.ref_render_set_camera_matrix(_p(None if camera_matrix is None else np.ascontiguousarray(camera_matrix, np.float32).reshape(16)))
    elif infinite is not None or camera_matrix is not None:
        raise RuntimeError("libpbrtref.so predates ref_render_set_infinite: rebuild oracle/ref")
    if hasattr(scene.L, "ref_render_set_crop"):
        scene.L.ref_render_set_crop.argtypes = [C.c_void_p]
        scene.L.ref_render_set_crop(_p(None if cropwindow is None else np.array(cropwindow, np.float32)))
    elif cropwindow is not None:
        raise RuntimeError("libpbrtref.so predates ref_render_set_crop: rebuild oracle/ref")
    if extra_lights:
        # more LightSource statements after the render's own light (dicts as b200rt.wavefront.make_light takes them: point /
        # distant / infinite) and Integrator "path" "string lightsamplestrategy"
        if not hasattr(scene.L, "ref_render_set_extra_lights"):
            raise RuntimeError("libpbrtref.so predates ref_render_set_extra_lights: rebuild oracle/ref")
        scene.L.ref_render_set_extra_lights.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        scene.L.ref_render_set_extra_lights(_p(extra_light_records(extra_lights)), len(extra_lights), lightsamplestrategy.encode())
    try:
        with tempfile.TemporaryDirectory() as td:   # the thesis's per-pixel stat dumps (film.cpp:170-264) land here
            pre = (tmp_prefix or os.path.join(td, "ref")).encode()
            if textures:
                # textures: list of (h, w, 3) uint8 images (row 0 = top).  They become .tga files the reference reads with its own
                # ReadImage; slots 18 / 19 of 20-float prim_mat records name them (1-based) as the Kd / opacity `imagemap`
                files = []
                for i, px in enumerate(textures):
                    files.append(os.path.join(td, "tex%d.tga" % i))
                    write_tga(files[-1], px)
                if texture_filter is not None:   # ("ewa" | "trilinear", maxanisotropy)
                    set_texture_filter(texture_filter[0] == "trilinear", texture_filter[1], scene.L)
                set_textures(files, scene.L)
            rc = scene.L.ref_render(scene.h, accel_kind.encode(), _p(cam), xres, yres, spp, maxdepth, _p(lp), _p(li), _p(k),
                                    sampler.encode(), n_threads, pre, _p(out), C.byref(nc), C.byref(ns), _p(hm), _p(pk), _p(el), C.byref(secs))
    finally:
        # the "for the NEXT renders" settings are process globals of the library (tests/hostplugin's driver reads them too):
        # leave them at their defaults, so that a later render in the same process is not lit by this one's infinite light
        for fn, args in (("ref_render_set_distant", (None,)), ("ref_render_set_infinite", (None, None)),
                         ("ref_render_set_camera_matrix", (None,)), ("ref_render_set_crop", (None,)), ("ref_render_set_mat_stride", (8,)),
                         ("ref_render_set_textures", (0, None)), ("ref_render_set_extra_lights", (None, 0, None))):
            if hasattr(scene.L, fn):
                getattr(scene.L, fn)(*args)
        if hasattr(scene.L, "ref_render_set_texture_filter"):
            set_texture_filter(True, 8.0, scene.L)
    if rc != 0:
        raise RuntimeError("ref_render failed (%d)" % rc)
    if timing:   # (rgb, n_closest, n_shadow, seconds spent inside Integrator::Render)
        return out, nc.value, ns.value, secs.value
    return (out, nc.value, ns.value, hm) if heat else (out, nc.value, ns.value)


def extra_light_records(lights):
    """24-float records of ref_render_set_extra_lights from make_light-style dicts."""
    recs = np.zeros((len(lights), 24), np.float32)
    for r, l in zip(recs, lights):
        if l["kind"] == "point":
            r[0], r[1:4], r[4:7] = 0, l["pos"], l["I"]
        elif l["kind"] == "distant":
            r[0], r[1:4], r[4:7] = 2, l["w"], l["L"]
        elif l["kind"] == "infinite":
            m = np.eye(4, dtype=np.float32)
            if l.get("light_to_world") is not None:
                m[:3, :3] = np.asarray(l["light_to_world"], np.float32)[:3, :3]
            r[0], r[4:7], r[7:23] = 3, l["L"], m.ravel()
        else:
            raise ValueError("extra light kind %r" % (l["kind"],))
    return recs


def last_light_powers(L=None):
    """Light::Power().y() of the lights of the last render's scene, in Scene::lights order."""
    L = L or lib()
    out = np.zeros(64, np.float32)
    L.ref_render_last_light_powers.argtypes = [C.c_void_p, C.c_int]
    n = L.ref_render_last_light_powers(_p(out), 64)
    return out[:n].copy()


def texture_pyramid(path):
    """The MIPMap the reference builds for the image file `path` (CreateImageSpectrumTexture, trilinear): list of (h, w, 3) levels."""
    L = lib()
    L.ref_texture_pyramid.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    w, h = C.c_int(), C.c_int()
    n = L.ref_texture_pyramid(path.encode(), 0, None, C.byref(w), C.byref(h))
    out = []
    for l in range(n):
        L.ref_texture_pyramid(path.encode(), l, None, C.byref(w), C.byref(h))
        a = np.zeros((h.value, w.value, 3), np.float32)
        L.ref_texture_pyramid(path.encode(), l, _p(a), C.byref(w), C.byref(h))
        out.append(a)
    return out


def texture_lookup(path, stw):
    """MIPMap::Lookup(st, width) of the reference's texture for (n, 3) rows (s, t, width)."""
    L = lib()
    L.ref_texture_lookup.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p]
    stw = np.ascontiguousarray(stw, np.float32).reshape(-1, 3)
    out = np.zeros((len(stw), 3), np.float32)
    L.ref_texture_lookup(path.encode(), _p(stw), len(stw), _p(out))
    return out


def set_texture_filter(trilinear=True, maxanisotropy=8.0, L=None):
    """"bool trilinear" / "float maxanisotropy" of the textures the NEXT renders and lookups make (harness default: trilinear)."""
    L = L or lib()
    L.ref_render_set_texture_filter.argtypes = [C.c_int, C.c_float]
    L.ref_render_set_texture_filter(int(bool(trilinear)), float(maxanisotropy))


def texture_lookup_diff(path, q):
    """MIPMap::Lookup(st, dst0, dst1) of the reference's texture for (n, 6) rows (s, t, dst0.x, dst0.y, dst1.x, dst1.y), under
    the filter set_texture_filter chose."""
    L = lib()
    L.ref_texture_lookup_diff.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p]
    q = np.ascontiguousarray(q, np.float32).reshape(-1, 6)
    out = np.zeros((len(q), 3), np.float32)
    L.ref_texture_lookup_diff(path.encode(), _p(q), len(q), _p(out))
    return out


def set_textures(files, L=None):
    """Image files the slots 18 / 19 of 20-float prim_mat records refer to (1-based) in the NEXT renders; [] clears."""
    L = L or lib()
    arr = (C.c_char_p * max(1, len(files)))(*[f.encode() for f in files])
    L.ref_render_set_textures.argtypes = [C.c_int, C.c_void_p]
    L.ref_render_set_textures(len(files), arr)


def write_tga(path, rgb8):
    """Uncompressed 24-bit TGA with top-to-bottom rows (descriptor 0x20) and BGR pixels, like the files of scenes/living-room."""
    import struct
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    h, w = rgb8.shape[:2]
    with open(path, "wb") as f:
        f.write(struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, w, h, 24, 0x20))
        f.write(np.ascontiguousarray(rgb8[..., ::-1]).tobytes())
