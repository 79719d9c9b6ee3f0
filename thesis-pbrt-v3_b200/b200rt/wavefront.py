"""Host mirror of the reference's integrator entry for the hot path: `Integrator::Render(scene)`
(src/core/integrator.h:57, SamplerIntegrator::Render src/core/integrator.cpp:230-360) with the
"path" integrator's parameters (src/integrators/path.cpp:206-240), over include/b200rt_wavefront.h."""
import ctypes as C
import numpy as np
from . import api


class Camera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("dir", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("xres", C.c_uint32), ("yres", C.c_uint32)]


class Light(C.Structure):
    """b200rt_light (include/b200rt_wavefront.h, ABI v10)."""
    _fields_ = [("kind", C.c_uint32), ("prim", C.c_uint32), ("pos", C.c_float * 3), ("I", C.c_float * 3), ("radius", C.c_float),
                ("to_world", C.c_float * 9), ("weight", C.c_float)]


def _unit_to_light(w):
    """wLight = Normalize(LightToWorld(from - to)) (lights/distant.cpp:44-47): v * (1 / Length), in float."""
    w = np.asarray(w, np.float32)
    inv = np.float32(1) / np.sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2], dtype=np.float32)
    return (w * inv).astype(np.float32)


def make_light(spec):
    """One entry of `lights=`: a dict with "kind" and that kind's parameters, "weight" (default 1) its entry in the light
    distribution:
      point    : pos, I                                   LightSource "point" "from" "I"
      sphere   : center, L, radius, prim                  AreaLightSource "area" "L" on the full Shape "sphere" number `prim`
      distant  : w (= from - to), L, world_radius         LightSource "distant"
      infinite : L, light_to_world (3x3 or None), world_radius   LightSource "infinite" without a readable map"""
    l = Light()
    kind = spec["kind"]
    l.weight = float(spec.get("weight", 1.0))
    f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v])
    if kind == "point":
        l.kind, l.pos, l.I = 0, f3(spec["pos"]), f3(spec["I"])
    elif kind == "sphere":
        l.kind, l.pos, l.I, l.radius, l.prim = 1, f3(spec["center"]), f3(spec["L"]), float(spec["radius"]), int(spec["prim"])
    elif kind == "distant":
        l.kind, l.pos, l.I, l.radius = 2, f3(_unit_to_light(spec["w"])), f3(spec["L"]), float(spec["world_radius"])
    elif kind == "infinite":
        m = np.eye(3) if spec.get("light_to_world") is None else np.asarray(spec["light_to_world"], np.float64)[:3, :3]
        l.kind, l.I, l.radius = 3, f3(spec["L"]), float(spec["world_radius"])
        l.to_world = (C.c_float * 9)(*[float(x) for x in m.ravel()])
    else:
        raise ValueError("unknown light kind %r" % (kind,))
    return l


class RenderDesc(C.Structure):
    _fields_ = [("camera", Camera), ("spp", C.c_uint32), ("max_depth", C.c_uint32), ("light_pos", C.c_float * 3),
                ("light_I", C.c_float * 3), ("kd", C.c_float * 3), ("tile_size", C.c_uint32), ("rank", C.c_uint32),
                ("world", C.c_uint32), ("seed", C.c_uint32), ("max_batch", C.c_uint32),
                ("light_kind", C.c_uint32), ("light_radius", C.c_float), ("light_prim", C.c_uint32),
                ("materials", C.c_void_p), ("n_materials", C.c_uint32), ("prim_material", C.c_void_p),
                ("crop", C.c_uint32 * 4), ("light_to_world", C.c_float * 9),
                ("sampler", C.c_uint32), ("camera_exact", C.c_uint32), ("raster_to_camera", C.c_float * 16),
                ("camera_to_world", C.c_float * 16), ("sobol_matrices32", C.c_void_p), ("sobol_vdc", C.c_void_p),
                ("sobol_vdc_inv", C.c_void_p), ("sobol_vdc_rows", C.c_uint32), ("material_stride", C.c_uint32),
                ("textures", C.c_void_p), ("n_textures", C.c_uint32), ("lights", C.c_void_p), ("n_lights", C.c_uint32),
                ("light_strategy", C.c_uint32)]


SAMPLERS = {"hash": 0, "halton": 1, "sobol": 2}


class RenderStats(C.Structure):
    _fields_ = [("camera_rays", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("launches", C.c_uint64), ("ms_closest", C.c_float), ("ms_any", C.c_float), ("ms_total", C.c_float),
                ("n_tiles", C.c_uint32), ("launches_closest", C.c_uint32), ("launches_any", C.c_uint32),
                ("mis_rays", C.c_uint64)]


_bound = False


def _lib():
    global _bound
    L = api.lib()
    if not _bound:
        L.b200rt_render.argtypes = [C.c_void_p, C.POINTER(RenderDesc), C.c_void_p, C.POINTER(RenderStats), C.c_void_p]
        L.b200rt_render_heatmaps.argtypes = [C.c_void_p, C.POINTER(RenderDesc), C.c_void_p, C.c_void_p, C.POINTER(RenderStats), C.c_void_p]
        L.b200rt_render_tiles.argtypes = [C.POINTER(RenderDesc), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
        L.b200rt_render_keep_queues.argtypes = [C.c_void_p, C.c_int]
        L.b200rt_render_queue.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]
        L.b200rt_render_queue_paths.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        L.b200rt_camera_perspective.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_uint32, C.c_uint32, C.POINTER(RenderDesc)]
        _bound = True
    return L


class WavefrontPathIntegrator:
    """Integrator "path" "integer maxdepth" [...] restructured as a wavefront (see module docstring).

    render(accel) -> film (yres, xres, 4) float32 = per-pixel (sum R, sum G, sum B, sum weight) for
    this rank's tiles (tile t belongs to rank t % world); stats as a dict."""

    def __init__(self, camera, spp=8, maxdepth=5, light_pos=(0, 0, 0), light_I=(1, 1, 1), kd=(.5, .5, .5),
                 rank=0, world=1, seed=0, tile_size=16, max_batch=0, sphere_light=None, materials=None,
                 prim_material=None, distant_light=None, crop=None, infinite_light=None, sampler="halton", sobol_tables=None,
                 textures=None, lights=None, lightsamplestrategy="uniform"):
        """sphere_light=(radius, prim): `AreaLightSource "area" "color L" light_I` on the full sphere primitive number
        `prim` of the aggregate, centred at light_pos (sampled with MIS like EstimateDirect).  materials: (n, 8) rows
        (kind, Kd.rgb, Ks.rgb, roughness) + prim_material: (n_prims,) uint8 index per primitive.
        distant_light=(from - to, world_radius): `LightSource "distant"` with radiance light_I (lights/distant.cpp);
        world_radius = the aggregate's bounding-sphere radius (see world_radius()).
        crop=(x0, x1, y0, y1): Film "cropwindow" in pixels (scenes.killeroo_feet()["crop"]); the film then holds
        (y1-y0, x1-x0) pixels, as the reference's Film does.
        infinite_light=(light_to_world 3x3 or None, world_radius): `LightSource "infinite"` with the constant radiance
        light_I (lights/infinite.cpp without a readable "mapname"), sampled with MIS.
        camera: either the pinhole frame dict (eye / dir / right / up / xres / yres) or a dict with "lookat" (9 floats:
        eye, look, up) or "w2c" (4x4 world-to-camera matrix) + "fov", "xres", "yres": the latter is turned into the
        reference camera's own matrices (b200rt_camera_perspective) and camera rays are bit-identical to the reference's.
        sampler: "hash" | "halton" | "sobol" (Sampler "halton" / "sobol" of the reference, same sample values and
        dimension order); sobol needs sobol_tables = (SobolMatrices32, VdCSobolMatrices, VdCSobolMatricesInv) arrays.
        materials: (n, 8) rows, or (n, 20) rows (needed for "uber", kind 6).
        textures: list of b200rt.textures.ImageTexture (`Texture ... "imagemap"`, trilinear); slot 18 of a 20-float material row =
        1 + index of the texture its Kd reads, slot 19 = 1 + index of uber's opacity texture.
        lights: a list of make_light() dicts = Scene::lights in the scene's order, REPLACING light_pos / light_I / sphere_light /
        distant_light / infinite_light; one of them is chosen per path vertex as UniformSampleOneLight does (integrator.cpp:86-107)
        from the distribution over their "weight"s: lightsamplestrategy "uniform" (all 1, the default here) or "power" (the
        caller's weights = Light::Power().y()) — or from the voxel's own distribution under "spatial", the reference's default
        for scenes with several lights (SpatialLightDistribution)."""
        d = RenderDesc()
        if "lookat" in camera or "w2c" in camera:
            la = None if camera.get("lookat") is None else np.ascontiguousarray(camera["lookat"], np.float32).reshape(9)
            m = None if camera.get("w2c") is None else np.ascontiguousarray(camera["w2c"], np.float32).reshape(16)
            api._check(_lib().b200rt_camera_perspective(None if la is None else la.ctypes.data, None if m is None else m.ctypes.data,
                                                        float(camera["fov"]), int(camera["xres"]), int(camera["yres"]), C.byref(d)))
        else:
            for k in ("eye", "dir", "right", "up"):
                setattr(d.camera, k, (C.c_float * 3)(*[float(x) for x in camera[k]]))
            d.camera.xres, d.camera.yres = int(camera["xres"]), int(camera["yres"])
        d.sampler = SAMPLERS[sampler]
        if sampler == "sobol":
            if sobol_tables is None:
                sobol_tables = load_sobol_tables()
            self._sobol = [np.ascontiguousarray(sobol_tables[0], np.uint32).reshape(-1),
                           np.ascontiguousarray(sobol_tables[1], np.uint64).reshape(-1, 52),
                           np.ascontiguousarray(sobol_tables[2], np.uint64).reshape(-1, 52)]
            d.sobol_matrices32 = self._sobol[0].ctypes.data
            d.sobol_vdc, d.sobol_vdc_inv = self._sobol[1].ctypes.data, self._sobol[2].ctypes.data
            d.sobol_vdc_rows = len(self._sobol[1])
        d.spp, d.max_depth = int(spp), int(maxdepth)
        d.light_pos = (C.c_float * 3)(*[float(x) for x in light_pos])
        d.light_I = (C.c_float * 3)(*[float(x) for x in light_I])
        d.kd = (C.c_float * 3)(*[float(x) for x in kd])
        d.tile_size, d.rank, d.world, d.seed, d.max_batch = tile_size, rank, world, seed, max_batch
        if sphere_light is not None:
            d.light_kind, d.light_radius, d.light_prim = 1, float(sphere_light[0]), int(sphere_light[1])
        if crop is not None:
            d.crop = (C.c_uint32 * 4)(*[int(c) for c in crop])
        if distant_light is not None:
            # wLight = Normalize(LightToWorld(from - to)) (lights/distant.cpp:44-47): v * (1 / Length), in float
            w = _unit_to_light(distant_light[0])
            d.light_kind, d.light_radius = 2, float(distant_light[1])
            d.light_pos = (C.c_float * 3)(*[float(x) for x in w])
        if infinite_light is not None:
            m = np.eye(3) if infinite_light[0] is None else np.asarray(infinite_light[0], np.float64)[:3, :3]
            d.light_kind, d.light_radius = 3, float(infinite_light[1])
            d.light_to_world = (C.c_float * 9)(*[float(x) for x in m.ravel()])
        if materials is not None:
            m = np.asarray(materials, np.float32)
            stride = 20 if m.ndim == 2 and m.shape[1] == 20 else 8
            self._mats = np.ascontiguousarray(m, np.float32).reshape(-1, stride)
            self._pm = np.ascontiguousarray(prim_material, np.uint8)
            d.materials, d.n_materials, d.material_stride = self._mats.ctypes.data, len(self._mats), stride
            d.prim_material = self._pm.ctypes.data
        if textures:
            from . import textures as tx
            self._textures = list(textures)
            self._tex_array = tx.texture_array(self._textures)
            d.textures, d.n_textures = C.cast(self._tex_array, C.c_void_p), len(self._textures)
        if lights:
            if lightsamplestrategy not in ("uniform", "power", "spatial"):
                raise ValueError("lightsamplestrategy %r: \"uniform\", \"power\" or \"spatial\"" % (lightsamplestrategy,))
            d.light_strategy = 1 if lightsamplestrategy == "spatial" and len(lights) > 1 else 0
            recs = [make_light(dict(l, weight=1.0) if lightsamplestrategy == "uniform" else l) for l in lights]
            self._lights = (Light * len(recs))(*recs)
            d.lights, d.n_lights = C.cast(self._lights, C.c_void_p), len(recs)
        self.desc = d

    @property
    def shape(self):
        c = self.desc.crop
        if c[0] or c[1] or c[2] or c[3]:
            return (c[3] - c[2], c[1] - c[0], 4)
        return (self.desc.camera.yres, self.desc.camera.xres, 4)

    def tiles(self):
        n, px = C.c_uint32(), C.c_uint64()
        api._check(_lib().b200rt_render_tiles(C.byref(self.desc), C.byref(n), C.byref(px)))
        return n.value, px.value

    def render_ptr(self, accel, film_ptr, stream=None):
        """film_ptr: device (or host) pointer to xres*yres*4 floats, accumulated into."""
        st = RenderStats()
        api._check(_lib().b200rt_render(accel._h, C.byref(self.desc), film_ptr, C.byref(st), stream))
        return {f: getattr(st, f) for f, _ in RenderStats._fields_}

    def Render(self, accel, stream=None):
        film = np.zeros(self.shape, np.float32)
        stats = self.render_ptr(accel, film.ctypes.data_as(C.c_void_p), stream)
        return film, stats


def world_radius(tri_verts):
    """Bounds3::BoundingSphere of the aggregate's world bound (geometry.h), as DistantLight::Preprocess uses it."""
    v = np.asarray(tri_verts, np.float32).reshape(-1, 3)
    lo, hi = v.min(0), v.max(0)
    c = (lo + hi) * (np.float32(1) / np.float32(2))   # Point3::operator/(2): multiplication by the reciprocal
    d = (hi - c).astype(np.float32)
    return float(np.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], dtype=np.float32))


HEAT_PLANES = ("primitiveIntersections", "primitiveIntersectionsP", "treeNodeTraversals", "treeNodeTraversalsP",
               "leafNodeTraversals", "leafNodeTraversalsP")


def render_heatmaps(integ, accel, stream=None):
    """Render + the thesis's per-pixel traversal statistics (Film::WriteGeneralStats, reference
    src/core/film.cpp:170-264): returns (film, heat (6, yres, xres) uint32 in HEAT_PLANES order, stats)."""
    film = np.zeros(integ.shape, np.float32)
    heat = np.zeros((len(HEAT_PLANES), integ.shape[0], integ.shape[1]), np.uint32)
    st = RenderStats()
    api._check(_lib().b200rt_render_heatmaps(accel._h, C.byref(integ.desc), film.ctypes.data_as(C.c_void_p),
                                             heat.ctypes.data_as(C.c_void_p), C.byref(st), stream))
    return film, heat, {f: getattr(st, f) for f, _ in RenderStats._fields_}


def set_streams(n):
    """Wavefront batches in flight per render call (b200rt_render_set_streams; default 2).  Use 1 to time kernels alone."""
    api._check(_lib().b200rt_render_set_streams(int(n)))


def keep_queues(accel, keep=True):
    api._check(_lib().b200rt_render_keep_queues(accel._h, int(keep)))


def queue_count(accel, q):
    n = C.c_uint64()
    api._check(_lib().b200rt_render_queue(accel._h, q, None, C.byref(n)))
    return n.value


def queue_rays(accel, q, ptr=None):
    """Rays of queue q (2*depth = closest, 2*depth+1 = shadow) of the last rendered batch.  With
    ptr=None returns a host (n,8) array; else copies into the given device pointer."""
    n = C.c_uint64(queue_count(accel, q))
    if ptr is not None:
        api._check(_lib().b200rt_render_queue(accel._h, q, ptr, C.byref(n)))
        return n.value
    out = np.zeros((n.value, 8), np.float32)
    if n.value:
        api._check(_lib().b200rt_render_queue(accel._h, q, out.ctypes.data_as(C.c_void_p), C.byref(n)))
    return out


def load_sobol_tables():
    """(SobolMatrices32, VdCSobolMatrices, VdCSobolMatricesInv) of the reference's Sampler "sobol" (data/sobol_tables.npz,
    exported from the compiled reference by tests/golden/make_sobol_tables.py)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "sobol_tables.npz"))
    return z["matrices32"], z["vdc"], z["vdc_inv"]


def queue_paths(accel, q, contrib=False):
    """Path slot of every entry of kept queue q (2*depth closest, 2*depth+1 shadow, 2*(maxdepth+1)+depth MIS); with
    contrib=True also the (n,4) radiance each shadow / MIS ray carries."""
    n = C.c_uint64(queue_count(accel, q))
    path = np.zeros(n.value, np.uint32)
    c = np.zeros((n.value, 4), np.float32) if contrib else None
    if n.value:
        api._check(_lib().b200rt_render_queue_paths(accel._h, q, path.ctypes.data_as(C.c_void_p),
                                                    None if c is None else c.ctypes.data_as(C.c_void_p), C.byref(n)))
    return (path, c) if contrib else path


def slot_pixels(desc, n_slots, start=0):
    """(x, y, s, valid) of path slots start .. start+n_slots-1 of a render: tiles of the crop window in row-major order
    (those with index % world == rank), pixels row-major inside a tile, samples innermost (wf_shade.cuh path_pixel)."""
    c = desc.crop
    xres, yres = (c[1] - c[0], c[3] - c[2]) if (c[0] or c[1] or c[2] or c[3]) else (desc.camera.xres, desc.camera.yres)
    tile = desc.tile_size or 16
    ntx, nty = (xres + tile - 1) // tile, (yres + tile - 1) // tile
    g = np.arange(start, start + n_slots, dtype=np.int64)
    pix, s = g // desc.spp, g % desc.spp
    tile_ord, in_tile = pix // (tile * tile), pix % (tile * tile)
    idx = desc.rank + tile_ord * desc.world
    tx, ty = idx % ntx, idx // ntx
    x, y = tx * tile + in_tile % tile, ty * tile + in_tile // tile
    return x, y, s, (ty < nty) & (x < xres) & (y < yres)


def film_to_rgb(film):
    """Film::WriteImage normalisation (src/core/film.cpp:266-305): rgb = sum / weight."""
    w = np.maximum(film[..., 3:4], 1e-20)
    return np.where(film[..., 3:4] > 0, film[..., :3] / w, 0.0).astype(np.float32)
