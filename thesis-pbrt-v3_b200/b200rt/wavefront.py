"""Host mirror of the reference's integrator entry for the hot path: `Integrator::Render(scene)`
(src/core/integrator.h:57, SamplerIntegrator::Render src/core/integrator.cpp:230-360) with the
"path" integrator's parameters (src/integrators/path.cpp:206-240), over include/b200rt_wavefront.h."""
import ctypes as C
import numpy as np
from . import api


class Camera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("dir", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("xres", C.c_uint32), ("yres", C.c_uint32)]


class RenderDesc(C.Structure):
    _fields_ = [("camera", Camera), ("spp", C.c_uint32), ("max_depth", C.c_uint32), ("light_pos", C.c_float * 3),
                ("light_I", C.c_float * 3), ("kd", C.c_float * 3), ("tile_size", C.c_uint32), ("rank", C.c_uint32),
                ("world", C.c_uint32), ("seed", C.c_uint32), ("max_batch", C.c_uint32),
                ("light_kind", C.c_uint32), ("light_radius", C.c_float), ("light_prim", C.c_uint32),
                ("materials", C.c_void_p), ("n_materials", C.c_uint32), ("prim_material", C.c_void_p),
                ("crop", C.c_uint32 * 4), ("light_to_world", C.c_float * 9)]


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
        _bound = True
    return L


class WavefrontPathIntegrator:
    """Integrator "path" "integer maxdepth" [...] restructured as a wavefront (see module docstring).

    render(accel) -> film (yres, xres, 4) float32 = per-pixel (sum R, sum G, sum B, sum weight) for
    this rank's tiles (tile t belongs to rank t % world); stats as a dict."""

    def __init__(self, camera, spp=8, maxdepth=5, light_pos=(0, 0, 0), light_I=(1, 1, 1), kd=(.5, .5, .5),
                 rank=0, world=1, seed=0, tile_size=16, max_batch=0, sphere_light=None, materials=None,
                 prim_material=None, distant_light=None, crop=None, infinite_light=None):
        """sphere_light=(radius, prim): `AreaLightSource "area" "color L" light_I` on the full sphere primitive number
        `prim` of the aggregate, centred at light_pos (sampled with MIS like EstimateDirect).  materials: (n, 8) rows
        (kind, Kd.rgb, Ks.rgb, roughness) + prim_material: (n_prims,) uint8 index per primitive.
        distant_light=(from - to, world_radius): `LightSource "distant"` with radiance light_I (lights/distant.cpp);
        world_radius = the aggregate's bounding-sphere radius (see world_radius()).
        crop=(x0, x1, y0, y1): Film "cropwindow" in pixels (scenes.killeroo_feet()["crop"]); the film then holds
        (y1-y0, x1-x0) pixels, as the reference's Film does.
        infinite_light=(light_to_world 3x3 or None, world_radius): `LightSource "infinite"` with the constant radiance
        light_I (lights/infinite.cpp without a readable "mapname"), sampled with MIS."""
        d = RenderDesc()
        for k in ("eye", "dir", "right", "up"):
            setattr(d.camera, k, (C.c_float * 3)(*[float(x) for x in camera[k]]))
        d.camera.xres, d.camera.yres = int(camera["xres"]), int(camera["yres"])
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
            w = np.asarray(distant_light[0], np.float32)
            w = w / np.sqrt((w * w).sum(dtype=np.float32))
            d.light_kind, d.light_radius = 2, float(distant_light[1])
            d.light_pos = (C.c_float * 3)(*[float(x) for x in w])
        if infinite_light is not None:
            m = np.eye(3) if infinite_light[0] is None else np.asarray(infinite_light[0], np.float64)[:3, :3]
            d.light_kind, d.light_radius = 3, float(infinite_light[1])
            d.light_to_world = (C.c_float * 9)(*[float(x) for x in m.ravel()])
        if materials is not None:
            self._mats = np.ascontiguousarray(materials, np.float32).reshape(-1, 8)
            self._pm = np.ascontiguousarray(prim_material, np.uint8)
            d.materials, d.n_materials = self._mats.ctypes.data, len(self._mats)
            d.prim_material = self._pm.ctypes.data
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
    c = (lo + hi) / np.float32(2)
    return float(np.sqrt(((hi - c) ** 2).sum(dtype=np.float32)))


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


def film_to_rgb(film):
    """Film::WriteImage normalisation (src/core/film.cpp:266-305): rgb = sum / weight."""
    w = np.maximum(film[..., 3:4], 1e-20)
    return np.where(film[..., 3:4] > 0, film[..., :3] / w, 0.0).astype(np.float32)
