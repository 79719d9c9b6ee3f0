"""TEST INFRASTRUCTURE ONLY.  Drives the host-compiled shade stage (tests/hostshade/libhostshade.so: the product's
wf_shade.cuh / wf_bsdf.cuh / wf_sampler.cuh compiled with g++) through the same wavefront loop the device driver runs,
with the CPU oracle answering the rays, and parses the event trace of a reference render (oracle/ref/ref_render.cpp) into
per-depth ray queues — so that sample values, camera rays and every spawned ray of the restatement can be compared with
the compiled reference in the CPU test suite."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HS_DIR = os.path.join(ROOT, "tests", "hostshade")
MISS = 0xFFFFFFFF
_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", HS_DIR])
        L = C.CDLL(os.path.join(HS_DIR, "libhostshade.so"))
        vp = C.c_void_p
        L.hs_create.restype = vp
        L.hs_create.argtypes = [vp, C.c_uint32, vp, vp, vp, vp, vp, vp]
        L.hs_destroy.argtypes = [vp]
        L.hs_last_error.restype = C.c_char_p
        L.hs_n_paths.restype = C.c_uint64
        L.hs_n_paths.argtypes = [vp]
        L.hs_sample_values.argtypes = [vp, C.c_uint64, vp, vp, vp, vp, vp]
        L.hs_camera.restype = C.c_uint32
        L.hs_camera.argtypes = [vp, C.c_uint64, C.c_uint32, vp, vp, vp, vp, vp]
        L.hs_shade.argtypes = [vp, C.c_uint32, C.c_uint64, C.c_uint32] + [vp] * 15
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def desc_has_area_light(d):
    """Some light of the description has area (sphere / infinite): EstimateDirect's BSDF-sampled rays exist."""
    if d.n_lights:
        from b200rt import wavefront as wf
        recs = C.cast(d.lights, C.POINTER(wf.Light))
        return any(recs[k].kind in (1, 3) for k in range(d.n_lights))
    return d.light_kind in (1, 3)


class HostWavefront:
    """The wavefront loop of csrc/wavefront.cu on the host: hs_camera / hs_shade are the kernels' bodies, `orc` (an
    oracle.orcbind.OracleAccel) traces the queues.  render() returns (film (yres, xres, 4), queues) with queues[q] =
    dict(rays (n, 8), path (n,), contrib (n, 4) for shadow / MIS queues) keyed like b200rt_render_queue."""

    def __init__(self, desc, tri_verts, orc, tri_uvs=None, tri_normals=None, prim_flags=None, prim_is_sphere=None):
        self.desc, self.orc = desc, orc
        self.tv = np.ascontiguousarray(tri_verts, np.float32).reshape(-1, 9)
        self.uv = None if tri_uvs is None else np.ascontiguousarray(tri_uvs, np.float32).reshape(-1, 6)
        self.tn = None if tri_normals is None else np.ascontiguousarray(tri_normals, np.float32).reshape(-1, 9)
        if self.tn is not None and prim_flags is None:
            prim_flags = np.full(len(self.tv), 2, np.uint8)
        self.pf = None if prim_flags is None else np.ascontiguousarray(prim_flags, np.uint8)
        self.sph = None if prim_is_sphere is None else np.ascontiguousarray(prim_is_sphere, np.uint8)
        self.bounds = np.ascontiguousarray(orc.bounds, np.float32)   # Scene::WorldBound() = the aggregate's
        self.h = lib().hs_create(C.byref(desc), len(self.tv), _p(self.tv), _p(self.uv), _p(self.tn), _p(self.pf), _p(self.sph), _p(self.bounds))
        if not self.h:
            raise RuntimeError(lib().hs_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            lib().hs_destroy(self.h)
            self.h = None

    def sample_values(self, px, py, s, dim):
        px, py = np.ascontiguousarray(px, np.int32), np.ascontiguousarray(py, np.int32)
        s, dim = np.ascontiguousarray(s, np.uint32), np.ascontiguousarray(dim, np.uint32)
        out = np.zeros(len(px), np.float32)
        lib().hs_sample_values(self.h, len(px), _p(px), _p(py), _p(s), _p(dim), _p(out))
        return out

    def _trace(self, rays):
        h = self.orc.intersect(rays, n_threads=8)
        rec = np.zeros((len(rays), 5), np.uint32)
        rec[:, 0] = h["prim"]
        rec[:, 1] = h["t"].view(np.uint32)
        for k, f in enumerate(("b0", "b1", "b2")):
            rec[:, 2 + k] = h[f].view(np.uint32)
        return rec

    def render(self):
        L = lib()
        d = self.desc
        n = int(L.hs_n_paths(self.h))
        mis = desc_has_area_light(d)
        maxd = d.max_depth
        rays = np.zeros((n, 8), np.float32)
        qpath = np.zeros(n, np.uint32)
        beta, Lp, pdim = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32), np.zeros(n, np.uint32)
        cnt = L.hs_camera(self.h, 0, n, _p(rays), _p(qpath), _p(beta), _p(Lp), _p(pdim))
        queues = {}
        stats = dict(closest_rays=0, shadow_rays=0, mis_rays=0, camera_rays=int(cnt))
        for depth in range(maxd + 1):
            cur_r, cur_p = rays[:cnt].copy(), qpath[:cnt].copy()
            queues[2 * depth] = dict(rays=cur_r, path=cur_p)
            stats["closest_rays"] += cnt
            hits = self._trace(cur_r) if cnt else np.zeros((0, 5), np.uint32)
            queues[2 * depth]["hits"] = hits
            if depth == maxd and not (mis and self._has_specular()):
                break
            nr, npth = np.zeros((max(cnt, 1), 8), np.float32), np.zeros(max(cnt, 1), np.uint32)
            sr, sp, sc = np.zeros((max(cnt, 1), 8), np.float32), np.zeros(max(cnt, 1), np.uint32), np.zeros((max(cnt, 1), 4), np.float32)
            mr, mp, mc = np.zeros((max(cnt, 1), 8), np.float32), np.zeros(max(cnt, 1), np.uint32), np.zeros((max(cnt, 1), 4), np.float32)
            nout = np.zeros(3, np.uint32)
            L.hs_shade(self.h, depth, 0, cnt, _p(cur_r), _p(cur_p), _p(hits), _p(beta), _p(pdim), _p(Lp), _p(nr), _p(npth),
                       _p(sr), _p(sp), _p(sc), _p(mr), _p(mp), _p(mc), _p(nout))
            if depth == maxd:
                break
            nn, ns, nm = (int(x) for x in nout)
            queues[2 * depth + 1] = dict(rays=sr[:ns].copy(), path=sp[:ns].copy(), contrib=sc[:ns].copy())
            stats["shadow_rays"] += ns
            if ns:
                occ = self.orc.intersect_p(sr[:ns], n_threads=8)
                free = occ == 0
                queues[2 * depth + 1]["occluded"] = occ
                np.add.at(Lp, (sp[:ns][free],), sc[:ns][free])   # one shadow ray per path per depth
            if mis:
                q = 2 * (maxd + 1) + depth
                queues[q] = dict(rays=mr[:nm].copy(), path=mp[:nm].copy(), contrib=mc[:nm].copy())
                stats["closest_rays"] += nm
                stats["mis_rays"] += nm
                if nm:
                    mh = self._trace(mr[:nm])
                    want = mc[:nm, 3].copy().view(np.uint32)   # the chosen light's primitive, or MISS for an infinite light
                    ok = mh[:, 0] == want
                    add = mc[:nm].copy()
                    add[:, 3] = 0
                    np.add.at(Lp, (mp[:nm][ok],), add[ok])
            rays[:nn], qpath[:nn], cnt = nr[:nn], npth[:nn], nn
        # k_film: per pixel, samples summed in order
        from b200rt import wavefront as wf
        x, y, s, valid = wf.slot_pixels(d, n)
        c = d.crop
        xres, yres = (c[1] - c[0], c[3] - c[2]) if (c[0] or c[1] or c[2] or c[3]) else (d.camera.xres, d.camera.yres)
        film = np.zeros((yres, xres, 4), np.float32)
        # FilmTile::AddSample with the box filter (film.h:130-169), as k_film / k_splat do it: own pixel in sample order,
        # then the contributions of boundary samples to the neighbouring pixels (clipped to the tile's one-pixel halo)
        x0, y0 = (c[0], c[2]) if (c[0] or c[1] or c[2] or c[3]) else (0, 0)
        f32 = np.float32
        u0 = self.sample_values(x + x0, y + y0, s, np.zeros(n, np.uint32))
        u1 = self.sample_values(x + x0, y + y0, s, np.ones(n, np.uint32))
        fx = ((x + x0).astype(f32) + u0) - f32(0.5)
        fy = ((y + y0).astype(f32) + u1) - f32(0.5)
        tile = d.tile_size or 16
        tx0, ty0 = (x // tile) * tile, (y // tile) * tile
        bx0, by0 = np.maximum(tx0 - 1, 0), np.maximum(ty0 - 1, 0)
        bx1 = np.minimum(np.minimum(tx0 + tile, xres) + 1, xres)
        by1 = np.minimum(np.minimum(ty0 + tile, yres) + 1, yres)
        p0x = np.maximum(np.ceil(fx - f32(0.5)).astype(np.int64) - x0, bx0)
        p1x = np.minimum(np.floor(fx + f32(0.5)).astype(np.int64) + 1 - x0, bx1)
        p0y = np.maximum(np.ceil(fy - f32(0.5)).astype(np.int64) - y0, by0)
        p1y = np.minimum(np.floor(fy + f32(0.5)).astype(np.int64) + 1 - y0, by1)
        for k in range(d.spp):
            m = valid & (s == k) & (p0x <= x) & (x < p1x) & (p0y <= y) & (y < p1y)
            film[y[m], x[m]] += Lp[m]
        nb = np.zeros((8, yres, xres, 4), np.float32)   # per source pixel and direction, summed in sample order
        for k in range(d.spp):
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    if dx == 0 and dy == 0:
                        continue
                    q = (dy + 1) * 3 + (dx + 1)
                    q = q if q < 4 else q - 1
                    m = valid & (s == k) & (p0x <= x + dx) & (x + dx < p1x) & (p0y <= y + dy) & (y + dy < p1y)
                    nb[q, y[m], x[m]] += Lp[m]
        self.boundary_samples = 0
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                if dx == 0 and dy == 0:
                    continue
                q = (dy + 1) * 3 + (dx + 1)
                q = q if q < 4 else q - 1
                ys, xs = np.nonzero(nb[q, ..., 3])
                film[ys + dy, xs + dx] += nb[q, ys, xs]
                self.boundary_samples += len(ys)
        self.L_paths = Lp
        return film, queues, stats

    def _has_specular(self):
        if not self.desc.materials:
            return False
        stride = self.desc.material_stride or 8
        m = np.ctypeslib.as_array(C.cast(self.desc.materials, C.POINTER(C.c_float)), (self.desc.n_materials, stride))
        k = m[:, 0].astype(np.uint32) & 15
        return bool(((k == 2) | (k == 3) | (k == 6)).any())


def parse_trace(rec):
    """Event trace of oracle/ref/ref_render.cpp -> list of samples, each a dict:
    pixel (x, y), sample index, events = list of (type, payload) in order, with type in
    {"1d", "2d", "closest", "any"}."""
    rec = np.asarray(rec, np.uint32).reshape(-1, 16)
    f = rec.view(np.float32)
    samples, cur, pix, sidx = [], None, None, 0
    for i in range(len(rec)):
        t = rec[i, 0]
        if t == 1:
            pix, sidx = (int(rec[i, 1]), int(rec[i, 2])), 0
            cur = dict(pixel=pix, s=0, events=[])
            samples.append(cur)
        elif t == 2:
            sidx += 1
            if rec[i, 1]:
                cur = dict(pixel=pix, s=sidx, events=[])
                samples.append(cur)
        elif t == 3:
            cur["events"].append(("1d", int(rec[i, 1]), f[i, 2:3].copy()))
        elif t == 4:
            cur["events"].append(("2d", int(rec[i, 1]), f[i, 2:4].copy()))
        elif t == 5:
            cur["events"].append(("closest", f[i, 1:8].copy(), int(rec[i, 8].view(np.int32)), f[i, 9], f[i, 10:16].copy()))
        elif t == 6:
            cur["events"].append(("any", f[i, 1:8].copy(), int(rec[i, 8])))
    return samples


def trace_queues(samples, maxdepth, mis):
    """Reorders a parsed reference trace into the wavefront's queues: {q: {(x, y, s): ray (7 floats o, d, tMax)}} with q as
    in b200rt_render_queue (2*depth closest path rays, 2*depth+1 shadow rays, 2*(maxdepth+1)+depth MIS rays).  Uses the
    order of calls inside PathIntegrator::Li: per bounce closest (path ray), then inside EstimateDirect the shadow ray and —
    for lights with area — the BSDF-sampled closest-hit ray."""
    out = {}
    for smp in samples:
        key = (smp["pixel"][0], smp["pixel"][1], smp["s"])
        depth = -1
        ev = smp["events"]
        for k, e in enumerate(ev):
            if e[0] == "closest":
                # a closest-hit call is the MIS ray when it follows this bounce's uLight / uScattering draws directly
                if mis and _after_uscattering(ev, k):
                    out.setdefault(2 * (maxdepth + 1) + depth, {})[key] = e
                else:
                    depth += 1
                    out.setdefault(2 * depth, {})[key] = e
            elif e[0] == "any":
                out.setdefault(2 * depth + 1, {})[key] = e
    return out


def _after_uscattering(ev, k):
    """True when event k (a closest call) follows the two 2D draws of UniformSampleOneLight with no 2D draw (the path's
    BSDF sample) in between, i.e. it is EstimateDirect's BSDF-sampled ray."""
    n2d = 0
    j = k - 1
    while j >= 0 and ev[j][0] != "1d":
        if ev[j][0] == "2d":
            n2d += 1
        if ev[j][0] == "closest":
            return False
        j -= 1
    return j >= 0 and n2d == 2


def reference_trace(case, spp, res, sampler="halton"):
    """Runs the reference's own renderer on shade_cases.build(case) single-threaded with the event trace on.
    Returns (built case, rgb image, (n_closest, n_shadow), parsed samples)."""
    import shade_cases
    from oracle import refbind as rb
    c = shade_cases.build(case, res)
    S = rb.RefScene(**c["ref"])
    r = dict(c["render"])
    rres = c.get("res", res)
    (img, nc, ns), rec = rb.trace_render(S, "bvh", r.pop("eye"), r.pop("look"), r.pop("up"), r.pop("fov"), rres, rres, spp,
                                         c["maxdepth"], r.pop("light_pos"), r.pop("light_I"), r.pop("kd"), sampler=sampler, **r)
    if c.get("power"):   # "lightsamplestrategy" "power": the weights are the reference's own Light::Power().y()
        for l, w in zip(c["integ"]["lights"], rb.last_light_powers(S.L)):
            l["weight"] = float(w)
    return c, img, (nc, ns), parse_trace(rec)


def ulp_diff(a, b):
    """Largest distance in units in the last place between two float32 arrays (same shape)."""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia)
    ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return np.abs(ia - ib)


def compare_queues(queues, ref_queues, desc, n_slots):
    """queues: {q: dict(rays, path)} of a wavefront render (host or device); ref_queues: trace_queues(...) of the reference.
    Returns {q: dict(n, n_ref, missing, origin_diff, dir_diff, max_ulp_dir)}: entries are matched by (pixel, sample)."""
    from b200rt import wavefront as wf
    x, y, s, _ = wf.slot_pixels(desc, n_slots)
    c = desc.crop
    x0, y0 = (c[0], c[2]) if (c[0] or c[1] or c[2] or c[3]) else (0, 0)
    out = {}
    for q in sorted(set(queues) | set(ref_queues)):
        Q = queues.get(q, dict(rays=np.zeros((0, 8), np.float32), path=np.zeros(0, np.uint32)))
        ref = ref_queues.get(q, {})
        n = len(Q["rays"])
        st = dict(n=n, n_ref=len(ref), missing=0, origin_diff=0, dir_diff=0, tmax_diff=0, max_ulp_dir=0, max_ulp_origin=0)
        if n:
            keys = list(zip((x[Q["path"]] + x0).tolist(), (y[Q["path"]] + y0).tolist(), s[Q["path"]].tolist()))
            want = np.zeros((n, 7), np.float32)
            have = np.ones(n, bool)
            for i, k in enumerate(keys):
                e = ref.get(k)
                if e is None:
                    have[i] = False
                else:
                    want[i] = e[1]
            st["missing"] = int((~have).sum())
            r = Q["rays"][have]
            w = want[have]
            uo, ud = ulp_diff(r[:, 0:3], w[:, 0:3]), ulp_diff(r[:, 4:7], w[:, 3:6])
            st["origin_diff"] = int((uo.max(1) > 0).sum()) if len(r) else 0
            st["dir_diff"] = int((ud.max(1) > 0).sum()) if len(r) else 0
            st["tmax_diff"] = int((r[:, 3].view(np.uint32) != w[:, 6].view(np.uint32)).sum())
            st["max_ulp_dir"] = int(ud.max()) if len(r) else 0
            st["max_ulp_origin"] = int(uo.max()) if len(r) else 0
        out[q] = st
    return out


def image_stats(rgb, ref):
    """Per-pixel comparison of two (h, w, 3) images: relative RMSE over the frame, and the pixels that differ by more than
    1e-3 of their own value (visibility / Russian-roulette flips show up here)."""
    rgb, ref = rgb.astype(np.float64), ref.astype(np.float64)
    d = rgb - ref
    rel_rmse = float(np.sqrt((d ** 2).mean()) / max(np.sqrt((ref ** 2).mean()), 1e-30))
    differ = np.abs(d).max(-1) > 1e-3 * np.maximum(np.abs(ref).max(-1), 1e-6)
    ok = ~differ
    rel_rmse_ok = float(np.sqrt((d[ok] ** 2).mean()) / max(np.sqrt((ref[ok] ** 2).mean()), 1e-30)) if ok.any() else 0.0
    return dict(rel_rmse=rel_rmse, differing=int(differ.sum()), frac_differing=float(differ.mean()), rel_rmse_excl=rel_rmse_ok)
