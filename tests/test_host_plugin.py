"""The C++ drop-in (`Accelerator "b200"`, thesis-pbrt-v3_b200/host/b200accel.cpp) driven through the REFERENCE'S
OWN Primitive interface, side by side with the reference's own accelerators built over the same
vector<shared_ptr<Primitive>> in the same process (tests/hostplugin builds the combined library into
oracle/_ref/, which travels to the GPU box).  Bar: isect.primitive, ray.tMax, SurfaceInteraction::uv and ::p
bit-identical; IntersectP identical."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
from conftest import ROOT

HOSTLIB = os.path.join(ROOT, "oracle", "_ref", "libpbrt_b200host.so")


@pytest.fixture(scope="module")
def host(built_lib):
    if os.path.isdir("/root/reference/src") and os.path.exists("/usr/bin/make"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle", "ref"), "-j8"])
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "hostplugin")])
    if not os.path.exists(HOSTLIB):
        pytest.skip("oracle/_ref/libpbrt_b200host.so not built (needs /root/reference)")
    from oracle import refbind as rb
    L = rb.bind(HOSTLIB)
    L.plugin_accel_create.restype = C.c_void_p
    L.plugin_accel_create.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.plugin_is_current.argtypes = [C.c_void_p]
    L.plugin_world_bound.argtypes = [C.c_void_p, C.c_void_p]
    L.plugin_intersect_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    L.plugin_intersect_p_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    return rb, L


class PluginAccel:
    """`Accelerator "b200" "string tree" tree "string builder" builder + the reference's own parameter names`."""

    def __init__(self, rb, L, scene, tree, builder="reference", int_params=None):
        ip = int_params or {}
        inames = (C.c_char_p * (len(ip) + 1))(*[k.encode() for k in ip], None)
        ivals = (C.c_int * max(1, len(ip)))(*ip.values())
        self.rb, self.L, self.scene = rb, L, scene
        self.h = L.plugin_accel_create(scene.h, tree.encode(), builder.encode(), inames, ivals, None, None)

    def intersect(self, rays):   # per-ray Aggregate::Intersect through the reference harness
        hits = np.zeros(len(rays), self.rb.REF_HIT)
        self.L.ref_intersect(self.h, rays.ctypes.data, len(rays), hits.ctypes.data, None, 1)
        return hits

    def intersect_p(self, rays):
        occ = np.zeros(len(rays), np.uint8)
        self.L.ref_intersect_p(self.h, rays.ctypes.data, len(rays), occ.ctypes.data, None, 1)
        return occ

    def intersect_batch(self, rays):
        hits = np.zeros(len(rays), self.rb.REF_HIT)
        assert self.L.plugin_intersect_batch(self.h, rays.ctypes.data, len(rays), hits.ctypes.data) == 0
        return hits

    def intersect_p_batch(self, rays):
        occ = np.zeros(len(rays), np.uint8)
        assert self.L.plugin_intersect_p_batch(self.h, rays.ctypes.data, len(rays), occ.ctypes.data) == 0
        return occ


def _scene(rb, L, n_tris=1500, seed=3, spheres=True):
    from b200rt import scenes
    v = scenes.soup(n_tris, seed=seed).reshape(-1, 3)
    idx = np.arange(len(v), dtype=np.int32).reshape(-1, 3)
    rng = np.random.default_rng(seed)
    uv = rng.random((len(v), 2), dtype=np.float32)
    sp = np.array([[0, 0, 0, 2.5], [6, -3, 2, 1.25]], np.float32) if spheres else None
    return rb.RefScene(v, idx, uv=uv, spheres=sp, L=L)


def _rays(n, seed):
    from b200rt import scenes
    r = scenes.random_rays([-12, -12, -12, 12, 12, 12], n, seed=seed)
    r[n // 2:, 3] = 9.0   # finite tMax half (shadow-style)
    return np.ascontiguousarray(r, np.float32)


def _same(a, b):
    assert np.array_equal(a["prim"], b["prim"])
    for f in ("t", "uv", "p"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


def test_host_library_exports(host):
    rb, L = host
    for s in ("plugin_accel_create", "plugin_intersect_batch", "plugin_intersect_p_batch", "ref_intersect",
              "ref_accel_create", "b200rt_scene_create"):
        assert hasattr(L, s), s


def test_no_device_means_null_like_an_unknown_accelerator(host):
    """Without a CUDA device CreateB200Accelerator reports through Error() and returns nullptr, which is what
    MakeAccelerator does for a name it does not know (api.cpp:1011-1014); there is no CPU path inside."""
    rb, L = host
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    sc = _scene(rb, L, n_tris=50)
    assert PluginAccel(rb, L, sc, "bvh").h is None


CASES = [("bvh", "reference", {}, "bvh"), ("bvh", "reference", {"maxnodeprims": 1}, "bvh"),
         ("kdtree", "reference", {}, "kdtree"), ("rbsp", "reference", {"nbDirections": 7}, "rbsp"),
         ("rbspkd", "reference", {"nbDirections": 13}, "rbspkd"),
         ("bsppaper", "reference", {}, "bsppaper"), ("bsppaperkd", "reference", {}, "bsppaperkd")]


@pytest.mark.gpu
@pytest.mark.parametrize("tree,builder,ip,refkind", CASES)
def test_plugin_matches_reference_accelerator(host, gpu, tree, builder, ip, refkind):
    rb, L = host
    sc = _scene(rb, L)
    ref = rb.RefAccel(sc, refkind, int_params=ip)
    acc = PluginAccel(rb, L, sc, tree, builder, ip)
    assert acc.h, "CreateB200Accelerator returned nullptr"
    assert L.plugin_is_current(acc.h) == 1
    b = np.zeros(6, np.float32)
    L.plugin_world_bound(acc.h, b.ctypes.data)
    assert np.array_equal(b, ref.export()["bounds"])
    rays = _rays(6000, seed=11)
    want, want_p = ref.intersect(rays), ref.intersect_p(rays)
    assert (want["prim"] >= 0).sum() > 500 and (want["prim"] >= len(sc.idx)).sum() > 20   # triangles and spheres
    _same(acc.intersect_batch(rays), want)
    assert np.array_equal(acc.intersect_p_batch(rays), want_p)
    # the per-ray virtual interface (batch of one per call)
    _same(acc.intersect(rays[:400]), want[:400])
    assert np.array_equal(acc.intersect_p(rays[:400]), want_p[:400])


@pytest.mark.gpu
@pytest.mark.parametrize("tree,ip", [("bvh", {}), ("kdtree", {}), ("rbsp", {"nbDirections": 7}), ("rbspkd", {"nbDirections": 13})])
def test_plugin_with_this_repos_builders(host, gpu, tree, ip):
    """`"string builder" "b200"`: the library's own builders — BVH / kd-tree on the host, RBSP on the device."""
    rb, L = host
    sc = _scene(rb, L, spheres=False)
    ref = rb.RefAccel(sc, tree, ip)
    acc = PluginAccel(rb, L, sc, tree, "b200", ip)
    assert acc.h
    rays = _rays(4000, seed=5)
    _same(acc.intersect_batch(rays), ref.intersect(rays))
    assert np.array_equal(acc.intersect_p_batch(rays), ref.intersect_p(rays))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["killeroo", "killeroo_full", "killeroo_glass", "killeroo_metal", "killeroo_oren", "feet", "feet_crop"])
def test_cpp_wavefront_integrator_matches_the_references_render(host, gpu, case):
    """`Integrator "b200wavefront"` (host/b200integrator.cpp) over `Accelerator "b200"`, set up exactly like the
    reference's own render of tests/golden/ref_render.npz (PerspectiveCamera from LookAt/fov, Film + BoxFilter,
    MatteMaterial / PlasticMaterial, PointLight or DiffuseAreaLight on a sphere): Integrator::Render ->
    b200rt_render -> Film::SetImage/WriteImage.  `killeroo` = point light + one matte Kd; `killeroo_full` =
    BASELINE config 1 as scenes/killeroo-simple describes it (sphere area light with MIS, matte ground, two plastic
    killeroos), i.e. the whole config rendered through the C++ drop-in; `feet` = BASELINE config 4's scene
    (scenes/killeroo-feet: LightSource "distant" + one plastic killeroo, whole frame).  Same statistical bar as
    tests/test_gpu_wavefront.py (the samplers differ), plus agreement with the Python mirror of the same driver
    (same seed; the only difference is the camera frame recovered through Camera::GenerateRay in float32)."""
    from conftest import load_golden
    from b200rt import scenes
    from b200rt.wavefront import film_to_rgb
    rb, L = host
    L.plugin_render.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_void_p, C.POINTER(C.c_uint64),
                                C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p]
    g = load_golden("ref_render.npz")
    glass = case == "killeroo_glass"   # MirrorMaterial / GlassMaterial objects read by the C++ integrator
    oren = case == "killeroo_oren"     # ... with MatteMaterial sigma 40 (Oren-Nayar) on the ground
    metal = case == "killeroo_metal" or oren   # MetalMaterial / SubstrateMaterial objects
    full = case == "killeroo_full" or glass or metal
    res, spp, maxdepth = (int(x) for x in g[case + "_params"])
    feet = case.startswith("feet")
    crop = case == "feet_crop"   # the scene's Film "cropwindow" [.3 .5 .4 .8] as well
    sc = scenes.killeroo_feet(xres=res, yres=res) if feet else scenes.killeroo_simple(xres=res, yres=res)
    tv = sc["tri_verts"]
    sl = sc.get("sphere_light")
    spheres = np.array([list(sl["center"]) + [sl["radius"]]], np.float32) if full else None
    S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3), spheres=spheres, L=L)
    r = scenes._rot_z(5.0)   # scenes/killeroo-simple:1-3 (LookAt + Rotate -5 about z)
    eye = scenes._xform(r, np.array([[400.0, 20, 30]]))[0]
    look = scenes._xform(r, np.array([[0.0, 63, -110]]))[0]
    cam = np.array(list(eye) + list(look) + [0, 0, 1, 39.0], np.float32)
    light = np.asarray(sc.get("light", (0, 0, 0)), np.float32)
    I, kd = g[case + "_I"].astype(np.float32), g[case + "_kd"].astype(np.float32)
    prim_mat = emitter = None
    if full:
        mt = scenes.specular_materials(sc) if glass else scenes.metal_materials(sc, oren_nayar=oren) if metal else sc["materials"]
        prim_mat = np.ascontiguousarray(np.concatenate([mt[sc["prim_material"]], np.zeros((1, 8), np.float32)]))
        emitter = np.ascontiguousarray(sl["L"], np.float32)
    L.ref_render_set_distant.argtypes = [C.c_void_p]
    if feet:
        prim_mat = np.ascontiguousarray(np.repeat(sc["materials"], len(tv), 0))
        ft = np.concatenate([sc["light_from"], sc["light_to"]]).astype(np.float32)
        L.ref_render_set_distant(ft.ctypes.data)
    else:
        L.ref_render_set_distant(None)
    L.ref_render_set_crop.argtypes = [C.c_void_p]
    L.ref_render_set_crop(np.array(sc["cropwindow"], np.float32).ctypes.data if crop else None)
    oh, ow = (sc["crop"][3] - sc["crop"][2], sc["crop"][1] - sc["crop"][0]) if crop else (res, res)
    out = np.zeros((oh, ow, 3), np.float32)
    nc, ns = C.c_uint64(), C.c_uint64()
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        rc = L.plugin_render(S.h, b"bvh", cam.ctypes.data, res, res, spp, maxdepth, light.ctypes.data, I.ctypes.data,
                             kd.ctypes.data, 9, 0, 1, os.path.join(td, "img").encode(), out.ctypes.data, C.byref(nc), C.byref(ns),
                             None if prim_mat is None else prim_mat.ctypes.data, None if emitter is None else emitter.ctypes.data)
    L.ref_render_set_distant(None)
    L.ref_render_set_crop(None)
    assert rc == 0
    ref = g[case + "_rgb"].astype(np.float64)
    img = out.astype(np.float64)

    def blocks(a, k=8):
        return a.reshape(oh // k, k, ow // k, k, 3).mean((1, 3))
    rnc, rns = (int(x) for x in g[case + "_rays"])
    assert abs(nc.value / rnc - 1) < 0.01 and abs(ns.value / rns - 1) < 0.01
    # python mirror with the same seed
    if full:
        verts, prim_kind, rows, pm, mats = scenes.with_sphere_light(sc, plastic=True)
        if glass:
            mats = scenes.specular_materials(sc)
        if metal:
            mats = scenes.metal_materials(sc, oren_nayar=oren)
        acc = gpu.B200Accel(gpu.build_bvh(verts), verts, prim_kind=prim_kind, spheres=rows)
        integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_pos=sl["center"], light_I=sl["L"], seed=9,
                                            sphere_light=(sl["radius"], len(verts) - 1), materials=mats, prim_material=pm)
    elif feet:
        from b200rt.wavefront import world_radius
        acc = gpu.MakeAccelerator("bvh", tv)
        integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_I=I, seed=9,
                                            distant_light=(sc["light_from"] - sc["light_to"], world_radius(tv)),
                                            materials=sc["materials"], prim_material=np.zeros(len(tv), np.uint8),
                                            crop=sc["crop"] if crop else None)
    else:
        acc = gpu.MakeAccelerator("bvh", tv)
        integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_pos=light, light_I=I, kd=kd, seed=9)
    film, st = integ.Render(acc)
    py = film_to_rgb(film).astype(np.float64)
    # mask the emitter and the pixels it covers partly (see tests/test_gpu_wavefront.py)
    near = (ref.max(-1) > 50) | (img.max(-1) > 50) | (py.max(-1) > 50)
    for _ in range(2):
        near = near | np.roll(near, 1, 0) | np.roll(near, -1, 0) | np.roll(near, 1, 1) | np.roll(near, -1, 1)
    img, ref, py = (np.where(near[..., None], 0, a) for a in (img, ref, py))
    if glass or metal:   # caustic fireflies: raw mean within 3 %, the tight bars on the images clamped at 1 (see tests/test_gpu_wavefront.py)
        assert abs(img.mean() / ref.mean() - 1) < 0.03
        img, ref, py = (np.minimum(a, 1.0) for a in (img, ref, py))
    assert abs(img.mean() / ref.mean() - 1) < 0.01
    cov = blocks(ref).max(-1) > 0   # `feet` covers ~10 % of the frame: RMSE over the covered blocks and their mean
    rm = blocks(ref)[cov].mean()
    assert np.sqrt(((blocks(img) - blocks(ref)) ** 2)[cov].mean()) / rm < (0.03 if case == "killeroo" else 0.085 if glass else 0.04)
    # the C++ drop-in and the Python mirror hand the device the same description (reference camera matrices, Halton stream,
    # materials, light): the same film up to the Film's RGB -> XYZ -> RGB round trip, and the same rays
    assert np.abs(img - py).max() <= 2e-6 * max(1.0, np.abs(py).max()), np.abs(img - py).max()
    assert (st["closest_rays"], st["shadow_rays"]) == (nc.value, ns.value)


@pytest.mark.gpu
def test_cpp_integrator_renders_living_room(host, gpu):
    """BASELINE config 5 through the C++ drop-in: scenes/living-room's geometry with its PLY vertex normals, the camera CTM
    (`Transform [...]`), the file's `Sampler "sobol"`, path maxdepth 65, `LightSource "infinite"` with the constant map the
    reference falls back to, the file's material table (matte, Oren-Nayar, substrate, metal, glass, mirror, uber; 20-float
    records) — EQUAL to the Python mirror, and per pixel against the reference's own render of the same objects."""
    import hostwf
    from b200rt import scenes, wavefront as wf
    from b200rt.wavefront import film_to_rgb
    rb, L = host
    L.plugin_render.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_void_p, C.POINTER(C.c_uint64),
                                C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p]
    xres, yres, spp, maxdepth = 160, 90, 16, 65
    sc = scenes.living_room(xres, yres)          # with its image textures (ABI v9)
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "living_room_mesh.npz"))
    tv, il = sc["tri_verts"], sc["infinite_light"]
    from b200rt import textures as tx
    import tempfile
    texdir = tempfile.mkdtemp()
    texfiles = []
    for n, px in zip(sc["texture_names"], sc["texture_pixels"]):
        texfiles.append(os.path.join(texdir, n + ".tga"))
        tx.write_tga(texfiles[-1], px)
    rb.set_textures(texfiles, L)
    S = rb.RefScene(z["P"], z["idx"], uv=z["UV"], normals=z["N"], L=L)
    l2w = np.eye(4, dtype=np.float32)
    l2w[:3, :3] = il["light_to_world"]
    w2c = np.ascontiguousarray(sc["camera"]["world_to_camera"], np.float32)
    Lrgb = np.ascontiguousarray(il["L"], np.float32)
    L.ref_render_set_infinite.argtypes = [C.c_void_p, C.c_void_p]
    L.ref_render_set_camera_matrix.argtypes = [C.c_void_p]
    L.ref_render_set_infinite(Lrgb.ctypes.data, l2w.ctypes.data)
    L.ref_render_set_camera_matrix(w2c.ctypes.data)
    L.ref_render_set_mat_stride(20)
    L.plugin_set_sampler(b"sobol")
    cam = np.array([0, 0, 0, 0, 0, 1, 0, 1, 0, 58.715508], np.float32)
    prim_mat = np.ascontiguousarray(sc["materials"][sc["prim_material"]])
    zero, kd = np.zeros(3, np.float32), np.full(3, .5, np.float32)
    out = np.zeros((yres, xres, 3), np.float32)
    nc, ns = C.c_uint64(), C.c_uint64()
    import tempfile
    try:
        with tempfile.TemporaryDirectory() as td:
            rc = L.plugin_render(S.h, b"bvh", cam.ctypes.data, xres, yres, spp, maxdepth, zero.ctypes.data, Lrgb.ctypes.data,
                                 kd.ctypes.data, 9, 0, 1, os.path.join(td, "img").encode(), out.ctypes.data, C.byref(nc), C.byref(ns),
                                 prim_mat.ctypes.data, None)
    finally:
        L.ref_render_set_infinite(None, None)
        L.ref_render_set_camera_matrix(None)
        L.ref_render_set_mat_stride(8)
        L.plugin_set_sampler(b"halton")
        rb.set_textures([], L)
    assert rc == 0
    # the Python mirror: same C ABI, same records (the harness's single mesh gives every triangle the has-normals flag)
    acc = gpu.B200Accel(gpu.build_bvh_device(tv), tv, tri_uvs=sc["tri_uvs"], tri_normals=sc["tri_normals"],
                        prim_flags=np.full(len(tv), 2, np.uint8))
    integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_I=il["L"], sampler="sobol",
                                        infinite_light=(il["light_to_world"], wf.world_radius(tv)),
                                        materials=sc["materials"], prim_material=sc["prim_material"],
                                        textures=[tx.ImageTexture(rgb8=px) for px in sc["texture_pixels"]])
    film, st = integ.Render(acc)
    py = film_to_rgb(film)
    # Film::WriteImage converts XYZ -> RGB on the way out (film.cpp:266-305): a rounding of the last bit, nothing else
    assert np.abs(out - py).max() <= 2e-6 * max(1.0, float(np.abs(py).max())), np.abs(out - py).max()
    assert (st["closest_rays"], st["shadow_rays"]) == (nc.value, ns.value)
    # the reference's own render of the same objects, same sample stream
    ref, rnc, rns = rb.render(S, "bvh", (0, 0, 0), (0, 0, 1), (0, 1, 0), 58.715508, xres, yres, spp, maxdepth, (0, 0, 0), il["L"],
                              (.5, .5, .5), sampler="sobol", n_threads=0, prim_mat=prim_mat, infinite=(il["L"], l2w),
                              camera_matrix=sc["camera"]["world_to_camera"], textures=sc["texture_pixels"])
    im = hostwf.image_stats(out, ref)
    print("\nliving-room through the C++ drop-in, %dx%d @ %d spp sobol: rays %d/%d ref %d/%d; %s" % (xres, yres, spp, nc.value, ns.value, rnc, rns, im))
    assert abs(nc.value - rnc) <= 1e-5 * rnc + 2 and abs(ns.value - rns) <= 1e-5 * rns + 2
    assert im["rel_rmse_excl"] <= 1e-4 and im["frac_differing"] < 1e-3, im


@pytest.mark.gpu
@pytest.mark.parametrize("tree", ["bvh", "kdtree"])
def test_cpp_integrator_writes_the_thesis_statistic_matrices(host, gpu, tree):
    """`Integrator "b200wavefront" "bool generalstats" "true"`: the per-pixel GeneralStats the thesis accumulates
    (integrator.cpp:327-328) come from the instrumented kernels, are stored in the reference's Film pixels and written by
    the reference's own Film::WriteGeneralStats (film.cpp:170-209) as `<image>-<counter>.txt` matrices.  Compared with
    the Python mirror of the same driver (same seed: totals within 0.2 %) and, for the BVH, with the matrices of the
    reference's own render of the frame (golden killeroo_heat; different sampler: totals 1.5 %, 8x8 blocks 4 %).
    The node counter lands in the field the tree's family owns: kdTreeNodeTraversals for the kd-tree (the BVH's own
    counter is not among the matrices the reference dumps)."""
    from conftest import load_golden
    from b200rt import scenes, wavefront as wf
    rb, L = host
    L.plugin_render.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_void_p, C.POINTER(C.c_uint64),
                                C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p]
    g = load_golden("ref_render.npz")
    res, spp, maxdepth = (int(x) for x in g["killeroo_params"])
    if tree != "bvh":
        spp = 16
    sc = scenes.killeroo_simple(xres=res, yres=res)
    tv = sc["tri_verts"]
    S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3), L=L)
    r = scenes._rot_z(5.0)
    eye = scenes._xform(r, np.array([[400.0, 20, 30]]))[0]
    look = scenes._xform(r, np.array([[0.0, 63, -110]]))[0]
    cam = np.array(list(eye) + list(look) + [0, 0, 1, 39.0], np.float32)
    light = np.asarray(sc["light"], np.float32)
    I, kd = g["killeroo_I"].astype(np.float32), g["killeroo_kd"].astype(np.float32)
    out = np.zeros((res, res, 3), np.float32)
    nc, ns = C.c_uint64(), C.c_uint64()
    import tempfile
    names = ["primitiveIntersections", "primitiveIntersectionsP", "kdTreeNodeTraversals", "kdTreeNodeTraversalsP",
             "leafNodeTraversals", "leafNodeTraversalsP", "bspTreeNodeTraversals"]
    mats = {}
    L.plugin_set_generalstats(1)
    try:
        with tempfile.TemporaryDirectory() as td:
            rc = L.plugin_render(S.h, tree.encode(), cam.ctypes.data, res, res, spp, maxdepth, light.ctypes.data, I.ctypes.data,
                                 kd.ctypes.data, 9, 0, 1, os.path.join(td, "img").encode(), out.ctypes.data, C.byref(nc), C.byref(ns),
                                 None, None)
            assert rc == 0
            for n in names:
                mats[n] = np.loadtxt(os.path.join(td, "img-%s.txt" % n), dtype=np.uint64).reshape(res, res)
    finally:
        L.plugin_set_generalstats(0)
    # builder "host": the any-hit counters depend on the order of the primitives INSIDE a kd leaf (IntersectP leaves at the first
    # hit), which the host builder shares with the reference's and the device builder (stable radix sort) does not
    acc = gpu.MakeAccelerator(tree, tv, builder="host")
    integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_pos=light, light_I=I, kd=kd, seed=9)
    _, heat, _ = wf.render_heatmaps(integ, acc)
    planes = {"primitiveIntersections": 0, "primitiveIntersectionsP": 1, "leafNodeTraversals": 4, "leafNodeTraversalsP": 5}
    if tree == "kdtree":
        planes.update(kdTreeNodeTraversals=2, kdTreeNodeTraversalsP=3)
        assert mats["kdTreeNodeTraversals"].sum() > 0
    else:
        assert mats["kdTreeNodeTraversals"].sum() == 0
    assert mats["bspTreeNodeTraversals"].sum() == 0
    ref = g["killeroo_heat"].astype(np.float64)
    print("\n%s: rays C++ %d/%d; %s" % (tree, nc.value, ns.value, {n: (int(mats[n].sum()), int(heat[k].sum()), int(ref[k].sum())) for n, k in planes.items()}))
    for n, k in planes.items():
        a, b = mats[n].astype(np.float64), heat[k].astype(np.float64)
        assert np.array_equal(a, b), (n, a.sum(), b.sum())   # same description, same Halton stream: the same rays
        if tree == "bvh":
            assert abs(a.sum() / ref[k].sum() - 1) < 0.015, (n, a.sum(), ref[k].sum())
            ab = a.reshape(res // 8, 8, res // 8, 8).mean((1, 3))
            bb = ref[k].reshape(res // 8, 8, res // 8, 8).mean((1, 3))
            assert np.sqrt(((ab - bb) ** 2).mean()) / bb.mean() < 0.04, n
