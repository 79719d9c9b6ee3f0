"""CPU: the wavefront shade stage's restatement of the reference (csrc/wf_*.cuh), compiled for the host by tests/hostshade
and driven through the wavefront loop by tests/hostwf.py with the CPU oracle tracing the rays, against the event trace of
the REFERENCE'S OWN single-threaded render (oracle/_ref: SamplerIntegrator::Render + PathIntegrator::Li with its own
HaltonSampler / SobolSampler, materials and lights):
  * every value the reference's sampler returned equals sampler_dim(sampler_index(pixel, sample), dimension) bit for bit;
  * b200rt_camera_perspective gives the reference camera's RasterToCamera / CameraToWorld matrices bit for bit;
  * every ray the reference handed to Scene::Intersect / IntersectP (camera rays, shadow rays, EstimateDirect's BSDF-sampled
    rays, bounce rays) is in the wavefront's queue of that depth with a bit-equal origin, direction and tMax, and the queue
    sizes are identical (glibc's sinf / cosf are the reference's, so on the host nothing is left to differ);
  * the film equals the reference's image to float rounding (the film's XYZ round trip).
The GPU tests (tests/test_gpu_halton.py) run the same comparison with the device build of the same text."""
import ctypes as C
import numpy as np
import pytest
from oracle import refbind as rb, orcbind as ob

pytestmark = pytest.mark.skipif(not rb.available(), reason="oracle/_ref is not built")


def _host_render(built_lib, c, spp, sampler="halton", sobol_tables=None):
    import hostwf
    integ = built_lib.WavefrontPathIntegrator(spp=spp, maxdepth=c["maxdepth"], sampler=sampler, sobol_tables=sobol_tables, **c["integ"])
    verts, akw = c["accel"]
    tree = built_lib.build_bvh(verts)
    orc = ob.OracleAccel(tree, verts, prim_kind=akw.get("prim_kind"), spheres=akw.get("spheres"))
    H = hostwf.HostWavefront(integ.desc, verts, orc, tri_uvs=akw.get("tri_uvs"), tri_normals=akw.get("tri_normals"),
                             prim_flags=akw.get("prim_flags"), prim_is_sphere=akw.get("prim_kind"))
    film, queues, st = H.render()
    return integ, H, film, queues, st


def test_camera_matrices_equal_the_reference_cameras(built_lib):
    from b200rt import wavefront as wf, scenes
    cams = [((-0.7, .45, .05), (0, .25, 0), (0, 1, 0), 42.0, 700, 700), ((400, 20, 30), (0, 63, -110), (0, 0, 1), 39.0, 640, 360),
            ((3.2, 1.1, .9), (0, 0, 0), (0, 0, 1), 40.0, 96, 128)]
    for eye, look, up, fov, xr, yr in cams:
        r2c, c2w = rb.camera_matrices(eye, look, up, fov, xr, yr)
        d = wf.WavefrontPathIntegrator(dict(lookat=list(eye) + list(look) + list(up), fov=fov, xres=xr, yres=yr)).desc
        assert np.array_equal(np.array(d.raster_to_camera, np.float32).view(np.uint32), r2c.ravel().view(np.uint32))
        assert np.array_equal(np.array(d.camera_to_world, np.float32).view(np.uint32), c2w.ravel().view(np.uint32))
    w2c = scenes.living_room(xres=64, yres=36)["camera"]["w2c"]   # `Transform [...]` CTM with determinant -1
    r2c, c2w = rb.camera_matrices((0, 0, 0), (0, 0, 1), (0, 1, 0), 58.715508, 1280, 720, camera_matrix=w2c)
    d = wf.WavefrontPathIntegrator(dict(w2c=w2c, fov=58.715508, xres=1280, yres=720)).desc
    assert np.array_equal(np.array(d.raster_to_camera, np.float32).view(np.uint32), r2c.ravel().view(np.uint32))
    assert np.array_equal(np.array(d.camera_to_world, np.float32).view(np.uint32), c2w.ravel().view(np.uint32))


@pytest.mark.parametrize("case", ["sponza_point", "killeroo_plastic", "killeroo_normals", "killeroo_glass", "killeroo_metal", "killeroo_oren", "killeroo_uber",
                                  "killeroo_infinite", "feet_distant", "normals", "normals_flipped", "living_room", "living_room_tex",
                                  "sponza_two_points", "killeroo_lights", "killeroo_lights_power", "killeroo_lights_spatial", "living_room_tex_ewa"])
def test_every_sample_and_every_ray_is_the_references(built_lib, case):
    _check_case(built_lib, case, "halton")


@pytest.mark.parametrize("case", ["sponza_point", "killeroo_metal", "feet_distant", "living_room", "living_room_tex", "living_room_tex_ewa"])
def test_sobol_stream_is_the_references(built_lib, case):
    """Sampler "sobol" (scenes/living-room/scene.pbrt:3): SobolIntervalToIndex + SobolSampleFloat over the reference's
    generator matrices (b200rt/data/sobol_tables.npz), incl. the pixel-relative remap of dimensions 0 / 1 and a crop window."""
    _check_case(built_lib, case, "sobol")


def _check_case(built_lib, case, sampler):
    import hostwf
    from b200rt import wavefront as wf
    res, spp = (32, 2) if case == "sponza_point" else (40, 2)
    c, img, (nc, ns), samples = hostwf.reference_trace(case, spp, res, sampler=sampler)
    integ, H, film, queues, st = _host_render(built_lib, c, spp, sampler=sampler)
    # --- sample values, by (pixel, sample, dimension)
    px, py, s_, dim, val = [], [], [], [], []
    for smp in samples:
        for e in smp["events"]:
            for k in range(1 if e[0] == "1d" else 2 if e[0] == "2d" else 0):
                px.append(smp["pixel"][0]); py.append(smp["pixel"][1]); s_.append(smp["s"]); dim.append(e[1] + k); val.append(e[2][k])
    got = H.sample_values(px, py, s_, dim)
    assert np.array_equal(got.view(np.uint32), np.array(val, np.float32).view(np.uint32))
    assert max(dim) > 10
    # --- rays per queue
    assert (st["closest_rays"], st["shadow_rays"]) == (nc, ns), (st, nc, ns)
    tq = hostwf.trace_queues(samples, c["maxdepth"], c["mis"])
    cmp = hostwf.compare_queues(queues, tq, integ.desc, int(hostwf.lib().hs_n_paths(H.h)))
    for q, r in cmp.items():
        assert r["n"] == r["n_ref"] and r["missing"] == 0, (q, r)
        assert r["origin_diff"] == 0 and r["dir_diff"] == 0 and r["tmax_diff"] == 0, (q, r)
    if c["mis"]:
        assert st["mis_rays"] > 0
    # --- image
    im = hostwf.image_stats(wf.film_to_rgb(film), img)
    assert im["rel_rmse"] < 2e-6 and im["differing"] == 0, im


def test_sinf_cosf_restatement_equals_this_libm():
    """wf_core.cuh restates glibc's sinf / cosf (the reference's std::sin / std::cos on floats) so that the DEVICE gets the
    reference's values; here the host build of that text is compared with the libm of the machine the reference runs on."""
    import hostwf
    L = hostwf.lib()
    L.hs_sincos_check.restype = C.c_uint64
    L.hs_sincos_check.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    # every 11th float in [2^-14, 8), both signs: identical (the samplers' angles are below 2 pi)
    assert L.hs_sincos_check(0x38800000, 0x41000000, 11) == 0
    # [8, 120): this libm's FMA build of the range reduction rounds a handful of arguments (~1e-6 of them) the other way
    assert L.hs_sincos_check(0x41000000, 0x42F00000, 11) <= 40
