"""GPU: matched-sample parity of the wavefront driver with the REFERENCE'S OWN renderer.

With `sampler="halton"` / `"sobol"` the device draws the reference sampler's values in PathIntegrator::Li's dimension order,
generates camera rays with the reference camera's matrices and shades with the reference's frame and float operations
(csrc/wf_*.cuh; the host build of the same text is pinned bit for bit in tests/test_shade_host.py).  What is left to differ
on the device is libm: sin / cos / acos are evaluated in double and rounded once (wf_core.cuh), which agrees with glibc's
float functions on all but ~1e-4 of the calls; a 1-ulp slip then moves everything downstream of it on that path by ulps.

  * test_device_rays_are_the_references: the reference's single-threaded render is traced event by event (oracle/ref/
    ref_render.cpp); every ray it handed to Scene::Intersect / IntersectP must be in the device queue of that depth
    (b200rt_render_queue / _queue_paths), matched by (pixel, sample): queue sizes identical, origins and directions bit-equal
    on >= 99 % of the rays and within a few ulp on the rest.
  * test_image_is_the_references_halton_render: the config scenes at their own resolution / spp against the reference's own
    multithreaded Halton render: identical ray counts, per-pixel relative RMSE <= 1e-4, < 1e-3 of the pixels off by more
    than 1e-3 (reported)."""
import numpy as np
import pytest
from oracle import refbind as rb

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not rb.available(), reason="oracle/_ref is not built")]


def _accel(gpu, c):
    verts, akw = c["accel"]
    return gpu.B200Accel(gpu.build_bvh(verts), verts, **akw)


def _device_queues(gpu, c, spp, sampler):
    from b200rt import wavefront as wf
    integ = gpu.WavefrontPathIntegrator(spp=spp, maxdepth=c["maxdepth"], sampler=sampler, **c["integ"])
    acc = _accel(gpu, c)
    wf.keep_queues(acc, True)
    film, st = integ.Render(acc)
    maxd = c["maxdepth"]
    queues = {}
    for q in range(2 * (maxd + 1) + (maxd + 1 if c["mis"] else 0)):
        if q < 2 * (maxd + 1) and q % 2 == 1 and q // 2 >= maxd:
            continue
        queues[q] = dict(rays=wf.queue_rays(acc, q), path=wf.queue_paths(acc, q))
    wf.keep_queues(acc, False)
    return integ, film, st, queues


@pytest.mark.parametrize("case,sampler", [("sponza_point", "halton"), ("killeroo_plastic", "halton"), ("killeroo_normals", "halton"), ("killeroo_glass", "halton"),
                                          ("killeroo_metal", "halton"), ("killeroo_oren", "halton"), ("killeroo_uber", "halton"), ("killeroo_infinite", "halton"),
                                          ("feet_distant", "halton"), ("normals", "halton"), ("normals_flipped", "halton"),
                                          ("killeroo_metal", "sobol"), ("feet_distant", "sobol"), ("living_room", "sobol"),
                                          ("living_room_tex", "sobol"), ("sponza_two_points", "halton"), ("killeroo_lights", "halton"),
                                          ("killeroo_lights_power", "halton"), ("killeroo_lights", "sobol"), ("killeroo_lights_spatial", "halton"), ("living_room_tex_ewa", "sobol")])
def test_device_rays_are_the_references(gpu, case, sampler):
    import hostwf
    from b200rt import wavefront as wf
    res, spp = 48, 2
    c, img, (nc, ns), samples = hostwf.reference_trace(case, spp, res, sampler=sampler)
    integ, film, st, queues = _device_queues(gpu, c, spp, sampler)
    tq = hostwf.trace_queues(samples, c["maxdepth"], c["mis"])
    n_slots = int(np.prod(integ.tiles()[0:1])) * 256 * spp
    cmp = hostwf.compare_queues(queues, tq, integ.desc, n_slots)
    total = differ = 0
    worst = 0
    for q, r in cmp.items():
        total += r["n"]
        differ += max(r["origin_diff"], r["dir_diff"]) + r["missing"] + abs(r["n"] - r["n_ref"])
        worst = max(worst, r["max_ulp_dir"], r["max_ulp_origin"])
    print("\n%s/%s: %d rays, %d not bit-equal (%.3g %%), worst %d ulp; counts dev %d/%d ref %d/%d" % (
        case, sampler, total, differ, 100.0 * differ / max(total, 1), worst, st["closest_rays"], st["shadow_rays"], nc, ns))
    # camera rays: no libm call on the way -> all bit-equal
    assert cmp[0]["n"] == cmp[0]["n_ref"] and cmp[0]["origin_diff"] == 0 and cmp[0]["dir_diff"] == 0, cmp[0]
    # first shadow rays: the shading point, OffsetRayOrigin and the light sample; trig only in the area / infinite light samples
    assert cmp[1]["n"] == cmp[1]["n_ref"] and cmp[1]["origin_diff"] == 0, cmp[1]
    assert differ <= 0.01 * total, cmp
    assert abs(st["closest_rays"] - nc) <= max(2, 1e-3 * nc) and abs(st["shadow_rays"] - ns) <= max(2, 1e-3 * ns)
    im = hostwf.image_stats(wf.film_to_rgb(film), img)
    print("image:", im)
    assert im["rel_rmse_excl"] < 1e-4 and im["frac_differing"] < 5e-3, im


IMAGE_CASES = {
    # BASELINE config 1: scenes/killeroo-simple as written — 700 x 700, Sampler "halton" 8 spp, path maxdepth 5, sphere area light, plastic
    "killeroo_plastic": dict(res=700, spp=8),
    # ... with the vertex normals of its loop-subdivided killeroos (interpolated shading normals, triangle.cpp:348-400)
    "killeroo_normals": dict(res=700, spp=8),
    # BASELINE config 2 (the bench workload): sponza stand-in, point light, path maxdepth 5, halton
    "sponza_point": dict(res=700, spp=4, full_detail=True),
    # BASELINE config 4: scenes/killeroo-feet through its crop window (res x 4 frame), distant light + plastic
    "feet_distant": dict(res=320, spp=8),
    "killeroo_glass": dict(res=256, spp=8),
    "killeroo_metal": dict(res=256, spp=8),
    "killeroo_uber": dict(res=256, spp=8),
    # BASELINE config 5: scenes/living-room as the reference resolves it (the file's Sampler "sobol", maxdepth 65, infinite light,
    # its material table incl. "uber", the PLY vertex normals), square frame
    "living_room": dict(res=256, spp=16, sampler="sobol"),
    # ... and with the image textures of the files that are in the reference tree (ABI v9: leaf.tga as Kd and opacity of the uber
    # Leaves, picture8.tga on the Painting; MIP level from the camera rays' differentials)
    "living_room_tex": dict(res=256, spp=16, sampler="sobol"),
}


@pytest.mark.parametrize("case", list(IMAGE_CASES))
def test_image_is_the_references_halton_render(gpu, case):
    import hostwf, shade_cases
    from b200rt import wavefront as wf, scenes
    p = IMAGE_CASES[case]
    c = shade_cases.build(case, p["res"])
    if p.get("full_detail"):   # the bench's own 262 136-triangle stand-in
        sc = scenes.sponza_standin(xres=p["res"], yres=p["res"])
        tv = sc["tri_verts"]
        c["ref"] = dict(P=tv.reshape(-1, 3), indices=np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
        c["accel"] = (tv, {})
    S = rb.RefScene(**c["ref"])
    r = dict(c["render"])
    rres = c.get("res", p["res"])
    img, nc, ns = rb.render(S, "bvh", r.pop("eye"), r.pop("look"), r.pop("up"), r.pop("fov"), rres, rres, p["spp"], c["maxdepth"],
                            r.pop("light_pos"), r.pop("light_I"), r.pop("kd"), sampler=p.get("sampler", "halton"), n_threads=0, **r)
    integ = gpu.WavefrontPathIntegrator(spp=p["spp"], maxdepth=c["maxdepth"], sampler=p.get("sampler", "halton"), **c["integ"])
    film, st = integ.Render(_accel(gpu, c))
    im = hostwf.image_stats(wf.film_to_rgb(film), img)
    print("\n%s %dx%d @ %d spp: rays dev %d/%d ref %d/%d; %s" % (case, img.shape[1], img.shape[0], p["spp"], st["closest_rays"],
                                                              st["shadow_rays"], nc, ns, im))
    assert abs(st["closest_rays"] - nc) <= 1e-5 * nc + 2 and abs(st["shadow_rays"] - ns) <= 1e-5 * ns + 2
    assert im["rel_rmse_excl"] <= 1e-4, im
    assert im["frac_differing"] < 1e-3, im
