"""BASELINE config 5 — scenes/living-room (the reference's own geometry, tests/golden/living_room_mesh.npz) against
the UNMODIFIED reference (tests/golden/ref_living_room.npz, made by tests/golden/make_golden_living_room.py):

CPU (`-m "not gpu"`):
 * this repo's BVH / kd-tree builders reproduce the reference's trees over the 143 163 triangles byte for byte
   (SHA-256 of the node and index arrays; 233 709 BVH nodes as SURVEY.md §6.2 measured with the reference binary);
 * the oracle answers the golden rays exactly as the reference's BVHAccel / KdTreeAccel did.
GPU (`-m gpu`):
 * the CUDA traversal kernels give the reference's prim / t / uv / occluded on the same rays, bit for bit;
 * the wavefront render (path maxdepth 65, constant infinite light with MIS, substituted materials) matches the
   reference's own render statistically (mean, channels, 8x8 blocks, ray counts);
 * tile sharding of the 1920x1080 frame over 8 ranks is bit-identical to one rank (window of the frame)."""
import hashlib
import numpy as np
import pytest
from conftest import load_golden
from oracle import orcbind as ob


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def room():
    from b200rt import scenes
    return scenes.living_room()


@pytest.fixture(scope="module")
def golden():
    return load_golden("ref_living_room.npz")


def _build(lib, kind, tv):
    return lib.build_bvh(tv) if kind == "bvh" else lib.build_kdtree(tv)


def test_fixture_is_the_reference_scene(room):
    assert room["tri_verts"].shape == (143163, 9)       # the reference's own count
    assert room["prim_material"].shape == (143163,)
    kinds = np.array(room["material_kinds"])[room["prim_material"]]
    assert (kinds == "matte").sum() == 88114 and (kinds == "glass").sum() == 11166


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_builders_reproduce_the_references_living_room_trees(built_lib, room, golden, kind):
    t = _build(built_lib, kind, room["tri_verts"])
    assert len(t["nodes"]) == int(golden[kind + "_n_nodes"])
    if kind == "bvh":
        assert len(t["nodes"]) == 233709
    assert _sha(t["nodes"]) == str(golden[kind + "_nodes_sha256"])
    assert len(t["indices"]) == int(golden[kind + "_n_indices"])
    assert _sha(t["indices"]) == str(golden[kind + "_indices_sha256"])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_device_builders_reproduce_the_references_living_room_trees(gpu, room, golden, kind):
    """The level-synchronous DEVICE builders on the reference's own 143 163 triangles: the node arrays hash to the reference's
    SHA-256 (BVH: primNumMapping too; kd-tree: 2 941 789 nodes, leaves hold the same primitives, their order inside a leaf
    follows the stable radix sort instead of std::sort's)."""
    tv = room["tri_verts"]
    t = gpu.build_bvh_device(tv) if kind == "bvh" else gpu.build_kdtree_device(tv)
    assert len(t["nodes"]) == int(golden[kind + "_n_nodes"])
    assert _sha(t["nodes"]) == str(golden[kind + "_nodes_sha256"])
    assert len(t["indices"]) == int(golden[kind + "_n_indices"])
    if kind == "bvh":
        assert _sha(t["indices"]) == str(golden[kind + "_indices_sha256"])
    else:
        h = gpu.build_kdtree(tv)
        w0, w1 = t["nodes"][:, 0].astype(np.int64), t["nodes"][:, 1].astype(np.int64)
        leaf = ((w1 & 3) == 3) & ((w1 >> 2) > 1)
        starts, counts = w0[leaf], (w1[leaf] >> 2)
        seg = np.repeat(np.arange(len(starts)), counts)
        at = np.repeat(starts, counts) + (np.arange(len(seg)) - np.repeat(np.cumsum(counts) - counts, counts))
        a, b = t["indices"][at], h["indices"][at]
        assert np.array_equal(a[np.lexsort((a, seg))], b[np.lexsort((b, seg))])      # every leaf: the same primitives
    print("\nliving-room %s on the B200: %d nodes in %.1f ms (%d levels)" % (kind, len(t["nodes"]), t["build"]["build_ms"], t["build"]["levels"]))


def _check_hits(hits, occ, g, kind):
    ref_prim = np.where(g[kind + "_prim"] < 0, ob.MISS, g[kind + "_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    m = ref_prim != ob.MISS
    assert m.mean() > 0.9
    assert np.array_equal(hits["t"].view(np.uint32)[m], g[kind + "_t"].view(np.uint32)[m])
    assert np.array_equal(hits["b2"][m], g[kind + "_uv"][m, 1])
    assert np.array_equal((hits["b1"] + hits["b2"])[m], g[kind + "_uv"][m, 0])
    assert np.array_equal(np.asarray(occ, np.uint8), g[kind + "_occ"])


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_oracle_matches_the_reference_on_living_room(built_lib, room, golden, kind):
    tv = room["tri_verts"]
    orc = ob.OracleAccel(_build(built_lib, kind, tv), tv)
    _check_hits(orc.intersect(golden["rays"], n_threads=4), orc.intersect_p(golden["rays"], n_threads=4), golden, kind)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_gpu_traversal_matches_the_reference_on_living_room(gpu, room, golden, kind):
    acc = gpu.MakeAccelerator(kind, room["tri_verts"])
    _check_hits(acc.Intersect(golden["rays"]), acc.IntersectP(golden["rays"]), golden, kind)


def _blocks(img, k):
    h, w = img.shape[0] // k * k, img.shape[1] // k * k
    return img[:h, :w].reshape(h // k, k, w // k, k, -1).mean((1, 3))


def scenes_mod():
    from b200rt import scenes
    return scenes


def _integrator(gpu, sc, spp, maxdepth, **kw):
    from b200rt import wavefront as wf
    il = sc["infinite_light"]
    return gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_I=il["L"],
                                       infinite_light=(il["light_to_world"], wf.world_radius(sc["tri_verts"])),
                                       materials=sc["materials"], prim_material=sc["prim_material"],
                                       textures=scenes_mod().living_room_textures(sc), **kw)


@pytest.mark.gpu
def test_living_room_render_matches_the_references_render(gpu):
    """Path maxdepth 65 under the constant infinite light (what the reference builds for the missing sky map), the file's
    Sampler "sobol" and material table, the PLY vertex normals: PER PIXEL against the reference's own SamplerIntegrator::Render
    of the same setup, run here through oracle/_ref with the same sample stream (tests/shade_cases.py::living_room; the
    ray-by-ray version of this comparison is tests/test_gpu_halton.py / tests/test_shade_host.py)."""
    import hostwf, shade_cases
    from oracle import refbind as rb
    from b200rt.wavefront import film_to_rgb
    res, spp = 192, 16
    c = shade_cases.build("living_room_tex", res)      # the scene with its image textures (ABI v9)
    S = rb.RefScene(**c["ref"])
    r = dict(c["render"])
    ref, nc, ns = rb.render(S, "bvh", r.pop("eye"), r.pop("look"), r.pop("up"), r.pop("fov"), res, res, spp, c["maxdepth"],
                            r.pop("light_pos"), r.pop("light_I"), r.pop("kd"), sampler="sobol", n_threads=0, **r)
    verts, akw = c["accel"]
    for kind in ("bvh", "kdtree"):
        acc = gpu.B200Accel(gpu.build_bvh_device(verts) if kind == "bvh" else gpu.build_kdtree_device(verts), verts, **akw)
        film, st = gpu.WavefrontPathIntegrator(spp=spp, maxdepth=c["maxdepth"], sampler="sobol", **c["integ"]).Render(acc)
        im = hostwf.image_stats(film_to_rgb(film), ref)
        print("\nliving-room %s %dx%d @ %d spp sobol: rays %d/%d ref %d/%d; %s" % (kind, res, res, spp, st["closest_rays"], st["shadow_rays"], nc, ns, im))
        assert abs(st["closest_rays"] - nc) <= 1e-5 * nc + 2 and abs(st["shadow_rays"] - ns) <= 1e-5 * ns + 2
        assert 0 < st["mis_rays"] < st["closest_rays"]
        assert im["rel_rmse_excl"] <= 1e-4 and im["frac_differing"] < 1e-3, im


@pytest.mark.gpu
def test_living_room_tile_sharding_is_bit_identical(gpu):
    """Config 5's sharding: the frame split over 8 ranks by tile (t % 8 == rank) sums to the 1-rank film bit for bit."""
    from b200rt import scenes
    sc = scenes.living_room(240, 135)
    acc = gpu.MakeAccelerator("bvh", sc["tri_verts"])
    whole, st = _integrator(gpu, sc, 4, 65, seed=3).Render(acc)
    parts = np.zeros_like(whole)
    rays = 0
    for r in range(8):
        f, s = _integrator(gpu, sc, 4, 65, seed=3, rank=r, world=8).Render(acc)
        parts += f
        rays += s["closest_rays"] + s["shadow_rays"]
    assert np.array_equal(parts.view(np.uint32), whole.view(np.uint32))
    assert rays == st["closest_rays"] + st["shadow_rays"]
