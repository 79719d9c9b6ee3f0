"""GPU: the CUDA path (through the C ABI) against the oracle on seeded inputs, against the committed
reference goldens, and through size-independent properties at full size.
Bar: primitive ids, t and barycentrics BIT-EXACT; any-hit exact; counters exact."""
import numpy as np
import pytest
from conftest import load_golden, golden_tree, SMALL, FEET, feet_tree
from oracle import orcbind as ob

pytestmark = pytest.mark.gpu


def _same_hits(h, o):
    assert np.array_equal(h["prim"], o["prim"])
    for f in ("t", "b0", "b1", "b2"):
        assert np.array_equal(h[f].view(np.uint32), o[f].view(np.uint32)), f


@pytest.mark.parametrize("name", SMALL)
def test_golden_small(gpu, name):
    g = load_golden("ref_small_%s.npz" % name)
    acc = gpu.B200Accel(golden_tree(g), g["tri_verts"])
    hits = acc.Intersect(g["rays"])
    ref_prim = np.where(g["hit_prim"] < 0, gpu.MISS, g["hit_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    assert np.array_equal(hits["t"].view(np.uint32), g["hit_t"].view(np.uint32))
    m = g["hit_prim"] >= 0
    assert np.array_equal(hits["b2"][m], g["hit_uv"][m, 1])
    assert np.array_equal((hits["b1"] + hits["b2"])[m], g["hit_uv"][m, 0])
    assert np.array_equal(acc.IntersectP(g["rays"]), g["occluded"])
    # instrumented kernels reproduce the reference's traversal counters (Ray::stats)
    hits2, c = acc.Intersect(g["rays"], counters=True)
    _same_hits(hits2, hits)
    assert c["rays"] == len(g["rays"])
    assert c["prim_tests"] == int(g["cnt_tests"].sum())
    assert c["leaf_visits"] == int(g["cnt_leaf"].sum())
    interior = c["nodes_passed"] - (c["leaf_visits"] if name == "bvh" else 0)
    assert interior == int(g["cnt_interior"].sum())
    occ2, cp = acc.IntersectP(g["rays"], counters=True)
    assert np.array_equal(occ2, g["occluded"])
    assert cp["prim_tests"] == int(g["cntp_tests"].sum())
    assert cp["leaf_visits"] == int(g["cntp_leaf"].sum())
    # PER-RAY statistics = the reference's Ray::stats after each call, ray by ray (what the thesis's heat maps sum)
    h3, rs = acc.IntersectStats(g["rays"])
    _same_hits(h3, hits)
    assert np.array_equal(rs["prim_tests"], g["cnt_tests"])
    assert np.array_equal(rs["interior_visits"], g["cnt_interior"])
    assert np.array_equal(rs["leaf_visits"], g["cnt_leaf"])
    assert int(rs["nodes_fetched"].sum()) == c["nodes_fetched"]
    o3, rsp = acc.IntersectStats(g["rays"], any_hit=True)
    assert np.array_equal(o3, g["occluded"])
    assert np.array_equal(rsp["prim_tests"], g["cntp_tests"])
    assert np.array_equal(rsp["interior_visits"], g["cntp_interior"])
    assert np.array_equal(rsp["leaf_visits"], g["cntp_leaf"])


@pytest.mark.parametrize("name", FEET)
def test_golden_feet(gpu, name):
    """BASELINE config 4 (scenes/killeroo-feet + scripts/run_rbsps.sh): reference-built RBSP 3/7/9/13,
    RBSPKd 13 and BVH over the feet crop, compared with the reference's own answers on the same rays."""
    g = load_golden("ref_feet.npz")
    acc = gpu.B200Accel(feet_tree(g, name), g["tri_verts"])
    hits = acc.Intersect(g["rays"])
    ref_prim = np.where(g[name + "_prim"] < 0, gpu.MISS, g[name + "_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    assert np.array_equal(hits["t"].view(np.uint32), g[name + "_t"].view(np.uint32))
    assert np.array_equal(acc.IntersectP(g["rays"]), g[name + "_occ"])
    # the instrumented (one-ray-per-thread) kernels agree with the persistent production kernels
    h2, c = acc.Intersect(g["rays"], counters=True)
    _same_hits(h2, hits)
    o2, _ = acc.IntersectP(g["rays"], counters=True)
    assert np.array_equal(o2, g[name + "_occ"])
    assert c["rays"] == len(g["rays"])


@pytest.mark.parametrize("name", ["bvh", "kdtree", "rbsp7"])
def test_golden_spheres(gpu, name):
    """Spheres in the aggregate (a16): reference-built trees over 300 triangles + 4 spheres."""
    from conftest import spheres_scene
    g = load_golden("ref_spheres.npz")
    tree, verts, kind, rows = spheres_scene(g, name)
    acc = gpu.B200Accel(tree, verts, prim_kind=kind, spheres=rows)
    hits = acc.Intersect(g["rays"])
    ref_prim = np.where(g[name + "_prim"] < 0, gpu.MISS, g[name + "_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    assert np.array_equal(hits["t"].view(np.uint32), g[name + "_t"].view(np.uint32))
    assert np.array_equal(acc.IntersectP(g["rays"]), g[name + "_occ"])
    h2, c = acc.Intersect(g["rays"], counters=True)   # instrumented one-ray-per-thread kernels
    _same_hits(h2, hits)
    assert c["prim_tests"] == int(g[name + "_tests"].sum())
    # malformed sphere tables are rejected
    with pytest.raises(gpu.B200RTError):
        gpu.B200Accel(tree, verts, prim_kind=kind, spheres=rows[:2])


def test_golden_triangles(gpu):
    """Single-triangle scenes: watertight vertex/edge rays, 1e+-8 coordinates, BadCases, degenerates."""
    g = load_golden("ref_triangle.npz")
    n = len(g["tris"])
    for i in range(n):   # every case of the fixture
        v = g["tris"][i:i + 1]
        acc = gpu.B200Accel(gpu.build_bvh(v), v)
        h = acc.Intersect(g["rays"][i:i + 1])
        assert (h["prim"][0] != gpu.MISS) == bool(g["hit"][i]), i
        assert acc.IntersectP(g["rays"][i:i + 1])[0] == g["occ"][i], i
        if g["hit"][i]:
            assert h["t"][0] == g["t"][i]
            assert h["b2"][0] == g["uv"][i, 1]
        acc.close()


def _rays(n, seed, extent=12.0):
    rng = np.random.default_rng(seed)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = rng.uniform(-extent, extent, (n, 3))
    d = rng.normal(size=(n, 3))
    r[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    r[:, 3] = np.inf
    r[1::3, 3] = rng.uniform(0.2, 15, len(r[1::3]))
    rows = np.arange(5, n, 50)
    r[rows, 4 + (np.arange(len(rows)) % 3)] = 0.0
    r[7::100, 4:7] *= 33.0
    return r


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
@pytest.mark.parametrize("ntri", [1, 2, 37, 5000])
def test_vs_oracle_own_builders(gpu, kind, ntri):
    from b200rt import scenes
    v = scenes.soup(ntri, seed=ntri)
    tree = gpu.build_bvh(v) if kind == "bvh" else gpu.build_kdtree(v)
    acc = gpu.B200Accel(tree, v)
    orc = ob.OracleAccel(tree, v)
    rays = _rays(20000, ntri + 1)
    _same_hits(acc.Intersect(rays), orc.intersect(rays, n_threads=4))
    assert np.array_equal(acc.IntersectP(rays), orc.intersect_p(rays, n_threads=4))
    h, c = acc.Intersect(rays, counters=True)
    _, oc = orc.intersect(rays, counters=True, n_threads=4)
    for f in ("nodes_fetched", "nodes_passed", "leaf_visits", "prim_tests", "index_reads"):
        assert c[f] == int(oc[f]), f
    o, c = acc.IntersectP(rays, counters=True)
    _, oc = orc.intersect_p(rays, counters=True, n_threads=4)
    for f in ("nodes_fetched", "nodes_passed", "leaf_visits", "prim_tests", "index_reads"):
        assert c[f] == int(oc[f]), f


def test_edge_cases(gpu):
    from b200rt import scenes
    v = scenes.soup(100, seed=3)
    acc = gpu.MakeAccelerator("bvh", v)
    # empty batch
    assert len(acc.Intersect(np.zeros((0, 8), np.float32))) == 0
    assert len(acc.IntersectP(np.zeros((0, 8), np.float32))) == 0
    # ragged sizes around the block size
    orc = ob.OracleAccel(acc.tree, v)
    for n in (1, 31, 127, 128, 129, 1000):
        rays = _rays(n, n)
        _same_hits(acc.Intersect(rays), orc.intersect(rays))
    # empty scene: every ray misses and keeps its tMax (bvh.cpp:355)
    empty = gpu.B200Accel(dict(kind="bvh", nodes=np.zeros((0, 8), np.uint32), indices=np.zeros(0, np.uint32),
                               directions=None, bounds=np.zeros(6, np.float32)), np.zeros((0, 9), np.float32))
    rays = _rays(100, 1)
    h = empty.Intersect(rays)
    assert (h["prim"] == gpu.MISS).all() and np.array_equal(h["t"], rays[:, 3])
    assert not empty.IntersectP(rays).any()
    # malformed trees are rejected, not traversed
    bad = dict(acc.tree)
    bad["nodes"] = acc.tree["nodes"].copy()
    bad["nodes"][0, 6] = 10 ** 6
    with pytest.raises(gpu.B200RTError):
        gpu.B200Accel(bad, v)
    # MakeAccelerator param names of the .pbrt Accelerator line
    acc2 = gpu.MakeAccelerator("bvh", v, {"maxnodeprims": 1, "intersectcost": 8, "traversalcost": 1})
    rays = _rays(5000, 9)
    a, b = acc.Intersect(rays), acc2.Intersect(rays)
    assert np.array_equal(a["t"], b["t"])  # different tree, same closest hits


def test_killeroo_camera_and_random_rays(gpu):
    """Config 0 (scenes/killeroo-simple, BVH): camera rays + incoherent rays vs the oracle; kd-tree on
    the same rays must give the same hits (cross-accelerator agreement)."""
    from b200rt import scenes
    sc = scenes.killeroo_simple(xres=200, yres=200)
    v = sc["tri_verts"]
    assert len(v) == 66532
    bvh = gpu.MakeAccelerator("bvh", v)
    kd = gpu.MakeAccelerator("kdtree", v)
    rays = np.concatenate([scenes.camera_rays(sc["camera"]), scenes.random_rays(bvh.WorldBound(), 40000)])
    orc = ob.OracleAccel(bvh.tree, v)
    hb = bvh.Intersect(rays)
    _same_hits(hb, orc.intersect(rays, n_threads=8))
    assert (hb["prim"] != gpu.MISS).mean() > 0.3
    hk = kd.Intersect(rays)
    same = hk["prim"] == hb["prim"]
    # documented epsilon ties: two primitives with bit-equal t may be picked differently by different trees
    assert np.array_equal(hk["t"][same], hb["t"][same])
    assert (~same).mean() < 1e-4
    assert np.array_equal(hk["t"], hb["t"]) or (hk["t"] != hb["t"]).mean() < 1e-4
    # shadow rays towards the light
    m = hb["prim"] != gpu.MISS
    p = rays[m, 0:3] + hb["t"][m, None] * rays[m, 4:7]
    sh = np.zeros((m.sum(), 8), np.float32)
    sh[:, 0:3] = p + 1e-2 * (sc["light"] - p) / np.linalg.norm(sc["light"] - p, axis=1, keepdims=True)
    sh[:, 4:7] = sc["light"] - sh[:, 0:3]
    sh[:, 3] = 1 - 1e-4  # ShadowEpsilon, interaction.h:68-72
    assert np.array_equal(bvh.IntersectP(sh), orc.intersect_p(sh, n_threads=8))
    assert np.array_equal(kd.IntersectP(sh), bvh.IntersectP(sh))


def test_full_size_properties(gpu):
    """BASELINE config 1 size (sponza stand-in, ~262 k triangles, 2^22 rays): properties that need no
    oracle run — tMax monotonicity, any-hit == (closest-hit exists) on the same rays, re-tracing with
    tMax = t_hit*(1-eps) never returns the same primitive at a larger t, idempotence."""
    from b200rt import scenes
    sc = scenes.sponza_standin()
    v = sc["tri_verts"]
    acc = gpu.MakeAccelerator("bvh", v)
    cam = dict(sc["camera"], xres=1024, yres=1024)
    rays = np.concatenate([scenes.camera_rays(cam), scenes.random_rays(acc.WorldBound(), 1 << 21)])
    h = acc.Intersect(rays)
    hit = h["prim"] != gpu.MISS
    assert hit.mean() > 0.9                              # closed atrium: almost everything hits
    assert (h["t"][hit] > 0).all() and (h["t"][hit] <= rays[hit, 3]).all()
    assert np.array_equal(acc.IntersectP(rays).astype(bool), hit)
    h2 = acc.Intersect(rays)
    _same_hits(h, h2)                                    # deterministic / idempotent
    b = h["b0"][hit] + h["b1"][hit] + h["b2"][hit]
    assert np.abs(b - 1).max() < 1e-5
    # shrinking tMax below the hit makes that primitive unreachable
    r2 = rays[hit].copy()
    r2[:, 3] = h["t"][hit] * np.float32(1 - 1e-3)
    h3 = acc.Intersect(r2)
    assert ((h3["prim"] == gpu.MISS) | (h3["t"] < h["t"][hit])).all()
    # sampled oracle check at full scene size
    orc = ob.OracleAccel(acc.tree, v)
    idx = np.random.default_rng(0).choice(len(rays), 200000, replace=False)
    _same_hits(h[idx], orc.intersect(rays[idx], n_threads=8))


@pytest.mark.parametrize("kind", ["bvh", "kdtree"])
def test_degenerate_rays(gpu, kind):
    """Rays the reference never rejects up front: zero direction, NaN / infinite components, tMax of 0, negative or
    NaN.  Whatever the reference's float arithmetic makes of them (NaN compares false, 0 * inf = NaN ...), the
    kernels make the same: compared with the oracle restatement bit for bit."""
    from b200rt import scenes
    v = scenes.soup(1500, seed=8)
    acc = gpu.MakeAccelerator(kind, v)
    orc = ob.OracleAccel(acc.tree, v)
    rays = _rays(4096, 77)
    n = len(rays)
    nan, inf = np.float32(np.nan), np.float32(np.inf)
    rays[0:n:16, 4:7] = 0                         # zero direction
    rays[1:n:16, 4] = nan                         # NaN direction component
    rays[2:n:16, 0] = nan                         # NaN origin component
    rays[3:n:16, 3] = 0                           # tMax = 0
    rays[4:n:16, 3] = -1                          # negative tMax
    rays[5:n:16, 3] = nan                         # NaN tMax
    rays[6:n:16, 5] = inf                         # infinite direction component
    rays[7:n:16, 1] = -inf                        # infinite origin component
    rays[8:n:16, 4:7] *= np.float32(1e-30)        # denormal-scale direction
    rays[9:n:16, 4:6] = 0                         # axis-aligned along z
    h, o = acc.Intersect(rays), orc.intersect(rays)
    assert np.array_equal(h["prim"], o["prim"])
    for f in ("t", "b0", "b1", "b2"):
        # a NaN stays a NaN on both sides (x86 and the GPU differ only in the NaN's sign/payload bits: 0xffc00000 vs
        # 0x7fffffff); every other value is bit-identical
        nn = np.isnan(o[f])
        assert np.array_equal(np.isnan(h[f]), nn), f
        assert np.array_equal(h[f].view(np.uint32)[~nn], o[f].view(np.uint32)[~nn]), f
    assert np.isnan(o["t"]).sum() > 100   # the reference does report "hits" at t = NaN for such rays; so do we
    assert np.array_equal(acc.IntersectP(rays), orc.intersect_p(rays))


def test_more_than_2pow30_rays(gpu):
    """Maximum sizes: one call with more than 2^30 rays (the persistent kernels index rays with 32 bits, so the
    library splits the launch; b200rt.cu launch_trace).  2^30 + 777 device-resident rays = a 2^20-ray block tiled
    1024 times + a tail, so the expected answer is the block's answer tiled."""
    import torch
    from b200rt import scenes
    v = scenes.soup(20000, seed=2)
    acc = gpu.MakeAccelerator("bvh", v)
    block = scenes.random_rays(acc.WorldBound(), 1 << 20, seed=4)
    want = acc.IntersectP(block)
    assert 0.05 < want.mean() < 0.95
    n = (1 << 30) + 777
    free, _ = torch.cuda.mem_get_info()
    if free < 40 * (1 << 30):
        pytest.skip("needs 40 GB of free HBM")
    d_block = torch.from_numpy(block).cuda()
    rays = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    rays[: 1 << 30].view(1024, 1 << 20, 8).copy_(d_block.unsqueeze(0).expand(1024, -1, -1))
    rays[1 << 30:] = d_block[:777]
    occ = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
    acc.intersect_p_ptr(rays.data_ptr(), n, occ.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    d_want = torch.from_numpy(want).cuda()
    assert torch.equal(occ[: 1 << 30].view(1024, 1 << 20), d_want.unsqueeze(0).expand(1024, -1))
    assert torch.equal(occ[1 << 30:], d_want[:777])
    del rays, occ
    torch.cuda.empty_cache()
