"""CPU: the oracle (oracle/oracle.c) against the golden vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py), and the host builders against the reference-built trees."""
import numpy as np
import pytest
from conftest import load_golden, golden_tree, SMALL, FEET, feet_tree
from oracle import orcbind as ob


@pytest.mark.parametrize("name", SMALL)
def test_oracle_matches_reference_golden(name):
    g = load_golden("ref_small_%s.npz" % name)
    acc = ob.OracleAccel(golden_tree(g), g["tri_verts"])
    hits, per = acc.intersect(g["rays"], per_ray=True)
    ref_prim = np.where(g["hit_prim"] < 0, ob.MISS, g["hit_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)                      # primitive ids: exact
    assert np.array_equal(hits["t"].view(np.uint32), g["hit_t"].view(np.uint32))  # t: bit-exact
    m = g["hit_prim"] >= 0
    # barycentrics through SurfaceInteraction::uv with default UVs (triangle.cpp:326): uv = (b1 + b2, b2)
    assert np.array_equal(hits["b2"][m], g["hit_uv"][m, 1])
    assert np.array_equal((hits["b1"] + hits["b2"])[m], g["hit_uv"][m, 0])
    # pHit = b0*p0 + b1*p1 + b2*p2 (triangle.cpp:325)
    v = g["tri_verts"].reshape(-1, 3, 3)[hits["prim"][m]]
    p = (hits["b0"][m, None] * v[:, 0] + hits["b1"][m, None] * v[:, 1]) + hits["b2"][m, None] * v[:, 2]
    assert np.array_equal(p.astype(np.float32), g["hit_p"][m])
    occ, perp = acc.intersect_p(g["rays"], per_ray=True)
    assert np.array_equal(occ, g["occluded"])
    # per-ray traversal counters = the reference's Ray::stats
    assert np.array_equal(per["prim_tests"], g["cnt_tests"])
    assert np.array_equal(per["leaf_visits"], g["cnt_leaf"])
    interior = per["nodes_passed"] - (per["leaf_visits"] if name == "bvh" else 0)
    assert np.array_equal(interior, g["cnt_interior"])
    assert np.array_equal(perp["prim_tests"], g["cntp_tests"])
    assert np.array_equal(perp["leaf_visits"], g["cntp_leaf"])


def test_oracle_triangle_cases():
    """Watertight / Reintersect / BadCases-style single-triangle cases (reference src/tests/shapes.cpp)."""
    g = load_golden("ref_triangle.npz")
    flags = ob.degenerate_flags(g["tris"])
    n_hit = 0
    for i in range(len(g["tris"])):
        hit, t, b0, b1, b2 = ob.triangle(g["rays"][i], g["tris"][i], flags[i])
        assert hit == bool(g["hit"][i]), i
        hitp = ob.triangle(g["rays"][i], g["tris"][i], 0)[0]
        assert hitp == bool(g["occ"][i]), i
        if hit:
            n_hit += 1
            assert np.float32(t) == g["t"][i]
            assert np.float32(b2) == g["uv"][i, 1]
            assert np.float32(b1) + np.float32(b2) == g["uv"][i, 0]
    assert n_hit > 1000
    # Triangle.BadCases: known miss
    assert not g["hit"][3200]


def test_builders_reproduce_reference_trees(built_lib):
    """b200rt_build_bvh / b200rt_build_kdtree vs the trees the reference built (node for node)."""
    g = load_golden("ref_small_bvh.npz")
    t = built_lib.build_bvh(g["tri_verts"])
    assert np.array_equal(t["nodes"], g["nodes"])
    assert np.array_equal(t["indices"], g["indices"])
    g = load_golden("ref_small_kdtree.npz")
    t = built_lib.build_kdtree(g["tri_verts"])
    assert np.array_equal(t["nodes"], g["nodes"])
    assert np.array_equal(t["indices"], g["indices"])
    assert np.array_equal(t["bounds"], g["bounds"])


def test_rbsp3_equals_kdtree_hits():
    """Cross-accelerator agreement (how the thesis itself validated kd/RBSP)."""
    a = load_golden("ref_small_kdtree.npz")
    for other in ("bvh", "rbsp3", "rbsp7", "rbsp13", "rbspkd13"):
        b = load_golden("ref_small_%s.npz" % other)
        assert np.array_equal(a["hit_prim"], b["hit_prim"])
        assert np.array_equal(a["hit_t"], b["hit_t"])
        assert np.array_equal(a["occluded"], b["occluded"])


@pytest.mark.parametrize("name", FEET)
def test_oracle_matches_reference_feet(name):
    """BASELINE config 4 (scenes/killeroo-feet, scripts/run_rbsps.sh): RBSP 3/7/9/13 + RBSPKd and the BVH
    on the same rays, trees and answers from the unmodified reference."""
    g = load_golden("ref_feet.npz")
    acc = ob.OracleAccel(feet_tree(g, name), g["tri_verts"])
    hits = acc.intersect(g["rays"], n_threads=4)
    ref_prim = np.where(g[name + "_prim"] < 0, ob.MISS, g[name + "_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    assert np.array_equal(hits["t"].view(np.uint32), g[name + "_t"].view(np.uint32))
    assert np.array_equal(acc.intersect_p(g["rays"], n_threads=4), g[name + "_occ"])
    # cross-accelerator agreement on identical rays: same t as the BVH
    assert np.array_equal(g[name + "_t"], g["bvh_t"])


@pytest.mark.parametrize("name", ["bvh", "kdtree", "rbsp7"])
def test_oracle_matches_reference_spheres(name):
    """Sphere::Intersect / IntersectP inside reference-built aggregates."""
    from conftest import spheres_scene
    g = load_golden("ref_spheres.npz")
    tree, verts, kind, rows = spheres_scene(g, name)
    acc = ob.OracleAccel(tree, verts, prim_kind=kind, spheres=rows)
    hits, per = acc.intersect(g["rays"], per_ray=True)
    ref_prim = np.where(g[name + "_prim"] < 0, ob.MISS, g[name + "_prim"]).astype(np.uint32)
    assert np.array_equal(hits["prim"], ref_prim)
    assert np.array_equal(hits["t"].view(np.uint32), g[name + "_t"].view(np.uint32))
    assert (hits["prim"][hits["prim"] != ob.MISS] >= len(g["tri_verts"])).sum() > 1000   # plenty of sphere hits
    assert np.array_equal(acc.intersect_p(g["rays"]), g[name + "_occ"])
    assert np.array_equal(per["prim_tests"], g[name + "_tests"])
