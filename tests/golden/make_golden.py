"""Generates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref/libpbrtref.so,
built from /root/reference by oracle/ref/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

Outputs (tests/golden/):
  killeroo_mesh.npz      scenes/geometry/killeroo.pbrt after the reference's own LoopSubdivide
                         (shapes/loopsubdiv.cpp), nlevels 1: P (float32), idx (uint16/uint32)
  ref_small_<kind>.npz   a 600-triangle soup, the reference-built tree (node arrays as exported from
                         the reference's memory), 4096 rays, and the reference's own Intersect /
                         IntersectP results + per-ray counters.  kinds: bvh, kdtree, rbsp3, rbsp7,
                         rbsp13, rbspkd13.
  ref_triangle.npz       Triangle.Watertight / Reintersect / BadCases style single-triangle cases
                         (reference src/tests/shapes.cpp:28-205,544-559) with the reference's answers.
"""
import os
import re
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind as rb  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def parse_killeroo():
    txt = open(os.path.join(REF, "scenes/geometry/killeroo.pbrt")).read()
    idx = re.search(r'"integer indices"\s*\[([^\]]*)\]', txt).group(1)
    P = re.search(r'"point P"\s*\[([^\]]*)\]', txt).group(1)
    lv = int(re.search(r'"integer nlevels"\s*\[\s*(\d+)', txt).group(1))
    return np.array(P.split(), np.float32).reshape(-1, 3), np.array(idx.split(), np.int32), lv


def gen_killeroo():
    P, idx, lv = parse_killeroo()
    oP, oI = rb.loopsubdiv(P, idx, lv)
    print("killeroo: control %d verts / %d tris -> %d verts / %d tris" % (len(P), len(idx) // 3, len(oP), len(oI)))
    it = np.uint16 if len(oP) < 65536 else np.uint32
    np.savez_compressed(os.path.join(OUT, "killeroo_mesh.npz"), P=oP.astype(np.float32), idx=oI.astype(it))


def soup(n, seed):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-10, 10, (n, 1, 3))
    return ((c + rng.normal(0, 0.8, (n, 3, 3))).reshape(-1, 3).astype(np.float32) + np.float32(0))


def rays_for(bounds, n, seed):
    rng = np.random.default_rng(seed)
    lo, hi = bounds[:3] - 1, bounds[3:] + 1
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = lo + rng.random((n, 3)) * (hi - lo)
    d = rng.normal(size=(n, 3))
    r[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    r[:, 3] = np.inf
    r[1::4, 3] = rng.uniform(0.5, 12, len(r[1::4]))      # finite tMax
    rows = np.arange(2, n, 16)
    r[rows, 4 + (np.arange(len(rows)) % 3)] = 0.0        # axis-parallel components (1/0 = inf paths)
    r[3::32, 4:7] *= 7.5                                   # unnormalised directions (shadow rays)
    return r


def gen_small():
    n = 600
    P = soup(n, 11)
    idx = np.arange(3 * n, dtype=np.int32).reshape(-1, 3)
    tv = P[idx].reshape(-1, 9)
    sc = rb.RefScene(P, idx)
    for name, kind, ip in [("bvh", "bvh", {}), ("kdtree", "kdtree", {}), ("rbsp3", "rbsp", {"nbDirections": 3}),
                           ("rbsp7", "rbsp", {"nbDirections": 7}), ("rbsp13", "rbsp", {"nbDirections": 13}),
                           ("rbspkd13", "rbspkd", {"nbDirections": 13})]:
        acc = rb.RefAccel(sc, kind, ip)
        ex = acc.export()
        rays = rays_for(np.array([-10, -10, -10, 10, 10, 10], np.float32), 4096, 5)
        hits, cnt = acc.intersect(rays, counters=True)
        occ, cntp = acc.intersect_p(rays, counters=True)
        np.savez_compressed(os.path.join(OUT, "ref_small_%s.npz" % name), kind=kind, tri_verts=tv, nodes=ex["nodes"],
                            indices=ex["indices"], directions=ex["directions"], bounds=ex["bounds"], rays=rays,
                            hit_prim=hits["prim"], hit_t=hits["t"], hit_uv=hits["uv"], hit_p=hits["p"], occluded=occ,
                            cnt_tests=cnt["prim_tests"], cnt_interior=cnt["interior_visits"],
                            cnt_leaf=cnt["leaf_visits"], cntp_tests=cntp["prim_tests"],
                            cntp_interior=cntp["interior_visits"], cntp_leaf=cntp["leaf_visits"])
        print(name, ex["nodes"].shape, "hits", int((hits["prim"] >= 0).sum()), "occ", int(occ.sum()))


def gen_triangle():
    """Single-triangle cases in the spirit of the reference's own triangle tests: rays through
    vertices and edges (watertightness), rays re-spawned near a hit, huge/small coordinates, the
    BadCases known answer, and degenerate triangles."""
    rng = np.random.default_rng(3)
    tris, rays = [], []
    for i in range(400):
        e = rng.integers(-8, 9)
        v = (rng.normal(size=(3, 3)) * 10.0 ** e).astype(np.float32)
        o = (rng.normal(size=3) * 10.0 ** e * 3).astype(np.float32)
        for j in range(8):
            if j < 3:
                tgt = v[j].astype(np.float64)                      # exactly at a vertex
            elif j < 6:
                a = rng.random()
                tgt = a * v[j - 3].astype(np.float64) + (1 - a) * v[(j - 2) % 3].astype(np.float64)  # on an edge
            else:
                b = rng.dirichlet((1, 1, 1))
                tgt = b @ v.astype(np.float64)
            d = (tgt - o).astype(np.float32)
            tris.append(v.reshape(9))
            rays.append(np.concatenate([o, [np.inf], d, [0]]).astype(np.float32))
    # Triangle.BadCases (src/tests/shapes.cpp:544-559)
    v = np.array([-1113.45459, -79.049614, -56.2431908, -1113.45459, -87.0922699, -56.2431908,
                  -1113.45459, -79.2090149, -56.2431908], np.float32)
    r = np.array([-1081.47925, 99.9999542, 87.7701111, 0.9999, -32.1072998, -183.355865, -144.607635, 0.9999],
                 np.float32)
    r[7] = 0
    tris.append(v)
    rays.append(r)
    # exactly degenerate triangles
    for _ in range(20):
        a, b = rng.normal(size=3), rng.normal(size=3)
        v = np.stack([a, a + b, a + 2 * b]).astype(np.float32)
        d = (v[1] - np.array([0, 0, 5], np.float32))
        tris.append(v.reshape(9))
        rays.append(np.concatenate([[0, 0, 5], [np.inf], d, [0]]).astype(np.float32))
    tris, rays = np.array(tris, np.float32) + np.float32(0), np.array(rays, np.float32)
    n = len(tris)
    sc = rb.RefScene(tris.reshape(-1, 3), np.arange(3 * n, dtype=np.int32).reshape(-1, 3))
    hit = np.zeros(n, np.int32)
    t = np.zeros(n, np.float32)
    uv = np.zeros((n, 2), np.float32)
    occ = np.zeros(n, np.uint8)
    for i in range(n):
        h = sc.prim_intersect(i, rays[i:i + 1])
        hit[i], t[i], uv[i] = h["prim"][0] >= 0, h["t"][0], h["uv"][0]
        occ[i] = sc.prim_intersect_p(i, rays[i:i + 1])[0]
    # Intersect rejects exactly-degenerate triangles in its tail (triangle.cpp:309-315); IntersectP has that
    # test only under an alpha mask (:533-556), so the two may differ on degenerate triangles.
    print("Intersect != IntersectP on", int((hit.astype(np.uint8) != occ).sum()), "cases")
    print("triangle cases:", n, "hits", int(hit.sum()))
    np.savez_compressed(os.path.join(OUT, "ref_triangle.npz"), tris=tris, rays=rays, hit=hit, t=t, uv=uv, occ=occ)


if __name__ == "__main__":
    gen_killeroo()
    gen_small()
    gen_triangle()
