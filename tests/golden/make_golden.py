"""Generates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref/libpbrtref.so,
built from /root/reference by oracle/ref/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

Outputs (tests/golden/):
  killeroo_mesh.npz      scenes/geometry/killeroo.pbrt: the control mesh (ctrl_*) and the mesh after the reference's own LoopSubdivide (P, idx, N)
                         (shapes/loopsubdiv.cpp), nlevels 1: P (float32), idx (uint16/uint32)
  ref_small_<kind>.npz   a 600-triangle soup, the reference-built tree (node arrays as exported from
                         the reference's memory), 4096 rays, and the reference's own Intersect /
                         IntersectP results + per-ray counters.  kinds: bvh, kdtree, rbsp3, rbsp7,
                         rbsp13, rbspkd13.
  ref_feet.npz           BASELINE config 4: the feet crop of scenes/killeroo-feet, reference-built BVH and
                         RBSP (3/7/9/13 directions) + RBSPKd(13) trees, the crop's camera rays + secondary
                         rays, and the reference's own answers per accelerator.
  ref_spheres.npz        300 triangles + 4 full spheres in ONE aggregate (BVH, kd-tree, RBSP-7 built by the
                         reference) and the reference's answers — pins Sphere::Intersect/IntersectP (a16).
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
    oP, oI, oN = rb.loopsubdiv(P, idx, lv, with_normals=True)
    print("killeroo: control %d verts / %d tris -> %d verts / %d tris" % (len(P), len(idx) // 3, len(oP), len(oI)))
    it = np.uint16 if len(oP) < 65536 else np.uint32
    np.savez_compressed(os.path.join(OUT, "killeroo_mesh.npz"), P=oP.astype(np.float32), idx=oI.astype(it), N=oN.astype(np.float32),
                        ctrl_P=P.astype(np.float32), ctrl_idx=idx.astype(np.uint16 if len(P) < 65536 else np.uint32), ctrl_levels=np.int32(lv))


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


SMALL_CASES = [("bvh", "bvh", {}), ("kdtree", "kdtree", {}), ("rbsp3", "rbsp", {"nbDirections": 3}),
               ("rbsp7", "rbsp", {"nbDirections": 7}), ("rbsp13", "rbsp", {"nbDirections": 13}),
               ("rbspkd13", "rbspkd", {"nbDirections": 13})]
# General-BSP families: Ize-et-al "paper" planes (deterministic) and k-means-of-normals
# clusters (seeded by std::random_device in the reference, F9 — the fixture pins the tree that was built).
BSP_CASES = [("bsppaper", "bsppaper", {}), ("bsppaperkd", "bsppaperkd", {}),
             ("bspcluster", "bspcluster", {"nbDirections": 5}), ("bspclusterfastkd", "bspclusterfastkd", {"nbDirections": 5})]


def gen_small(cases=SMALL_CASES):
    n = 600
    P = soup(n, 11)
    idx = np.arange(3 * n, dtype=np.int32).reshape(-1, 3)
    tv = P[idx].reshape(-1, 9)
    sc = rb.RefScene(P, idx)
    for name, kind, ip in cases:
        acc = rb.RefAccel(sc, kind, ip)
        ex = acc.export()
        kind = ex["kind"]
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


def gen_spheres():
    """Spheres inside the aggregate (scenes/killeroo-simple:19-25 puts its area-light
    sphere next to the triangles): a 300-triangle soup + 4 full spheres (pure Translate, one of them the
    killeroo-simple light's size), reference-built BVH / kd-tree / RBSP-7 over all 304 primitives, and
    the reference's answers.  Primitive order: triangles first, then the spheres."""
    n = 300
    P = soup(n, 17)
    idx = np.arange(3 * n, dtype=np.int32).reshape(-1, 3)
    tv = P[idx].reshape(-1, 9)
    cr = np.array([[1.5, 12.0, 2.0, 3.0], [-4, -3, 5, 1.0], [6, -7, -2, 2.5], [0.25, 0.5, -8, 0.125]], np.float32)
    sc = rb.RefScene(P, idx, spheres=cr)
    rays = rays_for(np.array([-10, -10, -10, 10, 10, 10], np.float32), 4096, 9)
    rng = np.random.default_rng(4)
    aim = np.zeros((2048, 8), np.float32)  # rays aimed at the spheres, from outside and from inside
    k = rng.integers(0, len(cr), len(aim))
    aim[:, 0:3] = cr[k, :3] + rng.normal(size=(len(aim), 3)) * cr[k, 3:4] * rng.choice([0.5, 4.0], size=(len(aim), 1))
    tgt = cr[k, :3] + rng.normal(size=(len(aim), 3)) * cr[k, 3:4] * 0.8
    aim[:, 4:7] = tgt - aim[:, 0:3]
    aim[:, 3] = np.inf
    aim[::5, 3] = rng.uniform(0.2, 1.5, len(aim[::5]))
    rays = np.concatenate([rays, aim]).astype(np.float32)
    out = dict(tri_verts=tv, spheres=cr, rays=rays)
    for name, kind, ip in [("bvh", "bvh", {}), ("kdtree", "kdtree", {}), ("rbsp7", "rbsp", {"nbDirections": 7})]:
        acc = rb.RefAccel(sc, kind, ip)
        ex = acc.export()
        hits, cnt = acc.intersect(rays, counters=True)
        occ = acc.intersect_p(rays)
        for f in ("nodes", "indices", "directions", "bounds"):
            out["%s_%s" % (name, f)] = ex[f]
        out["%s_prim" % name], out["%s_t" % name], out["%s_occ" % name] = hits["prim"], hits["t"], occ
        out["%s_tests" % name] = cnt["prim_tests"]
        print("  spheres/%-7s nodes %d hits %d (on spheres %d) occ %d" % (name, len(ex["nodes"]), int((hits["prim"] >= 0).sum()),
                                                                        int((hits["prim"] >= n).sum()), int(occ.sum())))
    np.savez_compressed(os.path.join(OUT, "ref_spheres.npz"), **out)


def gen_feet():
    """BASELINE config 4 (scenes/killeroo-feet + scripts/run_rbsps.sh): the feet crop of one killeroo,
    RBSP with nbDirections 3/7/9/13 (+ RBSPKd 13) against the BVH on the SAME rays.  To keep the
    fixture small only the triangles seen through the crop window (plus a margin) are kept."""
    sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
    from b200rt import scenes
    sc = scenes.killeroo_feet()
    tv_all, cam = sc["tri_verts"], sc["camera"]
    n_all = len(tv_all)
    full = rb.RefScene(tv_all.reshape(-1, 3), np.arange(3 * n_all, dtype=np.int32).reshape(-1, 3))
    bvh_full = rb.RefAccel(full, "bvh")
    x0, x1, y0, y1 = sc["crop"]
    probe = scenes.camera_rays(dict(cam, xres=cam["xres"] // 8, yres=cam["yres"] // 8), x0=x0 // 8, x1=x1 // 8, y0=y0 // 8, y1=y1 // 8)
    h = bvh_full.intersect(probe, n_threads=8)
    pts = h["p"][h["prim"] >= 0]
    c = tv_all.reshape(-1, 3, 3).mean(1)
    keep = np.sort(np.argsort(np.linalg.norm(c - pts.mean(0), axis=1))[:4096])  # the 4096 triangles around the feet
    tv = np.ascontiguousarray(tv_all[keep])
    n = len(tv)
    print("killeroo-feet: %d of %d triangles kept, %d/%d probe rays hit" % (n, n_all, len(pts), len(probe)))
    P, idx = tv.reshape(-1, 3), np.arange(3 * n, dtype=np.int32).reshape(-1, 3)
    rsc = rb.RefScene(P, idx)
    # rays: the crop window's camera rays on a coarse grid + path-like rays leaving the surface
    rng = np.random.default_rng(21)
    hh = rb.RefAccel(rsc, "bvh").intersect(probe, n_threads=8)
    m = hh["prim"] >= 0
    cam_r = np.concatenate([probe[m], probe[~m][:: max(1, (~m).sum() // 1024)]])  # every hitting ray + ~1024 missing ones
    o = hh["p"][m][:4096]
    d = rng.normal(size=o.shape)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.zeros((len(o), 8), np.float32)
    sec[:, 0:3], sec[:, 4:7], sec[:, 3] = o + 1e-3 * d, d, np.inf
    sec[::3, 3] = rng.uniform(1, 60, len(sec[::3]))
    rays = np.concatenate([cam_r, sec]).astype(np.float32)
    out = dict(tri_verts=tv, rays=rays)
    import time
    for name, kind, ip in [("bvh", "bvh", {}), ("rbsp3", "rbsp", {"nbDirections": 3}), ("rbsp7", "rbsp", {"nbDirections": 7}),
                           ("rbsp9", "rbsp", {"nbDirections": 9}), ("rbsp13", "rbsp", {"nbDirections": 13}),
                           ("rbspkd13", "rbspkd", {"nbDirections": 13})]:
        t0 = time.time()
        acc = rb.RefAccel(rsc, kind, ip)
        tb = time.time() - t0
        ex = acc.export()
        hits = acc.intersect(rays, n_threads=8)
        occ = acc.intersect_p(rays, n_threads=8)
        for k in ("nodes", "indices", "directions", "bounds"):
            out["%s_%s" % (name, k)] = ex[k]
        out["%s_prim" % name], out["%s_t" % name], out["%s_occ" % name] = hits["prim"], hits["t"], occ
        print("  %-9s build %.1fs nodes %d indices %d hits %d occ %d" % (name, tb, len(ex["nodes"]), len(ex["indices"]),
                                                                      int((hits["prim"] >= 0).sum()), int(occ.sum())))
    np.savez_compressed(os.path.join(OUT, "ref_feet.npz"), **out)
    print("ref_feet.npz: %.1f MB" % (os.path.getsize(os.path.join(OUT, "ref_feet.npz")) / 1e6))


if __name__ == "__main__":
    which = sys.argv[1:] or ["killeroo", "small", "triangle", "feet", "spheres"]
    if "spheres" in which: gen_spheres()
    if "killeroo" in which: gen_killeroo()
    if "small" in which: gen_small()
    if "bsp" in which: gen_small(BSP_CASES)
    if "triangle" in which: gen_triangle()
    if "feet" in which: gen_feet()
