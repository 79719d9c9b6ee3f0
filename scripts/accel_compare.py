"""BASELINE config 4 (scenes/killeroo-feet + scripts/run_rbsps.sh of the reference): the thesis's accelerator
comparison — BVH vs kd-tree vs RBSP with 3/7/9/13 directions (+ RBSPKd, BSPPaper) ON THE SAME RAYS — on the B200.

    python scripts/accel_compare.py --build     # HERE (needs /root/reference): reference builders -> oracle/_ref/feet_trees.npz
    python scripts/accel_compare.py             # on the GPU box: Mrays/s + the thesis's per-ray statistics per accelerator
    python scripts/accel_compare.py --device-build   # same, but BVH / kd-tree / RBSP / RBSPKd trees are BUILT on the B200 (round 2:
                                                     # csrc/rbsp_build.cuh, bvh_build.cuh) and the build time is printed; the BSPPaper
                                                     # families still come from the tree file when it is there

The tree file is git-ignored (oracle/_ref/) but travels to the GPU box with the snapshot.  The full killeroo of the
scene (33 264 triangles) is used; rays = the crop window's camera rays (1000 x 2000 px of the 5000^2 film) and the
diffuse bounce + shadow rays leaving the first hits.  Tuning / reporting aid (test infrastructure: it drives the
reference's builders through oracle/refbind.py); not the bench contract.
"""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
TREES = os.path.join(ROOT, "oracle", "_ref", "feet_trees.npz")
CASES = [("bvh", "bvh", {}), ("kdtree", "kdtree", {}), ("rbsp3", "rbsp", {"nbDirections": 3}),
         ("rbsp7", "rbsp", {"nbDirections": 7}), ("rbsp9", "rbsp", {"nbDirections": 9}),
         ("rbsp13", "rbsp", {"nbDirections": 13}), ("rbspkd13", "rbspkd", {"nbDirections": 13}),
         ("bsppaper", "bsppaper", {}), ("bsppaperkd", "bsppaperkd", {})]


def build():
    from oracle import refbind as rb
    from b200rt import scenes
    sc = scenes.killeroo_feet()
    tv = sc["tri_verts"]
    S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
    out = dict(tri_verts=tv)
    only = [a for a in sys.argv[1:] if not a.startswith("--")]
    if os.path.exists(TREES):
        out.update(dict(np.load(TREES)))
    for name, kind, ip in CASES:
        if (only and name not in only) or (not only and name + "_nodes" in out):
            continue
        t0 = time.time()
        acc = rb.RefAccel(S, kind, ip)
        tb = time.time() - t0
        ex = acc.export()
        for k in ("nodes", "indices", "directions", "bounds"):
            out["%s_%s" % (name, k)] = ex[k]
        out[name + "_kind"] = np.array(ex["kind"])
        out[name + "_build_s"] = np.array(tb)
        print("%-10s reference build %.1f s, %d nodes, %d indices" % (name, tb, len(ex["nodes"]), len(ex["indices"])), flush=True)
        np.savez(TREES, **out)


def timeit(fn, sync, reps=5):
    fn(); sync()
    import torch
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); sync()
        best = min(best, a.elapsed_time(b))
    return best


def run():
    import torch
    import b200rt
    from b200rt import scenes
    device_build = "--device-build" in sys.argv
    z = np.load(TREES) if os.path.exists(TREES) else None
    tv = z["tri_verts"] if z is not None else scenes.killeroo_feet()["tri_verts"]
    b200rt.init(0)
    b200rt.build_bvh_device(tv[:64])   # load the module before timing anything
    torch.zeros(1, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sc = scenes.killeroo_feet()
    x0, x1, y0, y1 = sc["crop"]
    cam = scenes.camera_rays(sc["camera"], x0=x0, x1=x1, y0=y0, y1=y1)   # the crop window at the scene's resolution
    names = [n for n, k, _ in CASES if (device_build and k in ("bvh", "kdtree", "rbsp", "rbspkd")) or (z is not None and n + "_nodes" in z.files)]
    ref = None
    accs = {}
    print("killeroo-feet: %d triangles, %d camera rays (crop %s of 5000^2)" % (len(tv), len(cam), (x0, x1, y0, y1)))
    print("%-10s %9s | %-28s | %-28s | %-28s" % ("accel", "nodes", "camera closest", "bounce closest", "shadow any-hit"))
    print("%-10s %9s | %8s %8s %9s | %8s %8s %9s | %8s %8s %9s" % (("", "") + ("Mrays/s", "nodes/r", "tests/r") * 3))
    builds, nodes_differ = {}, {}
    for name in names:
        kind, ip = [(k, p) for n_, k, p in CASES if n_ == name][0]
        if device_build and kind in ("bvh", "kdtree", "rbsp", "rbspkd"):
            tree = {"bvh": b200rt.build_bvh_device, "kdtree": b200rt.build_kdtree_device}[kind](tv) if kind in ("bvh", "kdtree") else \
                b200rt.build_rbsp(tv, nbDirections=ip["nbDirections"], kdtraversalcost=1 if kind == "rbspkd" else None)
            builds[name] = tree["build"]["build_ms"]
            if z is not None and name + "_nodes" in z.files:   # the reference's own tree of the same scene: same node array
                # same node array; the word that may differ is the primitive stored in a single-primitive leaf whose membership
                # the reference decides by std::sort's order among equal keys (tests/test_rbsp_builder.py)
                ne = (tree["nodes"] != z[name + "_nodes"])
                assert not ne[:, 1].any() and ne[:, 0].sum() <= 1e-5 * len(ne), (name, int(ne[:, 0].sum()))
                nodes_differ[name] = int(ne[:, 0].sum())
                builds[name] = (tree["build"]["build_ms"], float(z[name + "_build_s"]))
        else:
            tree = dict(kind=str(z[name + "_kind"]), nodes=z[name + "_nodes"], indices=z[name + "_indices"],
                        directions=z[name + "_directions"], bounds=z[name + "_bounds"], n_prims=len(tv))
        acc = b200rt.B200Accel(tree, tv)
        hits = acc.Intersect(cam)
        if ref is None:
            ref = hits
            m = hits["prim"] != b200rt.MISS
            rng = np.random.default_rng(3)
            p = cam[m, 0:3] + hits["t"][m, None] * cam[m, 4:7]
            v = tv.reshape(-1, 3, 3)[hits["prim"][m]]
            n = np.cross(v[:, 0] - v[:, 2], v[:, 1] - v[:, 2])
            n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)
            n = np.where((np.sum(n * -cam[m, 4:7], 1) < 0)[:, None], -n, n)
            d = rng.normal(size=p.shape)
            d /= np.linalg.norm(d, axis=1, keepdims=True)
            d = np.where((np.sum(d * n, 1) < 0)[:, None], -d, d)
            bounce = np.zeros((len(p), 8), np.float32)
            bounce[:, 0:3], bounce[:, 4:7], bounce[:, 3] = p + 1e-2 * n, d, np.inf
            shadow = bounce.copy()   # distant light along the view direction (scene's LightSource "distant")
            L = sc["light_dir"] / np.linalg.norm(sc["light_dir"])
            shadow[:, 4:7], shadow[:, 3] = L * 1000.0, 1 - 1e-4
        else:   # every accelerator answers the same rays identically (hit prim up to exact-t ties, t bit for bit)
            assert np.array_equal(hits["t"].view(np.uint32), ref["t"].view(np.uint32)), name
            assert (hits["prim"] != ref["prim"]).mean() < 1e-5, name
        row = "%-10s %9d |" % (name, len(tree["nodes"]))
        for rays, anyhit in ((cam, False), (bounce, False), (shadow, True)):
            dr = torch.from_numpy(rays).cuda()
            n = len(rays)
            out = torch.zeros(n * 5, dtype=torch.int32, device="cuda") if not anyhit else torch.zeros(n, dtype=torch.uint8, device="cuda")
            fn = (lambda: acc.intersect_p_ptr(dr.data_ptr(), n, out.data_ptr(), st)) if anyhit else \
                 (lambda: acc.intersect_ptr(dr.data_ptr(), n, out.data_ptr(), st))
            ms = timeit(fn, torch.cuda.synchronize)
            _, rs = acc.IntersectStats(rays[:: max(1, n // 200000)], any_hit=anyhit)
            row += " %8.0f %8.2f %9.2f |" % (n / ms / 1e3, (rs["interior_visits"] + rs["leaf_visits"]).mean(), rs["prim_tests"].mean())
        print(row, flush=True)
        accs[name] = acc
    if builds:
        print("\ntree builds on the B200 (ms)" + (" | the reference's builder on this container's CPU (s), node arrays asserted equal" if z is not None else "") + ":")
        for name, b in builds.items():
            print("%-10s %s" % (name, ("%9.1f ms | %7.1f s | node words differing: %d" % (b + (nodes_differ.get(name, 0),))) if isinstance(b, tuple) else "%9.1f ms" % b))
    # ---- the thesis's experiment itself: the scene file rendered per accelerator (scripts/run_rbsps.sh of the reference:
    # 5000^2 film, cropwindow [.3 .5 .4 .8], 8 spp, path maxdepth 5, distant light, plastic killeroo) ----
    from b200rt import wavefront as wf
    print("\nscenes/killeroo-feet as written (crop %d x %d px @ 8 spp, maxdepth 5), whole frame per accelerator:" % (x1 - x0, y1 - y0))
    print("%-10s %10s %12s %12s %10s %12s" % ("accel", "frame ms", "closest rays", "shadow rays", "Mrays/s", "film mean"))
    film = torch.zeros((y1 - y0, x1 - x0, 4), dtype=torch.float32, device="cuda")
    for name in names:
        integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=8, maxdepth=5, light_I=(1, 1, 1), seed=1, crop=sc["crop"],
                                               distant_light=(sc["light_from"] - sc["light_to"], wf.world_radius(tv)),
                                               materials=sc["materials"], prim_material=np.zeros(len(tv), np.uint8))
        stats = {}

        def go():
            film.zero_()
            stats.update(integ.render_ptr(accs[name], film.data_ptr(), st))
        ms = timeit(go, torch.cuda.synchronize, reps=3)
        f = film.cpu().numpy()
        rays = stats["closest_rays"] + stats["shadow_rays"]
        print("%-10s %10.2f %12d %12d %10.0f %12.6f" % (name, ms, stats["closest_rays"], stats["shadow_rays"], rays / ms / 1e3,
                                                         (f[..., :3] / np.maximum(f[..., 3:4], 1)).mean()), flush=True)


if __name__ == "__main__":
    build() if "--build" in sys.argv else run()
