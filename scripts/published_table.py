"""The reference's ONE published table (`killerroo-simple-times` at its repo root): scenes/killeroo-simple
(700x700, 8 spp, path maxdepth 5) rendered with kd / rbsp-3 / rbsp-7 / rbsp-9 / rbsp-13 — node traversals,
ray-triangle intersections, build time, render time — re-run on the B200 with the scene as written (sphere area light,
matte ground, two plastic killeroos; trees from the reference's own builders).

    python scripts/published_table.py --build [names...]   # HERE (needs /root/reference) -> oracle/_ref/ks_trees.npz
    python scripts/published_table.py                      # on the GPU box

The counters come from the instrumented kernels (b200rt_render_heatmaps = the thesis's per-pixel Ray::stats, bit-exact
per ray); they agree with the published ones statistically only (the table was made with the halton sampler, this
integrator draws counter-hash samples).  Reporting aid, not the bench contract.
"""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
TREES = os.path.join(ROOT, "oracle", "_ref", "ks_trees.npz")
CASES = [("bvh", "bvh", {}), ("kdtree", "kdtree", {}), ("rbsp3", "rbsp", {"nbDirections": 3}),
         ("rbsp7", "rbsp", {"nbDirections": 7}), ("rbsp9", "rbsp", {"nbDirections": 9}),
         ("rbsp13", "rbsp", {"nbDirections": 13})]
# /root/reference/killerroo-simple-times:2-5: closest-hit node traversals, ray-triangle intersections,
# build s, render s
PUBLISHED = {"kdtree": (449597065, 128156919, 2, 8), "rbsp3": (449614985, 128146776, 38, 27),
             "rbsp7": (536362041, 149309427, 126, 61), "rbsp9": (508118176, 126294832, 190, 79),
             "rbsp13": (509277939, 123089940, 319, 81)}


def build():
    from oracle import refbind as rb
    from b200rt import scenes
    sc = scenes.killeroo_simple()
    tv = sc["tri_verts"]
    sl = sc["sphere_light"]
    S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3),
                    spheres=np.array([list(sl["center"]) + [sl["radius"]]], np.float32))
    out = dict(np.load(TREES)) if os.path.exists(TREES) else {}
    only = [a for a in sys.argv[1:] if not a.startswith("--")]
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
        print("%-8s reference build %.1f s, %d nodes, %d indices" % (name, tb, len(ex["nodes"]), len(ex["indices"])), flush=True)
        np.savez(TREES, **out)


def run():
    import torch
    import b200rt
    from b200rt import scenes, wavefront as wf
    z = np.load(TREES)
    b200rt.init(0)
    torch.zeros(1, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sc = scenes.killeroo_simple()
    verts, prim_kind, rows, pm, mats = scenes.with_sphere_light(sc, plastic=True)
    sl = sc["sphere_light"]
    integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=8, maxdepth=5, light_pos=sl["center"], light_I=sl["L"], seed=1,
                                           sphere_light=(sl["radius"], len(verts) - 1), materials=mats, prim_material=pm)
    film = torch.zeros(integ.shape, dtype=torch.float32, device="cuda")
    print("scenes/killeroo-simple as written: %d primitives, 700x700 @ 8 spp, maxdepth 5, one B200" % len(verts))
    print("%-8s | %13s %13s %6s | %13s %13s %6s | %9s %9s | %9s %9s" %
          ("accel", "node trav.", "published", "ratio", "ray-tri", "published", "ratio", "build s", "(publ.)", "render ms", "(publ. s)"))
    for name, _, _ in CASES:
        if name + "_nodes" not in z.files:
            continue
        tree = dict(kind=str(z[name + "_kind"]), nodes=z[name + "_nodes"], indices=z[name + "_indices"],
                    directions=z[name + "_directions"], bounds=z[name + "_bounds"], n_prims=len(verts))
        acc = b200rt.B200Accel(tree, verts, prim_kind=prim_kind, spheres=rows)
        _, heat, _ = wf.render_heatmaps(integ, acc)
        nodes, tests = int(heat[2].astype(np.uint64).sum()), int(heat[0].astype(np.uint64).sum())
        best = 1e9
        for _ in range(4):
            film.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); integ.render_ptr(acc, film.data_ptr(), st); b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        pub = PUBLISHED.get(name)
        if pub:
            print("%-8s | %13d %13d %6.3f | %13d %13d %6.3f | %9.1f %9d | %9.2f %9d" %
                  (name, nodes, pub[0], nodes / pub[0], tests, pub[1], tests / pub[1], float(z[name + "_build_s"]), pub[2], best, pub[3]), flush=True)
        else:
            print("%-8s | %13d %13s %6s | %13d %13s %6s | %9.1f %9s | %9.2f %9s" %
                  (name, nodes, "-", "-", tests, "-", "-", float(z[name + "_build_s"]), "-", best, "-"), flush=True)


if __name__ == "__main__":
    build() if "--build" in sys.argv else run()
