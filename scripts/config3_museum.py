"""BASELINE config 3 on the B200: the thesis's kd-tree scene variants (`_k`: scenes/test_museum_local_k = scenes/museum
with `Accelerator "kdtree"`) on the museum STAND-IN (b200rt/scenes.py::museum_standin — the museum parts are absent from
the reference tree), rendered as the template says (path maxdepth 1, 64 spp, 700x700, point light) once per
accelerator, plus the kd traversal kernels alone on the frame's own ray queues.

    python scripts/config3_museum.py [--scene museum|conference] [--spp 64] [--res 700] [--cpu]

Prints one JSON line per accelerator: frame time (CUDA events), rays, whole-frame and per-kernel Mrays/s, film mean; the
kd-tree and BVH films must agree (cross-accelerator agreement is how the reference itself pins its kd-tree).
--cpu also times the reference's own KdTreeAccel render of the same scene on the host cores (oracle/_ref, bounded spp)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--res", type=int, default=700)
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--scene", default="museum", choices=["museum", "conference"])
    args = ap.parse_args()
    import numpy as np
    import b200rt
    from b200rt import scenes
    b200rt.init(0)
    sc = (scenes.museum_standin if args.scene == "museum" else scenes.conference_standin)(xres=args.res, yres=args.res)
    from b200rt import wavefront as wf
    wf.set_streams(1)   # per-kernel times below: one wavefront batch in flight
    tv = sc["tri_verts"]
    films = {}
    for accel in ("kdtree", "bvh"):
        acc = b200rt.MakeAccelerator(accel, tv)
        integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=args.spp, maxdepth=sc["maxdepth"], light_pos=sc["light"],
                                               light_I=(1, 1, 1), kd=(.5, .5, .5), seed=1)
        integ.Render(acc)   # warm-up
        best = None
        for _ in range(args.frames):
            film, st = integ.Render(acc)
            if best is None or st["ms_total"] < best["ms_total"]:
                best = st
        films[accel] = film
        rays = best["closest_rays"] + best["shadow_rays"]
        print(json.dumps({
            "workload": "%s stand-in (%d triangles, %s, %d nodes), %dx%d @ %d spp, path maxdepth %d, point light, matte"
                        % (args.scene, len(tv), accel, len(acc.tree["nodes"]), args.res, args.res, args.spp, sc["maxdepth"]),
            "ms_per_frame": best["ms_total"], "closest_rays": best["closest_rays"], "shadow_rays": best["shadow_rays"],
            "mrays_per_s": rays / best["ms_total"] / 1e3,
            "closest_kernel_mrays_per_s": best["closest_rays"] / max(best["ms_closest"], 1e-9) / 1e3,
            "any_kernel_mrays_per_s": best["shadow_rays"] / max(best["ms_any"], 1e-9) / 1e3,
            "film_mean": float(film[..., :3].sum() / max(3 * film[..., 3].sum(), 1))}))   # per channel, like the reference's img.mean()
    same = float((films["kdtree"] == films["bvh"]).all(-1).mean())
    print(json.dumps({"kdtree_vs_bvh": {"pixels_bit_identical": same,
                                        "max_abs_diff": float(np.abs(films["kdtree"] - films["bvh"]).max())}}))
    if args.cpu:
        from oracle import refbind as rb
        S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
        spp = max(1, args.spp // 16)
        for accel in ("kdtree", "bvh"):
            la = sc["camera"]["lookat"]
            img, nc, ns, secs = rb.render(S, accel, la[0:3], la[3:6], la[6:9], sc["camera"]["fov"], args.res, args.res, spp,
                                          sc["maxdepth"], sc["light"], (1, 1, 1), (.5, .5, .5), sampler="random", n_threads=0, timing=True)
            print(json.dumps({"reference_render": accel, "cores": os.cpu_count(), "spp": spp, "seconds": secs,
                              "mrays_per_s": (nc + ns) / secs / 1e6, "film_mean": float(img.mean())}))


if __name__ == "__main__":
    main()
