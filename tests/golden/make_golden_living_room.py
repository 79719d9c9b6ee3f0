"""Generates tests/golden/ref_living_room.npz — BASELINE config 5 (scenes/living-room) from the UNMODIFIED reference
(oracle/_ref/libpbrtref.so, built from /root/reference by oracle/ref/Makefile).  Run in the build container:

    python tests/golden/make_golden_living_room.py

Contents
* the reference's own BVH and kd-tree over the scene's 143 163 triangles: node/index counts + SHA-256 of the node and
  index arrays (the arrays are 7.5 / 11 MB; this repo's builders must reproduce them byte for byte);
* RAYS rays (camera rays of the scene's camera, rays between random surface points, random rays in the bounds) with the
  reference's answers for BVHAccel and KdTreeAccel: Intersect -> (prim, t, uv, p), IntersectP -> occluded;
* the reference's OWN render (SamplerIntegrator::Render + PathIntegrator maxdepth 65 + InfiniteAreaLight with the
  constant map it falls back to, since textures/Sky 19.pfm is absent + the material substitution documented in
  b200rt/scenes.py::living_room) at 320x180, 64 spp, box filter, with its ray counts.
"""
import hashlib
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
from oracle import refbind as rb   # noqa: E402
from b200rt import scenes          # noqa: E402

RES, SPP = (320, 180), 64


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def ray_set(sc, n_each=4096, seed=11):
    rng = np.random.default_rng(seed)
    tv = sc["tri_verts"].reshape(-1, 3, 3).astype(np.float64)
    cam = scenes.camera_rays(sc["camera"], spp=1, seed=3)
    cam = cam[rng.choice(len(cam), n_each, replace=False)]
    # segments between random points ON the surfaces (shadow-ray-like: unnormalised d, tMax just short of 1)
    def surf(k):
        t = tv[rng.integers(0, len(tv), k)]
        b = rng.dirichlet([1, 1, 1], k)
        return (b[:, :, None] * t).sum(1)
    a, b = surf(n_each), surf(n_each)
    seg = np.zeros((n_each, 8), np.float32)
    seg[:, 0:3], seg[:, 4:7], seg[:, 3] = a, b - a, 1 - 1e-4
    lo, hi = tv.reshape(-1, 3).min(0), tv.reshape(-1, 3).max(0)
    rnd = scenes.random_rays(list(lo) + list(hi), n_each, seed=seed + 1)
    return np.concatenate([cam, seg, rnd]).astype(np.float32)


def main():
    sc = scenes.living_room(*RES)
    tv = sc["tri_verts"]
    out = {}
    S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
    rays = ray_set(sc)
    out["rays"] = rays
    for kind in ("bvh", "kdtree"):
        acc = rb.RefAccel(S, kind)
        t = acc.export()
        out[kind + "_n_nodes"] = np.int64(len(t["nodes"]))
        out[kind + "_n_indices"] = np.int64(len(t["indices"]))
        out[kind + "_nodes_sha256"] = np.array(sha(t["nodes"]))
        out[kind + "_indices_sha256"] = np.array(sha(t["indices"]))
        out[kind + "_bounds"] = t["bounds"]
        h = acc.intersect(rays, n_threads=8)
        out[kind + "_prim"], out[kind + "_t"] = h["prim"].astype(np.int32), h["t"]
        out[kind + "_uv"], out[kind + "_p"] = h["uv"], h["p"]   # default uvs: u = b1 + b2, v = b2 (triangle.cpp:55-60)
        out[kind + "_occ"] = acc.intersect_p(rays, n_threads=8)
        print(kind, "nodes", len(t["nodes"]), "indices", len(t["indices"]), "hit fraction", (h["prim"] >= 0).mean(),
              "occluded", out[kind + "_occ"].mean())
    assert np.array_equal(out["bvh_t"], out["kdtree_t"])   # cross-accelerator agreement
    il = sc["infinite_light"]
    l2w = np.eye(4, dtype=np.float32)
    l2w[:3, :3] = il["light_to_world"]
    img, nc, ns = rb.render(S, "bvh", (0, 0, 0), (0, 0, 1), (0, 1, 0), 58.715508, RES[0], RES[1], SPP, sc["maxdepth"],
                            (0, 0, 0), il["L"], (.5, .5, .5), sampler="random", n_threads=0,
                            prim_mat=sc["materials"][sc["prim_material"]], infinite=(il["L"], l2w),
                            camera_matrix=sc["camera"]["world_to_camera"])
    print("render mean", img.mean(), "closest", nc, "shadow", ns)
    out["render_rgb"], out["render_rays"] = img, np.array([nc, ns], np.uint64)
    out["render_params"] = np.array([RES[0], RES[1], SPP, sc["maxdepth"]], np.int32)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_living_room.npz"), **out)


if __name__ == "__main__":
    main()
