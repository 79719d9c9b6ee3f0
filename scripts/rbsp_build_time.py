"""Device RBSP build times on the thesis's config-4 scene (scenes/killeroo-feet, 33 264 triangles) — the builds the thesis's own
table is dominated by (38-319 s on its CPU; this repo's round-1 log measured 48-316 s with the reference's builder on the box).
With --ref the reference's own builder (oracle/_ref) is timed beside it on the host and the node arrays are compared."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import b200rt
from b200rt import scenes

ap = argparse.ArgumentParser()
ap.add_argument("--dirs", default="3,7,9,13")
ap.add_argument("--ref", default="", help="comma list of nbDirections to also build with the reference's builder")
ap.add_argument("--scene", default="feet")
a = ap.parse_args()
b200rt.init(0)
sc = scenes.killeroo_feet() if a.scene == "feet" else scenes.killeroo_simple()
tv = np.ascontiguousarray(sc["tri_verts"], np.float32)
refs = [int(x) for x in a.ref.split(",") if x]
b200rt.build_rbsp(tv[:64], nbDirections=3)   # context + module load
for M in [int(x) for x in a.dirs.split(",")]:
    t0 = time.time()
    t = b200rt.build_rbsp(tv, nbDirections=M)
    wall = time.time() - t0
    rec = dict(scene=a.scene, triangles=len(tv), nbDirections=M, nodes=len(t["nodes"]), indices=len(t["indices"]), wall_s=wall, **t["build"])
    if M in refs:
        from oracle import refbind as rb
        S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
        t0 = time.time()
        ex = rb.RefAccel(S, "rbsp", {"nbDirections": M}).export()
        rec["reference_build_s"] = time.time() - t0
        rec["nodes_equal_reference"] = bool(ex["nodes"].shape == t["nodes"].shape and np.array_equal(ex["nodes"], t["nodes"]))
    print(json.dumps(rec), flush=True)
