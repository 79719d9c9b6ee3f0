"""Per-depth wavefront queues of the living-room render (BASELINE config 5) through the production kernels: for each
recorded queue (closest at depth d, shadow at depth d) the ray count, the kernel time (CUDA events, best of 3), the rate,
and the reference traversal's per-ray node fetches / triangle tests from the instrumented kernel -> algorithmic bytes
per ray and GB/s.  Also usable as an ncu target (`--only Q` runs one launch of queue Q).

    python scripts/living_room_queues.py [--spp 16] [--accel bvh] [--depths 8] [--only Q]"""
import argparse
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import torch
import b200rt
from b200rt import scenes, wavefront as wf

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--accel", default="bvh")
ap.add_argument("--depths", type=int, default=8)
ap.add_argument("--maxdepth", type=int, default=65)
ap.add_argument("--only", type=int, default=-1)
ap.add_argument("--xres", type=int, default=640)
ap.add_argument("--yres", type=int, default=360)
args = ap.parse_args()

if args.only >= 0:   # ncu target: the render that fills the queues runs on the baseline kernels, so that only the one
    os.environ["B200RT_KERNEL"] = "baseline"   # launch of interest matches `-k regex:persistent`
b200rt.init(0)
torch.zeros(1, device="cuda")
st = torch.cuda.current_stream().cuda_stream
sc = scenes.living_room(args.xres, args.yres)
tv, il = sc["tri_verts"], sc["infinite_light"]
acc = b200rt.MakeAccelerator(args.accel, tv)
integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=args.spp, maxdepth=args.maxdepth, light_I=il["L"], seed=1,
                                       infinite_light=(il["light_to_world"], wf.world_radius(tv)),
                                       materials=sc["materials"], prim_material=sc["prim_material"],
                                       textures=scenes.living_room_textures(sc))
film = torch.zeros(integ.shape, dtype=torch.float32, device="cuda")
wf.keep_queues(acc, True)
stats = integ.render_ptr(acc, film.data_ptr(), st)
torch.cuda.synchronize()
wf.keep_queues(acc, False)
if args.only >= 0:
    os.environ["B200RT_KERNEL"] = "persistent"
    b200rt.init(0)
print("render:", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in stats.items()})
node_bytes = 32 if args.accel == "bvh" else 8
qs = [args.only] if args.only >= 0 else range(2 * args.depths)
for q in qs:
    n = wf.queue_count(acc, q)
    if n == 0:
        continue
    buf = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    wf.queue_rays(acc, q, buf.data_ptr())
    out = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
    best = 1e30
    for _ in range(1 if args.only >= 0 else 3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        if q % 2:
            acc.intersect_p_ptr(buf.data_ptr(), n, out.data_ptr(), st)
        else:
            acc.intersect_ptr(buf.data_ptr(), n, out.data_ptr(), st)
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    line = "q%-3d %-7s depth %-2d n=%-9d %8.3f ms %7.0f Mrays/s" % (q, "shadow" if q % 2 else "closest", q // 2, n, best, n / best / 1e3)
    if args.only < 0:
        host = buf.cpu().numpy()
        _, c = (acc.IntersectP if q % 2 else acc.Intersect)(host, counters=True)
        f, t = c["nodes_fetched"] / c["rays"], c["prim_tests"] / c["rays"]
        bpr = 32 + (1 if q % 2 else 20) + node_bytes * f + 36 * t
        line += "  nodes/ray %6.1f tests/ray %5.2f  %6.0f B/ray -> %6.0f GB/s" % (f, t, bpr, bpr * n / best / 1e6)
    print(line, flush=True)
