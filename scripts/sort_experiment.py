"""Experiment: how much does ray ORDER matter for the production kernel?  Takes the depth-2 closest
queue (q4) and shadow queue (q5) of the bench workload and times the kernel on (a) the queue as the
wavefront produced it, (b) a stable partition by direction octant, (c) octant + 3D Morton cell of the
origin (full sort), (d) a random shuffle."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import torch
import b200rt
from b200rt import scenes, wavefront as wf

def timeit(fn, reps=7):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best

def part1by2(v):
    v = v.astype(np.uint32) & 0x3ff
    v = (v | (v << 16)) & 0x30000ff
    v = (v | (v << 8)) & 0x300f00f
    v = (v | (v << 4)) & 0x30c30c3
    v = (v | (v << 2)) & 0x9249249
    return v

b200rt.init(0)
torch.zeros(1, device="cuda")
st = torch.cuda.current_stream().cuda_stream
sc = scenes.sponza_standin()
acc = b200rt.MakeAccelerator("bvh", sc["tri_verts"])
integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=8, maxdepth=5, light_pos=sc["light"], seed=1)
film = torch.zeros(integ.shape, dtype=torch.float32, device="cuda")
wf.keep_queues(acc, True)
integ.render_ptr(acc, film.data_ptr(), st)
torch.cuda.synchronize()
wf.keep_queues(acc, False)
wb = acc.WorldBound()
for q in (4, 5):
    rays = wf.queue_rays(acc, q)
    n = len(rays)
    octant = (rays[:, 4] < 0).astype(np.uint32) | ((rays[:, 5] < 0).astype(np.uint32) << 1) | ((rays[:, 6] < 0).astype(np.uint32) << 2)
    cell = np.clip(((rays[:, 0:3] - wb[:3]) / (wb[3:] - wb[:3]) * 1024).astype(np.int64), 0, 1023)
    morton = part1by2(cell[:, 0]) | (part1by2(cell[:, 1]) << 1) | (part1by2(cell[:, 2]) << 2)
    cell6 = cell >> 4
    morton6 = part1by2(cell6[:, 0]) | (part1by2(cell6[:, 1]) << 1) | (part1by2(cell6[:, 2]) << 2)
    orders = {"as produced": np.arange(n), "octant (stable)": np.argsort(octant, kind="stable"),
              "morton18+octant": np.argsort((morton6.astype(np.uint64) << 3) | octant, kind="stable"),
              "octant+morton30": np.argsort((octant.astype(np.uint64) << 30) | morton, kind="stable"),
              "shuffled": np.random.default_rng(0).permutation(n)}
    for name, o in orders.items():
        d = torch.from_numpy(np.ascontiguousarray(rays[o])).cuda()
        out = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
        fn = (lambda: acc.intersect_p_ptr(d.data_ptr(), n, out.data_ptr(), st)) if q % 2 else (lambda: acc.intersect_ptr(d.data_ptr(), n, out.data_ptr(), st))
        ms = timeit(fn)
        print("q%d %-18s %.3f ms = %.0f Mrays/s" % (q, name, ms, n / ms / 1e3), flush=True)
