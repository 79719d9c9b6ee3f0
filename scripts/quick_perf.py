"""Ad-hoc kernel timing (device-resident rays), used while tuning.  Not the bench contract."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import torch
import b200rt
from b200rt import scenes

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best

def main():
    b200rt.init(0)
    torch.cuda.init(); torch.zeros(1, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    which = sys.argv[1:] or ["sponza", "killeroo"]
    for name in which:
        sc = scenes.sponza_standin() if name == "sponza" else scenes.killeroo_simple()
        v = sc["tri_verts"]
        cam = dict(sc["camera"], xres=2048, yres=2048)
        sets = {"camera": scenes.camera_rays(cam), "random": scenes.random_rays(None, 0) if False else None}
        for accname in ("bvh", "kdtree"):
            t0 = time.time(); acc = b200rt.MakeAccelerator(accname, v); tb = time.time() - t0
            if sets["random"] is None:
                sets["random"] = scenes.random_rays(acc.WorldBound(), 1 << 22)
            for rname, rays in sets.items():
                n = len(rays)
                d_r = torch.from_numpy(rays).cuda()
                d_h = torch.empty(n * 5, dtype=torch.int32, device="cuda")
                d_o = torch.empty(n, dtype=torch.uint8, device="cuda")
                ms_c = timeit(lambda: acc.intersect_ptr(d_r.data_ptr(), n, d_h.data_ptr(), st))
                ms_a = timeit(lambda: acc.intersect_p_ptr(d_r.data_ptr(), n, d_o.data_ptr(), st))
                _, c = acc.Intersect(rays[: 1 << 20], counters=True)
                hitrate = float((d_h.view(-1, 5)[:, 0] != -1).float().mean())
                print("%-9s %-6s %-6s n=%d build %.1fs closest %.2f ms = %.0f Mrays/s | any %.2f ms = %.0f Mrays/s | hit %.2f fetch/ray %.1f tests/ray %.2f"
                      % (name, accname, rname, n, tb, ms_c, n / ms_c / 1e3, ms_a, n / ms_a / 1e3, hitrate,
                         c["nodes_fetched"] / c["rays"], c["prim_tests"] / c["rays"]), flush=True)
            acc.close()

if __name__ == "__main__":
    main()
