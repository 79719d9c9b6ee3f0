"""A/B kernel timing on the REAL wavefront ray queues of the bench workload (sponza stand-in, BVH,
path maxdepth 5): baseline one-ray-per-thread kernel vs the persistent-warp kernel at several refill
thresholds.  Also checks that both give bit-identical results.  Tuning aid, not the bench contract.

    python scripts/ab_perf.py [spp] [thresh,thresh,...]
"""
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


def main():
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    threshes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 8, 16, 20, 24, 28]
    scene = sys.argv[3] if len(sys.argv) > 3 else "sponza"
    accname = sys.argv[4] if len(sys.argv) > 4 else "bvh"
    b200rt.init(0)
    torch.zeros(1, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sc = {"sponza": scenes.sponza_standin, "killeroo": scenes.killeroo_simple, "museum": scenes.museum_standin,
          "conference": scenes.conference_standin, "living": scenes.living_room}[scene]()
    maxdepth = int(os.environ.get("B200RT_AB_MAXDEPTH", sc.get("maxdepth", 5) if scene in ("museum", "conference") else 5))
    acc = b200rt.MakeAccelerator(accname, sc["tri_verts"])
    light = sc["light"] if "light" in sc else np.array([0.0, 1.5, 0.0], np.float32)   # (living-room: a point light stands in for its sky)
    integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_pos=light, seed=1)
    film = torch.zeros(integ.shape, dtype=torch.float32, device="cuda")
    wf.keep_queues(acc, True)
    integ.render_ptr(acc, film.data_ptr(), st)
    torch.cuda.synchronize()
    wf.keep_queues(acc, False)
    queues = {}
    for q in range(12):
        n = wf.queue_count(acc, q)
        if n:
            buf = torch.empty((n, 8), dtype=torch.float32, device="cuda")
            wf.queue_rays(acc, q, buf.data_ptr())
            queues[q] = buf
    # coherent camera rays + incoherent random rays as two extra sets
    cam = dict(sc["camera"], xres=2048, yres=2048)
    queues[100] = torch.from_numpy(scenes.camera_rays(cam)).cuda()
    queues[102] = torch.from_numpy(scenes.random_rays(acc.WorldBound(), 1 << 22)).cuda()
    configs = [("baseline", None)] + [("persist", t) for t in threshes]
    # extra env sweeps: B200RT_AB_EXTRA="NAME=VAL,NAME=VAL;NAME=VAL" -> one persistent config per ';' group
    extra = [g for g in os.environ.get("B200RT_AB_EXTRA", "").split(";") if g]
    configs += [("persist", g) for g in extra]
    results = {}
    ref_out = {}
    for name, t in configs:
        os.environ["B200RT_KERNEL"] = "baseline" if name == "baseline" else "persistent"
        if isinstance(t, str):
            for kv in t.split(","):
                k, v = kv.split("=")
                os.environ[k] = v
        elif t is not None:
            os.environ["B200RT_THRESH"] = str(t)
        b200rt.init(0)
        tot_c = tot_a = 0.0
        for q, buf in queues.items():
            n = len(buf)
            anyq = (q % 2 == 1)
            out = torch.zeros(n * 5, dtype=torch.int32, device="cuda") if not anyq else torch.zeros(n, dtype=torch.uint8, device="cuda")
            fn = (lambda: acc.intersect_p_ptr(buf.data_ptr(), n, out.data_ptr(), st)) if anyq else \
                 (lambda: acc.intersect_ptr(buf.data_ptr(), n, out.data_ptr(), st))
            ms = timeit(fn)
            results[(name, t, q)] = (n, ms)
            if name == "baseline":
                ref_out[q] = out.clone()
            else:
                if os.environ.get("B200RT_AB_ALLOW_MISMATCH"):
                    bad = int((out != ref_out[q]).view(-1, 5 if not anyq else 1).any(1).sum())
                    if bad:
                        print("   queue %d: %d of %d results differ from the baseline kernel (%s)" % (q, bad, n, t))
                else:
                    assert torch.equal(out, ref_out[q]), "MISMATCH vs baseline kernel on queue %d (thresh %s)" % (q, t)
            if q < 100:
                if anyq: tot_a += ms
                else: tot_c += ms
        nc = sum(len(b) for q, b in queues.items() if q < 100 and q % 2 == 0)
        na = sum(len(b) for q, b in queues.items() if q < 100 and q % 2 == 1)
        print("%-9s %-28s | path queues: closest %7.3f ms = %6.0f Mrays/s | any %7.3f ms = %6.0f Mrays/s | camera %6.0f | random %6.0f Mrays/s"
              % (name, t, tot_c, nc / tot_c / 1e3, tot_a, na / max(tot_a, 1e-9) / 1e3,
                 results[(name, t, 100)][0] / results[(name, t, 100)][1] / 1e3,
                 results[(name, t, 102)][0] / results[(name, t, 102)][1] / 1e3), flush=True)
    print("per-queue ms (n):")
    for q in queues:
        print(" q%-3d n=%8d " % (q, len(queues[q])) + " ".join("%s/%s=%.3f" % (nm[:4], t, results[(nm, t, q)][1]) for nm, t in configs))
    print("all persistent results bit-identical to the baseline kernel")


if __name__ == "__main__":
    main()
