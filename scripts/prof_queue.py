"""ncu target: one launch each of the production kernels on three real wavefront queues of the bench
workload (q0 = camera rays, q4 = closest-hit queue at depth 2, q5 = shadow queue at depth 2).

    ncu --set full -k regex:bvh_persistent -c 3 python scripts/prof_queue.py [spp] [bvh|kdtree] [sponza|museum|conference|killeroo]
The render that produces the queues runs on the baseline kernel so that only the three launches of
interest match the filter."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import torch
import b200rt
from b200rt import scenes, wavefront as wf

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
accname = sys.argv[2] if len(sys.argv) > 2 else "bvh"
os.environ["B200RT_KERNEL"] = "baseline"
b200rt.init(0)
torch.zeros(1, device="cuda")
st = torch.cuda.current_stream().cuda_stream
scene = sys.argv[3] if len(sys.argv) > 3 else "sponza"
sc = {"sponza": scenes.sponza_standin, "museum": scenes.museum_standin, "conference": scenes.conference_standin,
      "killeroo": scenes.killeroo_simple}[scene]()
acc = b200rt.MakeAccelerator(accname, sc["tri_verts"])
integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=5, light_pos=sc["light"], seed=1)
film = torch.zeros(integ.shape, dtype=torch.float32, device="cuda")
wf.keep_queues(acc, True)
integ.render_ptr(acc, film.data_ptr(), st)
torch.cuda.synchronize()
wf.keep_queues(acc, False)
os.environ["B200RT_KERNEL"] = "persistent"
b200rt.init(0)
for q in (0, 4, 5):
    n = wf.queue_count(acc, q)
    buf = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    wf.queue_rays(acc, q, buf.data_ptr())
    out = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    if q % 2:
        acc.intersect_p_ptr(buf.data_ptr(), n, out.data_ptr(), st)
    else:
        acc.intersect_ptr(buf.data_ptr(), n, out.data_ptr(), st)
    b.record()
    torch.cuda.synchronize()
    print("q%d n=%d %.3f ms = %.0f Mrays/s" % (q, n, a.elapsed_time(b), n / a.elapsed_time(b) / 1e3))
