"""Host-buffer (e2e) throughput of b200rt_intersect_closest/any vs pipeline chunk size, plus the raw
pinned H2D / D2H bandwidth of the box for reference.  Tuning aid."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import torch
import b200rt
from b200rt import scenes

b200rt.init(0)
torch.zeros(1, device="cuda")
st = torch.cuda.current_stream().cuda_stream
h = torch.empty(1 << 28, dtype=torch.uint8).pin_memory()
d = torch.empty(1 << 28, dtype=torch.uint8, device="cuda")
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); [fn() for _ in range(5)]; b.record(); torch.cuda.synchronize()
    print("%s pinned 256 MiB: %.1f GB/s" % (name, 5 * (1 << 28) / a.elapsed_time(b) / 1e6), flush=True)
sc = scenes.sponza_standin()
acc = b200rt.MakeAccelerator("bvh", sc["tri_verts"])
rays = torch.from_numpy(scenes.random_rays(acc.WorldBound(), 1 << 22)).pin_memory()
hits = torch.empty((1 << 22, 5), dtype=torch.int32).pin_memory()
occ = torch.empty(1 << 22, dtype=torch.uint8).pin_memory()
N = int(os.environ.get("E2E_N", 1 << 22))
for chunk, ramp in ((131072, 0), (131072, 1), (262144, 0), (262144, 1), (524288, 0), (524288, 1)):
    os.environ["B200RT_PIPE_CHUNK"] = str(chunk)
    os.environ["B200RT_PIPE_RAMP"] = str(ramp)
    b200rt.init(0)
    for nm, fn in (("closest", lambda: acc.intersect_ptr(rays.data_ptr(), N, hits.data_ptr(), st)),
                   ("any", lambda: acc.intersect_p_ptr(rays.data_ptr(), N, occ.data_ptr(), st))):
        fn(); fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); [fn() for _ in range(5)]; b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        print("n %8d chunk %8d ramp %d %-7s %.3f ms = %.0f Mrays/s (H2D %.1f GB/s)" % (N, chunk, ramp, nm, ms, N / ms / 1e3, 32 * N / ms / 1e6), flush=True)
