"""A/B on ONE B200: the share of the headline frame that one rank of an N-rank run renders (tiles t % N == 0), at several
wavefront batch sizes.  At N = 8 a rank's 3.9 M paths fit ONE batch of the then default size (2^22 paths), so the second set of queue buffers
(two batches in flight) has nothing to overlap with; this measures whether splitting the share into 2-4 smaller
batches buys the overlap back.

    python scripts/ab_rank_share_batches.py [--world 8] > gpurun_out/ab_rank_share_batches.log
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, nargs="*", default=[8, 4, 1])
    ap.add_argument("--frames", type=int, default=10)
    args = ap.parse_args()
    import torch
    import b200rt
    from b200rt import scenes
    import bench

    b200rt.init(0)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    sc = scenes.sponza_standin()
    accel = b200rt.MakeAccelerator("bvh", sc["tri_verts"])
    for world in args.world:
        ref = None
        for max_batch in (0, 1 << 21, 1 << 20, 1 << 19):
            integ = b200rt.WavefrontPathIntegrator(sc["camera"], spp=bench.SPP, maxdepth=bench.MAXDEPTH, light_pos=sc["light"],
                                                   light_I=bench.LIGHT_I, kd=bench.KD, rank=0, world=world, seed=1,
                                                   max_batch=max_batch)
            film = torch.zeros(integ.shape, dtype=torch.float32, device=dev)
            for _ in range(3):
                film.zero_()
                st = integ.render_ptr(accel, film.data_ptr(), stream)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.frames):
                film.zero_()
                st = integ.render_ptr(accel, film.data_ptr(), stream)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.frames
            f = film.cpu()
            if ref is None:
                ref = f
            rays = st["closest_rays"] + st["shadow_rays"]
            print(json.dumps({"world": world, "rank": 0, "max_batch": max_batch or "default", "ms_per_frame": ms,
                              "rays": rays, "mrays_per_s": rays / ms / 1e3, "film_equals_default": bool(torch.equal(f, ref))}),
                  flush=True)


if __name__ == "__main__":
    main()
