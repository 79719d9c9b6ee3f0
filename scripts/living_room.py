"""BASELINE config 5 on the B200: scenes/living-room (the reference's own 143 163-triangle geometry, camera, path
maxdepth 65, constant infinite light = what the reference falls back to for the missing sky map, materials substituted
as b200rt/scenes.py::living_room documents) tile-sharded over the ranks, film gathered to rank 0 over NCCL.

    python scripts/living_room.py --xres 1920 --yres 1080 --spp 1024                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \\
           --master-port 29511 scripts/living_room.py --xres 1920 --yres 1080 --spp 1024   # 8 GPUs

Prints one JSON line (rank 0): frame time (CUDA events, max over ranks), rays traced, whole-frame Mrays/s, and the
closest-hit / any-hit kernel rates of this rank's queues.  Not the headline bench (bench.py = config 2)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--xres", type=int, default=1920)
    ap.add_argument("--yres", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--maxdepth", type=int, default=65)
    ap.add_argument("--accel", default="bvh", choices=["bvh", "kdtree"])
    ap.add_argument("--frames", type=int, default=1)
    ap.add_argument("--warm-spp", type=int, default=4, help="spp of the untimed warm-up frame")
    ap.add_argument("--out", default="", help="rank 0 writes the frame's RGB as .npy here")
    ap.add_argument("--sampler", default="sobol", choices=["sobol", "halton", "hash"], help="the file asks for sobol")
    ap.add_argument("--no-textures", action="store_true", help="constants instead of the two image files (A/B)")
    ap.add_argument("--no-normals", action="store_true", help="geometric normals instead of the PLY vertex normals (A/B)")
    args = ap.parse_args()

    import numpy as np
    import torch
    import torch.distributed as dist
    import b200rt
    from b200rt import scenes, wavefront as wf
    from b200rt.film import FilmGather

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import datetime
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=600))
    b200rt.init(local)
    dev = torch.device("cuda", local)
    stream = torch.cuda.current_stream().cuda_stream
    sc = scenes.living_room(args.xres, args.yres, textures=not args.no_textures)
    tv = sc["tri_verts"]
    accel = b200rt.MakeAccelerator(args.accel, tv, tri_uvs=sc["tri_uvs"], tri_normals=None if args.no_normals else sc["tri_normals"], prim_flags=None if args.no_normals else sc["prim_flags"])
    il = sc["infinite_light"]

    textures = scenes.living_room_textures(sc)

    def integrator(spp):
        return b200rt.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=args.maxdepth, light_I=il["L"],
                                              infinite_light=(il["light_to_world"], wf.world_radius(tv)),
                                              materials=sc["materials"], prim_material=sc["prim_material"],
                                              textures=textures, rank=rank, world=world, sampler=args.sampler, seed=1)
    film = torch.zeros((args.yres, args.xres, 4), dtype=torch.float32, device=dev)
    gather = FilmGather(args.xres, args.yres, 16, rank, world, dev)

    def frame(integ):
        film.zero_()
        st = integ.render_ptr(accel, film.data_ptr(), stream)
        gather.gather(film)
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the warm-up frame must be large enough to make the library allocate its full-size wavefront buffers (2^25 paths per batch,
    # two batches in flight: several GB), or that allocation lands in the timed frame
    warm = max(args.warm_spp, -(-(1 << 26) // max(1, args.xres * args.yres // world)))
    warm = min(args.spp, 1 << (warm - 1).bit_length())   # Sampler "sobol" wants a power of two
    frame(integrator(warm))
    integ = integrator(args.spp)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    agg = {}
    for _ in range(args.frames):
        st = frame(integ)
        for k, v in st.items():
            agg[k] = agg.get(k, 0) + v
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1), float(agg["closest_rays"]), float(agg["shadow_rays"]), agg["ms_closest"],
                      agg["ms_any"], float(agg["mis_rays"]), float(agg["camera_rays"])], dtype=torch.float64, device=dev)
    tmax, tsum = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    if rank == 0:
        ms = float(tmax[0]) / args.frames
        rays = float(tsum[1] + tsum[2]) / args.frames
        rgb = wf.film_to_rgb(film.cpu().numpy())
        if args.out:
            np.save(args.out, rgb)
        print(json.dumps({
            "workload": "scenes/living-room (reference geometry, 143163 triangles, %s), %dx%d @ %d spp, path maxdepth %d, "
                        "Sampler %s, constant infinite light + MIS, the file's materials (matte, Oren-Nayar, substrate, metal, glass, mirror, uber)%s%s; "
                        "box filter" % (args.accel, args.xres, args.yres, args.spp, args.maxdepth, args.sampler,
                                        "" if args.no_textures else ", image textures", "" if args.no_normals else ", PLY vertex normals"),
            "n_gpus": world, "ms_per_frame": ms, "rays_per_frame": rays, "mrays_per_s": rays / ms / 1e3,
            "camera_rays": float(tsum[6]) / args.frames, "closest_rays": float(tsum[1]) / args.frames,
            "shadow_rays": float(tsum[2]) / args.frames, "mis_rays": float(tsum[5]) / args.frames,
            "rank0_closest_kernel_mrays_per_s": agg["closest_rays"] / agg["ms_closest"] / 1e3,
            "rank0_any_kernel_mrays_per_s": agg["shadow_rays"] / agg["ms_any"] / 1e3,
            "rank0_trace_share_of_frame": (agg["ms_closest"] + agg["ms_any"]) / float(t[0]),
            "film_mean_rgb": [float(x) for x in rgb.reshape(-1, 3).mean(0)],
            "film_gather_bytes": gather.bytes_per_gather if world > 1 else 0}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
