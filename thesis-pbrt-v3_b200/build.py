"""Builds libb200rt.so (CUDA kernels + C ABI + host builders) for sm_100a, in-tree.

    python thesis-pbrt-v3_b200/build.py [--force] [--verbose]

-fmad=false and IEEE division are part of the parity contract (csrc/rt_math.cuh)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "b200rt", "libb200rt.so")
OBJ = os.path.join(HERE, "_obj")
SOURCES = ["b200rt.cu", "wavefront.cu", "film_reduce.cu", "build_rbsp.cu", "no_alloc.cu", "build_bvh.cpp", "build_kd.cpp", "wf_sampler_host.cpp", "wf_camera_host.cpp", "wf_setup_host.cpp", "wf_texture_host.cpp"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
         "--shared", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-cudart", "static"]


def sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b200rt.h"),
                                                                 os.path.join(HERE, "..", "include", "b200rt_wavefront.h")]
    return all(os.path.getmtime(d) <= t for d in deps if os.path.exists(d))


def _compile_one(src, obj, verbose):
    cmd = [NVCC] + [f for f in FLAGS if f != "--shared"] + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)


def build(force=False, verbose=False):
    """One object per source (compiled in parallel, rebuilt only when the source or a header is newer), then one link."""
    if not force and up_to_date():
        return OUT
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))] + \
           [os.path.join(HERE, "..", "include", h) for h in ("b200rt.h", "b200rt_wavefront.h")]
    hdr_t = max(os.path.getmtime(h) for h in hdrs if os.path.exists(h))
    jobs, objs = [], []
    for src in sources():
        obj = os.path.join(OBJ, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t):
            jobs.append((src, obj))
    with ThreadPoolExecutor(max_workers=max(1, len(jobs))) as ex:
        for f in [ex.submit(_compile_one, s, o, verbose) for s, o in jobs]:
            f.result()
    subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "--shared", "-cudart", "static", "-o", OUT] + objs)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(OUT)
