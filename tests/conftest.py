import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


def golden_tree(g):
    return dict(kind=str(g["kind"]), nodes=g["nodes"], indices=g["indices"], directions=g["directions"],
                bounds=g["bounds"], n_prims=len(g["tri_verts"]))


SMALL = ["bvh", "kdtree", "rbsp3", "rbsp7", "rbsp13", "rbspkd13", "bsppaper", "bsppaperkd", "bspcluster", "bspclusterfastkd"]


@pytest.fixture(scope="session")
def built_lib():
    """Builds (if needed) and loads libb200rt.so; the build is a no-op when the .so is current."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("b200rt_build", os.path.join(ROOT, "thesis-pbrt-v3_b200", "build.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    if not os.path.exists(m.OUT) or (os.path.exists("/usr/local/cuda/bin/nvcc") and not m.up_to_date()):
        m.build()
    import b200rt
    return b200rt


@pytest.fixture(scope="session")
def gpu(built_lib):
    built_lib.init(0)
    return built_lib


FEET = ["bvh", "rbsp3", "rbsp7", "rbsp9", "rbsp13", "rbspkd13"]


def feet_tree(g, name):
    """Reference-built tree `name` of tests/golden/ref_feet.npz (BASELINE config 4) in the dict form
    B200Accel / OracleAccel take."""
    kind = "bvh" if name == "bvh" else ("rbspkd" if name.startswith("rbspkd") else "rbsp")
    return dict(kind=kind, nodes=g[name + "_nodes"], indices=g[name + "_indices"], directions=g[name + "_directions"],
                bounds=g[name + "_bounds"], n_prims=len(g["tri_verts"]))


def spheres_scene(g, name):
    """tests/golden/ref_spheres.npz -> (tree, tri_verts padded to all prims, prim_kind, sphere rows)."""
    nt, ns = len(g["tri_verts"]), len(g["spheres"])
    kind = "bvh" if name == "bvh" else ("kdtree" if name == "kdtree" else "rbsp")
    tree = dict(kind=kind, nodes=g[name + "_nodes"], indices=g[name + "_indices"], directions=g[name + "_directions"],
                bounds=g[name + "_bounds"], n_prims=nt + ns)
    verts = np.concatenate([g["tri_verts"], np.zeros((ns, 9), np.float32)])
    prim_kind = np.concatenate([np.zeros(nt, np.uint8), np.ones(ns, np.uint8)])
    cr = g["spheres"]
    rows = np.zeros((ns, 16), np.float32)  # WorldToObject = Translate(-c) (pure-Translate ObjectToWorld), radius
    rows[:, 0] = rows[:, 5] = rows[:, 10] = 1
    rows[:, 3], rows[:, 7], rows[:, 11], rows[:, 12] = -cr[:, 0], -cr[:, 1], -cr[:, 2], cr[:, 3]
    return tree, verts, prim_kind, rows
