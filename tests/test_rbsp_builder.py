"""The RBSP builder (reference src/accelerators/rbsp.cpp:181-403 + kDOPMesh.h).

The product builds the tree on the B200 (b200rt_build_rbsp, csrc/build_rbsp.cu): level-synchronous, every candidate plane of a
level cut and measured in parallel.  Its kernels are functors (csrc/rbsp_build.cuh) that tests/hostbuild also runs in host
loops, so the restatement is pinned here on the CPU against the reference's own trees — the committed goldens (made by
oracle/_ref, tests/golden/make_golden.py) and, where oracle/_ref is present, fresh reference builds with other parameters:

  * the NODE ARRAY (split direction, split position, above-child index, leaf sizes, leaf offsets / single primitives) equals
    the reference's byte for byte;
  * the primitives of every leaf are the reference's as a SET — except leaves whose membership in the reference hangs on
    std::sort's (unstable) order among bound edges with equal keys: bounded by 1e-3 of the leaves, same size, and reported.

On the GPU the device build must equal the host-run functors exactly (same text, same tie order), equal the goldens in the
sense above, and the tree must answer the goldens' rays bit-exactly through the traversal kernels."""
import ctypes as C
import os
import subprocess
import time
import numpy as np
import pytest
from conftest import ROOT, load_golden

HB_DIR = os.path.join(ROOT, "tests", "hostbuild")


@pytest.fixture(scope="module")
def hb():
    subprocess.check_call(["make", "-s", "-C", HB_DIR])
    L = C.CDLL(os.path.join(HB_DIR, "libhostbuild.so"))
    L.hb_build_rbsp.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.c_uint32), C.POINTER(C.POINTER(C.c_uint32)),
                                C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
    L.hb_free.argtypes = [C.c_void_p]

    def build(tv, nbDirections=3, intersectcost=80, traversalcost=5, emptybonus=0.0, maxprims=1, maxdepth=-1):
        tv = np.ascontiguousarray(tv, np.float32).reshape(-1, 9)
        nodes, idx = C.POINTER(C.c_uint32)(), C.POINTER(C.c_uint32)()
        nn, ni, cand = C.c_uint32(), C.c_uint32(), C.c_uint64()
        dirs, b = np.zeros((16, 3), np.float32), np.zeros(6, np.float32)
        rc = L.hb_build_rbsp(tv.ctypes.data, len(tv), nbDirections, intersectcost, traversalcost, emptybonus, maxprims, maxdepth,
                             C.byref(nodes), C.byref(nn), C.byref(idx), C.byref(ni), dirs.ctypes.data, b.ctypes.data, C.byref(cand))
        assert rc == 0, rc
        N = np.ctypeslib.as_array(nodes, (nn.value, 2)).copy()
        I = np.ctypeslib.as_array(idx, (max(ni.value, 1),)).copy()[:ni.value]
        L.hb_free(nodes)
        L.hb_free(idx)
        return dict(kind="rbsp", nodes=N, indices=I, directions=dirs[:nbDirections % 100].copy(), bounds=b, n_prims=len(tv),
                    candidates=cand.value)
    return build


def leaf_sets(nodes, indices, M):
    bo = int(M).bit_length()   # getBitOffset(M), kDOPMesh.h:229-236
    w0, w1 = nodes[:, 0].astype(np.int64), nodes[:, 1].astype(np.int64)
    leaf = (w1 & ((1 << bo) - 1)) == M
    out = []
    for a, n in zip(w0[leaf], (w1[leaf] >> bo)):
        out.append(tuple(sorted(indices[a:a + n].tolist())) if n > 1 else ((int(a),) if n == 1 else ()))
    return out


def compare_with_reference(tree, ref_nodes, ref_indices, M):
    """Node arrays byte for byte — the one word that may differ is the primitive a SINGLE-primitive leaf stores in the node
    itself (onePrimitive, rbsp.cpp:141-144), which is leaf membership; leaves as sets.  Returns the number of leaves whose
    members differ."""
    assert tree["nodes"].shape == ref_nodes.shape, (tree["nodes"].shape, ref_nodes.shape)
    bo = int(M).bit_length()
    single = ((ref_nodes[:, 1] & ((1 << bo) - 1)) == M) & ((ref_nodes[:, 1] >> bo) == 1)
    assert np.array_equal(tree["nodes"][:, 1], ref_nodes[:, 1])                       # kinds, directions, children, leaf sizes
    assert np.array_equal(tree["nodes"][~single, 0], ref_nodes[~single, 0])           # split positions, leaf offsets
    assert len(tree["indices"]) == len(ref_indices)
    a, r = leaf_sets(tree["nodes"], tree["indices"], M), leaf_sets(ref_nodes, ref_indices, M)
    bad = [i for i, (x, y) in enumerate(zip(a, r)) if x != y]
    assert all(len(a[i]) == len(r[i]) for i in bad)
    assert len(bad) <= 1e-3 * len(a), (len(bad), len(a))
    return len(bad), len(a)


@pytest.mark.parametrize("M", [3, 7, 13])
def test_host_run_kernels_reproduce_the_reference_tree_small(hb, M):
    g = load_golden("ref_small_rbsp%d.npz" % M)
    t = hb(g["tri_verts"], nbDirections=M)
    assert np.array_equal(t["directions"].view(np.uint32), g["directions"].view(np.uint32))   # getDirections, RBSPShared.h:30-76
    assert np.array_equal(t["bounds"], g["bounds"])
    bad, n = compare_with_reference(t, g["nodes"], g["indices"], M)
    assert bad == 0 and np.array_equal(t["indices"], g["indices"])   # a random soup has no tied bounds: identical to the last index


@pytest.mark.parametrize("M", [3, 7, 9, 13])
def test_host_run_kernels_reproduce_the_reference_tree_feet(hb, M):
    """scenes/killeroo-feet (config 4): a mesh, so neighbouring triangles share bounds and the reference's sort order among the
    ties is unspecified — node arrays still equal byte for byte."""
    g = load_golden("ref_feet.npz")
    t = hb(g["tri_verts"], nbDirections=M)
    bad, n = compare_with_reference(t, g["rbsp%d_nodes" % M], g["rbsp%d_indices" % M], M)
    print("\nfeet rbsp-%d: %d nodes byte-identical; %d of %d leaves differ in membership (std::sort ties)" % (M, len(t["nodes"]), bad, n))


def _ref():
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libpbrtref.so")):
        pytest.skip("oracle/_ref is not built")
    from oracle import refbind
    return refbind


@pytest.mark.parametrize("params", [dict(nbDirections=7, maxprims=4), dict(nbDirections=9, emptybonus=0.5),
                                    dict(nbDirections=13, intersectcost=20, traversalcost=1), dict(nbDirections=3, maxdepth=6),
                                    dict(nbDirections=7, maxprims=3, emptybonus=0.3, maxdepth=11)])
def test_parameters_against_a_fresh_reference_build(hb, params):
    """CreateRBSPTreeAccelerator's parameters (rbsp.cpp:549-570) other than the defaults, on a soup and on a small mesh."""
    rb = _ref()
    from b200rt import scenes
    soup = scenes.soup(300, seed=21)
    z = np.load(os.path.join(ROOT, "tests", "golden", "killeroo_mesh.npz"))
    for tv in (soup, load_golden("ref_feet.npz")["tri_verts"][:700]):
        tv = np.ascontiguousarray(tv, np.float32).reshape(-1, 9)
        S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3))
        ip = {k: v for k, v in params.items() if k != "emptybonus"}
        fp = {k: v for k, v in params.items() if k == "emptybonus"}
        ex = rb.RefAccel(S, "rbsp", ip, fp).export()
        t = hb(tv, **params)
        compare_with_reference(t, ex["nodes"], ex["indices"], params["nbDirections"])


def test_rbspkd_mode_reproduces_the_reference_tree(hb):
    """RBSPKd::buildTree (rbspKd.cpp:187-470): nbDirections + 100 in the host harness."""
    g = load_golden("ref_small_rbspkd13.npz")
    t = hb(g["tri_verts"], nbDirections=113)
    assert np.array_equal(t["nodes"], g["nodes"]) and np.array_equal(t["indices"], g["indices"])
    g = load_golden("ref_feet.npz")
    t = hb(g["tri_verts"], nbDirections=113)
    bad, n = compare_with_reference(t, g["rbspkd13_nodes"], g["rbspkd13_indices"], 13)
    assert bad == 0


def test_kd_mode_reproduces_the_reference_kdtree(hb, built_lib):
    """The builder's kd mode (nbDirections = 0: box areas in closed form instead of k-DOP cuts, KdTreeAccel::buildTree) against
    the reference-built golden and against this repo's host builder (itself byte-identical to the reference up to 2.9 M nodes,
    tests/test_living_room.py)."""
    g = load_golden("ref_small_kdtree.npz")
    t = hb(g["tri_verts"], nbDirections=0, traversalcost=1)
    assert np.array_equal(t["nodes"], g["nodes"]) and np.array_equal(t["indices"], g["indices"]) and np.array_equal(t["bounds"], g["bounds"])
    from b200rt import scenes
    tv = scenes.killeroo_simple()["tri_verts"][:20000]
    for kw in (dict(), dict(maxprims=4, emptybonus=0.5), dict(maxdepth=9, intersectcost=20, traversalcost=3)):
        h = built_lib.build_kdtree(tv, **{"traversalcost": 1, **kw})
        t = hb(tv, nbDirections=0, **{"traversalcost": 1, **kw})
        bad, n = compare_with_reference(t, h["nodes"], h["indices"], 3)
        assert bad == 0


@pytest.fixture(scope="module")
def hb_bvh(hb):
    L = C.CDLL(os.path.join(HB_DIR, "libhostbuild.so"))
    L.hb_build_bvh.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int, C.POINTER(C.POINTER(C.c_uint32)),
                               C.POINTER(C.c_uint32), C.POINTER(C.POINTER(C.c_uint32))]

    def build(tv, maxnodeprims=4, intersectcost=8, traversalcost=1):
        tv = np.ascontiguousarray(tv, np.float32).reshape(-1, 9)
        nodes, order, nn = C.POINTER(C.c_uint32)(), C.POINTER(C.c_uint32)(), C.c_uint32()
        assert L.hb_build_bvh(tv.ctypes.data, len(tv), maxnodeprims, intersectcost, traversalcost, C.byref(nodes), C.byref(nn), C.byref(order)) == 0
        N, O = np.ctypeslib.as_array(nodes, (nn.value, 8)).copy(), np.ctypeslib.as_array(order, (len(tv),)).copy()
        L.hb_free(nodes)
        L.hb_free(order)
        return N, O
    return build


def test_level_synchronous_bvh_build_is_byte_identical(hb_bvh, built_lib):
    """csrc/bvh_build.cuh (sorted once per axis, segmented box scans, std::partition's permutation from two scans) against the
    reference-built golden and against the serial host builder on meshes and soups, several parameter sets: the 32-byte nodes
    AND primNumMapping, byte for byte."""
    g = load_golden("ref_small_bvh.npz")
    N, O = hb_bvh(g["tri_verts"])
    assert np.array_equal(N, g["nodes"]) and np.array_equal(O, g["indices"])
    from b200rt import scenes
    cases = [(scenes.soup(3000, seed=3), {}), (scenes.killeroo_simple()["tri_verts"][:25000], {}),
             (scenes.killeroo_simple()["tri_verts"][:9000], dict(maxnodeprims=1)),
             (scenes.sponza_standin(detail=0.1)["tri_verts"], dict(maxnodeprims=8, intersectcost=20, traversalcost=3)),
             (np.tile(scenes.soup(1, seed=1), (40, 1)), {}),           # 40 identical triangles: every centroid ties, the primitive number decides
             (scenes.soup(1, seed=2), {}), (scenes.soup(2, seed=2), {})]
    for tv, kw in cases:
        h = built_lib.build_bvh(tv, **kw)
        N, O = hb_bvh(tv, **kw)
        assert np.array_equal(N, h["nodes"]) and np.array_equal(O, h["indices"]), (len(tv), kw)


def test_empty_and_tiny_inputs(hb):
    t = hb(np.zeros((0, 9), np.float32), nbDirections=7)
    assert t["nodes"].shape == (1, 2) and t["nodes"][0, 1] == 7 and len(t["indices"]) == 0      # one empty leaf: flags = M, np = 0
    one = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32)
    t = hb(one, nbDirections=3)
    assert t["nodes"].shape == (1, 2) and t["nodes"][0, 0] == 0 and t["nodes"][0, 1] == (3 | (1 << 2))


# ---- the product: the same build on the B200 ------------------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("M", [3, 7, 13])
def test_device_build_small(gpu, hb, M):
    g = load_golden("ref_small_rbsp%d.npz" % M)
    t = gpu.build_rbsp(g["tri_verts"], nbDirections=M)
    h = hb(g["tri_verts"], nbDirections=M)
    assert np.array_equal(t["nodes"], h["nodes"]) and np.array_equal(t["indices"], h["indices"])   # same kernels, same tie order
    assert np.array_equal(t["nodes"], g["nodes"]) and np.array_equal(t["indices"], g["indices"])   # = the reference's tree
    assert np.array_equal(t["directions"].view(np.uint32), g["directions"].view(np.uint32)) and np.array_equal(t["bounds"], g["bounds"])
    # and it traverses: the goldens' rays, the reference's answers
    acc = gpu.MakeAccelerator("rbsp", g["tri_verts"], {"nbDirections": M})
    hits = acc.Intersect(g["rays"])
    assert np.array_equal(hits["prim"], np.where(g["hit_prim"] < 0, gpu.MISS, g["hit_prim"]).astype(np.uint32))
    assert np.array_equal(hits["t"].view(np.uint32), g["hit_t"].view(np.uint32))
    assert np.array_equal(acc.IntersectP(g["rays"]), g["occluded"])


@pytest.mark.gpu
@pytest.mark.parametrize("M", [3, 7, 9, 13])
def test_device_build_feet(gpu, hb, M):
    g = load_golden("ref_feet.npz")
    t0 = time.time()
    t = gpu.build_rbsp(g["tri_verts"], nbDirections=M)
    wall = time.time() - t0
    h = hb(g["tri_verts"], nbDirections=M)
    assert np.array_equal(t["nodes"], h["nodes"]) and np.array_equal(t["indices"], h["indices"])
    bad, n = compare_with_reference(t, g["rbsp%d_nodes" % M], g["rbsp%d_indices" % M], M)
    print("\nfeet rbsp-%d on the B200: %d nodes in %.1f ms (%d levels, %d bound edges, %d launches; wall %.2f s); %d of %d leaves differ in membership"
          % (M, len(t["nodes"]), t["build"]["build_ms"], t["build"]["levels"], t["build"]["candidates"], t["build"]["kernel_launches"], wall, bad, n))
    acc = gpu.B200Accel(t, g["tri_verts"])
    hits = acc.Intersect(g["rays"])
    assert np.array_equal(hits["prim"], np.where(g["rbsp%d_prim" % M] < 0, gpu.MISS, g["rbsp%d_prim" % M]).astype(np.uint32))
    assert np.array_equal(hits["t"].view(np.uint32), g["rbsp%d_t" % M].view(np.uint32))
    assert np.array_equal(acc.IntersectP(g["rays"]), g["rbsp%d_occ" % M])


@pytest.mark.gpu
def test_device_build_rejects_what_it_does_not_do(gpu):
    with pytest.raises(gpu.B200RTError, match="nbDirections"):
        gpu.build_rbsp(np.zeros((4, 9), np.float32), nbDirections=5)
    t = gpu.build_rbsp(np.zeros((0, 9), np.float32), nbDirections=7)
    assert t["nodes"].shape == (1, 2) and t["nodes"][0, 1] == 7


@pytest.mark.gpu
def test_device_kdtree_build(gpu, hb):
    """b200rt_build_kdtree_device on scenes/killeroo-simple's 66 532 triangles: the host builder's (= the reference's) node
    array, the same leaves, and the same answers on rays."""
    from b200rt import scenes
    sc = scenes.killeroo_simple()
    tv = sc["tri_verts"]
    h = gpu.build_kdtree(tv)
    t = gpu.build_kdtree_device(tv)
    bad, n = compare_with_reference(t, h["nodes"], h["indices"], 3)
    assert bad == 0 and np.array_equal(t["bounds"], h["bounds"])
    print("\nkilleroo-simple kd-tree on the B200: %d nodes in %.1f ms (%d levels, %d launches)"
          % (len(t["nodes"]), t["build"]["build_ms"], t["build"]["levels"], t["build"]["kernel_launches"]))
    g = load_golden("ref_small_kdtree.npz")
    t = gpu.build_kdtree_device(g["tri_verts"])
    assert np.array_equal(t["nodes"], g["nodes"]) and np.array_equal(t["indices"], g["indices"])
    acc = gpu.B200Accel(t, g["tri_verts"])
    hits = acc.Intersect(g["rays"])
    assert np.array_equal(hits["prim"], np.where(g["hit_prim"] < 0, gpu.MISS, g["hit_prim"]).astype(np.uint32))
    assert np.array_equal(hits["t"].view(np.uint32), g["hit_t"].view(np.uint32))


@pytest.mark.gpu
def test_device_bvh_build(gpu):
    """b200rt_build_bvh_device: the host builder's (= the reference's) nodes and primNumMapping byte for byte, on the golden,
    on scenes/killeroo-simple and on the headline sponza stand-in."""
    from b200rt import scenes
    g = load_golden("ref_small_bvh.npz")
    t = gpu.build_bvh_device(g["tri_verts"])
    assert np.array_equal(t["nodes"], g["nodes"]) and np.array_equal(t["indices"], g["indices"])
    for name, tv, kw in [("killeroo-simple", scenes.killeroo_simple()["tri_verts"], {}),
                         ("sponza stand-in", scenes.sponza_standin()["tri_verts"], {}),
                         ("killeroo-simple, maxnodeprims 1", scenes.killeroo_simple()["tri_verts"], dict(maxnodeprims=1))]:
        t0 = time.time()
        h = gpu.build_bvh(tv, **kw)
        host_s = time.time() - t0
        t = gpu.build_bvh_device(tv, **kw)
        assert np.array_equal(t["nodes"], h["nodes"]) and np.array_equal(t["indices"], h["indices"]), name
        print("\n%s: %d triangles, %d nodes: %.1f ms on the B200 (%d levels, %d launches); serial host builder %.0f ms"
              % (name, len(tv), len(t["nodes"]), t["build"]["build_ms"], t["build"]["levels"], t["build"]["kernel_launches"], 1e3 * host_s))


@pytest.mark.gpu
def test_device_rbspkd_build(gpu):
    g = load_golden("ref_feet.npz")
    t = gpu.build_rbsp(g["tri_verts"], nbDirections=13, kdtraversalcost=1)
    bad, n = compare_with_reference(t, g["rbspkd13_nodes"], g["rbspkd13_indices"], 13)
    assert bad == 0 and t["kind"] == "rbspkd"
    acc = gpu.MakeAccelerator("rbspkd", g["tri_verts"], {"nbDirections": 13})
    hits = acc.Intersect(g["rays"])
    assert np.array_equal(hits["prim"], np.where(g["rbspkd13_prim"] < 0, gpu.MISS, g["rbspkd13_prim"]).astype(np.uint32))
    assert np.array_equal(hits["t"].view(np.uint32), g["rbspkd13_t"].view(np.uint32))
    assert np.array_equal(acc.IntersectP(g["rays"]), g["rbspkd13_occ"])
