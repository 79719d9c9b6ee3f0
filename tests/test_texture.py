"""`imagemap` textures (ABI v9): the product's MIPMap pyramid (b200rt_mipmap_build, host C++) and trilinear lookup (wf_texture.cuh,
compiled for the host by tests/hostshade) against the reference's own ImageTexture / MIPMap on the two image files scenes/living-room
ships, written back from the committed 8-bit fixture (tests/golden/living_room_textures.npz) into .tga files that the reference
reads with its own ReadImage.  Bit for bit."""
import ctypes as C
import os
import numpy as np
import pytest
from conftest import ROOT, load_golden


def _ref():
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libpbrtref.so")):
        pytest.skip("oracle/_ref is not built")
    from oracle import refbind
    return refbind


@pytest.mark.parametrize("name", ["leaf", "picture8"])
def test_pyramid_equals_the_references_mipmap(built_lib, tmp_path, name):
    rb = _ref()
    from b200rt import textures
    px = load_golden("living_room_textures.npz")[name]
    path = str(tmp_path / (name + ".tga"))
    textures.write_tga(path, px)
    assert np.array_equal(textures.read_tga(path), px)
    ref = rb.texture_pyramid(path)
    t = textures.ImageTexture(rgb8=px)
    assert t.n_levels == len(ref) and (t.height, t.width) == ref[0].shape[:2]
    for l, r in enumerate(ref):
        assert np.array_equal(t.level(l).view(np.uint32), r.view(np.uint32)), l      # every texel of every level
    # 365 x 271 -> 512 x 512, 1280 x 853 -> 2048 x 1024: the Lanczos resampling path (mipmap.h:123-176)
    assert (t.width, t.height) == ((512, 512) if name == "leaf" else (2048, 1024))


def test_missing_file_is_the_references_constant_grey(built_lib, tmp_path):
    """textures/wood.tga and wood5.tga are absent from the reference tree: ImageTexture::GetTexture replaces them by ONE texel 0.5
    (imagemap.cpp:67-72), which convertIn gamma-decodes — the constant scenes.living_room gives Floor and TableWood."""
    rb = _ref()
    from b200rt import textures, scenes
    ref = rb.texture_pyramid(str(tmp_path / "no_such_file.tga"))
    assert len(ref) == 1 and ref[0].shape == (1, 1, 3)
    t = textures.ImageTexture(rgb=np.full((1, 1, 3), 0.5, np.float32))
    assert np.array_equal(t.level(0).view(np.uint32), ref[0].view(np.uint32))
    lr = scenes.living_room(64, 36)
    names = lr["material_names"]
    for m in ("Floor", "TableWood"):
        assert np.array_equal(lr["materials"][names.index(m), 1:4].view(np.uint32), ref[0].reshape(3).view(np.uint32))


def test_trilinear_lookup_equals_the_references(built_lib, tmp_path):
    rb = _ref()
    import hostwf
    from b200rt import textures
    px = load_golden("living_room_textures.npz")["leaf"]
    path = str(tmp_path / "leaf.tga")
    textures.write_tga(path, px)
    t = textures.ImageTexture(rgb8=px)
    rng = np.random.default_rng(5)
    n = 20000
    stw = np.zeros((n, 3), np.float32)
    stw[:, :2] = rng.uniform(-1.5, 2.5, (n, 2))                      # Repeat wrap on both sides
    stw[:, 2] = np.where(rng.random(n) < 0.25, 0.0, 10.0 ** rng.uniform(-9, 0.5, n))   # width 0 (bounce hits), tiny .. beyond the coarsest level
    want = rb.texture_lookup(path, stw)
    L = hostwf.lib()
    L.hs_tex_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    got = np.zeros((n, 3), np.float32)
    rec = t.record()
    L.hs_tex_lookup(C.byref(rec), stw.ctypes.data, n, got.ctypes.data)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    L.hs_log_check.restype = C.c_uint64
    L.hs_log_check.argtypes = [C.c_uint32] * 3
    assert L.hs_log_check(0x00000001, 0x7f800000, 257) == 0          # log_ref == this libm's logf (2^23 samples of all positive floats)
    assert L.hs_log_check(0x30000000, 0x41800000, 11) == 0           # and densely over [1e-9, 16)


@pytest.mark.parametrize("maxaniso", [8.0, 2.0])
def test_ewa_lookup_equals_the_references(built_lib, tmp_path, maxaniso):
    """The ImageTexture default, "bool trilinear" false: MIPMap::Lookup(st, dst0, dst1) -> EWA over two levels (mipmap.h:277-353),
    with its eccentricity clamp ("float maxanisotropy"), its Gaussian weight table and the degenerate cases (a zero minor axis ->
    the bilinear `triangle`, ellipses wider than the coarsest level)."""
    rb = _ref()
    import hostwf
    from b200rt import api, textures
    px = load_golden("living_room_textures.npz")["picture8"]
    path = str(tmp_path / "picture8.tga")
    textures.write_tga(path, px)
    t = textures.ImageTexture(rgb8=px, trilinear=False, maxanisotropy=maxaniso)
    rng = np.random.default_rng(7)
    n = 20000
    q = np.zeros((n, 6), np.float32)
    q[:, :2] = rng.uniform(-1.5, 2.5, (n, 2))
    scale = 10.0 ** rng.uniform(-6, -0.3, (n, 1))                      # sub-texel .. a third of the image
    q[:, 2:] = rng.standard_normal((n, 4)) * scale * np.array([1, 1, 0.2, 0.2])   # often anisotropic beyond the clamp
    q[::17, 4:] = 0                                                     # minor axis exactly zero
    q[::19, 2:] = 0                                                     # no differentials at all (hits past the camera ray)
    rb.set_texture_filter(False, maxaniso)
    try:
        want = rb.texture_lookup_diff(path, q)
    finally:
        rb.set_texture_filter(True, 8.0)
    L = hostwf.lib()
    L.hs_tex_lookup_diff.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    got = np.zeros((n, 3), np.float32)
    rec = t.record()
    L.hs_tex_lookup_diff(C.byref(rec), q.ctypes.data, n, got.ctypes.data)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), int((got.view(np.uint32) != want.view(np.uint32)).any(1).sum())
    # the weight table itself, against the one the reference's MIPMap holds (exp through this machine's libm on both sides)
    lut = np.zeros(128, np.float32)
    api.lib().b200rt_ewa_weights.argtypes = [C.c_void_p]
    api.lib().b200rt_ewa_weights(lut.ctypes.data)
    assert lut[0] > 0.86 and lut[127] == 0.0 and np.all(np.diff(lut) < 0)
