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
