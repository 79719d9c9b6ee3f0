"""Host-side scene descriptions (no GPU): the stand-ins are deterministic, and the material tables the shade stage is given
follow the C ABI's record layout (include/b200rt_wavefront.h, ABI v7)."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
from b200rt import scenes  # noqa: E402

KINDS = {0: "matte", 1: "plastic", 2: "mirror", 3: "glass", 4: "metal", 5: "substrate", 6: "uber"}


def _check_records(m):
    m = np.asarray(m, np.float32)
    assert m.ndim == 2 and m.shape[1] in (8, 20)     # material_stride: the short record, or the long one "uber" needs
    for r in m:
        kf = int(r[0])
        assert kf == r[0] and 0 <= kf < 64 and (kf & 15) in KINDS, r
        kind = kf & 15
        if kind == 3:
            assert r[7] > 0          # index of refraction
        if kind in (4, 5):
            assert r[7] > 0          # roughness
        if kind in (0, 1, 2, 5):
            assert (r[1:4] >= 0).all() and (r[1:4] <= 1).all()
        if kind == 0:
            assert 0 <= r[4] <= 90   # sigma in degrees
        if kind == 6:
            assert len(r) == 20 and r[7] > 0 and r[17] > 0 and (r[8:17] >= 0).all() and (r[8:17] <= 1).all()   # Kr, Kt, opacity; eta


def test_museum_standin_is_deterministic_and_holds_the_templates_camera():
    a, b = scenes.museum_standin(detail=0.05, xres=64, yres=48), scenes.museum_standin(detail=0.05, xres=64, yres=48)
    assert np.array_equal(a["tri_verts"], b["tri_verts"]) and a["tri_verts"].dtype == np.float32
    assert np.isfinite(a["tri_verts"]).all() and len(a["tri_verts"]) > 5000
    lo, hi = a["tri_verts"].reshape(-1, 3).min(0), a["tri_verts"].reshape(-1, 3).max(0)
    eye = np.array(a["camera"]["eye"])
    assert (eye > lo).all() and (eye < hi).all() and (a["light"] > lo).all() and (a["light"] < hi).all()   # both inside the hall
    assert np.allclose(eye, [0.4, 0.06, -0.05]) and a["maxdepth"] == 1 and a["spp"] == 64   # scenes/museum:1-3,12,15
    d = np.array(a["camera"]["dir"])
    want = np.array([-0.2, 0.1, 0.05]) - eye
    assert np.allclose(d, want / np.linalg.norm(want), atol=1e-6)
    full = scenes.museum_standin()
    assert 200_000 < len(full["tri_verts"]) < 320_000 and "STAND-IN" in full["note"]
    assert hashlib.sha256(full["tri_verts"].tobytes()).hexdigest() == hashlib.sha256(scenes.museum_standin()["tri_verts"].tobytes()).hexdigest()


def test_material_tables_follow_the_abi_record_layout():
    sc = scenes.killeroo_simple(xres=16, yres=16)
    for m in (sc["materials"], scenes.specular_materials(sc), scenes.metal_materials(sc), scenes.metal_materials(sc, oren_nayar=True)):
        _check_records(m)
        assert len(m) == len(sc["materials"])
    sp = scenes.specular_materials(sc)
    assert int(sp[1, 0]) == 3 and sp[1, 7] == 1.5 and int(sp[2, 0]) == 2
    me = scenes.metal_materials(sc)
    assert int(me[1, 0]) == 4 + 32 and int(me[2, 0]) == 5 + 32 and int(me[0, 0]) == 5   # remaproughness false / true
    assert scenes.metal_materials(sc, oren_nayar=True)[0, 4] == 40


def test_living_room_materials_are_the_files_kinds():
    lr = scenes.living_room(64, 36)
    _check_records(lr["materials"])
    _check_records(lr["legacy_materials"])
    kinds = dict(zip(lr["material_names"], lr["materials"][:, 0].astype(int)))
    assert kinds["Mirror"] == 2 and kinds["FireplaceGlass"] == 2 and kinds["Glass"] == 3
    assert kinds["BrushedStainlessSteel"] == 4 + 32 and kinds["Paneling"] == 5 + 32 and kinds["Floor"] == 5 + 32
    assert kinds["Sofa"] == 0 and kinds["Leaves"] == 6      # scene.pbrt:29 "string type" [ "uber" ]
    names = lr["material_names"]
    # textures/wood.tga and wood5.tga are absent from the reference tree: the reference resolves their Kd to one grey texel 0.5,
    # gamma-decoded (imagemap.cpp:67-72; pinned against the reference in tests/test_texture.py)
    grey = np.array([0x3e5b2d9a], np.uint32).view(np.float32)[0]
    assert (lr["materials"][names.index("Floor"), 1:4] == grey).all() and (lr["materials"][names.index("TableWood"), 1:4] == grey).all()
    # the files that are there: picture8.tga on the Painting's Kd, leaf.tga on the Leaves' Kd and opacity (ABI v9 slots 18 / 19)
    assert lr["texture_names"] == ["leaf", "picture8"] and [p.shape for p in lr["texture_pixels"]] == [(271, 365, 3), (853, 1280, 3)]
    assert lr["materials"][names.index("Painting"), 18] == 2 and tuple(lr["materials"][names.index("Leaves"), 18:20]) == (1, 1)
    plain = scenes.living_room(64, 36, textures=False)
    assert (plain["materials"][:, 18:20] == 0).all() and plain["texture_pixels"] is None
    assert lr["tri_normals"].shape == (143163, 9) and np.bincount(lr["prim_flags"]).tolist() == [3008, 0, 140155]
    assert lr["materials"][names.index("PlantPot"), 4] == 1          # scene.pbrt:26 "float sigma" [1]
    assert np.allclose(lr["materials"][names.index("BrushedStainlessSteel"), 1:8],
                       [1.657460, 0.880369, 0.521229, 9.223869, 6.269523, 4.837001, .02])   # scene.pbrt:21
    assert set(np.unique(lr["legacy_materials"][:, 0])) <= {0.0, 1.0}
    assert lr["prim_material"].max() < len(lr["materials"]) and len(lr["prim_material"]) == 143163


def test_conference_standin_is_deterministic_and_sized():
    """BASELINE config 3 names conference next to museum; neither geometry is in the reference tree."""
    from b200rt import scenes
    a, b = scenes.conference_standin(xres=64, yres=64), scenes.conference_standin(xres=64, yres=64)
    assert np.array_equal(a["tri_verts"], b["tri_verts"])
    assert 250_000 < len(a["tri_verts"]) < 320_000 and np.isfinite(a["tri_verts"]).all()
    assert a["maxdepth"] == 1 and "STAND-IN" in a["note"]
