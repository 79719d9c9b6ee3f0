"""The attach mechanism of SURVEY.md §8(b), end to end: oracle/_ref/pbrt_b200 is the reference's whole scene-description
layer (parser, api, every plugin — compiled verbatim by oracle/ref/Makefile) with src/core/api.cpp passed through the scripted
two-hunk overlay (oracle/ref/make_api_overlay.py) and linked with the product's C++ drop-in (thesis-pbrt-v3_b200/host) and
libb200rt.so.  The tests write scenes/killeroo-simple as the reference tree has it (tests/pbrt_scenes.py), instantiate its
`Accelerator $acc` placeholder the way scripts/render_simple.sh does, and run the binary through pbrtParseFile
(src/core/parser.cpp:1094):

  * CPU: with `Accelerator "bvh"` / `Integrator "path"` the binary IS the reference renderer; `Accelerator "b200"` is a known
    name (no "unknown" warning) that fails loudly without a CUDA device and lets MakeScene fall back (api.cpp:1886);
  * GPU: `Accelerator "b200" "string tree" "bvh" | "kdtree"` + `Integrator "b200wavefront"` render the same file on the B200:
    the image equals the stock run's (the reference's own CPU render of the same parsed scene, Halton 8 spp, area light,
    plastic killeroos with their loop-subdivision vertex normals, uv-mapped ground) to <= 1e-4 relative RMSE per pixel with
    < 1e-3 of the pixels off, and unused parameters are reported as api.cpp:1013 does."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PBRT = os.path.join(ROOT, "oracle", "_ref", "pbrt_b200")
pytestmark = pytest.mark.skipif(not os.path.exists(PBRT), reason="oracle/_ref/pbrt_b200 is not built (make -C oracle/ref pbrt)")


def _run(scene, cwd, *args, quiet=True):
    p = subprocess.run([PBRT] + (["--quiet"] if quiet else []) + list(args) + [scene], cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return p.returncode, p.stdout


def test_stock_names_are_the_reference_and_b200_is_a_known_name(tmp_path):
    import pbrt_scenes as ps
    d = str(tmp_path)
    scene = ps.killeroo_simple(d, 'Accelerator "bvh" "integer nbDirections" [3]', xres=64, yres=64, spp=2, filename="stock.pfm")
    rc, out = _run(scene, d, "--nthreads", "4")
    assert rc == 0 and os.path.exists(os.path.join(d, "stock.pfm")), out
    img = ps.read_pfm(os.path.join(d, "stock.pfm"))
    assert img.shape == (64, 64, 3) and img.max() > 0
    # "Regular ray intersection tests" etc. come from the reference's own statistics block
    assert "Shapes created" in out and "66533" in out
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        scene = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"', xres=32, yres=32, spp=1, filename="nogpu.pfm")
        rc, out = _run(scene, d, "--nthreads", "2")
        assert 'Accelerator "b200" unknown' not in out          # the overlay's name branch is there ...
        assert "b200" in out and ("device" in out.lower() or "cuda" in out.lower()), out   # ... and fails loudly: no CPU path
    scene = ps.killeroo_simple(d, 'Accelerator "nosuch"', xres=32, yres=32, spp=1, filename="unk.pfm")
    rc, out = _run(scene, d, "--nthreads", "2", quiet=False)   # Warning() is silent under --quiet (error.cpp)
    assert 'Accelerator "nosuch" unknown' in out             # api.cpp:1011-1012 untouched


@pytest.mark.gpu
@pytest.mark.parametrize("tree", ["bvh", "kdtree"])
def test_parsed_scene_renders_on_the_b200(tmp_path, tree):
    import pbrt_scenes as ps
    import hostwf
    d = str(tmp_path)
    scene = ps.killeroo_simple(d, 'Accelerator "%s" "integer nbDirections" [3]' % tree, filename="stock.pfm")
    rc, out = _run(scene, d)
    assert rc == 0, out
    ref = ps.read_pfm(os.path.join(d, "stock.pfm"))
    scene = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "%s" "integer bogus" [1]' % tree,
                               integrator='Integrator "b200wavefront" "integer maxdepth" [5]', filename="b200.pfm")
    rc, out2 = _run(scene, d, quiet=False)
    assert rc == 0 and os.path.exists(os.path.join(d, "b200.pfm")), out2
    assert 'Parameter "bogus" not used' in out2, out2
    img = ps.read_pfm(os.path.join(d, "b200.pfm"))
    st = hostwf.image_stats(img, ref)
    print("\nparsed killeroo-simple, tree %s: %s" % (tree, st))
    assert img.shape == ref.shape == (700, 700, 3)
    assert st["rel_rmse_excl"] <= 1e-4 and st["frac_differing"] < 1e-3, st


@pytest.mark.gpu
def test_two_ranks_write_one_image(tmp_path):
    """`Integrator "b200wavefront" "integer world" [2]`: two pbrt processes (here sharing the one GPU of the box), each rendering
    the tiles t % 2 == rank; b200rt_film_reduce sums rank 1's film into rank 0's over CUDA IPC and rank 0 alone writes the
    image, which equals the one-process image (boundary-sample halos are the only pixels summed in another order)."""
    import pbrt_scenes as ps
    d = str(tmp_path)
    rdv = os.path.join(d, "rdv")
    os.makedirs(rdv)
    kw = dict(xres=256, yres=256, spp=4)
    one = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"', integrator='Integrator "b200wavefront"', filename="one.pfm",
                             scene_name="one.pbrt", **kw)
    rc, out = _run(one, d)
    assert rc == 0, out
    procs = []
    for r in (1, 0):
        sc = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"',
                                integrator='Integrator "b200wavefront" "integer rank" [%d] "integer world" [2] "string rendezvous" "%s"' % (r, rdv),
                                filename="two_rank%d.pfm" % r, scene_name="two_rank%d.pbrt" % r, **kw)
        procs.append(subprocess.Popen([PBRT, "--quiet", sc], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert os.path.exists(os.path.join(d, "two_rank0.pfm")) and not os.path.exists(os.path.join(d, "two_rank1.pfm")), outs
    a, b = ps.read_pfm(os.path.join(d, "one.pfm")), ps.read_pfm(os.path.join(d, "two_rank0.pfm"))
    same = np.all(a == b, axis=-1).mean()
    print("\ntwo ranks vs one: %.5f of the pixels bit-identical, max abs diff %.3g" % (same, np.abs(a - b).max()))
    assert same > 0.98 and np.allclose(a, b, rtol=2e-6, atol=1e-7)
    assert os.listdir(rdv) == []     # the rendezvous files are consumed

    # world > 1 without a rendezvous directory is an error, not a partial image
    sc = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"', integrator='Integrator "b200wavefront" "integer world" [2]',
                            filename="bad.pfm", scene_name="bad.pbrt", **kw)
    rc, out = _run(sc, d, quiet=False)
    assert "rendezvous" in out and not os.path.exists(os.path.join(d, "bad.pfm")), out


@pytest.mark.gpu
@pytest.mark.parametrize("trilinear", [True, False])
def test_parsed_scene_with_an_image_texture(tmp_path, trilinear):
    """`Texture ... "imagemap" ... "bool trilinear" "true"` + `Material "matte" "texture Kd"` through pbrtParseFile: the C++
    integrator hands the reference's own MIPMap pyramid to the device (ABI v9), the MIP level comes from the camera rays'
    differentials; against the stock run of the same file."""
    import pbrt_scenes as ps
    import hostwf
    from conftest import load_golden
    from b200rt import textures
    d = str(tmp_path)
    textures.write_tga(os.path.join(d, "leaf.tga"), load_golden("living_room_textures.npz")["leaf"])
    kw = dict(xres=384, yres=384, spp=8, ground_texture="leaf.tga", trilinear=trilinear)   # False: the ImageTexture default, EWA
    scene = ps.killeroo_simple(d, 'Accelerator "bvh"', filename="stock.pfm", **kw)
    rc, out = _run(scene, d)
    assert rc == 0, out
    ref = ps.read_pfm(os.path.join(d, "stock.pfm"))
    scene = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"', integrator='Integrator "b200wavefront" "integer maxdepth" [5]',
                               filename="b200.pfm", **kw)
    rc, out2 = _run(scene, d)
    assert rc == 0 and os.path.exists(os.path.join(d, "b200.pfm")), out2
    img = ps.read_pfm(os.path.join(d, "b200.pfm"))
    st = hostwf.image_stats(img, ref)
    plain = ps.read_pfm(os.path.join(d, "stock.pfm"))
    print("\nparsed killeroo-simple with a textured ground: %s" % st)
    assert st["rel_rmse_excl"] <= 1e-4 and st["frac_differing"] < 1e-3, st
    # and the texture is really there: the ground of the plain scene is a constant (.5 .5 .8)
    scene = ps.killeroo_simple(d, 'Accelerator "bvh"', filename="plain.pfm", xres=384, yres=384, spp=8)
    rc, _ = _run(scene, d)
    assert rc == 0 and np.abs(ps.read_pfm(os.path.join(d, "plain.pfm")) - ref).max() > 0.05


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", ["uniform", "power", "spatial", None])
def test_parsed_scene_with_four_lights(tmp_path, strategy):
    """Several LightSource statements (point, distant, constant infinite) next to the file's sphere area light, through
    pbrtParseFile: UniformSampleOneLight over `"string lightsamplestrategy" "uniform"` / `"power"` (ABI v10: the C++ integrator
    describes every light of Scene::lights and, for "power", weighs them by the reference's own Light::Power().y()), `"spatial"`
    and no strategy at all (= "spatial", the path integrator's default: one distribution per voxel of the scene's bound) against
    the stock run of the same file with the same strategy."""
    import pbrt_scenes as ps
    import hostwf
    d = str(tmp_path)
    kw = dict(xres=384, yres=384, spp=8, more_lights=ps.MORE_LIGHTS)
    integ = '"integer maxdepth" [5]' + ('' if strategy is None else ' "string lightsamplestrategy" "%s"' % strategy)
    scene = ps.killeroo_simple(d, 'Accelerator "bvh"', integrator='Integrator "path" ' + integ, filename="stock.pfm", **kw)
    rc, out = _run(scene, d)
    assert rc == 0, out
    ref = ps.read_pfm(os.path.join(d, "stock.pfm"))
    scene = ps.killeroo_simple(d, 'Accelerator "b200" "string tree" "bvh"', integrator='Integrator "b200wavefront" ' + integ,
                               filename="b200.pfm", **kw)
    rc, out2 = _run(scene, d)
    assert rc == 0 and os.path.exists(os.path.join(d, "b200.pfm")), out2
    st = hostwf.image_stats(ps.read_pfm(os.path.join(d, "b200.pfm")), ref)
    print("\nparsed killeroo-simple with four lights, %s: %s" % (strategy, st))
    assert st["rel_rmse_excl"] <= 1e-4 and st["frac_differing"] < 1e-3, st
    # the lights are really there: the file's own single light gives another picture
    scene = ps.killeroo_simple(d, 'Accelerator "bvh"', filename="plain.pfm", xres=384, yres=384, spp=8)
    rc, _ = _run(scene, d)
    assert rc == 0 and np.abs(ps.read_pfm(os.path.join(d, "plain.pfm")) - ref).max() > 0.05
