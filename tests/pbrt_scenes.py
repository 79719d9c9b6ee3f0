"""TEST INFRASTRUCTURE.  Writes the thesis's scene templates as .pbrt files a pbrt binary can parse on a machine that does
not have the reference tree: the directives of scenes/killeroo-simple and scenes/killeroo-feet as the files have them
(camera, film, sampler, lights, materials, transforms), the `$acc $accnr` placeholder of their Accelerator line instantiated
the way scripts/render_simple.sh:24-29 does it, and `Include`-ing scenes/geometry/killeroo.pbrt regenerated from the
committed fixture (tests/golden/killeroo_mesh.npz: the file's own `Shape "loopsubdiv"` control mesh)."""
import os
import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def write_killeroo_geometry(path):
    z = np.load(os.path.join(GOLDEN, "killeroo_mesh.npz"))
    with open(path, "w") as f:
        f.write('    Shape "loopsubdiv"  "integer nlevels" [%d ] "integer indices" [' % int(z["ctrl_levels"]))
        f.write(" ".join(str(int(i)) for i in z["ctrl_idx"].ravel()))
        f.write(' ] "point P" [')
        f.write(" ".join(repr(float(v)) for v in z["ctrl_P"].astype(np.float64).ravel()))
        f.write(" ]\n")


def killeroo_simple(dirname, accelerator, integrator='Integrator "path" "integer maxdepth" [5]', xres=700, yres=700, spp=8,
                    filename="killeroo-simple.pfm", extra_film="", scene_name="input.pbrt", ground_texture=None, more_lights="", trilinear=True):
    """scenes/killeroo-simple with its Accelerator line replaced by `accelerator` (a full directive) and, optionally, its
    Integrator line by `integrator`; `more_lights` = LightSource directives placed right after WorldBegin (before the file's own
    area light, so they come first in Scene::lights).  Returns the scene file's path."""
    geo = os.path.join(dirname, "killeroo.pbrt")
    if not os.path.exists(geo):
        write_killeroo_geometry(geo)
    ground = 'Material "matte" "color Kd" [.5 .5 .8]'
    if ground_texture is not None:
        # the ground (uv 0..5: five repeats) with an image texture on its Kd, the way scenes/sponza and scenes/living-room
        # declare theirs: `Texture "name" "spectrum" "imagemap" "string filename" ... "bool trilinear" "true"`
        # (trilinear=False: no "bool trilinear" at all, the way scenes/sponza declares its textures = the EWA filter)
        ground = ('Texture "groundtex" "spectrum" "imagemap" "string filename" "%s"%s\n'
                  '  Material "matte" "texture Kd" "groundtex"' % (ground_texture, ' "bool trilinear" "true"' if trilinear else ""))
    txt = '''LookAt 400 20 30   0 63 -110   0 0 1
Rotate -5 0 0 1
Camera "perspective" "float fov" [39]
Film "image"
"integer xresolution" [%d] "integer yresolution" [%d]
    "string filename" "%s" %s

Sampler "halton" "integer pixelsamples" [%d]
%s

%s

WorldBegin
%s
AttributeBegin
Material "matte" "color Kd" [0 0 0]
Translate 150 0  20
Translate 0 120 0
AreaLightSource "area"  "color L" [2000 2000 2000] "integer nsamples" [8]
Shape "sphere" "float radius" [3]
AttributeEnd


AttributeBegin
  %s
  Translate 0 0 -140
Shape "trianglemesh" "point P" [ -1000 -1000 0 1000 -1000 0 1000 1000 0 -1000 1000 0 ]
      "float uv" [ 0 0 5 0 5 5 0 5 ]
	"integer indices" [ 0 1 2 2 3 0]
Shape "trianglemesh" "point P" [ -400 -1000 -1000   -400 1000 -1000   -400 1000 1000 -400 -1000 1000 ]
      "float uv" [ 0 0 5 0 5 5 0 5 ]
        "integer indices" [ 0 1 2 2 3 0]
AttributeEnd

AttributeBegin
Scale .5 .5 .5
Rotate -60 0 0 1
    Material "plastic" "color Kd" [.4 .2 .2] "color Ks" [.5 .5 .5]
        "float roughness" [.025]
Translate 100 200 -140
    Include "%s"
    Material "plastic" "color Ks" [.3 .3 .3] "color Kd" [.4 .5 .4]
        "float roughness" [.15]
Translate -200 0 0
    Include "%s"

AttributeEnd
WorldEnd
''' % (xres, yres, filename, extra_film, spp, accelerator, integrator, more_lights, ground, geo, geo)
    path = os.path.join(dirname, scene_name)
    with open(path, "w") as f:
        f.write(txt)
    return path


# a point, a distant and a constant infinite light next to killeroo-simple's sphere area light: every light kind in one scene
MORE_LIGHTS = '''LightSource "point" "point from" [100 -150 200] "rgb I" [60000 50000 40000]
LightSource "distant" "point from" [-0.3 0.4 1] "point to" [0 0 0] "rgb L" [1.5 1.5 1.2]
AttributeBegin
Rotate 40 0 0 1
LightSource "infinite" "rgb L" [.3 .35 .5]
AttributeEnd
'''


def read_pfm(path):
    """(h, w, 3) float32, top row first (pbrt's WriteImagePFM writes the rows bottom-up, imageio.cpp:437-480)."""
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4")
    nc = 3 if kind == b"PF" else 1
    return data.reshape(h, w, nc)[::-1].astype(np.float32)
