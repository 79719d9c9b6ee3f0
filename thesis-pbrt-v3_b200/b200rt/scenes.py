"""Scene instantiation for the configs of BASELINE.json.

* killeroo-simple (reference scenes/killeroo-simple): the loop-subdivided killeroo comes from the
  committed fixture tests/golden/killeroo_mesh.npz (generated from the reference's own
  scenes/geometry/killeroo.pbrt by tests/golden/make_golden.py through the reference's
  LoopSubdivide); transforms, ground quads, camera and film follow the scene file.  The sphere area
  light (scenes/killeroo-simple:20-24) is not a triangle and is left out of the aggregate.
* sponza / museum / conference: their geometry is NOT in the reference tree, so these
  are documented PROCEDURAL STAND-INS (fixed seed, stated triangle count); every number reported on
  them says so.  Cameras/lights follow the scene templates (scenes/sponza:1-2,16).

Everything here is host-side numpy; output = world-space triangle soup (n,9) float32 + camera."""
import os
import numpy as np

_GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests", "golden")


# ---- tiny mesh toolkit -----------------------------------------------------------------------
def _grid(nu, nv):
    """Index list (two triangles per cell) of an (nu+1) x (nv+1) vertex grid."""
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * (nv + 1) + j).ravel()
    b, c, d = a + (nv + 1), a + (nv + 1) + 1, a + 1
    return np.stack([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], 1).reshape(-1, 3)


def _surface(fn, nu, nv):
    u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
    P = fn(u.ravel(), v.ravel())
    return np.asarray(P, np.float64).reshape(-1, 3)[_grid(nu, nv)].reshape(-1, 9)


def _box(lo, hi, n=1):
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    out = []
    for ax in range(3):
        a, b = (ax + 1) % 3, (ax + 2) % 3
        for side in (lo[ax], hi[ax]):
            def f(u, v, ax=ax, a=a, b=b, side=side):
                P = np.zeros((len(u), 3))
                P[:, ax] = side
                P[:, a] = lo[a] + u * (hi[a] - lo[a])
                P[:, b] = lo[b] + v * (hi[b] - lo[b])
                return P
            out.append(_surface(f, n, n))
    return np.concatenate(out)


def _lathe(profile, cx, cz, y0, nseg, nring):
    """Surface of revolution around the vertical axis; profile(v) -> (radius, height)."""
    def f(u, v):
        r, h = profile(v)
        ang = 2 * np.pi * u
        return np.stack([cx + r * np.cos(ang), y0 + h, cz + r * np.sin(ang)], 1)
    return _surface(f, nseg, nring)


def look_at_camera(eye, look, up, fov_deg, xres, yres):
    """Perspective camera of the scene files (pbrt LookAt + `Camera "perspective" "float fov"`):
    fov spans the SHORTER image axis (reference src/cameras/perspective.cpp:240-262)."""
    eye, look, up = (np.asarray(a, np.float64) for a in (eye, look, up))
    d = look - eye
    d /= np.linalg.norm(d)
    right = np.cross(up / np.linalg.norm(up), d)
    right /= np.linalg.norm(right)
    newup = np.cross(d, right)
    aspect = xres / yres
    t = np.tan(np.radians(fov_deg) / 2)
    sx, sy = (aspect, 1.0) if aspect > 1 else (1.0, 1 / aspect)
    # "lookat" / "fov": what WavefrontPathIntegrator turns into the reference camera's own matrices (camera rays bit-identical
    # to PerspectiveCamera::GenerateRayDifferential); eye / dir / right / up: the pinhole frame for host-side ray generators
    return dict(eye=eye.astype(np.float32), dir=d.astype(np.float32), right=(right * t * sx).astype(np.float32),
                up=(newup * t * sy).astype(np.float32), xres=int(xres), yres=int(yres),
                lookat=np.concatenate([eye, look, up]).astype(np.float32), fov=float(fov_deg))


def camera_rays(cam, spp=1, seed=0, x0=0, x1=None, y0=0, y1=None):
    """(n,8) float32 primary rays, row-major over the pixel window, spp jittered samples per pixel."""
    x1 = cam["xres"] if x1 is None else x1
    y1 = cam["yres"] if y1 is None else y1
    rng = np.random.default_rng(seed)
    px, py = np.meshgrid(np.arange(x0, x1), np.arange(y0, y1), indexing="xy")
    px = np.repeat(px.ravel(), spp).astype(np.float64)
    py = np.repeat(py.ravel(), spp).astype(np.float64)
    jx, jy = (rng.random(len(px)), rng.random(len(px))) if spp > 1 or seed else (0.5, 0.5)
    sx = 2 * (px + jx) / cam["xres"] - 1   # raster x grows along camera +x = `right` = up x dir (perspective.cpp:61-66; pinned by tests/golden/ref_render.npz)
    sy = 1 - 2 * (py + jy) / cam["yres"]
    d = cam["dir"][None, :] + sx[:, None] * cam["right"][None, :] + sy[:, None] * cam["up"][None, :]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((len(px), 8), np.float32)
    r[:, 0:3] = cam["eye"]
    r[:, 3] = np.inf
    r[:, 4:7] = d
    return r


def random_rays(bounds, n, seed=12345):
    """Rrand of SURVEY.md §8(d): origins uniform in the root bounds, directions uniform on the sphere."""
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(bounds[:3], np.float64), np.asarray(bounds[3:], np.float64)
    o = lo + rng.random((n, 3)) * (hi - lo)
    z = 1 - 2 * rng.random(n)
    phi = 2 * np.pi * rng.random(n)
    s = np.sqrt(np.maximum(0, 1 - z * z))
    d = np.stack([s * np.cos(phi), s * np.sin(phi), z], 1)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = o
    r[:, 3] = np.inf
    r[:, 4:7] = d
    return r


# ---- killeroo-simple -------------------------------------------------------------------------
def _rot_z(deg):
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    return np.array([[c, -s, 0, 0], [s, c, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], float)


def _translate(x, y, z):
    m = np.eye(4)
    m[:3, 3] = (x, y, z)
    return m


def _scale(x, y, z):
    return np.diag([x, y, z, 1.0])


def _xform(m, P):
    return P @ m[:3, :3].T + m[:3, 3]


def killeroo_normals():
    """Per-vertex limit-surface normals of the subdivided killeroo, as the reference's mesh stores them (object space;
    LoopSubdivide hands them to CreateTriangleMesh, shapes/loopsubdiv.cpp:395-398)."""
    z = np.load(os.path.join(_GOLDEN, "killeroo_mesh.npz"))
    return z["N"].astype(np.float64)


def _xform_normals(m, N):
    """Transform::operator()(Normal3f): multiplication by the inverse transpose of the matrix's 3x3 part."""
    return N @ np.linalg.inv(m[:3, :3])


def killeroo_mesh():
    f = os.path.join(_GOLDEN, "killeroo_mesh.npz")
    z = np.load(f)
    return z["P"].astype(np.float64), z["idx"].astype(np.int64)


def _killeroo_camera(xres, yres):
    """LookAt 400 20 30  0 63 -110  0 0 1 ; Rotate -5 0 0 1 ; Camera "perspective" fov 39
    (scenes/killeroo-simple:1-3, scenes/killeroo-feet:1-3).  The CTM at `Camera` is LookAt * Rotate(-5, z)
    = world-to-camera, i.e. the world turns by -5 deg about z before the LookAt: the same view as a
    LookAt whose eye and target are turned by +5 deg about the world z axis (up = z is invariant)."""
    r = _rot_z(5.0)
    eye, look = _xform(r, np.array([[400.0, 20, 30]]))[0], _xform(r, np.array([[0.0, 63, -110]]))[0]
    return look_at_camera(eye, look, (0, 0, 1), 39, xres, yres)


def killeroo_feet(xres=5000, yres=5000):
    """scenes/killeroo-feet (BASELINE config 4, scripts/run_rbsps.sh): one killeroo translated by
    (-200, 100, -100) (33 264 triangles), distant light along the view direction, 5000x5000 film with
    crop window [.3 .5] x [.4 .8] -> pixels x in [1500, 2500), y in [2000, 4000)."""
    P, idx = killeroo_mesh()
    tv = _xform(_translate(-200, 100, -100), P)[idx].reshape(-1, 9).astype(np.float32) + np.float32(0)
    cam = _killeroo_camera(xres, yres)
    f32 = np.float32   # Film::croppedPixelBounds (film.cpp:56-60): ceil(int resolution * float window), in float
    cropwindow = (.3, .5, .4, .8)
    crop = (int(np.ceil(f32(xres) * f32(.3))), int(np.ceil(f32(xres) * f32(.5))),
            int(np.ceil(f32(yres) * f32(.4))), int(np.ceil(f32(yres) * f32(.8))))
    # LightSource "distant" "point from" [400 20 30] "point to" [0 63 -110] (L = 1); Material "plastic" Kd (.4 .5 .4)
    # Ks .3 roughness .15 on every triangle (scenes/killeroo-feet:17-25)
    return dict(name="killeroo-feet", tri_verts=tv, camera=cam, crop=crop, cropwindow=cropwindow,
                light_dir=np.array([400 - 0, 20 - 63, 30 + 110], np.float32),
                light_from=np.array([400, 20, 30], np.float32), light_to=np.array([0, 63, -110], np.float32),
                materials=np.array([[1, .4, .5, .4, .3, .3, .3, .15]], np.float32))


def killeroo_simple(xres=700, yres=700):
    """scenes/killeroo-simple: 2 killeroos (2 x 33 264 tris) + 2 ground quads (4 tris) = 66 532
    triangles; camera LookAt 400 20 30 / 0 63 -110 / 0 0 1, fov 39, 700x700."""
    P, idx = killeroo_mesh()
    tris = []
    # AttributeBegin Translate 0 0 -140 : two quads (scenes/killeroo-simple:27-35)
    q1 = np.array([[-1000, -1000, 0], [1000, -1000, 0], [1000, 1000, 0], [-1000, 1000, 0]], float)
    q2 = np.array([[-400, -1000, -1000], [-400, 1000, -1000], [-400, 1000, 1000], [-400, -1000, 1000]], float)
    for q in (q1, q2):
        qw = _xform(_translate(0, 0, -140), q)
        tris.append(qw[[0, 1, 2, 2, 3, 0]].reshape(-1, 9))
    # Scale .5 .5 .5 ; Rotate -60 0 0 1 ; Translate 100 200 -140 ; killeroo ; Translate -200 0 0 ; killeroo (:38-49)
    ctm = _scale(.5, .5, .5) @ _rot_z(-60) @ _translate(100, 200, -140)
    tris.append(_xform(ctm, P)[idx].reshape(-1, 9))
    # world-space vertex normals (TriangleMesh::n = ObjectToWorld(N)); the ground quads have none (flag bit 1 clear)
    N = killeroo_normals()
    normals = [np.zeros((4, 9)), _xform_normals(ctm, N)[idx].reshape(-1, 9)]
    ctm = ctm @ _translate(-200, 0, 0)
    tris.append(_xform(ctm, P)[idx].reshape(-1, 9))
    normals.append(_xform_normals(ctm, N)[idx].reshape(-1, 9))
    tn = np.concatenate(normals).astype(np.float32)
    tv = np.concatenate(tris).astype(np.float32) + np.float32(0)
    cam = _killeroo_camera(xres, yres)
    nk = len(idx)
    # Material per primitive (scenes/killeroo-simple:27,40,44): ground matte Kd (.5 .5 .8); the two killeroos are
    # "plastic" Kd (.4 .2 .2) / (.4 .5 .4) — their diffuse part; index 3 = the emitter's matte Kd 0 (:19)
    prim_material = np.concatenate([np.zeros(4, np.uint8), np.full(nk, 1, np.uint8), np.full(nk, 2, np.uint8)])
    kds = np.array([[.5, .5, .8], [.4, .2, .2], [.4, .5, .4], [0, 0, 0]], np.float32)
    # the scene's full materials (kind 0 matte / 1 plastic, Kd, Ks, roughness), scenes/killeroo-simple:19,27,40-41,44-45
    materials = np.array([[0, .5, .5, .8, 0, 0, 0, 0], [1, .4, .2, .2, .5, .5, .5, .025], [1, .4, .5, .4, .3, .3, .3, .15],
                          [0, 0, 0, 0, 0, 0, 0, 0]], np.float32)
    return dict(name="killeroo-simple", tri_verts=tv, camera=cam, light=np.array([150, 120, 20], np.float32),
                sphere_light=dict(center=np.array([150, 120, 20], np.float32), radius=3.0, L=np.array([2000.0] * 3, np.float32)),
                prim_material=prim_material, material_kd=kds, materials=materials,
                tri_normals=tn, prim_flags=np.concatenate([np.zeros(4, np.uint8), np.full(2 * nk, 2, np.uint8)]),
                note="reference geometry (fixture); light = the scene's sphere area light (scenes/killeroo-simple:18-24) "
                     "or, for the point-light integrator, a point light at its centre")


def specular_materials(sc):
    """scenes/killeroo-simple with its two killeroos made perfectly specular — the first `Material "glass"` (Kr 1, Kt 1,
    index 1.5: materials/glass.cpp's defaults), the second `Material "mirror" "color Kr" [.9 .9 .9]` (mirror.cpp's
    default) — over the scene's own matte ground and sphere area light: the test scene of the specular path
    (specularBounce / etaScale in path.cpp, FresnelSpecular and SpecularReflection in reflection.cpp)."""
    m = sc["materials"].copy()
    m[1] = [3, 1, 1, 1, 1, 1, 1, 1.5]
    m[2] = [2, .9, .9, .9, 0, 0, 0, 0]
    return m


def metal_materials(sc, oren_nayar=False):
    """scenes/killeroo-simple with the microfacet materials scenes/living-room uses (scene.pbrt:14-21): ground =
    `Material "substrate"` Kd (.5 .5 .8) Ks .04 roughness .1 (remapped); first killeroo = `Material "metal"` with the
    file's BrushedStainlessSteel eta / k, roughness .1, "remaproughness" false (kind + 32); second = the file's
    "Paneling" substrate (Kd .8, Ks .04, roughness .2, "remaproughness" false)."""
    m = sc["materials"].copy()
    m[0] = [5, .5, .5, .8, .04, .04, .04, .1]
    if oren_nayar:   # ground = `Material "matte" "float sigma" [40]`: the Oren-Nayar lobe (living-room's PlantPot has sigma 1)
        m[0] = [0, .5, .5, .8, 40, 0, 0, 0]
    m[1] = [4 + 32, 1.657460, 0.880369, 0.521229, 9.223869, 6.269523, 4.837001, .1]
    m[2] = [5 + 32, .8, .8, .8, .04, .04, .04, .2]
    return m


def with_sphere_light(sc, plastic=False):
    """Aggregate of `sc` with its sphere area light appended as the LAST primitive, as pbrtShape adds it to the
    primitive list (scenes/killeroo-simple:18-24): returns (bounds_verts, prim_kind, sphere_rows, prim_material,
    materials).  bounds_verts gives the sphere primitive a stand-in triangle with the sphere's world bound, which is
    all the BVH / kd builders read; the device gets prim_kind/sphere_rows for the real test."""
    c, r = sc["sphere_light"]["center"].astype(np.float32), np.float32(sc["sphere_light"]["radius"])
    lo, hi = c - r, c + r
    fake = np.concatenate([lo, hi, lo]).astype(np.float32)[None]
    verts = np.concatenate([sc["tri_verts"], fake])
    n = len(verts)
    prim_kind = np.zeros(n, np.uint8)
    prim_kind[-1] = 1
    rows = np.zeros((1, 16), np.float32)   # WorldToObject = Translate(-c), radius
    rows[0, 0] = rows[0, 5] = rows[0, 10] = 1
    rows[0, 3], rows[0, 7], rows[0, 11], rows[0, 12] = -c[0], -c[1], -c[2], r
    pm = np.concatenate([sc["prim_material"], np.array([len(sc["material_kd"]) - 1], np.uint8)])
    mats = np.zeros((len(sc["material_kd"]), 8), np.float32)
    mats[:, 1:4] = sc["material_kd"]   # the diffuse part only, or (plastic=True) the scene's own materials
    if plastic:
        mats = sc["materials"].copy()
    return verts, prim_kind, rows, pm, mats


def matrix_camera(world_to_camera, fov_deg, xres, yres):
    """Camera of a scene file that sets the CTM with `Transform [...]` before `Camera "perspective"` (e.g.
    scenes/living-room/scene.pbrt:2,6): camera-to-world = inverse CTM, whose columns are the camera's x (raster right),
    y (up), z (view direction) axes and position (cameras/perspective.cpp:61-66).  Works for CTMs that swap
    handedness (living-room's has determinant -1)."""
    c2w = np.linalg.inv(np.asarray(world_to_camera, np.float64))
    right, up, d, eye = c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3]
    right, up, d = right / np.linalg.norm(right), up / np.linalg.norm(up), d / np.linalg.norm(d)
    aspect = xres / yres
    t = np.tan(np.radians(fov_deg) / 2)
    sx, sy = (aspect, 1.0) if aspect > 1 else (1.0, 1 / aspect)
    return dict(eye=eye.astype(np.float32), dir=d.astype(np.float32), right=(right * t * sx).astype(np.float32),
                up=(up * t * sy).astype(np.float32), xres=int(xres), yres=int(yres), world_to_camera=np.asarray(world_to_camera, np.float64),
                w2c=np.asarray(world_to_camera, np.float32).reshape(16), fov=float(fov_deg))


def living_room(xres=1280, yres=720, textures=True):
    """scenes/living-room/scene.pbrt (BASELINE config 5): the reference's OWN geometry — 65 `plymesh` shapes,
    143 163 triangles in file order, world space — from the committed fixture tests/golden/living_room_mesh.npz
    (made from the reference's scene by tests/golden/make_living_room.py); camera = the file's `Transform` + fov
    58.715508; film 1280x720 in the file (BASELINE quotes 1920x1080 @ 1024 spp).

    Light: `LightSource "infinite" "string mapname" "textures/Sky 19.pfm"` under the file's rotation; the map is NOT in
    the reference tree, so the reference itself falls back to a constant radiance L = 1 (infinite.cpp:52-60) — that is
    what `infinite_light` describes.
    Materials: the file uses matte (88 114 triangles), substrate, glass, metal, uber and mirror, some with image textures.
    `materials` (20-float records) follows the file's MakeNamedMaterial table (scene.pbrt:13-31) with the shade stage's own
    kinds: matte keeps its Kd (PlantPot: sigma 1, Oren-Nayar); substrate = "substrate" (its Kd, Ks .04, u = v roughness,
    "remaproughness" false); "metal" with the file's eta / k and roughness .02 unremapped; "glass" index 1.5 (Kr = Kt = 1,
    glass.cpp's defaults); "mirror" Kr .9 (mirror.cpp's default); "uber" (Leaves: Ks 0, uber.cpp's defaults for the rest).
    Image textures (scene.pbrt:8-12, all "bool trilinear" true): textures/wood.tga and wood5.tga are ABSENT from the reference tree,
    so the reference itself replaces them by one grey texel 0.5 (imagemap.cpp:67-72) which ImageTexture's convertIn gamma-decodes:
    Floor's and TableWood's Kd is the constant 0x3e5b2d9a = 0.21404114 here too.  The two files that are there — picture8.tga on
    the Painting's 2 triangles, leaf.tga as Kd and opacity of the Leaves' 3 264 — are `textures` (tests/golden/
    living_room_textures.npz holds their pixels), named by slots 18 / 19 of the material records; `textures=False` substitutes the
    constants .5 (Kd) and 1 (opacity) for them (the round-2 per-pixel goldens before ABI v9).
    `tri_normals` / `prim_flags`: the PLY files' vertex normals in world space (TriangleMesh applies ObjectToWorld to them,
    triangle.cpp:61-64; 140 155 triangles have them, 3 008 do not), interpolated shading normals as triangle.cpp:348-400.
    `legacy_materials` keeps the round-1 matte + plastic table (the numbers of profiles/r01_living_room_*.json).  The TRAVERSAL
    workload (tree, rays, hit parity) is the reference scene's as shipped."""
    z = np.load(os.path.join(_GOLDEN, "living_room_mesh.npz"))
    P, UV, idx = z["P"], z["UV"], z["idx"]
    tv = P[idx].reshape(-1, 9).astype(np.float32) + np.float32(0)
    uv = UV[idx].reshape(-1, 6).astype(np.float32)
    kinds, kd = [str(k) for k in z["material_kinds"]], z["material_kd"]
    rough = {"Floor": .15, "TableWood": .05, "Paneling": .2}   # scene.pbrt:13,14,18 ("remaproughness" false in the file)
    tex_files = ["leaf", "picture8"]   # textures/leaf.tga (Texture04, Texture05), textures/picture8.tga (Texture03)
    tex_of = {"Texture03": "picture8", "Texture04": "leaf", "Texture05": "leaf"}
    absent = {"Texture01", "Texture02"}   # textures/wood.tga, wood5.tga: not in the reference tree -> the constant 1
    mats = np.zeros((len(kinds), 20), np.float32)
    legacy = np.zeros((len(kinds), 8), np.float32)
    for i, (k, name) in enumerate(zip(kinds, [str(n) for n in z["material_names"]])):
        tex = str(z["material_tex"][i])
        grey = np.array([0x3e5b2d9a] * 3, np.uint32).view(np.float32)    # InverseGammaCorrect(0.5f): the reference's stand-in for a missing file
        c = kd[i] if kd[i][0] >= 0 else (grey if tex in absent else np.array([.5, .5, .5], np.float32))
        cl = kd[i] if kd[i][0] >= 0 else np.array([.5, .5, .5], np.float32)
        legacy[i] = [0, .5, .5, .5, 0, 0, 0, 0]
        if k == "substrate":
            mats[i, :8] = [5 + 32, c[0], c[1], c[2], .04, .04, .04, rough[name]]
            legacy[i] = [1, cl[0], cl[1], cl[2], .04, .04, .04, rough[name]]
        elif k == "matte":
            mats[i, :8] = [0, c[0], c[1], c[2], 0, 0, 0, 0]
            if textures and tex in tex_of:   # Painting: "texture Kd" Texture03
                mats[i, 18] = 1 + tex_files.index(tex_of[tex])
            legacy[i] = [0, cl[0], cl[1], cl[2], 0, 0, 0, 0]
            if name == "PlantPot":   # scene.pbrt:26 "float sigma" [1]: Oren-Nayar
                mats[i, :8] = [0, c[0], c[1], c[2], 1, 0, 0, 0]
        elif k == "metal":     # scene.pbrt:21 BrushedStainlessSteel
            mats[i, :8] = [4 + 32, 1.657460, 0.880369, 0.521229, 9.223869, 6.269523, 4.837001, .02]
        elif k == "glass":     # scene.pbrt:25
            mats[i, :8] = [3, 1, 1, 1, 1, 1, 1, 1.5]
        elif k == "mirror":    # scene.pbrt:20,27
            mats[i, :8] = [2, .9, .9, .9, 0, 0, 0, 0]
        elif k == "uber":      # scene.pbrt:29 Leaves: Ks 0; uber.cpp:102-128 defaults Kr 0, Kt 0, roughness .1, eta 1.5, opacity 1
            mats[i] = [6, c[0], c[1], c[2], 0, 0, 0, .1, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1.5, 0, 0]
            if textures:       # "texture Kd" Texture04, "texture opacity" Texture05: both textures/leaf.tga
                mats[i, 18] = mats[i, 19] = 1 + tex_files.index("leaf")
        else:
            raise ValueError("scenes/living-room material kind %r" % k)
    # TransformBegin Transform [-1 0 8.74228e-008 0  -8.74228e-008 0 -1 0  0 1 0 0  0 0 0 1] LightSource "infinite"
    # (scene.pbrt:163-166; column-major in the file) -> row-major LightToWorld rotation
    l2w = np.array([-1, 0, 8.74228e-008, 0, -8.74228e-008, 0, -1, 0, 0, 1, 0, 0, 0, 0, 0, 1], np.float64).reshape(4, 4).T[:3, :3]
    tn = z["N"][idx].reshape(-1, 9).astype(np.float32)
    tex_px = None
    if textures:
        zt = np.load(os.path.join(_GOLDEN, "living_room_textures.npz"))
        tex_px = [zt[n] for n in tex_files]      # 8-bit RGB, row 0 = top (what ReadImageTGA returns before / 255.f)
    return dict(name="living-room", texture_names=tex_files, texture_pixels=tex_px, tri_verts=tv, tri_uvs=uv, tri_normals=tn, prim_flags=z["tri_flags"].astype(np.uint8),
                camera=matrix_camera(z["world_to_camera"], 58.715508, xres, yres),
                materials=mats, legacy_materials=legacy, prim_material=z["tri_material"].astype(np.uint8), material_names=[str(n) for n in z["material_names"]],
                material_kinds=kinds, infinite_light=dict(L=np.array([1, 1, 1], np.float32), light_to_world=l2w),
                maxdepth=65, spp=64)


def living_room_textures(sc):
    """The `textures=` argument of WavefrontPathIntegrator for scenes.living_room(): the MIPMap pyramids of the scene's image
    files (None when the scene was made with textures=False)."""
    if not sc.get("texture_pixels"):
        return None
    from . import textures as tx
    return [tx.ImageTexture(rgb8=px) for px in sc["texture_pixels"]]


# ---- procedural stand-ins --------------------------------------------------------------------
def sponza_standin(seed=7, detail=0.885, xres=700, yres=700):
    """Procedural atrium standing in for Crytek Sponza (geometry absent from the reference tree).
    Two-storey colonnaded nave, arches, drapes, vases, displaced floor/walls, roof beams.
    the default detail gives ~262 k triangles.  Camera and point light from scenes/sponza:1-2,16."""
    rng = np.random.default_rng(seed)
    k = float(np.sqrt(detail))
    n = lambda x: max(2, int(round(x * k)))
    parts = []
    X0, X1, Z0, Z1, H = -1.0, 1.0, -0.45, 0.45, 0.8
    zc = 0.22  # column rows at z = +-zc
    # floor with bumpy tiles
    def floor(u, v):
        x, z = X0 + u * (X1 - X0), Z0 + v * (Z1 - Z0)
        y = 0.004 * np.sin(40 * x) * np.sin(40 * z)
        return np.stack([x, y, z], 1)
    parts.append(_surface(floor, n(220), n(100)))
    # outer walls with brick relief, two long + two short
    def wall_long(zpos, sgn):
        def f(u, v):
            x, y = X0 + u * (X1 - X0), v * H
            z = zpos + sgn * 0.003 * np.sign(np.sin(60 * x + 3 * np.floor(y * 25))) * (np.sin(25 * np.pi * y) > 0)
            return np.stack([x, y, z], 1)
        return f
    parts.append(_surface(wall_long(Z0, 1), n(180), n(70)))
    parts.append(_surface(wall_long(Z1, -1), n(180), n(70)))
    def wall_short(xpos):
        def f(u, v):
            z, y = Z0 + u * (Z1 - Z0), v * H
            return np.stack([np.full_like(z, xpos), y, z], 1)
        return f
    parts.append(_surface(wall_short(X0), n(60), n(50)))
    parts.append(_surface(wall_short(X1), n(60), n(50)))
    # ceiling + gallery floors
    parts.append(_box((X0, H, Z0), (X1, H + 0.02, Z1), n(24)))
    for s in (-1, 1):
        z0, z1 = (Z0, -zc) if s < 0 else (zc, Z1)
        parts.append(_box((X0, 0.38, z0), (X1, 0.40, z1), n(30)))
    # columns (two storeys, both sides) with entasis + fluting
    ncol = 10
    xs = np.linspace(X0 + 0.12, X1 - 0.12, ncol)
    def col_profile(h, r0):
        def p(v):
            r = r0 * (1 - 0.15 * v) * (1 + 0.25 * np.exp(-((v - 0.02) / 0.03) ** 2) + 0.3 * np.exp(-((v - 0.98) / 0.03) ** 2))
            return r, v * h
        return p
    for y0, h, r0 in ((0.0, 0.38, 0.028), (0.40, 0.36, 0.022)):
        for x in xs:
            for z in (-zc, zc):
                parts.append(_lathe(col_profile(h, r0), x, z, y0, n(28), n(36)))
    # arches between columns (half tori), both storeys
    def arch(xa, xb, z, ytop, rt):
        cx, R = 0.5 * (xa + xb), 0.5 * (xb - xa)
        def f(u, v):
            a, b = np.pi * u, 2 * np.pi * v
            rr = R + rt * np.cos(b)
            return np.stack([cx + rr * np.cos(a), ytop - R + rr * np.sin(a) * 0.6, z + rt * np.sin(b)], 1)
        return _surface(f, n(20), n(8))
    for ytop in (0.38, 0.76):
        for i in range(ncol - 1):
            for z in (-zc, zc):
                parts.append(arch(xs[i], xs[i + 1], z, ytop + 0.04, 0.012))
    # drapes hanging between upper columns
    for i in range(0, ncol - 1, 2):
        ph = rng.random() * 6
        z = zc if i % 4 == 0 else -zc
        def drape(u, v, i=i, ph=ph, z=z):
            x = xs[i] + u * (xs[i + 1] - xs[i])
            y = 0.70 - 0.28 * v - 0.03 * np.sin(np.pi * u)
            zz = z + 0.02 * np.sin(18 * u + ph) * (0.3 + v)
            return np.stack([x, y, zz], 1)
        parts.append(_surface(drape, n(64), n(64)))
    # vases along the nave
    def vase_profile(v):
        r = 0.02 + 0.025 * np.sin(np.pi * v) ** 2 + 0.01 * (v > 0.9)
        return r, 0.10 * v
    for x in np.linspace(X0 + 0.3, X1 - 0.3, 6):
        for z in (-0.09, 0.09):
            parts.append(_lathe(vase_profile, x, z, 0.0, n(32), n(24)))
    # roof beams
    for x in np.linspace(X0 + 0.1, X1 - 0.1, 12):
        parts.append(_box((x - 0.012, H - 0.05, Z0), (x + 0.012, H, Z1), n(4)))
    # lion-head-like blobs on the end wall
    def blob(cx, cy, cz, r):
        def f(u, v):
            th, ph = np.pi * v, 2 * np.pi * u
            rr = r * (1 + 0.15 * np.sin(5 * ph) * np.sin(4 * th))
            return np.stack([cx + rr * np.sin(th) * np.cos(ph), cy + rr * np.cos(th), cz + rr * np.sin(th) * np.sin(ph)], 1)
        return _surface(f, n(40), n(30))
    for zz in (-0.2, 0.0, 0.2):
        parts.append(blob(X1 - 0.03, 0.3, zz, 0.04))
    tv = np.concatenate(parts).astype(np.float32) + np.float32(0)
    cam = look_at_camera((-0.7, 0.45, 0.05), (0, 0.25, 0), (0, 1, 0), 42, xres, yres)
    return dict(name="sponza-standin", tri_verts=tv, camera=cam, light=np.array([0, .5, 0], np.float32),
                note="PROCEDURAL STAND-IN for scenes/sponza (geometry absent from the reference tree), seed=%d detail=%g, %d triangles"
                     % (seed, detail, len(tv)))


def museum_standin(seed=11, detail=1.45, xres=700, yres=700):
    """Procedural exhibition hall standing in for the thesis's museum scene (scenes/museum + its `_k` kd-tree variant
    scenes/test_museum_local_k: the geometry/museum/*.pbrt parts — bone, brushedmetals, floor, glass, painted, stone,
    wood... — are absent from the reference tree).  A hall with a relief floor, panelled walls, roof beams, pedestals
    carrying lathe-turned and blobby exhibits, display cases, paintings, benches and a rib-cage skeleton down the
    middle; ~250 k triangles at the default detail.  Camera, light and integrator settings are the template's
    (scenes/museum:1-3,12,15): `LookAt -0.4 0.06 0.05  0.2 0.1 -0.05  0 1 0` followed by `Rotate 180 0 1 0` (so the
    world-space eye is (0.4, 0.06, -0.05) looking at (-0.2, 0.1, 0.05)), fov 60, point light at (0, .5, 0),
    path maxdepth 1, 64 spp, 700x700."""
    rng = np.random.default_rng(seed)
    k = float(np.sqrt(detail))
    n = lambda x: max(2, int(round(x * k)))
    parts = []
    X0, X1, Z0, Z1, H = -0.7, 0.7, -0.4, 0.4, 0.62
    def floor(u, v):
        x, z = X0 + u * (X1 - X0), Z0 + v * (Z1 - Z0)
        y = 0.0015 * (np.abs(np.sin(24 * np.pi * x)) < 0.06) + 0.0015 * (np.abs(np.sin(24 * np.pi * z)) < 0.06)
        return np.stack([x, y, z], 1)
    parts.append(_surface(floor, n(200), n(120)))
    def wall_x(zpos, sgn):   # long walls with vertical panels
        def f(u, v):
            x, y = X0 + u * (X1 - X0), v * H
            z = zpos + sgn * 0.004 * (np.sin(30 * np.pi * u) > 0.8) + sgn * 0.006 * (v < 0.12)
            return np.stack([x, y, z], 1)
        return f
    parts.append(_surface(wall_x(Z0, 1), n(160), n(60)))
    parts.append(_surface(wall_x(Z1, -1), n(160), n(60)))
    for xpos in (X0, X1):
        def f(u, v, xpos=xpos):
            z, y = Z0 + u * (Z1 - Z0), v * H
            return np.stack([np.full_like(z, xpos), y, z], 1)
        parts.append(_surface(f, n(70), n(50)))
    parts.append(_box((X0, H, Z0), (X1, H + 0.02, Z1), n(20)))
    for x in np.linspace(X0 + 0.08, X1 - 0.08, 10):   # roof beams
        parts.append(_box((x - 0.01, H - 0.04, Z0), (x + 0.01, H, Z1), n(4)))
    # pedestals with exhibits along both sides
    def vase(v):
        return 0.012 + 0.022 * np.sin(np.pi * v) ** 1.5 + 0.006 * (v > 0.92), 0.09 * v
    def urn(v):
        return 0.02 * (1 + 0.6 * np.sin(2.5 * np.pi * v) * (1 - v)), 0.07 * v
    def blob(cx, cy, cz, r, a, b):
        def f(u, v):
            th, ph = np.pi * v, 2 * np.pi * u
            rr = r * (1 + 0.2 * np.sin(a * ph) * np.sin(b * th))
            return np.stack([cx + rr * np.sin(th) * np.cos(ph), cy + rr * np.cos(th), cz + rr * np.sin(th) * np.sin(ph)], 1)
        return _surface(f, n(44), n(32))
    px = np.linspace(X0 + 0.15, X1 - 0.15, 7)
    for i, x in enumerate(px):
        for z in (-0.27, 0.27):
            parts.append(_box((x - 0.035, 0, z - 0.035), (x + 0.035, 0.11, z + 0.035), n(6)))
            kind = int(rng.integers(0, 3))
            if kind == 0:
                parts.append(_lathe(vase, x, z, 0.11, n(36), n(28)))
            elif kind == 1:
                parts.append(_lathe(urn, x, z, 0.11, n(36), n(28)))
            else:
                parts.append(blob(x, 0.15, z, 0.035, int(rng.integers(3, 7)), int(rng.integers(2, 6))))
            if i % 2 == 0:   # display case around it (thin posts + top)
                for dx in (-0.045, 0.045):
                    for dz in (-0.045, 0.045):
                        parts.append(_box((x + dx - 0.003, 0.11, z + dz - 0.003), (x + dx + 0.003, 0.24, z + dz + 0.003), n(2)))
                parts.append(_box((x - 0.048, 0.24, z - 0.048), (x + 0.048, 0.245, z + 0.048), n(6)))
    # paintings on the long walls
    for x in np.linspace(X0 + 0.2, X1 - 0.2, 6):
        for z, sgn in ((Z0, 1), (Z1, -1)):
            w, h = 0.05 + 0.03 * rng.random(), 0.04 + 0.03 * rng.random()
            z0, z1 = sorted((z + sgn * 0.006, z + sgn * 0.014))
            parts.append(_box((x - w, 0.30 - h, z0), (x + w, 0.30 + h, z1), n(8)))
    # benches down the middle
    for x in (-0.35, 0.35):
        parts.append(_box((x - 0.08, 0.035, -0.03), (x + 0.08, 0.045, 0.03), n(8)))
        for dx in (-0.07, 0.07):
            parts.append(_box((x + dx - 0.005, 0, -0.025), (x + dx + 0.005, 0.035, 0.025), n(2)))
    # the skeleton: a spine with rib hoops, skull blob and tail
    def spine(u, v):
        t = -0.22 + 0.44 * u
        y = 0.17 + 0.03 * np.sin(np.pi * u)
        a = 2 * np.pi * v
        return np.stack([t, y + 0.006 * np.sin(a), 0.006 * np.cos(a)], 1)
    parts.append(_surface(spine, n(120), n(10)))
    for t in np.linspace(-0.12, 0.1, 14):
        R = 0.045 * (1 - 2.5 * (t + 0.01) ** 2 / 0.05)
        yc = 0.17 + 0.03 * np.sin(np.pi * (t + 0.22) / 0.44) - R * 0.85
        def rib(u, v, t=t, R=R, yc=yc):
            a, b = (0.15 + 1.7 * u) * np.pi + np.pi / 2 + 0.15 * np.pi, 2 * np.pi * v
            rr = R + 0.003 * np.cos(b)
            return np.stack([t + 0.003 * np.sin(b), yc + rr * np.sin(a) * 1.1, rr * np.cos(a)], 1)
        parts.append(_surface(rib, n(40), n(8)))
    parts.append(blob(-0.26, 0.2, 0.0, 0.03, 2, 3))
    for lx in (-0.1, 0.08):   # leg posts carrying it
        for lz in (-0.03, 0.03):
            parts.append(_lathe(lambda v: (0.005 + 0.002 * np.sin(6 * np.pi * v), 0.15 * v), lx, lz, 0.0, n(12), n(24)))
    tv = np.concatenate(parts).astype(np.float32) + np.float32(0)
    cam = look_at_camera((0.4, 0.06, -0.05), (-0.2, 0.1, 0.05), (0, 1, 0), 60, xres, yres)
    return dict(name="museum-standin", tri_verts=tv, camera=cam, light=np.array([0, .5, 0], np.float32), maxdepth=1, spp=64,
                note="PROCEDURAL STAND-IN for scenes/museum (geometry absent from the reference tree), seed=%d detail=%g, %d triangles"
                     % (seed, detail, len(tv)))


def conference_standin(seed=5, detail=1.65, xres=700, yres=700):
    """Procedural meeting room standing in for the thesis's conference scene (BASELINE config 3 names it next to the
    museum; `scenes/conference` and `scenes/test_conference_local(_k)` are listed in .MISSING_LARGE_BLOBS, so the reference
    tree holds neither the geometry nor the template).  A room with a coffered ceiling and panelled walls, a long oval
    table with a raised rim, two rows of chairs (seat, back of vertical slats, four turned legs, arm rests), ceiling lamps
    and a lectern; ~280 k triangles at the default detail (the classic conference model has 282 k).  No template exists, so
    camera and light follow the museum template's conventions: an eye at head height in a corner looking along the table,
    fov 60, a point light under the ceiling, path maxdepth 1, 64 spp, 700x700."""
    rng = np.random.default_rng(seed)
    k = float(np.sqrt(detail))
    n = lambda x: max(2, int(round(x * k)))
    parts = []
    X0, X1, Z0, Z1, H = -0.8, 0.8, -0.45, 0.45, 0.55
    def floor(u, v):
        x, z = X0 + u * (X1 - X0), Z0 + v * (Z1 - Z0)
        return np.stack([x, 0.001 * np.sin(40 * np.pi * u) * np.sin(40 * np.pi * v), z], 1)
    parts.append(_surface(floor, n(150), n(90)))
    def ceiling(u, v):   # coffers
        x, z = X0 + u * (X1 - X0), Z0 + v * (Z1 - Z0)
        y = H - 0.02 * ((np.abs(np.sin(8 * np.pi * u)) < 0.15) | (np.abs(np.sin(5 * np.pi * v)) < 0.15))
        return np.stack([x, y, z], 1)
    parts.append(_surface(ceiling, n(170), n(100)))
    for zpos, sgn in ((Z0, 1), (Z1, -1)):
        def wall(u, v, zpos=zpos, sgn=sgn):
            x, y = X0 + u * (X1 - X0), v * H
            return np.stack([x, y, zpos + sgn * 0.005 * (np.sin(22 * np.pi * u) > 0.85) + sgn * 0.008 * (v < 0.1)], 1)
        parts.append(_surface(wall, n(140), n(50)))
    for xpos in (X0, X1):
        def wall(u, v, xpos=xpos):
            z, y = Z0 + u * (Z1 - Z0), v * H
            return np.stack([np.full_like(z, xpos), y, z], 1)
        parts.append(_surface(wall, n(60), n(40)))
    # the table: an oval top with a raised rim on a central plinth
    def top(u, v):
        ang, r = 2 * np.pi * u, v
        y = 0.2 + 0.004 * (v > 0.93)
        return np.stack([0.5 * r * np.cos(ang), y, 0.16 * r * np.sin(ang)], 1)
    parts.append(_surface(top, n(220), n(40)))
    def rim(u, v):
        ang = 2 * np.pi * u
        return np.stack([0.5 * np.cos(ang), 0.185 + 0.02 * v, 0.16 * np.sin(ang)], 1)
    parts.append(_surface(rim, n(220), n(4)))
    parts.append(_box((-0.35, 0.0, -0.05), (0.35, 0.185, 0.05), n(10)))
    # chairs along both long sides
    for side in (-1, 1):
        for cx in np.linspace(-0.42, 0.42, 9):
            cz = side * 0.25
            parts.append(_box((cx - 0.035, 0.11, cz - 0.035), (cx + 0.035, 0.12, cz + 0.035), n(6)))      # seat
            for sx in np.linspace(-0.03, 0.03, 5):                                                       # back slats
                parts.append(_box((cx + sx - 0.004, 0.12, cz + side * 0.032), (cx + sx + 0.004, 0.24, cz + side * 0.037), n(3)))
            parts.append(_box((cx - 0.035, 0.235, cz + side * 0.03), (cx + 0.035, 0.25, cz + side * 0.04), n(4)))
            for ax in (-0.033, 0.033):                                                                     # arm rests
                parts.append(_box((cx + ax - 0.004, 0.16, cz - 0.03), (cx + ax + 0.004, 0.165, cz + 0.03), n(3)))
            for lx in (-0.03, 0.03):
                for lz in (-0.03, 0.03):
                    parts.append(_lathe(lambda v: (0.004 + 0.0015 * np.sin(5 * np.pi * v), 0.11 * v), cx + lx, cz + lz, 0.0, n(10), n(14)))
    # ceiling lamps and a lectern
    for lx in np.linspace(-0.5, 0.5, 5):
        parts.append(_lathe(lambda v: (0.03 * np.sin(np.pi * v * 0.5) + 0.002, -0.03 * v), lx, 0.0, H - 0.02, n(28), n(16)))
    parts.append(_box((0.62, 0.0, -0.06), (0.7, 0.22, 0.06), n(8)))
    parts.append(_lathe(lambda v: (0.012 + 0.004 * rng.random(), 0.05 * v), 0.66, 0.0, 0.22, n(16), n(8)))
    tv = np.concatenate(parts).astype(np.float32) + np.float32(0)
    cam = look_at_camera((0.7, 0.3, -0.38), (-0.3, 0.12, 0.1), (0, 1, 0), 60, xres, yres)
    return dict(name="conference-standin", tri_verts=tv, camera=cam, light=np.array([0, .5, 0], np.float32), maxdepth=1, spp=64,
                note="PROCEDURAL STAND-IN for scenes/conference (geometry AND template absent from the reference tree), seed=%d "
                     "detail=%g, %d triangles" % (seed, detail, len(tv)))


def soup(n, seed=1, extent=10.0, size=0.5):
    """Random triangle soup (unit tests)."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, (n, 1, 3))
    return (c + rng.normal(0, size, (n, 3, 3))).reshape(-1, 9).astype(np.float32) + np.float32(0)
