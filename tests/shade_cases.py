"""TEST INFRASTRUCTURE ONLY.  Scenes of the matched-sample parity tests, described once for both sides: the reference's
own renderer (oracle/refbind.render: SamplerIntegrator::Render + PathIntegrator::Li over reference Materials / Lights)
and the wavefront driver (b200rt.WavefrontPathIntegrator over the same geometry, light and material records)."""
import numpy as np


def _lookat(sc):
    la = sc["camera"]["lookat"]
    return la[0:3], la[3:6], la[6:9], sc["camera"]["fov"]


def _bumpy_sphere(n=24, seed=3):
    """A lumpy sphere mesh with smooth per-vertex normals and a (u, v) parameterisation, Triangle.Watertight-style
    (src/tests/shapes.cpp:28-60 builds the same kind of mesh)."""
    rng = np.random.default_rng(seed)
    u, v = np.meshgrid(np.linspace(0, 1, n + 1), np.linspace(0.02, 0.98, n + 1), indexing="ij")
    th, ph = v * np.pi, u * 2 * np.pi
    r = 1 + 0.15 * np.sin(3 * ph) * np.sin(4 * th) + 0.02 * rng.standard_normal(th.shape)
    r[-1] = r[0]
    P = np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), r * np.cos(th)], -1).reshape(-1, 3)
    N = np.stack([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)], -1).reshape(-1, 3)
    N += 0.1 * rng.standard_normal(N.shape)
    UV = np.stack([u, v], -1).reshape(-1, 2)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    a = (i * (n + 1) + j).ravel()
    idx = np.concatenate([np.stack([a, a + n + 1, a + n + 2], 1), np.stack([a, a + n + 2, a + 1], 1)])
    return P.astype(np.float32), N.astype(np.float32), UV.astype(np.float32), idx.astype(np.int32)


def build(name, res=48):
    """Returns a dict:
       ref      : kwargs of refbind.RefScene (P, indices, uv, spheres, normals, reverse_orientation)
       render   : kwargs of refbind.render after (scene, accel) — eye .. kd, plus extras
       integ    : kwargs of WavefrontPathIntegrator (without sampler / spp)
       accel    : (verts, kwargs) for B200Accel / OracleAccel (prim_kind, spheres, tri_uvs, tri_normals, prim_flags)
       maxdepth, mis (bool: the light has area)"""
    from b200rt import scenes
    out = dict(mis=False, maxdepth=5)
    if name == "sponza_point":
        sc = scenes.sponza_standin(detail=0.05, xres=res, yres=res)
        tv = sc["tri_verts"]
        eye, look, up, fov = _lookat(sc)
        I, kd = (1.0, .9, .8), (.6, .5, .4)
        out.update(ref=dict(P=tv.reshape(-1, 3), indices=np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3)),
                   render=dict(eye=eye, look=look, up=up, fov=fov, light_pos=sc["light"], light_I=I, kd=kd),
                   integ=dict(camera=sc["camera"], light_pos=sc["light"], light_I=I, kd=kd), accel=(tv, {}))
        return out
    if name == "sponza_two_points":
        # two LightSource "point" (delta lights only): UniformSampleOneLight's choice and the division by its probability
        sc = scenes.sponza_standin(detail=0.05, xres=res, yres=res)
        tv = sc["tri_verts"]
        eye, look, up, fov = _lookat(sc)
        I, kd = (1.0, .9, .8), (.6, .5, .4)
        second = dict(kind="point", pos=(-0.35, 0.3, 0.1), I=(.2, .5, .9))
        out.update(ref=dict(P=tv.reshape(-1, 3), indices=np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3)),
                   render=dict(eye=eye, look=look, up=up, fov=fov, light_pos=sc["light"], light_I=I, kd=kd, extra_lights=[second],
                               lightsamplestrategy="uniform"),
                   integ=dict(camera=sc["camera"], kd=kd, lights=[dict(kind="point", pos=sc["light"], I=I), second]), accel=(tv, {}))
        return out
    if name.startswith("killeroo_"):
        sc = scenes.killeroo_simple(xres=res, yres=res)
        eye, look, up, fov = _lookat(sc)
        verts, prim_kind, rows, pm, mats = scenes.with_sphere_light(sc, plastic=True)
        kind = name[len("killeroo_"):]
        if kind == "glass":
            mats = scenes.specular_materials(sc)
            out["maxdepth"] = 8
        elif kind == "metal":
            mats = scenes.metal_materials(sc)
        elif kind == "oren":
            mats = scenes.metal_materials(sc, oren_nayar=True)
        elif kind == "uber":
            # Material "uber" (materials/uber.cpp:45-100; scenes/living-room uses it): 20-float records.  Ground: every lobe at
            # once (diffuse + glossy + specular reflection + specular transmission), opaque; first killeroo: opacity .6, so the
            # pass-through SpecularTransmission(1 - opacity, 1, 1) lobe comes first and eta becomes 1; second: glossy + diffuse
            # with "remaproughness" false
            m20 = np.zeros((len(mats), 20), np.float32)
            m20[:, :8] = mats
            m20[0] = [6, .5, .5, .8, .25, .25, .25, .1, .1, .1, .1, .05, .05, .05, 1, 1, 1, 1.5, 0, 0]
            m20[1] = [6, .4, .2, .2, .3, .3, .3, .05, 0, 0, 0, .3, .3, .3, .6, .6, .6, 1.3, 0, 0]
            m20[2] = [6 + 32, .25, .25, .25, .25, .25, .25, .1, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1.5, 0, 0]
            mats = m20
            out["maxdepth"] = 8
        tv = sc["tri_verts"]
        sl = sc["sphere_light"]
        prim_mat = np.concatenate([mats[sc["prim_material"]], np.zeros((1, mats.shape[1]), np.float32)])
        spheres = np.array([list(sl["center"]) + [sl["radius"]]], np.float32)
        render = dict(eye=eye, look=look, up=up, fov=fov, light_pos=sl["center"], light_I=sl["L"], kd=(.5, .5, .5), prim_mat=prim_mat)
        integ = dict(camera=sc["camera"], light_pos=sl["center"], light_I=sl["L"], materials=mats, prim_material=pm)
        if kind == "infinite":
            # LightSource "infinite" with a constant map under a rotated LightToWorld, camera through a `Transform [...]` CTM
            # (what scenes/living-room does); the emitter sphere stays in the aggregate as a black matte primitive
            c, s_ = np.cos(0.7), np.sin(0.7)
            l2w = np.array([[c, 0, s_, 0], [0, 1, 0, 0], [-s_, 0, c, 0], [0, 0, 0, 1]], np.float32) @ \
                np.array([[1, 0, 0, 0], [0, 0, -1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], np.float32)
            from b200rt import wavefront as wf
            render.update(infinite=((1.0, 0.9, 0.8), l2w), light_I=(1.0, 0.9, 0.8))
            del integ["light_pos"]
            integ.update(light_I=(1.0, 0.9, 0.8), infinite_light=(l2w[:3, :3], wf.world_radius(verts)))
            out["mis"] = True
            out["infinite_scene"] = True
        elif kind in ("lights", "lights_power", "lights_spatial"):
            # Scene::lights = the file's sphere area light + a point, a distant and a constant infinite light: every light kind
            # in one distribution ("uniform", or "power": weights = the reference's own Light::Power().y(), filled in by
            # hostwf.reference_trace / the image test after the reference built its scene)
            from b200rt import wavefront as wf
            c, s_ = np.cos(0.7), np.sin(0.7)
            l2w = (np.array([[c, 0, s_], [0, 1, 0], [-s_, 0, c]], np.float32) @ np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32))
            wr = wf.world_radius(verts)
            extra = [dict(kind="point", pos=(100.0, -150.0, 200.0), I=(60000.0, 50000.0, 40000.0)),
                     dict(kind="distant", w=(-0.3, 0.4, 1.0), L=(1.5, 1.5, 1.2), world_radius=wr),
                     dict(kind="infinite", L=(0.3, 0.35, 0.5), light_to_world=l2w, world_radius=wr)]
            strategy = {"lights": "uniform", "lights_power": "power", "lights_spatial": "spatial"}[kind]
            render.update(emitter_L=sl["L"], extra_lights=extra, lightsamplestrategy=strategy)
            del integ["light_pos"], integ["light_I"]
            integ.update(lights=[dict(kind="sphere", center=sl["center"], L=sl["L"], radius=sl["radius"], prim=len(verts) - 1)] + extra,
                         lightsamplestrategy=strategy)
            out["mis"] = True
            out["power"] = strategy == "power"
        else:
            render["emitter_L"] = sl["L"]
            integ["sphere_light"] = (sl["radius"], len(verts) - 1)
            out["mis"] = True
        akw = dict(prim_kind=prim_kind, spheres=rows)
        ref = dict(P=tv.reshape(-1, 3), indices=np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3), spheres=spheres)
        if kind == "normals":
            # scenes/killeroo-simple with the vertex normals its loop-subdivided killeroos carry (the scene as written).  The
            # reference harness builds ONE mesh, so every triangle gets normals there: the four ground triangles get their
            # geometric normal at all three vertices (and the flag on this side: both take the mesh-with-normals path for them)
            tn = sc["tri_normals"].copy()
            e1, e2 = tv[:4, 0:3] - tv[:4, 6:9], tv[:4, 3:6] - tv[:4, 6:9]
            tn[:4] = np.tile(np.cross(e1.astype(np.float64), e2.astype(np.float64)), (1, 3)).astype(np.float32)
            ref["normals"] = tn.reshape(-1, 3)
            akw.update(tri_normals=np.concatenate([tn, np.zeros((1, 9), np.float32)]),
                       prim_flags=np.concatenate([np.full(len(tv), 2, np.uint8), np.zeros(1, np.uint8)]))
        out.update(ref=ref, render=render, integ=integ, accel=(verts, akw))
        return out
    if name == "feet_distant":
        sc = scenes.killeroo_feet(xres=res * 4, yres=res * 4)
        eye, look, up, fov = _lookat(sc)
        tv = sc["tri_verts"]
        from b200rt import wavefront as wf
        out.update(ref=dict(P=tv.reshape(-1, 3), indices=np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3)),
                   render=dict(eye=eye, look=look, up=up, fov=fov, light_pos=(0, 0, 0), light_I=(1, 1, 1), kd=(.5, .5, .5),
                               prim_mat=np.repeat(sc["materials"], len(tv), 0), distant=(sc["light_from"], sc["light_to"]),
                               cropwindow=sc["cropwindow"]),
                   integ=dict(camera=sc["camera"], light_I=(1, 1, 1), materials=sc["materials"], prim_material=np.zeros(len(tv), np.uint8),
                              distant_light=(sc["light_dir"], wf.world_radius(tv)), crop=sc["crop"]),
                   accel=(tv, {}), res=res * 4)
        return out
    if name in ("living_room", "living_room_tex", "living_room_tex_ewa"):
        # BASELINE config 5, scenes/living-room as the reference resolves it (b200rt.scenes.living_room): the file's camera CTM,
        # its infinite light (constant: the sky map is absent), maxdepth 65, the file's matte / Oren-Nayar / substrate / metal /
        # glass / mirror / uber materials and the PLY vertex normals.  The reference harness builds ONE mesh, so the 3 008
        # triangles whose PLY has no normals carry zero normals on both sides (ns falls back to n, triangle.cpp:352-355).
        import os
        from b200rt import wavefront as wf
        # "living_room_tex": with the image textures of the files that are in the reference tree (leaf.tga: Kd and opacity of the
        # Leaves; picture8.tga: Kd of the Painting) — the scene as the reference renders it, box filter aside
        # "living_room_tex_ewa": the same without `"bool trilinear" "true"`, i.e. under the ImageTexture default, the EWA filter
        sc = scenes.living_room(res, res, textures=name != "living_room")
        z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "living_room_mesh.npz"))
        tv = sc["tri_verts"]
        il = sc["infinite_light"]
        l2w = np.eye(4, dtype=np.float32)
        l2w[:3, :3] = il["light_to_world"]
        L = tuple(float(v) for v in il["L"])
        out.update(ref=dict(P=z["P"], indices=z["idx"], uv=z["UV"], normals=z["N"]),
                   render=dict(eye=(0, 0, 0), look=(0, 0, 1), up=(0, 1, 0), fov=sc["camera"]["fov"], light_pos=(0, 0, 0), light_I=L,
                               kd=(.5, .5, .5), prim_mat=sc["materials"][sc["prim_material"]], infinite=(L, l2w),
                               camera_matrix=sc["camera"]["world_to_camera"]),
                   integ=dict(camera=sc["camera"], light_I=L, infinite_light=(il["light_to_world"], wf.world_radius(tv)),
                              materials=sc["materials"], prim_material=sc["prim_material"]),
                   accel=(tv, dict(tri_uvs=sc["tri_uvs"], tri_normals=sc["tri_normals"], prim_flags=np.full(len(tv), 2, np.uint8))),
                   mis=True, infinite_scene=True, maxdepth=sc["maxdepth"])
        if name != "living_room":
            from b200rt import textures as tx
            ewa = name == "living_room_tex_ewa"
            out["render"]["textures"] = sc["texture_pixels"]
            if ewa:
                out["render"]["texture_filter"] = ("ewa", 8.0)
            out["integ"]["textures"] = [tx.ImageTexture(rgb8=px, trilinear=not ewa) for px in sc["texture_pixels"]]
        return out
    if name in ("normals", "normals_flipped"):
        # vertex normals (interpolated shading normal, triangle.cpp:348-400), a real (u, v) parameterisation (dpdu frame) and,
        # for normals_flipped, `ReverseOrientation`; two-lobe plastic so that the frame matters for both lobes
        P, N, UV, idx = _bumpy_sphere()
        flip = name == "normals_flipped"
        tv = P[idx].reshape(-1, 9)
        tn = N[idx].reshape(-1, 9)
        tuv = UV[idx].reshape(-1, 6)
        cam = scenes.look_at_camera((3.2, 1.1, 0.9), (0, 0, 0), (0, 0, 1), 40.0, res, res)
        eye, look, up, fov = cam["lookat"][0:3], cam["lookat"][3:6], cam["lookat"][6:9], cam["fov"]
        mats = np.array([[1, .5, .4, .3, .4, .4, .4, .12]], np.float32)
        light, I = np.array([4.0, -2.0, 3.0], np.float32), (40.0, 40.0, 40.0)
        flags = np.full(len(tv), 2 | (1 if flip else 0), np.uint8)
        out.update(ref=dict(P=P, indices=idx, uv=UV, normals=N, reverse_orientation=flip),
                   render=dict(eye=eye, look=look, up=up, fov=fov, light_pos=light, light_I=I, kd=(.5, .5, .5),
                               prim_mat=np.repeat(mats, len(tv), 0)),
                   integ=dict(camera=cam, light_pos=light, light_I=I, materials=mats, prim_material=np.zeros(len(tv), np.uint8)),
                   accel=(tv, dict(tri_uvs=tuv, tri_normals=tn, prim_flags=flags)))
        return out
    raise KeyError(name)


CASES = ["sponza_point", "killeroo_plastic", "killeroo_normals", "killeroo_glass", "killeroo_metal", "killeroo_oren", "killeroo_uber", "killeroo_infinite",
         "feet_distant", "normals", "normals_flipped", "living_room", "living_room_tex", "sponza_two_points", "killeroo_lights",
         "killeroo_lights_power", "killeroo_lights_spatial", "living_room_tex_ewa"]
