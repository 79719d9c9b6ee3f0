"""Generates tests/golden/ref_render.npz: images rendered by the REFERENCE'S OWN renderer
(SamplerIntegrator::Render + PathIntegrator::Li + Film/BoxFilter + RandomSampler + PointLight + MatteMaterial +
BVHAccel, compiled from /root/reference by oracle/ref/Makefile; driver oracle/ref/ref_render.cpp), plus the
number of rays that reached Aggregate::Intersect / IntersectP.  Run in the build container:

    python tests/golden/make_golden_render.py
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
from oracle import refbind as rb   # noqa: E402
from b200rt import scenes          # noqa: E402

CASES = {
    # BASELINE config 1: scenes/killeroo-simple geometry + camera (point light at the sphere light's centre)
    "killeroo": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=5, I=(50000.0,) * 3, kd=(.5, .5, .5)),
    # BASELINE config 2: the sponza stand-in the bench renders
    # BASELINE config 1 with the scene's own light and Kd's: sphere AreaLightSource L = 2000 (MIS in EstimateDirect),
    # ground matte (.5 .5 .8), killeroos with the diffuse part of their "plastic" (Kd .4 .2 .2 / .4 .5 .4)
    "killeroo_area": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=5, I=(2000.0,) * 3, kd=(.5, .5, .5), area=True),
    # BASELINE config 1 as the scene file has it: area light + matte ground + the two "plastic" killeroos
    "killeroo_full": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=5, I=(2000.0,) * 3, kd=(.5, .5, .5), area=True, plastic=True),
    # the same scene with the killeroos made perfectly specular (glass index 1.5 / mirror Kr .9): specularBounce, etaScale,
    # FresnelSpecular and SpecularReflection of the reference; maxdepth 8 so that Russian roulette sees etaScale
    "killeroo_glass": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=8, I=(2000.0,) * 3, kd=(.5, .5, .5), area=True, specular=True),
    # ... and with scenes/living-room's microfacet materials: substrate ground, metal (conductor Fresnel) and substrate killeroos
    "killeroo_metal": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=5, I=(2000.0,) * 3, kd=(.5, .5, .5), area=True, metal=True),
    # ... the same with an Oren-Nayar ground (Material "matte" "float sigma" [40])
    "killeroo_oren": dict(scene="killeroo_simple", res=128, spp=256, maxdepth=5, I=(2000.0,) * 3, kd=(.5, .5, .5), area=True, metal="oren"),
    # BASELINE config 4's scene as written (scenes/killeroo-feet: LightSource "distant" from the eye towards the look-at
    # point, Material "plastic" Kd (.4 .5 .4) Ks .3 roughness .15, path maxdepth 5), whole frame instead of the crop window
    "feet": dict(scene="killeroo_feet", res=192, spp=256, maxdepth=5, I=(1.0,) * 3, kd=(.5, .5, .5), distant=True),
    # ... and through its Film "cropwindow" [.3 .5 .4 .8] (the feet): 640^2 frame -> 128 x 256 pixel window
    "feet_crop": dict(scene="killeroo_feet", res=640, spp=64, maxdepth=5, I=(1.0,) * 3, kd=(.5, .5, .5), distant=True, crop=True),
    "sponza": dict(scene="sponza_standin", res=96, spp=256, maxdepth=5, I=(1.0,) * 3, kd=(.5, .5, .5)),
    # BASELINE config 3: the museum stand-in as scenes/test_museum_local_k renders it — Accelerator "kdtree", path maxdepth 1
    "museum": dict(scene="museum_standin", res=96, spp=256, maxdepth=1, I=(1.0,) * 3, kd=(.5, .5, .5), accel="kdtree"),
}


def lookat_of(name):
    """eye, look, up, fov of the scene templates (scenes/killeroo-simple:1-3, scenes/sponza:1-3)."""
    if name in ("killeroo_simple", "killeroo_feet"):
        r = scenes._rot_z(5.0)
        eye = scenes._xform(r, np.array([[400.0, 20, 30]]))[0]
        look = scenes._xform(r, np.array([[0.0, 63, -110]]))[0]
        return eye, look, (0, 0, 1), 39.0
    if name == "museum_standin":   # scenes/museum:1-3: LookAt -0.4 0.06 0.05  0.2 0.1 -0.05  0 1 0 ; Rotate 180 0 1 0 ; fov 60
        return (0.4, 0.06, -0.05), (-0.2, 0.1, 0.05), (0, 1, 0), 60.0
    return (-0.7, .45, .05), (0, .25, 0), (0, 1, 0), 42.0


def main():
    path = os.path.join(ROOT, "tests", "golden", "ref_render.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}
    for key, c in CASES.items():
        if sys.argv[1:] and key not in sys.argv[1:]:
            continue
        sc = getattr(scenes, c["scene"])(xres=c["res"], yres=c["res"])
        tv = sc["tri_verts"]
        extra = {}
        spheres = None
        if c.get("area"):
            sl = sc["sphere_light"]
            spheres = np.array([list(sl["center"]) + [sl["radius"]]], np.float32)
            mats = sc["materials"] if c.get("plastic") else np.pad(sc["material_kd"], ((0, 0), (1, 4)))
            if c.get("specular"):
                mats = scenes.specular_materials(sc)
            if c.get("metal"):
                mats = scenes.metal_materials(sc, oren_nayar=c["metal"] == "oren")
            extra = dict(prim_mat=np.concatenate([mats[sc["prim_material"]], np.zeros((1, 8), np.float32)]), emitter_L=sl["L"])
        if c.get("distant"):
            extra = dict(prim_mat=np.repeat(sc["materials"], len(tv), 0), distant=(sc["light_from"], sc["light_to"]))
            sc["light"] = np.zeros(3, np.float32)
        if c.get("crop"):
            extra["cropwindow"] = sc["cropwindow"]
        S = rb.RefScene(tv.reshape(-1, 3), np.arange(3 * len(tv), dtype=np.int32).reshape(-1, 3), spheres=spheres)
        eye, look, up, fov = lookat_of(c["scene"])
        img, nc, ns, heat = rb.render(S, c.get("accel", "bvh"), eye, look, up, fov, c["res"], c["res"], c["spp"], c["maxdepth"], sc["light"],
                                      c["I"], c["kd"], sampler="random", n_threads=0, heat=True, **extra)
        if key == "killeroo":   # the thesis's per-pixel traversal statistics for the same frame (BVH)
            out[key + "_heat"] = heat
        print(key, "mean", img.mean(), "closest", nc, "shadow", ns)
        out[key + "_rgb"] = img
        out[key + "_rays"] = np.array([nc, ns], np.uint64)
        out[key + "_params"] = np.array([c["res"], c["spp"], c["maxdepth"]], np.int32)
        out[key + "_I"] = np.array(c["I"], np.float32)
        out[key + "_kd"] = np.array(c["kd"], np.float32)
    np.savez_compressed(path, **out)


if __name__ == "__main__":
    main()
