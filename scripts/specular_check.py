"""Noise study for the killeroo_glass golden: renders the case with several seeds on the B200 and prints, next to the
reference render's figures, the masked mean, the clamped mean and the ray counts (how much of the difference to the
golden is Monte-Carlo noise of the caustic paths, how much is bias)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "thesis-pbrt-v3_b200"))
import b200rt as gpu
from b200rt import scenes
from b200rt.wavefront import film_to_rgb

g = np.load(os.path.join(ROOT, "tests", "golden", "ref_render.npz"))
case = sys.argv[1] if len(sys.argv) > 1 else "killeroo_glass"
res, spp, maxdepth = (int(x) for x in g[case + "_params"])
gpu.init(0)
sc = scenes.killeroo_simple(xres=res, yres=res)
verts, prim_kind, rows, pm, mats = scenes.with_sphere_light(sc, plastic=True)
if case == "killeroo_glass":
    mats = scenes.specular_materials(sc)
if case in ("killeroo_metal", "killeroo_oren"):
    mats = scenes.metal_materials(sc, oren_nayar=case == "killeroo_oren")
acc = gpu.B200Accel(gpu.build_bvh(verts), verts, prim_kind=prim_kind, spheres=rows)
sl = sc["sphere_light"]
ref = g[case + "_rgb"].astype(np.float64)


def figures(img):
    m = img.max(-1)
    near = m > 50
    for _ in range(2):
        near = near | np.roll(near, 1, 0) | np.roll(near, -1, 0) | np.roll(near, 1, 1) | np.roll(near, -1, 1)
    a = np.where(near[..., None], 0, img)
    return a.mean(), np.minimum(a, 1.0).mean(), np.minimum(a, 0.3).mean(), (a.max(-1) > 2).sum()


def blocks(img, k=8):
    return img.reshape(res // k, k, res // k, k, 3).mean((1, 3))


def clamped(img):
    m = img.max(-1)
    near = m > 50
    for _ in range(2):
        near = near | np.roll(near, 1, 0) | np.roll(near, -1, 0) | np.roll(near, 1, 1) | np.roll(near, -1, 1)
    return np.minimum(np.where(near[..., None], 0, img), 1.0)


print("ref      ", figures(ref), g[case + "_rays"])
imgs = []
for seed in range(1, 7):
    integ = gpu.WavefrontPathIntegrator(sc["camera"], spp=spp, maxdepth=maxdepth, light_pos=sl["center"], light_I=sl["L"],
                                        seed=seed, sphere_light=(sl["radius"], len(verts) - 1), materials=mats, prim_material=pm)
    film, st = integ.Render(acc)
    imgs.append(clamped(film_to_rgb(film).astype(np.float64)))
    print("seed", seed, figures(film_to_rgb(film).astype(np.float64)), st["closest_rays"], st["shadow_rays"], st["mis_rays"])

# 8x8-block RMSE (relative to the mean) of the clamped images: seed against seed = pure Monte-Carlo noise of this driver,
# seed against the reference = noise of both + any bias; and of the 6-seed average, whose own noise is 1/sqrt(6) of a seed's
rc = clamped(ref)
rm = rc.mean()
rmse = lambda a, b: np.sqrt(((blocks(a) - blocks(b)) ** 2).mean()) / rm
print("block rmse seed vs seed:", [round(rmse(imgs[i], imgs[i + 1]), 4) for i in range(5)])
print("block rmse seed vs ref :", [round(rmse(im, rc), 4) for im in imgs])
avg = sum(imgs) / len(imgs)
print("block rmse avg6 vs ref :", round(rmse(avg, rc), 4), " mean ratio avg6/ref:", avg.mean() / rm)
d = blocks(avg).mean(-1) - blocks(rc).mean(-1)
idx = np.argsort(-np.abs(d).ravel())[:8]
print("largest block differences (by, bx, avg6, ref):", [(int(i // d.shape[1]), int(i % d.shape[1]), round(float(blocks(avg).mean(-1).ravel()[i]), 4), round(float(blocks(rc).mean(-1).ravel()[i]), 4)) for i in idx])
