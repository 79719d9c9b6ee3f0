"""Generates tests/golden/living_room_mesh.npz from the reference's OWN scene (BASELINE config 5):
/root/reference/scenes/living-room/scene.pbrt + models/Mesh0NN.ply.  Run in the build container only
(/root/reference does not exist on the GPU box); the fixture is what travels.

What is restated from the reference (file:line):
* `Shape "plymesh"` (src/shapes/plymesh.cpp:100-147): binary little-endian PLY, vertex x y z [nx ny nz] [u v],
  faces of 3 or 4 indices; a quad (a b c d) becomes the triangles (a b c), (d a c).
* `Transform [16 floats]` (src/core/api.cpp pbrtTransform): column-major, so elements 12..14 are the translation;
  TransformBegin/End save and restore the CTM.  Triangle meshes are stored in WORLD space
  (src/shapes/triangle.cpp:46-48): p_world = ObjectToWorld(p), evaluated in float as
  ((m00*x + m01*y) + m02*z) + m03 (src/core/transform.h:183-192), which is what `xform32` does.
* primitives are appended in file order (api.cpp pbrtShape), so triangle k of this fixture is primitive k
  of the reference's aggregate: 143 163 triangles (SURVEY.md §6.2 measured 143 163 / 233 709 BVH nodes).
* `NamedMaterial` / `MakeNamedMaterial`: one material row per triangle; the table keeps the scene's type and
  colours (texture-mapped Kd noted by name)."""
import os
import re
import sys
import numpy as np

REF = "/root/reference/scenes/living-room"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_ply(path):
    raw = open(path, "rb").read()
    end = raw.index(b"end_header\n") + len(b"end_header\n")
    head = raw[:end].decode("ascii").splitlines()
    assert head[1].strip() == "format binary_little_endian 1.0", head[1]
    elems, cur = [], None
    for ln in head:
        t = ln.split()
        if t and t[0] == "element":
            cur = dict(name=t[1], n=int(t[2]), props=[])
            elems.append(cur)
        elif t and t[0] == "property":
            cur["props"].append(t[1:])
    off = end
    P = UV = faces = N = None
    for e in elems:
        if e["name"] == "vertex":
            names = [p[-1] for p in e["props"]]
            assert all(p[0] == "float" for p in e["props"])
            a = np.frombuffer(raw, "<f4", e["n"] * len(names), off).reshape(e["n"], len(names))
            off += a.nbytes
            P = a[:, [names.index(c) for c in "xyz"]].copy()
            UV = a[:, [names.index("u"), names.index("v")]].copy() if "u" in names else np.zeros((e["n"], 2), np.float32)
            N = a[:, [names.index(c) for c in ("nx", "ny", "nz")]].copy() if "nx" in names else None
        elif e["name"] == "face":
            assert e["props"] == [["list", "uint8", "int", "vertex_indices"]], e["props"]
            tris = []
            for _ in range(e["n"]):
                k = raw[off]
                f = np.frombuffer(raw, "<i4", k, off + 1)
                off += 1 + 4 * k
                if k == 3:
                    tris.append(f)
                elif k == 4:   # plymesh.cpp:137-146
                    tris.append(f[[0, 1, 2]])
                    tris.append(f[[3, 0, 2]])
            faces = np.asarray(tris, np.int32)
    assert off == len(raw), (path, off, len(raw))
    return P, UV, faces, N


def xform_normals(m, N):
    """TriangleMesh's `n[i] = ObjectToWorld(N[i])` (triangle.cpp:61-64) through the reference's own Transform (its Inverse()
    and operator()(Normal3f)); also returns Transform::SwapsHandedness()."""
    import ctypes as C
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import refbind as rb
    L = rb.lib()
    L.ref_transform_normals.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    m32 = np.ascontiguousarray(m, np.float32)
    N = np.ascontiguousarray(N, np.float32)
    out = np.zeros_like(N)
    swaps = L.ref_transform_normals(m32.ctypes.data, N.ctypes.data, len(N), out.ctypes.data)
    return out, bool(swaps)


def xform32(m, P):
    """Transform::operator()(Point3f) in float (transform.h:183-192); m is the row-major 4x4."""
    m = m.astype(np.float32)
    x, y, z = P[:, 0], P[:, 1], P[:, 2]
    out = np.empty_like(P)
    for r in range(3):
        out[:, r] = ((m[r, 0] * x + m[r, 1] * y) + m[r, 2] * z) + m[r, 3]
    w = ((m[3, 0] * x + m[3, 1] * y) + m[3, 2] * z) + m[3, 3]
    assert np.all(w == 1)
    return out


def main():
    txt = open(os.path.join(REF, "scene.pbrt")).read()
    world = txt[txt.index("WorldBegin"):]
    mats, order = {}, []
    for m in re.finditer(r'MakeNamedMaterial "([^"]+)" "string type" \[ "([^"]+)" \](.*)', world):
        name, kind, rest = m.group(1), m.group(2), m.group(3)
        kd = re.search(r'"rgb Kd" \[ ([\d.eE+-]+) ([\d.eE+-]+) ([\d.eE+-]+) \]', rest)
        tex = re.search(r'"texture Kd" \[ "([^"]+)" \]', rest)
        mats[name] = dict(kind=kind, kd=[float(v) for v in kd.groups()] if kd else None, tex=tex.group(1) if tex else "")
        order.append(name)
    ctm, stack, cur = np.eye(4), [], None
    Ps, UVs, Is, Ms, meshes, Ns, Fl = [], [], [], [], [], [], []
    base = 0
    for ln in world.splitlines():
        t = ln.strip()
        if t.startswith("TransformBegin"):
            stack.append(ctm.copy())
        elif t.startswith("TransformEnd"):
            ctm = stack.pop()
        elif t.startswith("Transform "):
            v = [float(s) for s in t[t.index("[") + 1:t.index("]")].split()]
            ctm = np.array(v).reshape(4, 4).T   # column-major in the file
        elif t.startswith("NamedMaterial"):
            cur = re.search(r'"([^"]+)"', t).group(1)
        elif t.startswith("Shape"):
            f = re.search(r'"string filename" \[ "([^"]+)" \]', t).group(1)
            P, UV, F, N = read_ply(os.path.join(REF, f))
            Nw, swaps = xform_normals(ctm, N if N is not None else np.zeros_like(P))
            Ns.append(Nw if N is not None else np.zeros_like(P))
            # bit 1: the mesh has normals; bit 0: transformSwapsHandedness (no ReverseOrientation in the file)
            Fl.append(np.full(len(F), (2 if N is not None else 0) | (1 if swaps else 0), np.uint8))
            Ps.append(xform32(ctm, P))
            UVs.append(UV)
            Is.append(F + base)
            Ms.append(np.full(len(F), order.index(cur), np.uint8))
            meshes.append(f)
            base += len(P)
    P, UV, I, M = np.concatenate(Ps), np.concatenate(UVs), np.concatenate(Is), np.concatenate(Ms)
    Nall, flags = np.concatenate(Ns), np.concatenate(Fl)
    print("meshes whose transform swaps handedness:", sum(int(f[0] & 1) for f in Fl), "of", len(Fl))
    print("meshes", len(meshes), "vertices", len(P), "triangles", len(I))
    assert len(I) == 143163, len(I)   # SURVEY.md §6.2 (the reference's own count for this scene)
    cam = re.search(r"^Transform \[([^\]]+)\]", txt, re.M).group(1).split()
    np.savez_compressed(os.path.join(HERE, "living_room_mesh.npz"), P=P, UV=UV, idx=I, tri_material=M, N=Nall, tri_flags=flags,
                        material_names=np.array(order), material_kinds=np.array([mats[n]["kind"] for n in order]),
                        material_kd=np.array([mats[n]["kd"] or [-1, -1, -1] for n in order], np.float32),
                        material_tex=np.array([mats[n]["tex"] for n in order]),
                        world_to_camera=np.array([float(c) for c in cam]).reshape(4, 4).T,
                        meshes=np.array(meshes))


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    main()
