"""Generates tests/golden/living_room_textures.npz from the reference's OWN scene (BASELINE config 5): the two image files of
scenes/living-room/textures that ARE in the reference tree — leaf.tga (Kd and opacity of the "Leaves" uber material, Texture04 /
Texture05) and picture8.tga (Kd of the "Painting" matte material, Texture03) — as 8-bit RGB arrays with row 0 = the top of the
image, i.e. the pixels ReadImageTGA (src/core/imageio.cpp:216-256) hands to ImageTexture before the division by 255.
(textures/wood.tga and wood5.tga, Texture01 / Texture02, are absent from the tree: the reference replaces them by a constant
grey 0.5, imagemap.cpp:67-72.)  Run in the build container only; tests write these arrays back into .tga files for the
reference harness (tests/textures.py)."""
import os
import struct
import sys
import numpy as np

REF = "/root/reference/scenes/living-room/textures"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_tga(path):
    raw = open(path, "rb").read()
    idlen, cmap, typ = raw[0], raw[1], raw[2]
    w, h, bpp, desc = struct.unpack("<HHBB", raw[12:18])
    assert typ == 2 and cmap == 0 and bpp == 24, (typ, cmap, bpp)      # uncompressed true colour
    px = np.frombuffer(raw, np.uint8, w * h * 3, 18 + idlen).reshape(h, w, 3)[..., ::-1]   # BGR -> RGB
    if not (desc & 0x20):
        px = px[::-1]          # bottom-to-top file: tga_flip_vert
    if desc & 0x10:
        px = px[:, ::-1]       # right-to-left file: tga_flip_horiz
    return np.ascontiguousarray(px)


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    out = {n: read_tga(os.path.join(REF, n + ".tga")) for n in ("leaf", "picture8")}
    for k, v in out.items():
        print(k, v.shape, v.dtype)
    np.savez_compressed(os.path.join(HERE, "living_room_textures.npz"), **out)
