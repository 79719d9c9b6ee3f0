"""`Texture "name" "spectrum" "imagemap"` for the wavefront driver (include/b200rt_wavefront.h, ABI v9): the MIPMap pyramid of an
image as the reference builds it (b200rt_mipmap_build: ImageTexture::GetTexture + MIPMap::MIPMap, textures/imagemap.cpp:52-97,
core/mipmap.h:113-209) and small TGA helpers (type 2, 24 bit: what scenes/living-room ships)."""
import ctypes as C
import struct
import numpy as np
from . import api


class Texture(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32), ("n_levels", C.c_uint32),
                ("su", C.c_float), ("sv", C.c_float), ("du", C.c_float), ("dv", C.c_float),
                ("filter", C.c_uint32), ("max_anisotropy", C.c_float)]


class ImageTexture:
    """rgb8: (h, w, 3) uint8 with row 0 = the top of the image (ReadImage's order), or rgb: float32 of the same shape.
    gamma: InverseGammaCorrect on the way in (the reference's default for .tga / .png); scale = "float scale".
    trilinear: "bool trilinear" (False, the reference's default, = the EWA filter with "float maxanisotropy")."""

    def __init__(self, rgb8=None, rgb=None, gamma=True, scale=1.0, su=1.0, sv=1.0, du=0.0, dv=0.0, trilinear=True, maxanisotropy=8.0):
        self.trilinear, self.maxanisotropy = bool(trilinear), float(maxanisotropy)
        if rgb is None:
            rgb = np.asarray(rgb8, np.uint8).astype(np.float32) / np.float32(255)     # c = src / 255.f (imageio.cpp:245-247)
        rgb = np.ascontiguousarray(rgb, np.float32)
        h, w = rgb.shape[:2]
        L = api.lib()
        L.b200rt_mipmap_dims.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.b200rt_mipmap_build.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_float, C.c_void_p]
        w0, h0, nl, nt = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint64()
        api._check(L.b200rt_mipmap_dims(w, h, C.byref(w0), C.byref(h0), C.byref(nl), C.byref(nt)))
        self.texels = np.zeros((nt.value, 3), np.float32)
        api._check(L.b200rt_mipmap_build(rgb.ctypes.data, w, h, int(bool(gamma)), float(scale), self.texels.ctypes.data))
        self.width, self.height, self.n_levels = w0.value, h0.value, nl.value
        self.su, self.sv, self.du, self.dv = su, sv, du, dv

    def level(self, l):
        off = sum(max(1, self.width >> k) * max(1, self.height >> k) for k in range(l))
        w, h = max(1, self.width >> l), max(1, self.height >> l)
        return self.texels[off:off + w * h].reshape(h, w, 3)

    def record(self):
        return Texture(self.texels.ctypes.data, self.width, self.height, self.n_levels, self.su, self.sv, self.du, self.dv,
                       0 if self.trilinear else 1, self.maxanisotropy)


def texture_array(textures):
    """ctypes array of b200rt_texture for RenderDesc.textures (keep `textures` alive while it is in use)."""
    arr = (Texture * max(1, len(textures)))()
    for i, t in enumerate(textures):
        arr[i] = t.record()
    return arr


def write_tga(path, rgb8):
    """Uncompressed 24-bit TGA, top-to-bottom rows (descriptor 0x20), BGR pixels — the layout of the reference's own files."""
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    h, w = rgb8.shape[:2]
    with open(path, "wb") as f:
        f.write(struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, w, h, 24, 0x20))
        f.write(np.ascontiguousarray(rgb8[..., ::-1]).tobytes())


def read_tga(path):
    raw = open(path, "rb").read()
    idlen, cmap, typ = raw[0], raw[1], raw[2]
    w, h, bpp, desc = struct.unpack("<HHBB", raw[12:18])
    if typ != 2 or cmap != 0 or bpp != 24:
        raise ValueError("only uncompressed 24-bit true-colour TGA files are read here")
    px = np.frombuffer(raw, np.uint8, w * h * 3, 18 + idlen).reshape(h, w, 3)[..., ::-1]
    if not (desc & 0x20):
        px = px[::-1]
    if desc & 0x10:
        px = px[:, ::-1]
    return np.ascontiguousarray(px)
