"""Writes thesis-pbrt-v3_b200/b200rt/data/sobol_tables.npz: the generator matrices of the reference's Sampler "sobol"
(SobolMatrices32 1024 x 52, VdCSobolMatrices / VdCSobolMatricesInv 25 x 52; src/core/sobolmatrices.cpp), read out of the
compiled reference (oracle/_ref/libpbrtref.so).  They are DATA the host hands to b200rt_render (b200rt_render_desc::
sobol_matrices32 ...), exactly like a host-built tree: the C++ drop-in passes the reference's own arrays, the Python mirror
loads this file.  Run in the build container:  python tests/golden/make_sobol_tables.py"""
import ctypes as C
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind as rb   # noqa: E402

L = rb.lib()
m32 = np.ctypeslib.as_array((C.c_uint32 * (1024 * 52)).in_dll(L, "_ZN4pbrt15SobolMatrices32E")).copy()
vdc = np.ctypeslib.as_array((C.c_uint64 * (25 * 52)).in_dll(L, "_ZN4pbrt16VdCSobolMatricesE")).copy().reshape(25, 52)
inv = np.ctypeslib.as_array((C.c_uint64 * (25 * 52)).in_dll(L, "_ZN4pbrt19VdCSobolMatricesInvE")).copy().reshape(25, 52)
out = os.path.join(ROOT, "thesis-pbrt-v3_b200", "b200rt", "data", "sobol_tables.npz")
np.savez_compressed(out, matrices32=m32, vdc=vdc, vdc_inv=inv)
print(out, os.path.getsize(out), "bytes; first words", m32[:3], vdc[0, :2], inv[0, :2])
