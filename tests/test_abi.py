"""CPU: the C-ABI library loads, exports every symbol include/*.h declares, and refuses to compute
without a CUDA device (no CPU fallback)."""
import ctypes
import glob
import os
import re
import numpy as np
import pytest
from conftest import ROOT


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        txt = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        txt = re.sub(r"//.*", "", txt)
        names |= set(re.findall(r"\b(b200rt_[a-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_exports_every_declared_symbol(built_lib):
    L = ctypes.CDLL(built_lib.lib_path())
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(L, s), "libb200rt.so does not export %s" % s


def test_version(built_lib):
    assert built_lib.lib().b200rt_version() == 10


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(built_lib.B200RTError) as e:
        built_lib.init(0)
    assert e.value.code == 2  # ENODEVICE
    v = np.random.default_rng(0).random((8, 9), dtype=np.float32)
    tree = built_lib.build_bvh(v)  # host builder works anywhere
    with pytest.raises(built_lib.B200RTError):
        built_lib.B200Accel(tree, v)


def test_make_accelerator_names(built_lib):
    with pytest.warns(UserWarning):
        assert built_lib.MakeAccelerator("nonsense", np.zeros((1, 9), np.float32)) is None


def test_product_never_imports_oracle():
    """The product package must not reference oracle/ (oracle = test infrastructure only)."""
    pkg = os.path.join(ROOT, "thesis-pbrt-v3_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.lower().replace("test oracle", ""), os.path.join(dirpath, f)
