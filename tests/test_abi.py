"""The C-ABI library loads and exports every symbol include/orb_gpu.h declares (not gpu; no
compute calls), and compute entry points fail loudly -- never fall back -- without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "orb_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orb_gpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(gpu):
    names = _declared()
    assert len(names) >= 20
    L = ctypes.CDLL(os.path.join(ROOT, "c_orb_slam_b200", "liborb_gpu.so"))
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert b"sm_100a" in gpu.lib().orb_gpu_version()
    assert set(gpu.api.EXPORTS) <= set(names)


def test_struct_layouts(gpu):
    assert gpu.KP_DTYPE.itemsize == 28 and gpu.PROJ_DTYPE.itemsize == 24
    assert gpu.ORBmatcher.TH_HIGH == 100 and gpu.ORBmatcher.TH_LOW == 50 and gpu.ORBmatcher.HISTO_LENGTH == 30


def test_argument_errors_and_no_fallback(gpu):
    import torch
    h = ctypes.c_void_p()
    assert gpu.lib().orb_gpu_create(0, 1.2, 8, 20, 7, 640, 480, 1, 0, ctypes.byref(h)) == -1   # nfeatures < 1
    assert gpu.lib().orb_gpu_create(1000, 1.0, 8, 20, 7, 640, 480, 1, 0, ctypes.byref(h)) == -1  # scale <= 1
    assert gpu.lib().orb_gpu_create(1000, 1.2, 8, 20, 7, 640, 480, 1, 0, None) == -1
    if not torch.cuda.is_available():
        with pytest.raises(gpu.OrbGpuError) as e:
            gpu.ORBextractor()
        assert e.value.code == -3                                                      # ERR_CUDA, loudly
        with pytest.raises(gpu.OrbGpuError):
            gpu.ORBmatcher().match_bruteforce(np.zeros((1, 32), np.uint8), np.zeros((4, 32), np.uint8))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "c_orb_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                text = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in text and "liborb_oracle" not in text and "oracle/" not in text.replace("oracle/ ", ""), f
