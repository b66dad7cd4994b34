"""The C-ABI library loads without a GPU, exports every symbol include/spray_gpu.h declares, and its
compute entry points fail loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os

import pytest

from spray_b200 import capi


def test_library_builds_and_exports_all_declared_symbols():
    capi.build()
    L = capi.lib()
    declared = capi.declared_symbols()
    assert len(declared) >= 26
    for name in declared:
        assert hasattr(L, name), "missing export: " + name
    assert set(capi._PROTOS) == set(declared)
    assert b"sm_100a" in L.spraygpu_version()


def test_sass_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out


def test_no_cpu_fallback():
    """Without a CUDA device device_create must fail with an error, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = capi.lib().spraygpu_device_create(0, C.byref(h))
    assert rc != 0 and not h.value
    assert b"no CPU fallback" in capi.lib().spraygpu_last_error()
    with pytest.raises(capi.SprayGpuError):
        capi.Device(0)


def test_product_never_touches_oracle():
    """Nothing under spray_b200/ or host/ may import, include or link the oracle."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for base in ("spray_b200", "host", "include"):
        for dp, _, files in os.walk(os.path.join(root, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", ".hpp", "Makefile")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    assert "pyoracle" not in txt and "spray_oracle" not in txt and "orc_" not in txt, os.path.join(dp, f)
