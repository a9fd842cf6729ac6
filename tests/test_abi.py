"""CPU: the C-ABI library loads and exports every symbol include/eagle_b200.h declares, and the
product path fails loudly without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from conftest import ROOT, load_package


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "eagle_b200.h")).read()
    return sorted(set(re.findall(r"\b(eb_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    eb = load_package()
    assert os.path.exists(eb.LIB_PATH), "run `make` / __graft_entry__.build() first"
    lib = C.CDLL(eb.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 45
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in include/eagle_b200.h but not exported"
    out = subprocess.check_output(["nm", "-D", "--defined-only", eb.LIB_PATH], text=True)
    exported = set(re.findall(r" T (eb_[a-z_0-9]+)", out))
    assert exported == set(decl), exported ^ set(decl)


def test_python_mirror_binds_every_symbol():
    eb = load_package()
    lib = eb.lib()
    for name in declared_symbols():
        assert getattr(lib, name).argtypes is not None or name in ("eb_last_error", "eb_version", "eb_device_sm_count"), name


def test_pure_host_entry_points():
    eb = load_package()
    lib = eb.lib()
    assert lib.eb_version() >= 100
    assert lib.eb_row_size(eb.Q4_K, 4096) == 16 * 144 and lib.eb_row_size(eb.Q6_K, 256) == 210
    assert lib.eb_row_size(eb.Q4_0, 64) == 36 and lib.eb_row_size(eb.Q8_0, 32) == 34 and lib.eb_row_size(eb.Q4_K, 100) == 0
    assert lib.eb_mul_mat_ws_size(eb.Q4_K, 4096, 4096, 65) > 65 * 4096


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    eb = load_package()
    with pytest.raises(eb.EagleB200Error):
        eb.init(0)
    with pytest.raises(eb.EagleB200Error):
        eb.DeviceBuffer(16)


def test_product_package_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "eagle-in-llama.cpp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in src and "eagle_oracle" not in src and "libref_harness" not in src, f
