"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/b200_nonbonded.h declares."""
import ctypes
import os
import re

import openmm_b200 as b

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_nonbonded.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_python_binding_agree():
    assert declared_symbols() == sorted(b.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(b.LIB_PATH), "run python -m openmm_b200.build"
    lib = ctypes.CDLL(b.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_no_cpu_fallback_without_gpu():
    # creating a context on a machine without a GPU must fail loudly, never fall back
    import numpy as np
    import pytest
    lib = b.load_library()
    if lib.b200_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(b.B200Error):
        b.Context(np.ones(8), [3, 0, 0, 0, 3, 0, 0, 0, 3])


def test_product_code_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "openmm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "liboracle" not in text and "nonbonded_oracle" not in text or f == "build.py", f


def test_legal_fft_dimension_host_helper():
    lib = b.load_library()
    for n in (1, 7, 11, 13, 97, 121, 197):
        assert lib.b200_find_legal_fft_dimension(n) == b.find_legal_fft_dimension(n)
