"""CPU tests of the host side: the C-ABI library loads and exports every declared symbol, the
header and the ctypes binding agree, and calls fail loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "netsdb_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nb200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from netsdb_b200 import build, capi

    build.build()
    L = ctypes.CDLL(capi.LIB_PATH)
    declared = _header_symbols()
    assert len(declared) >= 25
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, f"declared in include/netsdb_b200.h but not exported: {missing}"
    assert sorted(capi.EXPORTS) == declared, "capi.EXPORTS and the header disagree"


def test_calls_fail_loudly_without_init():
    from netsdb_b200 import capi

    p = ctypes.c_void_p()
    st = capi.lib().nb200_set_create(b"ff", b"x", 4, 4, 2, 2, ctypes.byref(p))
    if st == 0:  # a previous test in this process initialised the library on a GPU box
        capi.lib().nb200_set_destroy(p)
        pytest.skip("library already initialised")
    assert st == 5 and b"nb200_init" in capi.lib().nb200_last_error()


def test_no_cpu_fallback_when_no_gpu():
    import torch

    from netsdb_b200 import capi

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.Nb200Error):
        capi.init(0)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "netsdb_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "ff_oracle" not in text and "import oracle" not in text and "from oracle" not in text, f


def test_cpp_shim_drivers_compile():
    """The C++ host shim: our FFTest driver and -- where the reference tree is present -- the reference's own
    unmodified src/tests/source/FFTest.cc both compile and link against libnetsdb_b200.so."""
    from netsdb_b200 import build

    exes = build.build_host()
    assert os.path.exists(exes["FFTest"])
    if os.path.exists(build.REF_DRIVER):
        assert os.path.exists(exes["FFTest_netsdb"])
