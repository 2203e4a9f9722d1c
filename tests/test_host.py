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


def test_udf_plugin_libraries_export_the_netsdb_abi():
    """Every user type of the path as its own lib<Type>.so, built from the reference's unmodified one-line
    GET_V_TABLE(Type) sources against the shim classes: the three extern "C" symbols a netsDB catalog / worker binds
    (objectModel/headers/GetVTable.h:8-58) are there and answer."""
    import ctypes as C
    import glob

    from netsdb_b200 import build

    if os.path.exists("/root/reference"):
        build.build_udf_libraries()
    libs = sorted(glob.glob(os.path.join(build.HOST_LIBRARIES, "lib*.so")))
    if not libs:
        pytest.skip("UDF libraries were not built (no reference tree at build time)")
    expected = {t for types in build.REF_UDF_SOURCES.values() for t in types}
    assert {os.path.basename(p)[3:-3] for p in libs} == expected
    for path in libs:
        t = os.path.basename(path)[3:-3]
        lib = C.CDLL(path, mode=getattr(os, "RTLD_LOCAL", 0) | os.RTLD_LAZY)
        lib.getObjectTypeName.restype = C.c_char_p
        lib.getObjectVTable.restype = C.c_void_p
        lib.setAllGlobalVariables.argtypes = [C.c_void_p] * 4
        lib.setAllGlobalVariables(None, None, None, None)
        name = lib.getObjectTypeName().decode()
        assert name in (t, "conv2d_memory_fusion::" + t), (path, name)
        assert lib.getObjectVTable()


def test_l1_stream_wave_plan_covers_every_tile_once():
    """Host logic of the streamed layer-1 kernel (nb200_l1stream.cu: plan_l1_stream): every (tile, K block) is covered by
    exactly one segment, a tile's partial slots are 0..S-1 with S = tile_slots[tile], and no plan is much worse than the
    ideal tiles * K / clusters (the row-shard sizes of BASELINE config 3 included)."""
    import ctypes as C

    import numpy as np

    from netsdb_b200 import build

    L = C.CDLL(build.build())
    L.nb200_debug_plan_l1.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    for tiles, num_kb, G in ((20, 9337, 15), (40, 9337, 33), (10, 9337, 15), (5, 9337, 15), (3, 9337, 15), (5, 9337, 33),
                             (1, 65, 15), (7, 40, 4), (64, 100, 15), (1, 3, 8), (2, 9, 33)):
        out = np.zeros(5 * 4096, dtype=np.int32)
        slots = np.zeros(tiles, dtype=np.int32)
        n = L.nb200_debug_plan_l1(tiles, num_kb, G, out.ctypes.data, out.size, slots.ctypes.data)
        assert n > 0
        segs = out[:5 * n].reshape(n, 5)
        cover = np.zeros((tiles, num_kb), dtype=np.int32)
        per_cluster = np.zeros(G, dtype=np.int64)
        for g, t, k0, k1, slot in segs:
            assert 0 <= g < G and 0 <= t < tiles and 0 <= k0 < k1 <= num_kb and 0 <= slot < slots[t]
            cover[t, k0:k1] += 1
            per_cluster[g] += k1 - k0
        assert (cover == 1).all(), (tiles, num_kb, G)
        for t in range(tiles):
            assert sorted(s[4] for s in segs if s[1] == t) == list(range(slots[t]))
        ideal = tiles * num_kb / G
        assert per_cluster.max() <= max(1.35 * ideal + 110 * (1 + tiles // G), num_kb / max(1, min(G, num_kb // 8)) + 1), (tiles, num_kb, G, per_cluster.max(), ideal)


def test_reference_side_stub_compiles_against_the_reference_headers():
    """INTEGRATION.md section 2 / SURVEY 8f-1 (the reachable part): netsdb_b200/host/reference_side/FFGpuStage.h is a real
    file, written against netsDB's own FFMatrixBlock.h and the C ABI only; `make -C oracle ref` compiles it with the
    reference's unmodified header wherever the reference tree is mounted.  The object must reference exactly the entry
    points the stub claims to bind."""
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    obj = os.path.join(root, "oracle", "_ref", "ffgpustage_check.o")
    if os.path.exists("/root/reference/src/FF/headers/FFMatrixBlock.h"):
        subprocess.check_call(["make", "-C", os.path.join(root, "oracle"), "_ref/ffgpustage_check.o"], stdout=subprocess.DEVNULL)
    if not os.path.exists(obj):
        import pytest

        pytest.skip("no reference tree here and no prebuilt object")
    und = subprocess.run(["nm", "-u", obj], capture_output=True, text=True).stdout
    for sym in ("nb200_upload_block", "nb200_download_block", "nb200_set_info", "nb200_set_create", "nb200_set_lookup",
                "nb200_ff_inference_unit", "nb200_last_error"):
        assert sym in und, sym
    header = open(os.path.join(root, "include", "netsdb_b200.h")).read()
    stub = open(os.path.join(root, "netsdb_b200", "host", "reference_side", "FFGpuStage.h")).read()
    import re

    for sym in set(re.findall(r"\b(nb200_[a-z0-9_]+)\(", stub)):
        assert re.search(r"\b" + sym + r"\(", header), f"{sym} is not declared in the public header"
