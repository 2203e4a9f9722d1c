"""Builds libnetsdb_b200.so (sm_100a only) in-tree with nvcc.  No torch involved."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnetsdb_b200.so")
SOURCES = ["nb200_api.cu", "nb200_gemm.cu", "nb200_l1stream.cu", "nb200_kernels.cu", "nb200_pool.cu"]
HEADERS = ["nb200_internal.h", "nb200_ptx.cuh", os.path.join("..", "..", "include", "netsdb_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-ccbin", "/usr/bin/g++"]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    objs = []
    for src in SOURCES:
        obj = os.path.join(CSRC, src.replace(".cu", ".o"))
        cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        objs.append(obj)
    cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", "/usr/bin/g++",
           "-cudart", "shared"]
    subprocess.check_call(cmd)
    return LIB


HOST = os.path.join(HERE, "host")
HOST_BIN = os.path.join(HOST, "bin")
HOST_SRCS = [os.path.join(HOST, "src", f) for f in ("SimpleFF.cc", "FFMatrixUtil.cc", "ImageUtils.cc")]
REF_DRIVER = "/root/reference/src/tests/source/FFTest.cc"
REF_LSTM_DRIVER = "/root/reference/src/tests/source/LSTMTest.cc"
REF_LA_DRIVERS = [("TestLA04_netsdb", "/root/reference/src/tests/source/TestLA04_Multiply.cc"),
                  ("TestLA07_netsdb", "/root/reference/src/tests/source/TestLA07_TransposeMultiply_Gram.cc"),
                  ("Word2Vec_netsdb", "/root/reference/src/word2vec/source/Word2Vec.cc"),
                  ("PipelinedConv2dMemFuseTest_netsdb", "/root/reference/src/tests/source/PipelinedConv2dMemFuseTest.cc"),
                  ("FFTestWithoutDeduplication_netsdb", "/root/reference/src/tests/source/FFTestWithoutDeduplication.cc"),
                  ("FCProjTest_netsdb", "/root/reference/src/tests/source/FCProjTest.cc"),
                  ("TestSemanticClassifier_netsdb", "/root/reference/src/word2vec/source/TestSemanticClassifier.cc")]


def build_host(force: bool = False) -> dict:
    """Compile the C++ host shim drivers against libnetsdb_b200.so.

    FFTest          our driver in the shape of the reference's (adds a label dump for the parity tests)
    FFTest_netsdb   the reference's own, UNMODIFIED src/tests/source/FFTest.cc compiled in place against the
                    shim headers (only where /root/reference exists; the binary travels to the GPU box).
    LSTMTest / LSTMTest_netsdb   the same pair for the LSTM cell (src/tests/source/LSTMTest.cc).
    TestLA04_netsdb / TestLA07_netsdb   the LA DSL's multiply and Gram-matrix drivers (TestLA04_Multiply.cc,
                    TestLA07_TransposeMultiply_Gram.cc), unmodified.
    Word2Vec_netsdb the word2vec embedding driver (src/word2vec/source/Word2Vec.cc), unmodified.
    FCProjTest_netsdb / TestSemanticClassifier_netsdb   the drivers of the two network-in-a-SelectionComp UDFs
                    (src/tests/source/FCProjTest.cc, src/word2vec/source/TestSemanticClassifier.cc), unmodified."""
    build()
    os.makedirs(HOST_BIN, exist_ok=True)
    inc = ["-I" + os.path.join(HERE, "..", "include"), "-I" + os.path.join(HOST, "include")]
    common = ["/usr/bin/g++", "-std=c++17", "-O2", "-ffp-contract=off", *inc]
    link = ["-L" + HERE, "-lnetsdb_b200", "-Wl,-rpath,$ORIGIN/../.."]
    out = {}
    targets = [("FFTest", os.path.join(HOST, "src", "FFTest.cc")), ("LSTMTest", os.path.join(HOST, "src", "LSTMTest.cc")),
               ("Conv2dMemFuseTest", os.path.join(HOST, "src", "Conv2dMemFuseTest.cc"))]
    if os.path.exists(REF_DRIVER):
        targets.append(("FFTest_netsdb", REF_DRIVER))
    if os.path.exists(REF_LSTM_DRIVER):
        targets.append(("LSTMTest_netsdb", REF_LSTM_DRIVER))
    for name, src in REF_LA_DRIVERS:
        if os.path.exists(src):
            targets.append((name, src))
    for name, main_src in targets:
        exe = os.path.join(HOST_BIN, name)
        deps = [main_src, *HOST_SRCS, LIB] + [os.path.join(HOST, "include", f) for f in os.listdir(os.path.join(HOST, "include"))]
        if force or not os.path.exists(exe) or any(os.path.getmtime(d) > os.path.getmtime(exe) for d in deps):
            subprocess.check_call([*common, main_src, *HOST_SRCS, *link, "-o", exe])
        out[name] = exe
    return out


# The reference's one-line GET_V_TABLE(Type) sources, compiled unmodified into lib<Type>.so (the netsDB plugin ABI)
HOST_LIBRARIES = os.path.join(HOST, "libraries")
REF_UDF_SOURCES = {
    "/root/reference/src/FF/source": ["FFMatrixMeta", "FFMatrixData", "FFMatrixBlock", "FFMatrixBlockScanner", "FFMatrixWriter",
                                      "FFInputLayerJoin", "FFAggMatrix", "FFReluBiasSum", "FFTransposeMult", "FFTransposeBiasSum",
                                      "FFRowAggregate", "FFOutputLayer"],
    "/root/reference/src/LSTM/source": ["LSTMThreeWaySum", "LSTMTwoSum", "LSTMHiddenState"],
    "/root/reference/src/FF_proj/source": ["FullyConnectedNetwork"],
    "/root/reference/src/word2vec/source": ["SemanticClassifier"],
    "/root/reference/src/conv2d_memory_fusion/source": ["Image", "Kernel", "ImageChunk", "ImageBlock", "ImageToChunks", "ImageChunksToBlock",
                                                        "ImageBlockToMatrix", "KernelToChunks", "KernelBiasJoin", "ConvResultToChunks",
                                                        "ConvChunksToImage"],
    "/root/reference/src/sharedLibraries/source": ["MatrixMeta", "MatrixData", "MatrixBlock", "LAScanMatrixBlockSet", "LAWriteMatrixBlockSet",
                                                   "LASillyMultiply1Join", "LASillyTransposeMultiply1Join", "LASillyMultiply2Aggregate"],
}


def build_udf_libraries(force: bool = False) -> list:
    """lib<Type>.so for every type of the path, from the reference's own GET_V_TABLE sources (only where the reference
    tree is present; the libraries travel with the repository snapshot like the other built binaries)."""
    os.makedirs(HOST_LIBRARIES, exist_ok=True)
    inc = ["-I" + os.path.join(HERE, "..", "include"), "-I" + os.path.join(HOST, "include")]
    out = []
    for src_dir, types in REF_UDF_SOURCES.items():
        for t in types:
            src = os.path.join(src_dir, t + ".cc")
            lib = os.path.join(HOST_LIBRARIES, f"lib{t}.so")
            if os.path.exists(src) and (force or not os.path.exists(lib) or os.path.getmtime(src) > os.path.getmtime(lib)
                                        or os.path.getmtime(os.path.join(HOST, "include", "GetVTable.h")) > os.path.getmtime(lib)):
                subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-shared", "-fPIC", *inc, src, "-o", lib])
            if os.path.exists(lib):
                out.append(lib)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
    print(len(build_udf_libraries(force="--force" in sys.argv)), "UDF libraries")
