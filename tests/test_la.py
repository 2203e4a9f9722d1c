"""LA DSL multiply joins on MatrixBlock: LASillyMultiply1Join / LASillyTransposeMultiply1Join +
LASillyMultiply2Aggregate.  CPU: oracle vs numpy.  GPU: nb200_gemm_blocks(NB200_OP_NN / NB200_OP_TN) vs the oracle and
the reference's own, unmodified TestLA07 driver (Gram matrix) through the shim."""
import os
import re
import subprocess

import numpy as np
import pytest

TOL = {"fp64": 1e-12, "bf16x3": 1e-4, "f16": 2.5e-3, "bf16": 2e-2}


@pytest.mark.parametrize("shape", [(37, 23, 19, 10), (64, 64, 64, 32), (5, 3, 4, 7), (130, 70, 90, 25)])
def test_oracle_transpose_multiply(oracle, shape):
    K, I, J, bs = shape
    rng = np.random.default_rng(K)
    a, b = rng.normal(size=(K, I)), rng.normal(size=(K, J))
    r = oracle.la_transpose_multiply_agg(oracle.load_matrix(a, bs, bs), oracle.load_matrix(b, bs, bs), 3)
    assert r.totals() == (I, J)
    assert np.abs(r.dense() - a.T @ b).max() <= 1e-12 * np.abs(a.T @ b).max()
    with pytest.raises(RuntimeError):  # "Block dimemsions mismatch!": the row blocks of the operands differ in height
        oracle.la_transpose_multiply_agg(oracle.load_matrix(a, bs, bs), oracle.load_matrix(b, bs + 1, bs), 1)


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp64", "bf16x3", "f16", "bf16"])
@pytest.mark.parametrize("shape", [(370, 230, 190, 100), (64, 64, 64, 32), (5, 3, 4, 7), (2500, 300, 129, 100)])
def test_transpose_multiply_against_oracle(nb, oracle, prec, shape):
    K, I, J, bs = shape
    a, b = oracle.synth_dense(K, I, 31), oracle.synth_dense(K, J, 32)
    A = nb.DeviceSet.create("la", "a", K, I, bs, bs); A.upload_dense(a)
    B = nb.DeviceSet.create("la", "b", K, J, bs, bs); B.upload_dense(b)
    want = oracle.la_transpose_multiply_agg(oracle.load_matrix(a, bs, bs), oracle.load_matrix(b, bs, bs), 4).dense()
    C = nb.gemm_blocks(nb.OP_TN, A, B, nb.EPI_NONE, None, nb.PRECISIONS[prec], "la", "c")
    assert C.info() == (I, J, bs, bs)  # out meta: (in1.totalCols, in2.totalCols), LASillyTransposeMultiply1Join.h:48-53
    assert np.abs(C.download_dense() - want).max() <= TOL[prec] * np.abs(want).max()
    G = nb.gemm_blocks(nb.OP_TN, A, A, nb.EPI_NONE, None, nb.PRECISIONS[prec], "la", "gram")  # TestLA07: both inputs one set
    gw = a.T @ a
    assert np.abs(G.download_dense() - gw).max() <= TOL[prec] * np.abs(gw).max()
    with pytest.raises(nb.Nb200Error):  # no bias epilogue on this join
        nb.gemm_blocks(nb.OP_TN, A, B, nb.EPI_BIAS_RELU, None, nb.PRECISIONS[prec], "la", "c")
    [s.destroy() for s in (A, B, C, G)]


def _blocks_from_print(text):
    """Parse MatrixBlock::print() output: 'Block: (r,c), size: (R,C), length:N' followed by R lines of C values."""
    out = {}
    lines = text.splitlines()
    i = 0
    while i < len(lines):
        m = re.search(r"Block: \((\d+),(\d+)\), size: \((\d+),(\d+)\), length:(\d+)", lines[i])
        if m:
            r, c, R, C, _ = map(int, m.groups())
            vals = []
            j = i + 1
            while len(vals) < R * C and j < len(lines):
                vals += [float(x) for x in lines[j].split()] if re.fullmatch(r"[-0-9.e+ ]*", lines[j]) else []
                j += 1
            out[(r, c)] = np.array(vals[:R * C]).reshape(R, C)
            i = j
        else:
            i += 1
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp64", "bf16"])
def test_unmodified_reference_gram_driver(prec):
    """netsDB's own src/tests/source/TestLA07_TransposeMultiply_Gram.cc, compiled unmodified against the shim: loads a
    40 x 20 matrix in 10 x 10 blocks (value = i + j + ii + jj), computes A^T A through LASillyTransposeMultiply1Join +
    LASillyMultiply2Aggregate and prints the product blocks; small integers, so every precision must be exact."""
    from netsdb_b200 import build
    exe = os.path.join(build.HOST_BIN, "TestLA07_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    env = dict(os.environ, NB200_PRECISION=prec)
    r = subprocess.run([exe, "Y", "Y"], env=env, capture_output=True, text=True, timeout=180, cwd=os.path.dirname(exe))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "Transpose multiply output count:4" in r.stdout
    tail = r.stdout[r.stdout.index("Transpose multiply query results"):]
    got = _blocks_from_print(tail)
    a = np.zeros((40, 20))
    for i in range(4):
        for j in range(2):
            for ii in range(10):
                for jj in range(10):
                    a[10 * i + ii, 10 * j + jj] = i + j + ii + jj
    gram = a.T @ a
    assert set(got) == {(0, 0), (0, 1), (1, 0), (1, 1)}
    for (br, bc), blk in got.items():
        assert np.array_equal(blk, gram[10 * br:10 * br + 10, 10 * bc:10 * bc + 10]), (br, bc)


@pytest.mark.gpu
def test_reference_multiply_driver_builds():
    """TestLA04_Multiply.cc (LASillyMultiply1Join) compiles unmodified against the shim too; it is not run: its sizes
    are hard-coded to two 40 GB operands built in host memory."""
    from netsdb_b200 import build
    if not os.path.exists("/root/reference") and not os.path.exists(os.path.join(build.HOST_BIN, "TestLA04_netsdb")):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    assert os.path.exists(os.path.join(build.HOST_BIN, "TestLA04_netsdb"))
