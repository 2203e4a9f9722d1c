"""LSTM cell operators (LSTMThreeWaySum / LSTMTwoSum / LSTMHiddenState over FFMatrixBlock sets).

CPU part: the oracle's blocked restatement against hand-computed values and an independent dense numpy evaluation,
including the reference's padding leak for ragged extents (quirk Q9).  GPU part (-m gpu): the device operators
through the C ABI against the oracle.
"""
import numpy as np
import pytest

NAMES = ["w_f", "u_f", "b_f", "w_i", "u_i", "b_i", "w_o", "u_o", "b_o", "w_c", "u_c", "b_c", "x_t", "h_t_1", "c_t_1"]
OUT = ["f_t", "i_t", "o_t", "c_t_temp", "c_t", "h_t"]


def make_inputs(L, D, B, seed):
    rng = np.random.default_rng(seed)
    d = {}
    for g in "fioc":
        d["w_" + g] = rng.uniform(-0.1, 0.1, (L, D))
        d["u_" + g] = rng.uniform(-0.1, 0.1, (L, L))
        d["b_" + g] = rng.uniform(-0.5, 0.5, (L, B))
    d["x_t"] = rng.uniform(-1, 1, (D, B))
    d["h_t_1"] = rng.uniform(-1, 1, (L, B))
    d["c_t_1"] = rng.uniform(-1, 1, (L, B))
    return d


def test_three_way_sum_kat(oracle):
    """2 x 3 blocks by hand: sigmoid(0) = 0.5, tanh(ln 2) = 0.6, padding goes through the activation too."""
    a = np.array([[0.0, 1.0, -1.0], [2.0, 0.5, 0.25]])
    b = np.array([[0.0, -1.0, 1.0], [-2.0, 0.5, 0.25]])
    c = np.array([[0.0, np.log(2.0), 0.0], [0.0, 0.0, -0.5]])
    A, B, C = (oracle.load_matrix(m, 2, 2) for m in (a, b, c))
    sig = oracle.lstm_sum(A, B, C, oracle.SIGMOID)
    want = 1 / (1 + np.exp(-(a + b + c)))
    assert np.allclose(sig.dense(), want, rtol=0, atol=1e-16)
    assert sig.dense()[0, 0] == 0.5 and abs(sig.dense()[0, 1] - 2 / 3) < 1e-16
    th = oracle.lstm_sum(A, B, C, oracle.TANH).dense()
    assert abs(th[0, 1] - 0.6) < 1e-15 and th[0, 0] == 0.0
    # block (0, 1) is 2 x 2 with one padded column: the reference's loop runs over the padded block
    blocks = {(br, bc): arr.copy() for br, bc, _, _, arr in sig.blocks()}
    assert blocks[(0, 1)].shape == (2, 2) and np.all(blocks[(0, 1)][:, 1] == 0.5)
    two = oracle.lstm_sum(A, B).dense()
    assert np.array_equal(two, a + b)


def test_three_way_sum_dimension_mismatch(oracle):
    a = oracle.load_matrix(np.ones((4, 4)), 2, 2)
    b = oracle.load_matrix(np.ones((4, 4)), 4, 4)
    with pytest.raises(RuntimeError):
        oracle.lstm_sum(a, b, a, oracle.SIGMOID)


@pytest.mark.parametrize("shape", [(128, 256, 128, 32), (96, 200, 96, 32), (64, 64, 64, 64)])
def test_cell_oracle_vs_numpy(oracle, shape):
    L, D, B, bs = shape
    d = make_inputs(L, D, B, 5)
    r = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    want = oracle.lstm_cell_numpy(d)
    for k in OUT:
        assert np.abs(r[k].dense() - want[k]).max() < 1e-11, k


def test_cell_reference_driver_shape_leaks_padding(oracle):
    """LSTMTest.cc's own sizes (L 128, D 400, B 500, blocks of 100): gates and c_t follow the zero-padding model;
    h_t picks up the sigmoid(0) = 0.5 the reference writes into the padding of f_t / i_t / o_t (Q9)."""
    L, D, B, bs = 128, 400, 500, 100
    d = make_inputs(L, D, B, 6)
    r = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    as_written, zero_pad = oracle.lstm_cell_numpy(d, bs), oracle.lstm_cell_numpy(d)
    for k in OUT:
        assert np.abs(r[k].dense() - as_written[k]).max() < 1e-11, k
    for k in OUT[:-1]:
        assert np.abs(r[k].dense() - zero_pad[k]).max() < 1e-11, k
    assert np.abs(r["h_t"].dense() - zero_pad["h_t"]).max() > 1.0


# ---------------------------------------------------------------------------------------------- GPU
TOL = {"fp64": 1e-12, "bf16x3": 1e-4, "f16": 2.5e-3, "bf16": 2e-2}


def upload(nb, db, d, bs):
    sets = {}
    for k, v in d.items():
        s = nb.DeviceSet.create(db, k, v.shape[0], v.shape[1], bs, bs)
        s.upload_dense(np.ascontiguousarray(v))
        sets[k] = s
    return sets


@pytest.mark.gpu
@pytest.mark.parametrize("act", [0, 1, 2])
@pytest.mark.parametrize("n_in", [1, 2, 3])
def test_sum_blocks_against_oracle(nb, oracle, act, n_in):
    rng = np.random.default_rng(act * 3 + n_in)
    mats = [rng.uniform(-3, 3, (37, 53)) for _ in range(n_in)]
    mats[0][0, 0] = 800.0 if act == 2 else mats[0][0, 0]  # beyond the range where the reference's tanh quotient is finite
    sets = [nb.DeviceSet.create("t", f"s{i}", 37, 53, 10, 12) for i in range(n_in)]
    for s, m in zip(sets, mats):
        s.upload_dense(m)
    o = nb.sum_blocks(sets[0], sets[1] if n_in > 1 else None, sets[2] if n_in > 2 else None, act, "t", "sum_out")
    want = oracle.lstm_sum(*[oracle.load_matrix(m, 10, 12) for m in mats], *([None] * (3 - n_in)), activation=act).dense()
    got = o.download_dense()
    if act == 2 and n_in == 1:
        assert got[0, 0] == 1.0 and np.isnan(want[0, 0])  # reference: inf / inf
        got[0, 0] = want[0, 0] = 0.0
    else:
        finite = np.isfinite(want)
        got, want = np.where(finite, got, 0.0), np.where(finite, want, 0.0)
    assert np.abs(got - want).max() <= 4e-16 * max(1.0, np.abs(want).max())
    assert o.info() == (37, 53, 10, 12)
    o.destroy()
    [s.destroy() for s in sets]


@pytest.mark.gpu
def test_sum_blocks_errors(nb):
    a = nb.DeviceSet.create("t", "ea", 8, 8, 4, 4); a.upload_dense(np.ones((8, 8)))
    b = nb.DeviceSet.create("t", "eb", 8, 8, 8, 8); b.upload_dense(np.ones((8, 8)))
    with pytest.raises(nb.Nb200Error) as e:
        nb.sum_blocks(a, b, None, 0, "t", "eo")
    assert e.value.status == 3 and "Block dimemsions mismatch" in str(e.value)
    with pytest.raises(nb.Nb200Error):
        nb.sum_blocks(a, a, a, 7, "t", "eo")
    # the result may replace an input
    o = nb.sum_blocks(a, a, None, 0, "t", "ea")
    assert np.array_equal(o.download_dense(), 2 * np.ones((8, 8)))
    o.destroy(), b.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp64", "bf16x3", "f16", "bf16"])
@pytest.mark.parametrize("shape", [(128, 256, 128, 32), (96, 200, 96, 32), (256, 400, 256, 128)])
def test_lstm_cell_against_oracle(nb, oracle, prec, shape):
    from netsdb_b200 import lstm
    L, D, B, bs = shape
    d = make_inputs(L, D, B, 7)
    want = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    sets = upload(nb, "lstm", d, bs)
    r = lstm.lstm_cell("lstm", prec)
    for k in OUT:
        w, g = want[k].dense(), r[k].download_dense()
        s = max(np.abs(w).max(), 1.0)
        assert np.abs(g - w).max() <= TOL[prec] * s * (4 if k in ("c_t", "h_t") and prec != "fp64" else 1), (k, prec)
        assert r[k].info() == (L, B, bs, bs)
    [s.destroy() for s in sets.values()]


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["bf16x3", "bf16", "fp64", "f16"])
@pytest.mark.parametrize("shape", [(128, 256, 128, 32), (96, 200, 96, 32), (128, 192, 64, 32), (64, 100, 192, 64), (320, 700, 320, 64)])
def test_lstm_cell_single_entry_point(nb, oracle, prec, shape):
    """nb200_lstm_cell: K-concatenated gate GEMM + fused gate pass (bf16 / bf16x3), composed operators otherwise;
    L != B exercises the joins over differing K extents (contraction over min(B, L))."""
    from netsdb_b200 import lstm
    L, D, B, bs = shape
    d = make_inputs(L, D, B, 11)
    want = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    sets = upload(nb, "lstm", d, bs)
    for rep in range(2):  # the second call runs on the cached weight operand and the reused result sets
        r = lstm.lstm_cell_fused("lstm", prec)
        for k in OUT:
            w, g = want[k].dense(), r[k].download_dense()
            s = max(np.abs(w).max(), 1.0)
            assert np.abs(g - w).max() <= TOL[prec] * s * (4 if k in ("c_t", "h_t") and prec != "fp64" else 1), (k, prec, rep)
            assert r[k].info() == (L, B, bs, bs)
    # a changed weight set invalidates the cached stacked operand
    d["u_o"] = d["u_o"] * -1.0
    sets["u_o"].upload_dense(np.ascontiguousarray(d["u_o"]))
    want = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    r = lstm.lstm_cell_fused("lstm", prec)
    for k in ("o_t", "h_t"):
        w, g = want[k].dense(), r[k].download_dense()
        assert np.abs(g - w).max() <= TOL[prec] * max(np.abs(w).max(), 1.0) * 4, (k, prec)
    [s.destroy() for s in sets.values()]


@pytest.mark.gpu
def test_lstm_cell_block_mismatch_is_the_references_error(nb):
    """i_t (column blocks as wide as x_t's) joins c_t_temp (row blocks as tall as w's): unequal block sizes trip the
    per-pair dimension check of FFInputLayerJoin in the reference (exit(1)); same error here, on every precision."""
    from netsdb_b200 import lstm
    L, D, B = 64, 96, 64
    d = make_inputs(L, D, B, 12)
    sets = {}
    for k, v in d.items():
        # weights and biases in 32-row blocks, x_t / h_t_1 / c_t_1 columns in blocks of 16
        br = 32 if k[0] in "wub" else (32 if k != "x_t" else 32)
        bc = 16 if (k[0] == "b" or k in ("x_t", "h_t_1", "c_t_1")) else 32
        if k == "c_t_1":
            br = 16
        s = nb.DeviceSet.create("lstm", k, v.shape[0], v.shape[1], br, bc)
        s.upload_dense(np.ascontiguousarray(v))
        sets[k] = s
    for prec in ("bf16x3", "fp64"):
        with pytest.raises(nb.Nb200Error) as e:
            lstm.lstm_cell_fused("lstm", prec)
        assert "Block dimemsions mismatch" in str(e.value)
    [s.destroy() for s in sets.values()]


@pytest.mark.gpu
def test_lstm_reference_driver_shape(nb, oracle):
    """LSTMTest.cc's sizes: the K extents of f_t (128 x 500) and c_t_1 (128 x 500) differ and the join keeps the
    leading column blocks only.  Gates and c_t equal the reference as written; h_t equals the zero-padding model."""
    from netsdb_b200 import lstm
    L, D, B, bs = 128, 400, 500, 100
    d = make_inputs(L, D, B, 8)
    want = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=4)
    zero_pad = oracle.lstm_cell_numpy(d)
    sets = upload(nb, "lstm", d, bs)
    r = lstm.lstm_cell("lstm", "fp64")
    for k in OUT[:-1]:
        assert np.abs(r[k].download_dense() - want[k].dense()).max() < 1e-11, k
    assert np.abs(r["h_t"].download_dense() - zero_pad["h_t"]).max() < 1e-10
    [s.destroy() for s in sets.values()]


def _run(cmd, env, timeout):
    import os
    import subprocess
    e = dict(os.environ)
    e.update(env)
    return subprocess.run(cmd, env=e, capture_output=True, text=True, timeout=timeout)


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp64", "bf16x3"])
def test_cpp_lstm_driver(oracle, tmp_path, prec):
    """The C++ driver builds the reference's Computation graph (FFInputLayerJoin/FFAggMatrix/LSTMThreeWaySum/
    LSTMTwoSum/LSTMHiddenState objects, text-format loader) and the shim lowers it to the device."""
    import os
    from netsdb_b200 import build
    exe = os.path.join(build.HOST_BIN, "LSTMTest")
    assert os.path.exists(exe), "host drivers not built (python -m netsdb_b200.build)"
    L, D, B, bs = 64, 100, 64, 32
    d = make_inputs(L, D, B, 9)
    for k, v in d.items():
        with open(tmp_path / f"{k}.txt", "w") as f:
            f.write(f"{v.shape[0]}, {v.shape[1]}\n")
            np.savetxt(f, v, fmt="%.17g")
    r = _run([exe, str(bs), str(tmp_path)], {"NB200_PRECISION": prec}, timeout=120)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    want = oracle.lstm_cell({k: oracle.load_matrix(v, bs, bs) for k, v in d.items()}, threads=2)
    for k in OUT:
        got = np.loadtxt(tmp_path / f"{k}.out.txt", skiprows=1).reshape(L, B)
        w = want[k].dense()
        tol = TOL[prec] * max(np.abs(w).max(), 1.0) * (4 if k in ("c_t", "h_t") and prec != "fp64" else 1)
        assert np.abs(got - w).max() <= tol, (k, prec)


@pytest.mark.gpu
def test_unmodified_reference_lstm_driver():
    """netsDB's own src/tests/source/LSTMTest.cc compiled unmodified against the shim (only where the reference tree
    was present at build time).  It loads all-zero weights and checks nothing; exit status and the shapes it prints
    (print_stats) are what can be asserted: 128 x 500 results in 100 x 100 blocks = 2 x 5 blocks, which the driver's
    print_stats reports with rows and columns swapped (it counts blocks of block row 0 as "rows", LSTMTest.cc:44-51)."""
    import os
    from netsdb_b200 import build
    exe = os.path.join(build.HOST_BIN, "LSTMTest_netsdb")
    if not os.path.exists(exe):
        pytest.skip("reference driver was not built (no reference tree at build time)")
    r = _run([exe], {"NB200_PRECISION": "fp64"}, timeout=180)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    for s in ("f_t", "i_t", "o_t", "c_t_temp", "c_t", "h_t"):
        assert f"{s} (5 X 2) (10) : (500 x 200), Each block size: 100 x 100" in r.stdout, (s, r.stdout[-2500:])
