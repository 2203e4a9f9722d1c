"""One LSTM step over FFMatrixBlock sets, wired like the reference's driver src/tests/source/LSTMTest.cc (:204-543):
same set names, same operators, every join + aggregation pair lowered to one device call.  No arithmetic happens here.

Note that the driver combines gates and states with FFInputLayerJoin, i.e. as matrix products (:470-510), and that
is what this mirrors.  Sets whose extents are not multiples of the block size: the reference applies the gate
activation to the zero padding as well (LSTMThreeWaySum.h:76-95), which leaks sigmoid(0) = 0.5 into h_t; the device
keeps padding at zero, so results agree exactly when extents are block multiples.
"""
from __future__ import annotations

from . import capi

GATES = (("f_t", "w_f", "u_f", "b_f", capi.ACT_SIGMOID),       # LSTMTest.cc:204-256
         ("i_t", "w_i", "u_i", "b_i", capi.ACT_SIGMOID),       # :271-321
         ("o_t", "w_o", "u_o", "b_o", capi.ACT_SIGMOID),       # :336-386
         ("c_t_temp", "w_c", "u_c", "b_c", capi.ACT_TANH))     # :401-451


def lstm_cell_fused(db: str = "lstm", precision: str = "bf16x3") -> dict[str, capi.DeviceSet]:
    """The same step through the single entry point nb200_lstm_cell (K-concatenated gate GEMM, fused gate pass)."""
    return capi.lstm_cell(db, capi.PRECISIONS[precision])


def lstm_cell(db: str = "lstm", precision: str = "fp64") -> dict[str, capi.DeviceSet]:
    """Operator by operator, as the reference driver issues its jobs.  Reads sets w_*/u_*/b_* (* in f i o c), x_t, h_t_1, c_t_1 of `db`; writes f_t i_t o_t c_t_temp c_t h_t."""
    L = capi.DeviceSet.lookup
    prec = capi.PRECISIONS[precision]
    mm = lambda a, b, out: capi.gemm_blocks(capi.OP_NN, a, b, capi.EPI_NONE, None, prec, db, out)
    r = {}
    for out, w, u, b, act in GATES:
        wx = mm(L(db, w), L(db, "x_t"), "__wx")                 # FFInputLayerJoin + FFAggMatrix
        uh = mm(L(db, u), L(db, "h_t_1"), "__uh")
        r[out] = capi.sum_blocks(wx, uh, L(db, b), act, db, out)  # LSTMThreeWaySum
    fc = mm(r["f_t"], L(db, "c_t_1"), "__fc")                   # :470-480
    ic = mm(r["i_t"], r["c_t_temp"], "__ic")                    # :482-494
    r["c_t"] = capi.sum_blocks(fc, ic, None, capi.ACT_NONE, db, "c_t")  # LSTMTwoSum :496-498
    tc = capi.sum_blocks(r["c_t"], None, None, capi.ACT_TANH, db, "__tanh_c_t")  # LSTMHiddenState.h:44-47
    r["h_t"] = mm(r["o_t"], tc, "h_t")                          # LSTMHiddenState + FFAggMatrix :528-543
    return r
