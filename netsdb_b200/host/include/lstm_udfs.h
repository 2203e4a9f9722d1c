// lstm_udfs.h -- the LSTM cell UDFs by their netsDB names, base classes, constructor signatures and selection
// lambdas (src/LSTM/headers/LSTMThreeWaySum.h, LSTMTwoSum.h, LSTMHiddenState.h).  As with ff_udfs.h the projection
// lambdas are never run per tuple: executeComputations lowers each of these joins to a device call (ff_stage.h).
#pragma once
#include "ff_udfs.h"

enum class SumActivation { Sigmod = 1, Tanh };  // LSTMThreeWaySum.h:12 (spelling as in the reference)

// predicate: the three block keys are equal (LSTMThreeWaySum.h:39-42); out = act(in1 + in2 + in3) (:76-95)
class LSTMThreeWaySum : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    SumActivation modifier;
    LSTMThreeWaySum() : modifier(SumActivation::Sigmod) {}
    LSTMThreeWaySum(SumActivation modifierIn) : modifier(modifierIn) {}
    Lambda<bool> getSelection(Handle<FFMatrixBlock> in1, Handle<FFMatrixBlock> in2, Handle<FFMatrixBlock> in3) override {
        return ((makeLambdaFromMethod(in1, getKey) == makeLambdaFromMethod(in2, getKey)) &&
                (makeLambdaFromMethod(in1, getKey) == makeLambdaFromMethod(in3, getKey)));
    }
    Lambda<Handle<FFMatrixBlock>> getProjection(Handle<FFMatrixBlock> in1, Handle<FFMatrixBlock> in2,
                                                Handle<FFMatrixBlock> in3) override {
        return makeLambda(in1, in2, in3, [](Handle<FFMatrixBlock> &, Handle<FFMatrixBlock> &, Handle<FFMatrixBlock> &) {
            nb200_host::no_host_projection("LSTMThreeWaySum");
            return Handle<FFMatrixBlock>();
        });
    }
};
// predicate: equal block keys (LSTMTwoSum.h:24-27); out = in1 + in2 (:58-63)
class LSTMTwoSum : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LSTMTwoSum() {}
    NB200_JOIN_UDF(LSTMTwoSum, getKey, getKey)
};
// predicate in1.blockCol == in2.blockRow (LSTMHiddenState.h:24-27); out = in1 . tanh(in2) (:44-62)
class LSTMHiddenState : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LSTMHiddenState() {}
    NB200_JOIN_UDF(LSTMHiddenState, getBlockColIndex, getBlockRowIndex)
};
