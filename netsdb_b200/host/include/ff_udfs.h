// ff_udfs.h -- the FF UDF classes by their netsDB names, base classes, template arguments, constructor
// signatures and selection lambdas (so that a planner can still extract the join keys).  Their projection
// lambdas exist for API fidelity only: on this path a JoinComp + ClusterAggregateComp pair is never run per
// joined tuple -- executeComputations lowers it to ONE block GEMM on the device (ff_stage.h).  Invoking a
// projection on the host is therefore an error (there is no CPU fallback).
//
// Reference: src/FF/headers/FFTransposeMult.h, FFInputLayerJoin.h, FFAggMatrix.h, FFReluBiasSum.h,
// FFTransposeBiasSum.h, FFRowAggregate.h, FFOutputLayer.h, FFMatrixBlockScanner.h, FFMatrixWriter.h.
#pragma once
#include "FFMatrixBlock.h"

namespace nb200_host {
[[noreturn]] inline void no_host_projection(const char *udf) {
    std::cerr << udf << ": the per-pair projection is not executed on the host; this computation must be lowered to the "
              << "device by executeComputations (no CPU fallback)" << std::endl;
    exit(1);
}
}  // namespace nb200_host

#define NB200_JOIN_UDF(NAME, SEL1, SEL2)                                                                         \
    pdb::Lambda<bool> getSelection(pdb::Handle<FFMatrixBlock> in1, pdb::Handle<FFMatrixBlock> in2) override {    \
        return makeLambdaFromMethod(in1, SEL1) == makeLambdaFromMethod(in2, SEL2);                               \
    }                                                                                                            \
    pdb::Lambda<pdb::Handle<FFMatrixBlock>> getProjection(pdb::Handle<FFMatrixBlock> in1,                        \
                                                          pdb::Handle<FFMatrixBlock> in2) override {             \
        return pdb::makeLambda(in1, in2, [](pdb::Handle<FFMatrixBlock> &, pdb::Handle<FFMatrixBlock> &) {        \
            nb200_host::no_host_projection(#NAME);                                                               \
            return pdb::Handle<FFMatrixBlock>();                                                                 \
        });                                                                                                      \
    }

using namespace pdb;

// predicate in1.blockCol == in2.blockCol, out(I x J) = in1 . in2^T (FFTransposeMult.h:38-39, 80-92)
class FFTransposeMult : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFTransposeMult() = default;
    NB200_JOIN_UDF(FFTransposeMult, getBlockColIndex, getBlockColIndex)
};
// predicate in1.blockCol == in2.blockRow, out = in1 . in2 (FFInputLayerJoin.h:30-31, 72-86)
class FFInputLayerJoin : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFInputLayerJoin() {}
    NB200_JOIN_UDF(FFInputLayerJoin, getBlockColIndex, getBlockRowIndex)
};
// out = max(0, in1 + in2[row]) with optional dropout (FFReluBiasSum.h:20-22, 63-80)
class FFReluBiasSum : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFReluBiasSum() : dropout_rate(0) {}
    FFReluBiasSum(double dropout_rate_in) : dropout_rate(dropout_rate_in) {}
    double dropout_rate;
    NB200_JOIN_UDF(FFReluBiasSum, getBlockRowIndex, getBlockRowIndex)
};
// out = exp((in1 + in2[row])^T) (FFTransposeBiasSum.h:64-104, intended semantics)
class FFTransposeBiasSum : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFTransposeBiasSum() = default;
    NB200_JOIN_UDF(FFTransposeBiasSum, getBlockRowIndex, getBlockRowIndex)
};
// out = in1[i,j] / in2[i] (FFOutputLayer.h:21-22, 31-57, intended semantics)
class FFOutputLayer : public JoinComp<FFMatrixBlock, FFMatrixBlock, FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFOutputLayer() {}
    NB200_JOIN_UDF(FFOutputLayer, getBlockRowIndex, getBlockRowIndex)
};
// key = block meta, value = block data, summed by FFMatrixData::operator+ (FFAggMatrix.h:20-33)
class FFAggMatrix : public ClusterAggregateComp<FFMatrixBlock, FFMatrixBlock, Handle<FFMatrixMeta>, FFMatrixData> {
public:
    ENABLE_DEEP_COPY
    FFAggMatrix() {}
    Lambda<Handle<FFMatrixMeta>> getKeyProjection(Handle<FFMatrixBlock> aggMe) override { return makeLambdaFromMethod(aggMe, getKey); }
    Lambda<FFMatrixData> getValueProjection(Handle<FFMatrixBlock> aggMe) override {
        return makeLambda(aggMe, [](Handle<FFMatrixBlock> &b) { return b->getValue(); });
    }
};
// key = (blockRow, 0), value = row sums (FFRowAggregate.h:21-28)
class FFRowAggregate : public ClusterAggregateComp<FFMatrixBlock, FFMatrixBlock, Handle<FFMatrixMeta>, FFMatrixData> {
public:
    ENABLE_DEEP_COPY
    FFRowAggregate() {}
    Lambda<Handle<FFMatrixMeta>> getKeyProjection(Handle<FFMatrixBlock> aggMe) override { return makeLambdaFromMethod(aggMe, getRowKey); }
    Lambda<FFMatrixData> getValueProjection(Handle<FFMatrixBlock> aggMe) override { return makeLambdaFromMethod(aggMe, getRowSumValue); }
};
class FFMatrixBlockScanner : public ScanUserSet<FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFMatrixBlockScanner() {}
    FFMatrixBlockScanner(std::string dbName, std::string setName) { setDatabaseName(dbName); setSetName(setName); }
};
class FFMatrixWriter : public WriteUserSet<FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    FFMatrixWriter() {}
    FFMatrixWriter(std::string dbName, std::string setName) { this->setOutput(dbName, setName); }
};
