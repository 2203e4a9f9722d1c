// la_udfs.h -- the LA DSL's multiply operators on MatrixBlock by their netsDB names (sharedLibraries/headers/
// LASillyMultiply1Join.h, LASillyTransposeMultiply1Join.h, LASillyMultiply2Aggregate.h, LAScanMatrixBlockSet.h,
// LAWriteMatrixBlockSet.h).  Join + aggregate are lowered together to one block GEMM (ff_stage.h); the projection
// lambdas are never run per pair.
#pragma once
#include "MatrixBlock.h"
#include "ff_udfs.h"

#define NB200_LA_JOIN_UDF(NAME, SEL1, SEL2)                                                                  \
    pdb::Lambda<bool> getSelection(pdb::Handle<MatrixBlock> in1, pdb::Handle<MatrixBlock> in2) override {    \
        return makeLambdaFromMethod(in1, SEL1) == makeLambdaFromMethod(in2, SEL2);                           \
    }                                                                                                        \
    pdb::Lambda<pdb::Handle<MatrixBlock>> getProjection(pdb::Handle<MatrixBlock> in1,                        \
                                                        pdb::Handle<MatrixBlock> in2) override {             \
        return pdb::makeLambda(in1, in2, [](pdb::Handle<MatrixBlock> &, pdb::Handle<MatrixBlock> &) {        \
            nb200_host::no_host_projection(#NAME);                                                           \
            return pdb::Handle<MatrixBlock>();                                                               \
        });                                                                                                  \
    }

// predicate in1.blockCol == in2.blockRow, out = in1 . in2 (LASillyMultiply1Join.h:24-70)
class LASillyMultiply1Join : public pdb::JoinComp<MatrixBlock, MatrixBlock, MatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LASillyMultiply1Join() {}
    NB200_LA_JOIN_UDF(LASillyMultiply1Join, getBlockColIndex, getBlockRowIndex)
};
// predicate in1.blockRow == in2.blockRow, out = in1^T . in2 (LASillyTransposeMultiply1Join.h:24-70)
class LASillyTransposeMultiply1Join : public pdb::JoinComp<MatrixBlock, MatrixBlock, MatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LASillyTransposeMultiply1Join() {}
    NB200_LA_JOIN_UDF(LASillyTransposeMultiply1Join, getBlockRowIndex, getBlockRowIndex)
};
// key = block meta, value = block data flagged for summation (LASillyMultiply2Aggregate.h)
class LASillyMultiply2Aggregate : public pdb::ClusterAggregateComp<MatrixBlock, MatrixBlock, MatrixMeta, MatrixData> {
public:
    ENABLE_DEEP_COPY
    LASillyMultiply2Aggregate() {}
    pdb::Lambda<MatrixMeta> getKeyProjection(pdb::Handle<MatrixBlock> aggMe) override { return makeLambdaFromMethod(aggMe, getMultiplyKey); }
    pdb::Lambda<MatrixData> getValueProjection(pdb::Handle<MatrixBlock> aggMe) override { return makeLambdaFromMethod(aggMe, getMultiplyValue); }
};
class LAScanMatrixBlockSet : public pdb::ScanUserSet<MatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LAScanMatrixBlockSet() {}
    LAScanMatrixBlockSet(std::string dbName, std::string setName) { setDatabaseName(dbName); setSetName(setName); }
};
class LAWriteMatrixBlockSet : public pdb::WriteUserSet<MatrixBlock> {
public:
    ENABLE_DEEP_COPY
    LAWriteMatrixBlockSet() {}
    LAWriteMatrixBlockSet(std::string dbName, std::string setName) { this->setOutput(dbName, setName); }
};
