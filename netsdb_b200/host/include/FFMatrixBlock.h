// FFMatrixBlock -- the tensor-block record (mirrors the public surface of src/FF/headers/FFMatrixBlock.h:18-232:
// constructors, accessors, key/value helpers, hash).  The Eigen row-sum helper is replaced by a plain loop.
#pragma once
#include "FFMatrixData.h"
#include "FFMatrixMeta.h"

#ifndef EIGEN_CODE
#define EIGEN_CODE 0
#endif

class FFMatrixBlock : public pdb::Object {
public:
    FFMatrixData data;
    pdb::Handle<FFMatrixMeta> meta;
    bool partitionByCol = true;
    const static int librayCode = EIGEN_CODE;
    ENABLE_DEEP_COPY

    FFMatrixBlock() {}
    FFMatrixBlock(int br, int bc, int rows, int cols, bool byCol = true) : data(rows, cols), partitionByCol(byCol) {
        meta = pdb::makeObject<FFMatrixMeta>(br, bc);
    }
    FFMatrixBlock(int br, int bc, int rows, int cols, int totalRows, int totalCols, bool byCol = true)
        : data(rows, cols), partitionByCol(byCol) {
        meta = pdb::makeObject<FFMatrixMeta>(br, bc, totalRows, totalCols);
    }
    FFMatrixBlock(int br, int bc, int rows, int cols, int totalRows, int totalCols, int distinctBlockId, bool byCol = true)
        : data(rows, cols), partitionByCol(byCol) {
        meta = pdb::makeObject<FFMatrixMeta>(br, bc, totalRows, totalCols);
        meta->distinctBlockId = distinctBlockId;
    }
    FFMatrixBlock(int br, int bc, int rows, int cols, pdb::Handle<pdb::Vector<double>> raw, bool byCol = true)
        : data(rows, cols, raw), partitionByCol(byCol) {
        meta = pdb::makeObject<FFMatrixMeta>(br, bc);
    }
    FFMatrixBlock(int br, int bc, int rows, int cols, int totalRows, int totalCols, pdb::Handle<pdb::Vector<double>> raw,
                  bool byCol = true)
        : data(rows, cols, raw), partitionByCol(byCol) {
        meta = pdb::makeObject<FFMatrixMeta>(br, bc, totalRows, totalCols);
    }

    int getRowNums() { return data.rowNums; }
    int getColNums() { return data.colNums; }
    int rowIndexStart() { return getBlockRowIndex() * getRowNums(); }
    int rowIndexEnd() { return rowIndexStart() + getRowNums(); }
    int getBlockRowIndex() { return meta->blockRowIndex; }
    int getBlockColIndex() { return meta->blockColIndex; }
    int getTotalRowNums() { return meta->totalRows; }
    int getTotalColNums() { return meta->totalCols; }
    int getNumRowBlocks() { return static_cast<int>(std::ceil(getTotalRowNums() / static_cast<double>(getRowNums()))); }
    int getNumColBlocks() { return static_cast<int>(std::ceil(getTotalColNums() / static_cast<double>(getColNums()))); }
    bool isLastColBlock() { return getBlockColIndex() == getNumColBlocks() - 1; }
    bool isLastRowBlock() { return getBlockRowIndex() == getNumRowBlocks() - 1; }
    pdb::Handle<pdb::Vector<double>> &getRawDataHandle() { return data.rawData; }
    pdb::Handle<FFMatrixMeta> &getKey() { return meta; }
    FFMatrixMeta &getKeyRef() { return *meta; }
    FFMatrixData &getValue() { return data; }

    // row-wise helpers used by FFRowAggregate (FFMatrixBlock.h:116-142)
    pdb::Handle<FFMatrixMeta> getRowKey() { return pdb::makeObject<FFMatrixMeta>(meta->blockRowIndex, 0, meta->totalRows, 1); }
    FFMatrixData getRowSumValue() {
        FFMatrixData sums(data.rowNums, 1);
        double *out = sums.rawData->c_ptr(), *in = data.rawData->c_ptr();
        for (int i = 0; i < data.rowNums; i++) {
            double s = 0.0;
            for (int j = 0; j < data.colNums; j++) s += in[static_cast<long>(i) * data.colNums + j];
            out[i] = s;
        }
        return sums;
    }
    void print() {
        std::cout << "BlockRowIndex:" << getBlockRowIndex() << std::endl << "BlockColumnIndex:" << getBlockColIndex() << std::endl
                  << "TotalRowNums:" << getTotalRowNums() << std::endl << "TotalColNums:" << getTotalColNums() << std::endl;
    }
    size_t hash() const override {
        return std::hash<int>()(partitionByCol ? meta->blockColIndex : meta->blockRowIndex);
    }
};
