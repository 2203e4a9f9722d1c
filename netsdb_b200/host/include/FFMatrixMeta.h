// FFMatrixMeta -- block coordinates of an FFMatrixBlock (mirrors src/FF/headers/FFMatrixMeta.h:7-53:
// same members, defaults of -1, equality on (row, col) only, hash = 1e6*row + col).
#pragma once
#include "pdb_shim.h"

class FFMatrixMeta : public pdb::Object {
public:
    int blockRowIndex = -1, blockColIndex = -1, totalRows = -1, totalCols = -1, distinctBlockId = -1;
    ENABLE_DEEP_COPY
    FFMatrixMeta() {}
    FFMatrixMeta(int rowIndex, int colIndex) : blockRowIndex(rowIndex), blockColIndex(colIndex) {}
    FFMatrixMeta(int rowIndex, int colIndex, int rows, int cols)
        : blockRowIndex(rowIndex), blockColIndex(colIndex), totalRows(rows), totalCols(cols) {}
    bool operator==(const FFMatrixMeta &o) const { return blockRowIndex == o.blockRowIndex && blockColIndex == o.blockColIndex; }
    size_t hash() const override { return static_cast<size_t>(1000000) * blockRowIndex + blockColIndex; }
};
