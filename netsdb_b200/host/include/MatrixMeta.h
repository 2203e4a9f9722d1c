// MatrixMeta -- block coordinates of the LA DSL's MatrixBlock (mirrors sharedLibraries/headers/MatrixMeta.h: public
// members, totals default 0, equality on (row, col), hash = 10000 * row + col).
#pragma once
#include "pdb_shim.h"

class MatrixMeta : public pdb::Object {
public:
    ENABLE_DEEP_COPY
    int blockRowIndex = 0, blockColIndex = 0, totalRows = 0, totalCols = 0;
    MatrixMeta() {}
    bool operator==(const MatrixMeta &o) const { return blockRowIndex == o.blockRowIndex && blockColIndex == o.blockColIndex; }
    size_t hash() const override { return static_cast<size_t>(10000) * blockRowIndex + blockColIndex; }
};
