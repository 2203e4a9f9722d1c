// FFMatrixData -- payload of an FFMatrixBlock (mirrors src/FF/headers/FFMatrixData.h:11-58): row-major
// rowNums x colNums doubles in rawData, optional bias vector, operator+ = element-wise accumulate.
#pragma once
#include "pdb_shim.h"

class FFMatrixData : public pdb::Object {
public:
    int rowNums = 0, colNums = 0;
    pdb::Handle<pdb::Vector<double>> rawData;
    pdb::Handle<pdb::Vector<double>> bias;
    ENABLE_DEEP_COPY
    FFMatrixData() {}
    FFMatrixData(int rowNumsIn, int colNumsIn) : rowNums(rowNumsIn), colNums(colNumsIn) { allocate(); }
    // the reference ignores rawDataIn and allocates fresh zeros (FFMatrixData.h:26-31)
    FFMatrixData(int rowNumsIn, int colNumsIn, pdb::Handle<pdb::Vector<double>>) : rowNums(rowNumsIn), colNums(colNumsIn) { allocate(); }
    FFMatrixData &operator+(FFMatrixData &other) {
        double *mine = rawData->c_ptr(), *theirs = other.rawData->c_ptr();
        const long n = static_cast<long>(rowNums) * colNums;
        for (long i = 0; i < n; i++) mine[i] += theirs[i];
        if (bias != nullptr && other.bias != nullptr)
            for (size_t i = 0; i < other.bias->size(); i++) (*bias)[i] += (*other.bias)[i];
        return *this;
    }

private:
    void allocate() {
        const uint32_t n = static_cast<uint32_t>(rowNums) * static_cast<uint32_t>(colNums);
        rawData = pdb::makeObject<pdb::Vector<double>>(n, n);
    }
};
