// MatrixData -- payload of a MatrixBlock (mirrors sharedLibraries/headers/MatrixData.h: rawData / rowNums / colNums /
// flag, the flag setters, and operator+ dispatching on the flag: sum, or element-wise max / min for the row / column
// extremum aggregates; a shape mismatch clears the left operand, an unset flag reports "Operator not supported yet!").
#pragma once
#include "pdb_shim.h"

#define UNSETFLAG 0
#define MAXTRIXSUMFLAG 1
#define MATRIXROWMAXFLAG 2
#define MATRIXROWMINFLAG 3
#define MATRIXCOLMAXFLAG 4
#define MATRIXCOLMINFLAG 5

class MatrixData : public pdb::Object {
public:
    ENABLE_DEEP_COPY
    pdb::Handle<pdb::Vector<double>> rawData;
    int rowNums = 0, colNums = 0, flag = UNSETFLAG;
    MatrixData() {}
    void setSumFlag() { flag = MAXTRIXSUMFLAG; }
    void setRowMaxFlag() { flag = MATRIXROWMAXFLAG; }
    void setRowMinFlag() { flag = MATRIXROWMINFLAG; }
    void setColMaxFlag() { flag = MATRIXCOLMAXFLAG; }
    void setColMinFlag() { flag = MATRIXCOLMINFLAG; }
    void print() {
        std::cout << "Flag: " << flag << " Row: " << rowNums << " Col: " << colNums << " Buffer size:" << rawData->size() << " ";
        if (static_cast<int>(rawData->size()) != rowNums * colNums) { std::cout << "Matrix Error: size misMatch!" << std::endl; return; }
        for (int i = 0; i < static_cast<int>(rawData->size()); i++) {
            if (i % colNums == 0) std::cout << std::endl;
            std::cout << (*rawData)[i] << " ";
        }
        std::cout << std::endl;
    }
    MatrixData &operator+(MatrixData &o) {
        if (rowNums != o.rowNums || colNums != o.colNums) { rawData->clear(); return *this; }
        double *x = rawData->c_ptr(), *y = o.rawData->c_ptr();
        const int n = rowNums * colNums;
        if (flag == MAXTRIXSUMFLAG) for (int i = 0; i < n; i++) x[i] += y[i];
        else if (flag == MATRIXROWMAXFLAG || flag == MATRIXCOLMAXFLAG) { for (int i = 0; i < n; i++) if (x[i] < y[i]) x[i] = y[i]; }
        else if (flag == MATRIXROWMINFLAG || flag == MATRIXCOLMINFLAG) { for (int i = 0; i < n; i++) if (x[i] > y[i]) x[i] = y[i]; }
        else std::cerr << "Operator not supported yet!" << std::endl;
        return *this;
    }
    size_t hash() const override { return 0; }
};
