// MatrixBlock -- the LA DSL's tensor block (mirrors the part of sharedLibraries/headers/MatrixBlock.h the multiply
// drivers touch: the four constructors, accessors, key / value helpers for LASillyMultiply2Aggregate, print).  Same
// payload layout as FFMatrixBlock (row-major rowNums x colNums doubles), meta and data held by value; totals stay 0
// with the 4-argument constructor, as in the reference.
#pragma once
#include "MatrixData.h"
#include "MatrixMeta.h"

#ifndef EIGEN_CODE
#define EIGEN_CODE 0
#endif

class MatrixBlock : public pdb::Object {
private:
    MatrixMeta meta;
    MatrixData data;

public:
    const static int librayCode = EIGEN_CODE;
    ENABLE_DEEP_COPY
    MatrixBlock() {}
    MatrixBlock(int br, int bc, int rows, int cols) { init(br, bc, rows, cols, 0, 0, pdb::makeObject<pdb::Vector<double>>(rows * cols, rows * cols)); }
    MatrixBlock(int br, int bc, int rows, int cols, int totalRows, int totalCols) {
        init(br, bc, rows, cols, totalRows, totalCols, pdb::makeObject<pdb::Vector<double>>(rows * cols, rows * cols));
    }
    MatrixBlock(int br, int bc, int rows, int cols, pdb::Handle<pdb::Vector<double>> raw) { init(br, bc, rows, cols, 0, 0, raw); }
    MatrixBlock(int br, int bc, int rows, int cols, int totalRows, int totalCols, pdb::Handle<pdb::Vector<double>> raw) {
        init(br, bc, rows, cols, totalRows, totalCols, raw);
    }
    size_t hash() const override { return static_cast<size_t>(10000) * meta.blockRowIndex + meta.blockColIndex; }
    std::string toString() {
        return "Block: (" + std::to_string(meta.blockRowIndex) + "," + std::to_string(meta.blockColIndex) + ")";
    }
    void printMeta() {
        std::cout << "Block: (" << meta.blockRowIndex << "," << meta.blockColIndex << "), size: (" << data.rowNums << "," << data.colNums
                  << "), length:" << data.rawData->size() << " ";
    }
    void print() {
        printMeta();
        if (static_cast<int>(data.rawData->size()) != data.rowNums * data.colNums) { std::cout << "Matrix Error: size misMatch!" << std::endl; return; }
        for (int i = 0; i < static_cast<int>(data.rawData->size()); i++) {
            if (i % data.colNums == 0) std::cout << std::endl;
            std::cout << (*data.rawData)[i] << " ";
        }
        std::cout << std::endl;
    }
    int getBlockRowIndex() { return meta.blockRowIndex; }
    int getBlockColIndex() { return meta.blockColIndex; }
    int getRowNums() { return data.rowNums; }
    int getColNums() { return data.colNums; }
    int getTotalRowNums() { return meta.totalRows; }
    int getTotalColNums() { return meta.totalCols; }
    pdb::Handle<pdb::Vector<double>> &getRawDataHandle() { return data.rawData; }
    MatrixMeta &getKey() { return meta; }
    MatrixData &getValue() { return data; }
    MatrixMeta &getMultiplyKey() { return meta; }
    MatrixData &getMultiplyValue() { data.setSumFlag(); return data; }

private:
    void init(int br, int bc, int rows, int cols, int tr, int tc, pdb::Handle<pdb::Vector<double>> raw) {
        meta.blockRowIndex = br; meta.blockColIndex = bc; meta.totalRows = tr; meta.totalCols = tc;
        data.rowNums = rows; data.colNums = cols; data.rawData = raw;
    }
};
