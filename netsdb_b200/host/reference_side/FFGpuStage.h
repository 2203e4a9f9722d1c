// FFGpuStage.h -- the stub a netsDB maintainer drops into src/FF/headers/ (INTEGRATION.md section 2).
//
// It is written against netsDB's OWN headers (FFMatrixBlock.h, Handle.h, PDBVector.h) and the C ABI of
// include/netsdb_b200.h, nothing else.  `make -C oracle ref` compile-checks it against the reference's unmodified
// src/FF/headers/FFMatrixBlock.h (oracle/ref/ffgpustage_check.cc), so the member functions and the constructor it uses
// are the ones the reference really has (FFMatrixBlock.h:37-109).
//
// What it covers: the two places where FFMatrixBlocks cross between Pangea pages and the device, and the call that
// replaces ff::inference_unit's two executeComputations graphs (src/FF/source/SimpleFF.cc:331-424).  What it does not:
// the planner change that makes PipelineStage pick the fused stage by itself.
#ifndef FF_GPU_STAGE_H
#define FF_GPU_STAGE_H

#include <cstdlib>
#include <iostream>
#include <string>

#include "FFMatrixBlock.h"
#include "netsdb_b200.h"

namespace ff_gpu {

// the reference's error convention for UDFs: print and exit(1) (FFTransposeMult.h:61-65)
inline void check(int st) {
  if (st != NB200_OK) {
    std::cerr << nb200_last_error() << std::endl;
    exit(1);
  }
}

// PDBClient::createSet<FFMatrixBlock>(db, set, ...) on the device; totals as in FFMatrixMeta, block_* the nominal block
inline nb200_set *create(const std::string &db, const std::string &set, int totalRows, int totalCols, int blockRows,
                         int blockCols) {
  nb200_set *s = nullptr;
  check(nb200_set_create(db.c_str(), set.c_str(), totalRows, totalCols, blockRows, blockCols, &s));
  return s;
}

// called once per FFMatrixBlock as pages are scanned (ScanUserSet) or received (sendData)
inline void ship(nb200_set *set, pdb::Handle<FFMatrixBlock> &b) {
  check(nb200_upload_block(set, b->getBlockRowIndex(), b->getBlockColIndex(), b->getValue().rawData->c_ptr(),
                           b->getRowNums(), b->getColNums()));
}

// re-wrap one result block for WriteUserSet / getSetIterator.  The FFMatrixBlock is allocated FIRST: NotEnoughSpace must
// fire before any device work, so that Pipeline::run can rotate pages and re-invoke the lambda (Pipeline.h:253-280)
inline pdb::Handle<FFMatrixBlock> fetch(nb200_set *set, int blockRow, int blockCol) {
  int totalRows = 0, totalCols = 0, rows = 0, cols = 0;
  check(nb200_set_info(set, &totalRows, &totalCols, &rows, &cols));
  pdb::Handle<FFMatrixBlock> out = pdb::makeObject<FFMatrixBlock>(blockRow, blockCol, rows, cols, totalRows, totalCols, false);
  check(nb200_download_block(set, blockRow, blockCol, out->getValue().rawData->c_ptr(), rows, cols));
  return out;
}

// ff::inference_unit(pdbClient, db, w1, wo, inputs, b1, bo, output, dropout) on sets that were shipped before: creates
// `yo` and `output` on the device; dropout must be 0 (the reference's is seeded with time(NULL), FFReluBiasSum.h:67)
inline nb200_set *inference_unit(const std::string &db, const std::string &w1, const std::string &wo, const std::string &inputs,
                                 const std::string &b1, const std::string &bo, const std::string &output, int precision) {
  nb200_set *sw1 = nullptr, *swo = nullptr, *sx = nullptr, *sb1 = nullptr, *sbo = nullptr, *out = nullptr;
  check(nb200_set_lookup(db.c_str(), w1.c_str(), &sw1));
  check(nb200_set_lookup(db.c_str(), wo.c_str(), &swo));
  check(nb200_set_lookup(db.c_str(), inputs.c_str(), &sx));
  check(nb200_set_lookup(db.c_str(), b1.c_str(), &sb1));
  check(nb200_set_lookup(db.c_str(), bo.c_str(), &sbo));
  check(nb200_ff_inference_unit(sw1, swo, sx, sb1, sbo, precision, db.c_str(), output.c_str(), &out, nullptr, nullptr));
  return out;
}

}  // namespace ff_gpu
#endif
