// Compile check of netsdb_b200/host/reference_side/FFGpuStage.h against the reference's UNMODIFIED
// src/FF/headers/FFMatrixBlock.h (found through -I$(REF)/FF/headers) and the stand-in object model of oracle/ref/pdbmin.
// Test infrastructure; nothing here is linked into the product.
#include "FFGpuStage.h"

// every entry point of the stub, instantiated
void ffgpustage_uses_the_reference_api(nb200_set *set) {
  pdb::Handle<FFMatrixBlock> blk = pdb::makeObject<FFMatrixBlock>(0, 0, 4, 4, 8, 8, false);
  ff_gpu::ship(set, blk);
  pdb::Handle<FFMatrixBlock> back = ff_gpu::fetch(set, 0, 0);
  (void)back;
  nb200_set *s = ff_gpu::create("ff", "inputs", 8, 8, 4, 4);
  (void)s;
  nb200_set *out = ff_gpu::inference_unit("ff", "w1", "wo", "inputs", "b1", "bo", "output", NB200_PREC_F16);
  (void)out;
}
