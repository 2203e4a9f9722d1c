// ImageUtils.h -- image / kernel loaders by their netsDB names (src/conv2d_memory_fusion/headers/ImageUtils.h:14-24)
#pragma once
#include <string>

#include "pdb_shim.h"

namespace pdb {
class PDBClient;
}
namespace conv2d_memory_fusion {
// text file of img_count * channels * x * y values separated by blanks / commas (ImageUtils.cc:18-136); record i gets
// index i.  M = Image or Kernel.
template <typename M>
void load_imgs_from_file(pdb::PDBClient &pdbClient, std::string path, pdb::String dbName, pdb::String setName, int img_count,
                         int channels, int x, int y);
// Values with the distribution of the reference's generator (ImageUtils.cc:146-170: sign * U(1e-4, 0.5)), seeded from
// the set name so that runs are reproducible; image i gets index i (the reference gives every image index 0, :158,
// which makes its own ImageToChunks stack all images on the same rows).
void loadRandomImages(int width, int height, int channels, int numOfImages, pdb::PDBClient &pdbClient, pdb::String dbName,
                      pdb::String setName, int pageSize);
// extension: the same for any record type and an explicit seed; element e of the (record, channel, x, y) array =
// ff::synth_value(seed, e)
template <typename M>
void load_seeded(pdb::PDBClient &pdbClient, pdb::String dbName, pdb::String setName, int count, int channels, int x, int y,
                 uint64_t seed);
}  // namespace conv2d_memory_fusion
