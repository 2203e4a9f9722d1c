// FFMatrixUtil.h -- loaders by their netsDB names and signatures (src/FF/headers/FFMatrixUtil.h:14-40)
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace pdb {
class PDBClient;
class CatalogClient;
}  // namespace pdb
namespace ff {
// Same blocking and zero padding as the reference (FFMatrixUtil.cc:245-302).  The reference draws from an
// unseeded generator (:229-234); here each set name maps to a fixed seed (inputs 1, w1 2, b1 3, wo 4, bo 5,
// otherwise a hash of the name) and element (r, c) = synth(seed, r * totalY + c), the same counter-based
// generator as the device's nb200_set_fill_synth, so runs are reproducible and checkable.
void loadMatrix(pdb::PDBClient &pdbClient, std::string dbName, std::string setName, int totalX, int totalY, int blockX,
                int blockY, bool dont_pad_x, bool dont_pad_y, std::string &errMsg, int size = 128, bool partitionByCol = true);
// text format "rows cols" header then values (FFMatrixUtil.cc:11-134)
void load_matrix_data(pdb::PDBClient &pdbClient, std::string path, std::string dbName, std::string setName, int blockX,
                      int blockY, bool dont_pad_x, bool dont_pad_y, std::string &errMsg, int size = 128,
                      bool partitionByCol = true);
void load_matrix_from_file(std::string path, std::vector<std::vector<double>> &matrix);
void print_stats(pdb::PDBClient &pdbClient, std::string dbName, std::string setName);
void print(pdb::PDBClient &pdbClient, std::string dbName, std::string setName);
bool is_empty_set(pdb::PDBClient &pdbClient, pdb::CatalogClient &catalogClient, std::string dbName, std::string setName);
uint64_t seed_of_set(const std::string &setName);
double synth_value(uint64_t seed, uint64_t index);
}  // namespace ff
