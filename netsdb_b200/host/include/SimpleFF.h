// SimpleFF.h -- ff:: driver functions by their netsDB names and signatures (src/FF/headers/SimpleFF.h:14-52)
#pragma once
#include <string>

namespace pdb {
class PDBClient;
}
namespace ff {
void loadLibrary(pdb::PDBClient &pdbClient, std::string path);
void createSet(pdb::PDBClient &pdbClient, std::string dbName, std::string setName, std::string setName1, int size);
void createDatabase(pdb::PDBClient &pdbClient, std::string dbName);
void setup(pdb::PDBClient &pdbClient, std::string database);
// 3-layer variant (src/FF/source/SimpleFF.cc:94-329): y1 = relu(w1 . inputs^T + b1), y2 = relu(w2 . y1 + b2),
// yo = exp((wo . y2 + bo)^T), output = yo / rowsum(yo); writes sets y1, y2, yo and `output` like the reference.
void inference_compute(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string w2, std::string wo,
                       std::string inputs, std::string b1, std::string b2, std::string bo, double dropout_rate,
                       bool enablePartition = false);
void inference(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string w2, std::string wo,
               std::string inputs, std::string b1, std::string b2, std::string bo, std::string output, double dropout_rate,
               bool enablePartition = false);
void inference_unit(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string wo, std::string inputs,
                    std::string b1, std::string bo, std::string output, double dropout_rate, bool enablePartition = false);
void cleanup(pdb::PDBClient &pdbClient, std::string database);
}  // namespace ff
