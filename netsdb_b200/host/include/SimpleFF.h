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
void inference_unit(pdb::PDBClient &pdbClient, std::string database, std::string w1, std::string wo, std::string inputs,
                    std::string b1, std::string bo, std::string output, double dropout_rate, bool enablePartition = false);
void cleanup(pdb::PDBClient &pdbClient, std::string database);
}  // namespace ff
