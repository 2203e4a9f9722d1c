// Loaders over the shim (counterpart of src/FF/source/FFMatrixUtil.cc).
#include "FFMatrixUtil.h"

#include <fstream>

#include "PDBClient.h"

namespace ff {

static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
// sign * U(1e-4, 0.5): the distribution of the reference's generator (FFMatrixUtil.cc:229-237), counter based;
// built with -ffp-contract=off so that it matches the device generator bit for bit
double synth_value(uint64_t seed, uint64_t index) {
    const uint64_t h = splitmix64(splitmix64(seed) ^ (index * 0xD1342543DE82EF95ULL + 1ULL));
    const double u = static_cast<double>(h >> 11) * (1.0 / 9007199254740992.0);
    double mag = u * (0.5 - 0.0001);
    mag = mag + 0.0001;
    return (h & 1ULL) ? -mag : mag;
}
uint64_t seed_of_set(const std::string &n) {
    if (n == "inputs") return 1;
    if (n == "w1") return 2;
    if (n == "b1") return 3;
    if (n == "wo") return 4;
    if (n == "bo") return 5;
    if (n == "w2") return 6;
    if (n == "b2") return 7;
    return 1000 + std::hash<std::string>()(n) % 1000000;
}

static void send(pdb::PDBClient &pdbClient, const std::string &dbName, const std::string &setName,
                 pdb::Handle<pdb::Vector<pdb::Handle<FFMatrixBlock>>> &blocks) {
    std::string errMsg;
    if (!pdbClient.sendData<FFMatrixBlock>(std::pair<std::string, std::string>(setName, dbName), blocks, errMsg)) {
        std::cout << "Failed to send data to dispatcher server: " << errMsg << std::endl;
        exit(1);
    }
}

template <class Gen>
static void load_blocks(pdb::PDBClient &pdbClient, const std::string &dbName, const std::string &setName, int totalX, int totalY,
                        int blockX, int blockY, bool dont_pad_x, bool dont_pad_y, bool partitionByCol, Gen value_at) {
    const int numXBlocks = (totalX + blockX - 1) / blockX, numYBlocks = (totalY + blockY - 1) / blockY;
    int total = 0;
    for (int i = 0; i < numXBlocks; i++) {
        // one block row per sendData (the reference sends whenever its 128 MB allocation block fills up)
        auto batch = pdb::makeObject<pdb::Vector<pdb::Handle<FFMatrixBlock>>>();
        const int bx = dont_pad_x ? std::min(blockX, totalX - i * blockX) : blockX;
        for (int j = 0; j < numYBlocks; j++) {
            const int by = dont_pad_y ? std::min(blockY, totalY - j * blockY) : blockY;
            pdb::Handle<FFMatrixBlock> blk = pdb::makeObject<FFMatrixBlock>(i, j, bx, by, totalX, totalY, partitionByCol);
            double *d = blk->getRawDataHandle()->c_ptr();
            for (int ii = 0; ii < bx; ii++) {
                const int r = i * blockX + ii;
                if (r >= totalX) break;  // zero padding stays (Array.cc:84-100)
                for (int jj = 0; jj < by; jj++) {
                    const int c = j * blockY + jj;
                    if (c >= totalY) break;
                    d[static_cast<long>(ii) * by + jj] = value_at(r, c);
                }
            }
            batch->push_back(blk);
            total++;
        }
        send(pdbClient, dbName, setName, batch);
    }
    std::cout << setName << "(" << totalX << "x" << totalY << "): (" << numXBlocks << " x " << numYBlocks << ")" << total
              << " blocks = " << blockX << " x " << blockY << " each" << std::endl;
    std::string errMsg;
    pdbClient.flushData(errMsg);
}

void loadMatrix(pdb::PDBClient &pdbClient, std::string dbName, std::string setName, int totalX, int totalY, int blockX, int blockY,
                bool dont_pad_x, bool dont_pad_y, std::string &, int, bool partitionByCol) {
    if (totalX == 0 || totalY == 0 || blockX == 0 || blockY == 0) return;  // FFMatrixUtil.cc:221-225
    const uint64_t seed = seed_of_set(setName);
    load_blocks(pdbClient, dbName, setName, totalX, totalY, blockX, blockY, dont_pad_x, dont_pad_y, partitionByCol,
                [&](int r, int c) { return synth_value(seed, static_cast<uint64_t>(r) * totalY + c); });
}

void load_matrix_from_file(std::string path, std::vector<std::vector<double>> &matrix) {
    std::ifstream is(path);
    if (!is.is_open()) {
        std::cout << "Failed to open " << path << std::endl;
        exit(1);
    }
    int rows = 0, cols = 0;
    is >> rows;
    is.ignore(1, ',');  // "rows, cols" or "rows cols"
    is >> cols;
    matrix.assign(rows, std::vector<double>(cols));
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) {
            is >> matrix[i][j];
            if (is.peek() == ',') is.ignore();
        }
}

void load_matrix_data(pdb::PDBClient &pdbClient, std::string path, std::string dbName, std::string setName, int blockX, int blockY,
                      bool dont_pad_x, bool dont_pad_y, std::string &, int, bool partitionByCol) {
    std::vector<std::vector<double>> m;
    load_matrix_from_file(path, m);
    if (m.empty()) return;
    if (blockY == -1 || blockY > static_cast<int>(m[0].size())) blockY = static_cast<int>(m[0].size());
    load_blocks(pdbClient, dbName, setName, static_cast<int>(m.size()), static_cast<int>(m[0].size()), blockX, blockY, dont_pad_x,
                dont_pad_y, partitionByCol, [&](int r, int c) { return m[r][c]; });
}

void print_stats(pdb::PDBClient &pdbClient, std::string dbName, std::string setName) {
    int count = 0, rows = 0, cols = 0, totalRows = 0, totalCols = 0;
    auto it = pdbClient.getSetIterator<FFMatrixBlock>(dbName, setName);
    for (auto r : it) {
        count++;
        rows = r->getRowNums(); cols = r->getColNums(); totalRows = r->getTotalRowNums(); totalCols = r->getTotalColNums();
    }
    std::cout << setName << " ( " << totalRows << " X " << totalCols << " ) (" << count << " ) : Each block size: " << rows << " X "
              << cols << std::endl;
}

// FFMatrixUtil.cc:388-415 of the reference: the set's totals from its first block, then the number of blocks (the
// per-element dump is commented out there)
void print(pdb::PDBClient &pdbClient, std::string dbName, std::string setName) {
    int count = 0;
    auto it = pdbClient.getSetIterator<FFMatrixBlock>(dbName, setName);
    for (auto r : it) {
        if (count == 0) std::cout << "[STATS] " << setName << " : " << r->getTotalRowNums() << " X " << r->getTotalColNums() << std::endl;
        count++;
    }
    std::cout << "count = " << count << std::endl;
}

bool is_empty_set(pdb::PDBClient &pdbClient, pdb::CatalogClient &catalogClient, std::string dbName, std::string setName) {
    if (!catalogClient.setExists(dbName, setName)) return true;
    auto it = pdbClient.getSetIterator<FFMatrixBlock>(dbName, setName);
    for (auto r : it) { (void)r; return false; }
    return true;
}

}  // namespace ff
