// Image / kernel loaders over the shim (counterpart of src/conv2d_memory_fusion/source/ImageUtils.cc).
#include "ImageUtils.h"

#include <fstream>

#include "FFMatrixUtil.h"
#include "PDBClient.h"

namespace conv2d_memory_fusion {

template <typename M, class Gen>
static void load_records(pdb::PDBClient &pdbClient, const std::string &dbName, const std::string &setName, int count, int channels,
                         int x, int y, Gen next_value) {
    static_assert(std::is_base_of<Matrix3D, M>::value, "M must be a descendant of conv2d_memory_fusion::Matrix3D");
    std::string errMsg;
    pdb::Handle<pdb::Vector<pdb::Handle<M>>> records = pdb::makeObject<pdb::Vector<pdb::Handle<M>>>();
    for (int im = 0; im < count; im++) {
        pdb::Handle<M> rec = pdb::makeObject<M>(im, x, y, channels);
        for (int c = 0; c < channels; c++) {
            pdb::Handle<FFMatrixBlock> chan = pdb::makeObject<FFMatrixBlock>(0, 0, x, y, x, y);
            double *d = chan->getRawDataHandle()->c_ptr();
            for (int i = 0; i < x * y; i++) d[i] = next_value();
            rec->addChannel(chan);
        }
        records->push_back(rec);
    }
    if (!pdbClient.sendData(std::pair<std::string, std::string>(setName, dbName), records, errMsg)) {
        std::cout << "Failed to send data to dispatcher server" << std::endl;
        exit(1);
    }
    std::cout << "Sent total of " << count << " images." << std::endl;
    pdbClient.flushData(errMsg);
}

template <typename M>
void load_imgs_from_file(pdb::PDBClient &pdbClient, std::string path, pdb::String dbName, pdb::String setName, int img_count,
                         int channels, int x, int y) {
    if (path.empty()) throw std::runtime_error("Invalid filepath: " + path);
    std::ifstream is(path);
    if (!is.is_open()) {
        std::cout << "Failed to open " << path << std::endl;
        exit(1);
    }
    load_records<M>(pdbClient, dbName, setName, img_count, channels, x, y, [&]() {
        double v = 0.0;
        is >> v;
        while (is.peek() == ',' || is.peek() == ' ') is.ignore();
        return v;
    });
}

template <typename M>
void load_seeded(pdb::PDBClient &pdbClient, pdb::String dbName, pdb::String setName, int count, int channels, int x, int y,
                 uint64_t seed) {
    uint64_t e = 0;
    load_records<M>(pdbClient, dbName, setName, count, channels, x, y, [&]() { return ff::synth_value(seed, e++); });
}

void loadRandomImages(int width, int height, int channels, int numOfImages, pdb::PDBClient &pdbClient, pdb::String dbName,
                      pdb::String setName, int) {
    load_seeded<Image>(pdbClient, dbName, setName, numOfImages, channels, width, height, ff::seed_of_set(setName));
}

template void load_imgs_from_file<Image>(pdb::PDBClient &, std::string, pdb::String, pdb::String, int, int, int, int);
template void load_imgs_from_file<Kernel>(pdb::PDBClient &, std::string, pdb::String, pdb::String, int, int, int, int);
template void load_seeded<Image>(pdb::PDBClient &, pdb::String, pdb::String, int, int, int, int, uint64_t);
template void load_seeded<Kernel>(pdb::PDBClient &, pdb::String, pdb::String, int, int, int, int, uint64_t);

}  // namespace conv2d_memory_fusion
