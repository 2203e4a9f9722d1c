// conv_stage.h -- lowering of the conv2d_memory_fusion jobs (src/tests/source/PipelinedConv2dMemFuseTest.cc:136-273)
// to the device, and the host-side home of Image / Kernel sets.
//
//   kernel_bias_join: Write(kernel_flat) <- KernelBiasJoin(ImageBlockToMatrix(ImageChunksToBlock(KernelToChunks(Scan<Kernel>))), Scan(bias))
//                                                                              -> nb200_conv2d_kernel_flat
//   image_ops:        Write(image_flat)  <- ImageBlockToMatrix(ImageChunksToBlock(ImageToChunks(Scan<Image>)))
//                                                                              -> nb200_conv2d_image_flat
//   conv2d:           Write(result)      <- FFAggMatrix(FFTransposeMult(Scan(image_flat), Scan(kernel_flat)))   (ff_stage.h)
//   result-img:       Write<Image>(set)  <- ConvChunksToImage(ImageChunksToBlock(ConvResultToChunks(Scan(result))))
//                                                                              -> nb200_conv2d_result_to_images
// Image and Kernel records are pdb objects with one FFMatrixBlock per channel; the reference keeps them in Pangea pages,
// here they stay in host memory (a registry keyed by database and set) and are packed into one NCHW array when a job
// scans them.
#pragma once
#include <algorithm>
#include <stdexcept>

#include "conv_udfs.h"

namespace nb200_host {

template <class T>
inline std::map<std::pair<std::string, std::string>, std::vector<pdb::Handle<T>>> &host_sets() {
    static std::map<std::pair<std::string, std::string>, std::vector<pdb::Handle<T>>> m;
    return m;
}
template <class T>
inline bool host_set_exists(const std::string &db, const std::string &set) {
    return host_sets<T>().count(std::make_pair(db, set)) != 0;
}
inline void host_set_drop(const std::string &db, const std::string &set) {
    host_sets<conv2d_memory_fusion::Image>().erase(std::make_pair(db, set));
    host_sets<conv2d_memory_fusion::Kernel>().erase(std::make_pair(db, set));
}

// Pack a set of Matrix3D records as [record][channel][x][y], ordered by their index.  The reference places record i at
// rows i * windows of the flat matrix (ImageToChunks.h:52, KernelToChunks.h:49), so two records with one index would
// overwrite each other there; that is reported instead (its loadRandomImages gives every image index 0,
// ImageUtils.cc:158).
template <class T>
inline std::vector<double> pack_records(const std::string &db, const std::string &set, int *n, int *z, int *x, int *y) {
    auto it = host_sets<T>().find(std::make_pair(db, set));
    if (it == host_sets<T>().end() || it->second.empty()) throw std::runtime_error("set " + db + ":" + set + " holds no records");
    std::vector<pdb::Handle<T>> recs = it->second;
    std::sort(recs.begin(), recs.end(), [](const pdb::Handle<T> &a, const pdb::Handle<T> &b) { return a->index < b->index; });
    *n = static_cast<int>(recs.size()); *z = recs[0]->z; *x = recs[0]->x; *y = recs[0]->y;
    std::vector<double> out(static_cast<size_t>(*n) * *z * *x * *y);
    for (int i = 0; i < *n; i++) {
        T &r = *recs[i];
        if (r.index != i) throw std::runtime_error("set " + db + ":" + set + ": record indices must be 0..n-1 without repeats");
        if (r.x != *x || r.y != *y || r.z != *z || static_cast<int>(r.matrices.size()) != *z)
            throw std::runtime_error("set " + db + ":" + set + ": records differ in shape");
        for (int c = 0; c < *z; c++)
            memcpy(&out[((static_cast<size_t>(i) * *z + c) * *x) * *y], r.matrices[c]->getValue().rawData->c_ptr(),
                   static_cast<size_t>(*x) * *y * sizeof(double));
    }
    return out;
}

class ConvStage {
public:
    // returns true when `w` is one of the conv jobs (and has been run)
    static bool try_run(pdb::WriteSetBase *w) {
        using namespace conv2d_memory_fusion;
        pdb::Handle<pdb::Computation> top = w->getIthInput(0);
        const std::string db = w->getDatabaseName(), out = w->getSetName();
        if (auto *to_img = as<ConvChunksToImage>(top)) {
            auto *blk = chain<ImageChunksToBlock>(top);
            auto *chk = blk ? as<ConvResultToChunks>(blk->getIthInput(0)) : nullptr;
            auto *scan = chk ? as<pdb::ScanSetBase>(chk->getIthInput(0)) : nullptr;
            if (!scan) throw std::runtime_error("ConvChunksToImage over an unsupported graph (no CPU fallback)");
            nb200_set *res = nullptr;
            check(nb200_set_lookup(scan->getDatabaseName().c_str(), scan->getSetName().c_str(), &res));
            const int n = chk->num_images, x = to_img->x, y = to_img->y, z = to_img->z;
            std::vector<double> buf(static_cast<size_t>(n) * z * x * y);
            check(nb200_conv2d_result_to_images(res, n, x, y, z, buf.data()));
            auto &dst = host_sets<Image>()[std::make_pair(db, out)];
            dst.clear();
            for (int i = 0; i < n; i++) {
                pdb::Handle<Image> im = pdb::makeObject<Image>(i, x, y, z);
                for (int c = 0; c < z; c++) {
                    pdb::Handle<FFMatrixBlock> ch = pdb::makeObject<FFMatrixBlock>(0, 0, x, y, x, y);
                    memcpy(ch->getValue().rawData->c_ptr(), &buf[((static_cast<size_t>(i) * z + c) * x) * y], static_cast<size_t>(x) * y * sizeof(double));
                    im->addChannel(ch);
                }
                dst.push_back(im);
            }
            return true;
        }
        pdb::Handle<pdb::Computation> mat = top;
        nb200_set *bias = nullptr;
        if (as<KernelBiasJoin>(top)) {
            mat = top->getIthInput(0);
            auto *bscan = as<pdb::ScanSetBase>(top->getIthInput(1));
            if (!bscan) throw std::runtime_error("KernelBiasJoin: the second input must be a scanned bias set");
            check(nb200_set_lookup(bscan->getDatabaseName().c_str(), bscan->getSetName().c_str(), &bias));
        }
        auto *ibm = as<ImageBlockToMatrix>(mat);
        if (!ibm) return false;
        auto *blk = chain<ImageChunksToBlock>(mat);
        if (!blk) throw std::runtime_error("ImageBlockToMatrix needs ImageChunksToBlock below it (no CPU fallback)");
        pdb::Handle<pdb::Computation> chunks = blk->getIthInput(0);
        auto *scan = chunks == nullptr ? nullptr : as<pdb::ScanSetBase>(chunks->getIthInput(0));
        if (!scan) throw std::runtime_error("the chunking computation must read a scanned set");
        if (!ibm->padding) throw std::runtime_error("ImageBlockToMatrix: block_padding = false is not supported on the device");
        if (blk->chunk_size != ibm->block_x) throw std::runtime_error("ImageChunksToBlock(chunk_size) and ImageBlockToMatrix(block_x) disagree");
        int n = 0, z = 0, x = 0, y = 0;
        nb200_set *o = nullptr;
        if (auto *itc = as<ImageToChunks>(chunks)) {
            if (bias) throw std::runtime_error("KernelBiasJoin over an image pipeline");
            std::vector<double> img = pack_records<Image>(scan->getDatabaseName(), scan->getSetName(), &n, &z, &x, &y);
            check(nb200_conv2d_image_flat(img.data(), n, z, x, y, itc->kernel_size, itc->stride, itc->padding, ibm->block_x, ibm->block_y,
                                          db.c_str(), out.c_str(), &o));
        } else if (as<KernelToChunks>(chunks)) {
            std::vector<double> ker = pack_records<Kernel>(scan->getDatabaseName(), scan->getSetName(), &n, &z, &x, &y);
            if (x != y) throw std::runtime_error("kernels must be square");
            std::vector<double> b(n, 0.0);  // without KernelBiasJoin the placeholder column stays 0 (KernelToChunks.h:79)
            if (bias) {
                int tr = 0, tc = 0;
                nb200_set_info(bias, &tr, &tc, nullptr, nullptr);
                if (tr != n || tc != 1) throw std::runtime_error("KernelBiasJoin: bias must be (number of kernels) x 1");
                check(nb200_download_rect(bias, 0, 0, b.data(), n, 1, 1));
            }
            check(nb200_conv2d_kernel_flat(ker.data(), b.data(), n, z, x, ibm->block_x, ibm->block_y, db.c_str(), out.c_str(), &o));
        } else {
            throw std::runtime_error("ImageChunksToBlock over an unsupported chunking computation (no CPU fallback)");
        }
        int tr = 0, tc = 0;
        nb200_set_info(o, &tr, &tc, nullptr, nullptr);
        if (tr != ibm->batch_size || tc != ibm->total_features)
            std::cerr << "warning: ImageBlockToMatrix was given totals " << ibm->batch_size << " x " << ibm->total_features
                      << " but the set flattens to " << tr << " x " << tc << std::endl;
        return true;
    }

private:
    static void check(int st) {
        if (st != NB200_OK) throw std::runtime_error(nb200_last_error());
    }
    template <class T> static T *as(const pdb::Handle<pdb::Computation> &c) { return c == nullptr ? nullptr : dynamic_cast<T *>(c.ptr().get()); }
    template <class T> static T *chain(const pdb::Handle<pdb::Computation> &c) { return as<T>(c->getIthInput(0)); }
};

}  // namespace nb200_host
