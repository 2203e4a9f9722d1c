// Conv2dMemFuseTest -- the conv2d_memory_fusion pipeline wired with the same Computation objects and set names as the
// reference's driver (src/tests/source/PipelinedConv2dMemFuseTest.cc:136-273: kernel_bias_join, image_ops, conv2d and
// the commented-out result-img job), on generated data instead of the files under /home/ubuntu that driver reads, and
// with the `result` matrix and the re-assembled images written out as text for the parity tests.
//
//   Conv2dMemFuseTest n channels h w kout k stride padding block_x block_y dumpdir
//   images: seed 41, kernels: seed 42, bias: seed 43 (element e of the NCHW array = ff::synth_value(seed, e))
#include <cstdio>

#include "ConvChunksToImage.h"
#include "ConvResultToChunks.h"
#include "FFMatrixUtil.h"
#include "ImageBlockToMatrix.h"
#include "ImageChunksToBlock.h"
#include "ImageToChunks.h"
#include "ImageUtils.h"
#include "KernelBiasJoin.h"
#include "KernelToChunks.h"
#include "PDBClient.h"

using namespace conv2d_memory_fusion;
using pdb::Computation;
using pdb::Handle;
using pdb::makeObject;

static void run(pdb::PDBClient &client, const char *job, Handle<Computation> sink) {
    std::string errMsg;
    if (!client.executeComputations(errMsg, job, sink)) {
        std::cout << "Computation failed. Message was: " << errMsg << "\n";
        exit(1);
    }
}

int main(int argc, char *argv[]) {
    if (argc < 12) {
        std::cout << "Usage: n channels h w kout k stride padding block_x block_y dumpdir" << std::endl;
        return -1;
    }
    const int n = atoi(argv[1]), ch = atoi(argv[2]), h = atoi(argv[3]), w = atoi(argv[4]), kout = atoi(argv[5]), k = atoi(argv[6]);
    const int stride = atoi(argv[7]), padding = atoi(argv[8]), block_x = atoi(argv[9]), block_y = atoi(argv[10]);
    const std::string dir = argv[11], db = "conv2d";
    std::string errMsg;
    pdb::PDBLoggerPtr logger = std::make_shared<pdb::PDBLogger>("Conv2DclientLog");
    pdb::PDBClient client(8108, "localhost", logger, false, true);
    client.createDatabase(db, errMsg);
    for (const char *s : {"img", "kernel"}) client.createSet<Image>(db, s, errMsg, static_cast<size_t>(128) << 20, s);
    for (const char *s : {"bias", "kernel_flat", "image_flat", "result"}) client.createSet<FFMatrixBlock>(db, s, errMsg, static_cast<size_t>(128) << 20, s);
    load_seeded<Image>(client, db, "img", n, ch, h, w, 41);
    load_seeded<Kernel>(client, db, "kernel", kout, ch, k, k, 42);
    {   // bias: kout x 1 in blocks of block_x x 1 (PipelinedConv2dMemFuseTest.cc:181-182), seed 43
        auto blocks = makeObject<pdb::Vector<Handle<FFMatrixBlock>>>();
        for (int b = 0; b * block_x < kout; b++) {
            Handle<FFMatrixBlock> blk = makeObject<FFMatrixBlock>(b, 0, block_x, 1, kout, 1);
            for (int i = 0; i < block_x && b * block_x + i < kout; i++) blk->getRawDataHandle()->c_ptr()[i] = ff::synth_value(43, b * block_x + i);
            blocks->push_back(blk);
        }
        client.sendData<FFMatrixBlock>(std::make_pair(std::string("bias"), db), blocks, errMsg);
    }
    const int hout = (h + 2 * padding - k) / stride + 1, wout = (w + 2 * padding - k) / stride + 1, features = ch * k * k + 1;
    {   // kernel_bias_join
        Handle<Computation> chunk = makeObject<KernelToChunks>(block_y);
        chunk->setInput(makeObject<pdb::ScanUserSet<Kernel>>(db, "kernel"));
        Handle<Computation> block = makeObject<ImageChunksToBlock>(block_x);
        block->setInput(chunk);
        Handle<Computation> matrix = makeObject<ImageBlockToMatrix>(block_x, block_y, true, kout, features);
        matrix->setInput(block);
        Handle<Computation> join = makeObject<KernelBiasJoin>();
        join->setInput(0, matrix);
        join->setInput(1, makeObject<pdb::ScanUserSet<FFMatrixBlock>>(db, "bias"));
        Handle<Computation> sink = makeObject<pdb::WriteUserSet<FFMatrixBlock>>(db, "kernel_flat");
        sink->setInput(join);
        run(client, "kernel_bias_join", sink);
    }
    {   // image_ops
        Handle<Computation> chunk = makeObject<ImageToChunks>(block_y, k, stride, padding);
        chunk->setInput(makeObject<pdb::ScanUserSet<Image>>(db, "img"));
        Handle<Computation> block = makeObject<ImageChunksToBlock>(block_x);
        block->setInput(chunk);
        Handle<Computation> matrix = makeObject<ImageBlockToMatrix>(block_x, block_y, true, n * hout * wout, features);
        matrix->setInput(block);
        Handle<Computation> sink = makeObject<pdb::WriteUserSet<FFMatrixBlock>>(db, "image_flat");
        sink->setInput(matrix);
        run(client, "image_ops", sink);
    }
    {   // conv2d
        Handle<Computation> join = makeObject<FFTransposeMult>();
        join->setInput(0, makeObject<pdb::ScanUserSet<FFMatrixBlock>>(db, "image_flat"));
        join->setInput(1, makeObject<pdb::ScanUserSet<FFMatrixBlock>>(db, "kernel_flat"));
        Handle<Computation> agg = makeObject<FFAggMatrix>();
        agg->setInput(join);
        Handle<Computation> sink = makeObject<pdb::WriteUserSet<FFMatrixBlock>>(db, "result");
        sink->setInput(agg);
        run(client, "conv2d", sink);
    }
    {   // result-img
        Handle<Computation> chunk = makeObject<ConvResultToChunks>(n);
        chunk->setInput(makeObject<pdb::ScanUserSet<FFMatrixBlock>>(db, "result"));
        Handle<Computation> block = makeObject<ImageChunksToBlock>(hout * wout);
        block->setInput(chunk);
        Handle<Computation> images = makeObject<ConvChunksToImage>(hout, wout, kout);
        images->setInput(block);
        Handle<Computation> sink = makeObject<pdb::WriteUserSet<Image>>(db, "result_chunked2");
        sink->setInput(images);
        run(client, "result-img", sink);
    }
    ff::print_stats(client, db, "image_flat");
    ff::print_stats(client, db, "kernel_flat");
    ff::print_stats(client, db, "result");
    // images as text: n * kout * hout * wout values, NCHW
    FILE *f = fopen((dir + "/images.out.txt").c_str(), "w");
    int count = 0;
    for (auto im : client.getSetIterator<Image>(db, "result_chunked2")) {
        count++;
        for (int c = 0; c < im->z; c++) {
            const double *d = im->getMatrixAtIndex(c)->getValue().rawData->c_ptr();
            for (int i = 0; i < im->x * im->y; i++) fprintf(f, "%.17g\n", d[i]);
        }
    }
    fclose(f);
    std::cout << "images=" << count << std::endl;
    return 0;
}
