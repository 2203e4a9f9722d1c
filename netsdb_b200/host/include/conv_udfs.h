// conv_udfs.h -- the conv2d_memory_fusion types and computations by their netsDB names, base classes and constructor
// signatures (src/conv2d_memory_fusion/headers/Matrix3D.h, Image.h, Kernel.h, ImageChunk.h, ImageBlock.h,
// ImageToChunks.h, ImageChunksToBlock.h, ImageBlockToMatrix.h, KernelToChunks.h, KernelBiasJoin.h,
// ConvResultToChunks.h, ConvChunksToImage.h).  Image and Kernel records stay on the host until a job needs them; the
// three jobs of the pipeline (kernel_bias_join, image_ops, conv2d) and the result re-assembly are lowered as wholes to
// the device (conv_stage.h).  The per-record projections are therefore never executed here.
#pragma once
#include <cassert>

#include "ff_udfs.h"

namespace conv2d_memory_fusion {

// x * y * z tensor kept as z channel matrices (Matrix3D.h:15-36)
class Matrix3D : public pdb::Object {
public:
    int x = 0, y = 0, z = 0;
    pdb::Vector<pdb::Handle<FFMatrixBlock>> matrices;
    ENABLE_DEEP_COPY
    Matrix3D() {}
    Matrix3D(int x, int y, int z) : x(x), y(y), z(z) { matrices.resize(z); }
    void addChannel(pdb::Handle<FFMatrixBlock> data) { matrices.push_back(data); }
    void setMatrixAtIndex(int i, pdb::Handle<FFMatrixBlock> data) { assert(i < z); matrices[i] = data; }
    pdb::Handle<FFMatrixBlock> getMatrixAtIndex(int i) { assert(i < z); return matrices[i]; }
};

// Image.h:12-77: sliding-window geometry helpers; h_slides = (y + 2p - k) / stride + 1 in integer arithmetic
class Image : public Matrix3D {
public:
    int index = 0;
    ENABLE_DEEP_COPY
    Image() {}
    Image(int index, int x, int y, int z) : Matrix3D(x, y, z), index(index) {}
    int &getKey() { return index; }
    Image &getValue() { return *this; }
    int get_x(int padding) { return x + padding * 2; }
    int get_y(int padding) { return y + padding * 2; }
    int get_h_slides(int padding, int kernel_size, int stride) { return (get_y(padding) - kernel_size) / stride + 1; }
    int get_v_slides(int padding, int kernel_size, int stride) { return (get_x(padding) - kernel_size) / stride + 1; }
    // value under kernel tap (i, j) of window `window` (row-major over output positions); 0 outside the image, which
    // can only happen with padding > 0 (Image.h:36-57)
    double get_win_value(int window, int i, int j, int stride, int channel, int kernel_size, int padding) {
        const int h_slides = get_h_slides(padding, kernel_size, stride);
        const int xi = (window / h_slides) * stride + i - padding, yi = (window % h_slides) * stride + j - padding;
        if (padding > 0 && (xi < 0 || yi < 0 || xi >= x || yi >= y)) return 0;
        return matrices[channel]->getValue().rawData->c_ptr()[xi * y + yi];
    }
    int get_conv2d_matrix_rows(int kernel_size, int stride, int padding) {
        return get_v_slides(padding, kernel_size, stride) * get_h_slides(padding, kernel_size, stride);
    }
    int get_conv2d_matrix_cols(int kernel_size, int, int) { return z * kernel_size * kernel_size; }
    int get_conv2d_window_count(int kernel_size, int stride, int padding) { return get_conv2d_matrix_rows(kernel_size, stride, padding); }
    int get_num_channels() { return z; }
};

class Kernel : public Matrix3D {
public:
    int index = 0;
    ENABLE_DEEP_COPY
    Kernel() {}
    Kernel(int index, int x, int y, int z) : Matrix3D(x, y, z), index(index) {}
    int getKey() { return index; }
    int get_num_channels() { return z; }
};

// intermediate record types of the reference pipeline (ImageChunk.h, ImageBlock.h); never materialised here
class ImageChunk : public pdb::Object {
public:
    int block_row = 0, y_index = 0;
    pdb::Vector<double> chunk;
    int chunk_size = 0, y_size = 0, block_row_start = 0;
    ENABLE_DEEP_COPY
    ImageChunk() {}
    ImageChunk(int size, int y_size, int block_row, int y_index, int block_row_start)
        : block_row(block_row), y_index(y_index), chunk_size(size), y_size(y_size), block_row_start(block_row_start) { chunk.resize(size); }
    int getChunkActualRowIndex() { return block_row_start + block_row; }
    pdb::Vector<double> &getChunk() { return chunk; }
    int getSize() { return static_cast<int>(chunk.size()); }
};
class ImageBlock : public pdb::Object {
public:
    long key = 0;
    int block_x_index = 0, block_x_size = 0, block_y_index = 0, block_y_size = 0;
    pdb::Map<int, pdb::Vector<double>> chunks;
    ENABLE_DEEP_COPY
    ImageBlock() {}
    ImageBlock(ImageChunk &c, int block_x_size)
        : block_x_index(c.getChunkActualRowIndex() / block_x_size), block_x_size(block_x_size), block_y_index(c.y_index),
          block_y_size(c.y_size) { key = block_x_index * 10000 + block_y_index; }
    long &getKey() { return key; }
    ImageBlock &getValue() { return *this; }
    pdb::Map<int, pdb::Vector<double>> &getBlock() { return chunks; }
};

#define NB200_CONV_NOT_ON_HOST(NAME, OUT, ...)                                                   \
    return pdb::makeLambda(__VA_ARGS__, [](auto &...) {                                          \
        nb200_host::no_host_projection(#NAME);                                                   \
        return OUT();                                                                            \
    });

// Image -> row chunks of the im2col matrix (ImageToChunks.h:23-102)
class ImageToChunks : public pdb::MultiSelectionComp<ImageChunk, Image> {
public:
    ENABLE_DEEP_COPY
    int block_y = 0, kernel_size = 0, stride = 1, padding = 0;
    ImageToChunks() {}
    ImageToChunks(int block_y, int kernel_size, int stride, int padding = 0)
        : block_y(block_y), kernel_size(kernel_size), stride(stride), padding(padding) {}
    pdb::Lambda<bool> getSelection(pdb::Handle<Image> checkMe) override { return pdb::makeLambda(checkMe, [](pdb::Handle<Image> &) { return true; }); }
    pdb::Lambda<pdb::Vector<pdb::Handle<ImageChunk>>> getProjection(pdb::Handle<Image> checkMe) override {
        NB200_CONV_NOT_ON_HOST(ImageToChunks, pdb::Vector<pdb::Handle<ImageChunk>>, checkMe)
    }
};
// Kernel -> row chunks of kernel_flat (KernelToChunks.h:24-83)
class KernelToChunks : public pdb::MultiSelectionComp<ImageChunk, Kernel> {
public:
    ENABLE_DEEP_COPY
    int block_y = 0;
    KernelToChunks() {}
    KernelToChunks(int block_y) : block_y(block_y) {}
    pdb::Lambda<bool> getSelection(pdb::Handle<Kernel> checkMe) override { return pdb::makeLambda(checkMe, [](pdb::Handle<Kernel> &) { return true; }); }
    pdb::Lambda<pdb::Vector<pdb::Handle<ImageChunk>>> getProjection(pdb::Handle<Kernel> checkMe) override {
        NB200_CONV_NOT_ON_HOST(KernelToChunks, pdb::Vector<pdb::Handle<ImageChunk>>, checkMe)
    }
};
// chunks -> blocks keyed blockRow * 10000 + yIndex (ImageChunksToBlock.h:22-50)
class ImageChunksToBlock : public pdb::ClusterAggregateComp<ImageBlock, ImageChunk, long, ImageBlock> {
public:
    ENABLE_DEEP_COPY
    int chunk_size = 0;
    ImageChunksToBlock() {}
    ImageChunksToBlock(int chunk_size) : chunk_size(chunk_size) {}
    pdb::Lambda<long> getKeyProjection(pdb::Handle<ImageChunk> aggMe) override { NB200_CONV_NOT_ON_HOST(ImageChunksToBlock, long, aggMe) }
    pdb::Lambda<ImageBlock> getValueProjection(pdb::Handle<ImageChunk> aggMe) override { NB200_CONV_NOT_ON_HOST(ImageChunksToBlock, ImageBlock, aggMe) }
};
// block of chunks -> FFMatrixBlock, zero padded when `padding` (ImageBlockToMatrix.h:22-73)
class ImageBlockToMatrix : public pdb::SelectionComp<::FFMatrixBlock, ImageBlock> {
public:
    ENABLE_DEEP_COPY
    int block_x = 0, block_y = 0;
    bool padding = true;
    int batch_size = 0, total_features = 0;
    ImageBlockToMatrix() {}
    ImageBlockToMatrix(int block_x, int block_y, bool padding, int batch_size, int total_features)
        : block_x(block_x), block_y(block_y), padding(padding), batch_size(batch_size), total_features(total_features) {}
    pdb::Lambda<bool> getSelection(pdb::Handle<ImageBlock> checkMe) override { return pdb::makeLambda(checkMe, [](pdb::Handle<ImageBlock> &) { return true; }); }
    pdb::Lambda<pdb::Handle<::FFMatrixBlock>> getProjection(pdb::Handle<ImageBlock> checkMe) override {
        NB200_CONV_NOT_ON_HOST(ImageBlockToMatrix, pdb::Handle<::FFMatrixBlock>, checkMe)
    }
};
// writes the bias into the placeholder column of kernel_flat (KernelBiasJoin.h:14-52)
class KernelBiasJoin : public pdb::JoinComp<::FFMatrixBlock, ::FFMatrixBlock, ::FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    KernelBiasJoin() {}
    NB200_JOIN_UDF(KernelBiasJoin, getBlockRowIndex, getBlockRowIndex)
};
// result matrix -> column chunks per image (ConvResultToChunks.h:22-108)
class ConvResultToChunks : public pdb::MultiSelectionComp<ImageChunk, ::FFMatrixBlock> {
public:
    ENABLE_DEEP_COPY
    int num_images = 0;
    ConvResultToChunks() {}
    ConvResultToChunks(int num_images) : num_images(num_images) {}
    pdb::Lambda<bool> getSelection(pdb::Handle<::FFMatrixBlock> checkMe) override { return pdb::makeLambda(checkMe, [](pdb::Handle<::FFMatrixBlock> &) { return true; }); }
    pdb::Lambda<pdb::Vector<pdb::Handle<ImageChunk>>> getProjection(pdb::Handle<::FFMatrixBlock> checkMe) override {
        NB200_CONV_NOT_ON_HOST(ConvResultToChunks, pdb::Vector<pdb::Handle<ImageChunk>>, checkMe)
    }
};
// chunks of one image -> Image with z channels of x * y (ConvChunksToImage.h:22-72)
class ConvChunksToImage : public pdb::ClusterAggregateComp<Image, ImageBlock, long, Image> {
public:
    ENABLE_DEEP_COPY
    int x = 0, y = 0, z = 0;
    ConvChunksToImage() {}
    ConvChunksToImage(int x, int y, int z) : x(x), y(y), z(z) {}
    pdb::Lambda<long> getKeyProjection(pdb::Handle<ImageBlock> aggMe) override { NB200_CONV_NOT_ON_HOST(ConvChunksToImage, long, aggMe) }
    pdb::Lambda<Image> getValueProjection(pdb::Handle<ImageBlock> aggMe) override { NB200_CONV_NOT_ON_HOST(ConvChunksToImage, Image, aggMe) }
};

}  // namespace conv2d_memory_fusion
