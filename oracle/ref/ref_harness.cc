/* ref_harness.cc -- TEST INFRASTRUCTURE (part of the oracle).
 *
 * Runs the REFERENCE'S OWN UDF code for the path.  The headers included below are netsDB's, unmodified, compiled from
 * where they lie under /root/reference (oracle/Makefile, target `ref`; output oracle/_ref/libff_ref.so):
 *   src/FF/headers: FFMatrixBlock.h FFMatrixData.h FFMatrixMeta.h FFTransposeMult.h FFInputLayerJoin.h FFAggMatrix.h
 *                   FFReluBiasSum.h FFTransposeBiasSum.h FFRowAggregate.h FFOutputLayer.h
 *   src/LSTM/headers: LSTMThreeWaySum.h LSTMTwoSum.h LSTMHiddenState.h
 *   src/conv2d_memory_fusion/headers: Image.h Kernel.h ImageChunk.h ImageBlock.h ImageToChunks.h ImageChunksToBlock.h
 *                   ImageBlockToMatrix.h KernelToChunks.h KernelBiasJoin.h ConvResultToChunks.h ConvChunksToImage.h
 *   src/sharedLibraries/headers: MatrixBlock.h MatrixData.h MatrixMeta.h LASillyMultiply1Join.h
 *                   LASillyTransposeMultiply1Join.h LASillyMultiply2Aggregate.h
 *   src/FF_proj: headers/FullyConnectedNetwork.h + source/FullyConnectedNetwork.cc (sigmoid, relu, outLabel, randomGen);
 *   src/word2vec/headers: SemanticClassifier.h (its .cc defines the same three functions again, so only one is compiled)
 * Two of their dependencies are stood in for (see those files): netsDB's object model (pdbmin/pdb_min.h: handles,
 * vectors, lambdas evaluated in place) and Eigen 3 (minieigen: plain-loop products and row sums).
 *
 * What this file adds is only the execution a netsDB worker would do around the UDFs, single-threaded and in arrival
 * order: a JoinComp is "for every pair: if getSelection then getProjection" (JoinTuple.h:381-565), a
 * ClusterAggregateComp keeps one value per key, deep-copies the first and does `value = value + next` for the rest
 * (CombinedShuffleSink.h:112-144) and re-wraps (key, value) with a default-constructed output object
 * (AggOutProcessor.cc:78-80); a SelectionComp / MultiSelectionComp maps its projection over the set.
 *
 * Sets cross the C boundary as nbo_set (ff_oracle.h), so that the Python tests compare block by block -- meta,
 * padding and values -- with the restatement in ff_oracle.c.
 */
#include <cassert>
#include <vector>

#include "FFAggMatrix.h"
#include "FFInputLayerJoin.h"
#include "FFMatrixBlock.h"
#include "FFOutputLayer.h"
#include "FFReluBiasSum.h"
#include "FFRowAggregate.h"
#include "FFTransposeBiasSum.h"
#include "FFTransposeMult.h"
#include "LSTMHiddenState.h"
#include "LSTMThreeWaySum.h"
#include "LSTMTwoSum.h"
#include "ConvChunksToImage.h"
#include "ConvResultToChunks.h"
#include "ImageBlockToMatrix.h"
#include "ImageChunksToBlock.h"
#include "ImageToChunks.h"
#include "KernelBiasJoin.h"
#include "KernelToChunks.h"
#include "LASillyMultiply1Join.h"
#include "LASillyMultiply2Aggregate.h"
#include "LASillyTransposeMultiply1Join.h"
/* the network's weights are private members filled from std::random_device: the harness has to read them to hand the
 * same weights to the restatement, hence the keyword trick around this one include */
#include <functional>
#include <random>
#define private public
#include "FullyConnectedNetwork.cc"
#undef private
#include "SemanticClassifier.h"

extern "C" {
#include "../ff_oracle.h"
}

using pdb::Handle;
using pdb::makeObject;
template <class T> using Set = std::vector<Handle<T>>;

/* ---- the engine around the UDFs ------------------------------------------------------------------------ */
template <class Out, class J, class A, class B>
static Set<Out> run_join(J &j, Set<A> &as, Set<B> &bs) {
    Set<Out> out;
    for (auto &a : as)
        for (auto &b : bs)
            if (j.getSelection(a, b).eval()) out.push_back(j.getProjection(a, b).eval());
    return out;
}
template <class Out, class J, class A, class B, class C>
static Set<Out> run_join3(J &j, Set<A> &as, Set<B> &bs, Set<C> &cs) {
    Set<Out> out;
    for (auto &a : as)
        for (auto &b : bs)
            for (auto &c : cs)
                if (j.getSelection(a, b, c).eval()) out.push_back(j.getProjection(a, b, c).eval());
    return out;
}
template <class T> static bool key_eq(const Handle<T> &a, const Handle<T> &b) { return *a == *b; }
template <class T> static bool key_eq(const T &a, const T &b) { return a == b; }

template <class Out, class Key, class Val, class G, class In>
static Set<Out> run_agg(G &g, Set<In> &in) {
    std::vector<Key> keys;
    std::vector<Val> vals;
    for (auto &x : in) {
        Key k = g.getKeyProjection(x).eval();
        Val v = g.getValueProjection(x).eval();
        size_t i = 0;
        while (i < keys.size() && !key_eq(keys[i], k)) i++;
        if (i == keys.size()) { keys.push_back(k); vals.push_back(v); }
        else vals[i] = vals[i] + v;
    }
    Set<Out> out;
    for (size_t i = 0; i < keys.size(); i++) {
        Handle<Out> o = makeObject<Out>();
        o->getKey() = keys[i];
        o->getValue() = vals[i];
        out.push_back(o);
    }
    return out;
}
template <class Out, class S, class In>
static Set<Out> run_select(S &s, Set<In> &in) {
    Set<Out> out;
    for (auto &x : in)
        if (s.getSelection(x).eval()) out.push_back(s.getProjection(x).eval());
    return out;
}
template <class Out, class S, class In>
static Set<Out> run_multiselect(S &s, Set<In> &in) {
    Set<Out> out;
    for (auto &x : in)
        if (s.getSelection(x).eval()) {
            pdb::Vector<Handle<Out>> v = s.getProjection(x).eval();
            for (size_t i = 0; i < v.size(); i++) out.push_back(v[i]);
        }
    return out;
}

/* ---- nbo_set <-> the reference's block types -------------------------------------------------------------- */
static Set<FFMatrixBlock> ff_blocks(const nbo_set *s) {
    Set<FFMatrixBlock> out;
    for (int i = 0; i < s->n; i++) {
        const nbo_block *b = &s->blocks[i];
        Handle<FFMatrixBlock> h = makeObject<FFMatrixBlock>(b->block_row, b->block_col, b->rows, b->cols, b->total_rows, b->total_cols);
        memcpy(h->getValue().rawData->c_ptr(), b->data, (size_t)b->rows * b->cols * sizeof(double));
        out.push_back(h);
    }
    return out;
}
static nbo_set *to_nbo(Set<FFMatrixBlock> &blocks) {
    nbo_set *s = nbo_set_new();
    for (auto &h : blocks) {
        nbo_block *b = nbo_set_add_block(s, h->getBlockRowIndex(), h->getBlockColIndex(), h->getRowNums(), h->getColNums(),
                                         h->getTotalRowNums(), h->getTotalColNums());
        memcpy(b->data, h->getValue().rawData->c_ptr(), (size_t)b->rows * b->cols * sizeof(double));
    }
    return s;
}
static Set<MatrixBlock> la_blocks(const nbo_set *s) {
    Set<MatrixBlock> out;
    for (int i = 0; i < s->n; i++) {
        const nbo_block *b = &s->blocks[i];
        Handle<MatrixBlock> h = makeObject<MatrixBlock>(b->block_row, b->block_col, b->rows, b->cols, b->total_rows, b->total_cols);
        memcpy(h->getRawDataHandle()->c_ptr(), b->data, (size_t)b->rows * b->cols * sizeof(double));
        out.push_back(h);
    }
    return out;
}
static nbo_set *to_nbo(Set<MatrixBlock> &blocks) {
    nbo_set *s = nbo_set_new();
    for (auto &h : blocks) {
        nbo_block *b = nbo_set_add_block(s, h->getBlockRowIndex(), h->getBlockColIndex(), h->getRowNums(), h->getColNums(),
                                         h->getTotalRowNums(), h->getTotalColNums());
        memcpy(b->data, h->getRawDataHandle()->c_ptr(), (size_t)b->rows * b->cols * sizeof(double));
    }
    return s;
}
template <class M>
static Set<M> records(const double *data, int n, int c, int x, int y) {
    Set<M> out;
    for (int i = 0; i < n; i++) {
        Handle<M> r = makeObject<M>(i, x, y, c);
        for (int ch = 0; ch < c; ch++) {
            Handle<FFMatrixBlock> m = makeObject<FFMatrixBlock>(0, 0, x, y, x, y);
            memcpy(m->getValue().rawData->c_ptr(), data + ((size_t)i * c + ch) * x * y, (size_t)x * y * sizeof(double));
            r->addChannel(m);
        }
        out.push_back(r);
    }
    return out;
}

extern "C" {

/* FFTransposeMult | FFInputLayerJoin | LSTMHiddenState, each followed by FFAggMatrix */
nbo_set *ref_join_agg(int kind, const nbo_set *a, const nbo_set *b) {
    Set<FFMatrixBlock> as = ff_blocks(a), bs = ff_blocks(b), joined;
    if (kind == 0) { FFTransposeMult j; joined = run_join<FFMatrixBlock>(j, as, bs); }
    else if (kind == 1) { FFInputLayerJoin j; joined = run_join<FFMatrixBlock>(j, as, bs); }
    else { LSTMHiddenState j; joined = run_join<FFMatrixBlock>(j, as, bs); }
    FFAggMatrix g;
    Set<FFMatrixBlock> out = run_agg<FFMatrixBlock, Handle<FFMatrixMeta>, FFMatrixData>(g, joined);
    return to_nbo(out);
}
/* FFReluBiasSum(dropout_rate) | FFTransposeBiasSum | FFOutputLayer, as written */
nbo_set *ref_bias_join(int kind, const nbo_set *in, const nbo_set *other, double dropout_rate) {
    Set<FFMatrixBlock> as = ff_blocks(in), bs = ff_blocks(other), out;
    if (kind == 0) { FFReluBiasSum j(dropout_rate); out = run_join<FFMatrixBlock>(j, as, bs); }
    else if (kind == 1) { FFTransposeBiasSum j; out = run_join<FFMatrixBlock>(j, as, bs); }
    else { FFOutputLayer j; out = run_join<FFMatrixBlock>(j, as, bs); }
    return to_nbo(out);
}
nbo_set *ref_row_aggregate(const nbo_set *in) {
    Set<FFMatrixBlock> as = ff_blocks(in);
    FFRowAggregate g;
    Set<FFMatrixBlock> out = run_agg<FFMatrixBlock, Handle<FFMatrixMeta>, FFMatrixData>(g, as);
    return to_nbo(out);
}
/* LSTMThreeWaySum(activation 1 | 2) with c, LSTMTwoSum without */
nbo_set *ref_lstm_sum(const nbo_set *a, const nbo_set *b, const nbo_set *c, int activation) {
    Set<FFMatrixBlock> as = ff_blocks(a), bs = ff_blocks(b), out;
    if (c) {
        Set<FFMatrixBlock> cs = ff_blocks(c);
        LSTMThreeWaySum j(activation == 2 ? SumActivation::Tanh : SumActivation::Sigmod);
        out = run_join3<FFMatrixBlock>(j, as, bs, cs);
    } else {
        LSTMTwoSum j;
        out = run_join<FFMatrixBlock>(j, as, bs);
    }
    return to_nbo(out);
}
/* LASillyMultiply1Join | LASillyTransposeMultiply1Join, followed by LASillyMultiply2Aggregate (MatrixBlock) */
nbo_set *ref_la_multiply_agg(int transpose_left, const nbo_set *a, const nbo_set *b) {
    Set<MatrixBlock> as = la_blocks(a), bs = la_blocks(b), joined;
    if (transpose_left) { LASillyTransposeMultiply1Join j; joined = run_join<MatrixBlock>(j, as, bs); }
    else { LASillyMultiply1Join j; joined = run_join<MatrixBlock>(j, as, bs); }
    LASillyMultiply2Aggregate g;
    Set<MatrixBlock> out = run_agg<MatrixBlock, MatrixMeta, MatrixData>(g, joined);
    return to_nbo(out);
}

/* image_ops: ImageToChunks -> ImageChunksToBlock -> ImageBlockToMatrix (PipelinedConv2dMemFuseTest.cc:217-236) */
nbo_set *ref_image_flat(const double *images, int n, int c, int h, int w, int k, int stride, int padding, int block_x, int block_y) {
    using namespace conv2d_memory_fusion;
    Set<Image> imgs = records<Image>(images, n, c, h, w);
    ImageToChunks tc(block_y, k, stride, padding);
    Set<ImageChunk> chunks = run_multiselect<ImageChunk>(tc, imgs);
    ImageChunksToBlock tb(block_x);
    Set<ImageBlock> blocks = run_agg<ImageBlock, long, ImageBlock>(tb, chunks);
    const int windows = imgs[0]->get_conv2d_matrix_rows(k, stride, padding);
    ImageBlockToMatrix tm(block_x, block_y, true, n * windows, c * k * k + 1);
    Set<FFMatrixBlock> out = run_select<FFMatrixBlock>(tm, blocks);
    return to_nbo(out);
}
/* kernel_bias_join (PipelinedConv2dMemFuseTest.cc:186-214): with_join = 0 stops before KernelBiasJoin */
nbo_set *ref_kernel_flat(const double *kernels, const nbo_set *bias, int kout, int c, int k, int block_x, int block_y, int with_join) {
    using namespace conv2d_memory_fusion;
    Set<Kernel> ks = records<Kernel>(kernels, kout, c, k, k);
    KernelToChunks tc(block_y);
    Set<ImageChunk> chunks = run_multiselect<ImageChunk>(tc, ks);
    ImageChunksToBlock tb(block_x);
    Set<ImageBlock> blocks = run_agg<ImageBlock, long, ImageBlock>(tb, chunks);
    ImageBlockToMatrix tm(block_x, block_y, true, kout, c * k * k + 1);
    Set<FFMatrixBlock> flat = run_select<FFMatrixBlock>(tm, blocks);
    if (!with_join) return to_nbo(flat);
    Set<FFMatrixBlock> bs = ff_blocks(bias);
    KernelBiasJoin j;
    Set<FFMatrixBlock> out = run_join<FFMatrixBlock>(j, flat, bs);
    return to_nbo(out);
}
/* ConvResultToChunks -> ImageChunksToBlock -> ConvChunksToImage (the commented-out job of :276-296 and
 * Conv2dMemFuseTest.cc); out: n x z x (x*y), channels the pipeline left empty stay 0 */
int ref_conv_result_to_images(const nbo_set *result, int n, int x, int y, int z, double *out) {
    using namespace conv2d_memory_fusion;
    Set<FFMatrixBlock> rs = ff_blocks(result);
    ConvResultToChunks tc(n);
    Set<ImageChunk> chunks = run_multiselect<ImageChunk>(tc, rs);
    ImageChunksToBlock tb(x * y);
    Set<ImageBlock> blocks = run_agg<ImageBlock, long, ImageBlock>(tb, chunks);
    ConvChunksToImage ti(x, y, z);
    Set<Image> imgs = run_agg<Image, long, Image>(ti, blocks);
    memset(out, 0, (size_t)n * z * x * y * sizeof(double));
    for (auto &im : imgs) {
        const int i = im->getKey();
        if (i < 0 || i >= n) return -1;
        for (int c = 0; c < z; c++) {
            Handle<FFMatrixBlock> m = im->getMatrixAtIndex(c);
            if (m == nullptr) continue;
            memcpy(out + ((size_t)i * z + c) * x * y, m->getValue().rawData->c_ptr(), (size_t)x * y * sizeof(double));
        }
    }
    return static_cast<int>(imgs.size());
}

/* FullyConnectedNetwork (FCProjTest.cc:88-97): constructed as the driver does (random weights), run over the set; the
 * weights it drew are copied out (w0: hidden x features, b0, w1: labels x hidden, b1) */
nbo_set *ref_fully_connected(const nbo_set *in, int features, int labels, int hidden, int batch, double *w0, double *b0, double *w1,
                             double *b1) {
    FullyConnectedNetwork net(features, labels, hidden, (uint32_t)batch);
    memcpy(w0, net.weights1->c_ptr(), (size_t)features * hidden * sizeof(double));
    memcpy(b0, net.bias1->c_ptr(), (size_t)hidden * sizeof(double));
    memcpy(w1, net.weights2->c_ptr(), (size_t)hidden * labels * sizeof(double));
    memcpy(b1, net.bias2->c_ptr(), (size_t)labels * sizeof(double));
    Set<FFMatrixBlock> xs = ff_blocks(in);
    Set<FFMatrixBlock> out = run_select<FFMatrixBlock>(net, xs);
    return to_nbo(out);
}
/* SemanticClassifier (TestSemanticClassifier.cc): weights and biases are the all-ones vectors the projection builds */
nbo_set *ref_semantic_classifier(const nbo_set *in, int size_embed, int dense0, int dense1) {
    SemanticClassifier net((uint32_t)size_embed, (uint32_t)dense0, (uint32_t)dense1);
    Set<FFMatrixBlock> xs = ff_blocks(in);
    Set<FFMatrixBlock> out = run_select<FFMatrixBlock>(net, xs);
    return to_nbo(out);
}

}  /* extern "C" */
