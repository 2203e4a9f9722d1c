/*
 * ff_oracle.h -- CPU restatement (plain C, fp64) of netsDB's block-partitioned FF inference
 * path, the conv2d_memory_fusion im2col path and the word2vec block join.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (libnetsdb_b200.so) never links, loads or calls anything in oracle/.
 *
 * PARITY PINNED AGAINST THE REFERENCE'S OWN UDF CODE (not against a reference binary): the
 * reference ships no golden vectors, known-answer tests or fixtures for this path (SURVEY.md
 * section 4) and its build cannot run here (SCons-generated object model, Eigen 3 absent).  What
 * does compile is the part that carries the arithmetic: netsDB's UDF headers, unmodified, from
 * where they lie under /root/reference, around a stand-in object model and a plain-loop stand-in
 * for the five Eigen expressions they use (oracle/ref/, `make -C oracle ref` ->
 * oracle/_ref/libff_ref.so).  tests/test_oracle_pinned.py runs every operator of this file and
 * the reference's own lambda bodies on the same blocks and compares keys, meta, padding and
 * values (exact, or 1e-13 where a product's summation order differs); tests/golden/ref_pinned.npz
 * holds vectors recorded from that code for machines without the reference tree.  Not pinned:
 * Eigen's own dgemm summation order (third-party, unpinned version) and anything above the UDFs
 * (planner, shuffle, multi-threaded arrival order).
 *
 * Each function cites the reference file:line it follows (paths relative to the netsDB tree).
 * Where this fork's debugging edits changed results the
 * oracle implements the INTENDED semantics; debug prints (Q5) are omitted; dropout is 0 (Q3).
 */
#ifndef NB200_FF_ORACLE_H
#define NB200_FF_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One FFMatrixBlock: src/FF/headers/FFMatrixBlock.h:18-66, FFMatrixData.h:11-36,
 * FFMatrixMeta.h:7-24.  data is row-major rows x cols (FFMatrixData::rawData). */
typedef struct {
    int block_row, block_col;   /* FFMatrixMeta::blockRowIndex / blockColIndex */
    int rows, cols;             /* FFMatrixData::rowNums / colNums             */
    int total_rows, total_cols; /* FFMatrixMeta::totalRows / totalCols         */
    double *data;
} nbo_block;

/* A user set of FFMatrixBlocks (what ScanUserSet<FFMatrixBlock> iterates). */
typedef struct {
    int n, cap;
    nbo_block *blocks;
} nbo_set;

/* ---- generator ---------------------------------------------------------------------------
 * Seeded stand-in for ff::loadMatrix's value distribution (src/FF/source/FFMatrixUtil.cc:
 * 229-237, 280-282): sign Bernoulli(1/2), magnitude U(1e-4, 0.5).  Counter based so the CUDA
 * generator reproduces it bit for bit: value(seed, row * total_cols + col). */
double nbo_synth_value(uint64_t seed, uint64_t index);
void nbo_synth_dense(double *out, int64_t rows, int64_t cols, uint64_t seed);

/* ---- sets ------------------------------------------------------------------------------ */
nbo_set *nbo_set_new(void);
/* appends a zero-filled block and returns it */
nbo_block *nbo_set_add_block(nbo_set *s, int block_row, int block_col, int rows, int cols, int total_rows, int total_cols);
void nbo_set_free(nbo_set *s);
int nbo_set_size(const nbo_set *s);
const nbo_block *nbo_set_get(const nbo_set *s, int i);
const nbo_block *nbo_set_find(const nbo_set *s, int block_row, int block_col);

/* ff::loadMatrix blocking and zero padding (FFMatrixUtil.cc:245-302): dense row-major
 * total_rows x total_cols -> ceil() grid of full-size blocks; dont_pad_* as in the reference. */
nbo_set *nbo_load_matrix(const double *dense, int total_rows, int total_cols, int block_rows,
                         int block_cols, int dont_pad_rows, int dont_pad_cols);
/* Same, values straight from the generator (no dense copy). */
nbo_set *nbo_load_matrix_synth(uint64_t seed, int total_rows, int total_cols, int block_rows,
                               int block_cols);
/* Gather a set back into a dense total_rows x total_cols matrix (padding dropped). */
void nbo_set_to_dense(const nbo_set *s, double *dense, int total_rows, int total_cols);

/* ---- FF operators ---------------------------------------------------------------------- */
/* FFTransposeMult (FFTransposeMult.h:38-39 predicate, :58-92 projection) followed by
 * FFAggMatrix (FFAggMatrix.h:20-33; FFMatrixData::operator+ FFMatrixData.h:38-57) with the
 * two-level merge of CombinedShuffleSink.h:78-159 + AggregationProcessor.cc:131-156:
 * out(i,j) = sum_k A(i,k) * B(j,k)^T.  threads = pipeline threads (hash partitions). */
nbo_set *nbo_transpose_mult_agg(const nbo_set *a, const nbo_set *b, int threads);
/* FFInputLayerJoin (FFInputLayerJoin.h:30-31, :51-98) + FFAggMatrix:
 * out(i,j) = sum_k A(i,k) * B(k,j). */
nbo_set *nbo_input_layer_join_agg(const nbo_set *a, const nbo_set *b, int threads);
/* FFReluBiasSum (FFReluBiasSum.h:34-35, :46-71), dropout_rate = 0. */
nbo_set *nbo_relu_bias_sum(const nbo_set *in, const nbo_set *bias);
/* FFTransposeBiasSum, intended semantics (FFTransposeBiasSum.h:64-67 output meta, :84 the
 * surviving comment `currentMatrix1.colwise() + currentMatrix2`, :94-104 masked exp):
 * out = exp(in + bias[row])^T on the un-padded region, 0 elsewhere.  apply_exp = 0 returns the
 * pre-exp logits (used to define labels when exp overflows). */
nbo_set *nbo_transpose_bias_sum(const nbo_set *in, const nbo_set *bias, int apply_exp);
/* FullyConnectedNetwork (FF_proj/headers/FullyConnectedNetwork.h:62-123) / SemanticClassifier
 * (word2vec/headers/SemanticClassifier.h:43-141): per block, raw buffer viewed as features x batch;
 * out block = outLabel(sigmoid(w1 * relu(w0 * x + b0) + b1)), labels x batch, same block indices. */
nbo_set *nbo_udf_dense2(const nbo_set *in, int features, int batch, int hidden, int labels, const double *w0, const double *b0,
                        const double *w1, const double *b1, int with_totals);
/* FFRowAggregate (FFRowAggregate.h:21-28; FFMatrixBlock.h:116-142). */
nbo_set *nbo_row_aggregate(const nbo_set *in);
/* FFOutputLayer, intended semantics (FFOutputLayer.h:21-22, :31-57: in1[i,j] / in2[i]). */
nbo_set *nbo_output_layer(const nbo_set *in, const nbo_set *rowsum);

/* ff::inference_unit (src/FF/source/SimpleFF.cc:331-424).  Returns the `output` set
 * (softmax probabilities, batch x labels blocks); *logits_out (optional) receives the pre-exp
 * batch x labels set; labels (optional, length batch) = first row argmax of the logits. */
nbo_set *nbo_ff_inference_unit(const nbo_set *w1, const nbo_set *wo, const nbo_set *inputs,
                               const nbo_set *b1, const nbo_set *bo, int threads,
                               nbo_set **logits_out, int32_t *labels);

/* The same two UDFs AS WRITTEN in this fork (quirks Q1, Q2): for pinning against the reference's code and for
 * documentation only; see the comments at their definitions. */
nbo_set *nbo_transpose_bias_sum_as_written(const nbo_set *in, const nbo_set *bias);
nbo_set *nbo_output_layer_as_written(const nbo_set *in, const nbo_set *rowsum);

/* ---- LA DSL multiply joins on MatrixBlock (same block layout) --------------------------- */
/* LASillyMultiply1Join + LASillyMultiply2Aggregate == nbo_input_layer_join_agg (same predicate and product,
 * LASillyMultiply1Join.h:24-33, :44-70).  LASillyTransposeMultiply1Join + LASillyMultiply2Aggregate
 * (LASillyTransposeMultiply1Join.h:24-70): out(i,j) = sum_k A(k,i) * B(k,j). */
nbo_set *nbo_la_transpose_multiply_agg(const nbo_set *a, const nbo_set *b, int threads);

/* ---- LSTM cell operators (src/LSTM/headers, wired by src/tests/source/LSTMTest.cc:204-543) -------------- */
/* LSTMThreeWaySum (LSTMThreeWaySum.h:39-95): out = act(a + b + c) per block with equal (blockRow, blockCol);
 * activation 1 = Sigmod, 2 = Tanh (enum SumActivation, :12), 0 = none.  c == NULL: LSTMTwoSum (LSTMTwoSum.h:24-64).
 * b == c == NULL: the activation mapped over a (the tanh of LSTMHiddenState.h:44-47). */
nbo_set *nbo_lstm_sum(const nbo_set *a, const nbo_set *b, const nbo_set *c, int activation);
/* LSTMHiddenState (LSTMHiddenState.h:24-62) + FFAggMatrix: out(i,j) = sum_k o_t(i,k) * tanh(c_t)(k,j). */
nbo_set *nbo_lstm_hidden_state_agg(const nbo_set *o_t, const nbo_set *c_t, int threads);

/* ---- conv2d_memory_fusion -------------------------------------------------------------- */
/* Image -> image_flat: Image.h:26-57, ImageToChunks.h:43-102, ImageChunksToBlock.h:35-50,
 * ImageBlockToMatrix.h:46-73.  images: n x c x h x w row-major.  Row = img * windows + win,
 * col = ch*k*k + i*k + j, plus the trailing column of 1.0.  Blocks are zero padded
 * (block_padding = true in PipelinedConv2dMemFuseTest.cc:77). */
nbo_set *nbo_image_flat(const double *images, int n, int c, int h, int w, int k, int stride,
                        int padding, int block_rows, int block_cols);
/* Kernel -> kernel_flat with bias: KernelToChunks.h:39-83, KernelBiasJoin.h:42-46 (bias lands
 * in the last real column; Q6: callers keep (c*k*k+1) % block_cols != 0). */
nbo_set *nbo_kernel_flat(const double *kernels, const double *bias, int kout, int c, int k,
                         int block_rows, int block_cols);

/* ---- misc ------------------------------------------------------------------------------ */
/* ConvResultToChunks + ConvChunksToImage (ConvResultToChunks.h:42-103, ConvChunksToImage.h:34-72): result matrix
 * [(image, window), channel] -> num_images x z channels x (x * y) window-ordered values.  0 on success. */
int nbo_conv_result_to_images(const nbo_set *result, int num_images, int x, int y, int z, double *out);

/* plain fp64 GEMM used by the pair products: C(MxN) = A(MxK) * B^T (B is NxK) if b_trans else
 * A(MxK) * B(KxN); all row-major. */
void nbo_dgemm(int m, int n, int k, const double *a, int lda, const double *b, int ldb,
               int b_trans, double *c, int ldc);
int nbo_hw_threads(void);

#ifdef __cplusplus
}
#endif
#endif
