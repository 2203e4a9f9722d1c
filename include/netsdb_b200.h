/*
 * netsdb_b200.h -- C ABI of libnetsdb_b200.so, the B200 (sm_100a) implementation of netsDB's
 * in-database DNN inference hot path (src/FF, src/conv2d_memory_fusion, word2vec block join).
 *
 * Plain C types only: pointers, sizes, ints.  All host pointers are caller-owned and are never
 * retained after a call returns.  Every function returns 0 (NB200_OK) on success or an
 * nb200_status code; nb200_last_error() returns a thread-local description of the last failure,
 * which the host wrappers turn into the reference's `false + errMsg` / print-and-exit(1)
 * convention.
 *
 * There is no CPU fallback: if no sm_100 device is present nb200_init fails.
 *
 * Threads and memory: any number of devices per process (nb200_init / nb200_init_devices after the worker's fork(); a
 * thread picks its device with nb200_use_device).  Entry points may be called from any number of threads -- netsDB runs one
 * pipeline thread per hash partition --; calls onto the same device are served one at a time (a per-device lock: every call
 * already fills the device), calls onto different devices run concurrently.  Device memory comes from a caching allocator inside the
 * library; the fused stages (nb200_ff_inference_unit, nb200_lstm_cell) allocate nothing in steady state.  Uploads are
 * asynchronous: a source block must stay unmodified until nb200_synchronize() or the next call that returns data to the
 * host (see nb200_pin_pool / nb200_upload_commit).
 *
 * Paths in the citations are relative to the netsDB tree (venkate5hgunda/netsdb).
 */
#ifndef NETSDB_B200_H
#define NETSDB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    NB200_OK = 0,
    NB200_ERR_INVALID = 1, /* bad argument                                            */
    NB200_ERR_CUDA = 2,    /* a CUDA runtime/driver call failed                       */
    NB200_ERR_SHAPE = 3,   /* "Block dimemsions mismatch!" (FFTransposeMult.h:61-65)  */
    NB200_ERR_NOMEM = 4,
    NB200_ERR_STATE = 5,   /* not initialised / set missing a needed representation   */
    NB200_ERR_KERNEL = 6   /* a kernel reported an internal fault (e.g. pipeline wait timed out) */
} nb200_status;

/* Join flavour: which reference JoinComp the block GEMM replaces. */
typedef enum {
    NB200_OP_NT = 0, /* FFTransposeMult  (FFTransposeMult.h:38-39,80-92): C(i,j) = sum_k A(i,k) B(j,k)^T */
    NB200_OP_NN = 1, /* FFInputLayerJoin (FFInputLayerJoin.h:30-31,72-86): C(i,j) = sum_k A(i,k) B(k,j); also the LA DSL's
                        LASillyMultiply1Join (sharedLibraries/headers/LASillyMultiply1Join.h:24-70) */
    NB200_OP_TN = 2  /* LASillyTransposeMultiply1Join (LASillyTransposeMultiply1Join.h:24-70): predicate
                        in1.blockRow == in2.blockRow, C(i,j) = sum_k A(k,i) B(k,j); no epilogue */
} nb200_op;

/* What is fused after the aggregation. */
typedef enum {
    NB200_EPI_NONE = 0,
    NB200_EPI_BIAS_RELU = 1,     /* FFReluBiasSum (FFReluBiasSum.h:63-71): max(0, C[i,j] + bias[i]), dropout 0 */
    NB200_EPI_BIAS_TRANSPOSE = 2 /* FFTransposeBiasSum intended form (FFTransposeBiasSum.h:64-104):
                                    out = (C[i,j] + bias[i])^T; the exp() is applied by nb200_softmax_rows /
                                    on download of the `yo` view so that the device never overflows */
} nb200_epilogue;

/* Activation of nb200_sum_blocks = enum class SumActivation of the reference (LSTMThreeWaySum.h:12) plus "none". */
typedef enum {
    NB200_ACT_NONE = 0,    /* LSTMTwoSum (LSTMTwoSum.h:58-63): plain sum */
    NB200_ACT_SIGMOID = 1, /* SumActivation::Sigmod: 1 / (1 + exp(-s))          (LSTMThreeWaySum.h:81)    */
    NB200_ACT_TANH = 2     /* SumActivation::Tanh: (e^s - e^-s) / (e^s + e^-s)  (LSTMThreeWaySum.h:85-86) */
} nb200_activation;

/* Arithmetic of the block GEMM.  The reference is fp64 end to end (FFMatrixData.h:35-36). */
typedef enum {
    NB200_PREC_BF16 = 0,   /* one tcgen05 kind::f16 pass, bf16 operands, fp32 TMEM accumulation        */
    NB200_PREC_BF16X3 = 1, /* bf16 hi/lo split, 3 passes (hi*hi + hi*lo + lo*hi), ~fp32-grade products   */
    NB200_PREC_FP64 = 2,   /* exact-path DFMA kernel (also used to re-evaluate near-tie rows)            */
    NB200_PREC_F16 = 3     /* one tcgen05 kind::f16 pass on fp16 operands with per-matrix power-of-two scales:
                              3 more mantissa bits than BF16 at the same rate, so far fewer undecided rows     */
} nb200_precision;

/* Opaque handle on a device-resident user set of FFMatrixBlocks (db, set) -- the counterpart of
 * what ScanUserSet<FFMatrixBlock>(db, set) scans (builtInPDBObjects/headers/ScanUserSet.h:53-58).
 * On the device the blocks of a set are de-blocked into one dense matrix. */
typedef struct nb200_set nb200_set;

/* Counters filled by nb200_ff_inference_unit. */
typedef struct {
    float ms_convert;     /* fp64 -> bf16 plane conversion of `inputs`              */
    float ms_gemm1;       /* layer-1 block GEMM incl. split-K reduction + bias/relu */
    float ms_gemm2;       /* layer-2 block GEMM + bias                              */
    float ms_softmax;     /* row softmax / argmax / near-tie scan                   */
    float ms_refine;      /* fp64 re-evaluation of near-tie rows                    */
    float ms_total;
    float ms_gemm1_kernel; /* the tcgen05 kernel of layer 1 alone                   */
    int refined_rows;     /* rows re-evaluated in fp64                              */
    int relevel_rows;     /* rows re-run in bf16x3 by the single-pass variant       */
    int kernel_launches;  /* kernels launched by this call                          */
    int splits1;          /* split-K factor used by layer 1                         */
} nb200_ff_stats;

/* ---- lifetime ------------------------------------------------------------------------------ */
/* Create the library's context on CUDA device `device` (call after fork(), cf. mainServer/source/WorkerMain.cc:107) and make
 * the CALLING THREAD work on it.  Idempotent; may be called for several devices.
 *
 * One process can drive all GPUs of a box: netsDB's backend runs `numThreads` pipeline threads in one process
 * (conf/headers/Configuration.h:52-53, work/source/PDBWorkerQueue.cc:163).  Every device has its own context, sets (the
 * (db, set) namespace is per device), scratch, allocator pool and lock; a thread works on the device it chose last
 * (nb200_init / nb200_use_device; thread-local), and threads on different devices never wait for one another.  A set can
 * only be passed to entry points by a thread that works on the set's device.  For config C3 each thread uploads its row
 * shard of `inputs` (nb200_upload_rect) and the replicated weights to its own device and calls nb200_ff_inference_unit. */
int nb200_init(int device);
/* nb200_init for each of devices[0..n), then nb200_use_device(devices[0]). */
int nb200_init_devices(const int *devices, int n);
/* The calling thread works on `device` (already initialised) from now on. */
int nb200_use_device(int device);
int nb200_current_device(int *device);
int nb200_shutdown(void);
const char *nb200_last_error(void);
int nb200_device_count(int *count);
/* Launch everything on `cuda_stream` (a cudaStream_t) instead of the library's own stream. */
int nb200_set_stream(void *cuda_stream);
int nb200_synchronize(void);
/* Tuning switches (process wide).
 *  "fused_layer1": how layer 1 of nb200_ff_inference_unit(NB200_PREC_F16) runs on long feature vectors.  2 (default): the
 *      cluster kernel that converts the fp64 inputs inside the CTAs that consume them and splits K in waves, so that any
 *      number of rows (a row shard of 1,250 as well as 10,000) fills the device; 1: the round-1 kernel that passes the
 *      converted inputs through a ring in L2 (whole 256-row tiles only); 0: separate conversion and GEMM kernels;
 *      -1: back to the default.  Shapes a kernel does not cover fall through to the next one.
 *  "near_tie_tau_x10" (default 25): width of the near-tie band in tenths of eps_r, the norm-based estimate of a row's
 *      low-precision logit error.  Rows whose top-2 logit gap is below tau * eps_r are re-evaluated
 *      at the next precision level.  Measured on 102,400 rows (tests/test_gpu_fullsize.py): every raw fp16 label that
 *      differs from fp64 has gap < 0.77 eps_r, so 2.5 leaves a factor of three; 40 restores round 1's band.
 *  "l1_stream_mt": cluster shape of the default kernel, 1 (4 CTAs, 256 rows) or 2 (8 CTAs, 512 rows, W1 multicast); 0
 *      (default) picks by the padding the row count would need. */
int nb200_set_option(const char *key, int value);
/* Number of kernels this library has launched since nb200_init (monotonic). */
int64_t nb200_launch_count(void);

/* Page-lock a host range so block uploads/downloads run as async DMA -- the shared-memory
 * buffer pool of memory/source/SharedMem.cc:128 or any client allocation block. */
int nb200_pin_pool(void *base, size_t bytes);
int nb200_unpin_pool(void *base);

/* ---- sets ------------------------------------------------------------------------------------ */
/* PDBClient::createSet<FFMatrixBlock>(db, set, ...) (mainClient/headers/PDBClient.h:74-259).
 * total_* are FFMatrixMeta::totalRows/totalCols, block_* the nominal FFMatrixData::rowNums/colNums. */
int nb200_set_create(const char *db, const char *set, int total_rows, int total_cols, int block_rows,
                     int block_cols, nb200_set **out);
int nb200_set_lookup(const char *db, const char *set, nb200_set **out);
int nb200_set_destroy(nb200_set *s); /* PDBClient::removeSet */
int nb200_set_info(const nb200_set *s, int *total_rows, int *total_cols, int *block_rows, int *block_cols);

/* PDBClient::sendData<FFMatrixBlock> for one block: `host` is FFMatrixData::rawData->c_ptr(),
 * row-major rows x cols (full-size zero-padded edge blocks as ff::loadMatrix builds them,
 * FFMatrixUtil.cc:245-302, or trimmed ones).  Padding beyond totalRows/totalCols is dropped. */
int nb200_upload_block(nb200_set *s, int block_row, int block_col, const double *host, int rows, int cols);
/* Marks the end of the uploads queued so far into `set`.  Uploads are asynchronous (two copy streams); an entry point that
 * reads a set first makes the compute stream wait for them -- for ALL queued uploads by default.  After a commit,
 * nb200_ff_inference_unit and nb200_download_* wait only for the uploads of the sets they read, so a caller can queue the
 * next batch into a second set and have its transfer run under the computation of the current one (netsDB's pipeline
 * does the same with its page-circular buffer, storage/headers/PageCircularBuffer.h).  A later upload into the set
 * withdraws the commit. */
int nb200_upload_commit(nb200_set *set);
/* Arbitrary rectangle (row-granular sharding across GPUs): element (r,c) of the rectangle is
 * host[r*ld + c] and lands at (row0 + r, col0 + c) of the set. */
int nb200_upload_rect(nb200_set *s, int row0, int col0, const double *host, int rows, int cols, int64_t ld);
/* getSetIterator<FFMatrixBlock> for one block: fills a rows x cols row-major block, zero padded
 * outside the set's totals (queries/headers/SetIterator.h:12-68). */
int nb200_download_block(nb200_set *s, int block_row, int block_col, double *host, int rows, int cols);
int nb200_download_rect(nb200_set *s, int row0, int col0, double *host, int rows, int cols, int64_t ld);
/* Seeded device-side stand-in for ff::loadMatrix's value distribution (FFMatrixUtil.cc:229-237):
 * element (r,c) = synth(seed, (row_offset + r) * total_cols + c).  Bit-identical to the CPU oracle's
 * generator.  row_offset lets a row shard reproduce its slice of a larger matrix. */
int nb200_set_fill_synth(nb200_set *s, uint64_t seed, int64_t row_offset);
/* A read-only view of `src` under another name whose downloads apply exp() inside the un-padded region --
 * the reference's `yo` (FFTransposeBiasSum.h:94-104) over a set of pre-exp logits. */
int nb200_set_exp_view(nb200_set *src, const char *db, const char *set, nb200_set **out);
/* Re-register a set under another (db, set) name; an existing set of that name is replaced. */
int nb200_set_rename(nb200_set *s, const char *db, const char *set);
/* Drop the derived bf16 operand planes of a set (they are rebuilt from the fp64 data on next use). */
int nb200_set_invalidate_planes(nb200_set *s);

/* ---- fp16 side format ("a binary/bf16 side-format to cut the fp64 transfer and conversion cost") ----
 * An input set can be kept, next to its fp64 FFMatrixBlocks, as what the tensor cores consume: one fp16 operand plane
 * (row-major, row pitch ldk = total_cols rounded up to a multiple of 64, values x * 2^exp), the exponent and the per-row
 * sums of squares of the error model.  nb200_set_export_f16 produces it on the device from a resident fp64 set (done once,
 * at load time) and writes it to host memory; nb200_set_import_f16 brings a set into HBM from that format alone -- a
 * quarter of the bytes over PCIe and no conversion -- asynchronously on the upload streams (nb200_upload_commit applies).
 * nb200_ff_inference_unit(NB200_PREC_F16) accepts such a set as `inputs`; for the label-exact cascade it reads the fp64
 * rows of the few undecided inputs straight out of the host blocks (zero-copy over PCIe), so `host_blocks` -- the
 * row-major grid of pointers to the set's fp64 blocks, ceil(rows / block_rows) x ceil(cols / block_cols) entries -- must
 * point into pinned, device-visible memory (nb200_pin_pool / cudaHostAlloc) and stay valid while the set is in use.
 * edge_ld = row pitch of the blocks in the last block column (block_cols for zero-padded blocks, the remaining columns
 * for trimmed ones).  The caller's plane, row_sumsq and pointer arrays must stay unmodified until nb200_synchronize()
 * or the next call that returns data to the host. */
int nb200_set_export_f16(nb200_set *set, void *plane_host, int64_t ldk, int *exp_out, float *row_sumsq_host);
int nb200_set_import_f16(nb200_set *set, const void *plane_host, int64_t ldk, int exp, const float *row_sumsq_host,
                         const double *const *host_blocks, int edge_ld);

/* ---- operators ------------------------------------------------------------------------------- */
/* JoinComp + ClusterAggregateComp as ONE block GEMM: FFTransposeMult|FFInputLayerJoin followed by
 * FFAggMatrix (FFAggMatrix.h:11-35; ClusterAggregateComp.h:40-344), optionally with the next
 * bias join fused.  The K-block partial products are accumulated inside the kernel (fp32 TMEM /
 * split-K workspace), never materialised per pair.  Result set (db, out_set) is created (or
 * replaced); its meta follow the reference (FFTransposeMult.h:68-71, FFInputLayerJoin.h:64-67,
 * FFTransposeBiasSum.h:64-67).  `bias` must be (rows of C) x 1 (FFReluBiasSum.h:49-56). */
int nb200_gemm_blocks(int op, nb200_set *a, nb200_set *b, int epilogue, nb200_set *bias, int precision,
                      const char *db, const char *out_set, nb200_set **out);

/* FullyConnectedNetwork (src/FF_proj/headers/FullyConnectedNetwork.h:62-123; driver src/tests/source/FCProjTest.cc:88-97)
 * and SemanticClassifier (src/word2vec/headers/SemanticClassifier.h:43-141): a SelectionComp<FFMatrixBlock, FFMatrixBlock>
 * whose projection runs a two-layer network on every block of the input set.  The block's raw buffer (block_rows x
 * block_cols, zero-padded at the set's edges) is viewed as a row-major features x batch matrix x0 -- whatever the block's
 * own shape; it must hold at least features * batch values -- and
 *     y0 = relu(w0 x0 + b0)                   w0: hidden x features, row-major, b0: hidden
 *     out = outLabel(sigmoid(w1 y0 + b1))     w1: labels x hidden, b1: labels; 1 where 1 / (1 + exp(-z)) > 0.5, else 0
 * is written as a labels x batch block with the input block's (blockRow, blockCol): the output set has
 * ceil(rows / block_rows) * labels rows and ceil(cols / block_cols) * batch columns in labels x batch blocks.  (The
 * reference stamps each output block's totalRows / totalCols with labels / batch (SemanticClassifier) or leaves them at
 * -1 (FullyConnectedNetwork); a set here has one pair of totals.)  FullyConnectedNetwork draws its weights from
 * std::random_device in its constructor and SemanticClassifier uses all-ones: both are host arrays here.  Evaluated in
 * fp64 (tensor-core DMMA): the output is a thresholded sign.  The input may be an fp64 set or the fp32 result of a
 * tensor-core GEMM (the FFAggMatrix below SemanticClassifier in TestSemanticClassifier.cc:108-124). */
int nb200_udf_dense2(nb200_set *in, int features, int batch, int hidden, int labels, const double *w0_host, const double *b0_host,
                     const double *w1_host, const double *b1_host, const char *db, const char *out_set, nb200_set **out);

/* LSTMThreeWaySum / LSTMTwoSum (src/LSTM/headers/LSTMThreeWaySum.h:39-95, LSTMTwoSum.h:24-64): the join on equal
 * (blockRow, blockCol) keys followed by out = act(a + b + c) per element, for whole sets at once.  b and c may be
 * NULL (c NULL: LSTMTwoSum; both NULL: the activation mapped over a, which is the tanh LSTMHiddenState applies to its
 * second input, LSTMHiddenState.h:44-47 -- LSTMHiddenState + FFAggMatrix is this call followed by
 * nb200_gemm_blocks(NB200_OP_NN, o_t, tanh_c_t, ...)).  Shapes and block sizes of all inputs must be equal
 * ("Block dimemsions mismatch!", LSTMThreeWaySum.h:58-65).  Inputs may be fp64 sets or fp32 results of the
 * tensor-core GEMM; the sum and the activation are evaluated in fp64 and the result set holds fp64.  Out meta = a's
 * (LSTMThreeWaySum.h:67-70). */
int nb200_sum_blocks(nb200_set *a, nb200_set *b, nb200_set *c, int activation, const char *db, const char *out_set,
                     nb200_set **out);

/* One LSTM step as the reference's driver wires it (src/tests/source/LSTMTest.cc:204-543), in one call:
 *   f_t = sigmoid(w_f.x_t + u_f.h_t_1 + b_f), i_t and o_t alike, c_t_temp = tanh(w_c.x_t + u_c.h_t_1 + b_c)
 *       (FFInputLayerJoin + FFAggMatrix twice, LSTMThreeWaySum; :204-451)
 *   c_t = f_t . c_t_1 + i_t . c_t_temp     (matrix products through FFInputLayerJoin, LSTMTwoSum; :470-498)
 *   h_t = o_t . tanh(c_t)                  (LSTMHiddenState + FFAggMatrix; :528-543)
 * Reads the 15 sets of `in`, creates (or replaces) sets f_t, i_t, o_t, c_t_temp, c_t, h_t in `db` (the driver's
 * names), all fp64.  With BF16 / BF16X3 the eight gate products run as ONE tensor-core GEMM over K-concatenated
 * operands ([w_f u_f; w_i u_i; w_o u_o; w_c u_c] . [x_t; h_t_1], the stacked weight operand is cached across steps
 * until a weight set changes), the four LSTMThreeWaySum joins as one pass, and c_t's two products as one GEMM
 * ([f_t | i_t] . [c_t_1; c_t_temp]); no device memory is allocated in steady state.  FP64 and F16 compose the
 * per-operator entry points above.  Gate order in the arrays: forget, input, output, cell candidate. */
typedef struct nb200_lstm_sets {
    nb200_set *w[4], *u[4], *b[4];
    nb200_set *x_t, *h_t_1, *c_t_1;
} nb200_lstm_sets;
int nb200_lstm_cell(const nb200_lstm_sets *in, int precision, const char *db);

/* FFRowAggregate + FFOutputLayer (FFRowAggregate.h:11-29, FFMatrixBlock.h:116-142,
 * FFOutputLayer.h:31-57 intended form) on a set of pre-exp logits (batch x labels):
 * out[i,j] = exp(z[i,j]) / sum_j exp(z[i,j]), evaluated in the max-subtracted form.
 * labels_host (optional, length = rows) receives the first row argmax of z. */
int nb200_softmax_rows(nb200_set *logits, const char *db, const char *out_set, nb200_set **out,
                       int32_t *labels_host);

/* The whole of ff::inference_unit (FF/source/SimpleFF.cc:331-424) as one fused stage:
 * y1 = relu(w1 . inputs^T + b1); z = (wo . y1 + bo)^T; output = softmax_rows(z).
 * Creates sets (db, "yo") [a view: exp(z)] and (db, out_set).  With a tensor-core precision, rows
 * whose top-2 logit gap is inside the low-precision error band are re-evaluated in fp64 so that
 * labels equal the fp64 pipeline's.  labels_host (optional) gets one int32 per input row. */
int nb200_ff_inference_unit(nb200_set *w1, nb200_set *wo, nb200_set *inputs, nb200_set *b1, nb200_set *bo,
                            int precision, const char *db, const char *out_set, nb200_set **out,
                            int32_t *labels_host, nb200_ff_stats *stats);

/* ---- K-sharded variant (weights larger than one GPU's HBM) ---------------------------------------------
 * The reference hash-partitions blocks by blockColIndex (FFMatrixBlock.h:151-156), so each worker multiplies a
 * slice of K and the aggregation sums the partials across workers over TCP (PipelineStage.cc:120-205).  Here each
 * GPU holds a column shard of `inputs` and `w1`; nb200_ff_layer1_partial leaves that shard's fp32 partial of
 * w1 . inputs^T (stored batch x hidden, row-major, pitch ld, NO bias) in a caller-owned DEVICE buffer, the
 * caller sums the partials across GPUs with one NCCL reduce-scatter / all-reduce over NVLink, and
 * nb200_ff_finish_from_layer1 applies FFReluBiasSum and runs layer 2 and the softmax on the rows this GPU owns.
 * The label-exact cascade is not available in this variant (the fp64 rows are sharded too); the number of
 * undecided rows is returned so that the caller knows how many labels rest on the tensor-core precision.
 * For that estimate the partial buffer may carry one extra column (ld > hidden): column `hidden` receives the
 * shard's per-row sum of squares of the inputs, which the same reduction turns into the global one, and the
 * caller passes the maximum of nb200_set_maxabs(w1_shard) over all ranks as w1_maxabs (0 = skip the estimate). */
int nb200_ff_layer1_partial(nb200_set *w1_shard, nb200_set *inputs_shard, int precision, float *partial_dev, int64_t ld);
int nb200_ff_finish_from_layer1(const float *y1pre_dev, int64_t ld, int rows, nb200_set *b1, nb200_set *wo, nb200_set *bo,
                                int precision, float w1_maxabs, const char *db, const char *out_set, nb200_set **out,
                                int32_t *labels_host, int *undecided_rows);
/* max |x| over a set (computed with its operand planes). */
int nb200_set_maxabs(nb200_set *s, float *out);

/* ---- conv2d_memory_fusion ---------------------------------------------------------------------- */
/* Image set -> `image_flat` (Image.h:26-57, ImageToChunks.h:43-102, ImageChunksToBlock.h:35-50,
 * ImageBlockToMatrix.h:46-73): rows img*windows+win, cols ch*k*k+i*k+j plus a trailing 1.0 column.
 * images_host: n x c x h x w row-major fp64. */
int nb200_conv2d_image_flat(const double *images_host, int n, int c, int h, int w, int k, int stride, int padding,
                            int block_rows, int block_cols, const char *db, const char *out_set, nb200_set **out);
/* Kernel set (+ bias) -> `kernel_flat` (KernelToChunks.h:39-83, KernelBiasJoin.h:42-46). */
int nb200_conv2d_kernel_flat(const double *kernels_host, const double *bias_host, int kout, int c, int k,
                             int block_rows, int block_cols, const char *db, const char *out_set, nb200_set **out);
/* The `conv2d` job of PipelinedConv2dMemFuseTest.cc:249-273 (FFTransposeMult + FFAggMatrix on
 * image_flat x kernel_flat) without materialising image_flat: implicit-GEMM over the image.
 * result[(img, ho, wo), cout] = conv + bias, optional ReLU (config C4).  Stride 1. */
int nb200_conv2d_fused(const double *images_host, int n, int c, int h, int w, const double *kernels_host,
                       const double *bias_host, int kout, int k, int padding, int relu, int precision,
                       int block_rows, int block_cols, const char *db, const char *out_set, nb200_set **out);

/* The `img` set of the reference's driver (a set of Image records, conv2d_memory_fusion/headers/Image.h;
 * PipelinedConv2dMemFuseTest.cc:99-107 loads it, the `image_ops` job scans it) resident on the device: NCHW fp64 as sent,
 * plus the NHWC operand planes of the implicit GEMM, derived on first use and kept with the set. */
int nb200_conv2d_images_create(const double *images_host, int n, int c, int h, int w, const char *db, const char *set,
                               nb200_set **out);
/* nb200_conv2d_fused on a resident image set: nothing crosses PCIe but the kernels and the bias.  kernel_ms (optional)
 * receives the device time of the implicit-GEMM launch. */
int nb200_conv2d_fused_from_set(nb200_set *images, const double *kernels_host, const double *bias_host, int kout, int k,
                                int padding, int relu, int precision, int block_rows, int block_cols, const char *db,
                                const char *out_set, nb200_set **out, float *kernel_ms);

/* ConvResultToChunks + ConvChunksToImage (ConvResultToChunks.h:42-103, ConvChunksToImage.h:34-72; the last two
 * jobs of Conv2dMemFuseTest.cc): re-assemble the `result` matrix [(image, window), channel] into num_images images of
 * z channels, each an x * y matrix in window order.  images_host receives num_images * z * x * y doubles (NCHW).
 * The result set's rows must divide evenly among the images (the reference asserts it, :46) and x * y must equal
 * that quotient; channels beyond the set's columns (z > total columns) are zero. */
int nb200_conv2d_result_to_images(nb200_set *result, int num_images, int x, int y, int z, double *images_host);

/* ---- allocation-free variant ----------------------------------------------------------------------
 * This entry point has no counterpart in the reference.
 *
 * C[I x J] = A[I x K] . B[J x K]^T -- the FFTransposeMult + FFAggMatrix product -- for callers that are
 * not allowed to allocate device memory: a pipeline thread that lives inside an HBM arena the engine's
 * buffer pool sized up front, a call site between cudaStreamBeginCapture/EndCapture, or a serving
 * thread that cannot afford cudaMalloc's device-wide synchronisation.  The regular entry points grow a
 * split-K workspace and operand planes with cudaMalloc on demand; this one keeps that scratch in a
 * fixed array inside the library image instead and calls no cudaMalloc, cudaFree or allocating library
 * routine.  It does not need nb200_init.
 *
 * a_dev, b_dev, c_dev are DEVICE pointers to row-major fp64 matrices (pitches lda, ldb, ldc in
 * elements) on the current device; the work is queued on cuda_stream (a cudaStream_t, NULL = default
 * stream) and the call returns without waiting.  Arithmetic is the single-pass bf16 tensor-core block
 * GEMM (same as NB200_PREC_BF16, without the near-tie cascade).
 *
 * Limits and rules:
 *  - scratch needed = bf16 copies of A and B (K rounded up to 64) + split-K partial tiles; it must fit
 *    NB200_NOALLOC_SCRATCH_BYTES (64 MiB, held by every GPU that loads the library until exit).  Larger
 *    requests are refused with NB200_ERR_INVALID and a message stating need and limit.
 *  - the first call on a device looks up the scratch array's device address and sets kernel attributes;
 *    make that first call before any stream capture.
 *  - all calls share the one array, which still holds the previous call's data: calls whose GPU work
 *    can overlap (different streams, different threads) race and must be serialised by the caller.
 *  - the call returns before the work has run, so a fault inside the kernels (a pipeline barrier that timed
 *    out) cannot come back as a status: in that case every element of C is NaN once the stream has drained. */
#define NB200_NOALLOC_SCRATCH_BYTES (64u << 20)
int nb200_transpose_mult_noalloc(const double *a_dev, int64_t lda, const double *b_dev, int64_t ldb, int I, int J,
                                 int K, double *c_dev, int64_t ldc, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
