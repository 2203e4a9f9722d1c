// nb200_internal.h -- declarations shared by the translation units of libnetsdb_b200.so.
// Not part of the public ABI (that is include/netsdb_b200.h).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/netsdb_b200.h"

namespace nb200 {

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char *fmt, ...);
#define NB_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            ::nb200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return NB200_ERR_CUDA;                                                                 \
        }                                                                                          \
    } while (0)
#define NB_TRY(expr)                \
    do {                            \
        int _s = (expr);            \
        if (_s != NB200_OK) return _s; \
    } while (0)

// ---- context --------------------------------------------------------------------------------
struct Context {
    int device = -1;
    int sm_count = 0;
    cudaStream_t stream = nullptr;      // stream everything is launched on
    cudaStream_t own_stream = nullptr;  // created by nb200_init
    int *err_flag = nullptr;            // device int: nonzero = kernel-internal fault code
    int *err_flag_host = nullptr;       // pinned mirror
    float *ws = nullptr;                // split-K workspace (grown on demand)
    size_t ws_bytes = 0;
    double *stage = nullptr;            // staging for transposed / converting downloads
    size_t stage_bytes = 0;
    int64_t launches = 0;
    cudaEvent_t ev[8] = {};
    cudaEvent_t time_ev0 = nullptr, time_ev1 = nullptr;   // when set: recorded right around the next timed kernel launch
    // K-chunk pipelining of layer 1: conversion of chunk j+1 runs beside the GEMM of chunk j
    cudaStream_t aux[3] = {};        // [0] conversion, [1],[2] alternating GEMM streams
    cudaEvent_t chunk_ev[64] = {};   // conversion of chunk j done
    cudaEvent_t join_ev[4] = {};
    // block uploads: contiguous H2D into a staging slot, then a device-side de-blocking copy, on two alternating
    // side streams (a pitched 2-D H2D copy reaches 47 GB/s on this platform, a contiguous one 55 GB/s)
    cudaStream_t copy[2] = {};
    double *up_stage[2] = {};
    size_t up_bytes[2] = {};
    cudaEvent_t up_ev[2] = {};
    cudaEvent_t main_ev = nullptr;
    int up_next = 0;
    bool uploads_pending = false;  // work queued on the copy streams that the compute stream has not waited for
    uint64_t up_epoch = 1;         // advanced by every wait on ALL uploads: a set with set->up_epoch == up_epoch has
                                   // uploads the compute stream has not waited for yet
    bool main_dirty = true;        // compute-stream work queued since the copy streams last waited for it
};
int wait_uploads(Context *c);      // make the compute stream wait for every queued upload
int wait_set_uploads(Context *c, nb200_set *s);  // ... for the uploads of one set (its commit fence when it has one)
Context *ctx();
                                   // nullptr before nb200_init
std::recursive_mutex &api_mutex();  // held by every entry point for its whole duration
// caching device allocator (nb200_pool.cu): every device allocation of the library goes through these
cudaError_t pool_malloc(void **p, size_t bytes);
cudaError_t pool_free(void *p);
void pool_release_all();
template <class T> inline cudaError_t nb_malloc(T **p, size_t bytes) { return pool_malloc(reinterpret_cast<void **>(p), bytes); }
inline cudaError_t nb_free(void *p) { return pool_free(p); }
int ensure_ws(Context *c, size_t bytes);          // grows c->ws
int check_kernel_fault(Context *c, const char *what);  // syncs the stream and reads err_flag

inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// bf16 operand planes: K-major, rows x ldk with ldk a multiple of 64 and the padding zeroed.
struct Planes {
    __nv_bfloat16 *hi = nullptr;
    __nv_bfloat16 *lo = nullptr;  // residual plane (x - bf16(x)), only for BF16X3
    int64_t rows = 0, k = 0, ldk = 0;
    int64_t cap_rows = 0;         // allocated rows (>= rows); scratch planes are reused across calls
    float *row_sumsq = nullptr;   // per row sum of squares of the source values
    float *maxabs = nullptr;      // device scalar: max |value|
    bool valid = false, lo_valid = false;
    // fp16 plane (NB200_PREC_F16): f16 = half(x * 2^e) with e a per-plane power-of-two exponent held on the device
    __half *f16 = nullptr;
    int *f16_exp = nullptr;        // device int
    float *f16_maxscaled = nullptr; // device float: max |x * 2^e| seen by the conversion (overflow check)
    bool f16_valid = false;
};

// ---- block GEMM on tensor cores (nb200_gemm.cu) ------------------------------------------------
struct GemmEpilogue {
    // mode 0: every split writes its fp32 partial tile to the workspace; a reduce kernel follows.
    // mode 1: (splits == 1) bias/relu applied in the kernel, fp32 result written to `out`.
    int mode = 0;
    float *out = nullptr;
    int64_t out_ld = 0;
    const float *bias_m = nullptr;  // per row of P
    const float *bias_n = nullptr;  // per column of P
    int relu = 0;
    // launch only splits [split_begin, split_begin + split_count) of the `splits` the workspace is laid out for,
    // on `stream` (default: the context's stream)
    int split_begin = 0, split_count = -1;
    cudaStream_t stream = nullptr;
};
enum { kFmtBF16 = 0, kFmtF16 = 1 };
// P[M x N] = U[M x K] * V[N x K]^T with U,V bf16 K-major planes; nterms 1 (hi*hi) or 3.
// splits > 1 (or mode 0) leaves partials in ctx->ws laid out [split][Mpad][Npad].
int launch_gemm_tc(Context *c, const Planes &u, const Planes &v, int M, int N, int nterms, int splits,
                   const GemmEpilogue &epi, int *Mpad, int *Npad, int *splits_used, int fmt = kFmtBF16);
int choose_splits(Context *c, int M, int N, int64_t ldk, int nterms);
// fused layer 1 (fp64 inputs converted to fp16 inside the persistent GEMM through an L2-resident ring)
struct FusedScratch {
    __half *ring = nullptr;
    size_t ring_bytes = 0;
    int *flags = nullptr;
    int flag_cap = 0;
};
int launch_l1_fused(Context *c, const double *x, int64_t x_ld, int rows_total, int cols, int row_base, int n_mtiles,
                    const Planes &xp, const Planes &w1, int N, const float *bias_n, float *out, int64_t out_ld,
                    float *out_maxabs, FusedScratch *fs);
// streamed layer 1 (nb200_l1stream.cu): fp64 inputs converted inside the clusters that consume them, (M tile, K range)
// segments laid out in waves; fp32 partial tiles go to ctx->ws and launch_l1_stream_reduce sums them
struct L1StreamSeg { int tile, kb0, kb1, slot; };
struct L1StreamPlan {
    std::vector<L1StreamSeg> segs;   // [clusters][max_seg], each cluster's list ends at the first empty segment
    std::vector<int> tile_slots;     // partial tiles per cluster tile
    int clusters = 0, max_seg = 0, max_slots = 1;
};
struct L1StreamPart {                // one launch: rows [row0, row0 + rows) through the kernel with cluster shape mt
    L1StreamPlan plan;
    int plan_rows = -1, plan_kb = -1, plan_g = -1, plan_mt = -1;
    unsigned char *dev = nullptr;    // device copy of the plan
    size_t dev_bytes = 0;
    const L1StreamSeg *seg_dev = nullptr;
    const int *slots_dev = nullptr;
    int row0 = 0, rows = 0, mt = 0, tile_rows = 0;
    long long slot_stride = 0, ws_off = 0;   // in floats
};
struct L1StreamScratch {
    L1StreamPart part[2];
    int nparts = 0;
    int max_slots() const { int m = 1; for (int i = 0; i < nparts; i++) m = part[i].plan.max_slots > m ? part[i].plan.max_slots : m; return m; }
};
extern int g_l1_stream_mt;
int plan_l1_stream(int tiles, int num_kb, int clusters, L1StreamPlan *plan);
bool l1_stream_supported(const double *x, int64_t x_ld, int64_t cols, int N, const Planes &w1);
int launch_l1_stream(Context *c, const double *x, int64_t x_ld, int rows, int cols, const int *exp_x, float *row_sumsq,
                     float *maxscaled, const Planes &w1, int N, L1StreamScratch *sc);
int launch_l1_stream_reduce(Context *c, const L1StreamScratch &sc, int rows, int N, const float *bias_n, int relu,
                            const int *exp_u, const int *exp_v, float *out, int64_t out_ld, float *out_maxabs);
// conv2d as implicit GEMM: NHWC bf16 image (channels padded to cp, multiple of 64), kernel planes with K = (tap, channel)
int launch_conv_implicit(Context *c, const __nv_bfloat16 *img_hi, const __nv_bfloat16 *img_lo, int n, int h, int w, int cp,
                         const Planes &ker, int kout, int k, int pad, int nterms, const float *bias, int relu, float *out);
// fp64 NCHW images -> bf16 hi/lo NHWC with channels zero padded to cp; fp64 kernels [kout][c][k][k] -> planes (tap, channel)
int launch_nchw_to_nhwc_planes(Context *c, const double *img, int n, int ch, int h, int w, int cp, __nv_bfloat16 *hi,
                               __nv_bfloat16 *lo);
int launch_kernel_to_planes(Context *c, const double *ker, int kout, int ch, int k, int cp, Planes *p, bool want_lo);

// ---- other kernels (nb200_kernels.cu) -----------------------------------------------------------
// source (double or float, row-major rows x cols, ld) -> planes of the same orientation
int launch_convert_planes(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld,
                          Planes *p, bool want_lo);
// source rows x cols -> planes of the transpose (plane rows = cols, K = rows)
int launch_convert_planes_t(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld,
                            Planes *p, bool want_lo);
struct ReduceOut {
    float *f32 = nullptr;  // P orientation, ld = f32_ld
    int64_t f32_ld = 0;
    double *f64 = nullptr;
    int64_t f64_ld = 0;
    Planes *planes = nullptr;  // P orientation (rows = M, K = N)
    bool want_lo = false;
    // fp16 operands carry power-of-two scales: the partial sums are multiplied by 2^-(exp_u + exp_v)
    const int *exp_u = nullptr, *exp_v = nullptr;
    float *out_maxabs = nullptr;  // device float: max |result| (to pick the next layer's fp16 scale)
};
int launch_splitk_reduce(Context *c, int M, int N, int splits, int Mpad, int Npad, const float *bias_m,
                         const float *bias_n, int relu, const ReduceOut &o);
// fp16 plane of `src` (same orientation): picks the exponent from a row sample (sample_stride > 1) or exactly
int launch_convert_f16(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld, Planes *p,
                       int sample_stride, const float *known_maxabs);
// the same in two steps: pick the scale (and clear the statistics) now, convert column ranges later on any stream
int prepare_convert_f16(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld, Planes *p,
                        int sample_stride, const float *known_maxabs);
int launch_convert_f16_cols(Context *c, cudaStream_t stream, const void *src, bool src_is_f64, int64_t rows, int64_t cols,
                            int64_t ld, Planes *p, int64_t col_begin, int64_t col_count);
int launch_gather_convert_rows(Context *c, const double *src, int64_t ld, int64_t cols, const int32_t *list, int n,
                               Planes *p);
int launch_transpose_f32(Context *c, const float *src, int64_t rows, int64_t cols, int64_t ld, float *dst, int64_t dld);
int launch_f64_to_f32(Context *c, const double *src, int64_t n, float *dst);
struct EwiseSrc {  // one operand of an element-wise kernel: fp64 or fp32 storage with a row pitch
    const double *d;
    const float *f;
    long long ld;
};
int launch_result_to_images(Context *c, const void *src, bool src_is_f64, int64_t ld, int num_images, int group, int zsrc, int z, double *dst);
int launch_convert_sub(Context *c, const double *src, int64_t ld, int64_t rows, int64_t cols, bool transposed, Planes *p,
                       int64_t row_off, int64_t k_off, bool want_lo);
int alloc_planes_raw(Planes *p, int64_t rows, int64_t k, bool want_lo);
int launch_lstm_gates(Context *c, const float *g, int64_t g_ld, int L, int B, const double *bf, const double *bi,
                      const double *bo, const double *bc, double *f_t, double *i_t, double *o_t, double *c_temp);
// The fp64 blocks of a set as they lie in (pinned, device-visible) HOST memory: a row-major grid of block pointers.
// Kernels read single rows out of them over PCIe (zero-copy) for the label-exact cascade of a set that is resident on
// the device only in its fp16 side format.
struct HostBlocks {
    const double *const *blocks;  // device array of n_rb x n_cb host pointers
    int block_rows, block_cols;   // nominal block size
    int n_cb;                     // block columns
    int edge_ld;                  // row pitch of the blocks of the LAST block column (block_cols when padded, the
                                  // remaining columns when trimmed); all other blocks have pitch block_cols
};
int launch_gather_convert_rows_hb(Context *c, const HostBlocks &hb, int64_t cols, const int32_t *list, int n, Planes *p);
int launch_gather_rows_f64_hb(Context *c, const HostBlocks &hb, int64_t cols, const int32_t *list, int n, double *dst);
int alloc_f16_plane(Context *c, Planes *p, int64_t rows, int64_t cols);  // storage of an fp16 operand plane, no conversion
int launch_sum_blocks(Context *c, const EwiseSrc *src, int nsrc, int act, int64_t rows, int64_t cols, double *out);
int launch_bias_relu_f32(Context *c, float *p, int64_t rows, int64_t cols, int64_t ld, const float *bias_n);
int launch_synth_fill(Context *c, double *dst, int64_t rows, int64_t cols, int64_t ld, uint64_t seed,
                      int64_t row_offset, int64_t total_cols);

// fp64 GEMM P[M x N] = U * V^T, U(m,k) = u[m*u_sm + k*u_sk], V(n,k) = v[n*v_sn + k*v_sk]; split-K in fp64.
struct F64Epilogue {
    const double *bias_m = nullptr;
    const double *bias_n = nullptr;
    int relu = 0;
};
int launch_gemm_f64(Context *c, const double *u, int64_t u_sm, int64_t u_sk, const double *v, int64_t v_sn,
                    int64_t v_sk, int M, int N, int64_t K, double *out, int64_t out_ld, const F64Epilogue &e);

// row softmax / argmax / top-2 gap over logits (float or double).  probs may be null.
int launch_softmax_rows(Context *c, const void *logits, bool is_f64, int64_t rows, int64_t cols, int64_t ld,
                        double *probs, int64_t probs_ld, int32_t *labels, float *gap, const int32_t *row_list,
                        int nrows_list, int in_compact);
// flag rows whose top-2 gap is inside tau * eps_r; appends to list, counts in *count
int launch_flag_rows(Context *c, int64_t rows, const float *gap, const float *x_sumsq, const float *y_sumsq,
                     const float *w1_maxabs, const float *wo_maxabs, const float *wo_rownorm_max, float u, float tau,
                     int32_t *list, int *count, int max_list, const int32_t *src_list);
int launch_gather_convert_rows(Context *c, const double *src, int64_t ld, int64_t cols, const int32_t *list, int n,
                               Planes *p);
int launch_scatter_rows_f32(Context *c, const float *src, int64_t src_ld, int64_t cols, const int32_t *list, int n,
                            float *dst, int64_t dst_ld);
int launch_gather_rows_f64(Context *c, const double *src, int64_t ld, int64_t cols, const int32_t *list, int n,
                           double *dst);
int launch_max_rownorm(Context *c, const float *row_sumsq, int64_t rows, float *out_scalar);
int launch_extract_to_f64(Context *c, const void *src, bool src_is_f64, int64_t ld, int64_t row0, int64_t col0, int rows,
                          int cols, int valid_rows, int valid_cols, double *dst, int64_t dst_ld, int apply_exp);
int launch_scatter_rows(Context *c, const double *src, int64_t cols, const int32_t *list, int n, float *dst_f32,
                        int64_t dst_ld);
int launch_udf_gather(Context *c, const void *src, bool src_is_f64, int64_t ld, int total_rows, int total_cols, int br, int bc, int nbr, int nbc,
                      int features, int batch, double *xt);
int launch_udf_labels(Context *c, const double *zt, int labels, int batch, int nbr, int nbc, double *dst, int64_t dst_ld);
int launch_im2col(Context *c, const double *images, int n, int ch, int h, int w, int k, int stride, int padding,
                  double *out, int64_t out_ld);

}  // namespace nb200

// ---- device block set ---------------------------------------------------------------------------
// A user set of FFMatrixBlocks, de-blocked on the device.  Logical matrix total_rows x total_cols.
// Representations (any subset may be present; the others are derived on demand):
//   f64    logical row-major fp64 (what uploads write; the fp64 "master")
//   f32    logical row-major fp32 (tensor-core results)
//   rowp   bf16 planes with rows = logical rows, K = logical cols
//   colp   bf16 planes with rows = logical cols, K = logical rows
struct nb200_set {
    std::string db, name;
    int device = -1;               // the device the set lives on (a thread must work on it: nb200_use_device)
    int total_rows = 0, total_cols = 0, block_rows = 0, block_cols = 0;
    double *f64 = nullptr;
    float *f32 = nullptr;
    int64_t f32_ld = 0;
    nb200::Planes rowp, colp;
    float *wo_rownorm_max = nullptr;  // device scalar derived from rowp.row_sumsq (error model)
    uint64_t version = 0;          // bumped whenever the set's contents change (caches of derived operands)
    // uploads: epoch of the context when the set was last written by an upload, and (after nb200_upload_commit) events
    // that mark the end of its uploads on the two copy streams, so that a reader can wait for THIS set only
    uint64_t up_epoch = 0;
    cudaEvent_t up_fence[2] = {nullptr, nullptr};
    bool up_fenced = false;
    // fp16 side format (nb200_set_import_f16): no fp64 copy on the device, rows are fetched from the host blocks on demand
    const double **hb_dev = nullptr;
    std::vector<const double *> hb_host;  // the pointer table as given (kept: its upload is asynchronous)
    int hb_edge_ld = 0, hb_exp = 0;
    bool side_format = false;
    // an image set (nb200_conv2d_images_create): NCHW fp64 in f64, NHWC bf16 operand planes derived on first use
    int img_n = 0, img_c = 0, img_h = 0, img_w = 0;
    __nv_bfloat16 *img_hi = nullptr, *img_lo = nullptr;
    bool view_exp = false;         // downloads apply exp() (the `yo` view of the logits)
    nb200_set *view_of = nullptr;  // set whose storage a view reads
};
