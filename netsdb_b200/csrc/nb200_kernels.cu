// nb200_kernels.cu -- the HBM-bound kernels around the block GEMM:
//   * fp64 -> bf16 (hi / lo) operand-plane conversion, de-blocked K-major layout        (new, K8)
//   * split-K reduction fused with bias / ReLU / re-quantisation (FFAggMatrix's sum +
//     FFReluBiasSum.h:63-71 + the conversion for the next layer)                         (K3)
//   * row softmax / argmax / top-2 gap (FFRowAggregate + FFOutputLayer + label)          (K5)
//   * fp64 DFMA GEMM used for the exact path and for re-evaluating near-tie rows         (K9)
//   * seeded generator, transposes, im2col (ImageToChunks.h:43-102), gather/scatter.
#include "nb200_internal.h"

#include <cfloat>
#include <cstdlib>
#include <math_constants.h>

namespace nb200 {

#define NB_LAUNCH_CHECK(c)           \
    do {                             \
        (c)->launches++;             \
        NB_CUDA(cudaGetLastError()); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// bf16 hi/lo split
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void split_bf16(T x, __nv_bfloat16 &hi, __nv_bfloat16 &lo) {
    const float xf = static_cast<float>(x);
    hi = __float2bfloat16_rn(xf);
    // residual in the source precision, then rounded once
    const T res = x - static_cast<T>(__bfloat162float(hi));
    lo = __float2bfloat16_rn(static_cast<float>(res));
}

__device__ __forceinline__ float block_reduce_sum(float v, float *sh) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
    if (w == 0)
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    return v;  // valid in thread 0
}
__device__ __forceinline__ float block_reduce_max(float v, float *sh) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
    if (w == 0)
        for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    return v;
}

constexpr int kCvtThreads = 256;
constexpr int kCvtPerThread = 8;
constexpr int kCvtChunk = kCvtThreads * kCvtPerThread;  // elements of one row per block

// grid: (rows, ceil(ldk / kCvtChunk)).  Reads the row once, writes hi (and lo) planes, zero pads
// the K tail, accumulates the row's sum of squares and the global max |x| for the error model.
template <typename T, bool WANT_LO>
__global__ void __launch_bounds__(kCvtThreads)
k_convert_planes(const T *__restrict__ src, long long ld, long long cols, __nv_bfloat16 *__restrict__ hi,
                 __nv_bfloat16 *__restrict__ lo, long long ldk, float *__restrict__ row_sumsq,
                 float *__restrict__ maxabs, long long kext) {  // kext: K range written (a multiple of 4, <= ldk)
    __shared__ float sh[32];
    const long long row = blockIdx.x;
    const long long c0 = static_cast<long long>(blockIdx.y) * kCvtChunk + threadIdx.x * 4;
    const T *srow = src + row * ld;
    float ss = 0.f, mx = 0.f;
#pragma unroll
    for (int h = 0; h < kCvtPerThread / 4; h++) {
        const long long c = c0 + static_cast<long long>(h) * kCvtThreads * 4;
        if (c >= kext) continue;
        T v[4];
        if (c + 3 < cols && ((reinterpret_cast<uintptr_t>(srow + c) & (sizeof(T) * 2 - 1)) == 0)) {
            // two 16-byte (double) or one 16-byte (float) loads
            if (sizeof(T) == 8) {
                const double2 a = *reinterpret_cast<const double2 *>(srow + c);
                const double2 b = *reinterpret_cast<const double2 *>(srow + c + 2);
                v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
            } else {
                const float2 a = *reinterpret_cast<const float2 *>(srow + c);
                const float2 b = *reinterpret_cast<const float2 *>(srow + c + 2);
                v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) v[j] = (c + j < cols) ? srow[c + j] : T(0);
        }
        __nv_bfloat16 h4[4], l4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            split_bf16<T>(v[j], h4[j], l4[j]);
            const float f = static_cast<float>(v[j]);
            ss += f * f;
            mx = fmaxf(mx, fabsf(f));
        }
        *reinterpret_cast<uint2 *>(hi + row * ldk + c) = *reinterpret_cast<uint2 *>(h4);
        if (WANT_LO) *reinterpret_cast<uint2 *>(lo + row * ldk + c) = *reinterpret_cast<uint2 *>(l4);
    }
    ss = block_reduce_sum(ss, sh);
    mx = block_reduce_max(mx, sh);
    if (threadIdx.x == 0) {
        if (row_sumsq) atomicAdd(row_sumsq + row, ss);
        if (maxabs) atomicMax(reinterpret_cast<unsigned int *>(maxabs), __float_as_uint(mx));
    }
}

static int alloc_planes(Planes *p, int64_t rows, int64_t k, bool want_lo) {
    const int64_t ldk = round_up(k, 64);
    if ((p->hi || p->f16) && (p->cap_rows < rows || p->ldk != ldk)) {
        nb_free(p->hi); nb_free(p->lo); nb_free(p->row_sumsq); nb_free(p->maxabs);
        nb_free(p->f16); nb_free(p->f16_exp); nb_free(p->f16_maxscaled);
        *p = Planes();
    }
    if (!p->hi && !p->f16) p->cap_rows = rows;
    p->rows = rows; p->k = k; p->ldk = ldk;
    const size_t bytes = static_cast<size_t>(p->cap_rows ? p->cap_rows : 1) * ldk * sizeof(__nv_bfloat16);
    if (!p->hi) NB_CUDA(nb_malloc(&p->hi, bytes));
    if (want_lo && !p->lo) NB_CUDA(nb_malloc(&p->lo, bytes));
    if (!p->row_sumsq) NB_CUDA(nb_malloc(&p->row_sumsq, static_cast<size_t>(p->cap_rows ? p->cap_rows : 1) * sizeof(float)));
    if (!p->maxabs) NB_CUDA(nb_malloc(&p->maxabs, sizeof(float)));
    return NB200_OK;
}

int launch_convert_planes(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld,
                          Planes *p, bool want_lo) {
    NB_TRY(alloc_planes(p, rows, cols, want_lo));
    NB_CUDA(cudaMemsetAsync(p->row_sumsq, 0, static_cast<size_t>(rows) * sizeof(float), c->stream));
    NB_CUDA(cudaMemsetAsync(p->maxabs, 0, sizeof(float), c->stream));
    if (rows == 0 || cols == 0) { p->valid = true; p->lo_valid = want_lo; return NB200_OK; }
    dim3 grid(static_cast<unsigned>(rows), static_cast<unsigned>(ceil_div(p->ldk, kCvtChunk)));
    if (grid.y > 65535) { set_error("convert: K too large for one launch (%lld)", (long long)cols); return NB200_ERR_INVALID; }
    if (src_is_f64) {
        if (want_lo) k_convert_planes<double, true><<<grid, kCvtThreads, 0, c->stream>>>((const double *)src, ld, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
        else k_convert_planes<double, false><<<grid, kCvtThreads, 0, c->stream>>>((const double *)src, ld, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
    } else {
        if (want_lo) k_convert_planes<float, true><<<grid, kCvtThreads, 0, c->stream>>>((const float *)src, ld, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
        else k_convert_planes<float, false><<<grid, kCvtThreads, 0, c->stream>>>((const float *)src, ld, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
    }
    NB_LAUNCH_CHECK(c);
    p->valid = true;
    p->lo_valid = want_lo;
    return NB200_OK;
}


// ---------------------------------------------------------------------------------------------
// fp16 operand planes (NB200_PREC_F16).  fp16 has 3 more mantissa bits than bf16 at the same tensor
// rate but only 5 exponent bits, so every plane carries a power-of-two scale 2^e chosen from max |x|:
// with the exact maximum the largest element lands in [2^14, 2^15); with a maximum estimated from a
// sample of rows (the per-query inputs, where a full extra pass over 47.8 GB would cost as much as the
// conversion itself) it lands in [2^10, 2^11), leaving 32x headroom, and the conversion records the
// largest scaled magnitude so that an overflow is detected and the caller can fall back.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
k_max_sampled(const T *__restrict__ src, long long ld, long long rows, long long cols, int stride,
              float *__restrict__ maxabs) {
    // work item = (sampled row, 8192-column piece); grid-stride so that a handful of sampled rows still fills the GPU
    const long long nsamp = (rows + stride - 1) / stride;
    const long long pieces = (cols + 8191) / 8192;
    float mx = 0.f;
    for (long long w = blockIdx.x; w < nsamp * pieces; w += gridDim.x) {
        const T *srow = src + (w / pieces) * stride * ld;
        const long long c0 = (w % pieces) * 8192, c1 = (c0 + 8192 < cols) ? c0 + 8192 : cols;
        for (long long c = c0 + threadIdx.x; c < c1; c += 256) mx = fmaxf(mx, fabsf(static_cast<float>(srow[c])));
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned int *>(maxabs), __float_as_uint(mx));
}
__global__ void k_exp_from_max(const float *__restrict__ maxabs, int target_log2, int *__restrict__ exp_out) {
    int q = 0;
    const float m = *maxabs;
    if (m > 0.f && isfinite(m)) frexpf(m, &q);  // m = f * 2^q with f in [0.5, 1)  =>  m * 2^(target - q) < 2^target
    *exp_out = (m > 0.f && isfinite(m)) ? target_log2 - q : 0;
}
// Persistent: a bounded number of CTAs walks over (row, 4096-column piece) tiles of the column range, so that
// when this conversion runs beside the layer-1 GEMM (K-chunk pipelining) it leaves SM thread slots to the GEMM's
// CTAs instead of flooding the block scheduler with millions of tiny CTAs.  16 elements (4 x 32-byte loads) per
// thread are in flight to keep HBM busy at that low occupancy.
constexpr int kF16Piece = kCvtThreads * 16;
template <typename T>
__global__ void __launch_bounds__(kCvtThreads)
k_convert_f16(const T *__restrict__ src, long long ld, long long rows, long long cols, __half *__restrict__ dst,
              long long ldk, const int *__restrict__ exp_in, float *__restrict__ row_sumsq,
              float *__restrict__ maxscaled, long long col_begin, long long col_end) {
    __shared__ float sh[32];
    const float scale = exp2f(static_cast<float>(*exp_in));
    const long long pieces = (col_end - col_begin + kF16Piece - 1) / kF16Piece;
    const long long tiles = rows * pieces;
    float mx = 0.f;
    for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
        const long long row = t / pieces, piece = t % pieces;
        const T *srow = src + row * ld;
        const long long c0 = col_begin + piece * kF16Piece + threadIdx.x * 4;
        T v[4][4];
#pragma unroll
        for (int h = 0; h < 4; h++) {
            const long long c = c0 + static_cast<long long>(h) * kCvtThreads * 4;
            if (c + 3 < cols && c + 3 < col_end && ((reinterpret_cast<uintptr_t>(srow + c) & (sizeof(T) * 2 - 1)) == 0)) {
                if (sizeof(T) == 8) {
                    const double2 a = *reinterpret_cast<const double2 *>(srow + c);
                    const double2 b = *reinterpret_cast<const double2 *>(srow + c + 2);
                    v[h][0] = a.x; v[h][1] = a.y; v[h][2] = b.x; v[h][3] = b.y;
                } else {
                    const float2 a = *reinterpret_cast<const float2 *>(srow + c);
                    const float2 b = *reinterpret_cast<const float2 *>(srow + c + 2);
                    v[h][0] = a.x; v[h][1] = a.y; v[h][2] = b.x; v[h][3] = b.y;
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) v[h][j] = (c + j < cols && c + j < col_end) ? srow[c + j] : T(0);
            }
        }
        float ss = 0.f;
#pragma unroll
        for (int h = 0; h < 4; h++) {
            const long long c = c0 + static_cast<long long>(h) * kCvtThreads * 4;
            if (c >= col_end) continue;  // col_end is a multiple of 64, c of 4: whole vector in or out
            __half h4[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float f = static_cast<float>(v[h][j]);
                const float sc = f * scale;  // exact (power of two) unless it overflows
                h4[j] = __float2half_rn(sc);
                ss += f * f;
                mx = fmaxf(mx, fabsf(sc));
            }
            *reinterpret_cast<uint2 *>(dst + row * ldk + c) = *reinterpret_cast<uint2 *>(h4);
        }
        ss = block_reduce_sum(ss, sh);
        if (threadIdx.x == 0 && row_sumsq) atomicAdd(row_sumsq + row, ss);
    }
    mx = block_reduce_max(mx, sh);
    if (threadIdx.x == 0 && maxscaled) atomicMax(reinterpret_cast<unsigned int *>(maxscaled), __float_as_uint(mx));
}

int alloc_f16_plane(Context *c, Planes *p, int64_t rows, int64_t cols) {
    const int64_t ldk = round_up(cols, 64);
    if ((p->hi || p->f16) && (p->cap_rows < rows || p->ldk != ldk)) {
        nb_free(p->hi); nb_free(p->lo); nb_free(p->row_sumsq); nb_free(p->maxabs);
        nb_free(p->f16); nb_free(p->f16_exp); nb_free(p->f16_maxscaled);
        *p = Planes();
    }
    if (!p->hi && !p->f16) p->cap_rows = rows;
    p->rows = rows; p->k = cols; p->ldk = ldk;
    const size_t cap = static_cast<size_t>(p->cap_rows ? p->cap_rows : 1);
    if (!p->f16) NB_CUDA(nb_malloc(&p->f16, cap * ldk * sizeof(__half)));
    if (!p->f16_exp) NB_CUDA(nb_malloc(&p->f16_exp, sizeof(int)));
    if (!p->f16_maxscaled) NB_CUDA(nb_malloc(&p->f16_maxscaled, sizeof(float)));
    if (!p->row_sumsq) NB_CUDA(nb_malloc(&p->row_sumsq, cap * sizeof(float)));
    if (!p->maxabs) NB_CUDA(nb_malloc(&p->maxabs, sizeof(float)));
    (void)c;
    return NB200_OK;
}

int prepare_convert_f16(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld, Planes *p,
                        int sample_stride, const float *known_maxabs) {
    const int64_t ldk = round_up(cols, 64);
    if ((p->hi || p->f16) && (p->cap_rows < rows || p->ldk != ldk)) {
        nb_free(p->hi); nb_free(p->lo); nb_free(p->row_sumsq); nb_free(p->maxabs);
        nb_free(p->f16); nb_free(p->f16_exp); nb_free(p->f16_maxscaled);
        *p = Planes();
    }
    if (!p->hi && !p->f16) p->cap_rows = rows;
    p->rows = rows; p->k = cols; p->ldk = ldk;
    const size_t cap = static_cast<size_t>(p->cap_rows ? p->cap_rows : 1);
    if (!p->f16) NB_CUDA(nb_malloc(&p->f16, cap * ldk * sizeof(__half)));
    if (!p->f16_exp) NB_CUDA(nb_malloc(&p->f16_exp, sizeof(int)));
    if (!p->f16_maxscaled) NB_CUDA(nb_malloc(&p->f16_maxscaled, sizeof(float)));
    if (!p->row_sumsq) NB_CUDA(nb_malloc(&p->row_sumsq, cap * sizeof(float)));
    if (!p->maxabs) NB_CUDA(nb_malloc(&p->maxabs, sizeof(float)));
    NB_CUDA(cudaMemsetAsync(p->row_sumsq, 0, static_cast<size_t>(rows ? rows : 1) * sizeof(float), c->stream));
    NB_CUDA(cudaMemsetAsync(p->f16_maxscaled, 0, sizeof(float), c->stream));
    p->f16_valid = true;
    if (rows == 0 || cols == 0) {
        NB_CUDA(cudaMemsetAsync(p->f16_exp, 0, sizeof(int), c->stream));
        return NB200_OK;
    }
    const float *mx = known_maxabs;
    if (!mx) {
        NB_CUDA(cudaMemsetAsync(p->maxabs, 0, sizeof(float), c->stream));
        const int stride = sample_stride > 1 ? sample_stride : 1;
        const int64_t items = ceil_div(rows, stride) * ceil_div(cols, 8192);
        const unsigned blocks = static_cast<unsigned>(items < 148 * 16 ? items : 148 * 16);
        if (src_is_f64) k_max_sampled<double><<<blocks, 256, 0, c->stream>>>((const double *)src, ld, rows, cols, stride, p->maxabs);
        else k_max_sampled<float><<<blocks, 256, 0, c->stream>>>((const float *)src, ld, rows, cols, stride, p->maxabs);
        NB_LAUNCH_CHECK(c);
        mx = p->maxabs;
    }
    k_exp_from_max<<<1, 1, 0, c->stream>>>(mx, (!known_maxabs && sample_stride > 1) ? 11 : 15, p->f16_exp);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

int launch_convert_f16_cols(Context *c, cudaStream_t stream, const void *src, bool src_is_f64, int64_t rows, int64_t cols,
                            int64_t ld, Planes *p, int64_t col_begin, int64_t col_count) {
    if (rows == 0 || cols == 0 || col_count <= 0) return NB200_OK;
    const int64_t col_end = (col_begin + col_count < p->ldk) ? col_begin + col_count : p->ldk;
    if (col_end <= col_begin) return NB200_OK;
    const int64_t tiles = rows * ceil_div(col_end - col_begin, kF16Piece);
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    // beside the GEMM (side stream): 2 CTAs per SM; alone: 8 per SM
    static const int side_ctas = getenv("NB200_CVT_CTAS") ? atoi(getenv("NB200_CVT_CTAS")) : 4;
    const int64_t cap = static_cast<int64_t>(sms) * (stream == c->stream ? 8 : side_ctas);
    const unsigned grid = static_cast<unsigned>(tiles < cap ? tiles : cap);
    if (src_is_f64) k_convert_f16<double><<<grid, kCvtThreads, 0, stream>>>((const double *)src, ld, rows, cols, p->f16, p->ldk, p->f16_exp, p->row_sumsq, p->f16_maxscaled, col_begin, col_end);
    else k_convert_f16<float><<<grid, kCvtThreads, 0, stream>>>((const float *)src, ld, rows, cols, p->f16, p->ldk, p->f16_exp, p->row_sumsq, p->f16_maxscaled, col_begin, col_end);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

int launch_convert_f16(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld, Planes *p,
                       int sample_stride, const float *known_maxabs) {
    NB_TRY(prepare_convert_f16(c, src, src_is_f64, rows, cols, ld, p, sample_stride, known_maxabs));
    return launch_convert_f16_cols(c, c->stream, src, src_is_f64, rows, cols, ld, p, 0, p->ldk);
}

// Transposing conversion: plane[c][r] = src[r][c].  32 x 32 tiles through shared memory so both
// the read (along c) and the write (along r, the plane's K) are coalesced.
template <typename T, bool WANT_LO>
__global__ void __launch_bounds__(256)
k_convert_planes_t(const T *__restrict__ src, long long ld, long long rows, long long cols,
                   __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, long long ldk,
                   float *__restrict__ row_sumsq, float *__restrict__ maxabs, long long kext) {
    __shared__ T tile[32][33];
    const long long r0 = static_cast<long long>(blockIdx.y) * 32, c0 = static_cast<long long>(blockIdx.x) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int i = ty; i < 32; i += 8) {
        const long long r = r0 + i, cc = c0 + tx;
        tile[i][tx] = (r < rows && cc < cols) ? src[r * ld + cc] : T(0);
    }
    __syncthreads();
    float mx = 0.f;
    for (int i = ty; i < 32; i += 8) {
        const long long pc = c0 + i;  // plane row = source column
        const long long pr = r0 + tx; // plane K index = source row
        const T v = tile[tx][i];
        __nv_bfloat16 h, l;
        split_bf16<T>(v, h, l);
        const float f = static_cast<float>(v);
        float ss = f * f;
        mx = fmaxf(mx, fabsf(f));
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (pc < cols && pr < kext) {
            hi[pc * ldk + pr] = h;
            if (WANT_LO) lo[pc * ldk + pr] = l;
        }
        if (tx == 0 && pc < cols && row_sumsq) atomicAdd(row_sumsq + pc, ss);
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (tx == 0 && maxabs) atomicMax(reinterpret_cast<unsigned int *>(maxabs), __float_as_uint(mx));
}

int launch_convert_planes_t(Context *c, const void *src, bool src_is_f64, int64_t rows, int64_t cols, int64_t ld,
                            Planes *p, bool want_lo) {
    NB_TRY(alloc_planes(p, cols, rows, want_lo));
    NB_CUDA(cudaMemsetAsync(p->row_sumsq, 0, static_cast<size_t>(cols) * sizeof(float), c->stream));
    NB_CUDA(cudaMemsetAsync(p->maxabs, 0, sizeof(float), c->stream));
    if (rows == 0 || cols == 0) { p->valid = true; p->lo_valid = want_lo; return NB200_OK; }
    // the tile grid covers K up to ldk so that the zero padding is (re)written too
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 32)), static_cast<unsigned>(ceil_div(p->ldk, 32)));
    if (grid.y > 65535) { set_error("transposing convert: source has too many rows (%lld)", (long long)rows); return NB200_ERR_INVALID; }
    if (src_is_f64) {
        if (want_lo) k_convert_planes_t<double, true><<<grid, 256, 0, c->stream>>>((const double *)src, ld, rows, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
        else k_convert_planes_t<double, false><<<grid, 256, 0, c->stream>>>((const double *)src, ld, rows, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
    } else {
        if (want_lo) k_convert_planes_t<float, true><<<grid, 256, 0, c->stream>>>((const float *)src, ld, rows, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
        else k_convert_planes_t<float, false><<<grid, 256, 0, c->stream>>>((const float *)src, ld, rows, cols, p->hi, p->lo, p->ldk, p->row_sumsq, p->maxabs, p->ldk);
    }
    NB_LAUNCH_CHECK(c);
    p->valid = true;
    p->lo_valid = want_lo;
    return NB200_OK;
}


// Conversion of an fp64 matrix into a sub-range of wider operand planes (K-concatenated operands of the fused LSTM
// step): plane rows [row_off, row_off + rows), K range [k_off, k_off + round_up(cols, 64)).  `transposed` reads a
// (K x rows) source instead (plane[c][k_off + r] = src[r][c]).
int launch_convert_sub(Context *c, const double *src, int64_t ld, int64_t rows, int64_t cols, bool transposed, Planes *p,
                       int64_t row_off, int64_t k_off, bool want_lo) {
    if (rows == 0 || cols == 0) return NB200_OK;
    __nv_bfloat16 *hi = p->hi + row_off * p->ldk + k_off;
    __nv_bfloat16 *lo = want_lo ? p->lo + row_off * p->ldk + k_off : nullptr;
    if (!transposed) {
        const int64_t kext = round_up(cols, 64);
        dim3 grid(static_cast<unsigned>(rows), static_cast<unsigned>(ceil_div(kext, kCvtChunk)));
        if (want_lo) k_convert_planes<double, true><<<grid, kCvtThreads, 0, c->stream>>>(src, ld, cols, hi, lo, p->ldk, nullptr, nullptr, kext);
        else k_convert_planes<double, false><<<grid, kCvtThreads, 0, c->stream>>>(src, ld, cols, hi, lo, p->ldk, nullptr, nullptr, kext);
    } else {  // here `rows` is the source's row count = K extent, `cols` its column count = plane rows
        const int64_t kext = round_up(rows, 64);
        dim3 grid(static_cast<unsigned>(ceil_div(cols, 32)), static_cast<unsigned>(ceil_div(kext, 32)));
        if (want_lo) k_convert_planes_t<double, true><<<grid, 256, 0, c->stream>>>(src, ld, rows, cols, hi, lo, p->ldk, nullptr, nullptr, kext);
        else k_convert_planes_t<double, false><<<grid, 256, 0, c->stream>>>(src, ld, rows, cols, hi, lo, p->ldk, nullptr, nullptr, kext);
    }
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
int alloc_planes_raw(Planes *p, int64_t rows, int64_t k, bool want_lo) {
    NB_TRY(alloc_planes(p, rows, k, want_lo));
    p->valid = true;
    p->lo_valid = want_lo;
    return NB200_OK;
}

// The four LSTMThreeWaySum joins of one step in a single pass (src/LSTM/headers/LSTMThreeWaySum.h:76-95;
// LSTMTest.cc:240-247, :305-312, :370-377, :435-440): g holds the stacked pre-activations [f; i; o; c] (4L x B, fp32,
// W.x + U.h already summed by the K-concatenated GEMM), each gate adds its own bias matrix and applies sigmoid
// (f, i, o) or tanh (c).
__global__ void __launch_bounds__(256)
k_lstm_gates(const float *__restrict__ g, long long g_ld, int L, int B, const double *__restrict__ bf,
             const double *__restrict__ bi, const double *__restrict__ bo, const double *__restrict__ bc,
             double *__restrict__ f_t, double *__restrict__ i_t, double *__restrict__ o_t, double *__restrict__ c_temp) {
    const long long n = static_cast<long long>(L) * B;
    for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < n;
         e += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = e / B, col = e - r * B;
        const double sf = static_cast<double>(g[r * g_ld + col]) + bf[e];
        const double si = static_cast<double>(g[(r + L) * g_ld + col]) + bi[e];
        const double so = static_cast<double>(g[(r + 2ll * L) * g_ld + col]) + bo[e];
        const double sc = static_cast<double>(g[(r + 3ll * L) * g_ld + col]) + bc[e];
        f_t[e] = 1.0 / (1.0 + exp(-1.0 * sf));
        i_t[e] = 1.0 / (1.0 + exp(-1.0 * si));
        o_t[e] = 1.0 / (1.0 + exp(-1.0 * so));
        double t;
        if (fabs(sc) > 350.0) t = copysign(1.0, sc);
        else { const double ep = exp(sc), em = exp(-1.0 * sc); t = (ep - em) / (ep + em); }
        c_temp[e] = t;
    }
}
int launch_lstm_gates(Context *c, const float *g, int64_t g_ld, int L, int B, const double *bf, const double *bi,
                      const double *bo, const double *bc, double *f_t, double *i_t, double *o_t, double *c_temp) {
    const int64_t n = static_cast<int64_t>(L) * B;
    if (n == 0) return NB200_OK;
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    const int blocks = static_cast<int>(ceil_div(n, 256) < sms * 8 ? ceil_div(n, 256) : sms * 8);
    k_lstm_gates<<<blocks, 256, 0, c->stream>>>(g, g_ld, L, B, bf, bi, bo, bc, f_t, i_t, o_t, c_temp);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// Gather the listed rows of an fp64 matrix straight into compact bf16 hi/lo planes (the inputs of the
// second cascade level): plane row i = source row list[i].
__global__ void __launch_bounds__(kCvtThreads)
k_gather_convert_rows(const double *__restrict__ src, long long ld, long long cols, const int32_t *__restrict__ list,
                      __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, long long ldk) {
    const long long i = blockIdx.x;
    const double *srow = src + static_cast<long long>(list[i]) * ld;
    for (long long c = (static_cast<long long>(blockIdx.y) * kCvtThreads + threadIdx.x) * 4; c < ldk;
         c += static_cast<long long>(gridDim.y) * kCvtThreads * 4) {
        __nv_bfloat16 h4[4], l4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) split_bf16<double>((c + j < cols) ? srow[c + j] : 0.0, h4[j], l4[j]);
        *reinterpret_cast<uint2 *>(hi + i * ldk + c) = *reinterpret_cast<uint2 *>(h4);
        *reinterpret_cast<uint2 *>(lo + i * ldk + c) = *reinterpret_cast<uint2 *>(l4);
    }
}
int launch_gather_convert_rows(Context *c, const double *src, int64_t ld, int64_t cols, const int32_t *list, int n,
                               Planes *p) {
    NB_TRY(alloc_planes(p, n, cols, true));
    p->valid = p->lo_valid = true;
    if (n == 0) return NB200_OK;
    const int64_t chunks = ceil_div(p->ldk, kCvtThreads * 4);
    dim3 grid(n, static_cast<unsigned>(chunks < 64 ? chunks : 64));
    k_gather_convert_rows<<<grid, kCvtThreads, 0, c->stream>>>(src, ld, cols, list, p->hi, p->lo, p->ldk);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
// The same two gathers for a set whose fp64 data live in HOST blocks (side format): the listed rows are read straight
// out of pinned host memory over PCIe, one block-column segment at a time (each warp reads contiguous runs).
__device__ __forceinline__ double hb_at(const HostBlocks &hb, long long r, long long cidx) {
    const int bcol = static_cast<int>(cidx / hb.block_cols);
    const double *blk = hb.blocks[(r / hb.block_rows) * hb.n_cb + bcol];
    const long long ld = (bcol == hb.n_cb - 1) ? hb.edge_ld : hb.block_cols;
    return blk[(r % hb.block_rows) * ld + (cidx - static_cast<long long>(bcol) * hb.block_cols)];
}
__global__ void __launch_bounds__(kCvtThreads)
k_gather_convert_rows_hb(HostBlocks hb, long long cols, const int32_t *__restrict__ list, __nv_bfloat16 *__restrict__ hi,
                         __nv_bfloat16 *__restrict__ lo, long long ldk) {
    const long long i = blockIdx.x, r = list[i];
    for (long long c = (static_cast<long long>(blockIdx.y) * kCvtThreads + threadIdx.x) * 4; c < ldk;
         c += static_cast<long long>(gridDim.y) * kCvtThreads * 4) {
        __nv_bfloat16 h4[4], l4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) split_bf16<double>((c + j < cols) ? hb_at(hb, r, c + j) : 0.0, h4[j], l4[j]);
        *reinterpret_cast<uint2 *>(hi + i * ldk + c) = *reinterpret_cast<uint2 *>(h4);
        *reinterpret_cast<uint2 *>(lo + i * ldk + c) = *reinterpret_cast<uint2 *>(l4);
    }
}
int launch_gather_convert_rows_hb(Context *c, const HostBlocks &hb, int64_t cols, const int32_t *list, int n, Planes *p) {
    NB_TRY(alloc_planes(p, n, cols, true));
    p->valid = p->lo_valid = true;
    if (n == 0) return NB200_OK;
    const int64_t chunks = ceil_div(p->ldk, kCvtThreads * 4);
    dim3 grid(n, static_cast<unsigned>(chunks < 64 ? chunks : 64));
    k_gather_convert_rows_hb<<<grid, kCvtThreads, 0, c->stream>>>(hb, cols, list, p->hi, p->lo, p->ldk);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
__global__ void k_gather_rows_f64_hb(HostBlocks hb, long long cols, const int32_t *__restrict__ list, double *__restrict__ dst) {
    const long long r = list[blockIdx.y];
    for (long long j = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; j < cols;
         j += static_cast<long long>(gridDim.x) * blockDim.x)
        dst[static_cast<long long>(blockIdx.y) * cols + j] = hb_at(hb, r, j);
}
int launch_gather_rows_f64_hb(Context *c, const HostBlocks &hb, int64_t cols, const int32_t *list, int n, double *dst) {
    if (n == 0 || cols == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 256) < 64 ? ceil_div(cols, 256) : 64), n);
    k_gather_rows_f64_hb<<<grid, 256, 0, c->stream>>>(hb, cols, list, dst);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
// compact fp32 rows -> rows list[i] of a larger fp32 matrix
__global__ void k_scatter_rows_f32(const float *__restrict__ src, long long src_ld, long long cols,
                                   const int32_t *__restrict__ list, float *__restrict__ dst, long long dst_ld) {
    const long long r = list[blockIdx.y];
    for (long long j = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; j < cols;
         j += static_cast<long long>(gridDim.x) * blockDim.x)
        dst[r * dst_ld + j] = src[static_cast<long long>(blockIdx.y) * src_ld + j];
}
int launch_scatter_rows_f32(Context *c, const float *src, int64_t src_ld, int64_t cols, const int32_t *list, int n,
                            float *dst, int64_t dst_ld) {
    if (n == 0 || cols == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 256) < 64 ? ceil_div(cols, 256) : 64), n);
    k_scatter_rows_f32<<<grid, 256, 0, c->stream>>>(src, src_ld, cols, list, dst, dst_ld);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// ---------------------------------------------------------------------------------------------
// split-K reduction + FFReluBiasSum epilogue + re-quantisation
// ---------------------------------------------------------------------------------------------
// grid: (M, ceil(Ncover / (256*4))) where Ncover = planes ? ldk : N.  Splits are summed in index
// order, so the result does not depend on scheduling (the reference's aggregation order is
// arrival order, CombinedShuffleSink.h:112-144).
__global__ void __launch_bounds__(256)
k_splitk_reduce(const float *__restrict__ ws, long long split_stride, int ws_ld, int splits, int M, int N,
                const float *__restrict__ bias_m, const float *__restrict__ bias_n, int relu,
                float *__restrict__ of32, long long of32_ld, double *__restrict__ of64, long long of64_ld,
                __nv_bfloat16 *__restrict__ phi, __nv_bfloat16 *__restrict__ plo, long long ldk,
                float *__restrict__ row_sumsq, const int *__restrict__ exp_u, const int *__restrict__ exp_v,
                float *__restrict__ out_maxabs) {
    __shared__ float sh[32];
    const float oscale = exp_u ? exp2f(-static_cast<float>(*exp_u + *exp_v)) : 1.f;
    const long long m = blockIdx.x;
    const int n0 = (blockIdx.y * 256 + threadIdx.x) * 4;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    const bool in_ws = n0 < ws_ld;  // ws_ld is a multiple of 128, so all four or none
    if (in_ws && n0 < N) {
        const float *p = ws + m * ws_ld + n0;
        float4 acc = *reinterpret_cast<const float4 *>(p);
        for (int s = 1; s < splits; s++) {
            const float4 t = *reinterpret_cast<const float4 *>(p + s * split_stride);
            acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
        v[0] = acc.x * oscale; v[1] = acc.y * oscale; v[2] = acc.z * oscale; v[3] = acc.w * oscale;
        const float bm = bias_m ? bias_m[m] : 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (n0 + j < N) {
                v[j] += bm + (bias_n ? bias_n[n0 + j] : 0.f);
                if (relu) v[j] = fmaxf(v[j], 0.f);
            } else {
                v[j] = 0.f;
            }
        }
    }
    if (of32) {
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (n0 + j < N) of32[m * of32_ld + n0 + j] = v[j];
    }
    if (of64) {
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (n0 + j < N) of64[m * of64_ld + n0 + j] = static_cast<double>(v[j]);
    }
    float ss = 0.f;
    if (phi && n0 < ldk) {
        __nv_bfloat16 h4[4], l4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            split_bf16<float>(v[j], h4[j], l4[j]);
            ss += v[j] * v[j];
        }
        *reinterpret_cast<uint2 *>(phi + m * ldk + n0) = *reinterpret_cast<uint2 *>(h4);
        if (plo) *reinterpret_cast<uint2 *>(plo + m * ldk + n0) = *reinterpret_cast<uint2 *>(l4);
    }
    if (row_sumsq) {
        ss = block_reduce_sum(ss, sh);
        if (threadIdx.x == 0) atomicAdd(row_sumsq + m, ss);
    }
    if (out_maxabs) {
        float mx = fmaxf(fmaxf(fabsf(v[0]), fabsf(v[1])), fmaxf(fabsf(v[2]), fabsf(v[3])));
        mx = block_reduce_max(mx, sh);
        if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned int *>(out_maxabs), __float_as_uint(mx));
    }
}

// planes only: global max |v| of a freshly written plane is not needed (error model uses weights')
int launch_splitk_reduce(Context *c, int M, int N, int splits, int Mpad, int Npad, const float *bias_m,
                         const float *bias_n, int relu, const ReduceOut &o) {
    Planes *p = o.planes;
    if (o.out_maxabs) NB_CUDA(cudaMemsetAsync(o.out_maxabs, 0, sizeof(float), c->stream));
    if (p) {
        NB_TRY(alloc_planes(p, M, N, o.want_lo));
        NB_CUDA(cudaMemsetAsync(p->row_sumsq, 0, static_cast<size_t>(M) * sizeof(float), c->stream));
    }
    const int64_t cover = p ? p->ldk : N;
    dim3 grid(static_cast<unsigned>(M), static_cast<unsigned>(ceil_div(cover, 1024)));
    k_splitk_reduce<<<grid, 256, 0, c->stream>>>(c->ws, static_cast<long long>(Mpad) * Npad, Npad, splits, M, N, bias_m,
                                                 bias_n, relu, o.f32, o.f32_ld, o.f64, o.f64_ld, p ? p->hi : nullptr,
                                                 (p && o.want_lo) ? p->lo : nullptr, p ? p->ldk : 0,
                                                 p ? p->row_sumsq : nullptr, o.exp_u, o.exp_v, o.out_maxabs);
    NB_LAUNCH_CHECK(c);
    if (p) { p->valid = true; p->lo_valid = o.want_lo; }
    return NB200_OK;
}


// ---------------------------------------------------------------------------------------------
// conv2d operand preparation for the implicit GEMM
// ---------------------------------------------------------------------------------------------
// fp64 NCHW -> bf16 hi/lo NHWC, channels zero padded to cp.  One block per (image, row): reads are coalesced along w
// per channel, writes along (w, c) through shared memory.
__global__ void __launch_bounds__(256)
k_nchw_to_nhwc(const double *__restrict__ img, int ch, int h, int w, int cp, __nv_bfloat16 *__restrict__ hi,
               __nv_bfloat16 *__restrict__ lo) {
    const int y = blockIdx.x % h, n = blockIdx.x / h;
    const long long total = static_cast<long long>(w) * cp;
    for (long long e = threadIdx.x; e < total; e += 256) {
        const int x = static_cast<int>(e / cp), c = static_cast<int>(e % cp);
        const double v = c < ch ? img[((static_cast<long long>(n) * ch + c) * h + y) * w + x] : 0.0;
        __nv_bfloat16 a, b;
        split_bf16<double>(v, a, b);
        const long long o = ((static_cast<long long>(n) * h + y) * w + x) * cp + c;
        hi[o] = a;
        if (lo) lo[o] = b;
    }
}
int launch_nchw_to_nhwc_planes(Context *c, const double *img, int n, int ch, int h, int w, int cp, __nv_bfloat16 *hi,
                               __nv_bfloat16 *lo) {
    k_nchw_to_nhwc<<<n * h, 256, 0, c->stream>>>(img, ch, h, w, cp, hi, lo);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
// kernels [kout][ch][k][k] fp64 -> planes [kout][k*k*cp], K index = (i*k + j)*cp + c  (tap major, channel minor)
__global__ void k_kernel_to_planes(const double *__restrict__ ker, int kout, int ch, int k, int cp,
                                   __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, long long ldk) {
    const long long total = static_cast<long long>(kout) * ldk;
    for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < total;
         e += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int o = static_cast<int>(e / ldk);
        const long long kk = e % ldk;
        const int tap = static_cast<int>(kk / cp), c = static_cast<int>(kk % cp);
        double v = 0.0;
        if (tap < k * k && c < ch) v = ker[((static_cast<long long>(o) * ch + c) * k + tap / k) * k + tap % k];
        __nv_bfloat16 a, b;
        split_bf16<double>(v, a, b);
        hi[e] = a;
        if (lo) lo[e] = b;
    }
}
int launch_kernel_to_planes(Context *c, const double *ker, int kout, int ch, int k, int cp, Planes *p, bool want_lo) {
    NB_TRY(alloc_planes(p, kout, static_cast<int64_t>(k) * k * cp, want_lo));
    k_kernel_to_planes<<<148, 256, 0, c->stream>>>(ker, kout, ch, k, cp, p->hi, want_lo ? p->lo : nullptr, p->ldk);
    NB_LAUNCH_CHECK(c);
    p->valid = true;
    p->lo_valid = want_lo;
    return NB200_OK;
}

// ---------------------------------------------------------------------------------------------
// small utilities
// ---------------------------------------------------------------------------------------------
__global__ void k_transpose_f32(const float *__restrict__ src, long long rows, long long cols, long long ld,
                                float *__restrict__ dst, long long dld) {
    __shared__ float tile[32][33];
    const long long r0 = static_cast<long long>(blockIdx.y) * 32, c0 = static_cast<long long>(blockIdx.x) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 32; i += 8)
        if (r0 + i < rows && c0 + tx < cols) tile[i][tx] = src[(r0 + i) * ld + c0 + tx];
    __syncthreads();
    for (int i = ty; i < 32; i += 8)
        if (c0 + i < cols && r0 + tx < rows) dst[(c0 + i) * dld + r0 + tx] = tile[tx][i];
}
int launch_transpose_f32(Context *c, const float *src, int64_t rows, int64_t cols, int64_t ld, float *dst, int64_t dld) {
    if (rows == 0 || cols == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 32)), static_cast<unsigned>(ceil_div(rows, 32)));
    k_transpose_f32<<<grid, 256, 0, c->stream>>>(src, rows, cols, ld, dst, dld);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// y = max(0, x + bias[col]) in place on an fp32 matrix (FFReluBiasSum.h:68-71 on the reduced layer-1 partials)
__global__ void k_bias_relu_f32(float *__restrict__ p, long long rows, long long cols, long long ld,
                                const float *__restrict__ bias_n) {
    const long long n = rows * cols;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / cols, cc = i % cols;
        p[r * ld + cc] = fmaxf(p[r * ld + cc] + bias_n[cc], 0.f);
    }
}
int launch_bias_relu_f32(Context *c, float *p, int64_t rows, int64_t cols, int64_t ld, const float *bias_n) {
    const int64_t n = rows * cols;
    if (n == 0) return NB200_OK;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    k_bias_relu_f32<<<blocks, 256, 0, c->stream>>>(p, rows, cols, ld, bias_n);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

__global__ void k_f64_to_f32(const double *__restrict__ s, long long n, float *__restrict__ d) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
        d[i] = static_cast<float>(s[i]);
}
int launch_f64_to_f32(Context *c, const double *src, int64_t n, float *dst) {
    if (n == 0) return NB200_OK;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    k_f64_to_f32<<<blocks, 256, 0, c->stream>>>(src, n, dst);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// LSTMThreeWaySum / LSTMTwoSum (src/LSTM/headers/LSTMThreeWaySum.h:76-95, LSTMTwoSum.h:58-63): out = act(a + b + c),
// element-wise over equally shaped sets; every source is fp64 or fp32 (results of the tensor-core GEMMs), the sum and
// the activation are evaluated in fp64 with the reference's own formulas.  HBM-bound: (8 or 4) x sources + 8 bytes
// per element.
__device__ __forceinline__ double ew_load(const EwiseSrc &s, long long r, long long c) {
    return s.d ? s.d[r * s.ld + c] : static_cast<double>(s.f[r * s.ld + c]);
}
__global__ void __launch_bounds__(256)
k_sum_blocks(EwiseSrc a, EwiseSrc b, EwiseSrc c, int nsrc, int act, long long rows, long long cols, double *__restrict__ out) {
    const long long n = rows * cols;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / cols, col = i - r * cols;
        double sum = ew_load(a, r, col);
        if (nsrc > 1) sum = sum + ew_load(b, r, col);
        if (nsrc > 2) sum = sum + ew_load(c, r, col);
        double v = sum;
        if (act == NB200_ACT_SIGMOID) {
            v = 1.0 / (1.0 + exp(-1.0 * sum));                       // LSTMThreeWaySum.h:81
        } else if (act == NB200_ACT_TANH) {
            // :85-86 (exp(s) - exp(-s)) / (exp(s) + exp(-s)); that quotient is inf/inf beyond |s| = 709, where the
            // function it spells is +-1
            if (fabs(sum) > 350.0) v = copysign(1.0, sum);
            else { const double ep = exp(sum), em = exp(-1.0 * sum); v = (ep - em) / (ep + em); }
        }
        out[i] = v;
    }
}
int launch_sum_blocks(Context *c, const EwiseSrc *src, int nsrc, int act, int64_t rows, int64_t cols, double *out) {
    const int64_t n = rows * cols;
    if (n == 0) return NB200_OK;
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    const int blocks = static_cast<int>(ceil_div(n, 256) < sms * 8 ? ceil_div(n, 256) : sms * 8);
    EwiseSrc z{};
    k_sum_blocks<<<blocks, 256, 0, c->stream>>>(src[0], nsrc > 1 ? src[1] : z, nsrc > 2 ? src[2] : z, nsrc, act, rows, cols, out);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// Seeded generator; the test suite's CPU generator is defined to match it bit for bit (explicit _rn intrinsics
// keep the compiler from contracting the multiply-add, which would change the rounding).
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void k_synth_fill(double *__restrict__ dst, long long rows, long long cols, long long ld,
                             unsigned long long seed_mixed, long long row_offset, long long total_cols) {
    const long long n = rows * cols;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / cols, cc = i % cols;
        const unsigned long long idx = static_cast<unsigned long long>((row_offset + r) * total_cols + cc);
        const unsigned long long h = splitmix64(seed_mixed ^ (idx * 0xD1342543DE82EF95ull + 1ull));
        const double u = __dmul_rn(static_cast<double>(h >> 11), 1.0 / 9007199254740992.0);
        double mag = __dmul_rn(u, 0.5 - 0.0001);
        mag = __dadd_rn(mag, 0.0001);
        dst[r * ld + cc] = (h & 1ull) ? -mag : mag;
    }
}
static unsigned long long host_splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
int launch_synth_fill(Context *c, double *dst, int64_t rows, int64_t cols, int64_t ld, uint64_t seed,
                      int64_t row_offset, int64_t total_cols) {
    if (rows == 0 || cols == 0) return NB200_OK;
    k_synth_fill<<<148 * 8, 256, 0, c->stream>>>(dst, rows, cols, ld, host_splitmix64(seed), row_offset, total_cols);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// ---------------------------------------------------------------------------------------------
// fp64 GEMM (exact path): P[M x N] += U[M x K] . V^T, V(n,k) = v[n*v_sn + k*v_sk]
// 256 threads, 4 x 4 register micro-tiles, BK = 16, split-K with fp64 atomics into a zeroed P.
// ---------------------------------------------------------------------------------------------
template <int BM, int BN>
__global__ void __launch_bounds__(256)
k_gemm_f64(const double *__restrict__ u, long long u_sm, long long u_sk, const double *__restrict__ v, long long v_sn,
           long long v_sk, int M, int N, long long K, long long k_per_split, double *__restrict__ out, long long out_ld) {
    constexpr int BK = 16;
    static_assert(BM * BN == 4096, "256 threads x 16 outputs");
    __shared__ double su[BK][BM + 1];
    __shared__ double sv[BK][BN + 1];
    constexpr int TM = BM / 4;  // thread groups along m
    const int tid = threadIdx.x;
    const int tm = tid % TM, tn = tid / TM;  // micro-tile origin (tm*4, tn*4)
    const long long m0 = static_cast<long long>(blockIdx.y) * BM, n0 = static_cast<long long>(blockIdx.x) * BN;
    const long long kbeg = static_cast<long long>(blockIdx.z) * k_per_split;
    const long long kend = (kbeg + k_per_split < K) ? kbeg + k_per_split : K;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
    for (long long k0 = kbeg; k0 < kend; k0 += BK) {
        if (u_sk == 1) {  // k fastest in memory
            for (int e = tid; e < BM * BK; e += 256) {
                const int kk = e % BK, mm = e / BK;
                const long long gm = m0 + mm, gk = k0 + kk;
                su[kk][mm] = (gm < M && gk < kend) ? u[gm * u_sm + gk] : 0.0;
            }
        } else {  // m fastest in memory
            for (int e = tid; e < BM * BK; e += 256) {
                const int mm = e % BM, kk = e / BM;
                const long long gm = m0 + mm, gk = k0 + kk;
                su[kk][mm] = (gm < M && gk < kend) ? u[gm * u_sm + gk * u_sk] : 0.0;
            }
        }
        if (v_sk == 1) {
            for (int e = tid; e < BN * BK; e += 256) {
                const int kk = e % BK, nn = e / BK;
                const long long gn = n0 + nn, gk = k0 + kk;
                sv[kk][nn] = (gn < N && gk < kend) ? v[gn * v_sn + gk] : 0.0;
            }
        } else {  // n fastest in memory
            for (int e = tid; e < BN * BK; e += 256) {
                const int nn = e % BN, kk = e / BN;
                const long long gn = n0 + nn, gk = k0 + kk;
                sv[kk][nn] = (gn < N && gk < kend) ? v[gn * v_sn + gk * v_sk] : 0.0;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; kk++) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; i++) a[i] = su[kk][tm * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; j++) b[j] = sv[kk][tn * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const long long gm = m0 + tm * 4 + i, gn = n0 + tn * 4 + j;
            if (gm < M && gn < N) {
                if (gridDim.z == 1) out[gm * out_ld + gn] = acc[i][j];
                else out[(static_cast<long long>(blockIdx.z) * M + gm) * out_ld + gn] = acc[i][j];   // split slab (summed in order by k_f64_sum_splits)
            }
        }
}

// ---------------------------------------------------------------------------------------------
// fp64 GEMM on the DMMA tensor path (mma.sync.m8n8k4.f64): P[M x N] (+)= U[M x K] . V[N x K]^T with
// both operands K-contiguous.  Used to re-evaluate near-tie rows (K9): M is the handful of flagged
// rows, N the hidden width, K the full feature count, so the grid is (N tiles, M tiles, K splits)
// and every split stores its partial into its own slab, summed in split order by k_f64_sum_splits.
// Fragment layout of m8n8k4 (row.col): lane t holds A[t/4][t%4], B[t%4][t/4], C[t/4][2*(t%4) + {0,1}].
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// m16n8k4 (sm_90+): lane t, g = t / 4, q = t % 4 holds A[g][q], A[g+8][q]; B[q][g]; C[g][2q + {0,1}], C[g+8][2q + {0,1}].
__device__ __forceinline__ void dmma_m16n8k4(double (&cfrag)[4], double a0, double a1, double b) {
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
                 : "+d"(cfrag[0]), "+d"(cfrag[1]), "+d"(cfrag[2]), "+d"(cfrag[3])
                 : "d"(a0), "d"(a1), "d"(b));
}

// 128 threads = 4 warps arranged WARPS_M x WARPS_N, each warp computes a (8*MT) x 32 tile.
template <int MT, int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(128)
k_gemm_f64_mma(const double *__restrict__ u, long long u_ld, const double *__restrict__ v, long long v_ld, int M, int N,
               long long K, long long k_per_split, double *__restrict__ out, long long out_ld) {
    constexpr int BM = 8 * MT * WARPS_M, BN = 32 * WARPS_N, BK = 16, LDS = BK + 4;  // +4: conflict-free fragment reads
    static_assert(WARPS_M * WARPS_N == 4, "4 warps");
    __shared__ double su[2][BM * LDS];
    __shared__ double sv[2][BN * LDS];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp / WARPS_N, wn = warp % WARPS_N;
    const long long m0 = static_cast<long long>(blockIdx.y) * BM, n0 = static_cast<long long>(blockIdx.x) * BN;
    const long long kbeg = static_cast<long long>(blockIdx.z) * k_per_split;
    const long long kend = (kbeg + k_per_split < K) ? kbeg + k_per_split : K;
    double acc[MT][4][2];
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    constexpr int U_ELEMS = BM * BK / 128, V_ELEMS = BN * BK / 128;  // doubles per thread per stage
    double ru[U_ELEMS], rv[V_ELEMS];
    auto gload = [&](long long k0) {
#pragma unroll
        for (int e = 0; e < U_ELEMS; e++) {
            const int idx = e * 128 + tid, r = idx / BK, kk = idx % BK;
            const long long gm = m0 + r, gk = k0 + kk;
            ru[e] = (gm < M && gk < kend) ? u[gm * u_ld + gk] : 0.0;
        }
#pragma unroll
        for (int e = 0; e < V_ELEMS; e++) {
            const int idx = e * 128 + tid, r = idx / BK, kk = idx % BK;
            const long long gn = n0 + r, gk = k0 + kk;
            rv[e] = (gn < N && gk < kend) ? v[gn * v_ld + gk] : 0.0;
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int e = 0; e < U_ELEMS; e++) {
            const int idx = e * 128 + tid;
            su[buf][(idx / BK) * LDS + idx % BK] = ru[e];
        }
#pragma unroll
        for (int e = 0; e < V_ELEMS; e++) {
            const int idx = e * 128 + tid;
            sv[buf][(idx / BK) * LDS + idx % BK] = rv[e];
        }
    };
    int buf = 0;
    if (kbeg < kend) {
        gload(kbeg);
        sstore(0);
    }
    __syncthreads();
    for (long long k0 = kbeg; k0 < kend; k0 += BK) {
        const bool more = k0 + BK < kend;
        if (more) gload(k0 + BK);  // global loads in flight while the tensor pipe works
        const double *pu = su[buf] + (wm * 8 * MT + lane / 4) * LDS + lane % 4;
        const double *pv = sv[buf] + (wn * 32 + lane / 4) * LDS + lane % 4;
#pragma unroll
        for (int ks = 0; ks < BK; ks += 4) {
            double a[MT], b[4];
#pragma unroll
            for (int i = 0; i < MT; i++) a[i] = pu[i * 8 * LDS + ks];
#pragma unroll
            for (int j = 0; j < 4; j++) b[j] = pv[j * 8 * LDS + ks];
#pragma unroll
            for (int i = 0; i < MT; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) dmma_m8n8k4(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        if (more) sstore(buf ^ 1);
        __syncthreads();
        buf ^= 1;
    }
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const long long gm = m0 + wm * 8 * MT + i * 8 + lane / 4;
                const long long gn = n0 + wn * 32 + j * 8 + 2 * (lane % 4) + e;
                if (gm < M && gn < N) {
                    if (gridDim.z == 1) out[gm * out_ld + gn] = acc[i][j][e];
                    else out[(static_cast<long long>(blockIdx.z) * M + gm) * out_ld + gn] = acc[i][j][e];
                }
            }
}

// Skinny variant for the last cascade level (a handful of rows against all of W1): the kernel is a stream over V, so
// what matters is bytes in flight.  16 x 128 tile, operands brought in by 16-byte cp.async into a 3-stage
// shared-memory ring (zero fill past the ends), no register staging; sized for 3 CTAs per SM.  Needs 16-byte aligned rows (even pitches).
__device__ __forceinline__ void cp_async16_zfill(void *smem_dst, const void *gsrc, int src_bytes) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
template <int STAGES, int BK, int BN>
__global__ void __launch_bounds__(128)
k_gemm_f64_mma_skinny(const double *__restrict__ u, long long u_ld, const double *__restrict__ v, long long v_ld, int M, int N,
                      long long K, long long k_per_split, double *__restrict__ out, long long out_ld) {
    constexpr int BM = 16, LDS = BK + 4, PIECES = BK / 2;  // 16-byte pieces per row
    constexpr int NT = BN / 32;                            // n8 tiles per warp (4 warps side by side)
    extern __shared__ __align__(16) double dyn_smem[];
    double *su = dyn_smem;                         // [STAGES][BM * LDS]
    double *sv = dyn_smem + STAGES * BM * LDS;     // [STAGES][BN * LDS]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long n0 = static_cast<long long>(blockIdx.x) * BN, m0 = static_cast<long long>(blockIdx.y) * BM;
    const long long kbeg = static_cast<long long>(blockIdx.z) * k_per_split;
    const long long kend = (kbeg + k_per_split < K) ? kbeg + k_per_split : K;
    const int nk = kend > kbeg ? static_cast<int>((kend - kbeg + BK - 1) / BK) : 0;
    double acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) acc[j][e] = 0.0;
    // one stage = (BM + BN) rows x PIECES pieces of 16 bytes
    auto issue = [&](int kb, int stage) {
        const long long k0 = kbeg + static_cast<long long>(kb) * BK;
#pragma unroll
        for (int e = 0; e < ((BM + BN) * PIECES + 127) / 128; e++) {
            const int idx = e * 128 + tid, r = idx / PIECES, piece = idx % PIECES;
            if (r >= BM + BN) break;
            const long long gk = k0 + piece * 2;
            int bytes = 0;
            if (gk < kend) bytes = (gk + 1 < kend) ? 16 : 8;
            if (r < BM) {
                const long long gm = m0 + r;
                if (gm >= M) bytes = 0;
                cp_async16_zfill(su + (stage * BM + r) * LDS + piece * 2, u + (bytes ? gm * u_ld + gk : 0), bytes);
            } else {
                const long long gn = n0 + (r - BM);
                if (gn >= N) bytes = 0;
                cp_async16_zfill(sv + (stage * BN + (r - BM)) * LDS + piece * 2, v + (bytes ? gn * v_ld + gk : 0), bytes);
            }
        }
    };
    for (int s = 0; s < STAGES - 1; s++) {
        if (s < nk) issue(s, s);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int kb = 0; kb < nk; kb++) {
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2) : "memory");
        __syncthreads();  // stage kb has landed for everyone; stage kb-1's readers are done
        if (kb + STAGES - 1 < nk) issue(kb + STAGES - 1, (kb + STAGES - 1) % STAGES);
        asm volatile("cp.async.commit_group;" ::: "memory");
        const int st = kb % STAGES;
        const double *pu = su + (st * BM + lane / 4) * LDS + lane % 4;
        const double *pv = sv + (st * BN + warp * (8 * NT) + lane / 4) * LDS + lane % 4;
#pragma unroll
        for (int ks = 0; ks < BK; ks += 4) {
            const double a0 = pu[ks], a1 = pu[8 * LDS + ks];
            double b[NT];
#pragma unroll
            for (int j = 0; j < NT; j++) b[j] = pv[j * 8 * LDS + ks];
#pragma unroll
            for (int j = 0; j < NT; j++) dmma_m16n8k4(acc[j], a0, a1, b[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < NT; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const long long gm = m0 + (e >> 1) * 8 + lane / 4;
            const long long gn = n0 + warp * (8 * NT) + j * 8 + 2 * (lane % 4) + (e & 1);
            if (gm < M && gn < N) {
                if (gridDim.z == 1) out[gm * out_ld + gn] = acc[j][e];
                else out[(static_cast<long long>(blockIdx.z) * M + gm) * out_ld + gn] = acc[j][e];
            }
        }
}

// The same product with the operands brought in through REGISTERS: 256-bit loads (ld.global.nc.L1::no_allocate.v4.f64,
// a warp instruction = 256 contiguous bytes of four rows) land in registers one stage ahead and are stored to a
// double-buffered shared-memory tile after the multiplies of the current stage -- the access pattern the layer-1 stream kernel
// reads its fp64 inputs with (6.1-6.4 TB/s in tools/hbm_pattern_bench.cu), where the cp.async version stops at 4.7 TB/s.
// 16 x 128 tile, 32 doubles of K per stage, 8 warps (each 16 columns).  Needs rows whose pitch is a multiple of 4 doubles,
// 32-byte aligned bases and K % 4 == 0.
// Measured on the Amazon shape (1,000 x 597,540): with all 16 rows of the tile the kernel sits on the DMMA pipe -- 1.007 ms =
// 19.4 TFLOP/s of DMMA.8x8x4, the only fp64 tensor primitive of sm_100 (m16n8k4 lowers to two of them) -- like the cp.async
// version (1.025 ms); HALF (at most 8 rows: one DMMA.8x8x4 per fragment and step) takes 0.838 ms = 5.7 TB/s of W1, which the
// cp.async version never reached whatever the DMMA work (4.7 TB/s).
template <bool HALF>
__global__ void __launch_bounds__(256, 2)
k_gemm_f64_mma_skinny_ldg(const double *__restrict__ u, long long u_ld, const double *__restrict__ v, long long v_ld, int M, int N,
                          long long K, long long k_per_split, double *__restrict__ out, long long out_ld) {
    constexpr int BM = 16, BN = 128, BK = 32, LDS = BK + 4;
    extern __shared__ __align__(16) double dyn_smem[];
    double *su = dyn_smem;                       // [2][BM * LDS]
    double *sv = dyn_smem + 2 * BM * LDS;        // [2][BN * LDS]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long n0 = static_cast<long long>(blockIdx.x) * BN;
    const long long kbeg = static_cast<long long>(blockIdx.z) * k_per_split;
    const long long kend = (kbeg + k_per_split < K) ? kbeg + k_per_split : K;
    const int nk = kend > kbeg ? static_cast<int>((kend - kbeg + BK - 1) / BK) : 0;
    // loads: lane = (row in group of 4, 32-byte piece of the 256-byte row segment)
    const int lrow = lane >> 3, piece = lane & 7;
    const double *vsrc[4];
    uint32_t vdst[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int r = warp * 16 + i * 4 + lrow;
        const long long gn = n0 + r;
        vsrc[i] = gn < N ? v + gn * v_ld + kbeg + piece * 4 : nullptr;
        vdst[i] = static_cast<uint32_t>(__cvta_generic_to_shared(sv + r * LDS + piece * 4));
    }
    const int ur = warp * 4 + lrow;              // warps 0..3 also bring in the U tile
    const double *usrc = (warp < 4 && ur < M) ? u + static_cast<long long>(ur) * u_ld + kbeg + piece * 4 : nullptr;
    const uint32_t udst = static_cast<uint32_t>(__cvta_generic_to_shared(su + ur * LDS + piece * 4));
    double rv[4][4], ru[4];
    auto load = [&](int kb) {
        const long long koff = static_cast<long long>(kb) * BK;
        const bool kin = kbeg + koff + piece * 4 < kend;   // K % 4 == 0: a piece is inside or outside as a whole
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (vsrc[i] && kin)
                asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0, %1, %2, %3}, [%4];"
                             : "=d"(rv[i][0]), "=d"(rv[i][1]), "=d"(rv[i][2]), "=d"(rv[i][3]) : "l"(vsrc[i] + koff));
            else rv[i][0] = rv[i][1] = rv[i][2] = rv[i][3] = 0.0;
        }
        if (usrc && kin)
            asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(ru[0]), "=d"(ru[1]), "=d"(ru[2]), "=d"(ru[3]) : "l"(usrc + koff));
        else ru[0] = ru[1] = ru[2] = ru[3] = 0.0;
    };
    auto store = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t d = vdst[i] + buf * (BN * LDS * 8);
            asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(d), "d"(rv[i][0]), "d"(rv[i][1]) : "memory");
            asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(d + 16), "d"(rv[i][2]), "d"(rv[i][3]) : "memory");
        }
        if (warp < 4) {
            const uint32_t d = udst + buf * (BM * LDS * 8);
            asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(d), "d"(ru[0]), "d"(ru[1]) : "memory");
            asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(d + 16), "d"(ru[2]), "d"(ru[3]) : "memory");
        }
    };
    double acc[2][4];
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) acc[j][e] = 0.0;
    if (nk > 0) {
        load(0);
        store(0);
        if (nk > 1) load(1);
    }
    __syncthreads();
    for (int kb = 0; kb < nk; kb++) {
        const int buf = kb & 1;
        const double *pu = su + (buf * BM + lane / 4) * LDS + lane % 4;
        const double *pv = sv + (buf * BN + warp * 16 + lane / 4) * LDS + lane % 4;
#pragma unroll
        for (int ks = 0; ks < BK; ks += 4) {
            const double a0 = pu[ks], a1 = HALF ? 0.0 : pu[8 * LDS + ks];
            const double b0 = pv[ks], b1 = pv[8 * LDS + ks];
            if (HALF) {
                dmma_m8n8k4(acc[0][0], acc[0][1], a0, b0);
                dmma_m8n8k4(acc[1][0], acc[1][1], a0, b1);
            } else {
                dmma_m16n8k4(acc[0], a0, a1, b0);
                dmma_m16n8k4(acc[1], a0, a1, b1);
            }
        }
        if (kb + 1 < nk) {
            store(buf ^ 1);                 // stage kb + 1 (in registers since the previous iteration) -> the other buffer,
            if (kb + 2 < nk) load(kb + 2);  // whose readers finished before the barrier that ended iteration kb - 1
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const long long gm = (e >> 1) * 8 + lane / 4;
            const long long gn = n0 + warp * 16 + j * 8 + 2 * (lane % 4) + (e & 1);
            if (gm < M && gn < N) {
                if (gridDim.z == 1) out[gm * out_ld + gn] = acc[j][e];
                else out[(static_cast<long long>(blockIdx.z) * M + gm) * out_ld + gn] = acc[j][e];
            }
        }
}

// Split-K of the fp64 GEMMs: every split stores its partial tile into its own slab ws[split][M][N]; this kernel sums the
// slabs in split order (the result does not depend on which CTA finished first, as fp64 atomics would make it) and applies
// the bias / ReLU epilogue.
__global__ void k_f64_sum_splits(const double *__restrict__ ws, int splits, int M, int N, double *__restrict__ out, long long out_ld,
                                 const double *__restrict__ bias_m, const double *__restrict__ bias_n, int relu) {
    const long long n = static_cast<long long>(M) * N;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        double x = 0.0;
        for (int s = 0; s < splits; s++) x += ws[s * n + i];
        const long long r = i / N, cc = i % N;
        if (bias_m) x += bias_m[r];
        if (bias_n) x += bias_n[cc];
        if (relu) x = x > 0.0 ? x : 0.0;
        out[r * out_ld + cc] = x;
    }
}
static int sum_splits(Context *c, double *ws, int splits, int M, int N, double *out, int64_t out_ld, const F64Epilogue &e) {
    const int64_t n = static_cast<int64_t>(M) * N;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    k_f64_sum_splits<<<blocks, 256, 0, c->stream>>>(ws, splits, M, N, out, out_ld, e.bias_m, e.bias_n, e.relu);
    NB_LAUNCH_CHECK(c);
    nb_free(ws);   // stream-ordered reuse: the pool hands the block out again only to work queued after this kernel
    return NB200_OK;
}
__global__ void k_bias_relu_f64(double *__restrict__ p, int M, int N, long long ld, const double *__restrict__ bias_m,
                                const double *__restrict__ bias_n, int relu) {
    const long long n = static_cast<long long>(M) * N;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / N, cc = i % N;
        double x = p[r * ld + cc];
        if (bias_m) x += bias_m[r];
        if (bias_n) x += bias_n[cc];
        if (relu) x = x > 0.0 ? x : 0.0;
        p[r * ld + cc] = x;
    }
}
int launch_gemm_f64(Context *c, const double *u, int64_t u_sm, int64_t u_sk, const double *v, int64_t v_sn,
                    int64_t v_sk, int M, int N, int64_t K, double *out, int64_t out_ld, const F64Epilogue &e) {
    if (M == 0 || N == 0) return NB200_OK;
    if (u_sk == 1 && v_sk == 1) {  // both K-contiguous: tensor (DMMA) path
        const bool small = M <= 16;
        const int BMm = small ? 16 : 64, BNm = small ? 128 : 64;
        const int64_t tiles = ceil_div(M, BMm) * ceil_div(N, BNm);
        const int sms = c->sm_count > 0 ? c->sm_count : 148;
        int splits = 1;
        if (tiles < 4 * sms && K >= 4096) {
            splits = static_cast<int>(ceil_div(4 * sms, tiles));
            const int64_t maxs = K / 2048;
            if (splits > maxs) splits = static_cast<int>(maxs);
            if (splits < 1) splits = 1;
        }
        // skinny + aligned: the cp.async streaming variant, 2 CTAs per SM, as many K splits as fill them
        static const bool skinny_env = !(getenv("NB200_F64_SKINNY") && atoi(getenv("NB200_F64_SKINNY")) == 0);
        const bool aligned = (u_sm % 2 == 0) && (v_sn % 2 == 0) && ((reinterpret_cast<uintptr_t>(u) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
        if (small && aligned && skinny_env && K >= 8192) {
            static const int cfg = getenv("NB200_F64_CFG") ? atoi(getenv("NB200_F64_CFG")) : 1;
            bool epilogue_done = false;
            static const bool ldg_env = !(getenv("NB200_F64_LDG") && atoi(getenv("NB200_F64_LDG")) == 0);   // default on
            if (ldg_env && (u_sm % 4 == 0) && (v_sn % 4 == 0) && (K % 4 == 0) &&
                ((reinterpret_cast<uintptr_t>(u) | reinterpret_cast<uintptr_t>(v)) & 31) == 0) {
                const size_t smem = static_cast<size_t>(2) * (16 + 128) * 36 * sizeof(double);
                NB_CUDA(cudaFuncSetAttribute(k_gemm_f64_mma_skinny_ldg<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
                NB_CUDA(cudaFuncSetAttribute(k_gemm_f64_mma_skinny_ldg<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
                NB_CUDA(cudaFuncSetAttribute(k_gemm_f64_mma_skinny_ldg<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 74));
                NB_CUDA(cudaFuncSetAttribute(k_gemm_f64_mma_skinny_ldg<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 74));
                const int64_t tl = ceil_div(N, 128);
                int sp = static_cast<int>(ceil_div(static_cast<int64_t>(2) * sms, tl));
                if (sp > K / 2048) sp = static_cast<int>(K / 2048);
                if (sp < 1) sp = 1;
                const int64_t kp = round_up(ceil_div(K, sp), 32);
                sp = static_cast<int>(ceil_div(K, kp));
                double *ws = nullptr;
                if (sp > 1) NB_CUDA(nb_malloc(&ws, static_cast<size_t>(sp) * M * N * sizeof(double)));
                dim3 g2(static_cast<unsigned>(tl), 1, sp);
                if (M <= 8) k_gemm_f64_mma_skinny_ldg<true><<<g2, 256, smem, c->stream>>>(u, u_sm, v, v_sn, M, N, K, kp, ws ? ws : out, ws ? N : out_ld);
                else k_gemm_f64_mma_skinny_ldg<false><<<g2, 256, smem, c->stream>>>(u, u_sm, v, v_sn, M, N, K, kp, ws ? ws : out, ws ? N : out_ld);
                NB_LAUNCH_CHECK(c);
                if (ws) return sum_splits(c, ws, sp, M, N, out, out_ld, e);
                if (e.bias_m || e.bias_n || e.relu) {
                    const int64_t n = static_cast<int64_t>(M) * N;
                    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
                    k_bias_relu_f64<<<blocks, 256, 0, c->stream>>>(out, M, N, out_ld, e.bias_m, e.bias_n, e.relu);
                    NB_LAUNCH_CHECK(c);
                }
                return NB200_OK;
            }
            auto run = [&](auto kern, int BNs, int BKs, int ST, int ctas_per_sm) -> int {
                const size_t smem = static_cast<size_t>(ST) * (16 + BNs) * (BKs + 4) * sizeof(double);
                NB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
                // What the resident CTAs leave of the SM's 228 KB is the L1, and the L1 is where global loads land: three
                // 69 KB CTAs leave none and the stream over W1 stops at ~4.2 TB/s.  Two CTAs per SM (carve-out 164 KB,
                // 64 KB of L1) stream faster than three.
                static const int carve = getenv("NB200_F64_CARVEOUT") ? atoi(getenv("NB200_F64_CARVEOUT")) : 72;
                if (carve > 0) NB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
                static const int cps_env = getenv("NB200_F64_CTAS") ? atoi(getenv("NB200_F64_CTAS")) : 0;
                if (cps_env > 0) ctas_per_sm = cps_env;
                else if (carve > 0) ctas_per_sm = static_cast<int>((228 * 1024 * carve / 100) / (smem + 1024));
                const int64_t tl = ceil_div(M, 16) * ceil_div(N, BNs);
                int sp = static_cast<int>(ceil_div(static_cast<int64_t>(ctas_per_sm) * sms, tl));
                if (sp > K / 2048) sp = static_cast<int>(K / 2048);
                if (sp < 1) sp = 1;
                const int64_t kp = round_up(ceil_div(K, sp), BKs);
                sp = static_cast<int>(ceil_div(K, kp));
                double *ws = nullptr;
                if (sp > 1) NB_CUDA(nb_malloc(&ws, static_cast<size_t>(sp) * M * N * sizeof(double)));
                dim3 g2(static_cast<unsigned>(ceil_div(N, BNs)), static_cast<unsigned>(ceil_div(M, 16)), sp);
                kern<<<g2, 128, smem, c->stream>>>(u, u_sm, v, v_sn, M, N, K, kp, ws ? ws : out, ws ? N : out_ld);
                NB_LAUNCH_CHECK(c);
                if (ws) { epilogue_done = true; return sum_splits(c, ws, sp, M, N, out, out_ld, e); }
                return NB200_OK;
            };
            // measured on the Amazon shape (15 x 1000 x 597,540, tools/f64_skinny_bench.py, profiles/r02_f64_skinny.txt):
            // with three CTAs per SM and no L1 left 16 x 64 / BK = 32 takes 1.13 ms, 16 x 128 / BK = 16 1.15 ms; with a
            // 72 % carve-out (two CTAs per SM, 64 KB of L1) 16 x 128 / BK = 16 takes 1.01 ms = 4.7 TB/s of W1 and
            // 19 TFLOP/s of DMMA.8x8x4 (m16n8k4 lowers to two of them); the HBM roofline is 0.73 ms.
            if (cfg == 1) NB_TRY(run(k_gemm_f64_mma_skinny<3, 16, 128>, 128, 16, 3, 3));
            else NB_TRY(run(k_gemm_f64_mma_skinny<3, 32, 64>, 64, 32, 3, 3));
            if (!epilogue_done && (e.bias_m || e.bias_n || e.relu)) {
                const int64_t n = static_cast<int64_t>(M) * N;
                const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
                k_bias_relu_f64<<<blocks, 256, 0, c->stream>>>(out, M, N, out_ld, e.bias_m, e.bias_n, e.relu);
                NB_LAUNCH_CHECK(c);
            }
            return NB200_OK;
        }
        const int64_t kper = round_up(ceil_div(K, splits), 16);
        splits = static_cast<int>(ceil_div(K, kper));
        double *ws = nullptr;
        if (splits > 1) NB_CUDA(nb_malloc(&ws, static_cast<size_t>(splits) * M * N * sizeof(double)));
        dim3 grid(static_cast<unsigned>(ceil_div(N, BNm)), static_cast<unsigned>(ceil_div(M, BMm)), splits);
        if (small) k_gemm_f64_mma<2, 1, 4><<<grid, 128, 0, c->stream>>>(u, u_sm, v, v_sn, M, N, K, kper, ws ? ws : out, ws ? N : out_ld);
        else k_gemm_f64_mma<4, 2, 2><<<grid, 128, 0, c->stream>>>(u, u_sm, v, v_sn, M, N, K, kper, ws ? ws : out, ws ? N : out_ld);
        NB_LAUNCH_CHECK(c);
        if (ws) return sum_splits(c, ws, splits, M, N, out, out_ld, e);
        if (e.bias_m || e.bias_n || e.relu) {
            const int64_t n = static_cast<int64_t>(M) * N;
            const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
            k_bias_relu_f64<<<blocks, 256, 0, c->stream>>>(out, M, N, out_ld, e.bias_m, e.bias_n, e.relu);
            NB_LAUNCH_CHECK(c);
        }
        return NB200_OK;
    }
    const bool skinny = M <= 16;
    const int BM = skinny ? 16 : 64, BN = skinny ? 256 : 64;
    const int64_t tiles = ceil_div(M, BM) * ceil_div(N, BN);
    int splits = 1;
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    if (tiles < 2 * sms && K >= 2048) {
        splits = static_cast<int>(ceil_div(2 * sms, tiles));
        const int64_t maxs = K / 1024;
        if (splits > maxs) splits = static_cast<int>(maxs);
        if (splits < 1) splits = 1;
    }
    int64_t kper = round_up(ceil_div(K, splits), 16);
    splits = static_cast<int>(ceil_div(K, kper));
    double *ws = nullptr;
    if (splits > 1) NB_CUDA(nb_malloc(&ws, static_cast<size_t>(splits) * M * N * sizeof(double)));
    dim3 grid(static_cast<unsigned>(ceil_div(N, BN)), static_cast<unsigned>(ceil_div(M, BM)), splits);
    if (skinny) k_gemm_f64<16, 256><<<grid, 256, 0, c->stream>>>(u, u_sm, u_sk, v, v_sn, v_sk, M, N, K, kper, ws ? ws : out, ws ? N : out_ld);
    else k_gemm_f64<64, 64><<<grid, 256, 0, c->stream>>>(u, u_sm, u_sk, v, v_sn, v_sk, M, N, K, kper, ws ? ws : out, ws ? N : out_ld);
    NB_LAUNCH_CHECK(c);
    if (ws) return sum_splits(c, ws, splits, M, N, out, out_ld, e);
    if (e.bias_m || e.bias_n || e.relu) {
        const int64_t n = static_cast<int64_t>(M) * N;
        const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
        k_bias_relu_f64<<<blocks, 256, 0, c->stream>>>(out, M, N, out_ld, e.bias_m, e.bias_n, e.relu);
        NB_LAUNCH_CHECK(c);
    }
    return NB200_OK;
}

// ---------------------------------------------------------------------------------------------
// row softmax / argmax / top-2 gap
// ---------------------------------------------------------------------------------------------
struct Top2 {
    float m1, m2;
    int i1;
};
__device__ __forceinline__ Top2 top2_merge(Top2 a, Top2 b) {
    // first maximum wins on ties (smaller index), as a sequential scan would
    Top2 r;
    if (b.m1 > a.m1 || (b.m1 == a.m1 && b.i1 < a.i1)) {
        r.m1 = b.m1; r.i1 = b.i1; r.m2 = fmaxf(a.m1, b.m2);
    } else {
        r.m1 = a.m1; r.i1 = a.i1; r.m2 = fmaxf(a.m2, b.m1);
    }
    return r;
}
template <typename T> struct TopD { T m1, m2; int i1; };

// One block per row.  T = float (tensor-core logits) or double (exact path).
// out[i,j] = exp(z - max) / sum_j exp(z - max)   (== exp(z)/sum exp(z) of FFOutputLayer, stable form)
template <typename T>
__global__ void __launch_bounds__(256)
k_softmax_rows(const T *__restrict__ z, long long ld, int cols, double *__restrict__ probs, long long probs_ld,
               int32_t *__restrict__ labels, float *__restrict__ gap, const int32_t *__restrict__ row_list,
               int in_compact) {
    __shared__ T s_m1[8], s_m2[8];
    __shared__ int s_i1[8];
    __shared__ double s_sum[8];
    __shared__ T b_m1;
    __shared__ double b_sum;
    // row_list: only those rows are (re)computed; in_compact: z holds just the listed rows, packed
    const long long row = row_list ? row_list[blockIdx.x] : blockIdx.x;
    const T *zr = z + (in_compact ? static_cast<long long>(blockIdx.x) : row) * ld;
    const T NEG = static_cast<T>(-CUDART_INF);
    T m1 = NEG, m2 = NEG;
    int i1 = 0x7fffffff;
    int bad = 0;  // a NaN / inf logit: comparisons skip it, so the gap would look healthy -- report the row as undecided
    for (int j = threadIdx.x; j < cols; j += 256) {
        const T v = zr[j];
        bad |= !(v < static_cast<T>(CUDART_INF));   // NaN or +inf
        if (v > m1) { m2 = m1; m1 = v; i1 = j; }
        else if (v > m2) m2 = v;
    }
    // warp merge
    for (int o = 16; o > 0; o >>= 1) {
        const T om1 = __shfl_xor_sync(0xffffffffu, m1, o), om2 = __shfl_xor_sync(0xffffffffu, m2, o);
        const int oi1 = __shfl_xor_sync(0xffffffffu, i1, o);
        if (om1 > m1 || (om1 == m1 && oi1 < i1)) { m2 = (m1 > om2) ? m1 : om2; m1 = om1; i1 = oi1; }
        else { m2 = (m2 > om1) ? m2 : om1; }
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { s_m1[w] = m1; s_m2[w] = m2; s_i1[w] = i1; }
    const int any_bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        T a1 = s_m1[0], a2 = s_m2[0];
        int ai = s_i1[0];
        for (int k = 1; k < 8; k++) {
            const T om1 = s_m1[k], om2 = s_m2[k];
            const int oi1 = s_i1[k];
            if (om1 > a1 || (om1 == a1 && oi1 < ai)) { a2 = (a1 > om2) ? a1 : om2; a1 = om1; ai = oi1; }
            else { a2 = (a2 > om1) ? a2 : om1; }
        }
        b_m1 = a1;
        if (labels) labels[row] = ai;
        if (gap) gap[row] = any_bad ? __int_as_float(0x7fc00000) : static_cast<float>(a1 - a2);
    }
    __syncthreads();
    if (!probs) return;
    const T mx = b_m1;
    double sum = 0.0;
    for (int j = threadIdx.x; j < cols; j += 256) sum += exp(static_cast<double>(zr[j] - mx));
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (l == 0) s_sum[w] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < 8; k++) t += s_sum[k];
        b_sum = t;
    }
    __syncthreads();
    const double inv = 1.0 / b_sum;
    double *pr = probs + row * probs_ld;
    for (int j = threadIdx.x; j < cols; j += 256) pr[j] = exp(static_cast<double>(zr[j] - mx)) * inv;
}

// Register-resident variant for fp32 logits with at most 16,384 columns (pitch a multiple of 4): the row is read
// from HBM exactly once as float4, kept in registers for the max / argmax / top-2 scan, the exp-sum and the
// normalisation, and the probabilities leave as 16-byte stores.  exp is evaluated in fp32 (the logits are fp32
// tensor-core results; their own error dominates) and widened to the fp64 output type of the reference.
// Persistent: each CTA walks over rows (stride gridDim.x) and loads the NEXT row's float4s before it starts the
// reductions of the current one, so the three block-wide phases of a row overlap with the following row's HBM reads.
template <int THREADS, bool PREFETCH, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
k_softmax_rows_reg(const float *__restrict__ z, long long ld, int cols, double *__restrict__ probs, long long probs_ld,
                   int32_t *__restrict__ labels, float *__restrict__ gap, const int32_t *__restrict__ row_list,
                   int in_compact, long long nrows) {
    constexpr int NV = 4096 / THREADS;  // float4 per thread (covers 16,384 columns)
    constexpr int NW = THREADS / 32;
    __shared__ float s_m1[NW], s_m2[NW];
    __shared__ double s_sum[NW];
    __shared__ int s_i1[NW];
    __shared__ float b_m1;
    __shared__ double b_sum;
    const int nvec = cols >> 2;  // cols % 4 == 0 guaranteed by the launcher
    const float4 ninf = make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    auto load_row = [&](long long it, float4 (&dst)[NV]) {
        const long long row = row_list ? row_list[it] : it;
        const float4 *zr = reinterpret_cast<const float4 *>(z + (in_compact ? it : row) * ld);
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const int q = threadIdx.x + i * THREADS;
            dst[i] = q < nvec ? __ldg(zr + q) : ninf;
        }
    };
    float4 v[NV], nv[PREFETCH ? NV : 1];
    long long it = blockIdx.x;
    if (it < nrows) load_row(it, v);
    for (; it < nrows; it += gridDim.x) {
        const long long row = row_list ? row_list[it] : it;
        const bool has_next = it + gridDim.x < nrows;
        if (PREFETCH) { if (has_next) load_row(it + gridDim.x, reinterpret_cast<float4 (&)[NV]>(nv)); }
        float m1 = -CUDART_INF_F, m2 = -CUDART_INF_F;
        int i1 = 0x7fffffff;
        int bad = 0;   // a NaN / +inf logit: comparisons skip a NaN, so the gap would look healthy -- report the row as undecided
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const int base = (threadIdx.x + i * THREADS) * 4;
            const float e[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                bad |= !(e[j] < CUDART_INF_F);
                if (e[j] > m1) { m2 = m1; m1 = e[j]; i1 = base + j; }
                else if (e[j] > m2) m2 = e[j];
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float om1 = __shfl_xor_sync(0xffffffffu, m1, o), om2 = __shfl_xor_sync(0xffffffffu, m2, o);
            const int oi1 = __shfl_xor_sync(0xffffffffu, i1, o);
            if (om1 > m1 || (om1 == m1 && oi1 < i1)) { m2 = fmaxf(m1, om2); m1 = om1; i1 = oi1; }
            else m2 = fmaxf(m2, om1);
        }
        if (l == 0) { s_m1[w] = m1; s_m2[w] = m2; s_i1[w] = i1; }
        const int any_bad = __syncthreads_or(bad);
        if (threadIdx.x == 0) {
            float a1 = s_m1[0], a2 = s_m2[0];
            int ai = s_i1[0];
            for (int k = 1; k < NW; k++) {
                const float om1 = s_m1[k], om2 = s_m2[k];
                const int oi1 = s_i1[k];
                if (om1 > a1 || (om1 == a1 && oi1 < ai)) { a2 = fmaxf(a1, om2); a1 = om1; ai = oi1; }
                else a2 = fmaxf(a2, om1);
            }
            b_m1 = a1;
            if (labels) labels[row] = ai;
            if (gap) gap[row] = any_bad ? __int_as_float(0x7fc00000) : a1 - a2;
        }
        __syncthreads();
        if (probs) {
            const float mx = b_m1;
            double sum = 0.0;  // summed in fp64 so that the probabilities of a row add up to 1 to fp64 rounding
#pragma unroll
            for (int i = 0; i < NV; i++) {  // exp(-inf) = 0 for the out-of-range slots
                v[i].x = __expf(v[i].x - mx); v[i].y = __expf(v[i].y - mx); v[i].z = __expf(v[i].z - mx); v[i].w = __expf(v[i].w - mx);
                sum += (static_cast<double>(v[i].x) + static_cast<double>(v[i].y)) + (static_cast<double>(v[i].z) + static_cast<double>(v[i].w));
            }
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (l == 0) s_sum[w] = sum;
            __syncthreads();
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int k = 0; k < NW; k++) t += s_sum[k];
                b_sum = t;
            }
            __syncthreads();
            const double inv = 1.0 / b_sum;
            double *pr = probs + row * probs_ld;
            const bool aligned = ((reinterpret_cast<uintptr_t>(pr) & 15) == 0);
#pragma unroll
            for (int i = 0; i < NV; i++) {
                const int q = threadIdx.x + i * THREADS;
                if (q < nvec) {
                    const double a = v[i].x * inv, b = v[i].y * inv, cc = v[i].z * inv, d = v[i].w * inv;
                    if (aligned) {
                        *reinterpret_cast<double2 *>(pr + q * 4) = make_double2(a, b);
                        *reinterpret_cast<double2 *>(pr + q * 4 + 2) = make_double2(cc, d);
                    } else {
                        pr[q * 4] = a; pr[q * 4 + 1] = b; pr[q * 4 + 2] = cc; pr[q * 4 + 3] = d;
                    }
                }
            }
        }
        __syncthreads();  // b_m1 / b_sum / s_* are reused by the next row
        if (PREFETCH) {
#pragma unroll
            for (int i = 0; i < NV; i++) v[i] = nv[i];
        } else if (has_next) {
            load_row(it + gridDim.x, v);  // several CTAs per SM hide this latency instead of a register prefetch
        }
    }
}

int launch_softmax_rows(Context *c, const void *logits, bool is_f64, int64_t rows, int64_t cols, int64_t ld,
                        double *probs, int64_t probs_ld, int32_t *labels, float *gap, const int32_t *row_list,
                        int nrows_list, int in_compact) {
    const int64_t n = row_list ? nrows_list : rows;
    if (n == 0 || cols == 0) return NB200_OK;
    if (!is_f64 && cols <= 16384 && (cols & 3) == 0 && (ld & 3) == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0) {
        const int sms = c->sm_count > 0 ? c->sm_count : 148;
        static const int variant = getenv("NB200_SOFTMAX_VARIANT") ? atoi(getenv("NB200_SOFTMAX_VARIANT")) : 0;
        const float *zf = (const float *)logits;
        auto grid_for = [&](int per_sm) { return static_cast<unsigned>(n < static_cast<int64_t>(per_sm) * sms ? n : static_cast<int64_t>(per_sm) * sms); };
        // 512 threads, two CTAs per SM, no register prefetch: one CTA's stores overlap the other's loads and reductions
        // (0.80 -> 0.50 ms on 10,000 x 14,588; the 1024-thread variant with a register prefetch of the next row keeps one
        // CTA per SM and is kept behind NB200_SOFTMAX_VARIANT=1)
        if (variant == 1) k_softmax_rows_reg<1024, true, 1><<<grid_for(2), 1024, 0, c->stream>>>(zf, ld, (int)cols, probs, probs_ld, labels, gap, row_list, in_compact, n);
        else k_softmax_rows_reg<512, false, 2><<<grid_for(2), 512, 0, c->stream>>>(zf, ld, (int)cols, probs, probs_ld, labels, gap, row_list, in_compact, n);
        NB_LAUNCH_CHECK(c);
        return NB200_OK;
    }
    if (is_f64) k_softmax_rows<double><<<static_cast<unsigned>(n), 256, 0, c->stream>>>((const double *)logits, ld, (int)cols, probs, probs_ld, labels, gap, row_list, in_compact);
    else k_softmax_rows<float><<<static_cast<unsigned>(n), 256, 0, c->stream>>>((const float *)logits, ld, (int)cols, probs, probs_ld, labels, gap, row_list, in_compact);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// eps_r = u * (||y1_r|| * max|wo| + ||x_r|| * max|w1| * max_l ||wo_l||): a one-sigma-style estimate
// of the low-precision logit error of row r (DESIGN.md "error model"); rows with gap < tau*eps_r are
// re-evaluated in fp64.
__global__ void k_flag_rows(long long rows, const float *__restrict__ gap, const float *__restrict__ x_sumsq,
                            const float *__restrict__ y_sumsq, const float *__restrict__ w1_maxabs,
                            const float *__restrict__ wo_maxabs, const float *__restrict__ wo_rownorm_max, float u,
                            float tau, int32_t *__restrict__ list, int *__restrict__ count, int max_list,
                            const int32_t *__restrict__ src_list) {
    const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (i >= rows) return;
    const long long r = src_list ? src_list[i] : i;  // src_list: only re-examine those rows
    const float eps = u * (sqrtf(y_sumsq[r]) * (*wo_maxabs) + sqrtf(x_sumsq[r]) * (*w1_maxabs) * (*wo_rownorm_max));
    if (!(gap[r] >= tau * eps)) {  // also catches NaN
        const int slot = atomicAdd(count, 1);
        if (slot < max_list) list[slot] = static_cast<int32_t>(r);
    }
}
int launch_flag_rows(Context *c, int64_t rows, const float *gap, const float *x_sumsq, const float *y_sumsq,
                     const float *w1_maxabs, const float *wo_maxabs, const float *wo_rownorm_max, float u, float tau,
                     int32_t *list, int *count, int max_list, const int32_t *src_list) {
    if (rows == 0) return NB200_OK;
    k_flag_rows<<<static_cast<unsigned>(ceil_div(rows, 256)), 256, 0, c->stream>>>(rows, gap, x_sumsq, y_sumsq, w1_maxabs, wo_maxabs, wo_rownorm_max, u, tau, list, count, max_list, src_list);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

__global__ void k_max_rownorm(const float *__restrict__ row_sumsq, long long rows, float *__restrict__ out) {
    float mx = 0.f;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < rows;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
        mx = fmaxf(mx, sqrtf(row_sumsq[i]));
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned int *>(out), __float_as_uint(mx));
}
int launch_max_rownorm(Context *c, const float *row_sumsq, int64_t rows, float *out_scalar) {
    NB_CUDA(cudaMemsetAsync(out_scalar, 0, sizeof(float), c->stream));
    if (rows == 0) return NB200_OK;
    k_max_rownorm<<<static_cast<unsigned>(ceil_div(rows, 256) < 256 ? ceil_div(rows, 256) : 256), 256, 0, c->stream>>>(row_sumsq, rows, out_scalar);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

__global__ void k_gather_rows_f64(const double *__restrict__ src, long long ld, long long cols,
                                  const int32_t *__restrict__ list, double *__restrict__ dst) {
    const long long r = list[blockIdx.y];
    for (long long j = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; j < cols;
         j += static_cast<long long>(gridDim.x) * blockDim.x)
        dst[static_cast<long long>(blockIdx.y) * cols + j] = src[r * ld + j];
}
int launch_gather_rows_f64(Context *c, const double *src, int64_t ld, int64_t cols, const int32_t *list, int n,
                           double *dst) {
    if (n == 0 || cols == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 256) < 64 ? ceil_div(cols, 256) : 64), n);
    k_gather_rows_f64<<<grid, 256, 0, c->stream>>>(src, ld, cols, list, dst);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
// refined fp64 logits of the listed rows back into the fp32 logits matrix
__global__ void k_scatter_rows(const double *__restrict__ src, long long cols, const int32_t *__restrict__ list,
                               float *__restrict__ dst, long long dst_ld) {
    const long long r = list[blockIdx.y];
    for (long long j = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; j < cols;
         j += static_cast<long long>(gridDim.x) * blockDim.x)
        dst[r * dst_ld + j] = static_cast<float>(src[static_cast<long long>(blockIdx.y) * cols + j]);
}
int launch_scatter_rows(Context *c, const double *src, int64_t cols, const int32_t *list, int n, float *dst_f32,
                        int64_t dst_ld) {
    if (n == 0 || cols == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(cols, 256) < 64 ? ceil_div(cols, 256) : 64), n);
    k_scatter_rows<<<grid, 256, 0, c->stream>>>(src, cols, list, dst_f32, dst_ld);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// fp32 master -> fp64 block staging (rows x cols, zero padded outside valid), optional exp (the
// `yo` view: FFTransposeBiasSum.h:94-104 applies exp only inside the un-padded region)
template <typename T>
__global__ void k_extract_to_f64(const T *__restrict__ src, long long ld, long long row0, long long col0,
                                     int rows, int cols, int valid_rows, int valid_cols, double *__restrict__ dst,
                                     long long dst_ld, int apply_exp) {
    const long long n = static_cast<long long>(rows) * cols;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / cols), cc = static_cast<int>(i % cols);
        double v = 0.0;
        if (r < valid_rows && cc < valid_cols) {
            v = static_cast<double>(src[(row0 + r) * ld + col0 + cc]);
            if (apply_exp) v = exp(v);
        }
        dst[r * dst_ld + cc] = v;
    }
}
int launch_extract_to_f64(Context *c, const void *src, bool src_is_f64, int64_t ld, int64_t row0, int64_t col0, int rows,
                          int cols, int valid_rows, int valid_cols, double *dst, int64_t dst_ld, int apply_exp) {
    const int64_t n = static_cast<int64_t>(rows) * cols;
    if (n == 0) return NB200_OK;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    if (src_is_f64) k_extract_to_f64<double><<<blocks, 256, 0, c->stream>>>((const double *)src, ld, row0, col0, rows, cols, valid_rows, valid_cols, dst, dst_ld, apply_exp);
    else k_extract_to_f64<float><<<blocks, 256, 0, c->stream>>>((const float *)src, ld, row0, col0, rows, cols, valid_rows, valid_cols, dst, dst_ld, apply_exp);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// FullyConnectedNetwork / SemanticClassifier (FF_proj/headers/FullyConnectedNetwork.h:62-123,
// word2vec/headers/SemanticClassifier.h:43-141) view the raw buffer of a (zero-padded) block as a row-major
// features x batch matrix, whatever the block's own shape.  k_udf_gather writes that view, for every block of the set,
// as rows of a K-contiguous operand: xt[(b * batch + n) * features + f] = raw_b[f * batch + n], where raw_b[i] is element
// (i / block_cols, i % block_cols) of block b = (bi, bj) of the de-blocked set (0 in the padding).
template <typename T>
__global__ void k_udf_gather(const T *__restrict__ src, long long ld, int total_rows, int total_cols, int br, int bc, int nbc,
                             int features, int batch, long long n_total, double *__restrict__ xt) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n_total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int f = static_cast<int>(i % features);
        const long long bn = i / features;
        const int n = static_cast<int>(bn % batch), b = static_cast<int>(bn / batch);
        const long long raw = static_cast<long long>(f) * batch + n;
        const int r = static_cast<int>(raw / bc), cc = static_cast<int>(raw % bc);
        const long long gr = static_cast<long long>(b / nbc) * br + r, gc = static_cast<long long>(b % nbc) * bc + cc;
        xt[i] = (gr < total_rows && gc < total_cols) ? static_cast<double>(src[gr * ld + gc]) : 0.0;
    }
}
// y1 = outLabel(sigmoid(z)) written as the labels x batch block (bi, bj) of the output set (FullyConnectedNetwork.cc:13-24:
// sigmoid(x) = 1 / (1 + exp(-x)), outLabel(x) = x > 0.5 ? 1 : 0, both evaluated in fp64 as the reference does)
__global__ void k_udf_labels(const double *__restrict__ zt, int labels, int batch, int nbc, long long n_total, double *__restrict__ dst,
                             long long dst_ld) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n_total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int l = static_cast<int>(i % labels);
        const long long bn = i / labels;
        const int n = static_cast<int>(bn % batch), b = static_cast<int>(bn / batch);
        const double sg = 1 / (1 + exp(-zt[i]));
        dst[(static_cast<long long>(b / nbc) * labels + l) * dst_ld + static_cast<long long>(b % nbc) * batch + n] = sg > 0.5 ? 1.0 : 0.0;
    }
}
int launch_udf_gather(Context *c, const void *src, bool src_is_f64, int64_t ld, int total_rows, int total_cols, int br, int bc, int nbr, int nbc,
                      int features, int batch, double *xt) {
    const int64_t n = static_cast<int64_t>(nbr) * nbc * batch * features;
    if (n == 0) return NB200_OK;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    if (src_is_f64) k_udf_gather<double><<<blocks, 256, 0, c->stream>>>(static_cast<const double *>(src), ld, total_rows, total_cols, br, bc, nbc, features, batch, n, xt);
    else k_udf_gather<float><<<blocks, 256, 0, c->stream>>>(static_cast<const float *>(src), ld, total_rows, total_cols, br, bc, nbc, features, batch, n, xt);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}
int launch_udf_labels(Context *c, const double *zt, int labels, int batch, int nbr, int nbc, double *dst, int64_t dst_ld) {
    const int64_t n = static_cast<int64_t>(nbr) * nbc * batch * labels;
    if (n == 0) return NB200_OK;
    const int blocks = static_cast<int>(ceil_div(n, 256) < 148 * 8 ? ceil_div(n, 256) : 148 * 8);
    k_udf_labels<<<blocks, 256, 0, c->stream>>>(zt, labels, batch, nbc, n, dst, dst_ld);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// ConvResultToChunks + ConvChunksToImage (conv2d_memory_fusion/headers/ConvResultToChunks.h:42-103,
// ConvChunksToImage.h:34-72): result[(image, window), channel] -> images[image][channel][window], a batched
// transpose through a padded shared-memory tile (both sides coalesced); channels at or beyond zsrc (the set's columns)
// are written as zero.  HBM-bound: 4 or 8 bytes in, 8 out per element.
template <typename T>
__global__ void __launch_bounds__(256)
k_result_to_images(const T *__restrict__ src, long long ld, int group, int zsrc, int z, double *__restrict__ dst) {
    __shared__ double tile[32][33];
    const int img = blockIdx.z;
    const int w0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int w = w0 + r, c = c0 + tx;
        tile[r][tx] = (w < group && c < zsrc) ? static_cast<double>(src[(static_cast<long long>(img) * group + w) * ld + c]) : 0.0;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int c = c0 + r, w = w0 + tx;
        if (c < z && w < group) dst[(static_cast<long long>(img) * z + c) * group + w] = tile[tx][r];
    }
}
int launch_result_to_images(Context *c, const void *src, bool src_is_f64, int64_t ld, int num_images, int group, int zsrc, int z, double *dst) {
    if (num_images == 0 || group == 0 || z == 0) return NB200_OK;
    dim3 grid(static_cast<unsigned>(ceil_div(group, 32)), static_cast<unsigned>(ceil_div(z, 32)), num_images);
    if (src_is_f64) k_result_to_images<double><<<grid, 256, 0, c->stream>>>((const double *)src, ld, group, zsrc, z, dst);
    else k_result_to_images<float><<<grid, 256, 0, c->stream>>>((const float *)src, ld, group, zsrc, z, dst);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

// im2col with the reference's column order and trailing 1.0 column
// (Image.h:36-57 get_win_value, ImageToChunks.h:52,67,75-84,98)
__global__ void k_im2col(const double *__restrict__ img, int n, int ch, int h, int w, int k, int stride, int padding,
                         int v_slides, int h_slides, double *__restrict__ out, long long out_ld) {
    const long long windows = static_cast<long long>(v_slides) * h_slides;
    const int cols = ch * k * k + 1;
    const long long total = static_cast<long long>(n) * windows * cols;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long row = i / cols;
        const int col = static_cast<int>(i % cols);
        double v;
        if (col == cols - 1) v = 1.0;
        else {
            const int im = static_cast<int>(row / windows), win = static_cast<int>(row % windows);
            const int c = col / (k * k), rem = col % (k * k), ki = rem / k, kj = rem % k;
            const int x = (win / h_slides) * stride + ki - padding, y = (win % h_slides) * stride + kj - padding;
            if (padding > 0 && (x < 0 || y < 0 || x >= h || y >= w)) v = 0.0;
            else v = img[((static_cast<long long>(im) * ch + c) * h + x) * w + y];
        }
        out[row * out_ld + col] = v;
    }
}
int launch_im2col(Context *c, const double *images, int n, int ch, int h, int w, int k, int stride, int padding,
                  double *out, int64_t out_ld) {
    const int vs = (h + 2 * padding - k) / stride + 1, hs = (w + 2 * padding - k) / stride + 1;
    k_im2col<<<148 * 8, 256, 0, c->stream>>>(images, n, ch, h, w, k, stride, padding, vs, hs, out, out_ld);
    NB_LAUNCH_CHECK(c);
    return NB200_OK;
}

}  // namespace nb200
