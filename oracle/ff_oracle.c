/*
 * ff_oracle.c -- see ff_oracle.h.  TEST INFRASTRUCTURE ONLY.  Parity: pinned against the reference's own UDF
 * headers compiled unmodified (oracle/ref, tests/test_oracle_pinned.py), not against a reference binary.
 *
 * Build: gcc -O3 -fopenmp -ffp-contract=off -shared -fPIC (oracle/Makefile).  The GEMM micro
 * kernels carry their own target attributes (AVX2+FMA / AVX-512) and are picked at run time,
 * so the .so built in the container also runs on the GPU box's host CPU.
 */
#include "ff_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ generator ------------ */

static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

/* FFMatrixUtil.cc:229-237: distp = U(0.0001, 0.5), distn = U(-0.5, -0.0001), coin flip picks. */
double nbo_synth_value(uint64_t seed, uint64_t index) {
    uint64_t h = splitmix64(splitmix64(seed) ^ (index * 0xD1342543DE82EF95ULL + 1ULL));
    double u = (double)(h >> 11) * (1.0 / 9007199254740992.0); /* [0,1), exact */
    double mag = u * (0.5 - 0.0001);                            /* one rounding  */
    mag = mag + 0.0001;                                         /* one rounding  */
    return (h & 1ULL) ? -mag : mag;
}

void nbo_synth_dense(double *out, int64_t rows, int64_t cols, uint64_t seed) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; r++)
        for (int64_t c = 0; c < cols; c++)
            out[r * cols + c] = nbo_synth_value(seed, (uint64_t)(r * cols + c));
}

int nbo_hw_threads(void) {
#ifdef _OPENMP
    return omp_get_num_procs();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------ sets ----------------- */

nbo_set *nbo_set_new(void) {
    nbo_set *s = (nbo_set *)calloc(1, sizeof(nbo_set));
    return s;
}

void nbo_set_free(nbo_set *s) {
    if (!s) return;
    for (int i = 0; i < s->n; i++) free(s->blocks[i].data);
    free(s->blocks);
    free(s);
}

int nbo_set_size(const nbo_set *s) { return s->n; }
const nbo_block *nbo_set_get(const nbo_set *s, int i) { return &s->blocks[i]; }

const nbo_block *nbo_set_find(const nbo_set *s, int br, int bc) {
    for (int i = 0; i < s->n; i++)
        if (s->blocks[i].block_row == br && s->blocks[i].block_col == bc) return &s->blocks[i];
    return NULL;
}

/* FFMatrixBlock(int,int,int,int,int,int,bool): data zeroed at construction (Array.cc:84-100). */
static nbo_block *set_add(nbo_set *s, int br, int bc, int rows, int cols, int tr, int tc) {
    if (s->n == s->cap) {
        s->cap = s->cap ? s->cap * 2 : 16;
        s->blocks = (nbo_block *)realloc(s->blocks, (size_t)s->cap * sizeof(nbo_block));
    }
    nbo_block *b = &s->blocks[s->n++];
    b->block_row = br; b->block_col = bc; b->rows = rows; b->cols = cols;
    b->total_rows = tr; b->total_cols = tc;
    b->data = (double *)calloc((size_t)rows * (size_t)cols, sizeof(double));
    if (!b->data) { fprintf(stderr, "nbo: out of memory\n"); abort(); }
    return b;
}

/* exported for oracle/ref/ref_harness.cc, which converts the reference's FFMatrixBlocks into sets of this library */
nbo_block *nbo_set_add_block(nbo_set *s, int br, int bc, int rows, int cols, int tr, int tc) { return set_add(s, br, bc, rows, cols, tr, tc); }

static int ceil_div(int a, int b) { return (a + b - 1) / b; }

/* FFMatrixUtil.cc:245-302 */
nbo_set *nbo_load_matrix(const double *dense, int total_rows, int total_cols, int block_rows,
                         int block_cols, int dont_pad_rows, int dont_pad_cols) {
    nbo_set *s = nbo_set_new();
    if (!total_rows || !total_cols || !block_rows || !block_cols) return s; /* :221-225 */
    int nbr = ceil_div(total_rows, block_rows), nbc = ceil_div(total_cols, block_cols);
    for (int i = 0; i < nbr; i++) {
        int abr = dont_pad_rows ? (block_rows < total_rows - i * block_rows ? block_rows : total_rows - i * block_rows) : block_rows;
        for (int j = 0; j < nbc; j++) {
            int abc = dont_pad_cols ? (block_cols < total_cols - j * block_cols ? block_cols : total_cols - j * block_cols) : block_cols;
            nbo_block *b = set_add(s, i, j, abr, abc, total_rows, total_cols);
            for (int ii = 0; ii < abr; ii++) {
                int r = i * block_rows + ii;
                if (r >= total_rows) break;
                for (int jj = 0; jj < abc; jj++) {
                    int c = j * block_cols + jj;
                    if (c >= total_cols) break;
                    b->data[(size_t)ii * abc + jj] = dense[(size_t)r * total_cols + c];
                }
            }
        }
    }
    return s;
}

nbo_set *nbo_load_matrix_synth(uint64_t seed, int total_rows, int total_cols, int block_rows,
                               int block_cols) {
    nbo_set *s = nbo_set_new();
    if (!total_rows || !total_cols || !block_rows || !block_cols) return s;
    int nbr = ceil_div(total_rows, block_rows), nbc = ceil_div(total_cols, block_cols);
    for (int i = 0; i < nbr; i++)
        for (int j = 0; j < nbc; j++) set_add(s, i, j, block_rows, block_cols, total_rows, total_cols);
#pragma omp parallel for schedule(dynamic)
    for (int bi = 0; bi < s->n; bi++) {
        nbo_block *b = &s->blocks[bi];
        for (int ii = 0; ii < b->rows; ii++) {
            int64_t r = (int64_t)b->block_row * block_rows + ii;
            if (r >= total_rows) break;
            for (int jj = 0; jj < b->cols; jj++) {
                int64_t c = (int64_t)b->block_col * block_cols + jj;
                if (c >= total_cols) break;
                b->data[(size_t)ii * b->cols + jj] = nbo_synth_value(seed, (uint64_t)(r * total_cols + c));
            }
        }
    }
    return s;
}

void nbo_set_to_dense(const nbo_set *s, double *dense, int total_rows, int total_cols) {
    for (int bi = 0; bi < s->n; bi++) {
        const nbo_block *b = &s->blocks[bi];
        /* block origin: all blocks of a set share the nominal block size of block (0,0) */
        const nbo_block *b0 = nbo_set_find(s, 0, 0);
        int nr = b0 ? b0->rows : b->rows, nc = b0 ? b0->cols : b->cols;
        for (int ii = 0; ii < b->rows; ii++) {
            int64_t r = (int64_t)b->block_row * nr + ii;
            if (r >= total_rows) break;
            for (int jj = 0; jj < b->cols; jj++) {
                int64_t c = (int64_t)b->block_col * nc + jj;
                if (c >= total_cols) break;
                dense[r * total_cols + c] = b->data[(size_t)ii * b->cols + jj];
            }
        }
    }
}

/* ------------------------------------------------------------------ fp64 GEMM ------------ */
/* Stand-in for the Eigen 3 products at FFTransposeMult.h:92 and FFInputLayerJoin.h:86 (Eigen is
 * an un-vendored, unpinned third-party dependency: `apt-get install libeigen3-dev`,
 * SConstruct:22).  Classic packed-panel GEMM; summation order differs from Eigen's, which is
 * itself unspecified, so agreement is to fp64 rounding (~1e-13 relative), not bitwise. */

#define KC 256
#define MC 96
#define NC 4096

typedef double v4d __attribute__((vector_size(32)));
typedef double v8d __attribute__((vector_size(64)));

#define DEFINE_MICRO(NAME, VEC, VL, TARGET)                                                     \
    __attribute__((target(TARGET), optimize("-ffp-contract=fast"))) static void NAME(int kc, const double *ap, const double *bp, \
                                                    double *c, int ldc, int mr, int nr, int acc) \
    {                                                                                          \
        VEC c00 = {0}, c01 = {0}, c10 = {0}, c11 = {0}, c20 = {0}, c21 = {0}, c30 = {0}, c31 = {0}; \
        for (int p = 0; p < kc; p++) {                                                         \
            VEC b0, b1;                                                                        \
            memcpy(&b0, bp + (size_t)p * 2 * VL, sizeof(VEC));                                  \
            memcpy(&b1, bp + (size_t)p * 2 * VL + VL, sizeof(VEC));                             \
            double a0 = ap[p * 4 + 0], a1 = ap[p * 4 + 1], a2 = ap[p * 4 + 2], a3 = ap[p * 4 + 3]; \
            c00 += a0 * b0; c01 += a0 * b1;                                                    \
            c10 += a1 * b0; c11 += a1 * b1;                                                    \
            c20 += a2 * b0; c21 += a2 * b1;                                                    \
            c30 += a3 * b0; c31 += a3 * b1;                                                    \
        }                                                                                      \
        double t[4][2 * VL];                                                                   \
        memcpy(&t[0][0], &c00, sizeof(VEC)); memcpy(&t[0][VL], &c01, sizeof(VEC));             \
        memcpy(&t[1][0], &c10, sizeof(VEC)); memcpy(&t[1][VL], &c11, sizeof(VEC));             \
        memcpy(&t[2][0], &c20, sizeof(VEC)); memcpy(&t[2][VL], &c21, sizeof(VEC));             \
        memcpy(&t[3][0], &c30, sizeof(VEC)); memcpy(&t[3][VL], &c31, sizeof(VEC));             \
        for (int i = 0; i < mr; i++)                                                           \
            for (int j = 0; j < nr; j++) {                                                     \
                if (acc) c[(size_t)i * ldc + j] += t[i][j];                                    \
                else c[(size_t)i * ldc + j] = t[i][j];                                         \
            }                                                                                  \
    }

DEFINE_MICRO(micro_avx2, v4d, 4, "avx2,fma")
DEFINE_MICRO(micro_avx512, v8d, 8, "avx512f,fma")

static int g_use512 = -1;

static void dgemm_impl(int m, int n, int k, const double *a, int lda, const double *b, int ldb,
                       int b_trans, double *c, int ldc, int use512) {
    const int MR = 4, NR = use512 ? 16 : 8;
    double *ap = (double *)aligned_alloc(64, (size_t)MC * KC * sizeof(double));
    int ncmax = n < NC ? n : NC;
    int ncpad = ((ncmax + NR - 1) / NR) * NR;
    double *bp = (double *)aligned_alloc(64, (((size_t)ncpad * KC * sizeof(double)) + 63) / 64 * 64);
    if (k == 0) {
        for (int i = 0; i < m; i++) memset(c + (size_t)i * ldc, 0, (size_t)n * sizeof(double));
    }
    for (int pc = 0; pc < k; pc += KC) {
        int kc = k - pc < KC ? k - pc : KC;
        for (int jc = 0; jc < n; jc += NC) {
            int nc = n - jc < NC ? n - jc : NC;
            /* pack B: panels of NR columns, each panel kc x NR row-major */
            for (int jr = 0; jr < nc; jr += NR) {
                int nr = nc - jr < NR ? nc - jr : NR;
                double *dst = bp + (size_t)jr * kc;
                for (int p = 0; p < kc; p++)
                    for (int j = 0; j < NR; j++) {
                        double v = 0.0;
                        if (j < nr)
                            v = b_trans ? b[(size_t)(jc + jr + j) * ldb + (pc + p)]
                                        : b[(size_t)(pc + p) * ldb + (jc + jr + j)];
                        dst[(size_t)p * NR + j] = v;
                    }
            }
            for (int ic = 0; ic < m; ic += MC) {
                int mc = m - ic < MC ? m - ic : MC;
                for (int ir = 0; ir < mc; ir += MR) {
                    int mr = mc - ir < MR ? mc - ir : MR;
                    double *dst = ap + (size_t)ir * kc;
                    for (int p = 0; p < kc; p++)
                        for (int i = 0; i < MR; i++)
                            dst[(size_t)p * MR + i] = i < mr ? a[(size_t)(ic + ir + i) * lda + (pc + p)] : 0.0;
                }
                for (int jr = 0; jr < nc; jr += NR) {
                    int nr = nc - jr < NR ? nc - jr : NR;
                    for (int ir = 0; ir < mc; ir += MR) {
                        int mr = mc - ir < MR ? mc - ir : MR;
                        double *ct = c + (size_t)(ic + ir) * ldc + (jc + jr);
                        if (use512) micro_avx512(kc, ap + (size_t)ir * kc, bp + (size_t)jr * kc, ct, ldc, mr, nr, pc > 0);
                        else micro_avx2(kc, ap + (size_t)ir * kc, bp + (size_t)jr * kc, ct, ldc, mr, nr, pc > 0);
                    }
                }
            }
        }
    }
    free(ap);
    free(bp);
}

void nbo_dgemm(int m, int n, int k, const double *a, int lda, const double *b, int ldb,
               int b_trans, double *c, int ldc) {
    if (g_use512 < 0) {
        __builtin_cpu_init();
        g_use512 = __builtin_cpu_supports("avx512f") ? 1 : 0;
        if (!__builtin_cpu_supports("avx2") || !__builtin_cpu_supports("fma")) {
            fprintf(stderr, "nbo: host CPU lacks AVX2/FMA\n");
            abort();
        }
    }
    dgemm_impl(m, n, k, a, lda, b, ldb, b_trans, c, ldc, g_use512);
}

/* ------------------------------------------------------------------ join + aggregate ----- */

typedef struct { int i, j; double *data; int rows, cols; } partial;

/* Shared body of FFTransposeMult/FFInputLayerJoin followed by FFAggMatrix.
 * mode 0: predicate a.blockCol == b.blockCol, out = A * B^T   (FFTransposeMult.h:38-39, :92)
 * mode 1: predicate a.blockCol == b.blockRow, out = A * B     (FFInputLayerJoin.h:30-31, :86)
 *
 * Pseudo-cluster analogue: `threads` pipeline threads.  The join's hash partition is the join
 * key (FFMatrixBlock::hash() = Hasher(blockColIndex), FFMatrixBlock.h:151-156), so thread t
 * probes the pairs whose key k satisfies k % threads == t.  Every pair product is materialised
 * into a fresh zeroed I x J block (makeObject<FFMatrixBlock>, FFTransposeMult.h:68-71) and then
 * combined into that thread's map by (i,j) (CombinedShuffleSink.h:112-144: first key keeps the
 * value, later ones `+`).  The per-thread maps are merged per key in a second phase
 * (AggregationProcessor.cc:131-156). */
static nbo_set *join_agg(const nbo_set *a, const nbo_set *b, int mode, int threads) {
    if (threads < 1) threads = 1;
    /* output grid */
    int ni = 0, nj = 0;
    for (int x = 0; x < a->n; x++) if (a->blocks[x].block_row + 1 > ni) ni = a->blocks[x].block_row + 1;
    for (int y = 0; y < b->n; y++) {
        int jj = mode == 0 ? b->blocks[y].block_row : b->blocks[y].block_col;
        if (jj + 1 > nj) nj = jj + 1;
    }
    size_t ncell = (size_t)ni * nj;
    int nkeys = 0;
    for (int x = 0; x < a->n; x++) if (a->blocks[x].block_col + 1 > nkeys) nkeys = a->blocks[x].block_col + 1;
    const int few_keys = nkeys < threads;
    /* per-thread combiner maps: direct-indexed by (i,j) */
    double **maps = (double **)calloc((size_t)threads * ncell, sizeof(double *));
    int *cell_rows = (int *)calloc(ncell, sizeof(int)), *cell_cols = (int *)calloc(ncell, sizeof(int));
    int *cell_tr = (int *)calloc(ncell, sizeof(int)), *cell_tc = (int *)calloc(ncell, sizeof(int));
    int bad = 0;

#pragma omp parallel num_threads(threads)
    {
#ifdef _OPENMP
        int t = omp_get_thread_num();
#else
        int t = 0;
#endif
        for (int x = 0; x < a->n; x++) {
            const nbo_block *ba = &a->blocks[x];
            int key = ba->block_col;
            if (!few_keys && key % threads != t) continue;
            for (int y = 0; y < b->n; y++) {
                const nbo_block *bb = &b->blocks[y];
                int bkey = mode == 0 ? bb->block_col : bb->block_row;
                if (bkey != key) continue;
                /* fewer join keys than threads (layer 2: one K block): the reference would run the
                 * whole join on one hash partition; spread the pairs instead so that the CPU
                 * baseline is not handicapped (it only gets faster than the reference this way) */
                if (few_keys && (x + y) % threads != t) continue;
                int I = ba->rows, K = ba->cols;
                int J = mode == 0 ? bb->rows : bb->cols;
                int Kb = mode == 0 ? bb->cols : bb->rows;
                if (K != Kb) { /* "Block dimemsions mismatch!" exit(1): FFTransposeMult.h:61-65 */
#pragma omp atomic write
                    bad = 1;
                    continue;
                }
                int oi = ba->block_row, oj = mode == 0 ? bb->block_row : bb->block_col;
                double *out = (double *)calloc((size_t)I * J, sizeof(double));
                nbo_dgemm(I, J, K, ba->data, ba->cols, bb->data, bb->cols, mode == 0, out, J);
                size_t cell = (size_t)oi * nj + oj;
                double **slot = &maps[(size_t)t * ncell + cell];
                if (!*slot) {
                    *slot = out;
                } else { /* FFMatrixData::operator+ (FFMatrixData.h:38-47) */
                    double *acc = *slot;
                    for (size_t e = 0; e < (size_t)I * J; e++) acc[e] += out[e];
                    free(out);
                }
                /* meta: FFTransposeMult.h:68-71 / FFInputLayerJoin.h:64-67 (every pair of one
                 * cell writes the same four values) */
                cell_rows[cell] = I; cell_cols[cell] = J;
                cell_tr[cell] = ba->total_rows;
                cell_tc[cell] = mode == 0 ? bb->total_rows : bb->total_cols;
            }
        }
    }
    if (bad) {
        fprintf(stderr, "Block dimemsions mismatch! \n");
        for (size_t z = 0; z < (size_t)threads * ncell; z++) free(maps[z]);
        free(maps); free(cell_rows); free(cell_cols); free(cell_tr); free(cell_tc);
        return NULL;
    }
    /* second merge, one owner per key */
    nbo_set *out = nbo_set_new();
    for (size_t cell = 0; cell < ncell; cell++) {
        int present = 0;
        for (int t = 0; t < threads; t++) if (maps[(size_t)t * ncell + cell]) present = 1;
        if (!present) continue;
        set_add(out, (int)(cell / nj), (int)(cell % nj), cell_rows[cell], cell_cols[cell], cell_tr[cell], cell_tc[cell]);
    }
#pragma omp parallel for schedule(dynamic) num_threads(threads)
    for (int bi = 0; bi < out->n; bi++) {
        nbo_block *ob = &out->blocks[bi];
        size_t cell = (size_t)ob->block_row * nj + ob->block_col;
        size_t ne = (size_t)ob->rows * ob->cols;
        int first = 1;
        for (int t = 0; t < threads; t++) {
            double *p = maps[(size_t)t * ncell + cell];
            if (!p) continue;
            if (first) { memcpy(ob->data, p, ne * sizeof(double)); first = 0; }
            else for (size_t e = 0; e < ne; e++) ob->data[e] += p[e];
            free(p);
        }
    }
    free(maps); free(cell_rows); free(cell_cols); free(cell_tr); free(cell_tc);
    return out;
}

nbo_set *nbo_transpose_mult_agg(const nbo_set *a, const nbo_set *b, int threads) { return join_agg(a, b, 0, threads); }
nbo_set *nbo_input_layer_join_agg(const nbo_set *a, const nbo_set *b, int threads) { return join_agg(a, b, 1, threads); }

/* FFReluBiasSum.h:34-35 predicate in1.blockRow == in2.blockRow; :49-56 bias must be I x 1;
 * :68-71 out = max(0, in1 + in2[i]); out meta = in1's (:58-61). */
nbo_set *nbo_relu_bias_sum(const nbo_set *in, const nbo_set *bias) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        for (int y = 0; y < bias->n; y++) {
            const nbo_block *bb = &bias->blocks[y];
            if (bb->block_row != a->block_row) continue;
            if (a->rows != bb->rows || bb->cols != 1) { fprintf(stderr, "Block dimemsions mismatch!\n"); nbo_set_free(out); return NULL; }
            nbo_block *o = set_add(out, a->block_row, a->block_col, a->rows, a->cols, a->total_rows, a->total_cols);
            for (int i = 0; i < a->rows; i++)
                for (int j = 0; j < a->cols; j++) {
                    double v = a->data[(size_t)i * a->cols + j] + bb->data[i];
                    o->data[(size_t)i * a->cols + j] = v > 0.0 ? v : 0.0;
                }
        }
    }
    return out;
}

/* FullyConnectedNetwork (FF_proj/headers/FullyConnectedNetwork.h:62-123) and SemanticClassifier
 * (word2vec/headers/SemanticClassifier.h:43-141): a SelectionComp that runs a whole two-layer network on every block.
 * The block's raw buffer is viewed as a row-major features x batch matrix x0 (whatever the block's own row/column
 * counts say: :75-77 / :77-79), then
 *     y0 = relu(w0 * x0 + b0)            w0: hidden x features, row-major             (:90-92 / :95-97)
 *     y1 = outLabel(sigmoid(w1 * y0 + b1))   w1: labels x hidden; sigmoid(x) = 1 / (1 + exp(-x)), outLabel(x) = x > 0.5
 *                                                                                      (:105-108 / :114-117; .cc:13-24)
 * and the result is a labels x batch block with the input's block indices (:111-113 / :120-123).  Its totals are what the
 * FFMatrixBlock constructor used leaves: the block's own size in SemanticClassifier (7 arguments, with_totals = 1), the
 * FFMatrixMeta defaults -1 / -1 in FullyConnectedNetwork (5 arguments, FFMatrixMeta.h:11-12; with_totals = 0).  Products accumulate over k in
 * increasing order.  A block with fewer than features * batch values is refused (the UDF would read past it). */
nbo_set *nbo_udf_dense2(const nbo_set *in, int features, int batch, int hidden, int labels, const double *w0, const double *b0,
                        const double *w1, const double *b1, int with_totals) {
    nbo_set *out = nbo_set_new();
    double *y0 = (double *)malloc((size_t)hidden * batch * sizeof(double));
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        if ((size_t)a->rows * a->cols < (size_t)features * batch) { fprintf(stderr, "block smaller than features x batch\n"); free(y0); nbo_set_free(out); return NULL; }
        for (int h = 0; h < hidden; h++)
            for (int n = 0; n < batch; n++) {
                double s = 0.0;
                for (int f = 0; f < features; f++) s += w0[(size_t)h * features + f] * a->data[(size_t)f * batch + n];
                s += b0[h];
                y0[(size_t)h * batch + n] = s > 0.0 ? s : 0.0;
            }
        nbo_block *o = set_add(out, a->block_row, a->block_col, labels, batch, with_totals ? labels : -1, with_totals ? batch : -1);
        for (int l = 0; l < labels; l++)
            for (int n = 0; n < batch; n++) {
                double s = 0.0;
                for (int h = 0; h < hidden; h++) s += w1[(size_t)l * hidden + h] * y0[(size_t)h * batch + n];
                s += b1[l];
                const double sg = 1 / (1 + exp(-s));
                o->data[(size_t)l * batch + n] = sg > 0.5 ? 1.0 : 0.0;
            }
    }
    free(y0);
    return out;
}

/* FFTransposeBiasSum, intended semantics (Q1): out block (in.blockCol, in.blockRow) of J x I,
 * totals swapped (FFTransposeBiasSum.h:64-67); value = in[i,j] + bias[i] (the `colwise() +`
 * of :84) written transposed; exp() only where the un-padded coordinates are in range
 * (:94-104), padded entries stay at the bias-sum of zero padding, which the mask leaves as is.
 * With zero-padded weights and biases that value is 0. */
nbo_set *nbo_transpose_bias_sum(const nbo_set *in, const nbo_set *bias, int apply_exp) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        for (int y = 0; y < bias->n; y++) {
            const nbo_block *bb = &bias->blocks[y];
            if (bb->block_row != a->block_row) continue;
            if (a->rows != bb->rows || bb->cols != 1) { fprintf(stderr, "Block dimemsions mismatch!\n"); nbo_set_free(out); return NULL; }
            int I = a->rows, J = a->cols;
            nbo_block *o = set_add(out, a->block_col, a->block_row, J, I, a->total_cols, a->total_rows);
            for (int i = 0; i < I; i++) {
                int act_x = a->block_row * I + i;
                for (int j = 0; j < J; j++) {
                    int act_y = a->block_col * J + j;
                    double v = a->data[(size_t)i * J + j] + bb->data[i];
                    if (apply_exp && act_x < a->total_rows && act_y < a->total_cols) v = exp(v);
                    o->data[(size_t)j * I + i] = v;
                }
            }
        }
    }
    return out;
}

/* FFRowAggregate.h:21-28: key = (blockRow, 0, totalRows, 1) (FFMatrixBlock.h:116-119), value =
 * rowwise().sum() as an I x 1 block (:122-142), summed by FFMatrixData::operator+. */
nbo_set *nbo_row_aggregate(const nbo_set *in) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        nbo_block *o = (nbo_block *)nbo_set_find(out, a->block_row, 0);
        if (!o) o = set_add(out, a->block_row, 0, a->rows, 1, a->total_rows, 1);
        for (int i = 0; i < a->rows; i++) {
            double s = 0.0;
            for (int j = 0; j < a->cols; j++) s += a->data[(size_t)i * a->cols + j];
            o->data[i] += s;
        }
    }
    return out;
}

/* FFOutputLayer, intended semantics (Q2): out[i,j] = in1[i,j] / in2[i] with stride J
 * (FFOutputLayer.h:31-57; the as-written `i*I+j` and `x/(1+x)` are this fork's edits). */
nbo_set *nbo_output_layer(const nbo_set *in, const nbo_set *rowsum) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        for (int y = 0; y < rowsum->n; y++) {
            const nbo_block *rs = &rowsum->blocks[y];
            if (rs->block_row != a->block_row) continue;
            if (a->rows != rs->rows || rs->cols != 1) { fprintf(stderr, "[FFOutputLayer] Cannot divide! Dimensions mismatch!\n"); nbo_set_free(out); return NULL; }
            nbo_block *o = set_add(out, a->block_row, a->block_col, a->rows, a->cols, a->total_rows, a->total_cols);
            for (int i = 0; i < a->rows; i++)
                for (int j = 0; j < a->cols; j++)
                    o->data[(size_t)i * a->cols + j] = a->data[(size_t)i * a->cols + j] / rs->data[i];
        }
    }
    return out;
}

/* SimpleFF.cc:331-424 */
nbo_set *nbo_ff_inference_unit(const nbo_set *w1, const nbo_set *wo, const nbo_set *inputs,
                               const nbo_set *b1, const nbo_set *bo, int threads,
                               nbo_set **logits_out, int32_t *labels) {
    nbo_set *y1pre = nbo_transpose_mult_agg(w1, inputs, threads);       /* :340-351 */
    if (!y1pre) return NULL;
    nbo_set *y1 = nbo_relu_bias_sum(y1pre, b1);                         /* :353-359 */
    nbo_set_free(y1pre);
    if (!y1) return NULL;
    nbo_set *zpre = nbo_input_layer_join_agg(wo, y1, threads);          /* :362-373 */
    nbo_set_free(y1);
    if (!zpre) return NULL;
    nbo_set *yo = nbo_transpose_bias_sum(zpre, bo, 1);                  /* :375-385 */
    nbo_set *logits = nbo_transpose_bias_sum(zpre, bo, 0);
    nbo_set_free(zpre);
    if (!yo || !logits) return NULL;
    nbo_set *rowsum = nbo_row_aggregate(yo);                            /* :398-402 */
    nbo_set *out = nbo_output_layer(yo, rowsum);                        /* :404-406 */
    nbo_set_free(rowsum);
    nbo_set_free(yo);
    if (labels && logits->n) {
        int batch = logits->blocks[0].total_rows, nlab = logits->blocks[0].total_cols;
        double *best = (double *)malloc((size_t)batch * sizeof(double));
        for (int r = 0; r < batch; r++) { best[r] = -INFINITY; labels[r] = -1; }
        const nbo_block *b0 = nbo_set_find(logits, 0, 0);
        /* row argmax, first maximum wins: scan column blocks in increasing order */
        int nbc = 0;
        for (int x = 0; x < logits->n; x++) if (logits->blocks[x].block_col + 1 > nbc) nbc = logits->blocks[x].block_col + 1;
        for (int bc = 0; bc < nbc; bc++)
            for (int x = 0; x < logits->n; x++) {
                const nbo_block *b = &logits->blocks[x];
                if (b->block_col != bc) continue;
                for (int i = 0; i < b->rows; i++) {
                    int r = b->block_row * b0->rows + i;
                    if (r >= batch) break;
                    for (int j = 0; j < b->cols; j++) {
                        int c = bc * b0->cols + j;
                        if (c >= nlab) break;
                        double v = b->data[(size_t)i * b->cols + j];
                        if (v > best[r]) { best[r] = v; labels[r] = c; }
                    }
                }
            }
        free(best);
    }
    if (logits_out) *logits_out = logits; else nbo_set_free(logits);
    return out;
}

/* ------------------------------------------------------------------ as written ----------- */
/* The two UDFs whose bodies in this fork do not do what their declarations say, restated AS WRITTEN.
 * They exist to pin this file against the reference's code where it is compiled (oracle/ref) and to document the
 * quirks; the pipeline above uses the intended forms.
 *
 * FFTransposeBiasSum.h:64-104 as written: the output block is declared transposed -- (blockCol, blockRow), J x I,
 * totals swapped -- and zero filled; only its first I elements receive in1[p] + in2[p] (both operands are mapped as
 * I x 1, :73-84); then exp() runs over i * J + j for every (i, j) whose un-padded coordinate lies inside in1's totals
 * (:94-104), so almost every in-range entry becomes exp(0) = 1. */
nbo_set *nbo_transpose_bias_sum_as_written(const nbo_set *in, const nbo_set *bias) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        for (int y = 0; y < bias->n; y++) {
            const nbo_block *bb = &bias->blocks[y];
            if (bb->block_row != a->block_row) continue;
            if (a->rows != bb->rows || bb->cols != 1) { fprintf(stderr, "Block dimemsions mismatch!\n"); nbo_set_free(out); return NULL; }
            const int I = a->rows, J = a->cols;
            nbo_block *o = set_add(out, a->block_col, a->block_row, J, I, a->total_cols, a->total_rows);
            for (int p = 0; p < I; p++) o->data[p] = a->data[p] + bb->data[p];
            for (int i = 0; i < I; i++)
                for (int j = 0; j < J; j++)
                    if (a->block_row * I + i < a->total_rows && a->block_col * J + j < a->total_cols)
                        o->data[(size_t)i * J + j] = exp(o->data[(size_t)i * J + j]);
        }
    }
    return out;
}
/* FFOutputLayer.h:31-57 as written: out[i * I + j] = in1[i * I + j] / (1 + in1[i * I + j]) (stride I instead of J, and
 * the row sum in2 is not used: "used to be in2Data[i]").  With I > J the reference writes past the end of the block;
 * this restatement refuses such shapes. */
nbo_set *nbo_output_layer_as_written(const nbo_set *in, const nbo_set *rowsum) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < in->n; x++) {
        const nbo_block *a = &in->blocks[x];
        for (int y = 0; y < rowsum->n; y++) {
            const nbo_block *rs = &rowsum->blocks[y];
            if (rs->block_row != a->block_row) continue;
            if (a->rows != rs->rows || rs->cols != 1 || a->rows > a->cols) { fprintf(stderr, "[FFOutputLayer] Cannot divide! Dimensions mismatch!\n"); nbo_set_free(out); return NULL; }
            const int I = a->rows, J = a->cols;
            nbo_block *o = set_add(out, a->block_row, a->block_col, I, J, a->total_rows, a->total_cols);
            for (int i = 0; i < I; i++)
                for (int j = 0; j < J; j++) {
                    const size_t pos = (size_t)i * I + j;
                    o->data[pos] = a->data[pos] / (1 + a->data[pos]);
                }
        }
    }
    return out;
}

/* ------------------------------------------------------------------ LA DSL --------------- */
/* LASillyTransposeMultiply1Join (sharedLibraries/headers/LASillyTransposeMultiply1Join.h:24-33 predicate
 * in1.blockRow == in2.blockRow; :44-47 row counts must agree; :48-53 out block (in1.blockCol, in2.blockCol) of
 * in1.colNums x in2.colNums, totals (in1.totalCols, in2.totalCols); product in1^T * in2) followed by
 * LASillyMultiply2Aggregate (LASillyMultiply2Aggregate.h: key = block meta, value summed by MatrixData::operator+
 * with the sum flag, MatrixData.h SumAggregate).  Restated as the block-wise transpose of `a` followed by the
 * LASillyMultiply1Join shape (in1.blockCol == in2.blockRow, in1 * in2), which is FFInputLayerJoin's. */
nbo_set *nbo_la_transpose_multiply_agg(const nbo_set *a, const nbo_set *b, int threads) {
    nbo_set *at = nbo_set_new();
    for (int x = 0; x < a->n; x++) {
        const nbo_block *ba = &a->blocks[x];
        nbo_block *o = set_add(at, ba->block_col, ba->block_row, ba->cols, ba->rows, ba->total_cols, ba->total_rows);
        for (int i = 0; i < ba->rows; i++)
            for (int j = 0; j < ba->cols; j++) o->data[(size_t)j * ba->rows + i] = ba->data[(size_t)i * ba->cols + j];
    }
    nbo_set *out = join_agg(at, b, 1, threads);
    nbo_set_free(at);
    return out;
}

/* ------------------------------------------------------------------ LSTM ----------------- */
/* The two activations as LSTMThreeWaySum.h:79-90 writes them. */
static double lstm_act(double s, int activation) {
    if (activation == 1) return 1 / (1 + exp(-1 * s));                                /* Sigmod, :81 */
    if (activation == 2) return (exp(s) - exp(-1 * s)) / (exp(s) + exp(-1 * s));      /* Tanh, :85-86 */
    return s;                                                                         /* LSTMTwoSum.h:62 */
}

/* LSTMThreeWaySum (LSTMThreeWaySum.h:39-42 predicate: the three FFMatrixMeta keys are equal, i.e. same block row
 * and column, FFMatrixMeta.h:34-37; :58-65 block dimensions must agree; :67-70 out meta = in1's; :76-95 the sum and
 * its activation over the whole padded block) and, with c == NULL and activation 0, LSTMTwoSum (LSTMTwoSum.h:24-27
 * same predicate on two inputs, :38-64).  With a alone it maps the activation over a set (the in-place tanh of
 * LSTMHiddenState.h:44-47). */
nbo_set *nbo_lstm_sum(const nbo_set *a, const nbo_set *b, const nbo_set *c, int activation) {
    nbo_set *out = nbo_set_new();
    for (int x = 0; x < a->n; x++) {
        const nbo_block *ba = &a->blocks[x];
        const nbo_block *bb = b ? nbo_set_find(b, ba->block_row, ba->block_col) : NULL;
        const nbo_block *bc = c ? nbo_set_find(c, ba->block_row, ba->block_col) : NULL;
        if ((b && !bb) || (c && !bc)) continue;                       /* inner join: no partner, no output */
        if ((bb && (bb->rows != ba->rows || bb->cols != ba->cols)) || (bc && (bc->rows != ba->rows || bc->cols != ba->cols))) {
            fprintf(stderr, "Block dimemsions mismatch!\n");
            nbo_set_free(out);
            return NULL;
        }
        nbo_block *o = set_add(out, ba->block_row, ba->block_col, ba->rows, ba->cols, ba->total_rows, ba->total_cols);
        const size_t n = (size_t)ba->rows * ba->cols;
        for (size_t i = 0; i < n; i++) {
            double sum = ba->data[i];
            if (bb) sum = sum + bb->data[i];
            if (bc) sum = sum + bc->data[i];
            o->data[i] = lstm_act(sum, activation);
        }
    }
    return out;
}

/* LSTMHiddenState (LSTMHiddenState.h:24-27 predicate in1.blockCol == in2.blockRow; :44-47 tanh over every element
 * of in2, padding included (tanh(0) = 0); :49-62 out = in1 * tanh(in2); out meta :35-38) followed by FFAggMatrix
 * (LSTMTest.cc:538-543).  The reference overwrites in2's page in place and says so ("Does this modify the data if
 * used elsewhere?"); a block probed by several in1 blocks would then be tanh'ed more than once.  The oracle rules
 * for the intended single application and leaves in2 untouched (quirk Q8 in DESIGN.md). */
nbo_set *nbo_lstm_hidden_state_agg(const nbo_set *o_t, const nbo_set *c_t, int threads) {
    nbo_set *t = nbo_lstm_sum(c_t, NULL, NULL, 2);
    if (!t) return NULL;
    nbo_set *out = nbo_input_layer_join_agg(o_t, t, threads);
    nbo_set_free(t);
    return out;
}

/* ------------------------------------------------------------------ conv2d --------------- */

/* Image::get_win_value (Image.h:36-57); h_slides/v_slides integer division (Image.h:26-32). */
static double win_value(const double *img, int c, int h, int w, int window, int i, int j, int stride,
                        int ch, int k, int padding) {
    int h_slides = (w + 2 * padding - k) / stride + 1;
    int r = window / h_slides, cc = window % h_slides;
    int x_index = r * stride + i - padding, y_index = cc * stride + j - padding;
    if (padding > 0 && (x_index < 0 || y_index < 0 || x_index >= h || y_index >= w)) return 0.0;
    (void)c;
    return img[((size_t)ch * h + x_index) * w + y_index];
}

nbo_set *nbo_image_flat(const double *images, int n, int c, int h, int w, int k, int stride,
                        int padding, int block_rows, int block_cols) {
    int v_slides = (h + 2 * padding - k) / stride + 1, h_slides = (w + 2 * padding - k) / stride + 1;
    int windows = v_slides * h_slides;
    int total_rows = n * windows, total_cols = c * k * k + 1; /* trailing 1.0: ImageToChunks.h:67,98 */
    nbo_set *s = nbo_set_new();
    int nbr = ceil_div(total_rows, block_rows), nbc = ceil_div(total_cols, block_cols);
    for (int bi = 0; bi < nbr; bi++)
        for (int bj = 0; bj < nbc; bj++) set_add(s, bi, bj, block_rows, block_cols, total_rows, total_cols);
#pragma omp parallel for schedule(dynamic)
    for (int x = 0; x < s->n; x++) {
        nbo_block *b = &s->blocks[x];
        for (int ii = 0; ii < block_rows; ii++) {
            int row = b->block_row * block_rows + ii; /* ImageToChunks.h:52: img * windows + w */
            if (row >= total_rows) break;
            int img = row / windows, win = row % windows;
            const double *ip = images + (size_t)img * c * h * w;
            for (int jj = 0; jj < block_cols; jj++) {
                int col = b->block_col * block_cols + jj;
                if (col >= total_cols) break;
                double v;
                if (col == total_cols - 1) v = 1.0;
                else { /* col = ch*k*k + i*k + j: ImageToChunks.h:75-84 */
                    int ch = col / (k * k), rem = col % (k * k);
                    v = win_value(ip, c, h, w, win, rem / k, rem % k, stride, ch, k, padding);
                }
                b->data[(size_t)ii * block_cols + jj] = v;
            }
        }
    }
    return s;
}

nbo_set *nbo_kernel_flat(const double *kernels, const double *bias, int kout, int c, int k,
                         int block_rows, int block_cols) {
    int total_cols = c * k * k + 1; /* placeholder 0 column: KernelToChunks.h:79 */
    nbo_set *s = nbo_set_new();
    int nbr = ceil_div(kout, block_rows), nbc = ceil_div(total_cols, block_cols);
    for (int bi = 0; bi < nbr; bi++)
        for (int bj = 0; bj < nbc; bj++) {
            nbo_block *b = set_add(s, bi, bj, block_rows, block_cols, kout, total_cols);
            for (int ii = 0; ii < block_rows; ii++) {
                int row = bi * block_rows + ii;
                if (row >= kout) break;
                for (int jj = 0; jj < block_cols; jj++) {
                    int col = bj * block_cols + jj;
                    if (col >= total_cols) break;
                    /* KernelBiasJoin.h:42-46: bias into the last real column of the last column block */
                    b->data[(size_t)ii * block_cols + jj] =
                        col == total_cols - 1 ? bias[row] : kernels[(size_t)row * c * k * k + col];
                }
            }
        }
    return s;
}

/* ConvResultToChunks + ConvChunksToImage (conv2d_memory_fusion/headers/ConvResultToChunks.h:42-103,
 * ConvChunksToImage.h:34-72): the `result` matrix (rows = image * group + window, columns = output channel) is cut
 * column by column into chunks of at most `group` consecutive rows; a chunk goes to image floor(row / group), channel
 * = its column, at offset row % group; ConvChunksToImage lays each channel out as an x * y matrix in window order.
 * out: num_images x z x (x*y) doubles.  Padding rows / columns of the edge blocks are stripped (:62-67). */
int nbo_conv_result_to_images(const nbo_set *result, int num_images, int x, int y, int z, double *out) {
    if (result->n == 0) return -1;
    const int total_rows = result->blocks[0].total_rows, total_cols = result->blocks[0].total_cols;
    if (num_images <= 0 || total_rows % num_images != 0) return -1;                 /* the assert of :46 */
    const int group = total_rows / num_images;
    if (group != x * y || total_cols > z) return -1;
    for (int b = 0; b < result->n; b++) {
        const nbo_block *blk = &result->blocks[b];
        const int origin_x = blk->block_row * blk->rows, origin_y = blk->block_col * blk->cols;
        int real_x = total_rows - origin_x, real_y = total_cols - origin_y;
        if (real_x > blk->rows) real_x = blk->rows;
        if (real_y > blk->cols) real_y = blk->cols;
        int image_id = origin_x / group, done = origin_x % group;
        int left = group - done < real_x ? group - done : real_x;
        int i = 0;
        while (i < real_x) {
            for (int j = 0; j < real_y; j++)
                for (int pos = 0; pos < left; pos++)
                    out[((size_t)image_id * z + (origin_y + j)) * group + done + pos] = blk->data[(size_t)(i + pos) * blk->cols + j];
            i += left;
            left = real_x - i < group ? real_x - i : group;
            image_id++;
            done = 0;
        }
    }
    return 0;
}
