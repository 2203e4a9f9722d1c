// nb200_gemm.cu -- the JoinComp + ClusterAggregateComp pair as ONE tensor-core block GEMM.
//
// Replaces: FFTransposeMult / FFInputLayerJoin projection lambdas (one Eigen fp64 GEMM per joined
// block pair, FF/headers/FFTransposeMult.h:80-92, FFInputLayerJoin.h:72-86) followed by the
// aggregation sink's FFMatrixData::operator+ over materialised partial blocks
// (queryExecution/headers/CombinedShuffleSink.h:112-144, AggregationProcessor.cc:131-156).
//
// Here the K-block partial products never exist as blocks: a CTA owns one 128 x BN tile of the
// output and a contiguous range of 64-wide K blocks; TMA streams the bf16 operand tiles into a
// 128B-swizzled shared-memory ring, one elected thread issues tcgen05.mma (kind::f16, bf16 in,
// fp32 accumulate in TMEM) and four epilogue warps drain TMEM.  With split-K the fp32 partial
// tiles go to a workspace that nb200_kernels.cu reduces in a fixed order (deterministic), fused
// with bias / ReLU / re-quantisation for the next layer.
//
// P[M x N] = U[M x K] . V[N x K]^T, U and V K-major bf16 planes (row pitch ldk, multiple of 64,
// zero padded).  nterms = 3 runs the bf16 hi/lo split: hi*hi + hi*lo + lo*hi.
#include "nb200_internal.h"
#include "nb200_ptx.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>

namespace nb200 {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                    // 64 bf16 = 128 bytes = one swizzle row
constexpr int kUmmaK = 16;
constexpr int kThreads = 192;                  // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;

struct GemmArgs {
    int M, N;
    int tiles_n;
    int num_kb;        // 64-wide K blocks in total
    int kb_per_split;
    int nterms;
    int mode;          // 0: partial -> ws, 1: direct epilogue
    float *ws;
    long long ws_split_stride;  // Mpad * Npad
    int ws_ld;                  // Npad
    float *out;
    long long out_ld;
    const float *bias_m;
    const float *bias_n;
    int relu;
    int *err_flag;
    unsigned int idesc;  // tcgen05 instruction descriptor (operand format bf16 or fp16, tile shape)
    const int *exp_u;    // fp16 operand scales (device ints, may be null): result *= 2^-(exp_u + exp_v)
    const int *exp_v;
    int split_offset;    // blockIdx.y + split_offset = global split index (partial launches)
};

template <int BN, int STAGES>
struct SmemLayout {
    static constexpr uint32_t kBBytes = BN * kBlockK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;
    static constexpr uint32_t kBarOffset = STAGES * kStageBytes;
    static constexpr uint32_t kTotal = kBarOffset + (2 * STAGES + 1) * 8 + 16 + 1024 /* alignment slack */;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
k_block_gemm(const __grid_constant__ CUtensorMap map_u_hi, const __grid_constant__ CUtensorMap map_u_lo,
             const __grid_constant__ CUtensorMap map_v_hi, const __grid_constant__ CUtensorMap map_v_lo,
             const GemmArgs g) {
    using L = SmemLayout<BN, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B tiles need 1024-byte alignment
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + L::kBarOffset);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full_bar = empty_bar + STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tile_m = blockIdx.x / g.tiles_n;
    const int tile_n = blockIdx.x % g.tiles_n;
    const int split = blockIdx.y + g.split_offset;
    const int kb0 = split * g.kb_per_split;
    const int kb1 = min(kb0 + g.kb_per_split, g.num_kb);
    const int n_iter = max(kb1 - kb0, 0) * g.nterms;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_u_hi);
        tma_prefetch_desc(&map_v_hi);
        if (g.nterms > 1) {
            tma_prefetch_desc(&map_u_lo);
            tma_prefetch_desc(&map_v_lo);
        }
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, BN);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int it = 0; it < n_iter; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&empty_bar[s], ph ^ 1)) { atomicExch(g.err_flag, 101); break; }
                const int kb = kb0 + it / g.nterms;
                const int t = it % g.nterms;  // 0: hi*hi, 1: hi*lo, 2: lo*hi
                const CUtensorMap *mu = (t == 2) ? &map_u_lo : &map_u_hi;
                const CUtensorMap *mv = (t == 1) ? &map_v_lo : &map_v_hi;
                uint8_t *sa = smem + s * L::kStageBytes;
                uint8_t *sb = sa + kABytes;
                mbar_arrive_expect_tx(&full_bar[s], L::kStageBytes);
                tma_load_2d(mu, &full_bar[s], sa, kb * kBlockK, tile_m * kBlockM, kEvictNormal);
                tma_load_2d(mv, &full_bar[s], sb, kb * kBlockK, tile_n * BN, kEvictLast);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            const uint32_t idesc = g.idesc;
            bool ok = true;
            for (int it = 0; it < n_iter && ok; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&full_bar[s], ph)) { atomicExch(g.err_flag, 102); ok = false; break; }
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + s * L::kStageBytes);
                const uint32_t sb = sa + kABytes;
                const uint64_t da = umma_desc_k_sw128(sa);
                const uint64_t db = umma_desc_k_sw128(sb);
#pragma unroll
                for (int k = 0; k < kBlockK / kUmmaK; k++) {
                    // advance 32 bytes (16 bf16) inside the 128B swizzle row: +2 in 16-byte units
                    umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);  // frees the smem slot when these MMAs retire
            }
            umma_commit(tmem_full_bar);      // accumulator complete
        }
    } else {
        // ===== epilogue warps: TMEM -> registers -> global =====
        const int q = warp & 3;              // TMEM lane quarter this warp may access
        const int row_in_tile = q * 32 + lane;
        const long long row = static_cast<long long>(tile_m) * kBlockM + row_in_tile;
        bool ok = mbar_wait(tmem_full_bar, 0);
        if (!ok && lane == 0) atomicExch(g.err_flag, 103);
        ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
        tc_fence_after();
        if (ok) {
            const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
            const float bm = (g.mode == 1 && g.bias_m != nullptr && row < g.M) ? g.bias_m[row] : 0.f;
            const float oscale = (g.mode == 1 && g.exp_u != nullptr) ? exp2f(-static_cast<float>(*g.exp_u + *g.exp_v)) : 1.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t r[32];
                if (n_iter > 0) {
                    tmem_ld_32x32b_x32(lane_addr + c * 32, r);
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int j = 0; j < 32; j++) r[j] = 0u;
                }
                const int col0 = tile_n * BN + c * 32;
                if (g.mode == 0) {
                    float *dst = g.ws + split * g.ws_split_stride + row * g.ws_ld + col0;
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<uint4 *>(dst + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                } else if (row < g.M) {
                    float *dst = g.out + row * g.out_ld + col0;
                    if (col0 + 32 <= g.N && (g.out_ld & 3) == 0 && (reinterpret_cast<uintptr_t>(g.out) & 15) == 0) {
                        // whole chunk in range: 8 x 16-byte stores, the thread's 128 bytes fill one line of its row
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float v[4];
#pragma unroll
                            for (int e = 0; e < 4; e++) {
                                v[e] = __uint_as_float(r[j + e]) * oscale + bm;
                                if (g.bias_n != nullptr) v[e] += __ldg(g.bias_n + col0 + j + e);
                                if (g.relu) v[e] = fmaxf(v[e], 0.f);
                            }
                            *reinterpret_cast<float4 *>(dst + j) = make_float4(v[0], v[1], v[2], v[3]);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; j++) {
                            if (col0 + j < g.N) {
                                float v = __uint_as_float(r[j]) * oscale + bm;
                                if (g.bias_n != nullptr) v += g.bias_n[col0 + j];
                                if (g.relu) v = fmaxf(v, 0.f);
                                dst[j] = v;
                            }
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}


// ---------------------------------------------------------------------------------------------
// 2-SM variant (tcgen05 cta_group::2).  A CTA pair (cluster of 2) owns one 256 x BN output tile:
// each CTA loads its own 128 rows of U and HALF of the V tile; the leader's single MMA thread issues
// M = 256 instructions that read both CTAs' shared memory and write both CTAs' TMEM.  Per SM and
// 64-wide K step that is 16 KB (U) + BN/2 x 128 B (V) of shared-memory fill for BN/256 x 512 MMA
// cycles: 64 B/cycle at BN = 256 and 47 B/cycle at BN = 512, against 96 B/cycle for the 1-SM
// 128 x 256 tile -- the 1-SM kernel is bound by L2 -> SM bandwidth at about half of tensor peak
// (profiles/r01_*).  BN = 512 uses all 512 TMEM columns (two N = 256 accumulators).
//
// Column bookkeeping: CTA r holds V rows [r*BN/2, (r+1)*BN/2) of the tile as NSUB chunks of 128;
// MMA `sub` consumes chunk `sub` of both CTAs, so TMEM column t = sub*256 + j maps to tile column
// (j < 128 ? sub*128 + j : BN/2 + sub*128 + j - 128).
// ---------------------------------------------------------------------------------------------
template <int NSUB, int STAGES>
struct SmemLayout2 {
    static constexpr int BN = 256 * NSUB;
    static constexpr uint32_t kBBytes = (BN / 2) * kBlockK * 2;
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;
    static constexpr uint32_t kBarOffset = STAGES * kStageBytes;
    static constexpr uint32_t kTotal = kBarOffset + (2 * STAGES + 1) * 8 + 16 + 1024;
};

// MC = true: clusters of 4 = two CTA pairs working on two adjacent M tiles of the SAME N tile; each CTA loads only half of
// its share of the V tile and multicasts it to its counterpart in the other pair, which halves the V traffic from L2
// (the L2 -> SM fabric saturates near 9.5 TB/s on this part and bounds the kernels that stream operands).
template <int NSUB, int STAGES, bool MC>
__global__ void __cluster_dims__(MC ? 4 : 2, 1, 1) __launch_bounds__(kThreads, 1)
k_block_gemm_2sm(const __grid_constant__ CUtensorMap map_u_hi, const __grid_constant__ CUtensorMap map_u_lo,
                 const __grid_constant__ CUtensorMap map_v_hi, const __grid_constant__ CUtensorMap map_v_lo,
                 const GemmArgs g) {
    using L = SmemLayout2<NSUB, STAGES>;
    constexpr int BN = L::BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + L::kBarOffset);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full_bar = empty_bar + STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();     // 0..1, or 0..3 with MC
    const uint32_t rank = crank & 1;              // rank inside the CTA pair
    const uint32_t pair_in_cluster = crank >> 1;  // 0 without MC
    const bool leader = rank == 0;
    int tile_m, tile_n;
    if (MC) {
        const int cl = blockIdx.x >> 2;           // cluster index: (tile_m / 2, tile_n)
        tile_n = cl % g.tiles_n;
        tile_m = (cl / g.tiles_n) * 2 + static_cast<int>(pair_in_cluster);
    } else {
        const int pair = blockIdx.x >> 1;
        tile_m = pair / g.tiles_n;
        tile_n = pair % g.tiles_n;
    }
    const int split = blockIdx.y + g.split_offset;
    const int kb0 = split * g.kb_per_split;
    const int kb1 = min(kb0 + g.kb_per_split, g.num_kb);
    const int n_iter = max(kb1 - kb0, 0) * g.nterms;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_u_hi);
        tma_prefetch_desc(&map_v_hi);
        if (g.nterms > 1) {
            tma_prefetch_desc(&map_u_lo);
            tma_prefetch_desc(&map_v_lo);
        }
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 2);            // leader's arrive.expect_tx + the peer's remote arrive
            mbar_init(&empty_bar[s], MC ? 2 : 1);  // one multicast commit per pair that writes into this CTA's stage
        }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {  // one warp of EACH CTA of the pair takes part in the 2-SM allocation
        tmem_alloc_2sm(tmem_slot, BN);
        tmem_relinquish_2sm();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            for (int it = 0; it < n_iter; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&empty_bar[s], ph ^ 1)) { atomicExch(g.err_flag, 201); break; }
                const int kb = kb0 + it / g.nterms;
                const int t = it % g.nterms;
                const CUtensorMap *mu = (t == 2) ? &map_u_lo : &map_u_hi;
                const CUtensorMap *mv = (t == 1) ? &map_v_lo : &map_v_hi;
                uint8_t *sa = smem + s * L::kStageBytes;
                uint8_t *sb = sa + kABytes;
                if (leader) mbar_arrive_expect_tx(&full_bar[s], 2 * L::kStageBytes);
                else mbar_arrive_remote(&full_bar[s], crank & ~1u);  // the leader of MY pair
                tma_load_2d_2sm(mu, &full_bar[s], sa, kb * kBlockK, tile_m * 256 + static_cast<int>(rank) * kBlockM,
                                kEvictNormal);
                if (!MC) {
                    tma_load_2d_2sm(mv, &full_bar[s], sb, kb * kBlockK, tile_n * BN + static_cast<int>(rank) * (BN / 2),
                                    kEvictLast);
                } else {
                    // my share of V is rows [rank*BN/2, +BN/2); I fetch one half of it (pair 0 the first, pair 1 the second)
                    // and it lands in my stage and in the stage of the same-rank CTA of the other pair
                    constexpr int kHalfRows = BN / 4;
                    const uint16_t mask = static_cast<uint16_t>((1u << rank) | (1u << (rank + 2)));
                    tma_load_2d_2sm_mc(mv, &full_bar[s], sb + pair_in_cluster * (kHalfRows * kBlockK * 2), kb * kBlockK,
                                       tile_n * BN + static_cast<int>(rank) * (BN / 2) + static_cast<int>(pair_in_cluster) * kHalfRows,
                                       mask, kEvictLast);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread of the leader CTA =====
        if (leader && lane == 0) {
            const uint32_t idesc = g.idesc;
            bool ok = true;
            for (int it = 0; it < n_iter && ok; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&full_bar[s], ph)) { atomicExch(g.err_flag, 202); ok = false; break; }
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + s * L::kStageBytes);
                const uint32_t sb = sa + kABytes;
                const uint64_t da = umma_desc_k_sw128(sa);
#pragma unroll
                for (int k = 0; k < kBlockK / kUmmaK; k++) {
#pragma unroll
                    for (int sub = 0; sub < NSUB; sub++) {
                        const uint64_t db = umma_desc_k_sw128(sb + sub * (128 * kBlockK * 2));
                        umma_f16_2sm(tmem_base + sub * 256, da + 2 * k, db + 2 * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    }
                }
                // frees the slot in both CTAs of the pair -- and, with MC, tells the other pair (which writes half of
                // our V tile) as well
                umma_commit_2sm(&empty_bar[s], MC ? 0xF : 0x3);
            }
            umma_commit_2sm(tmem_full_bar, MC ? static_cast<uint16_t>(0x3u << (crank & ~1u)) : 0x3);
        }
    } else {
        // ===== epilogue warps (both CTAs drain their own 128 TMEM lanes) =====
        const int q = warp & 3;
        const long long row = static_cast<long long>(tile_m) * 256 + rank * kBlockM + q * 32 + lane;
        bool ok = mbar_wait(tmem_full_bar, 0);
        if (!ok && lane == 0) atomicExch(g.err_flag, 203);
        ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
        tc_fence_after();
        if (ok) {
            const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
            const float bm = (g.mode == 1 && g.bias_m != nullptr && row < g.M) ? g.bias_m[row] : 0.f;
            const float oscale = (g.mode == 1 && g.exp_u != nullptr) ? exp2f(-static_cast<float>(*g.exp_u + *g.exp_v)) : 1.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(lane_addr + c * 32, r);
                tmem_ld_wait();
                const int sub = (c * 32) / 256, j = (c * 32) % 256;
                const int tcol = (j < 128) ? sub * 128 + j : BN / 2 + sub * 128 + (j - 128);
                const int col0 = tile_n * BN + tcol;
                if (g.mode == 0) {
                    float *dst = g.ws + split * g.ws_split_stride + row * g.ws_ld + col0;
#pragma unroll
                    for (int jj = 0; jj < 32; jj += 4)
                        *reinterpret_cast<uint4 *>(dst + jj) = make_uint4(r[jj], r[jj + 1], r[jj + 2], r[jj + 3]);
                } else if (row < g.M) {
                    float *dst = g.out + row * g.out_ld + col0;
                    if (col0 + 32 <= g.N && (g.out_ld & 3) == 0 && (reinterpret_cast<uintptr_t>(g.out) & 15) == 0) {
#pragma unroll
                        for (int jj = 0; jj < 32; jj += 4) {
                            float v[4];
#pragma unroll
                            for (int e = 0; e < 4; e++) {
                                v[e] = __uint_as_float(r[jj + e]) * oscale + bm;
                                if (g.bias_n != nullptr) v[e] += __ldg(g.bias_n + col0 + jj + e);
                                if (g.relu) v[e] = fmaxf(v[e], 0.f);
                            }
                            *reinterpret_cast<float4 *>(dst + jj) = make_float4(v[0], v[1], v[2], v[3]);
                        }
                    } else {
#pragma unroll
                        for (int jj = 0; jj < 32; jj++) {
                            if (col0 + jj < g.N) {
                                float v = __uint_as_float(r[jj]) * oscale + bm;
                                if (g.bias_n != nullptr) v += g.bias_n[col0 + jj];
                                if (g.relu) v = fmaxf(v, 0.f);
                                dst[jj] = v;
                            }
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    cluster_sync_all();  // the peer's TMEM / smem must not go away while the leader's MMAs use them
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2sm(tmem_base, BN);
    }
}

// ---- host side -----------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2D bf16 K-major plane: dim0 = K (padded pitch), dim1 = rows; box = 64 x box_rows, SWIZZLE_128B.
static int make_plane_map(CUtensorMap *m, const __nv_bfloat16 *base, int64_t rows, int64_t ldk, int box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return NB200_ERR_CUDA;
    }
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(ldk), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(ldk) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16 *>(base), dims, strides, box,
                     estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld ldk=%lld box_rows=%d)", (int)r,
                  (long long)rows, (long long)ldk, box_rows);
        return NB200_ERR_CUDA;
    }
    return NB200_OK;
}

// Split-K factor: the skinny FF shapes (Amazon-14K layer 1: 79 x 4 tiles, K = 597,540) have far
// fewer output tiles than SMs x waves, so K is cut until the grid is a whole number of waves and
// every CTA still runs a long (>= 32 K-block) main loop.
static int pick_kernel(int M, int N, int64_t ldk);

int choose_splits(Context *c, int M, int N, int64_t ldk, int nterms) {
    (void)nterms;
    const int kind = pick_kernel(M, N, ldk);
    const int BN = kind == 2 ? 512 : (kind == 1 ? 256 : ((N <= 128) ? 128 : 256));
    const int BM = kind == 0 ? kBlockM : 256;
    const int64_t tiles = ceil_div(M, BM) * ceil_div(N, BN);
    const int64_t num_kb = ldk / kBlockK;
    int units = c->sm_count > 0 ? c->sm_count : 148;  // concurrently resident tiles
    if (kind != 0) units /= 2;
    if (tiles >= 4 * units || num_kb < 64) return 1;
    int best = 1;
    double best_cost = 1e30;
    const int max_s = static_cast<int>(num_kb / 32 > 0 ? (num_kb / 32 < 256 ? num_kb / 32 : 256) : 1);
    for (int s = 1; s <= max_s; s++) {
        const int64_t per = ceil_div(num_kb, s);
        if (ceil_div(num_kb, per) != s) continue;  // would leave an empty split
        const int64_t waves = ceil_div(tiles * s, units);
        // time ~ waves * (main loop + fixed per-CTA prologue/epilogue) + reduction traffic
        const double cost = waves * (per + 8.0) + 0.02 * s;
        if (cost < best_cost) { best_cost = cost; best = s; }
    }
    return best;
}

template <int NSUB, int STAGES, bool MC>
static int launch_impl_2sm(Context *c, const CUtensorMap &mu_hi, const CUtensorMap &mu_lo, const CUtensorMap &mv_hi,
                           const CUtensorMap &mv_lo, const GemmArgs &g, int pair_tiles, int splits, cudaStream_t stream) {
    using L = SmemLayout2<NSUB, STAGES>;
    static bool attr_set_dev[16] = {};   // function attributes are per device
    bool &attr_set = attr_set_dev[c->device & 15];
    if (!attr_set) {
        NB_CUDA(cudaFuncSetAttribute(k_block_gemm_2sm<NSUB, STAGES, MC>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    dim3 grid(2 * pair_tiles, splits);  // with MC pair_tiles is even: clusters of 4 CTAs
    k_block_gemm_2sm<NSUB, STAGES, MC><<<grid, kThreads, L::kTotal, stream>>>(mu_hi, mu_lo, mv_hi, mv_lo, g);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}

// Which kernel / tile for a shape.  NB200_GEMM_KERNEL=1sm|2sm256|2sm512 overrides (experiments).
static int pick_kernel(int M, int N, int64_t ldk) {
    static int forced = -2;
    if (forced == -2) {
        const char *e = getenv("NB200_GEMM_KERNEL");
        forced = -1;
        if (e) {
            if (!strcmp(e, "1sm")) forced = 0;
            else if (!strcmp(e, "2sm256")) forced = 1;
            else if (!strcmp(e, "2sm512")) forced = 2;
        }
    }
    if (forced >= 0) return forced;
    if (M <= 128 || N <= 128) return 0;          // a pair tile would be mostly padding
    if (ldk >= 4096) return N > 256 ? 2 : 1;     // long K: the L2 -> SM fill rate decides
    static int shortk = -2;
    if (shortk == -2) {
        const char *e = getenv("NB200_GEMM_KERNEL_SHORTK");
        shortk = e ? atoi(e) : 0;
    }
    return (M >= 1024 && N >= 1024) ? shortk : 0;
}

template <int BN, int STAGES>
static int launch_impl(Context *c, const CUtensorMap &mu_hi, const CUtensorMap &mu_lo, const CUtensorMap &mv_hi,
                       const CUtensorMap &mv_lo, const GemmArgs &g, int tiles, int splits, cudaStream_t stream) {
    using L = SmemLayout<BN, STAGES>;
    static bool attr_set_dev[16] = {};   // function attributes are per device
    bool &attr_set = attr_set_dev[c->device & 15];
    if (!attr_set) {
        NB_CUDA(cudaFuncSetAttribute(k_block_gemm<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    dim3 grid(tiles, splits);
    k_block_gemm<BN, STAGES><<<grid, kThreads, L::kTotal, stream>>>(mu_hi, mu_lo, mv_hi, mv_lo, g);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}

int launch_gemm_tc(Context *c, const Planes &u, const Planes &v, int M, int N, int nterms, int splits,
                   const GemmEpilogue &epi, int *Mpad_out, int *Npad_out, int *splits_used, int fmt) {
    const bool f16 = fmt == kFmtF16;
    if (f16 && (nterms != 1 || !u.f16_valid || !v.f16_valid)) {
        set_error("block GEMM: fp16 operands missing (or asked for a split-precision pass)");
        return NB200_ERR_STATE;
    }
    if ((!f16 && (!u.valid || !v.valid)) || u.ldk % kBlockK != 0 || v.ldk % kBlockK != 0) {
        set_error("block GEMM: operand planes missing or K pitch not a multiple of %d (%lld, %lld)", kBlockK, (long long)u.ldk,
                  (long long)v.ldk);
        return NB200_ERR_STATE;
    }
    // operands whose K extents differ (a join that matches only the leading column blocks, see gemm_blocks): the
    // planes are zero beyond their own extent, so the product over the shorter pitch is the product over min(Ka, Kb)
    const int64_t ldk_eff = u.ldk < v.ldk ? u.ldk : v.ldk;
    if (nterms == 3 && (!u.lo_valid || !v.lo_valid)) {
        set_error("block GEMM: BF16X3 needs the residual planes");
        return NB200_ERR_STATE;
    }
    const int kind = pick_kernel(M, N, ldk_eff);  // 0: 1-SM 128 x {128,256}; 1: 2-SM 256 x 256; 2: 2-SM 256 x 512
    const int BN = kind == 2 ? 512 : (kind == 1 ? 256 : ((N <= 128) ? 128 : 256));
    const int BM = kind == 0 ? kBlockM : 256;
    int tiles_m = static_cast<int>(ceil_div(M, BM));
    const int tiles_n = static_cast<int>(ceil_div(N, BN));
    // V-tile multicast across two M tiles (clusters of 4) for the 2-SM kernels whenever there are at least two M tiles
    static const bool mc_env = !(getenv("NB200_GEMM_MC") && atoi(getenv("NB200_GEMM_MC")) == 0);
    // (an odd tile count gets an all-padding partner tile: only worth it when that is a small fraction of the work)
    const bool mc = mc_env && kind != 0 && tiles_m >= 2 && (tiles_m % 2 == 0 || tiles_m >= 9);
    if (mc) tiles_m = (tiles_m + 1) & ~1;
    const int Mpad = tiles_m * BM, Npad = tiles_n * BN;
    if (Mpad_out) *Mpad_out = Mpad;
    if (Npad_out) *Npad_out = Npad;
    const int num_kb = static_cast<int>(ldk_eff / kBlockK);
    if (splits < 1) splits = 1;
    if (splits > num_kb) splits = num_kb > 0 ? num_kb : 1;
    int per = static_cast<int>(ceil_div(num_kb, splits));
    splits = static_cast<int>(ceil_div(num_kb, per > 0 ? per : 1));
    if (splits < 1) splits = 1;
    if (splits_used) *splits_used = splits;
    if (epi.mode == 1 && splits != 1) {
        set_error("block GEMM: direct epilogue requires splits == 1");
        return NB200_ERR_INVALID;
    }

    GemmArgs g{};
    g.M = M; g.N = N; g.tiles_n = tiles_n; g.num_kb = num_kb; g.kb_per_split = per; g.nterms = nterms;
    g.mode = epi.mode;
    if (epi.mode == 0) {
        const size_t need = static_cast<size_t>(splits) * Mpad * Npad * sizeof(float);
        NB_TRY(ensure_ws(c, need));
        g.ws = c->ws;
        g.ws_split_stride = static_cast<long long>(Mpad) * Npad;
        g.ws_ld = Npad;
    } else {
        g.out = epi.out; g.out_ld = epi.out_ld; g.bias_m = epi.bias_m; g.bias_n = epi.bias_n; g.relu = epi.relu;
    }
    g.err_flag = c->err_flag;
    // operand format field of the instruction descriptor: 1 = BF16, 0 = F16 (bits 7-9 and 10-12)
    g.idesc = umma_idesc_bf16(kind == 0 ? kBlockM : 256, kind == 0 ? BN : 256);
    if (f16) {
        g.idesc &= ~((7u << 7) | (7u << 10));
        g.exp_u = u.f16_exp;
        g.exp_v = v.f16_exp;
    }
    const __nv_bfloat16 *u_hi = f16 ? reinterpret_cast<const __nv_bfloat16 *>(u.f16) : u.hi;  // TMA only moves 16-bit words
    const __nv_bfloat16 *v_hi = f16 ? reinterpret_cast<const __nv_bfloat16 *>(v.f16) : v.hi;

    CUtensorMap mu_hi, mu_lo, mv_hi, mv_lo;
    const int v_box = kind == 0 ? BN : (mc ? BN / 4 : BN / 2);  // 2-SM: half of the V tile per CTA; with multicast a quarter
    NB_TRY(make_plane_map(&mu_hi, u_hi, u.rows, u.ldk, kBlockM));
    NB_TRY(make_plane_map(&mv_hi, v_hi, v.rows, v.ldk, v_box));
    if (nterms == 3) {
        NB_TRY(make_plane_map(&mu_lo, u.lo, u.rows, u.ldk, kBlockM));
        NB_TRY(make_plane_map(&mv_lo, v.lo, v.rows, v.ldk, v_box));
    } else {
        mu_lo = mu_hi;
        mv_lo = mv_hi;
    }
    // partial launch: only splits [split_begin, split_begin + split_count)
    int launch_splits = splits;
    if (epi.split_count >= 0) {
        g.split_offset = epi.split_begin;
        launch_splits = epi.split_count;
        if (epi.split_begin + launch_splits > splits) launch_splits = splits - epi.split_begin;
        if (launch_splits <= 0) return NB200_OK;
    }
    cudaStream_t stream = epi.stream ? epi.stream : c->stream;
    if (kind == 2 && mc) return launch_impl_2sm<2, 4, true>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
    if (kind == 1 && mc) return launch_impl_2sm<1, 6, true>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
    if (kind == 2) return launch_impl_2sm<2, 4, false>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
    if (kind == 1) return launch_impl_2sm<1, 6, false>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
    if (BN == 256) return launch_impl<256, 4>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
    return launch_impl<128, 6>(c, mu_hi, mu_lo, mv_hi, mv_lo, g, tiles_m * tiles_n, launch_splits, stream);
}



// ---------------------------------------------------------------------------------------------
// Fused layer 1: fp64 inputs -> fp16 -> tcgen05 in ONE persistent kernel.
//
// The two-kernel path reads X (fp64, 47.8 GB on the Amazon shape), writes its fp16 plane (12 GB) and reads the plane
// back in the GEMM: 9.2 ms + 8.8 ms, both at their own rooflines (HBM, tensor).  Here every CTA pair owns ONE
// 256 x 512 output tile for the whole K (accumulators stay in TMEM, no split-K), all pairs sweep K together, and
// eight extra "converter" warps per CTA convert the next K slabs of X into a small ring of fp16 slabs that lives in
// L2 (tens of MB) and is consumed by the TMA producers a few slabs later.  The fp16 copy of X never goes to HBM, so
// the phase costs max(read X once from HBM, tensor time) instead of their sum.
//
//   converter warps:  wait done[S - ring] == pairs      (slot free)      | consumer (TMA thread of each CTA):
//                     convert rows x 256 columns of super-slab S -> ring |   wait ready[S] == CTAs, fence.proxy.async,
//                     fence, ++ready[S]                                  |   TMA ring slab + W1 tile -> smem ring
//   MMA leader:       after the last K block of S has landed in smem: ++done[S]
//
// All CTAs must be resident at once (grid = 2 x pairs <= SM count, 1 CTA / SM); every wait is bounded.
// ---------------------------------------------------------------------------------------------
constexpr int kCvtWarps = 12;
constexpr int kFusedThreads = kThreads + kCvtWarps * 32;  // 448
// K blocks (of 64) per conversion super-slab are a run-time parameter (FusedArgs::super_kb)

struct FusedArgs {
    const double *x;        // [rows_total x cols], pitch x_ld
    long long x_ld;
    int rows_total, cols;   // B, F
    int row_base;           // first row of this launch
    int rows_fused;         // rows covered by this launch (multiple of 256, may exceed rows_total - row_base)
    int tiles_n;            // N tiles of 512
    int num_kb;             // ldk / 64
    int super_kb;           // K blocks per conversion super-slab (multiple of 4: converters work in 256-column pieces)
    int ring_kb;            // ring length in K blocks (multiple of super_kb)
    __half *ring;           // [rows_fused][ring_kb * 64]
    const int *exp_x;       // fp16 scale exponent of X
    const int *exp_v;       // ... of W1
    int *ready;             // [num_super]
    int *done;              // [num_super]
    float *row_sumsq;       // [rows_total], zeroed by the host
    float *maxscaled;       // largest |x * 2^e| seen (overflow check)
    float *out;             // y1 fp32 [rows_total x N], pitch out_ld, receives relu(acc * scale + bias)
    long long out_ld;
    const float *bias_n;
    int N;
    float *out_maxabs;
    int *err_flag;
    unsigned int idesc;
    int debug;              // experiments: bit 0 = converters skip the data movement, bit 1 = no TMA/MMA work
};

__device__ __forceinline__ int ld_acquire_gpu(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool wait_flag_ge(const int *p, int target) {
    if (ld_acquire_gpu(p) >= target) return true;
    const long long t0 = clock64();
    while (ld_acquire_gpu(p) < target) {
        __nanosleep(64);
        if (clock64() - t0 > (1ll << 31)) return false;
    }
    return true;
}


template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFusedThreads, 1)
k_l1_fused(const __grid_constant__ CUtensorMap map_ring, const __grid_constant__ CUtensorMap map_v, const FusedArgs g) {
    using L = SmemLayout2<2, STAGES>;
    constexpr int BN = 512;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + L::kBarOffset);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full_bar = empty_bar + STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair = blockIdx.x >> 1;
    const int tile_m = pair / g.tiles_n, tile_n = pair % g.tiles_n;
    const int kSuperKb = g.super_kb;
    const int num_super = (g.num_kb + kSuperKb - 1) / kSuperKb;
    const int ring_super = g.ring_kb / kSuperKb;
    const int ring_ld = g.ring_kb * kBlockK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_ring);
        tma_prefetch_desc(&map_v);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 2);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_2sm(tmem_slot, BN);
        tmem_relinquish_2sm();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs): ring slab of X (fp16) + W1 tile =====
        if (lane == 0) {
            for (int kb = 0; kb < g.num_kb; kb++) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                if (kb % kSuperKb == 0) {
                    // only the 2 * tiles_n CTAs that share this M tile convert its rows
                    if (!wait_flag_ge(g.ready + tile_m * num_super + kb / kSuperKb, 2 * g.tiles_n)) { atomicExch(g.err_flag, 404); break; }
                    asm volatile("fence.proxy.async;" ::: "memory");  // converters wrote the slab through the generic proxy
                }
                if (g.debug & 2) continue;
                if (!mbar_wait(&empty_bar[s], ph ^ 1)) { atomicExch(g.err_flag, 401); break; }
                uint8_t *sa = smem + s * L::kStageBytes;
                uint8_t *sb = sa + kABytes;
                if (leader) mbar_arrive_expect_tx(&full_bar[s], 2 * L::kStageBytes);
                else mbar_arrive_remote(&full_bar[s], 0);
                tma_load_2d_2sm(&map_ring, &full_bar[s], sa, (kb % g.ring_kb) * kBlockK,
                                tile_m * 256 + static_cast<int>(rank) * kBlockM, kEvictNormal);
                tma_load_2d_2sm(&map_v, &full_bar[s], sb, kb * kBlockK, tile_n * BN + static_cast<int>(rank) * (BN / 2), kEvictLast);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader) =====
        if (leader && lane == 0) {
            for (int kb = 0; kb < g.num_kb; kb++) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                if (g.debug & 2) {  // experiment: consume the flags only
                    if (kb % kSuperKb == 0 && !wait_flag_ge(g.ready + tile_m * num_super + kb / kSuperKb, 2 * g.tiles_n)) break;
                    if (kb % kSuperKb == kSuperKb - 1 || kb == g.num_kb - 1) atomicAdd(g.done + tile_m * num_super + kb / kSuperKb, 1);
                    continue;
                }
                if (!mbar_wait(&full_bar[s], ph)) { atomicExch(g.err_flag, 402); break; }
                tc_fence_after();
                // both CTAs' loads of this K block have landed: once the super-slab's last block is in, its ring slot is free
                if (kb % kSuperKb == kSuperKb - 1 || kb == g.num_kb - 1) atomicAdd(g.done + tile_m * num_super + kb / kSuperKb, 1);
                const uint32_t sa = smem_u32(smem + s * L::kStageBytes);
                const uint32_t sb = sa + kABytes;
                const uint64_t da = umma_desc_k_sw128(sa);
#pragma unroll
                for (int k = 0; k < kBlockK / kUmmaK; k++) {
#pragma unroll
                    for (int sub = 0; sub < 2; sub++) {
                        const uint64_t db = umma_desc_k_sw128(sb + sub * (128 * kBlockK * 2));
                        umma_f16_2sm(tmem_base + sub * 256, da + 2 * k, db + 2 * k, g.idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    }
                }
                umma_commit_2sm(&empty_bar[s], 0x3);
            }
            umma_commit_2sm(tmem_full_bar, 0x3);
        }
    } else if (warp < 6) {
        // ===== epilogue: y1 = relu(acc * 2^-(ex+ev) + b1), fp32, and its maximum =====
        const int q = warp & 3;
        const long long row = static_cast<long long>(g.row_base) + tile_m * 256 + rank * kBlockM + q * 32 + lane;
        bool ok = mbar_wait(tmem_full_bar, 0);
        if (!ok && lane == 0) atomicExch(g.err_flag, 403);
        ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
        tc_fence_after();
        if (ok) {
            const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
            const float oscale = exp2f(-static_cast<float>(*g.exp_x + *g.exp_v));
            float mx = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(lane_addr + c * 32, r);
                tmem_ld_wait();
                const int sub = (c * 32) / 256, j = (c * 32) % 256;
                const int tcol = (j < 128) ? sub * 128 + j : BN / 2 + sub * 128 + (j - 128);
                const int col0 = tile_n * BN + tcol;
                if (row < g.rows_total) {
                    float *dst = g.out + row * g.out_ld + col0;
#pragma unroll
                    for (int jj = 0; jj < 32; jj++) {
                        if (col0 + jj < g.N) {
                            const float v = fmaxf(__uint_as_float(r[jj]) * oscale + g.bias_n[col0 + jj], 0.f);
                            dst[jj] = v;
                            mx = fmaxf(mx, v);
                        }
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            if (lane == 0) atomicMax(reinterpret_cast<unsigned int *>(g.out_maxabs), __float_as_uint(mx));
        }
    } else {
        // ===== converter warps: this CTA's share of the rows, one 256-column super-slab at a time =====
        const int cw = warp - 6;                        // 0..kCvtWarps-1
        const int ct = threadIdx.x - kThreads;          // 0..kCvtWarps*32-1
        // the 256 rows of an M tile are converted by the 2 * tiles_n CTAs that consume it, so every handshake is
        // local to that small group and the groups sweep K independently of one another
        const int rpc = 256 / (2 * g.tiles_n);               // rows per CTA (64 or 128)
        const int r_begin = tile_m * 256 + (tile_n * 2 + static_cast<int>(rank)) * rpc;
        const float scale = exp2f(static_cast<float>(*g.exp_x));
        __shared__ int s_abort;                          // written by one thread, read by all after the barrier
        if (ct == 0) s_abort = 0;
        asm volatile("bar.sync 1, %0;" ::"n"(kCvtWarps * 32) : "memory");
        uint32_t ovf = 0;                                // any converted value overflowed fp16
        float ss_acc[16];                                // per-warp rows: rpc / 8 <= 16
#pragma unroll
        for (int i = 0; i < 16; i++) ss_acc[i] = 0.f;
        bool alive = true;
        for (int S = 0; S < num_super && alive; S++) {
            if (S >= ring_super) {
                if (ct == 0 && !wait_flag_ge(g.done + tile_m * num_super + S - ring_super, g.tiles_n)) { atomicExch(g.err_flag, 405); s_abort = 1; }
                asm volatile("bar.sync 1, %0;" ::"n"(kCvtWarps * 32) : "memory");
                if (*reinterpret_cast<volatile int *>(&s_abort) != 0) alive = false;  // same value for all 256 threads
                if (!alive) break;
            }
            // (row, 256-column piece) items of this warp; per item a lane loads 4 x 16 bytes at a 512-byte stride so
            // that every load instruction of the warp is one contiguous 512-byte run (streaming, evict-first)
            const int pieces = kSuperKb / 4;
            const int my_rows = (rpc - cw + kCvtWarps - 1) / kCvtWarps;  // rows cw, cw+8, ... below rpc
            const int items = my_rows * pieces;
            // software pipelined by hand in groups of GI items: all loads of a group are issued before any of its
            // conversions, so GI x 4 x 16 bytes per lane are in flight (the compiler will not hoist loads over the
            // ring stores by itself)
            constexpr int GI = 4;
            for (int it0 = 0; it0 < ((g.debug & 1) ? 0 : items); it0 += GI) {
                double2 v[GI][4];
#pragma unroll
                for (int u = 0; u < GI; u++) {
                    const int it = it0 + u;
                    const int i = it / pieces, q = it % pieces;
                    const int rl = r_begin + cw + i * kCvtWarps;
                    const long long row = static_cast<long long>(g.row_base) + rl;
                    const long long cbase = (static_cast<long long>(S) * kSuperKb + q * 4) * kBlockK;
                    const double *p = g.x + row * g.x_ld + cbase + lane * 2;
                    const bool in = it < items && rl < g.rows_fused && row < g.rows_total;
                    if (in && cbase + 255 < g.cols && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
#pragma unroll
                        for (int e = 0; e < 4; e++) v[u][e] = __ldcs(reinterpret_cast<const double2 *>(p + e * 64));
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const long long cc = cbase + e * 64 + lane * 2;
                            v[u][e].x = (in && cc < g.cols) ? g.x[row * g.x_ld + cc] : 0.0;
                            v[u][e].y = (in && cc + 1 < g.cols) ? g.x[row * g.x_ld + cc + 1] : 0.0;
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < GI; u++) {
                    const int it = it0 + u;
                    const int i = it / pieces, q = it % pieces;
                    const int rl = r_begin + cw + i * kCvtWarps;
                    if (it >= items || rl >= g.rows_fused) continue;
                    const int ring_col = ((S % ring_super) * kSuperKb + q * 4) * kBlockK;
                    __half *dst = g.ring + static_cast<long long>(rl) * ring_ld + ring_col + lane * 2;
                    float ss = 0.f;
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const float f0 = static_cast<float>(v[u][e].x), f1 = static_cast<float>(v[u][e].y);
                        const __half2 h2 = __floats2half2_rn(f0 * scale, f1 * scale);
                        *reinterpret_cast<__half2 *>(dst + e * 64) = h2;
                        ss = fmaf(f0, f0, fmaf(f1, f1, ss));
                        // overflow check on the packed result: an all-ones fp16 exponent (inf) in either half
                        const uint32_t hb = *reinterpret_cast<const uint32_t *>(&h2);
                        ovf |= ((hb & 0x7C00u) == 0x7C00u) | ((hb & 0x7C000000u) == 0x7C000000u);
                    }
                    ss_acc[i] += ss;
                }
            }
            asm volatile("fence.proxy.async;" ::: "memory");
            __threadfence();
            asm volatile("bar.sync 1, %0;" ::"n"(kCvtWarps * 32) : "memory");
            if (ct == 0) atomicAdd(g.ready + tile_m * num_super + S, 1);
        }
        // per-row sum of squares (error model) and the overflow check value
#pragma unroll
        for (int i = 0; i < 16; i++) {
            float ss = ss_acc[i];
            for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            const int rl = r_begin + cw + i * kCvtWarps;
            const long long row = static_cast<long long>(g.row_base) + rl;
            if (lane == 0 && i * kCvtWarps + cw < rpc && rl < g.rows_fused && row < g.rows_total) atomicAdd(g.row_sumsq + row, ss);
        }
        // the host compares maxscaled with 65504: report +inf when anything overflowed
        if (__any_sync(0xffffffffu, ovf != 0) && lane == 0)
            atomicMax(reinterpret_cast<unsigned int *>(g.maxscaled), 0x7F800000u);
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2sm(tmem_base, BN);
    }
}

// Host side.  Returns NB200_ERR_STATE (without launching) when the shape does not suit the fused kernel.
int launch_l1_fused(Context *c, const double *x, int64_t x_ld, int rows_total, int cols, int row_base, int n_mtiles,
                    const Planes &xp, const Planes &w1, int N, const float *bias_n, float *out, int64_t out_ld,
                    float *out_maxabs, FusedScratch *fs) {
    const int tiles_n = static_cast<int>(ceil_div(N, 512));
    const int npairs = n_mtiles * tiles_n;
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    if (npairs < 1 || 2 * npairs > sms || !w1.f16_valid || w1.ldk % kBlockK != 0) {
        set_error("fused layer 1: shape not supported");
        return NB200_ERR_STATE;
    }
    const int num_kb = static_cast<int>(w1.ldk / kBlockK);
    static int super_env = getenv("NB200_SUPER_KB") ? atoi(getenv("NB200_SUPER_KB")) : 16;
    const int kSuperKb = (super_env / 4 > 0 ? super_env / 4 : 1) * 4;
    const int num_super = (num_kb + kSuperKb - 1) / kSuperKb;
    const int rows_fused = n_mtiles * 256;
    const int nctas = 2 * npairs;
    if (tiles_n > 2) { set_error("fused layer 1: more than 2 N tiles are not supported"); return NB200_ERR_STATE; }
    static int ring_kb_env = getenv("NB200_RING_KB") ? atoi(getenv("NB200_RING_KB")) : 32;
    const int ring_kb = (ring_kb_env / kSuperKb > 1 ? ring_kb_env / kSuperKb : 2) * kSuperKb;
    const size_t ring_bytes = static_cast<size_t>(rows_fused) * ring_kb * kBlockK * sizeof(__half);
    if (fs->ring_bytes < ring_bytes) {
        nb_free(fs->ring);
        fs->ring = nullptr;
        fs->ring_bytes = 0;
        NB_CUDA(nb_malloc(&fs->ring, ring_bytes));
        fs->ring_bytes = ring_bytes;
    }
    NB_CUDA(cudaMemsetAsync(fs->ring, 0, ring_bytes, c->stream));  // rows beyond the batch must read as zero
    const int nflags = 2 * n_mtiles * num_super;  // ready[tile_m][S], done[tile_m][S]
    if (fs->flag_cap < nflags) {
        nb_free(fs->flags);
        fs->flags = nullptr;
        fs->flag_cap = 0;
        NB_CUDA(nb_malloc(&fs->flags, static_cast<size_t>(nflags) * sizeof(int)));
        fs->flag_cap = nflags;
    }
    NB_CUDA(cudaMemsetAsync(fs->flags, 0, static_cast<size_t>(nflags) * sizeof(int), c->stream));
    // out_maxabs is NOT cleared here: the caller may already have folded the left-over rows' maximum into it

    FusedArgs g{};
    g.x = x; g.x_ld = x_ld; g.rows_total = rows_total; g.cols = cols; g.row_base = row_base; g.rows_fused = rows_fused;
    g.tiles_n = tiles_n; g.num_kb = num_kb; g.super_kb = kSuperKb; g.ring_kb = ring_kb; g.ring = fs->ring;
    g.exp_x = xp.f16_exp; g.exp_v = w1.f16_exp; g.ready = fs->flags; g.done = fs->flags + n_mtiles * num_super;
    g.row_sumsq = xp.row_sumsq; g.maxscaled = xp.f16_maxscaled; g.out = out; g.out_ld = out_ld; g.bias_n = bias_n; g.N = N;
    g.out_maxabs = out_maxabs; g.err_flag = c->err_flag;
    g.idesc = umma_idesc_bf16(256, 256) & ~((7u << 7) | (7u << 10));  // fp16 operands
    g.debug = getenv("NB200_FUSED_DEBUG") ? atoi(getenv("NB200_FUSED_DEBUG")) : 0;
    CUtensorMap mring, mv;
    NB_TRY(make_plane_map(&mring, reinterpret_cast<const __nv_bfloat16 *>(fs->ring), rows_fused, static_cast<int64_t>(ring_kb) * kBlockK, kBlockM));
    NB_TRY(make_plane_map(&mv, reinterpret_cast<const __nv_bfloat16 *>(w1.f16), w1.rows, w1.ldk, 256));
    using L = SmemLayout2<2, 4>;
    static bool attr_set_dev[16] = {};   // function attributes are per device
    bool &attr_set = attr_set_dev[c->device & 15];
    if (!attr_set) {
        NB_CUDA(cudaFuncSetAttribute(k_l1_fused<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    k_l1_fused<4><<<nctas, kFusedThreads, L::kTotal, c->stream>>>(mring, mv, g);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}

// ---------------------------------------------------------------------------------------------
// conv2d_memory_fusion as an implicit GEMM (K6).  The reference materialises the im2col matrix
// `image_flat` through ImageToChunks -> ImageChunksToBlock -> ImageBlockToMatrix
// (conv2d_memory_fusion/headers/ImageToChunks.h:43-102, ImageBlockToMatrix.h:46-73) and multiplies it
// with `kernel_flat` by FFTransposeMult + FFAggMatrix (tests/source/PipelinedConv2dMemFuseTest.cc:249-262).
// Here image_flat never exists: the image is kept NHWC in bf16 (channels padded to 64), a CTA owns one
// run of up to 128 output positions of one output row, and for every filter tap (i, j) a 4-D TMA box
// {64 channels, 128 pixels, 1, 1} at (c0, wo0 + j - pad, ho + i - pad, n) lands in shared memory exactly
// as a K-major 128-byte-swizzled operand tile (out-of-image pixels are zero filled by TMA = padding).
// K is ordered (tap, channel); the bias column of kernel_flat becomes the epilogue bias.
// result[(n, ho, wo), cout] = conv + bias (+ ReLU).  Stride 1.
// ---------------------------------------------------------------------------------------------
struct ConvArgs {
    int n, h, w, hout, wout, k, pad, cblocks;  // cblocks = padded channels / 64
    int kout;
    int wsegs;                                  // ceil(wout / 128)
    int nterms;
    float *out;                                 // [(n*hout + ho)*wout + wo][kout]
    const float *bias;
    int relu;
    int *err_flag;
    unsigned int idesc;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
k_conv_implicit(const __grid_constant__ CUtensorMap map_img_hi, const __grid_constant__ CUtensorMap map_img_lo,
                const __grid_constant__ CUtensorMap map_ker_hi, const __grid_constant__ CUtensorMap map_ker_lo,
                const ConvArgs g) {
    using L = SmemLayout<BN, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + L::kBarOffset);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full_bar = empty_bar + STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_full_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // tile -> (image, output row, 128-wide segment of that row)
    const int seg = blockIdx.x % g.wsegs;
    const int ho = (blockIdx.x / g.wsegs) % g.hout;
    const int img = blockIdx.x / (g.wsegs * g.hout);
    const int wo0 = seg * kBlockM;
    const int taps = g.k * g.k;
    const int n_iter = taps * g.cblocks * g.nterms;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_img_hi);
        tma_prefetch_desc(&map_ker_hi);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, BN < 32 ? 32 : BN);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < n_iter; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&empty_bar[s], ph ^ 1)) { atomicExch(g.err_flag, 301); break; }
                const int t = it % g.nterms;
                const int kb = it / g.nterms;           // = tap * cblocks + cb
                const int tap = kb / g.cblocks, cb = kb % g.cblocks;
                const int fi = tap / g.k, fj = tap % g.k;
                const CUtensorMap *mi = (t == 2) ? &map_img_lo : &map_img_hi;
                const CUtensorMap *mk = (t == 1) ? &map_ker_lo : &map_ker_hi;
                uint8_t *sa = smem + s * L::kStageBytes;
                uint8_t *sb = sa + kABytes;
                mbar_arrive_expect_tx(&full_bar[s], L::kStageBytes);
                tma_load_4d(mi, &full_bar[s], sa, cb * kBlockK, wo0 + fj - g.pad, ho + fi - g.pad, img, kEvictNormal);
                tma_load_2d(mk, &full_bar[s], sb, kb * kBlockK, 0, kEvictLast);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            for (int it = 0; it < n_iter; it++) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                if (!mbar_wait(&full_bar[s], ph)) { atomicExch(g.err_flag, 302); break; }
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + s * L::kStageBytes);
                const uint64_t da = umma_desc_k_sw128(sa), db = umma_desc_k_sw128(sa + kABytes);
#pragma unroll
                for (int k = 0; k < kBlockK / kUmmaK; k++)
                    umma_f16(tmem_base, da + 2 * k, db + 2 * k, g.idesc, (it > 0 || k > 0) ? 1u : 0u);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        const int q = warp & 3;
        const int wo = wo0 + q * 32 + lane;
        bool ok = mbar_wait(tmem_full_bar, 0);
        if (!ok && lane == 0) atomicExch(g.err_flag, 303);
        ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
        tc_fence_after();
        if (ok) {
            const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
            const long long row = (static_cast<long long>(img) * g.hout + ho) * g.wout + wo;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(lane_addr + c * 32, r);
                tmem_ld_wait();
                if (wo < g.wout) {
                    float *dst = g.out + row * g.kout + c * 32;
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const int col = c * 32 + j;
                        if (col < g.kout) {
                            float v = __uint_as_float(r[j]) + (g.bias ? g.bias[col] : 0.f);
                            if (g.relu) v = fmaxf(v, 0.f);
                            dst[j] = v;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN < 32 ? 32 : BN);
    }
}

// ---------------------------------------------------------------------------------------------
// Persistent form of the implicit GEMM (single-term precisions).  One CTA per SM walks a contiguous run of
// output rows (tile = up to 128 output positions of one output row; order image, 128-wide segment, row), and:
//   * the filter planes (taps x channel blocks x BN x 128 B) are loaded into shared memory ONCE per CTA;
//   * each INPUT row is loaded once per CTA and used for all k x k taps: a ring slot holds 128 + k - 1 pixels of
//     one input row (64 channels = one 128-byte swizzle row per pixel), the k horizontal taps are k operand
//     descriptors whose start address is moved by 128 B per tap (the 128-byte swizzle is a function of the
//     shared-memory address bits, so a view shifted by whole rows stays consistent with what TMA wrote: measured,
//     the descriptor's base-offset field stays 0), and the k vertical taps are the k most recent slots.  Moving
//     to the next output row loads one new input row: the image crosses L2 -> SM about once instead of k*k times;
//   * accumulators rotate through up to four TMEM slots so the epilogue of one tile runs under the MMAs of the
//     next ones; two groups of four epilogue warps take alternate 32-column chunks of a tile: bias (+ ReLU),
//     swizzled staging in shared memory, one TMA store per chunk (rows past wout and columns past kout are
//     clipped by the tensor map).
// The MMA warp runs its loop warp-uniformly (one elected lane issues): descriptors are built with 32-bit adds on
// their low word, ring slots and phases are tracked incrementally, no division in the loop -- a first version
// whose single thread spent ~5,000 cycles per tile on address arithmetic was bound by exactly that.
// ---------------------------------------------------------------------------------------------
struct ConvPArgs {
    int hout, wout, k, pad, cblocks, kout, wsegs;
    int tiles;        // n * wsegs * hout
    int rb;           // ring depth in row units (one unit = one input row x 64 channels)
    int row_bytes;    // bytes TMA delivers per unit = (128 + k - 1) * 128
    int row_stride;   // ring slot pitch (row_bytes rounded up to 1024)
    int ker_bytes;    // taps * cblocks * BN * 128
    const float *bias;
    int relu;
    int *err_flag;
    unsigned int idesc;
    int debug;        // development: 1 no result stores, 2 no MMAs
};
constexpr int kConvStageBytes = 128 * 128;   // one 32-column fp32 chunk of a tile
constexpr int kConvMaxRing = 16;
constexpr int kConvThreads = 352;            // warp 0 TMA, warps 1-2 MMA, warps 3..6 and 7..10 the two epilogue groups

__device__ __forceinline__ void bar_sync_named(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ bool bar_red_or_named(int id, int nthreads, bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "setp.ne.u32 q, %3, 0;\n\t"
        "bar.red.or.pred p, %1, %2, q;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(r)
        : "r"(id), "r"(nthreads), "r"(static_cast<uint32_t>(pred))
        : "memory");
    return r != 0;
}

// tcgen05.mma with both descriptors given as (low word, shared high word); executed by the whole (converged) warp,
// one elected lane issues -- the operands then stay in uniform registers
__device__ __forceinline__ void umma_f16_lohi(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                              uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\t.reg .b64 da, db;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

template <int BN>
__global__ void __launch_bounds__(kConvThreads, 1)
k_conv_persist(const __grid_constant__ CUtensorMap map_img, const __grid_constant__ CUtensorMap map_ker,
               const __grid_constant__ CUtensorMap map_out, const ConvPArgs g) {
    constexpr int NACC = (512 / BN) < 4 ? (512 / BN) : 4;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *s_ker = smem;
    uint8_t *s_ring = s_ker + g.ker_bytes;
    uint8_t *s_stage = s_ring + g.rb * g.row_stride;
    float *s_bias = reinterpret_cast<float *>(s_stage + 2 * kConvStageBytes);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(s_bias + BN);
    uint64_t *empty_bar = full_bar + kConvMaxRing;
    uint64_t *ker_bar = empty_bar + kConvMaxRing;
    uint64_t *tmem_full = ker_bar + 1;
    uint64_t *tmem_empty = tmem_full + NACC;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + NACC);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = static_cast<int>(static_cast<long long>(blockIdx.x) * g.tiles / gridDim.x);
    const int t1 = static_cast<int>(static_cast<long long>(blockIdx.x + 1) * g.tiles / gridDim.x);
    const int taps = g.k * g.k;
    const int units = g.k * g.cblocks;   // row units one tile reads

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_img);
        tma_prefetch_desc(&map_ker);
        tma_prefetch_desc(&map_out);
        for (int s = 0; s < g.rb; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(ker_bar, 1);
        for (int a = 0; a < NACC; a++) {
            mbar_init(&tmem_full[a], 1);
            mbar_init(&tmem_empty[a], 256);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(ker_bar, static_cast<uint32_t>(g.ker_bytes));
            for (int kb = 0; kb < taps * g.cblocks; kb++)
                tma_load_2d(&map_ker, ker_bar, s_ker + static_cast<size_t>(kb) * BN * 128, kb * kBlockK, 0, kEvictLast);
            int slot = 0;
            uint32_t ph = 0;
            bool ok = true;
            for (int t = t0; t < t1 && ok; t++) {
                const int ho = t % g.hout, seg = (t / g.hout) % g.wsegs, img = t / (g.hout * g.wsegs);
                const bool cont = t > t0 && ho > 0;
                for (int fi = cont ? g.k - 1 : 0; fi < g.k && ok; fi++)
                    for (int cb = 0; cb < g.cblocks; cb++) {
                        if (!mbar_wait(&empty_bar[slot], ph ^ 1)) { atomicExch(g.err_flag, 311); ok = false; break; }
                        mbar_arrive_expect_tx(&full_bar[slot], static_cast<uint32_t>(g.row_bytes));
                        tma_load_4d(&map_img, &full_bar[slot], s_ring + static_cast<size_t>(slot) * g.row_stride, cb * kBlockK,
                                    seg * kBlockM - g.pad, ho - g.pad + fi, img, kEvictFirst);
                        if (++slot == g.rb) { slot = 0; ph ^= 1; }
                    }
            }
        }
    } else if (warp == 1 || warp == 2) {
        // two MMA warps take alternate tiles (a single warp spends ~60 cycles of its own instruction stream per
        // tcgen05.mma, twice what the tensor pipe needs for a 128 x 64 x 16 step).  Warp-uniform loop, one elected lane issues.
        const int mw = warp - 1;
        bool ok = mbar_wait(ker_bar, 0);
        if (!ok && lane == 0) atomicExch(g.err_flag, 312);
        tc_fence_after();
        constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B, descriptor version 1, SWIZZLE_128B
        const uint32_t ker_lo = ((smem_u32(s_ker) >> 4) & 0x3FFFu) | (1u << 16);
        const uint32_t ring_lo = ((smem_u32(s_ring) >> 4) & 0x3FFFu) | (1u << 16);
        const uint32_t slot_step = static_cast<uint32_t>(g.row_stride) >> 4;
        constexpr uint32_t kb_step = (BN * 128) >> 4;
        const uint32_t tap_step = g.cblocks * kb_step;
        int slot = 0;           // ring slot and phase of the tile's first row unit
        uint32_t ph = 0;
        int ho = t0 % g.hout;
        int acc = 0;
        uint32_t aph = 0;
        for (int t = t0; t < t1 && ok; t++) {
            const bool last_row = ho + 1 == g.hout;
            {   // both warps wait for every row unit once, in load order (a parity wait on a slot is only meaningful while
                // the slot's previous use is known to be complete), whichever warp multiplies the tile
                const int first_new = (t > t0 && ho > 0) ? units - g.cblocks : 0;
                int us = slot;
                uint32_t up = ph;
                for (int u = 0; u < units; u++) {
                    if (u >= first_new && !mbar_wait(&full_bar[us], up)) { if (lane == 0) atomicExch(g.err_flag, 314); ok = false; break; }
                    if (++us == g.rb) { us = 0; up ^= 1; }
                }
                if (!ok) break;
            }
            if (((t - t0) & 1) == mw) {
                if (!mbar_wait(&tmem_empty[acc], aph ^ 1)) { if (lane == 0) atomicExch(g.err_flag, 313); break; }
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                uint32_t accum = 0;
                int us = slot;
                uint32_t b_lo = ker_lo;
                if (!(g.debug & 2)) {
                    for (int fi = 0; fi < g.k; fi++) {
                        for (int cb = 0; cb < g.cblocks; cb++) {
                            uint32_t aj = ring_lo + us * slot_step;
                            uint32_t bj = b_lo + cb * kb_step;
                            for (int fj = 0; fj < g.k; fj++) {
#pragma unroll
                                for (int kk = 0; kk < kBlockK / kUmmaK; kk++) {
                                    umma_f16_lohi(d_tmem, aj + 2 * kk, bj + 2 * kk, desc_hi, g.idesc, accum);
                                    accum = 1;
                                }
                                aj += 8;          // next horizontal tap: one pixel = 128 B further
                                bj += tap_step;
                            }
                            if (++us == g.rb) us = 0;
                        }
                        b_lo += g.k * tap_step;
                    }
                }
                if (lane == 0) umma_commit(&tmem_full[acc]);
                __syncwarp();
            }
            slot += last_row ? units : g.cblocks;
            if (slot >= g.rb) { slot -= g.rb; ph ^= 1; }
            ho = last_row ? 0 : ho + 1;
            if (++acc == NACC) { acc = 0; aph ^= 1; }
        }
    } else {
        const int ew = warp - 3;                  // epilogue warp 0..7
        const int grp = ew >> 2;                  // epilogue group: chunks grp, grp + 2, ...
        const int gt = (ew & 3) * 32 + lane;      // thread within the group
        const int q = warp & 3;                   // the TMEM lane quarter this warp may read
        const int r = q * 32 + lane;              // row of the tile = output position wo0 + r
        for (int cidx = ew * 32 + lane; cidx < BN; cidx += 256) s_bias[cidx] = (g.bias && cidx < g.kout) ? g.bias[cidx] : 0.f;
        bar_sync_named(3, 256);
        const int nchunks = (g.kout + 31) / 32;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        uint8_t *stage = s_stage + grp * kConvStageBytes;
        const uint32_t row_u = smem_u32(stage) + r * 128;
        const int bar_id = 1 + grp;
        int ho = t0 % g.hout, seg = (t0 / g.hout) % g.wsegs, img = t0 / (g.hout * g.wsegs);
        int acc = 0;
        uint32_t aph = 0;
        int slot = 0;
        for (int t = t0; t < t1; t++) {
            bool ok = mbar_wait(&tmem_full[acc], aph);
            if (!ok) atomicExch(g.err_flag, 315);
            ok = !bar_red_or_named(bar_id, 128, !ok);   // the whole group leaves together if any thread timed out
            tc_fence_after();
            if (!ok) break;
            // this tile and, seen earlier in order, every tile before it are complete: the row units no later tile reads go
            // back to the producer (the MMA warps cannot do it with tcgen05.commit, which only covers the issuing warp's MMAs)
            const int step = (ho + 1 == g.hout) ? units : g.cblocks;
            if (grp == 0 && gt == 0 && t + 1 < t1) {
                int rs = slot;
                for (int u = 0; u < step; u++) {
                    mbar_arrive(&empty_bar[rs]);
                    if (++rs == g.rb) rs = 0;
                }
            }
            slot += step;
            if (slot >= g.rb) slot -= g.rb;
            for (int c = grp; c < nchunks; c += 2) {
                uint32_t v[32];
                tmem_ld_32x32b_x32(lane_addr + acc * BN + c * 32, v);
                tmem_ld_wait();
                float f[32];
                const float4 *b4 = reinterpret_cast<const float4 *>(s_bias + c * 32);
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float4 bb = b4[j];
                    f[4 * j] = __uint_as_float(v[4 * j]) + bb.x;
                    f[4 * j + 1] = __uint_as_float(v[4 * j + 1]) + bb.y;
                    f[4 * j + 2] = __uint_as_float(v[4 * j + 2]) + bb.z;
                    f[4 * j + 3] = __uint_as_float(v[4 * j + 3]) + bb.w;
                }
                if (g.relu) {
#pragma unroll
                    for (int j = 0; j < 32; j++) f[j] = fmaxf(f[j], 0.f);
                }
                if (gt == 0) bulk_wait_group_read<0>();   // the group's previous store is done reading the staging buffer
                bar_sync_named(bar_id, 128);
#pragma unroll
                for (int j = 0; j < 8; j++)
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row_u + ((j ^ (r & 7)) << 4)), "f"(f[4 * j]),
                                 "f"(f[4 * j + 1]), "f"(f[4 * j + 2]), "f"(f[4 * j + 3])
                                 : "memory");
                bar_sync_named(bar_id, 128);
                if (gt == 0 && !(g.debug & 1)) {
                    fence_proxy_async_smem();   // by the issuing thread, after the barrier that ordered the group's writes before it
                    tma_store_3d(&map_out, stage, c * 32, seg * kBlockM, img * g.hout + ho);
                    bulk_commit_group();
                }
            }
            tc_fence_before();
            mbar_arrive(&tmem_empty[acc]);
            if (++acc == NACC) { acc = 0; aph ^= 1; }
            if (++ho == g.hout) { ho = 0; if (++seg == g.wsegs) { seg = 0; img++; } }
        }
        if (gt == 0) bulk_wait_group<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// NHWC bf16 image: dims {Cp, W, H, N}; box {64, 128, 1, 1}; out-of-range coordinates (the padding) read as zero
static int make_image_map(CUtensorMap *m, const __nv_bfloat16 *base, int n, int h, int w, int cp, int box_pixels = kBlockM) {
    EncodeTiledFn enc = get_encode();
    if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return NB200_ERR_CUDA; }
    cuuint64_t dims[4] = {(cuuint64_t)cp, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)cp * 2, (cuuint64_t)w * cp * 2, (cuuint64_t)h * w * cp * 2};
    cuuint32_t box[4] = {(cuuint32_t)kBlockK, (cuuint32_t)box_pixels, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<__nv_bfloat16 *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (image) failed with CUresult %d", (int)r); return NB200_ERR_CUDA; }
    return NB200_OK;
}

// fp32 result as a 3-D tensor {kout, wout, n*hout}: box {32 columns, 128 positions, 1 output row}, 128-byte swizzle
static int make_conv_out_map(CUtensorMap *m, float *out, int kout, int wout, long long rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return NB200_ERR_CUDA; }
    cuuint64_t dims[3] = {(cuuint64_t)kout, (cuuint64_t)wout, (cuuint64_t)rows};
    cuuint64_t strides[2] = {(cuuint64_t)kout * 4, (cuuint64_t)wout * kout * 4};
    cuuint32_t box[3] = {32, (cuuint32_t)kBlockM, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (conv result) failed with CUresult %d", (int)r); return NB200_ERR_CUDA; }
    return NB200_OK;
}

template <int BN>
static int launch_conv_persist_impl(Context *c, const CUtensorMap &im, const CUtensorMap &km, const CUtensorMap &om, const ConvPArgs &g,
                                    size_t smem_bytes, int grid) {
    static size_t attr_bytes_dev[16] = {};   // function attributes are per device
    size_t &attr_bytes = attr_bytes_dev[c->device & 15];
    if (attr_bytes < smem_bytes) {
        NB_CUDA(cudaFuncSetAttribute(k_conv_persist<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_bytes)));
        attr_bytes = smem_bytes;
    }
    if (c->time_ev0) cudaEventRecord(c->time_ev0, c->stream);
    k_conv_persist<BN><<<grid, kConvThreads, smem_bytes, c->stream>>>(im, km, om, g);
    if (c->time_ev1) cudaEventRecord(c->time_ev1, c->stream);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}

// returns NB200_OK and *done = true when the persistent kernel took the job
static int try_conv_persist(Context *c, const __nv_bfloat16 *img_hi, int n, int h, int w, int cp, const Planes &ker, int kout, int k, int pad,
                            const float *bias, int relu, float *out, int BN, bool *done) {
    *done = false;
    static const int mode = [] { const char *e = getenv("NB200_CONV_PERSIST"); return e ? atoi(e) : 1; }();   // 0: tile-per-CTA kernel
    const int hout = h + 2 * pad - k + 1, wout = w + 2 * pad - k + 1;
    if (!mode || k > 8 || (kout & 3) || (reinterpret_cast<uintptr_t>(out) & 15)) return NB200_OK;
    ConvPArgs g{};
    g.hout = hout; g.wout = wout; g.k = k; g.pad = pad; g.cblocks = cp / 64; g.kout = kout;
    g.wsegs = static_cast<int>(ceil_div(wout, kBlockM));
    const long long tiles = static_cast<long long>(n) * g.wsegs * hout;
    if (tiles > (1ll << 30)) return NB200_OK;
    g.tiles = static_cast<int>(tiles);
    const int box_pixels = kBlockM + k - 1;
    g.row_bytes = box_pixels * 128;
    g.row_stride = static_cast<int>(round_up(g.row_bytes, 1024));
    g.ker_bytes = k * k * g.cblocks * BN * 128;
    const size_t fixed = static_cast<size_t>(g.ker_bytes) + 2 * kConvStageBytes + BN * 4 + (2 * kConvMaxRing + 1 + 8) * 8 + 16 + 1024;
    const size_t budget = 227 * 1024;
    const int units = k * g.cblocks;
    if (fixed + static_cast<size_t>(units + g.cblocks) * g.row_stride > budget) return NB200_OK;   // filter or row window too large
    g.rb = static_cast<int>(std::min<size_t>(kConvMaxRing, (budget - fixed) / g.row_stride));
    g.rb = std::min(g.rb, units + 4 * g.cblocks);   // a window plus four rows ahead is plenty
    g.bias = bias; g.relu = relu; g.err_flag = c->err_flag;
    g.idesc = umma_idesc_bf16(kBlockM, BN);
    static const int dbg = [] { const char *e = getenv("NB200_CONV_DEBUG"); return e ? atoi(e) : 0; }();
    g.debug = dbg;
    CUtensorMap im, km, om;
    NB_TRY(make_image_map(&im, img_hi, n, h, w, cp, box_pixels));
    NB_TRY(make_plane_map(&km, ker.hi, ker.rows, ker.ldk, BN));
    NB_TRY(make_conv_out_map(&om, out, kout, wout, static_cast<long long>(n) * hout));
    const size_t smem_bytes = fixed + static_cast<size_t>(g.rb) * g.row_stride;
    const int grid = static_cast<int>(std::min<long long>(tiles, c->sm_count > 0 ? c->sm_count : 148));
    *done = true;
    if (BN == 64) return launch_conv_persist_impl<64>(c, im, km, om, g, smem_bytes, grid);
    if (BN == 128) return launch_conv_persist_impl<128>(c, im, km, om, g, smem_bytes, grid);
    return launch_conv_persist_impl<256>(c, im, km, om, g, smem_bytes, grid);
}

template <int BN, int STAGES>
static int launch_conv_impl(Context *c, const CUtensorMap &ih, const CUtensorMap &il, const CUtensorMap &kh,
                            const CUtensorMap &kl, const ConvArgs &g, int tiles) {
    using L = SmemLayout<BN, STAGES>;
    static bool attr_set_dev[16] = {};   // function attributes are per device
    bool &attr_set = attr_set_dev[c->device & 15];
    if (!attr_set) {
        NB_CUDA(cudaFuncSetAttribute(k_conv_implicit<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
        attr_set = true;
    }
    if (c->time_ev0) cudaEventRecord(c->time_ev0, c->stream);
    k_conv_implicit<BN, STAGES><<<tiles, kThreads, L::kTotal, c->stream>>>(ih, il, kh, kl, g);
    if (c->time_ev1) cudaEventRecord(c->time_ev1, c->stream);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}

// img_hi/lo: NHWC bf16 [n][h][w][cp]; ker: planes [kout][k*k*cp] with K ordered (tap, channel)
int launch_conv_implicit(Context *c, const __nv_bfloat16 *img_hi, const __nv_bfloat16 *img_lo, int n, int h, int w, int cp,
                         const Planes &ker, int kout, int k, int pad, int nterms, const float *bias, int relu, float *out) {
    const int hout = h + 2 * pad - k + 1, wout = w + 2 * pad - k + 1;
    if (hout <= 0 || wout <= 0 || kout > 256 || cp % 64 != 0) {
        set_error("conv2d implicit GEMM: unsupported shape (kout %d > 256 or kernel larger than image)", kout);
        return NB200_ERR_SHAPE;
    }
    ConvArgs g{};
    g.n = n; g.h = h; g.w = w; g.hout = hout; g.wout = wout; g.k = k; g.pad = pad; g.cblocks = cp / 64; g.kout = kout;
    g.wsegs = static_cast<int>(ceil_div(wout, kBlockM));
    g.nterms = nterms; g.out = out; g.bias = bias; g.relu = relu; g.err_flag = c->err_flag;
    const int BN = kout <= 64 ? 64 : (kout <= 128 ? 128 : 256);
    g.idesc = umma_idesc_bf16(kBlockM, BN);
    if (nterms == 1) {
        bool done = false;
        NB_TRY(try_conv_persist(c, img_hi, n, h, w, cp, ker, kout, k, pad, bias, relu, out, BN, &done));
        if (done) return NB200_OK;
    }
    CUtensorMap ih, il, kh, kl;
    NB_TRY(make_image_map(&ih, img_hi, n, h, w, cp));
    NB_TRY(make_plane_map(&kh, ker.hi, ker.rows, ker.ldk, BN));
    if (nterms == 3) {
        NB_TRY(make_image_map(&il, img_lo, n, h, w, cp));
        NB_TRY(make_plane_map(&kl, ker.lo, ker.rows, ker.ldk, BN));
    } else { il = ih; kl = kh; }
    const int tiles = n * hout * g.wsegs;
    if (BN == 64) return launch_conv_impl<64, 6>(c, ih, il, kh, kl, g, tiles);
    if (BN == 128) return launch_conv_impl<128, 6>(c, ih, il, kh, kl, g, tiles);
    return launch_conv_impl<256, 4>(c, ih, il, kh, kl, g, tiles);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free FFTransposeMult + FFAggMatrix product (nb200_transpose_mult_noalloc).
//
// Some callers may not allocate: a pipeline thread confined to an arena the buffer pool sized up
// front, code between stream-capture begin/end, a serving thread that must not trigger cudaMalloc's
// device-wide synchronisation.  For them the scratch the regular path grows on demand (bf16 operand
// planes + split-K partial tiles) lives in this fixed array, which is part of the library image: every
// GPU that loads the library holds it until exit.
//
// On the host the array's NAME is only a stand-in: passing it to a kernel would hand the device a host
// address.  The device address comes from cudaGetSymbolAddress, once per device with that device
// current (first call = set-up; do it before any stream capture).  Device code may name the array
// directly, as k_na_reduce does.  The array is shared by all calls and keeps the last call's contents:
// nothing in it needs to start at zero except the fault word in its last 256 bytes (cleared per call), because
// every split overwrites its whole partial tile and the conversion rewrites the K padding; calls whose GPU work
// overlaps race on it.  A kernel-internal fault (a pipeline barrier that timed out) sets the fault word and the
// reduction then fills C with NaN instead of summing stale partial tiles: the caller sees it without a query.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[NB200_NOALLOC_SCRATCH_BYTES];

// fp64 row-major -> bf16 K-major plane with the K tail zeroed (grid: rows x chunks)
__global__ void __launch_bounds__(256)
k_na_convert(const double *__restrict__ src, long long ld, int cols, __nv_bfloat16 *__restrict__ dst, int ldk) {
    const long long row = blockIdx.x;
    for (int c = blockIdx.y * 256 + threadIdx.x; c < ldk; c += gridDim.y * 256)
        dst[row * ldk + c] = __float2bfloat16_rn(c < cols ? static_cast<float>(src[row * ld + c]) : 0.f);
}

// sum the split-K partial tiles (fixed order) into the caller's fp64 output; reads the scratch by name
__global__ void __launch_bounds__(256)
k_na_reduce(unsigned long long partial_off, long long split_stride, int ws_ld, int splits, int M, int N,
            double *__restrict__ out, long long out_ld) {
    const float *ws = reinterpret_cast<const float *>(g_noalloc_scratch + partial_off);
    const int fault = *reinterpret_cast<const int *>(g_noalloc_scratch + (NB200_NOALLOC_SCRATCH_BYTES - 256));
    const long long m = blockIdx.x;
    for (int n = blockIdx.y * 256 + threadIdx.x; n < N; n += gridDim.y * 256) {
        float acc = 0.f;
        for (int s = 0; s < splits; s++) acc += ws[s * split_stride + m * ws_ld + n];
        // the GEMM skipped its stores if a barrier wait timed out: never hand back sums of stale tiles
        out[m * out_ld + n] = fault != 0 ? __longlong_as_double(0x7FF8000000000000ll) : static_cast<double>(acc);
    }
}

}  // namespace nb200

extern "C" int nb200_transpose_mult_noalloc(const double *a_dev, int64_t lda, const double *b_dev, int64_t ldb, int I,
                                            int J, int K, double *c_dev, int64_t ldc, void *cuda_stream) {
    using namespace nb200;
    if (!a_dev || !b_dev || !c_dev || I <= 0 || J <= 0 || K <= 0 || lda < K || ldb < K || ldc < J) {
        set_error("transpose_mult_noalloc: invalid arguments");
        return NB200_ERR_INVALID;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    int dev = 0;
    NB_CUDA(cudaGetDevice(&dev));
    // ---- once per device: device address of the scratch array, kernel attributes ----
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static unsigned char *base_of[kMaxDev] = {};
    static int sms_of[kMaxDev] = {};
    if (dev < 0 || dev >= kMaxDev) { set_error("transpose_mult_noalloc: device index %d not supported", dev); return NB200_ERR_INVALID; }
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!base_of[dev]) {
            unsigned char *p = nullptr;
            NB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_scratch));
            NB_CUDA(cudaFuncSetAttribute(k_block_gemm<256, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SmemLayout<256, 4>::kTotal));
            NB_CUDA(cudaFuncSetAttribute(k_block_gemm<128, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, SmemLayout<128, 6>::kTotal));
            int sms = 0;
            NB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            sms_of[dev] = sms > 0 ? sms : 148;
            base_of[dev] = p;
        }
    }
    unsigned char *base = base_of[dev];
    // ---- carve the scratch: [A plane][B plane][split-K partials][fault word] ----
    const int64_t ldk = round_up(K, kBlockK);
    const int BN = (J <= 128) ? 128 : 256;
    const int tiles_m = static_cast<int>(ceil_div(I, kBlockM)), tiles_n = static_cast<int>(ceil_div(J, BN));
    const int Mpad = tiles_m * kBlockM, Npad = tiles_n * BN;
    const int num_kb = static_cast<int>(ldk / kBlockK);
    int splits = static_cast<int>(ceil_div(2 * sms_of[dev], static_cast<int64_t>(tiles_m) * tiles_n));
    if (splits > num_kb / 8) splits = num_kb / 8;  // at least 8 K blocks per split
    if (splits < 1) splits = 1;
    const int per = static_cast<int>(ceil_div(num_kb, splits));
    splits = static_cast<int>(ceil_div(num_kb, per));
    const uint64_t a_bytes = round_up(static_cast<int64_t>(I) * ldk * 2, 1024);
    const uint64_t b_bytes = round_up(static_cast<int64_t>(J) * ldk * 2, 1024);
    const uint64_t p_bytes = static_cast<uint64_t>(splits) * Mpad * Npad * sizeof(float);
    const uint64_t need = a_bytes + b_bytes + p_bytes + 256;
    if (need > NB200_NOALLOC_SCRATCH_BYTES) {
        set_error("transpose_mult_noalloc: %d x %d x %d needs %llu bytes of scratch, the fixed array holds %u "
                  "(NB200_NOALLOC_SCRATCH_BYTES); use nb200_gemm_blocks for larger products",
                  I, J, K, static_cast<unsigned long long>(need), NB200_NOALLOC_SCRATCH_BYTES);
        return NB200_ERR_INVALID;
    }
    __nv_bfloat16 *pa = reinterpret_cast<__nv_bfloat16 *>(base);
    __nv_bfloat16 *pb = reinterpret_cast<__nv_bfloat16 *>(base + a_bytes);
    float *partial = reinterpret_cast<float *>(base + a_bytes + b_bytes);
    int *fault = reinterpret_cast<int *>(base + (NB200_NOALLOC_SCRATCH_BYTES - 256));  // fixed place: k_na_reduce reads it by name
    NB_CUDA(cudaMemsetAsync(fault, 0, sizeof(int), stream));  // the one word that must start at zero
    const unsigned chunks = static_cast<unsigned>(ceil_div(ldk, 256) < 32 ? ceil_div(ldk, 256) : 32);
    k_na_convert<<<dim3(I, chunks), 256, 0, stream>>>(a_dev, lda, K, pa, static_cast<int>(ldk));
    k_na_convert<<<dim3(J, chunks), 256, 0, stream>>>(b_dev, ldb, K, pb, static_cast<int>(ldk));
    NB_CUDA(cudaGetLastError());
    CUtensorMap ma, mb;
    NB_TRY(make_plane_map(&ma, pa, I, ldk, kBlockM));
    NB_TRY(make_plane_map(&mb, pb, J, ldk, BN));
    GemmArgs g{};
    g.M = I; g.N = J; g.tiles_n = tiles_n; g.num_kb = num_kb; g.kb_per_split = per; g.nterms = 1; g.mode = 0;
    g.ws = partial; g.ws_split_stride = static_cast<long long>(Mpad) * Npad; g.ws_ld = Npad; g.err_flag = fault;
    g.idesc = umma_idesc_bf16(kBlockM, BN);
    const dim3 grid(tiles_m * tiles_n, splits);
    if (BN == 256) k_block_gemm<256, 4><<<grid, kThreads, SmemLayout<256, 4>::kTotal, stream>>>(ma, ma, mb, mb, g);
    else k_block_gemm<128, 6><<<grid, kThreads, SmemLayout<128, 6>::kTotal, stream>>>(ma, ma, mb, mb, g);
    NB_CUDA(cudaGetLastError());
    const unsigned nchunks = static_cast<unsigned>(ceil_div(J, 256) < 32 ? ceil_div(J, 256) : 32);
    k_na_reduce<<<dim3(I, nchunks), 256, 0, stream>>>(a_bytes + b_bytes, g.ws_split_stride, Npad, splits, I, J, c_dev, ldc);
    NB_CUDA(cudaGetLastError());
    return NB200_OK;
}
