// nb200_l1stream.cu -- layer 1 of ff::inference_unit as ONE kernel: FFTransposeMult + FFAggMatrix over the fp64
// `inputs` blocks as they lie in HBM, converted to fp16 on the way into the tensor cores.
//
// Replaces, for the first join of FF/source/SimpleFF.cc:331-364: the per-pair Eigen GEMM of
// FF/headers/FFTransposeMult.h:80-92 and the keyed sum of queryExecution/headers/CombinedShuffleSink.h:112-144 /
// AggregationProcessor.cc:131-156 (the reference hash-partitions the pair products by blockColIndex,
// FFMatrixBlock.h:151-156, i.e. it splits K across workers and merges; the K split below is the same idea inside
// one GPU).
//
// Why a second fused kernel: the first one (k_l1_fused, nb200_gemm.cu) passes the fp16 copy of X through a ring in
// global memory, so every X element crosses the L2 -> SM fabric three times (fp64 once, fp16 once per N half) and the
// kernel sits on that fabric's limit (~6.4 KB / clock, profiles/r01_ncu_final_summary.txt).  Here a CLUSTER owns a
// 256*MT x 1024 output tile: every X element is read from HBM once, converted once, and handed to the CTA that
// needs the same rows for the other half of the hidden layer through distributed shared memory; with MT = 2 the
// W1 tile is multicast to the two M tiles of the cluster (W1's fp16 plane lives in this GPU's memory).  Fabric bytes
// per launch on the Amazon shape: 47.8 GB (X) + 24 GB (W1, MT = 2) against 117 GB before.
//
// Cluster of 4*MT CTAs, rank = 4*mt + 2*nh + c:
//   pair (mt, nh) = CTAs {c = 0, 1} issues tcgen05.mma.cta_group::2, M = 256 rows of X, N = 2 x 256 of hidden half nh
//   CTA (mt, nh, c) holds A rows [128c, 128c + 128) of M tile mt.  It converts rows [64nh, 64nh + 64) of those itself
//   (8 converter warps: LDG.256 fp64 -> fp16 -> 128B-swizzled st.shared) and receives the other 64 from
//   CTA (mt, 1 - nh, c) as one 8 KB cp.async.bulk smem -> DSMEM copy per K block.
//   W1: CTA (mt, nh, c) owns rows nh*512 + c*256 + {0..255}; with MT = 2 it fetches the 128-row chunk `mt` and
//   multicasts it to CTA (1 - mt, nh, c).
//
// Work is split over clusters as (M tile, K range) segments laid out in waves (host: plan_l1_stream) so that all
// clusters of a wave sweep the same K ranges together -- W1 comes out of L2, not HBM -- and no SM idles when the
// tile count is not a multiple of the cluster count (row shards of 5,000 / 2,500 / 1,250 rows keep every cluster
// busy).  A segment's accumulator (TMEM, fp32) is written to workspace slot `slot`; k_l1s_reduce sums the slots of
// a tile in index order (deterministic) and applies 2^-(ex+ev), bias and ReLU (FFReluBiasSum.h:63-71).
//
// Every wait is bounded: a protocol error becomes err_flag != 0 (NB200_ERR_KERNEL), never a hung GPU.
#include "nb200_internal.h"
#include "nb200_ptx.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace nb200 {
namespace {

constexpr int kBK = 64;             // K block: 64 fp16 = one 128-byte swizzle row
constexpr int kUK = 16;             // K of one tcgen05.mma kind::f16
// Shared-memory slots: SA for X tiles (chain: convert, DSMEM hand-off, relay), SB for W1 tiles (TMA out of L2: ~2 us from
// slot free to tile landed).  4 + 4 slots = 192 KB and NOT more: what the CTA leaves of the SM's 228 KB is the L1, and the L1
// is where the converters' global loads land.  Beyond a 196 KB carve-out nothing is left, and a CTA per SM then streams
// this access pattern at 3.9 TB/s instead of 6.2 TB/s (tools/hbm_pattern_bench.cu, profiles/r02_hbm_pattern.txt).
constexpr int kCvtWarps = 16;
constexpr int kWarpCvt0 = 8;        // warps: 0 W1 producer, 1 MMA, 2..5 epilogue, 6 relay, 7 manager, 8.. converters
constexpr int kThreadsL1S = (kWarpCvt0 + kCvtWarps) * 32;  // 768
constexpr uint32_t kATile = 128 * kBK * 2;   // 16 KB: 128 rows of X
constexpr uint32_t kBTile = 256 * kBK * 2;   // 32 KB: two 128-row chunks of W1
constexpr uint32_t kXHalf = 64 * kBK * 2;    // 8 KB: the 64 rows one CTA converts
constexpr uint32_t smem_total(int sa, int sb) { return sa * kATile + sb * kBTile + (4 * sa + 2 * sb + 2) * 8 + 16 + 1024; }
constexpr int kWsLd = 1024;         // fp32 partial tiles: [slot][row][1024]

struct Seg { int tile, kb0, kb1, slot; };
__constant__ long long c_wait_limit = 1ll << 30;   // cycles a bounded wait may take (profilers replay kernels far slower)

struct L1SArgs {
    const double *x;        // [rows_total x cols] fp64, pitch x_ld (multiple of 4, base 32-byte aligned)
    long long x_ld;
    int rows_total, cols;
    const int *exp_x;       // fp16 scale exponent of X (device)
    float *row_sumsq;       // [rows_total], accumulated with atomics (error model of the near-tie cascade); may be null
    float *maxscaled;       // receives +inf if a converted value overflowed fp16
    float *ws;              // fp32 partial tiles
    long long ws_slot_stride;
    const Seg *sched;       // [clusters][max_seg], a cluster's list ends at the first kb0 >= kb1
    int max_seg;
    int *err_flag;
    unsigned int idesc;
    int debug;              // experiments: 2 = no MMAs, 4 = no DSMEM copy, 8 = no proxy fence
    unsigned long long *prof;   // experiments: per-CTA cycle counters [cta][16] (null = off)
};

// All shared-memory addresses in this kernel are 32-bit shared-window addresses computed once from one base: a generic
// pointer would be converted again (S2UR + ULEA) wherever the compiler chose to rematerialise it, inside the hot loops.
__device__ __forceinline__ bool try_wait_u(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool try_wait_cluster_u(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool wait_u(uint32_t bar, uint32_t parity) {  // bounded, CTA scope
    if (try_wait_u(bar, parity)) return true;
    const long long t0 = clock64();
    while (!try_wait_u(bar, parity)) {
        if (clock64() - t0 > c_wait_limit) return false;
    }
    return true;
}
__device__ __forceinline__ bool wait_cluster_u(uint32_t bar, uint32_t parity) {  // bounded, cluster-scope acquire
    if (try_wait_cluster_u(bar, parity)) return true;
    const long long t0 = clock64();
    while (!try_wait_cluster_u(bar, parity)) {
        if (clock64() - t0 > c_wait_limit) return false;
    }
    return true;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void init_u(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void arrive_u(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void arrive_expect_tx_u(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void arrive_remote_u(uint32_t bar, uint32_t cta) {   // same offset in CTA `cta` (no fence emitted)
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(mapa_u32(bar, cta)) : "memory");
}
__device__ __forceinline__ void arrive_remote_release_u(uint32_t bar, uint32_t cta) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(mapa_u32(bar, cta)) : "memory");
}
// bytes (a multiple of 16) from this CTA's shared memory to the same offset of CTA `cta`; they complete on the mbarrier at
// `bar`'s offset in THAT CTA (a barrier in a third CTA never sees them: measured)
__device__ __forceinline__ void dsmem_bulk_copy_u(uint32_t addr, uint32_t bytes, uint32_t bar, uint32_t cta) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(mapa_u32(addr, cta)), "r"(addr), "r"(bytes), "r"(mapa_u32(bar, cta))
                 : "memory");
}
__device__ __forceinline__ void tma_w_2sm_u(const CUtensorMap *m, uint32_t bar, uint32_t dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "l"(kEvictLast)
        : "memory");
}
__device__ __forceinline__ void tma_w_2sm_mc_u(const CUtensorMap *m, uint32_t bar, uint32_t dst, int c0, int c1, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster.L2::cache_hint "
        "[%0], [%1, {%3, %4}], [%2], %5, %6;"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "h"(mask), "l"(kEvictLast)
        : "memory");
}
__device__ __forceinline__ void commit_2sm_u(uint32_t bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(cta_mask) : "memory");
}
__device__ __forceinline__ void ldg256(const double *p, double (&v)[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0, %1, %2, %3}, [%4];"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3])
                 : "l"(p));
}

template <int MT, int kSA, int kSB, bool PROF>
__global__ void __cluster_dims__(4 * MT, 1, 1) __launch_bounds__(kThreadsL1S, 1)
k_l1_stream(const __grid_constant__ CUtensorMap map_w, const L1SArgs g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint32_t sbase, tid;
    asm volatile("mov.u32 %0, %1;" : "=r"(sbase) : "r"(smem_u32(smem)));   // opaque: never rematerialised
    // the thread index, taken once THROUGH A SHUFFLE: ptxas rematerialises a plain read of %tid.x as an S2R wherever it is
    // short of registers, and inside the hot loops an S2R queues behind the loads in flight (measured: a third of the
    // converters' time)
    asm volatile("{\n\t.reg .u32 t, l;\n\tmov.u32 t, %%tid.x;\n\tmov.u32 l, %%laneid;\n\tshfl.sync.idx.b32 %0, t, l, 0x1f, 0xffffffff;\n\t}" : "=r"(tid));
    const uint32_t a_ring = sbase;                      // kSA x 16 KB: X tiles (fp16, 128B-swizzled K-major)
    const uint32_t b_ring = sbase + kSA * kATile;       // kSB x 32 KB: W1 tiles
    const uint32_t a_full = sbase + kSA * kATile + kSB * kBTile;  // leader: both CTAs' A tiles of K block `it` complete
    const uint32_t a_empty = a_full + 8 * kSA;    // every CTA: both pairs that share these X rows have consumed A slot s
    const uint32_t xfull_bar = a_empty + 8 * kSA; // every CTA: the partner's 64 rows have landed (the relay tells the leader)
    const uint32_t cdone_bar = xfull_bar + 8 * kSA;  // every CTA: this CTA's converters have written their 64 rows
    const uint32_t b_full = cdone_bar + 8 * kSA;  // leader: W1 tile landed in both CTAs
    const uint32_t b_empty = b_full + 8 * kSB;    // every CTA: all pairs that share this W1 tile have consumed B slot s
    const uint32_t tmem_full_bar = b_empty + 8 * kSB;
    const uint32_t tmem_empty_bar = tmem_full_bar + 8;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + kSA * kATile + kSB * kBTile + (4 * kSA + 2 * kSB + 2) * 8);

    const int warp = static_cast<int>(tid >> 5), lane = static_cast<int>(tid & 31);
    const uint32_t crank = cluster_ctarank();
    const uint32_t c = crank & 1, nh = (crank >> 1) & 1, mt = crank >> 2;
    const uint32_t leader_rank = crank & ~1u;
    const bool leader = c == 0;
    const int cluster_id = blockIdx.x / (4 * MT);
    const Seg *my = g.sched + static_cast<long long>(cluster_id) * g.max_seg;
    // cycle accounting (experiments): slot 0 producer b_empty, 1 mma b_full, 2 mma a_full, 3 relay xfull, 4 manager cdone,
    // 5 converter warp 0 a_empty, 6 converter warp 0 total, 7 converter warp 8 a_empty, 8 converter warp 8 total, 9 mma total
    unsigned long long *const prof = PROF ? g.prof + static_cast<long long>(blockIdx.x) * 16 : nullptr;
    unsigned long long t_acc = 0, t_acc2 = 0, t_cvt_issue = 0;
#define L1S_TIMED(acc, expr) [&]() { const long long _t0 = PROF ? clock64() : 0; const bool _r = (expr); if (PROF) acc += clock64() - _t0; return _r; }()

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        for (int s = 0; s < kSA; s++) {
            // per K block: the manager of each CTA of the pair (own 64 rows in place) and its relay (partner's 64 rows)
            init_u(a_full + 8 * s, (g.debug & 4) ? 2 : 4);
            init_u(a_empty + 8 * s, 2);          // the pairs (mt, 0) and (mt, 1)
            init_u(xfull_bar + 8 * s, 1);
            init_u(cdone_bar + 8 * s, kCvtWarps / 2);   // the 8 converter warps that take this K block
        }
        for (int s = 0; s < kSB; s++) {
            init_u(b_full + 8 * s, 2);           // leader's expect_tx arrive + the peer producer
            init_u(b_empty + 8 * s, MT);         // the pairs (0, nh) .. (MT - 1, nh)
        }
        init_u(tmem_full_bar, 1);
        init_u(tmem_empty_bar, 8);           // 4 epilogue warps of each CTA of the pair
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_2sm(tmem_slot, 512);
        tmem_relinquish_2sm();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== W1 producer (one thread per CTA) =====
        if (lane == 0) {
            uint32_t it = 0;
            bool ok = true;
            for (int si = 0; si < g.max_seg && ok; si++) {
                const Seg sg = my[si];
                if (sg.kb0 >= sg.kb1) break;
                for (int kb = sg.kb0; kb < sg.kb1; kb++, it++) {
                    const uint32_t s = it % kSB, ph = (it / kSB) & 1;
                    if (!L1S_TIMED(t_acc, wait_u(b_empty + 8 * s, ph ^ 1))) { atomicExch(g.err_flag, 501); ok = false; break; }
                    const uint32_t sb = b_ring + s * kBTile;
                    if (leader) arrive_expect_tx_u(b_full + 8 * s, 2 * kBTile);
                    else arrive_remote_u(b_full + 8 * s, leader_rank);
                    const int wrow = static_cast<int>(nh) * 512 + static_cast<int>(c) * 256;
                    if (MT == 1) {
                        tma_w_2sm_u(&map_w, b_full + 8 * s, sb, kb * kBK, wrow);  // box of 256 rows
                    } else {
                        const uint16_t mask = static_cast<uint16_t>((1u << crank) | (1u << (crank ^ 4u)));
                        tma_w_2sm_mc_u(&map_w, b_full + 8 * s, sb + mt * (128 * kBK * 2), kb * kBK, wrow + static_cast<int>(mt) * 128, mask);  // box of 128 rows
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread of each pair's leader =====
        if (leader && lane == 0) {
            const long long t_mma0 = PROF ? clock64() : 0;
            uint32_t it = 0;
            bool ok = true;
            // who shares my X rows (same mt) and who shares my W1 tile (same nh)
            const uint16_t mask_a = static_cast<uint16_t>(0xFu << (4 * mt));
            const uint16_t mask_b = static_cast<uint16_t>((MT == 1 ? 0x3u : 0x33u) << (2 * nh));
            for (int si = 0; si < g.max_seg && ok; si++) {
                const Seg sg = my[si];
                if (sg.kb0 >= sg.kb1) break;
                if (si > 0) {  // the epilogue of the previous segment must have drained TMEM
                    if (!wait_cluster_u(tmem_empty_bar, (si - 1) & 1)) { atomicExch(g.err_flag, 502); ok = false; break; }
                    tc_fence_after();
                }
                for (int kb = sg.kb0; kb < sg.kb1; kb++, it++) {
                    const uint32_t sa_i = it % kSA, pha = (it / kSA) & 1, sb_i = it % kSB, phb = (it / kSB) & 1;
                    if (!L1S_TIMED(t_acc, wait_cluster_u(b_full + 8 * sb_i, phb))) { atomicExch(g.err_flag, 503); ok = false; break; }
                    if (!L1S_TIMED(t_acc2, wait_cluster_u(a_full + 8 * sa_i, pha))) { atomicExch(g.err_flag, 508); ok = false; break; }
                    tc_fence_after();
                    if (!(g.debug & 2)) {
                        const uint32_t sa = a_ring + sa_i * kATile;
                        const uint32_t sb = b_ring + sb_i * kBTile;
                        const uint64_t da = umma_desc_k_sw128(sa);
#pragma unroll
                        for (int k = 0; k < kBK / kUK; k++) {
#pragma unroll
                            for (int sub = 0; sub < 2; sub++) {
                                const uint64_t db = umma_desc_k_sw128(sb + sub * (128 * kBK * 2));
                                umma_f16_2sm(tmem_base + sub * 256, da + 2 * k, db + 2 * k, g.idesc,
                                             (kb > sg.kb0 || k > 0) ? 1u : 0u);
                            }
                        }
                    }
                    commit_2sm_u(b_empty + 8 * sb_i, mask_b);
                    commit_2sm_u(a_empty + 8 * sa_i, mask_a);
                }
                if (ok) commit_2sm_u(tmem_full_bar, static_cast<uint16_t>(0x3u << leader_rank));
            }
            if (PROF) prof[9] = clock64() - t_mma0;
        }
    } else if (warp < 6) {
        // ===== epilogue: TMEM -> fp32 partial tile in the workspace =====
        const int q = warp & 3;
        bool ok = true;
        for (int si = 0; si < g.max_seg && ok; si++) {
            const Seg sg = my[si];
            if (sg.kb0 >= sg.kb1) break;
            ok = wait_u(tmem_full_bar, si & 1);
            if (!ok && lane == 0) atomicExch(g.err_flag, 504);
            ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
            if (!ok) break;
            tc_fence_after();
            const long long row = static_cast<long long>(sg.tile) * (256 * MT) + mt * 256 + c * 128 + q * 32 + lane;
            float *dst_row = g.ws + sg.slot * g.ws_slot_stride + row * kWsLd + nh * 512;
            const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
            for (int cc = 0; cc < 16; cc++) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(lane_addr + cc * 32, r);
                tmem_ld_wait();
                const int sub = cc >> 3, j = (cc & 7) * 32;
                const int tcol = (j < 128) ? sub * 128 + j : 256 + sub * 128 + (j - 128);
                float *dst = dst_row + tcol;
#pragma unroll
                for (int jj = 0; jj < 32; jj += 4)
                    *reinterpret_cast<uint4 *>(dst + jj) = make_uint4(r[jj], r[jj + 1], r[jj + 2], r[jj + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (leader) arrive_u(tmem_empty_bar);
                else arrive_remote_release_u(tmem_empty_bar, leader_rank);
            }
        }
    } else if (warp == 6) {
        // ===== relay: the rows the partner shipped in -> the pair leader's barrier.  A bulk copy into shared memory can
        // only signal a barrier of the CTA it writes to (measured: with the barrier in a third CTA the bytes never
        // complete), and the MMA thread can only wait on its own CTA's barriers. =====
        if (lane == 0 && !(g.debug & 4)) {
            uint32_t it = 0;
            bool ok = true;
            for (int si = 0; si < g.max_seg && ok; si++) {
                const Seg sg = my[si];
                if (sg.kb0 >= sg.kb1) break;
                for (int kb = sg.kb0; kb < sg.kb1; kb++, it++) {
                    const uint32_t s = it % kSA, ph = (it / kSA) & 1;
                    arrive_expect_tx_u(xfull_bar + 8 * s, kXHalf);
                    if (!L1S_TIMED(t_acc, wait_cluster_u(xfull_bar + 8 * s, ph))) { atomicExch(g.err_flag, 506); ok = false; break; }
                    if (leader) arrive_u(a_full + 8 * s);
                    else arrive_remote_u(a_full + 8 * s, leader_rank);
                }
            }
        }
    } else if (warp == 7) {
        // ===== manager: once this CTA's converters have written their 64 rows of a K block, make them visible to the
        // async proxy (tcgen05.mma and the bulk copy read shared memory through it), ship them to the CTA that needs
        // the same rows for the other half of the hidden layer, and tell the pair leader.  The fence lives here and
        // not in the converter warps: there it would wait for the fp64 loads they keep in flight. =====
        if (lane == 0) {
            uint32_t it = 0;
            bool ok = true;
            const uint32_t partner = crank ^ 2u;
            for (int si = 0; si < g.max_seg && ok; si++) {
                const Seg sg = my[si];
                if (sg.kb0 >= sg.kb1) break;
                for (int kb = sg.kb0; kb < sg.kb1; kb++, it++) {
                    const uint32_t s = it % kSA, ph = (it / kSA) & 1;
                    if (!L1S_TIMED(t_acc, wait_u(cdone_bar + 8 * s, ph))) { atomicExch(g.err_flag, 505); ok = false; break; }
                    if (!(g.debug & 8)) fence_proxy_async_smem();
                    if (!(g.debug & 4)) dsmem_bulk_copy_u(a_ring + s * kATile + nh * kXHalf, kXHalf, xfull_bar + 8 * s, partner);
                    if (leader) arrive_u(a_full + 8 * s);
                    else arrive_remote_u(a_full + 8 * s, leader_rank);
                }
            }
        }
    } else {
        // ===== converters: rows [64nh, 64nh + 64) of this CTA's 128, fp64 -> fp16, 128B-swizzled K-major.
        // Two groups of 8 warps take alternate K blocks; a warp owns 8 rows (one 1 KB swizzle atom), a lane the elements
        // [4c, 4c + 4) and [32 + 4c, 32 + 4c + 4) of rows rl and rl + 4 (c = lane & 7): each warp-wide 256-bit load then
        // covers 256 contiguous bytes of 4 rows -- 8 full lines, half the tag lookups of a 64-bytes-per-lane split.  The
        // registers are the landing zone of the HBM stream (128 bytes per lane, 64 KB per SM), so a buffer is re-issued
        // the moment it has been converted; many light warps rather than few heavy ones because the loop is a chain of
        // dependent conversions (measured: 8 warps sat in dependency stalls). =====
        const int cw = warp - kWarpCvt0;               // 0..15
        const int grp = cw >> 3, wq = cw & 7;
        const int chunk = lane & 7;
        const int rl = lane >> 3;
        const float scale = exp2f(static_cast<float>(*g.exp_x));
        const int row_in_tile = static_cast<int>(mt * 256 + c * 128 + nh * 64) + wq * 8 + rl;
        struct Cur { int si, kb, kb1, in; const double *p0; };   // valid while kb < kb1; in: bit h = row rl + 4h exists
        auto open = [&](Cur &u) {
            u.kb = 0; u.kb1 = 0;
            if (u.si >= g.max_seg) return;
            const Seg sg = my[u.si];
            if (sg.kb0 >= sg.kb1) return;
            u.kb = sg.kb0; u.kb1 = sg.kb1;
            const int row0 = sg.tile * (256 * MT) + row_in_tile;
            u.in = (row0 < g.rows_total ? 1 : 0) | (row0 + 4 < g.rows_total ? 2 : 0);
            u.p0 = g.x + static_cast<long long>(u.in ? row0 : 0) * g.x_ld + chunk * 4;
        };
        auto advance = [&](Cur &u) {
            if (u.kb >= u.kb1) return;
            if (++u.kb >= u.kb1) { u.si++; open(u); }
        };
        auto load_row = [&](const double *p, bool in, int kb, double (&v)[8]) {   // v[0..3]: k = 4c.., v[4..7]: k = 32 + 4c..
            const long long col = static_cast<long long>(kb) * kBK;
            if (in && col + kBK <= g.cols) {
                double a[4], b[4];
                ldg256(p + col, a);
                ldg256(p + col + 32, b);
#pragma unroll
                for (int e = 0; e < 4; e++) { v[e] = a[e]; v[4 + e] = b[e]; }
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    v[e] = (in && col + chunk * 4 + e < g.cols) ? p[col + e] : 0.0;
                    v[4 + e] = (in && col + 32 + chunk * 4 + e < g.cols) ? p[col + 32 + e] : 0.0;
                }
            }
        };
        auto issue = [&](const Cur &u, double (&v)[2][8]) {
            if (u.kb >= u.kb1) return;
            load_row(u.p0, (u.in & 1) != 0, u.kb, v[0]);
            load_row(u.p0 + 4 * g.x_ld, (u.in & 2) != 0, u.kb, v[1]);
        };
        Cur lc{0, 0, 0, 0, nullptr};
        open(lc);
        if (grp == 1) advance(lc);
        Cur cc = lc;
        double buf[2][8];
        issue(lc, buf);
        advance(lc); advance(lc);
        uint32_t it = grp;
        __half2 hmax = __floats2half2_rn(0.f, 0.f);   // running max |x * 2^e| (fp16 overflow check)
        float ss[2] = {0.f, 0.f};
        bool ok = true;
        const long long t_loop0 = PROF ? clock64() : 0;
        while (cc.kb < cc.kb1 && ok) {
            const uint32_t s = it % kSA, ph = (it / kSA) & 1;
            const long long tA = PROF ? clock64() : 0;
            uint32_t pk[2][4];
#pragma unroll
            for (int h = 0; h < 2; h++) {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float f0 = static_cast<float>(buf[h][2 * e]), f1 = static_cast<float>(buf[h][2 * e + 1]);
                    ss[h] = fmaf(f0, f0, fmaf(f1, f1, ss[h]));
                    const __half2 h2 = __floats2half2_rn(f0 * scale, f1 * scale);
                    hmax = __hmax2(hmax, __habs2(h2));
                    pk[h][e] = *reinterpret_cast<const uint32_t *>(&h2);
                }
            }
            const long long tB = PROF ? clock64() : 0;
            issue(lc, buf);            // this group's next K block goes in flight before anything else happens
            advance(lc); advance(lc);      // (measured: issuing after the stores instead costs 4 %)
            const long long tC = PROF ? clock64() : 0;
            if (PROF) { t_acc2 += tB - tA; t_cvt_issue += tC - tB; }
            if (!L1S_TIMED(t_acc, wait_u(a_empty + 8 * s, ph ^ 1))) {   // every lane polls: no broadcast needed
                if (lane == 0) atomicExch(g.err_flag, 507);
                ok = false;
                break;
            }
            const uint32_t atom = a_ring + s * kATile + nh * kXHalf + wq * 1024;   // rows [8wq, 8wq + 8) of my 64
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int ra = rl + 4 * h;
                // k = 4c.. sits in 16-byte chunk c / 2 of the row, k = 32 + 4c.. in chunk 4 + c / 2 (128B swizzle: chunk ^ row)
                const uint32_t row_addr = atom + ra * 128 + ((chunk & 1) << 3);
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(row_addr + (((chunk >> 1) ^ ra) << 4)), "r"(pk[h][0]), "r"(pk[h][1]) : "memory");
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(row_addr + ((((chunk >> 1) + 4) ^ ra) << 4)), "r"(pk[h][2]), "r"(pk[h][3]) : "memory");
            }
            __syncwarp();
            if (lane == 0) arrive_u(cdone_bar + 8 * s);   // (release: the warp's stores are ordered before it)
            const int si_before = cc.si;
            advance(cc); advance(cc);
            it += 2;
            if (cc.kb >= cc.kb1 || cc.si != si_before) {
                // per-row sum of squares of the K blocks this group converted in that segment (8 chunk lanes per row)
                if (g.row_sumsq != nullptr) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        float v = ss[h];
                        v += __shfl_xor_sync(0xffffffffu, v, 1);
                        v += __shfl_xor_sync(0xffffffffu, v, 2);
                        v += __shfl_xor_sync(0xffffffffu, v, 4);
                        const int row = my[si_before].tile * (256 * MT) + row_in_tile + 4 * h;
                        if (chunk == 0 && row < g.rows_total) atomicAdd(g.row_sumsq + row, v);
                        ss[h] = 0.f;
                    }
                }
            }
        }
        if (PROF && lane == 0 && (warp == kWarpCvt0 || warp == kWarpCvt0 + 8)) prof[warp == kWarpCvt0 ? 6 : 8] = clock64() - t_loop0;
        // an fp16 infinity anywhere means the sampled scale let an input overflow: report +inf
        const uint32_t hm = *reinterpret_cast<const uint32_t *>(&hmax);
        const bool ovf = (hm & 0x7FFFu) >= 0x7C00u || ((hm >> 16) & 0x7FFFu) >= 0x7C00u;
        if (__any_sync(0xffffffffu, ovf) && lane == 0 && g.maxscaled != nullptr)
            atomicMax(reinterpret_cast<unsigned int *>(g.maxscaled), 0x7F800000u);
    }
    if (PROF && lane == 0) {
        if (warp == 0) prof[0] = t_acc;
        if (warp == 1) { prof[1] = t_acc; prof[2] = t_acc2; }
        if (warp == 6) prof[3] = t_acc;
        if (warp == 7) prof[4] = t_acc;
        if (warp == kWarpCvt0) { prof[5] = t_acc; prof[10] = t_acc2; prof[11] = t_cvt_issue; }
        if (warp == kWarpCvt0 + 8) prof[7] = t_acc;
    }
    tc_fence_before();
    cluster_sync_all();   // nobody leaves while a peer may still read its shared memory or TMEM
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2sm(tmem_base, 512);
    }
}

// out[row][col] = relu?(sum over the tile's slots * 2^-(ex+ev) + bias[col]); block = one row, 256 threads x 4 columns
__global__ void __launch_bounds__(256)
k_l1s_reduce(const float *__restrict__ ws, long long slot_stride, const int *__restrict__ tile_slots, int tile_rows, int N,
             const float *__restrict__ bias_n, int relu, const int *__restrict__ exp_u, const int *__restrict__ exp_v,
             float *__restrict__ out, long long out_ld, float *__restrict__ out_maxabs) {
    __shared__ float sh[8];
    const long long m = blockIdx.x;
    const int n0 = threadIdx.x * 4;
    const int slots = tile_slots[m / tile_rows];
    const float oscale = exp_u ? exp2f(-static_cast<float>(*exp_u + *exp_v)) : 1.f;
    float mx = 0.f;
    if (n0 < N) {
        const float *p = ws + m * kWsLd + n0;
        float4 acc = *reinterpret_cast<const float4 *>(p);
        for (int s = 1; s < slots; s++) {
            const float4 t = *reinterpret_cast<const float4 *>(p + s * slot_stride);
            acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
        float v[4] = {acc.x * oscale, acc.y * oscale, acc.z * oscale, acc.w * oscale};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (n0 + j < N) {
                if (bias_n) v[j] += bias_n[n0 + j];
                if (relu) v[j] = fmaxf(v[j], 0.f);
                out[m * out_ld + n0 + j] = v[j];
                mx = fmaxf(mx, fabsf(v[j]));
            }
        }
    }
    if (out_maxabs) {
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < 8; w++) mx = fmaxf(mx, sh[w]);
            atomicMax(reinterpret_cast<unsigned int *>(out_maxabs), __float_as_uint(mx));
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int make_w_map(CUtensorMap *m, const void *base, int64_t rows, int64_t ldk, int box_rows) {
    static EncodeTiledFn enc = nullptr;
    if (!enc) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess) {
            set_error("cuTensorMapEncodeTiled not available from the driver");
            return NB200_ERR_CUDA;
        }
        enc = reinterpret_cast<EncodeTiledFn>(p);
    }
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(ldk), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(ldk) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kBK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (l1 stream, rows=%lld ldk=%lld)", (int)r, (long long)rows,
                  (long long)ldk);
        return NB200_ERR_CUDA;
    }
    return NB200_OK;
}

template <int MT, int SA, int SB>
int max_clusters(int *out) {
    constexpr bool PROF = false;
    static int cached_dev[16] = {-1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1};   // per device: attributes are
    int dev = 0;
    cudaGetDevice(&dev);
    int &cached = cached_dev[dev & 15];
    if (cached < 0) {
        NB_CUDA(cudaFuncSetAttribute(k_l1_stream<MT, SA, SB, PROF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(SA, SB)));
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(4 * MT * 64);
        cfg.blockDim = dim3(kThreadsL1S);
        cfg.dynamicSmemBytes = smem_total(SA, SB);
        cudaLaunchAttribute at{};
        at.id = cudaLaunchAttributeClusterDimension;
        at.val.clusterDim.x = 4 * MT; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at;
        cfg.numAttrs = 1;
        int n = 0;
        NB_CUDA(cudaOccupancyMaxActiveClusters(&n, k_l1_stream<MT, SA, SB, PROF>, &cfg));
        cached = n;
    }
    *out = cached;
    return NB200_OK;
}

}  // namespace

// ---- the wave plan ---------------------------------------------------------------------------------
// `tiles` cluster tiles x `num_kb` K blocks over `G` clusters.  A wave takes T tiles and gives each S = G / T
// clusters, each a contiguous K range; cost of a wave = ceil(num_kb / S) + a fixed part (pipeline fill, TMEM flush).
// The split (T per wave) minimising the sum is found by dynamic programming over the remaining tile count.
int plan_l1_stream(int tiles, int num_kb, int G, L1StreamPlan *plan) {
    plan->segs.clear();
    plan->tile_slots.assign(tiles > 0 ? tiles : 1, 1);
    plan->clusters = 0;
    plan->max_seg = 0;
    plan->max_slots = 1;
    if (tiles <= 0 || num_kb <= 0 || G <= 0) return NB200_OK;
    const int max_split = std::max(1, num_kb / 8);  // a K range shorter than 8 blocks is all prologue
    const double fixed = 100.0;                     // K-block equivalents per wave: pipeline fill, TMEM flush, W1 re-read
    std::vector<double> best(tiles + 1, 0.0);
    std::vector<int> take(tiles + 1, 0);
    for (int R = 1; R <= tiles; R++) {
        best[R] = 1e30;
        for (int T = 1; T <= std::min(R, G); T++) {
            const int S = std::min(G / T, max_split);
            const double cost = static_cast<double>((num_kb + S - 1) / S) + fixed + best[R - T];
            if (cost < best[R] - 1e-9) { best[R] = cost; take[R] = T; }
        }
    }
    std::vector<std::vector<L1StreamSeg>> per(G);
    int t0 = 0;
    for (int R = tiles; R > 0;) {
        const int T = take[R];
        int S = std::min(G / T, max_split);
        const int per_kb = (num_kb + S - 1) / S;
        S = (num_kb + per_kb - 1) / per_kb;          // no empty K range
        for (int j = 0; j < T; j++) {
            plan->tile_slots[t0 + j] = S;
            for (int sp = 0; sp < S; sp++) {
                // split-major cluster order: the T clusters of one K range are neighbours and sweep it together
                const int gidx = sp * T + j;
                per[gidx].push_back(L1StreamSeg{t0 + j, sp * per_kb, std::min((sp + 1) * per_kb, num_kb), sp});
            }
        }
        plan->max_slots = std::max(plan->max_slots, S);
        t0 += T;
        R -= T;
    }
    for (int gi = 0; gi < G; gi++)
        if (!per[gi].empty()) {
            plan->clusters = gi + 1;
            plan->max_seg = std::max(plan->max_seg, static_cast<int>(per[gi].size()));
        }
    plan->max_seg += 1;  // terminator
    plan->segs.assign(static_cast<size_t>(plan->clusters) * plan->max_seg, L1StreamSeg{0, 0, 0, 0});
    for (int gi = 0; gi < plan->clusters; gi++)
        for (size_t j = 0; j < per[gi].size(); j++) plan->segs[static_cast<size_t>(gi) * plan->max_seg + j] = per[gi][j];
    return NB200_OK;
}

static_assert(sizeof(L1StreamSeg) == sizeof(Seg), "plan segments are handed to the kernel as they are");

int g_l1_stream_mt = 0;   // nb200_set_option("l1_stream_mt"): 0 = choose by padding, 1 / 2 = force the cluster shape

// Hidden sizes in (256, 512] leave the second hidden half of every cluster multiplying zero rows (the TMA boxes of W1 rows
// >= N are zero-filled): wasted tensor work, but the kernel is bound by streaming X and still beats the separate conversion
// + GEMM kernels there (word2vec, dim 500: 16.7 ms against 23.7 ms per 10,000 x 1,000,000 slab).  NB200_L1S_MIN_N moves
// the lower bound for measurements.
bool l1_stream_supported(const double *x, int64_t x_ld, int64_t cols, int N, const Planes &w1) {
    static const int min_n = [] { const char *e = getenv("NB200_L1S_MIN_N"); return e ? atoi(e) : 256; }();
    return N > min_n && N <= 1024 && w1.f16_valid && w1.ldk % kBK == 0 && w1.ldk >= cols && (x_ld & 3) == 0 &&
           (reinterpret_cast<uintptr_t>(x) & 31) == 0 && cols >= 8 * kBK;
}

// One part of a launch: `rows` rows starting at row `row0` of the batch, through the kernel with cluster shape MT.
template <int MT>
static int prepare_part(Context *c, L1StreamPart *pt, int row0, int rows, int num_kb, int G) {
    const int tile_rows = 256 * MT;
    const int tiles = static_cast<int>(ceil_div(rows, tile_rows));
    if (pt->plan_rows != rows || pt->plan_kb != num_kb || pt->plan_g != G || pt->plan_mt != MT) {
        NB_TRY(plan_l1_stream(tiles, num_kb, G, &pt->plan));
        const size_t sb = pt->plan.segs.size() * sizeof(L1StreamSeg), tb = pt->plan.tile_slots.size() * sizeof(int);
        if (pt->dev_bytes < sb + tb) {
            nb_free(pt->dev);
            pt->dev = nullptr;
            pt->dev_bytes = 0;
            NB_CUDA(nb_malloc(&pt->dev, sb + tb));
            pt->dev_bytes = sb + tb;
        }
        // (pageable source: the copy is staged before the call returns, so the vectors may change afterwards)
        NB_CUDA(cudaMemcpyAsync(pt->dev, pt->plan.segs.data(), sb, cudaMemcpyHostToDevice, c->stream));
        NB_CUDA(cudaMemcpyAsync(pt->dev + sb, pt->plan.tile_slots.data(), tb, cudaMemcpyHostToDevice, c->stream));
        pt->seg_dev = reinterpret_cast<const L1StreamSeg *>(pt->dev);
        pt->slots_dev = reinterpret_cast<const int *>(pt->dev + sb);
        pt->plan_rows = rows; pt->plan_kb = num_kb; pt->plan_g = G; pt->plan_mt = MT;
        static const bool verbose = getenv("NB200_L1S_VERBOSE") != nullptr;
        if (verbose)
            fprintf(stderr, "[l1 stream] MT=%d rows=%d (from row %d) tiles=%d num_kb=%d clusters=%d used=%d max_seg=%d max_slots=%d\n", MT,
                    rows, row0, tiles, num_kb, G, pt->plan.clusters, pt->plan.max_seg - 1, pt->plan.max_slots);
    }
    pt->row0 = row0; pt->rows = rows; pt->mt = MT;
    pt->tile_rows = tile_rows;
    pt->slot_stride = static_cast<long long>(tiles) * tile_rows * kWsLd;
    return NB200_OK;
}

template <int MT, int SA, int SB>
static int launch_part(Context *c, const L1StreamPart &pt, const double *x, int64_t x_ld, int cols, const int *exp_x, float *row_sumsq,
                       float *maxscaled, const Planes &w1, int debug, cudaStream_t stream) {
    L1SArgs g{};
    g.x = x + static_cast<long long>(pt.row0) * x_ld; g.x_ld = x_ld; g.rows_total = pt.rows; g.cols = cols; g.exp_x = exp_x;
    g.row_sumsq = row_sumsq ? row_sumsq + pt.row0 : nullptr; g.maxscaled = maxscaled;
    g.ws = c->ws + pt.ws_off; g.ws_slot_stride = pt.slot_stride; g.sched = reinterpret_cast<const Seg *>(pt.seg_dev); g.max_seg = pt.plan.max_seg;
    g.err_flag = c->err_flag;
    g.idesc = umma_idesc_bf16(256, 256) & ~((7u << 7) | (7u << 10));  // fp16 operands
    g.debug = debug;
    static bool limit_set = false;
    if (!limit_set) {
        limit_set = true;
        if (const char *e = getenv("NB200_WAIT_LOG2")) {
            const long long lim = 1ll << atoi(e);
            NB_CUDA(cudaMemcpyToSymbol(c_wait_limit, &lim, sizeof(lim)));
        }
    }
    static unsigned long long *prof_dev = nullptr;
    static const bool prof_on = getenv("NB200_L1S_PROF") != nullptr;
    const int nctas = pt.plan.clusters * 4 * MT;
    if (prof_on && !prof_dev) NB_CUDA(cudaMalloc(&prof_dev, 1024 * 16 * sizeof(unsigned long long)));
    if (prof_on) NB_CUDA(cudaMemsetAsync(prof_dev, 0, 1024 * 16 * sizeof(unsigned long long), stream));
    g.prof = prof_on ? prof_dev : nullptr;
    CUtensorMap mw;
    NB_TRY(make_w_map(&mw, w1.f16, w1.rows, w1.ldk, MT == 1 ? 256 : 128));
    static bool attr_set_dev[16] = {};   // function attributes are per device
    bool &attr_set = attr_set_dev[c->device & 15];
    if (!attr_set) {
        NB_CUDA(cudaFuncSetAttribute(k_l1_stream<MT, SA, SB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(SA, SB)));
        NB_CUDA(cudaFuncSetAttribute(k_l1_stream<MT, SA, SB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(SA, SB)));
        attr_set = true;
    }
    if (prof_on) k_l1_stream<MT, SA, SB, true><<<nctas, kThreadsL1S, smem_total(SA, SB), stream>>>(mw, g);
    else k_l1_stream<MT, SA, SB, false><<<nctas, kThreadsL1S, smem_total(SA, SB), stream>>>(mw, g);
    c->launches++;
    NB_CUDA(cudaGetLastError());
    if (prof_on) {   // experiments only: synchronises
        std::vector<unsigned long long> h(static_cast<size_t>(nctas) * 16);
        NB_CUDA(cudaMemcpyAsync(h.data(), prof_dev, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        NB_CUDA(cudaStreamSynchronize(stream));
        static const char *names[12] = {"producer wait b_empty", "mma wait b_full", "mma wait a_full", "relay wait xfull", "manager wait cdone",
                                        "cvt w0 wait a_empty", "cvt w0 loop total", "cvt w8 wait a_empty", "cvt w8 loop total", "mma loop total",
                                        "cvt w0 load stall+convert", "cvt w0 issue loads"};
        for (int k = 0; k < 12; k++) {
            double sum = 0, mx = 0; int n = 0;
            for (int b = 0; b < nctas; b++) { const double v = static_cast<double>(h[static_cast<size_t>(b) * 16 + k]); if (v > 0) { sum += v; n++; if (v > mx) mx = v; } }
            fprintf(stderr, "[l1 stream prof] %-24s mean %.3f Mcycles over %d CTAs, max %.3f\n", names[k], n ? sum / n / 1e6 : 0.0, n, mx / 1e6);
        }
    }
    return NB200_OK;
}

// The batch goes through one launch.  Clusters of 8 (MT = 2: W1 multicast, least fabric traffic) only fit 15 times on the
// 148 SMs (GPC sizes), which leaves 28 SMs idle.  Tried and off by default (NB200_L1S_DUAL=1): sending the tail of the batch
// to a second, concurrent launch of the 4-CTA kernel sized for those SMs -- the two launches share nothing, but the second
// one is not placed beside the 8-CTA clusters and runs after them: 16.1 ms against 11.0 ms (profiles/r02_l1stream_dual.txt).
int launch_l1_stream(Context *c, const double *x, int64_t x_ld, int rows, int cols, const int *exp_x, float *row_sumsq,
                     float *maxscaled, const Planes &w1, int N, L1StreamScratch *sc) {
    if (!l1_stream_supported(x, x_ld, cols, N, w1)) { set_error("l1 stream: shape not supported"); return NB200_ERR_STATE; }
    static const int mt_env = getenv("NB200_L1S_MT") ? atoi(getenv("NB200_L1S_MT")) : 0;
    static const int debug = getenv("NB200_L1S_DEBUG") ? atoi(getenv("NB200_L1S_DEBUG")) : 0;
    static const int g_env = getenv("NB200_L1S_CLUSTERS") ? atoi(getenv("NB200_L1S_CLUSTERS")) : 0;
    static const int dual_env = getenv("NB200_L1S_DUAL") ? atoi(getenv("NB200_L1S_DUAL")) : 0;
    int mt = (g_l1_stream_mt == 1 || g_l1_stream_mt == 2) ? g_l1_stream_mt : mt_env;
    const bool forced = mt == 1 || mt == 2;
    if (!forced) {
        // MT = 2 halves the W1 traffic but works in 512-row tiles: take it unless its padding costs more than that
        const int64_t pad1 = ceil_div(rows, 256) * 256, pad2 = ceil_div(rows, 512) * 512;
        mt = (pad2 * 10 <= pad1 * 12) ? 2 : 1;
    }
    const int num_kb = static_cast<int>(w1.ldk / kBK);
    int G2 = 0, G1 = 0;
    NB_TRY((max_clusters<2, 4, 4>(&G2)));
    NB_TRY((max_clusters<1, 4, 4>(&G1)));
    if (g_env > 0) { G2 = std::min(G2, g_env); G1 = std::min(G1, g_env); }
    if ((mt == 2 && G2 < 1) || (mt == 1 && G1 < 1)) { set_error("l1 stream: no cluster of %d CTAs fits this device", 4 * mt); return NB200_ERR_STATE; }
    const int sms = c->sm_count > 0 ? c->sm_count : 148;
    const int spare = (sms - 8 * G2) / 4;   // 4-CTA clusters that fit beside the 8-CTA ones
    sc->nparts = 1;
    if (mt == 2 && !forced && dual_env && spare >= 2 && rows >= 4096) {
        // rows in proportion to the SMs each launch gets; the first part in whole 512-row tiles
        int rows_b = static_cast<int>(static_cast<int64_t>(rows) * (4 * spare) / (8 * G2 + 4 * spare));
        int rows_a = static_cast<int>(ceil_div(rows - rows_b, 512) * 512);
        if (rows_a >= rows) rows_a = rows;
        rows_b = rows - rows_a;
        NB_TRY(prepare_part<2>(c, &sc->part[0], 0, rows_a, num_kb, G2));
        if (rows_b > 0) {
            NB_TRY(prepare_part<1>(c, &sc->part[1], rows_a, rows_b, num_kb, spare));
            sc->nparts = 2;
        }
    } else if (mt == 2) {
        NB_TRY(prepare_part<2>(c, &sc->part[0], 0, rows, num_kb, G2));
    } else {
        NB_TRY(prepare_part<1>(c, &sc->part[0], 0, rows, num_kb, G1));
    }
    long long ws_floats = 0;
    for (int i = 0; i < sc->nparts; i++) {
        sc->part[i].ws_off = ws_floats;
        ws_floats += static_cast<long long>(sc->part[i].plan.max_slots) * sc->part[i].slot_stride;
    }
    NB_TRY(ensure_ws(c, static_cast<size_t>(ws_floats) * sizeof(float)));
    if (sc->nparts == 2) NB_CUDA(cudaEventRecord(c->join_ev[3], c->stream));   // fork point: everything queued so far
    if (sc->part[0].mt == 2) NB_TRY((launch_part<2, 4, 4>(c, sc->part[0], x, x_ld, cols, exp_x, row_sumsq, maxscaled, w1, debug, c->stream)));
    else NB_TRY((launch_part<1, 4, 4>(c, sc->part[0], x, x_ld, cols, exp_x, row_sumsq, maxscaled, w1, debug, c->stream)));
    if (sc->nparts == 2) {
        cudaStream_t side = c->aux[1];
        NB_CUDA(cudaStreamWaitEvent(side, c->join_ev[3], 0));
        NB_TRY((launch_part<1, 4, 4>(c, sc->part[1], x, x_ld, cols, exp_x, row_sumsq, maxscaled, w1, debug, side)));
        NB_CUDA(cudaEventRecord(c->chunk_ev[63], side));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->chunk_ev[63], 0));            // join: the reduction needs both
    }
    return NB200_OK;
}

int launch_l1_stream_reduce(Context *c, const L1StreamScratch &sc, int rows, int N, const float *bias_n, int relu,
                            const int *exp_u, const int *exp_v, float *out, int64_t out_ld, float *out_maxabs) {
    (void)rows;
    if (out_maxabs) NB_CUDA(cudaMemsetAsync(out_maxabs, 0, sizeof(float), c->stream));
    for (int i = 0; i < sc.nparts; i++) {
        const L1StreamPart &pt = sc.part[i];
        k_l1s_reduce<<<pt.rows, 256, 0, c->stream>>>(c->ws + pt.ws_off, pt.slot_stride, pt.slots_dev, pt.tile_rows, N, bias_n, relu, exp_u,
                                                     exp_v, out + static_cast<long long>(pt.row0) * out_ld, out_ld, out_maxabs);
        c->launches++;
        NB_CUDA(cudaGetLastError());
    }
    return NB200_OK;
}

}  // namespace nb200

// Test hook (not part of the public ABI, needs no device): the wave plan as flat ints, 5 per segment
// (cluster, tile, kb0, kb1, slot); returns the segment count, or -1 when `cap` ints do not hold them.
extern "C" int nb200_debug_plan_l1(int tiles, int num_kb, int clusters, int *out, int cap, int *tile_slots) {
    nb200::L1StreamPlan plan;
    if (nb200::plan_l1_stream(tiles, num_kb, clusters, &plan) != NB200_OK) return -1;
    int n = 0;
    for (int g = 0; g < plan.clusters; g++)
        for (int j = 0; j < plan.max_seg; j++) {
            const nb200::L1StreamSeg &sg = plan.segs[static_cast<size_t>(g) * plan.max_seg + j];
            if (sg.kb0 >= sg.kb1) break;
            if (5 * (n + 1) > cap) return -1;
            out[5 * n] = g; out[5 * n + 1] = sg.tile; out[5 * n + 2] = sg.kb0; out[5 * n + 3] = sg.kb1; out[5 * n + 4] = sg.slot;
            n++;
        }
    if (tile_slots)
        for (int t = 0; t < tiles; t++) tile_slots[t] = plan.tile_slots[t];
    return n;
}
