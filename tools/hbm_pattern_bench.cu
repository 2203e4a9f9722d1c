// How fast can one CTA per SM stream its 64 rows x 512-byte segments of a row-major fp64 matrix (pitch 4.78 MB), the
// access pattern of the layer-1 converters?  Strategies: LDG.256 into registers at several depths, the same behind an
// L2 prefetch (bulk or per line), and TMA box loads into shared-memory slots.  nvcc -arch=sm_100a, run on one B200.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ void ldg256(const double *p, double (&v)[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0, %1, %2, %3}, [%4];"
                 : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// mode 0: registers only; 1: + cp.async.bulk.prefetch.L2 of 512 B per row per block; 2: + 4 KB bulk prefetch per row every
// 8 blocks; 3: + prefetch.global.L2 per 128-byte line
template <int ROWS_PER_LANE>
__global__ void __launch_bounds__(256, 1) k_regs(const double *x, long long ld, int num_kb, int rows_per_cta, int mode, int pf,
                                                 double *out) {
    extern __shared__ uint8_t pad[];   // forces 1 CTA / SM
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = lane & 7, rl = lane >> 3;
    const int rpw = 4 * ROWS_PER_LANE;                       // rows per warp and item
    const int groups = (8 * rpw) / rows_per_cta;             // warp groups taking alternate K blocks
    const int wq = warp % (8 / groups), grp = warp / (8 / groups);
    const long long row0 = static_cast<long long>(blockIdx.x) * rows_per_cta + wq * rpw + rl;
    double acc = 0.0;
    double buf[ROWS_PER_LANE][8];
    auto issue = [&](int kb) {
        if (kb >= num_kb) return;
        if (pf > 0 && kb + pf < num_kb) {
            if (mode == 1 && chunk == 0) {
#pragma unroll
                for (int h = 0; h < ROWS_PER_LANE; h++)
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(x + (row0 + 4 * h) * ld + (long long)(kb + pf) * 64), "r"(512u) : "memory");
            } else if (mode == 2 && chunk == 0 && ((kb / groups) % 8) == 0) {
#pragma unroll
                for (int h = 0; h < ROWS_PER_LANE; h++)
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(x + (row0 + 4 * h) * ld + (long long)(kb + pf) * 64), "r"(512u * 8 * groups) : "memory");
            } else if (mode == 3 && (chunk & 1) == 0) {
#pragma unroll
                for (int h = 0; h < ROWS_PER_LANE; h++)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(x + (row0 + 4 * h) * ld + (long long)(kb + pf) * 64 + chunk * 8) : "memory");
            }
        }
#pragma unroll
        for (int h = 0; h < ROWS_PER_LANE; h++) {
            const double *p = x + (row0 + 4 * h) * ld + (long long)kb * 64 + chunk * 8;
            double a[4], b[4];
            ldg256(p, a);
            ldg256(p + 4, b);
#pragma unroll
            for (int e = 0; e < 4; e++) { buf[h][e] = a[e]; buf[h][4 + e] = b[e]; }
        }
    };
    issue(grp);
    for (int kb = grp; kb < num_kb; kb += groups) {
#pragma unroll
        for (int h = 0; h < ROWS_PER_LANE; h++)
#pragma unroll
            for (int e = 0; e < 8; e++) acc += buf[h][e];
        issue(kb + groups);
    }
    if (acc == 123.456) out[0] = acc;
}

// TMA: `slots` shared-memory slots of box_rows x 64 doubles; one producer thread, 8 consumer warps that read a slot and release it
__global__ void __launch_bounds__(288, 1) k_tma(const __grid_constant__ CUtensorMap map, int num_kb, int rows_per_cta, int box_rows,
                                                int slots, double *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + slots * box_rows * 512);
    uint64_t *empty = full + slots;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < slots; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(1));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&empty[s])), "r"(8));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int per_kb = rows_per_cta / box_rows;
    const long long items = (long long)num_kb * per_kb;
    auto wait = [&](uint64_t *bar, uint32_t parity) {
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    };
    if (warp == 8) {
        if (lane == 0) {
            for (long long it = 0; it < items; it++) {
                const int s = it % slots;
                const uint32_t ph = (it / slots) & 1;
                wait(&empty[s], ph ^ 1);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(box_rows * 512) : "memory");
                const int kb = it / per_kb, part = it % per_kb;
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             ::"r"(smem_u32(smem + s * box_rows * 512)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(smem_u32(&full[s])),
                             "r"(kb * 64), "r"(blockIdx.x * rows_per_cta + part * box_rows) : "memory");
            }
        }
    } else {
        double acc = 0.0;
        for (long long it = 0; it < items; it++) {
            const int s = it % slots;
            const uint32_t ph = (it / slots) & 1;
            wait(&full[s], ph);
            const int rows_w = box_rows / 8;   // rows of this slot per warp
            for (int r = 0; r < rows_w; r++) {
                const double2 v = *reinterpret_cast<const double2 *>(smem + s * box_rows * 512 + (warp * rows_w + r) * 512 + lane * 16);
                acc += v.x + v.y;
            }
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        if (acc == 123.456) out[0] = acc;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char **argv) {
    const int ctas = argc > 1 ? atoi(argv[1]) : 148;
    const long long ld = 597540;
    const int rows_per_cta = 64, num_kb = 9336;
    const long long rows = (long long)ctas * rows_per_cta;
    double *x, *out;
    CK(cudaMalloc(&x, rows * ld * 8));
    CK(cudaMalloc(&out, 8));
    CK(cudaMemset(x, 0, rows * ld * 8));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const double gb = (double)rows * num_kb * 512 / 1e9;
    const int pad = (argc > 2 ? atoi(argv[2]) : 160) * 1024;   // dynamic shared memory per CTA: what is left is the L1
    const bool regs_only = argc > 3;
    CK(cudaFuncSetAttribute(k_regs<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, pad));
    CK(cudaFuncSetAttribute(k_regs<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, pad));
    CK(cudaFuncSetAttribute(k_regs<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, pad));
    auto report = [&](const char *name, float ms) { printf("%-58s %7.3f ms  %6.0f GB/s\n", name, ms, gb / ms * 1e3); fflush(stdout); };
    for (int rpl : {2, 4, 8}) {
        for (int mode = 0; mode < (regs_only ? 1 : 4); mode++) {
            for (int pf : {8, 32}) {
                if (mode == 0 && pf != 8) continue;
                float best = 1e9;
                for (int rep = 0; rep < 2; rep++) {
                    cudaEventRecord(a);
                    if (rpl == 2) k_regs<2><<<ctas, 256, pad>>>(x, ld, num_kb, rows_per_cta, mode, pf, out);
                    if (rpl == 4) k_regs<4><<<ctas, 256, pad>>>(x, ld, num_kb, rows_per_cta, mode, pf, out);
                    if (rpl == 8) k_regs<8><<<ctas, 256, pad>>>(x, ld, num_kb, rows_per_cta, mode, pf, out);
                    cudaEventRecord(b);
                    CK(cudaEventSynchronize(b));
                    float ms; cudaEventElapsedTime(&ms, a, b);
                    best = ms < best ? ms : best;
                }
                char nm[128];
                snprintf(nm, sizeof nm, "regs rows/lane=%d (%d B/lane in flight) mode=%d pf=%d", rpl, rpl * 64, mode, pf);
                report(nm, best);
            }
        }
    }
    if (regs_only) return 0;
    // TMA
    void *fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(fp);
    CK(cudaFuncSetAttribute(k_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int box_rows : {64, 32, 16, 8}) {
        for (int kbytes : {64, 96, 128}) {
            const int slots = kbytes * 1024 / (box_rows * 512);
            CUtensorMap m;
            cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
            cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
            cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
            cuuint32_t es[2] = {1, 1};
            CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, x, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode failed %d (box_rows %d)\n", (int)r, box_rows); continue; }
            float best = 1e9;
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(a);
                k_tma<<<ctas, 288, 160 * 1024 + 1024>>>(m, num_kb, rows_per_cta, box_rows, slots, out);
                cudaEventRecord(b);
                CK(cudaEventSynchronize(b));
                float ms; cudaEventElapsedTime(&ms, a, b);
                best = ms < best ? ms : best;
            }
            char nm[128];
            snprintf(nm, sizeof nm, "tma box %dx64 fp64, %d slots (%d KB staging)", box_rows, slots, kbytes);
            report(nm, best);
        }
    }
    return 0;
}
