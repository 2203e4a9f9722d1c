// micro-benchmarks behind DESIGN.md's converter numbers: per-SM throughput of the fp64 -> fp16 conversion variants
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
template <int V>
__global__ void __launch_bounds__(256) k(const double* in, uint32_t* out, int iters, float scale) {
    double v[8];
    for (int j = 0; j < 8; j++) v[j] = in[threadIdx.x * 8 + j];
    uint32_t acc = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            uint32_t p;
            if (V == 0) {  // F2F.F32.F64 + FMUL + F2FP
                const __half2 h = __floats2half2_rn(static_cast<float>(v[j]) * scale, static_cast<float>(v[j + 1]) * scale);
                p = *reinterpret_cast<const uint32_t*>(&h);
            } else if (V == 1) {  // cvt.rn.f16.f64
                unsigned short a, b;
                asm volatile("cvt.rn.f16.f64 %0, %1;" : "=h"(a) : "d"(v[j]));
                asm volatile("cvt.rn.f16.f64 %0, %1;" : "=h"(b) : "d"(v[j + 1]));
                p = a | (uint32_t(b) << 16);
            } else {  // integer: fp64 bits -> fp32 bits (round to odd), then FMUL + F2FP
                uint32_t f[2];
                for (int q = 0; q < 2; q++) {
                    const uint32_t hi = __double2hiint(v[j + q]), lo = __double2loint(v[j + q]);
                    uint32_t t = __funnelshift_l(lo, hi, 3) - 0xC0000000u;
                    t = (t & 0x7fffffffu) | (hi & 0x80000000u) | ((lo & 0x1fffffffu) != 0);
                    const uint32_t ah = hi & 0x7fffffffu;
                    f[q] = (ah < (897u << 20)) ? (hi & 0x80000000u) : ((ah >= (1151u << 20)) ? ((hi & 0x80000000u) | 0x7f800000u) : t);
                }
                const __half2 h = __floats2half2_rn(__uint_as_float(f[0]) * scale, __uint_as_float(f[1]) * scale);
                p = *reinterpret_cast<const uint32_t*>(&h);
            }
            acc ^= p;
            v[j] = __hiloint2double(__double2hiint(v[j]) ^ (acc & 1), __double2loint(v[j]));  // keep the loop alive
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
    double* in; uint32_t* out;
    cudaMalloc(&in, 256 * 8 * 8); cudaMalloc(&out, 148 * 256 * 4);
    double h[2048]; for (int i = 0; i < 2048; i++) h[i] = 0.001 * (i + 1);
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 20000;
    for (int v = 0; v < 3; v++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(a);
            if (v == 0) k<0><<<148, 256>>>(in, out, iters, 4.f);
            if (v == 1) k<1><<<148, 256>>>(in, out, iters, 4.f);
            if (v == 2) k<2><<<148, 256>>>(in, out, iters, 4.f);
            cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep) printf("variant %d: %.3f ms, %.2f conversions / ns / SM  (%s)\n", v, ms, 256.0 * 8 * iters / (ms * 1e6),
                            cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
