#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
// throughput microbenchmark: per-thread chains of independent conversions
template <int MODE>
__global__ void k(const uint32_t* in, float* out, int iters) {
    uint32_t w0 = in[threadIdx.x], w1 = in[threadIdx.x + 32], w2 = in[threadIdx.x + 64], w3 = in[threadIdx.x + 96];
    float acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (MODE == 0) {  // I2F.U8 with byte selector
                acc0 += (float)((w0 >> 8) & 0xFF); acc1 += (float)((w1 >> 16) & 0xFF); acc2 += (float)(w2 & 0xFF); acc3 += (float)(w3 >> 24);
            } else if (MODE == 1) {  // PRMT + FADD magic
                acc0 += __uint_as_float(__byte_perm(w0, 0x4B000000u, 0x7441)) - 8388608.0f;
                acc1 += __uint_as_float(__byte_perm(w1, 0x4B000000u, 0x7442)) - 8388608.0f;
                acc2 += __uint_as_float(__byte_perm(w2, 0x4B000000u, 0x7440)) - 8388608.0f;
                acc3 += __uint_as_float(__byte_perm(w3, 0x4B000000u, 0x7443)) - 8388608.0f;
            } else if (MODE == 2) {  // half -> float
                __half2 h0 = *reinterpret_cast<__half2*>(&w0), h1 = *reinterpret_cast<__half2*>(&w1);
                __half2 h2 = *reinterpret_cast<__half2*>(&w2), h3 = *reinterpret_cast<__half2*>(&w3);
                acc0 += __low2float(h0); acc1 += __high2float(h1); acc2 += __low2float(h2); acc3 += __high2float(h3);
            } else {  // baseline: just the adds
                acc0 += __uint_as_float(w0); acc1 += __uint_as_float(w1); acc2 += __uint_as_float(w2); acc3 += __uint_as_float(w3);
            }
            w0 += 0x01010101u; w1 += 0x01010101u; w2 ^= 0x00010001u; w3 += 0x01000001u;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + acc2 + acc3;
}
template <int MODE> void run(const char* name, uint32_t* in, float* out) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 4096, blocks = 148 * 8, threads = 256;
    k<MODE><<<blocks, threads>>>(in, out, 16);
    cudaEventRecord(a); k<MODE><<<blocks, threads>>>(in, out, iters); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double ops = (double)blocks * threads * iters * 8 * 4;
    printf("%-12s %.3f ms  %.1f Gconv/s  = %.1f lanes/clk/SM @1.9GHz\n", name, ms, ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.9);
}
int main() {
    uint32_t* in; float* out; cudaMalloc(&in, 4096); cudaMemset(in, 1, 4096); cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<3>("baseline", in, out); run<0>("I2F.U8", in, out); run<1>("PRMT+FADD", in, out); run<2>("HADD2.F32", in, out);
    return 0;
}
