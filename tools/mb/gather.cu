// Microbenchmark: what a divergent gather costs in the L1 data stage.  Every thread of a warp reads W bytes from its own
// 128-byte line of an L1-resident table (the access pattern of the per-ray BVH node fetch).  Prints warp-level loads per
// clock per SM and bytes per clock per SM for W = 4, 8, 16, 32 (one load) and for a whole line as 8 x 16 B and 4 x 32 B.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mb/gather tools/mb/gather.cu && tools/mb/gather
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ void ldg256(const uint4* p, uint4& a, uint4& b) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}

// MODE 0: LDG.32, 1: LDG.64, 2: LDG.128, 3: LDG.256, 4: line as 8 x LDG.128, 5: line as 4 x LDG.256, 6: 3 x LDG.128 (48 B),
// 7: LDG.256 + LDG.128 (48 B)
template <int MODE>
__global__ void k(const uint4* table, uint32_t lines, uint32_t* out, int iters) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int i = 0; i < iters; i++) {
        s = s * 1664525u + 1013904223u;
        const uint4* p = table + (size_t)((s >> 8) % lines) * 8u;  // a random 128-byte line per thread
        if (MODE == 0) acc += __ldg(reinterpret_cast<const uint32_t*>(p));
        if (MODE == 1) { const uint2 v = __ldg(reinterpret_cast<const uint2*>(p)); acc += v.x ^ v.y; }
        if (MODE == 2) { const uint4 v = __ldg(p); acc += v.x ^ v.y ^ v.z ^ v.w; }
        if (MODE == 3) { uint4 a, b; ldg256(p, a, b); acc += a.x ^ a.w ^ b.x ^ b.w; }
        if (MODE == 4) {
#pragma unroll
            for (int j = 0; j < 8; j++) { const uint4 v = __ldg(p + j); acc += v.x ^ v.w; }
        }
        if (MODE == 5) {
#pragma unroll
            for (int j = 0; j < 4; j++) { uint4 a, b; ldg256(p + 2 * j, a, b); acc += a.x ^ b.w; }
        }
        if (MODE == 6) {
#pragma unroll
            for (int j = 0; j < 3; j++) { const uint4 v = __ldg(p + j); acc += v.x ^ v.w; }
        }
        if (MODE == 7) { uint4 a, b; ldg256(p, a, b); const uint4 v = __ldg(p + 2); acc += a.x ^ b.w ^ v.x; }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

static FILE* g_json = nullptr;  // optional: tools/mb/gather out.json -> the measured peaks bench.py's roofline reads
static bool g_first = true;
static double g_clock_hz = 1.965e9;

template <int MODE>
void run(const char* name, const uint4* table, uint32_t lines, uint32_t* out, int loads, int bytes) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 2048, blocks = 148 * 8, threads = 128;
    k<MODE><<<blocks, threads>>>(table, lines, out, 64);
    cudaEventRecord(a);
    k<MODE><<<blocks, threads>>>(table, lines, out, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    const double warp_iters = (double)blocks * threads / 32 * iters;
    const double clk = ms * 1e-3 * g_clock_hz * 148;  // SM-cycles at the bench clock
    printf("%-28s lines %6u: %.3f ms  %.1f cycles per warp-iteration per SM, %.3f warp-loads/clk/SM, %.1f B/clk/SM\n", name, lines, ms,
           clk / warp_iters, warp_iters * loads / clk, warp_iters * 32 * bytes / clk);
    if (g_json) {
        fprintf(g_json, "%s\n  {\"pattern\": \"%s\", \"table_bytes\": %llu, \"ms\": %.4f, \"cycles_per_warp_iteration\": %.2f, "
                        "\"thread_loads_per_s\": %.4e, \"warp_iterations_per_s\": %.4e}",
                g_first ? "" : ",", name, (unsigned long long)lines * 128ull, ms, clk / warp_iters, warp_iters * 32.0 * loads / (ms * 1e-3),
                warp_iters / (ms * 1e-3));
        g_first = false;
    }
}

int main(int argc, char** argv) {
    if (argc > 1) g_json = fopen(argv[1], "w");
    {
        int khz = 0;
        if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0) == cudaSuccess && khz > 0) g_clock_hz = khz * 1e3;
    }
    if (g_json) fprintf(g_json, "{\"what\": \"divergent gather: every thread of a warp reads from its own random 128-byte line (tools/mb/gather.cu)\", "
                                "\"sm_clock_hz\": %.4e, \"sms\": 148, \"blocks\": %d, \"threads\": 128, \"rows\": [", g_clock_hz, 148 * 8);
    uint4* table;
    uint32_t* out;
    const uint32_t max_lines = 1u << 20;  // 128 MB
    cudaMalloc(&table, (size_t)max_lines * 128);
    cudaMemset(table, 1, (size_t)max_lines * 128);
    cudaMalloc(&out, 148 * 8 * 128 * 4);
    for (uint32_t lines : {512u, 131072u}) {  // 64 KB (L1-resident) and 16 MB (L2-resident, like the BVH)
        run<0>("LDG.32", table, lines, out, 1, 4);
        run<1>("LDG.64", table, lines, out, 1, 8);
        run<2>("LDG.128", table, lines, out, 1, 16);
        run<3>("LDG.256", table, lines, out, 1, 32);
        run<6>("48 B = 3 x LDG.128", table, lines, out, 3, 48);
        run<7>("48 B = LDG.256 + LDG.128", table, lines, out, 2, 48);
        run<4>("128 B = 8 x LDG.128", table, lines, out, 8, 128);
        run<5>("128 B = 4 x LDG.256", table, lines, out, 4, 128);
    }
    if (g_json) { fprintf(g_json, "\n]}\n"); fclose(g_json); }
    return 0;
}
