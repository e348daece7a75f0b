// Microbenchmark 2: is the divergent gather of tools/mb/gather.cu limited by tags (one line per cycle) or by banks
// (every thread reads the SAME offset of its own line, so all 32 accesses fall on the same banks)?
// Every thread reads from its own random 128-byte line; what changes is WHERE in the line.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mb/gather2 tools/mb/gather2.cu && tools/mb/gather2 [out.json]
#include <cstdint>
#include <cstdio>

__device__ __forceinline__ void ldg256(const uint4* p, uint4& a, uint4& b) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}

// MODE 0: LDG.128 at offset 0 of the line (the old pattern)        1: LDG.128 at a random 16-byte offset
//      2: LDG.256 at offset 0                                       3: LDG.256 at a random 32-byte offset
//      4: whole line, 4 x LDG.256 in order 0,1,2,3 (the node fetch) 5: whole line, 4 x LDG.256 starting at chunk (line & 3)
//      6: whole line, 4 x LDG.256 starting at chunk (lane & 3)      7: 48-byte record at a random 48-byte multiple, 3 x LDG.128
//      8: the same record as LDG.256 + LDG.128 by parity            9: 64-byte record (random 64-byte multiple), 2 x LDG.256
//     10: LDG.128 at offset 16 * (lane & 7): perfect spread        11: 8 threads share a line (32 B each), 4 lines per warp
template <int MODE>
__global__ void k(const uint4* table, uint32_t lines, uint32_t* out, int iters) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lmask = lines - 1u;  // lines is a power of two
    for (int i = 0; i < iters; i++) {
        s = s * 1664525u + 1013904223u;
        const uint32_t line = (s >> 8) & lmask;
        const uint32_t r = s >> 28;  // 4 more random bits
        const uint4* p = table + (size_t)line * 8u;
        if (MODE == 0) { const uint4 v = __ldg(p); acc += v.x ^ v.w; }
        if (MODE == 1) { const uint4 v = __ldg(p + (r & 7u)); acc += v.x ^ v.w; }
        if (MODE == 2) { uint4 a, b; ldg256(p, a, b); acc += a.x ^ b.w; }
        if (MODE == 3) { uint4 a, b; ldg256(p + 2u * (r & 3u), a, b); acc += a.x ^ b.w; }
        if (MODE == 4 || MODE == 5 || MODE == 6) {
            const uint32_t st = MODE == 4 ? 0u : (MODE == 5 ? (line & 3u) : (lane & 3u));
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) { uint4 a, b; ldg256(p + 2u * ((j + st) & 3u), a, b); acc += a.x ^ b.w; }
        }
        if (MODE == 7 || MODE == 8) {
            const uint32_t rec = ((s >> 8) & (lmask * 2u + 1u));  // records of 48 B: 8 per 3 lines; stay inside the table
            const uint4* q = table + (size_t)rec * 3u;
            if (MODE == 7) {
#pragma unroll
                for (int j = 0; j < 3; j++) { const uint4 v = __ldg(q + j); acc += v.x ^ v.w; }
            } else {
                uint4 a, b, c;
                if (rec & 1u) { c = __ldg(q); ldg256(q + 1, a, b); }
                else { ldg256(q, a, b); c = __ldg(q + 2); }
                acc += a.x ^ b.w ^ c.x;
            }
        }
        if (MODE == 9) {
            const uint4* q = table + (size_t)((s >> 8) & (lmask * 2u + 1u)) * 4u;
            uint4 a, b, c, d; ldg256(q, a, b); ldg256(q + 2, c, d); acc += a.x ^ b.w ^ c.x ^ d.w;
        }
        if (MODE == 10) { const uint4 v = __ldg(p + (lane & 7u)); acc += v.x ^ v.w; }
        if (MODE == 11) {
            const uint32_t l8 = __shfl_sync(0xFFFFFFFFu, line, lane & ~7u);
            uint4 a, b; ldg256(table + (size_t)l8 * 8u + 2u * (lane & 3u), a, b); acc += a.x ^ b.w;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

static FILE* g_json = nullptr;
static bool g_first = true;
static double g_clock_hz = 1.965e9;

template <int MODE>
void run(const char* name, const uint4* table, uint32_t lines, uint32_t* out) {
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
    const double clk = ms * 1e-3 * g_clock_hz * 148;
    printf("%-58s lines %6u: %.3f ms  %6.1f cycles per warp-iteration per SM\n", name, lines, ms, clk / warp_iters);
    if (g_json) {
        fprintf(g_json, "%s\n  {\"pattern\": \"%s\", \"table_bytes\": %llu, \"ms\": %.4f, \"cycles_per_warp_iteration\": %.2f}", g_first ? "" : ",",
                name, (unsigned long long)lines * 128ull, ms, clk / warp_iters);
        g_first = false;
    }
}

int main(int argc, char** argv) {
    if (argc > 1) g_json = fopen(argv[1], "w");
    int khz = 0;
    if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0) == cudaSuccess && khz > 0) g_clock_hz = khz * 1e3;
    if (g_json) fprintf(g_json, "{\"what\": \"divergent gather, offset within the line varied (tools/mb/gather2.cu)\", \"sm_clock_hz\": %.4e, \"rows\": [", g_clock_hz);
    uint4* table;
    uint32_t* out;
    cudaMalloc(&table, (size_t)(1u << 20) * 128);
    cudaMemset(table, 1, (size_t)(1u << 20) * 128);
    cudaMalloc(&out, 148 * 8 * 128 * 4);
    for (uint32_t lines : {256u, 131072u}) {  // 32 KB (L1-resident; modes 7-9 touch twice that) and 16 MB (L2-resident)
        run<0>("LDG.128 @0", table, lines, out);
        run<1>("LDG.128 @random 16 B", table, lines, out);
        run<10>("LDG.128 @16*(lane&7)", table, lines, out);
        run<2>("LDG.256 @0", table, lines, out);
        run<3>("LDG.256 @random 32 B", table, lines, out);
        run<11>("LDG.256, 8 lanes per line, chunk lane&3", table, lines, out);
        run<4>("line = 4 x LDG.256 in order", table, lines, out);
        run<5>("line = 4 x LDG.256 from chunk line&3", table, lines, out);
        run<6>("line = 4 x LDG.256 from chunk lane&3", table, lines, out);
        run<7>("48 B record = 3 x LDG.128", table, lines, out);
        run<8>("48 B record = LDG.256 + LDG.128 by parity", table, lines, out);
        run<9>("64 B record = 2 x LDG.256", table, lines, out);
    }
    if (g_json) { fprintf(g_json, "\n]}\n"); fclose(g_json); }
    return 0;
}
