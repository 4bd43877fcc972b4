// Integer-pipe micro-benchmark for sm_100a: sustained per-SM throughput of LOP3, POPC, IADD3, IMAD and the
// mixes the pair-count kernel issues.  These are the DENOMINATORS of the "INT issue" roofline named by
// BASELINE.json's north_star (the HBM stream cannot bind once a thread block scores >= 16 queries per
// reference word).  Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes int_pipes.cu
// Output: one JSON object on stdout.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); return 1; } } while (0)

constexpr int NCH = 8;          // independent chains per thread
constexpr int ITER = 4096;

// one "unit" = the instruction group named by MODE, applied to every chain
template <int MODE>
__device__ __forceinline__ void unit(uint32_t (&x)[NCH], uint32_t (&y)[NCH], uint32_t k0, uint32_t k1)
{
    // every result is consumed by a later instruction of the same chain, so ptxas cannot drop or merge anything
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        uint32_t p;
        if (MODE == 0) {            // LOP3
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
        } else if (MODE == 1) {     // POPC (dependent chain; value collapses, the instruction still issues)
            asm volatile("popc.b32 %0, %0;" : "+r"(x[i]));
        } else if (MODE == 2) {     // IADD (x+=y; y+=x)
            asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(y[i]));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
        } else if (MODE == 3) {     // IMAD
            asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
        } else if (MODE == 4) {     // LOP3 + POPC (1:1)
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(p), "r"(k1));
        } else if (MODE == 5) {     // LOP3 x2 + POPC
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(p), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
        } else if (MODE == 12) {    // LOP3 x3 + POPC
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(p), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
        } else if (MODE == 6) {     // LOP3 x4 + POPC
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(p), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
        } else if (MODE == 7) {     // LOP3 + IMAD (1:1)
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(x[i]), "r"(k0));
        } else if (MODE == 8) {     // LOP3 x2 + POPC + IMAD accumulate
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(p), "r"(k0));
        } else if (MODE == 9) {     // LOP3 x2 + POPC + IADD accumulate
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(p));
        } else if (MODE == 10) {    // IADD + IMAD (1:1)
            asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(y[i]));
            asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(y[i]) : "r"(k0), "r"(x[i]));
        } else if (MODE == 11) {    // POPC + IMAD (1:1)
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(y[i]));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(p), "r"(k0));
        } else if (MODE == 13) {    // LOP3 x6 + POPC x1 + IMAD x1 : the mix of the pair-count kernel's CSA variant
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(p), "r"(k0));
        } else if (MODE == 14) {    // LOP3 x5 + POPC x2 + IMAD x2 : the plain (no CSA) fast path
            uint32_t p2;
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(y[i]), "r"(k1));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
            asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(x[i]));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x1e;" : "+r"(x[i]) : "r"(k1), "r"(k0));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(k0), "r"(y[i]));
            asm volatile("popc.b32 %0, %1;" : "=r"(p2) : "r"(x[i]));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(p), "r"(k0));
            asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(y[i]) : "r"(p2), "r"(k1));
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(1024, 2) bench(uint32_t* out, unsigned long long* cycles, uint32_t k0, uint32_t k1)
{
    uint32_t x[NCH], y[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) { x[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x; y[i] = i; }
    __syncthreads();
    unsigned long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it += 4) {
        unit<MODE>(x, y, k0, k1); unit<MODE>(x, y, k0, k1); unit<MODE>(x, y, k0, k1); unit<MODE>(x, y, k0, k1);
    }
    unsigned long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < NCH; ++i) s += x[i] ^ y[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

struct Res { double ms, inst_per_clk_sm, ginst_s; };

template <int MODE>
int run(const char* name, int per_unit, int nsm, uint32_t* out, unsigned long long* cyc, bool last)
{
    const int blocks = nsm * 2, threads = 1024;
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int w = 0; w < 2; ++w) bench<MODE><<<blocks, threads>>>(out, cyc, 0x9e3779b9u, 0x7f4a7c15u);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    const int reps = 5;
    for (int w = 0; w < reps; ++w) bench<MODE><<<blocks, threads>>>(out, cyc, 0x9e3779b9u, 0x7f4a7c15u);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); ms /= reps;
    static unsigned long long h[4096];
    CK(cudaMemcpy(h, cyc, sizeof(unsigned long long) * blocks, cudaMemcpyDeviceToHost));
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += (double)h[i]; mean /= blocks;
    // two CTAs of 1024 threads are co-resident per SM: thread-instructions per SM-clock
    double thread_inst_sm = 2.0 * threads * (double)NCH * ITER * per_unit;
    double per_clk = thread_inst_sm / mean;
    double total = (double)blocks * threads * (double)NCH * ITER * per_unit;
    printf("  \"%s\": {\"inst_per_unit\": %d, \"ms\": %.4f, \"thread_inst_per_clk_per_sm\": %.2f, \"gthread_inst_per_s\": %.1f, \"eff_mhz\": %.0f}%s\n",
           name, per_unit, ms, per_clk, total / (ms * 1e6), (total / (ms * 1e-3)) / (per_clk * nsm) / 1e6, last ? "" : ",");
    return 0;
}

int main()
{
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    uint32_t* out; unsigned long long* cyc;
    CK(cudaMalloc(&out, sizeof(uint32_t) * nsm * 2 * 1024)); CK(cudaMalloc(&cyc, sizeof(unsigned long long) * 4096));
    printf("{\n  \"device\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d,\n", p.name, nsm, p.clockRate);
    if (run<0>("lop3", 1, nsm, out, cyc, false)) return 1;
    if (run<1>("popc", 1, nsm, out, cyc, false)) return 1;
    if (run<2>("iadd", 2, nsm, out, cyc, false)) return 1;
    if (run<3>("imad", 1, nsm, out, cyc, false)) return 1;
    if (run<4>("lop3_popc_1_1", 2, nsm, out, cyc, false)) return 1;
    if (run<5>("lop3_popc_2_1", 3, nsm, out, cyc, false)) return 1;
    if (run<12>("lop3_popc_3_1", 4, nsm, out, cyc, false)) return 1;
    if (run<6>("lop3_popc_4_1", 5, nsm, out, cyc, false)) return 1;
    if (run<7>("lop3_imad_1_1", 2, nsm, out, cyc, false)) return 1;
    if (run<8>("lop3_popc_imad_2_1_1", 4, nsm, out, cyc, false)) return 1;
    if (run<9>("lop3_popc_iadd_2_1_1", 4, nsm, out, cyc, false)) return 1;
    if (run<10>("iadd_imad_1_1", 2, nsm, out, cyc, false)) return 1;
    if (run<11>("popc_imad_1_1", 2, nsm, out, cyc, false)) return 1;
    if (run<13>("lop3_popc_imad_6_1_1", 8, nsm, out, cyc, false)) return 1;
    if (run<14>("lop3_popc_imad_5_2_2", 9, nsm, out, cyc, true)) return 1;
    printf("}\n");
    return 0;
}
