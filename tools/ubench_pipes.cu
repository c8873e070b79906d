// ubench_pipes.cu -- which integer instructions of the Myers column update can leave the ALU pipe?
//
// Measures, on all SMs, the lane-instruction rate of: LOP3 alone, IMAD alone, IMAD.WIDE alone, and
// LOP3 mixed with IMAD / IMAD.WIDE at several ratios (do the FMA-pipe instructions run in the shadow
// of the ALU pipe?).  Multipliers come from a kernel argument so that ptxas cannot strength-reduce
// them into shifts.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes ubench_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define LOP(a, b, c) asm volatile("lop3.b32 %0, %0, %1, %2, 0x6a;" : "+r"(a) : "r"(b), "r"(c))
#define IMAD(a, b, c) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
// 64-bit result {hi,lo} = a*b + {chi,clo}
#define IMADHI(a, b, c) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define IMADW(lo, hi, a, b) asm volatile("{ .reg .u64 t, u; mov.b64 u, {%0,%1}; mad.wide.u32 t, %2, %3, u; mov.b64 {%0,%1}, t; }" : "+r"(lo), "+r"(hi) : "r"(a), "r"(b))

// MODE: 0 LOP3 only | 1 IMAD only | 2 IMAD.WIDE only | 3 LOP3:IMAD 4:1 | 4 LOP3:IMAD.WIDE 4:1
//       5 LOP3:IMAD.WIDE 3:1 | 6 LOP3:IMAD.WIDE 2:1 | 7 LOP3:IMAD 2:1 | 8 LOP3:IMAD 1:1
//       9 IMAD.HI only | 10 LOP3:IMAD.HI 1:1 | 11 LOP3:IMAD:IMAD.HI 2:1:1
template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t *out, int iters, uint32_t two, uint32_t k1)
{
    uint32_t a[8], b[8], h[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        a[j] = threadIdx.x * 2654435761u + j * 40503u + blockIdx.x;
        b[j] = a[j] ^ 0x9e3779b9u;
        h[j] = j;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (MODE == 0) { LOP(a[j], k1, a[(j + 1) & 7]); }
                if (MODE == 1) { IMAD(b[j], two, k1); }
                if (MODE == 2) { IMADW(b[j], h[j], b[j], two); }
                if (MODE == 3) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); LOP(a[j], k1, a[(j + 3) & 7]); LOP(a[j], k1, a[(j + 4) & 7]); IMAD(b[j], two, k1); }
                if (MODE == 4) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); LOP(a[j], k1, a[(j + 3) & 7]); LOP(a[j], k1, a[(j + 4) & 7]); IMADW(b[j], h[j], b[j], two); }
                if (MODE == 5) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); LOP(a[j], k1, a[(j + 3) & 7]); IMADW(b[j], h[j], b[j], two); }
                if (MODE == 6) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); IMADW(b[j], h[j], b[j], two); }
                if (MODE == 7) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); IMAD(b[j], two, k1); }
                if (MODE == 8) { LOP(a[j], k1, a[(j + 1) & 7]); IMAD(b[j], two, k1); }
                if (MODE == 9) { IMADHI(b[j], k1, two); }
                if (MODE == 10) { LOP(a[j], k1, a[(j + 1) & 7]); IMADHI(b[j], k1, two); }
                if (MODE == 11) { LOP(a[j], k1, a[(j + 1) & 7]); LOP(a[j], k1, a[(j + 2) & 7]); IMAD(h[j], two, k1); IMADHI(b[j], k1, two); }
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) r ^= a[j] ^ b[j] ^ h[j];
    if (r == 0x12345678u) out[0] = r;
}

struct Mode { const char *name; int lop, fma; };
static const Mode modes[12] = {
    {"LOP3 only", 1, 0}, {"IMAD only", 0, 1}, {"IMAD.WIDE only", 0, 1}, {"LOP3:IMAD 4:1", 4, 1},
    {"LOP3:IMAD.WIDE 4:1", 4, 1}, {"LOP3:IMAD.WIDE 3:1", 3, 1}, {"LOP3:IMAD.WIDE 2:1", 2, 1},
    {"LOP3:IMAD 2:1", 2, 1}, {"LOP3:IMAD 1:1", 1, 1}, {"IMAD.HI only", 0, 1}, {"LOP3:IMAD.HI 1:1", 1, 1},
    {"LOP3:IMAD:IMAD.HI 2:1:1", 2, 2}};

template <int MODE>
static void run(uint32_t *out, int sms)
{
    const int iters = 4096, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(out, 64, 2, 0x55aa55aa);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<MODE><<<blocks, 256>>>(out, iters, 2, 0x55aa55aa);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double groups = (double)blocks * 256 * iters * 32; // (lop x LOP3 + fma x IMAD) groups per launch
    const double lop = groups * modes[MODE].lop / (best * 1e-3) / 1e12;
    const double fma = groups * modes[MODE].fma / (best * 1e-3) / 1e12;
    printf("%-22s %8.3f ms   LOP3 %7.3f T/s   FMA-pipe op %7.3f T/s   total %7.3f T/s\n", modes[MODE].name, best, lop, fma, lop + fma);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    uint32_t *out;
    cudaMalloc(&out, 64);
    run<0>(out, p.multiProcessorCount);
    run<1>(out, p.multiProcessorCount);
    run<2>(out, p.multiProcessorCount);
    run<3>(out, p.multiProcessorCount);
    run<4>(out, p.multiProcessorCount);
    run<5>(out, p.multiProcessorCount);
    run<6>(out, p.multiProcessorCount);
    run<7>(out, p.multiProcessorCount);
    run<8>(out, p.multiProcessorCount);
    run<9>(out, p.multiProcessorCount);
    run<10>(out, p.multiProcessorCount);
    run<11>(out, p.multiProcessorCount);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
