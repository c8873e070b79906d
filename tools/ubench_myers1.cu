// ubench_myers1.cu -- which instruction selection runs the one-word (A = 1) Myers column update fastest?
//
// The cut-off kernel (libspm_b200/csrc/myers_cut.cuh) spends its time in an 8-column chunk on one 32-bit
// word: per column 7 LOP3 + SHF (ALU pipe), 3 IMAD.IADD + 2 IMAD (FMA-heavy pipe), 1 LDS.  ncu shows the
// ALU and FMA-heavy pipes adding up instead of overlapping.  Variants of the same arithmetic, timed on all
// SMs at 6 and 12 warps per scheduler:
//   0  compiler's choice (shifts and add as IMAD.IADD, symbol pop SHF + IMAD, address IMAD)
//   1  add and shifts forced onto the ALU pipe (add.cc -> IADD3)
//   2  symbol address without multiplies: SHF by a constant + LOP3 (x & 0x30 | base)
//   3  2 + 1
//   4  two independent rows per thread (ILP 2), compiler's choice
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_myers1 ubench_myers1.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lds32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t add_alu(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); // carry-out form: IADD3, never IMAD.IADD
    return r;
}

template <int V>
__device__ __forceinline__ void column(uint32_t &vp, uint32_t &vn, uint32_t eq)
{
    const uint32_t x = eq | vn;
    const uint32_t t = eq & vp;
    const uint32_t s = (V == 1 || V == 3) ? add_alu(t, vp) : t + vp;
    const uint32_t d0 = (s ^ vp) | x;
    const uint32_t hn = vp & d0;
    const uint32_t hp = vn | ~(vp | d0);
    const uint32_t hps = (V == 1 || V == 3) ? add_alu(hp, hp) : hp << 1;
    const uint32_t hns = (V == 1 || V == 3) ? add_alu(hn, hn) : hn << 1;
    vn = hps & d0;
    vp = hns | ~(hps | d0);
}

template <int V>
__device__ __forceinline__ void chunk8(uint32_t &vp, uint32_t &vn, uint32_t &w, uint32_t base, uint32_t mul4, uint32_t stride)
{
    if (V == 2 || V == 3) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            // symbol j of this chunk sits at bits [31-2j, 30-2j]; byte offset of its Eq row = sym * 16
            const int sh = 26 - 2 * j;
            const uint32_t r = sh >= 0 ? (w >> sh) : (w << -sh);
            uint32_t addr;
            asm("lop3.b32 %0, %1, 0x30, %2, 0xea;" : "=r"(addr) : "r"(r), "r"(base)); // (r & 0x30) | base
            column<V>(vp, vn, lds32(addr));
        }
        w <<= 16;
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t c = w >> 30;
            w *= mul4;
            column<V>(vp, vn, lds32(c * stride + base));
        }
    }
}

template <int V>
__global__ void __launch_bounds__(128) k(uint32_t *out, const uint32_t *text, int blocks32, uint32_t mul4, uint32_t stride, int kk)
{
    __shared__ __align__(64) uint32_t s_eq[16];
    if (threadIdx.x < 16) s_eq[threadIdx.x] = 0x9e3779b9u * (threadIdx.x + 1) ^ (threadIdx.x << 7);
    __syncthreads();
    uint32_t base;
    asm volatile("mov.u32 %0, %1;" : "=r"(base) : "r"((uint32_t)__cvta_generic_to_shared(s_eq)));
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t vp = 0xffffffffu, vn = 0, vp2 = 0xffffffffu, vn2 = 0;
    int grown = 0;
    const uint2 *tp = reinterpret_cast<const uint2 *>(text) + gid;
    const uint32_t n_thr = gridDim.x * blockDim.x;
    uint2 nxt = __ldg(tp);
    for (int b = 0; b < blocks32; ++b) {
        const uint2 cur = nxt;
        if (b + 1 < blocks32) nxt = __ldg(tp + (size_t)(b + 1) * n_thr);
        uint32_t w = cur.x, w2 = cur.y ^ 0x5a5a5a5au;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (q == 2) { w = cur.y; w2 = cur.x ^ 0xa5a5a5a5u; }
            if (V == 4) {
                // two rows interleaved by the scheduler (same basic block)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t c = w >> 30, c2 = w2 >> 30;
                    w *= mul4; w2 *= mul4;
                    const uint32_t e = lds32(c * stride + base), e2 = lds32(c2 * stride + base);
                    column<0>(vp, vn, e);
                    column<0>(vp2, vn2, e2);
                }
                const bool g = (int)__popc(vp & 0xffffffu) - (int)__popc(vn) <= kk || (int)__popc(vp2 & 0xffffffu) - (int)__popc(vn2) <= kk;
                if (__any_sync(0xffffffffu, g)) ++grown;
            } else {
                chunk8<V>(vp, vn, w, base, mul4, stride);
                const bool g = (int)__popc(vp & 0xffffffu) - (int)__popc(vn) <= kk;
                if (__any_sync(0xffffffffu, g)) ++grown;
            }
        }
    }
    out[gid] = vp ^ vn ^ vp2 ^ vn2 ^ grown;
}

template <int V>
static void run(const char *name, int warps_per_sched, uint32_t *out, const uint32_t *text, int blocks32, int sms, double ghz_hint)
{
    const int blocks = sms * warps_per_sched; // 4 warps per block, 4 schedulers per SM
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<V><<<blocks, 128>>>(out, text, blocks32, 4, 16, -100);
    cudaEventRecord(e0);
    k<V><<<blocks, 128>>>(out, text, blocks32, 4, 16, -100);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double rows = (double)blocks * 128 * (V == 4 ? 2 : 1);
    const double cols = rows * blocks32 * 32;
    const double warp_cols_per_sched = cols / 32 / (sms * 4);
    printf("%-44s %2d warps/sched  %8.3f ms  %7.1f Gcol/s  %6.2f cycles per warp-column per scheduler (at %.3f GHz)\n", name,
           warps_per_sched, ms, cols / ms / 1e6, ms * 1e-3 * ghz_hint * 1e9 / warp_cols_per_sched, ghz_hint);
}

int main()
{
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz / 1e6;
    printf("%s, %d SMs, %.3f GHz nominal\n", pr.name, pr.multiProcessorCount, ghz);
    const int sms = pr.multiProcessorCount, blocks32 = 256;
    const size_t n_thr_max = (size_t)sms * 12 * 128;
    uint32_t *text, *out;
    cudaMalloc(&text, n_thr_max * 8 * blocks32);
    cudaMalloc(&out, n_thr_max * 4);
    {
        // pseudo-random symbols
        uint32_t *h = (uint32_t *)malloc(n_thr_max * 8 * blocks32);
        uint64_t x = 88172645463325252ull;
        for (size_t i = 0; i < n_thr_max * 2 * blocks32; ++i) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; h[i] = (uint32_t)(x >> 16); }
        cudaMemcpy(text, h, n_thr_max * 8 * blocks32, cudaMemcpyHostToDevice);
        free(h);
    }
    for (int wps : {6, 12}) {
        run<0>("0 compiler's choice", wps, out, text, blocks32, sms, ghz);
        run<1>("1 add+shifts on the ALU pipe", wps, out, text, blocks32, sms, ghz);
        run<2>("2 address = SHF + LOP3, no IMAD", wps, out, text, blocks32, sms, ghz);
        run<3>("3 = 2 + 1 (everything on the ALU pipe)", wps, out, text, blocks32, sms, ghz);
        run<4>("4 two rows per thread (ILP 2)", wps, out, text, blocks32, sms, ghz);
    }
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
