// ubench_read.cu -- what does a pure READ stream reach on this part?  (the exact kernel's ceiling)
// 1) LDG.128 grid-stride XOR-reduce, several unrolls / CTA shapes; 2) TMA bulk copies into a 3-stage ring,
// consumers only wait/arrive (no compute).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_read.bin ubench_read.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int U>
__global__ void __launch_bounds__(256) ldg_kernel(const uint4 *__restrict__ p, uint64_t n16, uint32_t *out)
{
    uint32_t acc = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n16; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = __ldg(p + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) out[0] = acc;
}

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mwait(uint32_t bar, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" :: "r"(bar), "r"(parity) : "memory");
}

template <int STAGES, int STAGE_BYTES>
__global__ void __launch_bounds__(64, 1) tma_kernel(const unsigned char *__restrict__ g, uint64_t bytes, uint32_t *out)
{
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm + STAGES * STAGE_BYTES);
    const uint64_t n_tiles = bytes / STAGE_BYTES;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(bars + s)), "r"(1));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(bars + STAGES + s)), "r"(32));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t stage = 0, phase = 0;
    if (threadIdx.x >= 32) {
        if (threadIdx.x == 32) {
            for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                mwait(s32(bars + STAGES + stage), phase ^ 1u);
                const uint32_t full = s32(bars + stage);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(full), "r"(STAGE_BYTES) : "memory");
                for (uint32_t o = 0; o < STAGE_BYTES; o += 8192)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(s32(sm + stage * STAGE_BYTES + o)), "l"(g + t * STAGE_BYTES + o), "r"(8192), "r"(full) : "memory");
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
        return;
    }
    uint32_t acc = 0;
    for (uint64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        mwait(s32(bars + stage), phase);
        acc ^= reinterpret_cast<const uint32_t *>(sm + stage * STAGE_BYTES)[threadIdx.x];
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(s32(bars + STAGES + stage)) : "memory");
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <class F>
static void timeit(const char *name, uint64_t bytes, F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 10; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    printf("%-44s %8.3f us  %8.1f GB/s  (%s)\n", name, best * 1e3, bytes / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    const int sms = pr.multiProcessorCount;
    const uint64_t bytes = 800ull << 20;
    unsigned char *d; uint32_t *out;
    cudaMalloc(&d, bytes + (1 << 20)); cudaMalloc(&out, 64);
    cudaMemset(d, 1, bytes);
    const uint64_t n16 = bytes / 16;
    printf("%s, %d SMs, %llu MB read per launch\n", pr.name, sms, (unsigned long long)(bytes >> 20));
    timeit("LDG.128 unroll 4, 8 CTAs/SM x 256", bytes, [&] { ldg_kernel<4><<<sms * 8, 256>>>((const uint4 *)d, n16, out); });
    timeit("LDG.128 unroll 8, 8 CTAs/SM x 256", bytes, [&] { ldg_kernel<8><<<sms * 8, 256>>>((const uint4 *)d, n16, out); });
    timeit("LDG.128 unroll 16, 4 CTAs/SM x 256", bytes, [&] { ldg_kernel<16><<<sms * 4, 256>>>((const uint4 *)d, n16, out); });
    timeit("LDG.128 unroll 8, 16 CTAs/SM x 256", bytes, [&] { ldg_kernel<8><<<sms * 16, 256>>>((const uint4 *)d, n16, out); });
    cudaFuncSetAttribute(tma_kernel<3, 73728>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 73728 + 64);
    timeit("TMA bulk 3 x 72 KB stages, 1 CTA/SM", bytes, [&] { tma_kernel<3, 73728><<<sms, 64, 3 * 73728 + 64>>>(d, bytes, out); });
    cudaFuncSetAttribute(tma_kernel<6, 32768>, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768 + 128);
    timeit("TMA bulk 6 x 32 KB stages, 1 CTA/SM", bytes, [&] { tma_kernel<6, 32768><<<sms, 64, 6 * 32768 + 128>>>(d, bytes, out); });
    cudaFuncSetAttribute(tma_kernel<4, 16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 16384 + 64);
    timeit("TMA bulk 4 x 16 KB stages, 3 CTA/SM", bytes, [&] { tma_kernel<4, 16384><<<sms * 3, 64, 4 * 16384 + 64>>>(d, bytes, out); });
    // the copy figure MEASURED_PEAKS.json quotes, for reference
    unsigned char *d2; cudaMalloc(&d2, bytes);
    timeit("cudaMemcpy D2D (read+write bytes)", 2 * bytes, [&] { cudaMemcpyAsync(d2, d, bytes, cudaMemcpyDeviceToDevice); });
    return 0;
}
