// host_bw.c -- host memory read bandwidth vs threads (the ceiling of the upload pipeline's host packing):
//   gcc -O3 -mavx2 -fopenmp -o tools/host_bw.bin tools/host_bw.c -L libspm_b200/lib -lspmb   (pack mode links the library)
// prints GB/s for (a) a plain AVX2 read-and-add sweep, (b) spmb_host_pack_chunks into a per-thread 1 MB buffer.
#include <immintrin.h>
#include <omp.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

void spmb_host_pack_chunks(const uint8_t *ranks, uint64_t n, uint64_t *out);

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

int main(int argc, char **argv)
{
    const size_t n = argc > 1 ? strtoull(argv[1], 0, 10) : (size_t)3200000000ull;
    uint8_t *buf = aligned_alloc(4096, n);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i += 4096) memset(buf + i, (int)(i >> 12) & 3, n - i < 4096 ? n - i : 4096);
    const int tmax = omp_get_max_threads();
    const int ts[] = {1, 2, 4, 8, 12, 15, 16, 24, 32};
    for (unsigned k = 0; k < sizeof ts / sizeof *ts; ++k) {
        const int T = ts[k];
        if (T > tmax) break;
        double best_r = 0, best_p = 0;
        for (int rep = 0; rep < 3; ++rep) {
            long long sink = 0;
            double t0 = now();
#pragma omp parallel num_threads(T) reduction(+ : sink)
            {
                const int t = omp_get_thread_num();
                const size_t lo = n / T * t / 32 * 32, hi = (t == T - 1) ? n / 32 * 32 : n / T * (t + 1) / 32 * 32;
                __m256i a = _mm256_setzero_si256(), b = a;
                for (size_t i = lo; i + 64 <= hi; i += 64) {
                    a = _mm256_add_epi64(a, _mm256_load_si256((const __m256i *)(buf + i)));
                    b = _mm256_add_epi64(b, _mm256_load_si256((const __m256i *)(buf + i + 32)));
                }
                a = _mm256_add_epi64(a, b);
                sink += _mm256_extract_epi64(a, 0) + _mm256_extract_epi64(a, 2);
            }
            double dt = now() - t0;
            if (n / dt / 1e9 > best_r) best_r = n / dt / 1e9;
            t0 = now();
#pragma omp parallel num_threads(T) reduction(+ : sink)
            {
                const int t = omp_get_thread_num();
                uint64_t *out = aligned_alloc(64, 1 << 20);
                const size_t lo = n / T * t / 32 * 32, hi = (t == T - 1) ? n / 32 * 32 : n / T * (t + 1) / 32 * 32;
                for (size_t i = lo; i < hi; i += (4u << 20)) {
                    spmb_host_pack_chunks(buf + i, hi - i < (4u << 20) ? hi - i : (4u << 20), out);
                    sink += out[7];
                }
                free(out);
            }
            dt = now() - t0;
            if (n / dt / 1e9 > best_p) best_p = n / dt / 1e9;
            if (sink == 42) puts("");
        }
        printf("threads %2d: read %.1f GB/s   pack %.1f GB/s (input bytes)\n", T, best_r, best_p);
    }
    return 0;
}
