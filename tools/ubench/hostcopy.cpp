// Host microbenchmark for the packer's copy phase: T threads copy runs of `run` bytes (a region's reads of one plane) from
// pageable memory into a page-locked arena, with memcpy and with streaming (non-temporal) stores.
// build: g++ -O2 -std=c++17 -pthread -mavx2 tools/ubench/hostcopy.cpp -o /tmp/hostcopy -I/usr/local/cuda/include -L/usr/local/cuda/lib64 -lcudart
#include <cuda_runtime.h>
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
static void copy_nt(uint8_t *d, const uint8_t *s, size_t n) {
    size_t head = (32 - ((uintptr_t)d & 31)) & 31;
    if (head > n) head = n;
    std::memcpy(d, s, head); d += head; s += head; n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        __m256i a = _mm256_loadu_si256((const __m256i *)(s + i)), b = _mm256_loadu_si256((const __m256i *)(s + i + 32)),
                c = _mm256_loadu_si256((const __m256i *)(s + i + 64)), e = _mm256_loadu_si256((const __m256i *)(s + i + 96));
        _mm256_stream_si256((__m256i *)(d + i), a); _mm256_stream_si256((__m256i *)(d + i + 32), b);
        _mm256_stream_si256((__m256i *)(d + i + 64), c); _mm256_stream_si256((__m256i *)(d + i + 96), e);
    }
    std::memcpy(d + i, s + i, n - i);
}
int main(int argc, char **argv) {
    const size_t total = 512u << 20, run = argc > 1 ? atoi(argv[1]) : 30000;
    uint8_t *src = (uint8_t *)malloc(total), *dst = nullptr;
    if (cudaHostAlloc((void **)&dst, total, cudaHostAllocDefault) != cudaSuccess) { std::puts("cudaHostAlloc failed"); return 1; }
    std::memset(src, 1, total); std::memset(dst, 2, total);
    for (int T : {4, 8, 16, 32})
        for (int nt = 0; nt < 2; ++nt) {
            double best = 1e9;
            for (int rep = 0; rep < 4; ++rep) {
                auto t0 = std::chrono::steady_clock::now();
                std::vector<std::thread> th;
                for (int t = 0; t < T; ++t) th.emplace_back([&, t] {
                    const size_t per = total / T, b = t * per;
                    for (size_t o = 0; o + run <= per; o += run) {
                        if (nt) copy_nt(dst + b + o + 7, src + b + o, run - 7); else std::memcpy(dst + b + o + 7, src + b + o, run - 7);
                    }
                    if (nt) _mm_sfence();
                });
                for (auto &x : th) x.join();
                best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
            }
            std::printf("T=%2d run=%zu %s: %.1f GB/s\n", T, run, nt ? "streaming stores" : "memcpy          ", total / best / 1e9);
        }
    return 0;
}
