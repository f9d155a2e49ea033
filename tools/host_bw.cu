// host_bw.cu -- what bounds the host-buffer (e2e) leg on this box: PCIe D2H vs host-memory fill bandwidth.
//   nvcc -O3 -std=c++17 -o /tmp/host_bw tools/host_bw.cu -lpthread && /tmp/host_bw
// Prints: D2H GB/s (pinned), memset GB/s for T threads into pinned memory (plain and non-temporal stores),
// and both running at once (do they share a bottleneck?).
#include <cuda_runtime.h>
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static void fill_nt(char *p, size_t n)
{
    __m128i z = _mm_setzero_si128();
    for (size_t i = 0; i + 16 <= n; i += 16) _mm_stream_si128((__m128i *)(p + i), z);
    _mm_sfence();
}

static double run_fill(char *buf, size_t bytes, int threads, bool nt, int reps)
{
    double t0 = now();
    for (int r = 0; r < reps; r++) {
        std::vector<std::thread> th;
        size_t chunk = (bytes / threads) & ~(size_t)63;
        for (int t = 0; t < threads; t++)
            th.emplace_back([=] {
                char *p = buf + (size_t)t * chunk;
                size_t n = (t == threads - 1) ? bytes - (size_t)t * chunk : chunk;
                if (nt) fill_nt(p, n); else memset(p, 0, n);
            });
        for (auto &x : th) x.join();
    }
    return bytes * (double)reps / (now() - t0) / 1e9;
}

int main()
{
    const size_t bytes = (size_t)1 << 30;
    char *h = nullptr, *d = nullptr;
    if (cudaHostAlloc(&h, bytes, cudaHostAllocDefault) != cudaSuccess || cudaMalloc(&d, bytes) != cudaSuccess) {
        printf("alloc failed\n");
        return 1;
    }
    memset(h, 1, bytes);
    cudaMemset(d, 0, bytes);
    cudaStream_t s;
    cudaStreamCreate(&s);
    printf("hardware threads: %u\n", std::thread::hardware_concurrency());
    for (int i = 0; i < 2; i++) {
        double t0 = now();
        for (int r = 0; r < 4; r++) cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s);
        cudaStreamSynchronize(s);
        printf("D2H pinned: %.1f GB/s\n", 4.0 * bytes / (now() - t0) / 1e9);
    }
    {
        double t0 = now();
        for (int r = 0; r < 4; r++) cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s);
        cudaStreamSynchronize(s);
        printf("H2D pinned: %.1f GB/s\n", 4.0 * bytes / (now() - t0) / 1e9);
    }
    for (int nt = 0; nt < 2; nt++)
        for (int t : {1, 2, 4, 8, 12, 16, 24, 32}) {
            if (t > (int)std::thread::hardware_concurrency()) break;
            printf("fill %s threads=%d: %.1f GB/s\n", nt ? "non-temporal" : "memset", t, run_fill(h, bytes, t, nt, 4));
        }
    // both at once: D2H into the first half while 8 threads fill the second half
    for (int t : {4, 8, 16}) {
        if (t > (int)std::thread::hardware_concurrency()) break;
        double t0 = now();
        for (int r = 0; r < 4; r++) cudaMemcpyAsync(h, d, bytes / 2, cudaMemcpyDeviceToHost, s);
        double f = run_fill(h + bytes / 2, bytes / 2, t, true, 4);
        double tf = now() - t0;
        cudaStreamSynchronize(s);
        double tc = now() - t0;
        printf("concurrent: fill(nt,%d thr) %.1f GB/s (%.3f s) while D2H %.1f GB/s (%.3f s)\n", t, f, tf, 2.0 * bytes / tc / 1e9, tc);
    }
    return 0;
}
