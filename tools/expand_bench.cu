// expand_bench.cu -- times the host half of the compact read-back alone (no GPU work), on pinned memory.
//   nvcc -O3 -std=c++17 -o tools/expand_bench tools/expand_bench.cu mv3d_b200/csrc/host_expand.cu -lpthread
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "../mv3d_b200/csrc/host_expand.h"
using namespace mv3d;
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char **argv)
{
    const size_t cells = 563200; const int planes = 27; const int sweeps = 32;
    float *dst = nullptr;
    if (cudaHostAlloc(&dst, cells * planes * 4 * sweeps, cudaHostAllocDefault) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    const unsigned n = 25600;
    std::vector<unsigned> idx(n), vals((size_t)n * planes, 0x3f800000u);
    std::mt19937 rng(1);
    // clustered like a sweep: runs of touched cells
    for (unsigned i = 0; i < n; i++) idx[i] = (unsigned)((rng() % (cells / 4)) * 4 + (i & 3)) % cells;
    for (int threads : {4, 8, 12, 16}) {
        HostPool pool(threads);
        for (unsigned nn : {0u, n}) {
            for (int rep = 0; rep < 2; rep++) {
                double t0 = now();
                for (int s = 0; s < sweeps; s++) expand_compact(pool, dst + (size_t)s * cells * planes, cells, planes, idx.data(), vals.data(), nn, nullptr);
                double dt = (now() - t0) / sweeps;
                if (rep) printf("threads=%d entries=%u: %.3f ms/sweep  %.1f GB/s\n", threads, nn, dt * 1e3, cells * planes * 4 / dt / 1e9);
            }
        }
    }
    return 0;
}
