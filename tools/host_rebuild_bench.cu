// host_rebuild_bench.cu -- the host half of the compact read-back alone, on a realistic stream, against the plain
// streaming zero-fill of the same bytes by the same threads. Needs no GPU (pageable memory).
//   nvcc -O3 -std=c++17 -o /tmp/host_rebuild_bench tools/host_rebuild_bench.cu mv3d_b200/csrc/host_expand.cu -lpthread
//   python tools/host_rebuild_list.py /tmp/list.bin && /tmp/host_rebuild_bench /tmp/list.bin [sweeps=16]
#include <sys/mman.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include "../mv3d_b200/csrc/host_expand.h"
using namespace mv3d;
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main(int argc, char **argv)
{
    if (argc < 2) { printf("usage: %s list.bin [sweeps]\n", argv[0]); return 2; }
    FILE *f = fopen(argv[1], "rb");
    if (!f) { printf("cannot open %s\n", argv[1]); return 1; }
    unsigned hdr[3];
    if (fread(hdr, 4, 3, f) != 3) return 1;
    const unsigned n = hdr[0], tiles = hdr[1], tile_cells = hdr[2];
    std::vector<unsigned> table(2 * tiles + 1), words(n);
    if (fread(table.data(), 4, 2 * tiles + 1, f) != 2 * tiles + 1 || fread(words.data(), 4, n, f) != n) return 1;
    fclose(f);
    const size_t cells = 563200; const int planes = 27;
    const int sweeps = argc > 2 ? atoi(argv[2]) : 16;
    const size_t frame = cells * planes;
    float *dst = nullptr;
    const bool huge = getenv("HUGE") && atoi(getenv("HUGE")) != 0;  // HUGE=1: 2 MB pages (madvise) for the destination
    {
        const size_t bytes = frame * 4 * sweeps, two_mb = (size_t)2 << 20;
        char *raw = (char *)mmap(nullptr, bytes + two_mb, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (raw == MAP_FAILED) return 1;
        dst = (float *)(((uintptr_t)raw + two_mb - 1) & ~(uintptr_t)(two_mb - 1));
        printf("destination: %s pages (madvise rc %d)\n", huge ? "2 MB" : "4 KB", madvise(dst, bytes, huge ? MADV_HUGEPAGE : MADV_NOHUGEPAGE));
    }
    memset(dst, 1, frame * 4 * sweeps);
    std::vector<std::vector<unsigned>> copies(sweeps, words);  // one stream per sweep, as cold as a DMA would leave it
    const unsigned hw = std::thread::hardware_concurrency();
    for (int threads : {1, 2, 4, 8, 16, 32}) {
        if ((unsigned)threads > hw) break;
        HostPool pool(threads);
        double best_z = 1e9, best_s = 1e9, best_l = 1e9;
        for (int rep = 0; rep < 3; rep++) {
            double t0 = now();
            for (int s = 0; s < sweeps; s++) zero_fill(pool, dst + (size_t)s * frame, frame);
            best_z = std::min(best_z, (now() - t0) / sweeps);
            t0 = now();
            for (int s = 0; s < sweeps; s++) {
                zero_fill(pool, dst + (size_t)s * frame, frame);
                scatter_stream(pool, dst + (size_t)s * frame, cells, planes, copies[s].data(), n, table.data(), (int)tiles, (int)tile_cells, nullptr);
            }
            best_s = std::min(best_s, (now() - t0) / sweeps);
            t0 = now();
            for (int s = 0; s < sweeps; s++)
                expand_sorted(pool, dst + (size_t)s * frame, cells, planes, copies[s].data(), n, table.data(), (int)tiles, (int)tile_cells, nullptr);
            best_l = std::min(best_l, (now() - t0) / sweeps);
        }
        printf("threads=%d zero_fill %.3f ms/sweep %.1f GB/s | fill+scatter %.3f ms %.1f GB/s | expand_sorted %.3f ms %.1f GB/s (%.2f x fill)\n",
               threads, best_z * 1e3, frame * 4 / best_z / 1e9, best_s * 1e3, frame * 4 / best_s / 1e9, best_l * 1e3,
               frame * 4 / best_l / 1e9, best_l / best_z);
    }
    return 0;
}
