// anchor_kernels.h -- launchers of the empty-anchor filter kernels (internal; public ABI: include/mv3d_anchors.h).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mv3d {

constexpr int kAnchorChunk = 1024;       // anchors per compaction block
constexpr int kAnchorSumWarps = 8;       // anchors per block of the exact-sum kernel
constexpr int kAnchorMaxPlanes = 1024;   // the reference launches one thread per channel (block=(C,1,1))

struct AnchorGeometry {
    int xn, yn, planes;
    unsigned cells;            // xn * yn
    unsigned words;            // ceil(cells / 32): occupancy words per frame
    unsigned wpr;              // ceil(yn / 32): words per row of the row-aligned copy
    unsigned chunks;           // ceil(n / kAnchorChunk)
    int n;                     // anchors
    int max_lines;             // rows of a box that count (50 in the reference)
};

// Each launcher enqueues on `s` and returns the number of kernels it launched.
int launch_anchor_occupancy(const float *views, const AnchorGeometry &g, int frames, unsigned *occupancy,
                            unsigned *flags, int sm_count, cudaStream_t s);
int launch_anchor_test(const int4 *boxes, const AnchorGeometry &g, int frames, const unsigned *occupancy,
                       const unsigned *flags, unsigned *rows, unsigned *groups, unsigned char *keep,
                       unsigned *chunk_count, cudaStream_t s);
int launch_anchor_sum(const float *views, const int4 *boxes, const AnchorGeometry &g, int frames, float limit,
                      const unsigned *flags, bool always, unsigned char *keep, unsigned *chunk_count, float *totals,
                      int sm_count, cudaStream_t s);
int launch_anchor_compact(const AnchorGeometry &g, int frames, const unsigned char *keep, const unsigned *chunk_count,
                          int *keep_idx, int *count, cudaStream_t s);

} // namespace mv3d
