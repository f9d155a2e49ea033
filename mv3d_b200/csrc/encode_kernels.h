// encode_kernels.h -- launchers of the view-encoder kernels (internal; the public ABI is include/mv3d_encode.h).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mv3d {

struct ViewParams;
struct FrameOffsets;
struct PyFrontParams;

// Where the tile kernel reports cell occupancy for the empty-anchor filter (anchor_kernels.cu): one bit per cell in
// a PRE-ZEROED map of words_per_frame words per frame, and a per-frame flag "an intensity is negative, -0 or NaN".
struct OccupancySink {
    unsigned *words = nullptr;
    unsigned *flags = nullptr;
    unsigned words_per_frame = 0;
};

// Each launcher enqueues on `s` and returns the number of kernels it launched.
size_t tile_smem_bytes(const ViewParams &p);
// Opts every tile-kernel variant of this semantics into tile_smem_bytes(p) of dynamic shared memory and reports how
// many CTAs of the dense kernel an SM holds. Callers cache the answer (mv3d_abi.cu: tile_blocks_per_sm).
cudaError_t tile_kernels_configure(const ViewParams &p, int *blocks_per_sm);

int launch_classify(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                    uint2 *keys, unsigned *tile_count, unsigned *fv_winner, cudaStream_t s);
int launch_tile_scan(unsigned *tile_count, unsigned *tile_off, int tiles_per_frame, int batch, cudaStream_t s);
// fv_winner / front non-NULL: the same launch also finishes the C routine's front view (front_finalize_body)
int launch_bin(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
               const uint2 *keys, const unsigned *tile_off, uint4 *binned, cudaStream_t s,
               unsigned *fv_winner = nullptr, float *front = nullptr);
int launch_tile_build(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s,
                      const OccupancySink *sink = nullptr);
int launch_tile_compact(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                        const float *lut, int batch, int grid_blocks, unsigned *compact, unsigned compact_cap,
                        unsigned *compact_table, cudaStream_t s);
int launch_tile_patch(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s);
// SORT strategy (sort_strategy.cu)
unsigned sort_invalid_key(const ViewParams &p, int batch);
size_t sort_temp_bytes(size_t n, unsigned invalid_key);
int launch_sort_keys(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                     unsigned *keys, unsigned *vals, unsigned *fv_winner, unsigned invalid_key, cudaStream_t s);
int launch_sort_pairs(void *temp, size_t temp_bytes, const unsigned *keys_in, unsigned *keys_out, const unsigned *vals_in,
                      unsigned *vals_out, size_t n, unsigned invalid_key, cudaStream_t s);
int launch_sort_reduce(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const float *lut,
                       const unsigned *skeys, const unsigned *svals, size_t n, unsigned invalid_key, float *top,
                       cudaStream_t s);
int launch_front_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch,
                          unsigned *fv_winner, float *front, cudaStream_t s);
int launch_pyfront_link(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const PyFrontParams &fp,
                        int batch, int max_points, unsigned *head, unsigned *next, cudaStream_t s);
int launch_pyfront_reduce(const float4 *pts, const FrameOffsets &offs, const PyFrontParams &fp, int batch, unsigned *head,
                          const unsigned *next, float *front, cudaStream_t s);
int launch_point_keys(const float4 *pts, int n, const ViewParams &p, int *cell, int *pix, int *pix_exact, cudaStream_t s);
int launch_fill_zero(float *dst, size_t count, int sm_count, cudaStream_t s);
int launch_atomic_scatter(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                          float *top, unsigned long long *cellkey, unsigned *fv_winner, cudaStream_t s);
int launch_atomic_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const float *lut,
                           int batch, int max_points, float *top, unsigned long long *cellkey, cudaStream_t s);

// Violation bits of the -DMV3D_SELF_CHECK build (selfcheck.cuh), read and cleared; 0 in a normal build.
unsigned selfcheck_encode();
unsigned selfcheck_sort();
void selfcheck_trip(cudaStream_t s);  // one deliberate violation (bit MV3D_BAD_ON_PURPOSE)

} // namespace mv3d
