// selfcheck.cuh -- bounds checks of the library's own, for a pool where compute-sanitizer is not available.
// Built with -DMV3D_SELF_CHECK every index a kernel computes into a scratch buffer, a shared-memory tile or an output is
// tested before it is used; a violation sets a bit in a per-translation-unit device word (and the access is skipped
// where that is cheap) which mv3d_debug_selfcheck() reads and clears. The GPU test suite, run on such a build, asserts
// after every test that the word is zero (tests/conftest.py). Without the macro the checks compile to nothing.
#pragma once
#include <cuda_runtime.h>

#ifdef MV3D_SELF_CHECK
static __device__ unsigned mv3d_tu_violations = 0;
#define MV3D_CHECK(cond, bit) \
    do { if (!(cond)) atomicOr(&mv3d_tu_violations, 1u << (bit)); } while (0)
#define MV3D_SELF_CHECK_READER(name)                                                 \
    unsigned name()                                                                  \
    {                                                                                \
        unsigned v = 0, zero = 0;                                                    \
        if (cudaMemcpyFromSymbol(&v, mv3d_tu_violations, 4) != cudaSuccess) return 0x80000000u; \
        cudaMemcpyToSymbol(mv3d_tu_violations, &zero, 4);                            \
        return v;                                                                    \
    }
#define MV3D_SELF_CHECK_ENABLED 1
#else
#define MV3D_CHECK(cond, bit) do { } while (0)
#define MV3D_SELF_CHECK_READER(name) unsigned name() { return 0u; }
#define MV3D_SELF_CHECK_ENABLED 0
#endif

// Bits of the encode translation unit (encode_kernels.cu); the sort arm uses 16.., the anchor filter 24.. of the word
// mv3d_debug_selfcheck() assembles.
enum {
    MV3D_BAD_TILE = 0,         // classify / bin: tile index beyond the sweep's tiles
    MV3D_BAD_PIXEL = 1,        // front-view pixel beyond the image
    MV3D_BAD_BIN_SLOT = 2,     // bin: record position beyond the sweep's points
    MV3D_BAD_LOCAL_CELL = 3,   // tile kernels: cell-in-tile beyond the tile
    MV3D_BAD_SLICE = 4,        // tile / atomic kernels: height slice beyond the grid
    MV3D_BAD_DIRTY = 5,        // tile kernels: more touched cells than the tile has
    MV3D_BAD_STREAM = 6,       // compact tile kernel: entry beyond the read-back stream
    MV3D_BAD_CELL = 7,         // atomic kernels: output cell beyond the grid
    MV3D_BAD_WINNER = 8,       // front view: winner index beyond the sweep
    MV3D_BAD_BIN_RANGE = 9,    // tile kernels: a tile's bin range outside the sweep's records
    MV3D_BAD_ON_PURPOSE = 15,  // set by mv3d_debug_selfcheck_trip(): proves that a violation is seen and cleared
};
