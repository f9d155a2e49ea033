// tma_store_bench.cu -- how fast can shared->global bulk stores (cp.async.bulk, the TMA non-tensor form) write HBM,
// against plain vector stores and cudaMemset, for a buffer the size of one step's output (3.9 GB)?
// The tile kernel's stores are of this kind; this separates "limit of the store path" from "limit of the kernel".
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_store_bench tools/tma_store_bench.cu
//   tools/tma_store_bench            (prints one line per configuration)
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int PENDING>
__global__ void tma_fill(char *dst, size_t bytes, uint32_t chunk, int hint)
{
    extern __shared__ __align__(128) unsigned char zeros[];
    for (uint32_t w = threadIdx.x; w < chunk / 4; w += blockDim.x) reinterpret_cast<unsigned *>(zeros)[w] = 0u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        const size_t nchunk = bytes / chunk;
        for (size_t c = blockIdx.x; c < nchunk; c += gridDim.x) {
            char *g = dst + c * chunk;
            if (hint)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                             :: "l"(g), "r"(smem_addr(zeros)), "r"(chunk), "l"(pol) : "memory");
            else
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             :: "l"(g), "r"(smem_addr(zeros)), "r"(chunk) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (PENDING >= 0) asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(PENDING < 0 ? 0 : PENDING) : "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

__global__ void lsu_fill(float4 *dst, size_t n16)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) dst[i] = make_float4(0, 0, 0, 0);
}

// each CTA copies `chunk`-byte pieces from its shared memory with vector stores (the tile kernel's LSU variant)
__global__ void lsu_tile_fill(char *dst, size_t bytes, uint32_t chunk)
{
    extern __shared__ __align__(128) unsigned char zeros[];
    for (uint32_t w = threadIdx.x; w < chunk / 4; w += blockDim.x) reinterpret_cast<unsigned *>(zeros)[w] = 0u;
    __syncthreads();
    const size_t nchunk = bytes / chunk;
    for (size_t c = blockIdx.x; c < nchunk; c += gridDim.x) {
        uint4 *g = reinterpret_cast<uint4 *>(dst + c * chunk);
        const uint4 *s = reinterpret_cast<const uint4 *>(zeros);
        for (uint32_t w = threadIdx.x; w < chunk / 16; w += blockDim.x) g[w] = s[w];
    }
}

template <typename F>
static float timed(F &&launch, int reps = 10)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; i++) launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; i++) launch();
    CK(cudaEventRecord(b));
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

int main(int argc, char **argv)
{
    const size_t bytes = (size_t)64 * 704 * 800 * 27 * 4;
    char *dst;
    CK(cudaMalloc(&dst, bytes));
    int sms;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    if (argc == 4) {  // one configuration, for a profiler: <chunk bytes> <CTAs per SM> <threads per CTA>
        const uint32_t chunk = (uint32_t)atoi(argv[1]);
        const int k = atoi(argv[2]), threads = atoi(argv[3]);
        CK(cudaFuncSetAttribute(tma_fill<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
        float m = timed([&] { tma_fill<-1><<<sms * k, threads, chunk>>>(dst, bytes, chunk, 0); }, 5);
        printf("tma_fill chunk %u, %d CTA/SM x %d threads, never waiting: %.4f ms %7.1f GB/s\n", chunk, k, threads, m,
               bytes / m / 1e6);
        return 0;
    }
    float ms = timed([&] { cudaMemsetAsync(dst, 0, bytes); });
    printf("cudaMemset                                   %.4f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    ms = timed([&] { lsu_fill<<<sms * 8, 256>>>(reinterpret_cast<float4 *>(dst), bytes / 16); });
    printf("lsu_fill 8x256/SM                            %.4f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    ms = timed([&] { lsu_fill<<<sms * 16, 128>>>(reinterpret_cast<float4 *>(dst), bytes / 16); });
    printf("lsu_fill 16x128/SM                           %.4f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    const uint32_t chunks[] = {4320, 8640, 17280, 36720, 73440};  // multiples of one 27-plane cell (108 B), 16-B aligned
    const int per_sm[] = {1, 2, 4, 5, 8};
    for (uint32_t chunk : chunks) {
        for (int k : per_sm) {
            if ((size_t)chunk * k > 200 * 1024) continue;
            CK(cudaFuncSetAttribute(tma_fill<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
            CK(cudaFuncSetAttribute(tma_fill<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
            CK(cudaFuncSetAttribute(tma_fill<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
            CK(cudaFuncSetAttribute(tma_fill<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
            for (int hint = 0; hint < 2; hint++) {
                float m0 = timed([&] { tma_fill<0><<<sms * k, 32, chunk>>>(dst, bytes, chunk, hint); }, 5);
                float m1 = timed([&] { tma_fill<1><<<sms * k, 32, chunk>>>(dst, bytes, chunk, hint); }, 5);
                float m4 = timed([&] { tma_fill<4><<<sms * k, 32, chunk>>>(dst, bytes, chunk, hint); }, 5);
                float m16 = timed([&] { tma_fill<-1><<<sms * k, 32, chunk>>>(dst, bytes, chunk, hint); }, 5);
                printf("tma_fill chunk %6u  %d CTA/SM hint %d   wait for smem reads down to 0/1/4 pending, never: %7.1f %7.1f %7.1f %7.1f GB/s\n", chunk, k,
                       hint, bytes / m0 / 1e6, bytes / m1 / 1e6, bytes / m4 / 1e6, bytes / m16 / 1e6);
            }
        }
    }
    for (uint32_t chunk : {(uint32_t)17280, (uint32_t)36720}) {
        CK(cudaFuncSetAttribute(lsu_tile_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chunk));
        for (int k : {2, 4, 5}) {
            ms = timed([&] { lsu_tile_fill<<<sms * k, 256, chunk>>>(dst, bytes, chunk); }, 5);
            printf("lsu_tile_fill chunk %6u %d CTA/SM x256      %.4f ms  %7.1f GB/s\n", chunk, k, ms, bytes / ms / 1e6);
        }
    }
    return 0;
}
