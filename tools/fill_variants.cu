// fill_variants.cu -- which store instruction / access order lets a KERNEL write HBM as fast as cudaMemset does?
// 3.89 GB zero-fill (one step's output of the view encoder). cudaMemset reaches 7.3-7.5 TB/s on a B200, a plain
// float4 fill 6.1, shared->global bulk stores 6.2-6.4 (tools/tma_store_bench.cu). This tries the LSU store flavours
// (cache operators, 256-bit stores, L2 eviction hints) and two orders (grid-stride by 512 B per warp, or one
// contiguous slab per CTA walked front to back).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fill_variants tools/fill_variants.cu
//   tools/fill_variants            all variants       tools/fill_variants memset     only cudaMemset (for ncu)
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

enum Flavour { kPlain, kCs, kCg, kWt, kEvictFirst, kEvictLast, kNoAlloc, kV8, kV8EvictFirst, kFlavours };
static const char *kNames[] = {"st.v4", "st.cs.v4", "st.cg.v4", "st.wt.v4", "st.v4 L2::evict_first", "st.v4 L2::evict_last",
                               "st.v4 L1::no_allocate", "st.v8", "st.v8 L2::evict_first"};

template <int F>
__device__ __forceinline__ void store32(char *p, uint64_t pol)  // 32 bytes of zeros at p
{
    const float z = 0.0f;
    if (F == kPlain) {
        asm volatile("st.global.v4.f32 [%0], {%1,%1,%1,%1};\n\tst.global.v4.f32 [%0+16], {%1,%1,%1,%1};" :: "l"(p), "f"(z) : "memory");
    } else if (F == kCs) {
        asm volatile("st.global.cs.v4.f32 [%0], {%1,%1,%1,%1};\n\tst.global.cs.v4.f32 [%0+16], {%1,%1,%1,%1};" :: "l"(p), "f"(z) : "memory");
    } else if (F == kCg) {
        asm volatile("st.global.cg.v4.f32 [%0], {%1,%1,%1,%1};\n\tst.global.cg.v4.f32 [%0+16], {%1,%1,%1,%1};" :: "l"(p), "f"(z) : "memory");
    } else if (F == kWt) {
        asm volatile("st.global.wt.v4.f32 [%0], {%1,%1,%1,%1};\n\tst.global.wt.v4.f32 [%0+16], {%1,%1,%1,%1};" :: "l"(p), "f"(z) : "memory");
    } else if (F == kEvictFirst || F == kEvictLast) {
        asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%1,%1,%1}, %2;\n\tst.global.L2::cache_hint.v4.f32 [%0+16], {%1,%1,%1,%1}, %2;"
                     :: "l"(p), "f"(z), "l"(pol) : "memory");
    } else if (F == kNoAlloc) {
        asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%1,%1,%1};\n\tst.global.L1::no_allocate.v4.f32 [%0+16], {%1,%1,%1,%1};"
                     :: "l"(p), "f"(z) : "memory");
    } else if (F == kV8) {
        asm volatile("st.global.v8.f32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" :: "l"(p), "f"(z) : "memory");
    } else {
        asm volatile("st.global.L2::cache_hint.v8.f32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1}, %2;" :: "l"(p), "f"(z), "l"(pol) : "memory");
    }
}

// order 0: grid-stride, a warp writes 1 KB (32 lanes x 32 B) per step, consecutive warps consecutive KBs
// order 1: each CTA owns one contiguous slab of the buffer and walks it front to back
template <int F>
__global__ void fill(char *dst, size_t bytes, int order)
{
    uint64_t pol = 0;
    if (F == kEvictFirst || F == kV8EvictFirst) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    if (F == kEvictLast) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    const size_t n32 = bytes / 32;
    if (order == 0) {
        const size_t stride = (size_t)gridDim.x * blockDim.x;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += stride) store32<F>(dst + i * 32, pol);
    } else {
        const size_t per = (n32 + gridDim.x - 1) / gridDim.x;
        const size_t beg = per * blockIdx.x, end = beg + per < n32 ? beg + per : n32;
        for (size_t i = beg + threadIdx.x; i < end; i += blockDim.x) store32<F>(dst + i * 32, pol);
    }
}

template <typename Fn>
static float timed(Fn &&launch, int reps = 10)
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

template <int F>
static void run(char *dst, size_t bytes, int sms)
{
    for (int order = 0; order < 2; order++)
        for (int k : {2, 4, 8, 16})
            for (int threads : {256, 1024}) {
                if (k * threads > 2048) continue;
                float ms = timed([&] { fill<F><<<sms * k, threads>>>(dst, bytes, order); }, 5);
                printf("%-24s order %d  %2d CTA/SM x %4d   %.4f ms  %7.1f GB/s\n", kNames[F], order, k, threads, ms, bytes / ms / 1e6);
            }
}

int main(int argc, char **argv)
{
    const size_t bytes = (size_t)64 * 704 * 800 * 27 * 4;
    char *dst;
    CK(cudaMalloc(&dst, bytes));
    int sms;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    float ms = timed([&] { cudaMemsetAsync(dst, 0, bytes); });
    printf("cudaMemset(0)                                         %.4f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    ms = timed([&] { cudaMemsetAsync(dst, 0x5a, bytes); });
    printf("cudaMemset(0x5a)                                      %.4f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    ms = timed([&] { cudaMemsetAsync(dst + 16, 0, bytes / 2); });
    printf("cudaMemset(0), half the buffer                        %.4f ms  %7.1f GB/s\n", ms, bytes / 2 / ms / 1e6);
    if (argc > 1 && strcmp(argv[1], "memset") == 0) return 0;
    run<kPlain>(dst, bytes, sms);
    run<kCs>(dst, bytes, sms);
    run<kCg>(dst, bytes, sms);
    run<kWt>(dst, bytes, sms);
    run<kEvictFirst>(dst, bytes, sms);
    run<kEvictLast>(dst, bytes, sms);
    run<kNoAlloc>(dst, bytes, sms);
    run<kV8>(dst, bytes, sms);
    run<kV8EvictFirst>(dst, bytes, sms);
    return 0;
}
