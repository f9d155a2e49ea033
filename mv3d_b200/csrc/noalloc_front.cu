// noalloc_front.cu -- mv3d_front_view_noalloc: the front view of a few sweeps WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The front-view path needs
// one scratch word per pixel (the index of the winning point: "last point in input order wins",
// LidarTopAndFrontPreprocess.c:176-193). Everywhere else in this library that scratch is a context-owned cudaMalloc
// buffer that grows on demand; here it is a file-scope device array of a fixed, modest size
// (MV3D_NOALLOC_MAX_PIXELS words = 1 MiB, held by every GPU that loads the library until the process exits).
//
// This translation unit is self-contained on purpose: the array, the one-time cudaGetSymbolAddress per device, the two
// kernels and the entry point all live here (in host code the array's NAME is only a host-side stand-in, not a device
// address; the kernels get the address cudaGetSymbolAddress returned for the current device). The arithmetic is the
// shared header view_math.cuh, so results are bit-identical to mv3d_encode_batch_dev's front view. The entry point
// calls no cudaMalloc / cudaFree / allocating routine and needs no mv3d_ctx (creating one allocates).
//
// All calls on one GPU share the array: every call first clears the words it is going to use (stream-ordered), so a
// call never sees an earlier call's contents -- but two calls that OVERLAP on the GPU (different streams, no ordering
// between them) race on it. Order them on one stream or with events.
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

using namespace mv3d;

// The scratch: winning point index + 1 per pixel and sweep (0 = no point), [batch][R * C].
static __device__ __align__(256) unsigned g_noalloc_winners[MV3D_NOALLOC_MAX_PIXELS];

namespace {

constexpr int kThreads = 256;
constexpr int kMaxDevices = 64;

// Pass 1: gate every point by the bird's-eye-view box, compute its pixel, elect the last point per pixel.
__global__ void __launch_bounds__(kThreads)
noalloc_front_elect_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                           const __grid_constant__ ViewParams p, unsigned *__restrict__ winners_all)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *winners = winners_all + (size_t)f * p.front_pixels;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 pt = load_point(pts + base + i);
        const CellHit c = classify_as<false>(pt, p);
        if (!c.keep) continue;
        const int pix = front_pixel(pt, p);
        if (pix >= 0) atomicMax(winners + pix, (unsigned)i + 1u);
    }
}

// Pass 2: the whole [R][C][F] image from the winners (zeros where no point landed).
__global__ void __launch_bounds__(kThreads)
noalloc_front_write_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                           const __grid_constant__ ViewParams p, const unsigned *__restrict__ winners_all,
                           float *__restrict__ front)
{
    const int f = blockIdx.y;
    const unsigned *winners = winners_all + (size_t)f * p.front_pixels;
    float *out = front + (size_t)f * p.front_frame_floats;
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < p.front_pixels; pix += gridDim.x * blockDim.x) {
        const unsigned w = winners[pix];
        float v0 = 0.0f, v1 = 0.0f, v2 = 0.0f;
        if (w) front_values(pts[offs.off[f] + (int)(w - 1u)], p, v0, v1, v2);
        float *o = out + (size_t)pix * p.fn;
        o[0] = v0; o[1] = v1; o[2] = v2;
        for (int c = 3; c < p.fn; c++) o[c] = 0.0f;
    }
}

// Device address of the scratch array on the CURRENT device: looked up once per device, with that device current.
int scratch_address(unsigned **out)
{
    static std::mutex mu;
    static unsigned *addr[kMaxDevices] = {nullptr};
    int device = 0;
    CUDA_TRY(cudaGetDevice(&device));
    if (device < 0 || device >= kMaxDevices) return fail("mv3d_front_view_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        unsigned *ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_noalloc_winners));
        if (!ptr) return fail("mv3d_front_view_noalloc: cudaGetSymbolAddress returned NULL");
        addr[device] = ptr;
    }
    *out = addr[device];
    return 0;
}

float fast_reciprocal(float div)  // as in mv3d_abi.cu: NaN sends quantise() through the IEEE division
{
    const float inv = 1.0f / div;
    return (div > 0.0f && std::isfinite(inv) && inv >= 1.17549435e-38f) ? inv : std::nanf("");
}

bool integral(float v) { return std::isfinite(v) && v == std::floor(v); }

} // namespace

extern "C" int mv3d_front_view_noalloc(const float *points_dev, const int *offsets_host, int batch, float *front_dev,
                                       const mv3d_top_params *tp, const mv3d_front_params *fp, int semantics,
                                       void *stream)
{
    const char *who = "mv3d_front_view_noalloc";
    if (batch < 0) return fail("%s: negative batch %d", who, batch);
    if (batch == 0) return 0;
    if (!offsets_host || !front_dev || !tp || !fp) return fail("%s: NULL argument", who);
    if (batch > kMaxBatchPerLaunch) return fail("%s: at most %d sweeps per call (got %d)", who, kMaxBatchPerLaunch, batch);
    if (semantics != MV3D_SEM_TOP && semantics != MV3D_SEM_TOPFRONT) return fail("%s: unknown semantics %d", who, semantics);
    if (tp->flags) return fail("%s: top flags are not supported here (got 0x%x)", who, tp->flags);
    if (tp->xn < 0 || tp->yn < 0 || tp->zn < 0) return fail("%s: negative grid size %d x %d x %d", who, tp->xn, tp->yn, tp->zn);
    if (!integral(fp->c_size) || !integral(fp->r_size) || !integral(fp->f_size) || fp->c_size < 1 || fp->r_size < 1 ||
        fp->f_size < 3)
        return fail("%s: front view sizes must be integral with C,R >= 1 and F >= 3 (got C=%g R=%g F=%g)", who,
                    fp->c_size, fp->r_size, fp->f_size);
    if ((double)fp->c_size * fp->r_size * fp->f_size >= 16777216.0)
        return fail("%s: front view of %g floats exceeds the reference's exact float index range", who,
                    (double)fp->c_size * fp->r_size * fp->f_size);
    const long long pixels = (long long)fp->c_size * (long long)fp->r_size;
    if (pixels * batch > (long long)MV3D_NOALLOC_MAX_PIXELS)
        return fail("%s: %d sweeps x %lld pixels = %lld scratch words, the static scratch holds %d "
                    "(MV3D_NOALLOC_MAX_PIXELS); use mv3d_encode_batch_dev, which allocates", who, batch, pixels,
                    pixels * batch, (int)MV3D_NOALLOC_MAX_PIXELS);
    FrameOffsets offs;
    memset(&offs, 0, sizeof(offs));
    if (offsets_host[0] < 0) return fail("%s: offsets[0] = %d is negative", who, offsets_host[0]);
    int max_points = 0;
    for (int i = 0; i < batch; i++) {
        if (offsets_host[i + 1] < offsets_host[i])
            return fail("%s: offsets decrease at sweep %d (%d -> %d)", who, i, offsets_host[i], offsets_host[i + 1]);
        max_points = std::max(max_points, offsets_host[i + 1] - offsets_host[i]);
    }
    for (int i = 0; i <= batch; i++) offs.off[i] = offsets_host[i] - offsets_host[0];
    if (offs.off[batch] > 0 && !points_dev) return fail("%s: points_dev is NULL", who);
    if (offs.off[batch] > 0 && (reinterpret_cast<uintptr_t>(points_dev) & 15u)) return fail("%s: points_dev must be 16-byte aligned", who);

    // the kernel parameter block: the subset of make_params (mv3d_abi.cu) the front view needs
    ViewParams p;
    memset(&p, 0, sizeof(p));
    p.x_min = tp->x_min; p.y_min = tp->y_min; p.z_min = tp->z_min;
    p.x_div = tp->x_div; p.y_div = tp->y_div; p.z_div = tp->z_div;
    p.x_inv = fast_reciprocal(tp->x_div); p.y_inv = fast_reciprocal(tp->y_div); p.z_inv = fast_reciprocal(tp->z_div);
    p.xn = tp->xn; p.yn = tp->yn; p.zn = tp->zn; p.planes = tp->zn + 2;
    p.quant_floor = (semantics == MV3D_SEM_TOPFRONT) ? 1 : 0;
    p.want_front = 1;
    p.z_span = tp->z_max - tp->z_min;        // float subtraction, LidarTopAndFrontPreprocess.c:185
    p.rad_to_deg = (float)(180 / 3.141596);  // sic, :36
    p.phi_div = fp->phi_div; p.theta_div = fp->theta_div;
    p.phi_scale = p.rad_to_deg / fp->phi_div; p.theta_scale = p.rad_to_deg / fp->theta_div;
    p.phi_eps = std::fabs(p.phi_scale) * 2.5e-6f; p.theta_eps = std::fabs(p.theta_scale) * 2.5e-6f;
    p.r_size_f = fp->r_size; p.c_size_f = fp->c_size;
    p.rn = (int)fp->r_size; p.cn = (int)fp->c_size; p.fn = (int)fp->f_size;
    p.center_c = (int)(fp->c_size / 2);  // :37
    p.center_r = (int)(fp->r_size / 2);  // :38
    p.front_pixels = p.rn * p.cn;
    p.front_frame_floats = (long long)p.front_pixels * p.fn;

    unsigned *scratch = nullptr;
    if (scratch_address(&scratch)) return -1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // the array keeps whatever the last call (or an aborted one) left: clear the words this call uses
    CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)batch * (size_t)p.front_pixels * sizeof(unsigned), s));
    const float4 *pts = reinterpret_cast<const float4 *>(points_dev) + offsets_host[0];
    if (max_points > 0) {
        int bx = (max_points + kThreads - 1) / kThreads;
        bx = bx > 1024 ? 1024 : bx;
        noalloc_front_elect_kernel<<<dim3(bx, batch), kThreads, 0, s>>>(pts, offs, p, scratch);
    }
    int bx = (p.front_pixels + kThreads - 1) / kThreads;
    bx = bx > 1024 ? 1024 : bx;
    noalloc_front_write_kernel<<<dim3(bx, batch), kThreads, 0, s>>>(pts, offs, p, scratch, front_dev);
    CUDA_TRY(cudaGetLastError());
    return 0;
}
