// host_expand.h -- host side of the compact read-back (internal).
//
// The bird's-eye-view tensor is ~95 % zeros, so the host-buffer entry points do not drag the dense tensor over
// PCIe: the GPU reports only the touched cells ({cell index, planes values}, fully computed on the device) and
// the host rebuilds the dense layout the reference ABI promises -- a zero-fill plus a scatter of finished values.
// No encoding arithmetic happens here; this is the delivery of a result, not a CPU implementation of the path.
#pragma once
#include <stddef.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace mv3d {

// A small fork-join pool: parallel_for(parts, fn) runs fn(0..parts-1) on the workers and the calling thread.
class HostPool {
public:
    // `cpus`: if non-empty, worker i is pinned to cpus[i % size] (ranks that share a host take disjoint core ranges,
    // see host_cpu_share); the calling thread is never re-pinned.
    explicit HostPool(int threads, const std::vector<int> &cpus = std::vector<int>());
    ~HostPool();
    int size() const { return (int)workers_.size() + 1; }
    // `caller_first`, if given, runs on the calling thread after the workers have been released and before the
    // caller joins them.
    void parallel_for(int parts, const std::function<void(int)> &fn, const std::function<void()> *caller_first = nullptr);

private:
    void worker();
    void drain();
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_work_, cv_done_;
    const std::function<void(int)> *fn_ = nullptr;
    int parts_ = 0;
    std::atomic<int> next_{0};
    std::atomic<int> done_{0};
    std::atomic<unsigned long long> generation_{0};
    std::atomic<bool> stop_{false};
};

// The CPUs this process may run on (sched_getaffinity), cut into `sharers` equal consecutive ranges; returns range
// number `index`. One process per GPU under torchrun: sharers = LOCAL_WORLD_SIZE, index = LOCAL_RANK.
std::vector<int> host_cpu_share(int index, int sharers);

// Gather n source ranges into one destination buffer with streaming stores (no read-for-ownership of the destination,
// which a DMA reads next): source i goes to dst + dst_offsets[i], bytes[i] long. Cut into 256 KB pieces over the pool.
void copy_fill(HostPool &pool, void *dst, const void *const *srcs, const size_t *dst_offsets, const size_t *bytes, int n);

// The dense view dst[cells][planes] from what tile_compact_kernel reports for one sweep (see the layout at the top of
// the stream section in host_expand.cu): `words` is the stream of n_words 32-bit words -- per touched cell its index,
// ceil(planes / 32) plane masks and the values of the planes whose bit is set -- and `table` holds
// {first word, touched cells} per tile (tile t = tile_cells consecutive cells, its cells in ascending order).
// The destination is written exactly once, front to back, with streaming stores: one linear walk per part, no sorting.
// Whatever the stream and table hold, nothing is written outside dst: out-of-range or out-of-order cells are skipped,
// entries past the end of the stream end their tile.
void expand_sorted(HostPool &pool, float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                   const unsigned *table, int tiles, int tile_cells, const std::function<void()> *caller_first = nullptr);

// The same result in two steps, for a call that has nothing else to overlap with (one sweep per call, the reference
// symbols): zero_fill runs while the GPU is still encoding, scatter_stream places the finished values afterwards.
void zero_fill(HostPool &pool, float *dst, size_t floats);
void scatter_stream(HostPool &pool, float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                    const unsigned *table, int tiles, int tile_cells, const std::function<void()> *caller_first = nullptr);

} // namespace mv3d
