// host_expand.cu -- fork-join pool and the dense rebuild of a compact read-back (see host_expand.h).
// Host code only; compiled by nvcc together with the kernels so the library stays one translation toolchain.
#include "host_expand.h"

#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <chrono>

#ifdef __linux__
#include <pthread.h>
#include <sched.h>
#endif

#if defined(__SSE2__) || defined(__x86_64__)
#include <emmintrin.h>
#include <xmmintrin.h>
#define MV3D_HAVE_NT_STORES 1
#endif

namespace mv3d {

static inline void cpu_relax()
{
#ifdef MV3D_HAVE_NT_STORES
    _mm_pause();
#endif
}

// Spin for a short while (jobs of one host call follow each other within a few hundred microseconds, far less
// than a futex sleep/wake round trip on a virtualised host), then give up and let the caller block.
template <typename Pred>
static bool spin_until(Pred pred)
{
    using clock = std::chrono::steady_clock;
    const auto deadline = clock::now() + std::chrono::microseconds(300);
    for (int i = 0;; i++) {
        if (pred()) return true;
        cpu_relax();
        if ((i & 63) == 63 && clock::now() > deadline) return pred();
    }
}

std::vector<int> host_cpu_share(int index, int sharers)
{
    std::vector<int> all;
#ifdef __linux__
    cpu_set_t set;
    CPU_ZERO(&set);
    if (sched_getaffinity(0, sizeof(set), &set) == 0)
        for (int c = 0; c < CPU_SETSIZE; c++)
            if (CPU_ISSET(c, &set)) all.push_back(c);
#endif
    if (sharers <= 1 || index < 0 || index >= sharers || (int)all.size() < sharers) return all;
    const size_t per = all.size() / (size_t)sharers;
    return std::vector<int>(all.begin() + (size_t)index * per, all.begin() + ((size_t)index + 1) * per);
}

HostPool::HostPool(int threads, const std::vector<int> &cpus)
{
    if (threads < 1) threads = 1;
    for (int i = 0; i < threads - 1; i++) {
        workers_.emplace_back([this] { worker(); });
#ifdef __linux__
        if (!cpus.empty()) {  // worker 0 leaves the first CPU of the range to the calling thread
            cpu_set_t set;
            CPU_ZERO(&set);
            CPU_SET(cpus[(size_t)(i + 1) % cpus.size()], &set);
            pthread_setaffinity_np(workers_.back().native_handle(), sizeof(set), &set);
        }
#endif
    }
}

HostPool::~HostPool()
{
    {
        std::lock_guard<std::mutex> lock(mu_);
        stop_.store(true, std::memory_order_release);
    }
    cv_work_.notify_all();
    for (auto &t : workers_) t.join();
}

// Take parts of the current job until none are left.
void HostPool::drain()
{
    for (;;) {
        const int part = next_.fetch_add(1, std::memory_order_relaxed);
        if (part >= parts_) break;
        (*fn_)(part);
    }
}

void HostPool::worker()
{
    unsigned long long seen = 0;
    auto ready = [&] { return stop_.load(std::memory_order_acquire) || generation_.load(std::memory_order_acquire) != seen; };
    for (;;) {
        if (!spin_until(ready)) {
            std::unique_lock<std::mutex> lock(mu_);
            cv_work_.wait(lock, ready);
        }
        if (stop_.load(std::memory_order_acquire)) return;
        seen = generation_.load(std::memory_order_acquire);
        drain();
        done_.fetch_add(1, std::memory_order_release);
        { std::lock_guard<std::mutex> lock(mu_); }  // pairs with the caller's predicate check: no lost wake-up
        cv_done_.notify_one();
    }
}

void HostPool::parallel_for(int parts, const std::function<void(int)> &fn, const std::function<void()> *caller_first)
{
    if (parts <= 0) {
        if (caller_first) (*caller_first)();
        return;
    }
    if (workers_.empty()) {
        if (caller_first) (*caller_first)();
        for (int i = 0; i < parts; i++) fn(i);
        return;
    }
    fn_ = &fn;
    parts_ = parts;
    next_.store(0, std::memory_order_relaxed);
    done_.store(0, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> lock(mu_);
        generation_.fetch_add(1, std::memory_order_release);  // publishes fn_/parts_/next_/done_
    }
    cv_work_.notify_all();
    if (caller_first) (*caller_first)();  // the caller's own errand (e.g. queueing GPU work) overlaps the workers
    drain();
    // every worker reports once per generation; wait for all of them so fn_ is never used after we return
    auto all_done = [&] { return done_.load(std::memory_order_acquire) == (int)workers_.size(); };
    if (!spin_until(all_done)) {
        std::unique_lock<std::mutex> lock(mu_);
        cv_done_.wait(lock, all_done);
    }
    fn_ = nullptr;
}

// ---- streaming writers -----------------------------------------------------------------------------------------
// The destination is written exactly once, front to back, with non-temporal stores on whole 64-byte lines (no
// read-for-ownership, no cache pollution); only the first and last partial line of a run use ordinary stores.

// Zero [p, p + bytes).
static void zero_stream(char *p, size_t bytes)
{
#ifdef MV3D_HAVE_NT_STORES
    size_t head = (size_t)((64 - ((uintptr_t)p & 63)) & 63);
    if (head > bytes) head = bytes;
    memset(p, 0, head);
    p += head;
    bytes -= head;
    const __m128i z = _mm_setzero_si128();
    size_t i = 0;
    for (; i + 64 <= bytes; i += 64) {
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i), z);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 16), z);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 32), z);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 48), z);
    }
    memset(p + i, 0, bytes - i);
#else
    memset(p, 0, bytes);
#endif
}

// Copy [src, src + bytes) to p.
static void copy_stream(char *p, const char *src, size_t bytes)
{
#ifdef MV3D_HAVE_NT_STORES
    size_t head = (size_t)((64 - ((uintptr_t)p & 63)) & 63);
    if (head > bytes) head = bytes;
    memcpy(p, src, head);
    p += head;
    src += head;
    bytes -= head;
    size_t i = 0;
    for (; i + 64 <= bytes; i += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 32));
        const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 48));
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i), a);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 32), c);
        _mm_stream_si128(reinterpret_cast<__m128i *>(p + i + 48), d);
    }
    memcpy(p + i, src + i, bytes - i);
#else
    memcpy(p, src, bytes);
#endif
}

static void store_fence()
{
#ifdef MV3D_HAVE_NT_STORES
    _mm_sfence();
#endif
}

void copy_fill(HostPool &pool, void *dst, const void *const *srcs, const size_t *dst_offsets, const size_t *bytes, int n)
{
    if (!dst || !srcs || n <= 0) return;
    // pieces of at most 256 KB, so that a few long sources still spread over every thread
    constexpr size_t kPiece = (size_t)256 << 10;
    std::vector<size_t> first(n + 1, 0);
    for (int i = 0; i < n; i++) first[i + 1] = first[i] + (bytes[i] + kPiece - 1) / kPiece;
    const size_t pieces = first[n];
    if (pieces == 0) return;
    const int parts = (int)std::min<size_t>(pieces, (size_t)pool.size() * 8);
    const std::function<void(int)> body = [&](int part) {
        const size_t p0 = pieces * (size_t)part / parts, p1 = pieces * ((size_t)part + 1) / parts;
        int i = (int)(std::upper_bound(first.begin(), first.end(), p0) - first.begin()) - 1;
        for (size_t p = p0; p < p1; p++) {
            while (p >= first[i + 1]) i++;
            const size_t off = (p - first[i]) * kPiece, len = std::min(kPiece, bytes[i] - off);
            copy_stream(static_cast<char *>(dst) + dst_offsets[i] + off, static_cast<const char *>(srcs[i]) + off, len);
        }
        store_fence();
    };
    pool.parallel_for(parts, body);
}

void zero_fill(HostPool &pool, float *dst, size_t floats)
{
    if (!dst || floats == 0) return;
    const int parts = pool.size() * 4;
    const size_t bytes = floats * 4;
    const size_t per = ((bytes + parts - 1) / parts + 63) & ~(size_t)63;
    const std::function<void(int)> body = [&](int part) {
        const size_t b0 = (size_t)part * per;
        if (b0 >= bytes) return;
        zero_stream(reinterpret_cast<char *>(dst) + b0, (b0 + per < bytes) ? per : bytes - b0);
        store_fence();
    };
    pool.parallel_for(parts, body);
}

// ---- the read-back stream ---------------------------------------------------------------------------------------
// What tile_compact_kernel writes for one sweep: a stream of 32-bit words, per touched cell
//   { cell index, plane masks [G = ceil(planes / 32)], the values of the planes whose mask bit is set, in plane order }
// and a table { first word, touched cells } per tile (tile t = tile_cells consecutive cells); a tile's cells are in
// ASCENDING order. Everything read from the stream is checked before it is used: a cell outside its tile's range or
// out of order is skipped, an entry that runs past the stream ends its tile -- whatever the stream holds, nothing is
// written outside the part that owns the tile.
namespace {
struct CellEntry {
    size_t cell;
    const unsigned *masks;   // G words
    const unsigned *values;  // the non-zero planes
    unsigned next;           // first word after this entry
};
inline int popcount32(unsigned v) { return __builtin_popcount(v); }

// Decode the entry at word w (false: the stream ends here).
inline bool read_entry(const unsigned *words, unsigned n_words, unsigned w, int groups, CellEntry &e)
{
    if (w >= n_words || n_words - w < 1u + (unsigned)groups) return false;
    e.cell = words[w];
    e.masks = words + w + 1;
    unsigned nnz = 0;
    for (int g = 0; g < groups; g++) nnz += (unsigned)popcount32(e.masks[g]);
    if (n_words - w - 1u - (unsigned)groups < nnz) return false;
    e.values = e.masks + groups;
    e.next = w + 1u + (unsigned)groups + nnz;
    return true;
}

// Place the entry's values into a zeroed cell image of `planes` floats (mask bits at or beyond `planes` are ignored).
inline void place_values(unsigned *cellw, int planes, int groups, const CellEntry &e)
{
    const unsigned *v = e.values;
    for (int g = 0; g < groups; g++) {
        unsigned m = e.masks[g];
        while (m) {
            const int k = g * 32 + __builtin_ctz(m);
            m &= m - 1u;
            if (k < planes) cellw[k] = *v;
            v++;
        }
    }
}
}  // namespace

// Scatter the stream's values into a destination that is zero already (the second step of the split rebuild).
void scatter_stream(HostPool &pool, float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                    const unsigned *table, int tiles, int tile_cells, const std::function<void()> *caller_first)
{
    if (!dst || cells == 0 || planes <= 0 || !table || tiles <= 0 || tile_cells <= 0 || n_words == 0) {
        if (caller_first) (*caller_first)();
        return;
    }
    const int groups = (planes + 31) / 32;
    int parts = pool.size() * 4;
    if (parts > tiles) parts = tiles;
    const size_t tiles_per_part = ((size_t)tiles + parts - 1) / parts;
    const std::function<void(int)> body = [&](int part) {
        const size_t t0 = (size_t)part * tiles_per_part;
        const size_t t1 = (t0 + tiles_per_part < (size_t)tiles) ? t0 + tiles_per_part : (size_t)tiles;
        for (size_t t = t0; t < t1; t++) {
            const size_t lo = t * (size_t)tile_cells;
            const size_t hi = (lo + (size_t)tile_cells < cells) ? lo + (size_t)tile_cells : cells;
            size_t floor_cell = lo;
            unsigned w = table[2 * t];
            CellEntry e;
            for (unsigned c = 0; c < table[2 * t + 1] && read_entry(words, n_words, w, groups, e); c++, w = e.next) {
                if (e.cell < floor_cell || e.cell >= hi) continue;  // a tile's cells are disjoint from every other tile's
                floor_cell = e.cell + 1;
                place_values(reinterpret_cast<unsigned *>(dst) + e.cell * (size_t)planes, planes, groups, e);
            }
        }
    };
    pool.parallel_for(parts, body, caller_first);
}

// ---- expand_sorted: the rebuild as ONE linear walk ---------------------------------------------------------------
// A part (a run of whole tiles = a contiguous byte range of the destination) walks its tiles' entries front to back,
// streams zeros over the gaps and composes only the few 64-byte lines a cell (or a run of neighbouring cells) touches
// in a small window. No sorting, no per-block tables, every byte of the destination written once.
void expand_sorted(HostPool &pool, float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                   const unsigned *table, int tiles, int tile_cells, const std::function<void()> *caller_first)
{
    if (!dst || cells == 0 || planes <= 0) {
        if (caller_first) (*caller_first)();
        return;
    }
    if (!table || tiles <= 0 || tile_cells <= 0) {  // nothing to place
        if (caller_first) (*caller_first)();
        zero_fill(pool, dst, cells * (size_t)planes);
        return;
    }
    constexpr size_t kWin = 1024;  // window of composed lines; a cell must fit with a line to spare on either side
    const size_t cell_bytes = (size_t)planes * 4;
    const bool windowed = cell_bytes + 192 <= kWin;  // very deep grids: zero the part, then place the values
    const int groups = (planes + 31) / 32;
    // Parts are claimed dynamically; several per thread, so that the caller's errand and a slow core cost a fraction
    // of a part, not a whole share.
    int parts = pool.size() * 6;
    if (parts > tiles) parts = tiles;
    const size_t tiles_per_part = ((size_t)tiles + parts - 1) / parts;
    const std::function<void(int)> body = [&](int part) {
        const size_t t0 = (size_t)part * tiles_per_part;
        if (t0 >= (size_t)tiles) return;
        const size_t t1 = (t0 + tiles_per_part < (size_t)tiles) ? t0 + tiles_per_part : (size_t)tiles;
        const size_t c0 = t0 * (size_t)tile_cells;
        if (c0 >= cells) return;
        const size_t c1 = (t1 * (size_t)tile_cells < cells && t1 < (size_t)tiles) ? t1 * (size_t)tile_cells : cells;
        char *const base = reinterpret_cast<char *>(dst);
        const uintptr_t A0 = reinterpret_cast<uintptr_t>(base + c0 * cell_bytes);  // this part = addresses [A0, A1)
        const uintptr_t A1 = reinterpret_cast<uintptr_t>(base + c1 * cell_bytes);
        if (!windowed) {
            zero_stream(reinterpret_cast<char *>(A0), (size_t)(A1 - A0));
            store_fence();  // the values below land in the same lines with ordinary stores
        }
        alignas(64) char win[kWin];
        uintptr_t done = A0;        // [A0, done) has been emitted
        uintptr_t wlo = 0, whi = 0; // the open window holds [wlo, whi), wlo == done; whi is line-aligned or A1
        bool open = false;
        size_t floor_cell = c0;     // the next accepted cell must be >= this: ascending order, and memory safety
        for (size_t t = t0; t < t1; t++) {
            if (floor_cell < t * (size_t)tile_cells) floor_cell = t * (size_t)tile_cells;
            const size_t tile_end = ((t + 1) * (size_t)tile_cells < c1) ? (t + 1) * (size_t)tile_cells : c1;
            unsigned w = table[2 * t];
            CellEntry e;
            for (unsigned c = 0; c < table[2 * t + 1] && read_entry(words, n_words, w, groups, e); c++, w = e.next) {
                if (e.cell < floor_cell || e.cell >= tile_end) continue;  // out of contract: never written
                floor_cell = e.cell + 1;
#ifdef MV3D_HAVE_NT_STORES
                _mm_prefetch(reinterpret_cast<const char *>(words + e.next + 64), _MM_HINT_T0);
#endif
                const uintptr_t a = reinterpret_cast<uintptr_t>(base + e.cell * cell_bytes);
                if (!windowed) {
                    place_values(reinterpret_cast<unsigned *>(a), planes, groups, e);
                    continue;
                }
                uintptr_t end = (a + cell_bytes + 63) & ~(uintptr_t)63;
                if (end > A1) end = A1;
                if (open && a <= whi) {                              // touches the window's last line: extend it
                    if (end - wlo > kWin) {                          // full: emit the lines in front of this cell
                        const uintptr_t f = a & ~(uintptr_t)63;      // > wlo here, and every byte below `a` is final
                        copy_stream(reinterpret_cast<char *>(wlo), win, (size_t)(f - wlo));
                        memmove(win, win + (f - wlo), (size_t)(whi - f));
                        wlo = done = f;
                    }
                    if (end > whi) {
                        memset(win + (whi - wlo), 0, (size_t)(end - whi));
                        whi = end;
                    }
                } else {
                    if (open) {
                        copy_stream(reinterpret_cast<char *>(wlo), win, (size_t)(whi - wlo));
                        done = whi;
                    }
                    uintptr_t lo = a & ~(uintptr_t)63;
                    if (lo < done) lo = done;                        // only the part's first line can be cut
                    if (lo > done) zero_stream(reinterpret_cast<char *>(done), (size_t)(lo - done));
                    wlo = done = lo;
                    whi = end;
                    open = true;
                    memset(win, 0, (size_t)(whi - wlo));
                }
                place_values(reinterpret_cast<unsigned *>(win + (a - wlo)), planes, groups, e);
            }
        }
        if (windowed) {
            if (open) {
                copy_stream(reinterpret_cast<char *>(wlo), win, (size_t)(whi - wlo));
                done = whi;
            }
            if (A1 > done) zero_stream(reinterpret_cast<char *>(done), (size_t)(A1 - done));
        }
        store_fence();
    };
    pool.parallel_for(parts, body, caller_first);
}

} // namespace mv3d
