// sweep_io.cu -- on-disk formats either side of the view encoder (include/mv3d_io.h). Host code only.
//
//   .bin sweeps   np.fromfile(path, float32).reshape((-1, 4))  src/utils/batch_loading.py:656, src/raw_data.py:137-139
//   .npy views    np.save / np.load                           src/preprocess_for_mv3d.py:36-37, batch_loading.py:659-667
//   .npz views    np.savez_compressed(path, top_view=top)     src/data.py:521, read at batch_loading.py:51
//
// The .npy header is numpy's format 1.0 restated (numpy/lib/format.py `_write_array_header` + `_wrap_header`):
// magic, version, little-endian uint16 header length, a Python dict literal with sorted keys, spare blanks so the
// first axis can grow to 21 digits in place, blank padding to a multiple of 64 bytes, a newline.
#include "../../include/mv3d_io.h"
#include "ctx.h"
#include "host_expand.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Fd {
    int fd = -1;
    ~Fd() { if (fd >= 0) close(fd); }
};

// fn(i) for i in [0, n) on up to `threads` threads; returns the first failing index's message ("" if none).
template <typename F>
std::string run_parallel(int n, int threads, F &&fn)
{
    if (threads <= 0) threads = 16;
    if (threads > n) threads = n;
    std::atomic<int> next{0};
    std::vector<std::string> errors(threads > 0 ? threads : 1);
    auto work = [&](int t) {
        for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
            std::string e = fn(i);
            if (!e.empty() && errors[t].empty()) errors[t] = e;
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work, t);
    if (n > 0) work(0);
    for (auto &th : pool) th.join();
    for (auto &e : errors)
        if (!e.empty()) return e;
    return "";
}

std::string errno_text(const char *what, const char *path)
{
    return std::string(what) + " " + path + ": " + strerror(errno);
}

std::string read_exact(int fd, void *dst, size_t bytes, off_t offset, const char *path)
{
    char *p = static_cast<char *>(dst);
    while (bytes > 0) {
        const ssize_t got = pread(fd, p, bytes, offset);
        if (got < 0) {
            if (errno == EINTR) continue;
            return errno_text("read", path);
        }
        if (got == 0) return std::string("read ") + path + ": file is shorter than expected";
        p += got;
        offset += got;
        bytes -= (size_t)got;
    }
    return "";
}

std::string write_all(int fd, const void *src, size_t bytes, const char *path)
{
    const char *p = static_cast<const char *>(src);
    while (bytes > 0) {
        const ssize_t put = write(fd, p, bytes);
        if (put < 0) {
            if (errno == EINTR) continue;
            return errno_text("write", path);
        }
        p += put;
        bytes -= (size_t)put;
    }
    return "";
}

// ---- .npy ------------------------------------------------------------------------------------------------
std::string shape_repr(const long long *shape, int ndim)
{
    std::string s = "(";
    for (int i = 0; i < ndim; i++) {
        s += std::to_string(shape[i]);
        if (ndim == 1) s += ",";
        else if (i + 1 < ndim) s += ", ";
    }
    return s + ")";
}

// The complete file prefix np.save writes before the data (format 1.0).
std::string npy_prefix(const long long *shape, int ndim)
{
    std::string dict = "{'descr': '<f4', 'fortran_order': False, 'shape': " + shape_repr(shape, ndim) + ", }";
    if (ndim > 0) {  // spare room so that the first axis can be rewritten in place (GROWTH_AXIS_MAX_DIGITS = 21)
        const int digits = (int)std::to_string(shape[0]).size();
        if (digits < 21) dict.append((size_t)(21 - digits), ' ');
    }
    const size_t hlen = dict.size() + 1;
    const size_t padlen = 64 - ((8 + 2 + hlen) % 64);  // 64, not 0, when already aligned -- as numpy computes it
    dict.append(padlen, ' ');
    dict.push_back('\n');
    std::string out("\x93NUMPY\x01\x00", 8);
    const uint16_t len = (uint16_t)dict.size();
    out.push_back((char)(len & 0xff));
    out.push_back((char)(len >> 8));
    return out + dict;
}

int check_shape(const char *who, const long long *shape, int ndim, size_t *count)
{
    if (ndim < 0 || ndim > MV3D_NPY_MAX_DIMS || (ndim > 0 && !shape)) return fail("%s: ndim = %d", who, ndim);
    size_t n = 1;
    for (int i = 0; i < ndim; i++) {
        if (shape[i] < 0) return fail("%s: negative dimension %lld", who, shape[i]);
        n *= (size_t)shape[i];
    }
    *count = n;
    return 0;
}

std::string write_npy_file(const char *path, const float *data, const long long *shape, int ndim, size_t count)
{
    Fd f;
    f.fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (f.fd < 0) return errno_text("open", path);
    const std::string prefix = npy_prefix(shape, ndim);
    std::string e = write_all(f.fd, prefix.data(), prefix.size(), path);
    if (e.empty()) e = write_all(f.fd, data, count * sizeof(float), path);
    return e;
}

struct NpyInfo {
    long long shape[MV3D_NPY_MAX_DIMS];
    int ndim = 0;
    long long data_offset = 0;
    size_t count = 1;
};

std::string parse_npy(int fd, const char *path, NpyInfo *info)
{
    unsigned char head[12];
    std::string e = read_exact(fd, head, 10, 0, path);
    if (!e.empty()) return e;
    if (memcmp(head, "\x93NUMPY", 6) != 0) return std::string(path) + ": not a .npy file";
    const int major = head[6];
    size_t hlen, pos;
    if (major == 1) {
        hlen = (size_t)head[8] | ((size_t)head[9] << 8);
        pos = 10;
    } else if (major == 2 || major == 3) {
        e = read_exact(fd, head + 10, 2, 10, path);
        if (!e.empty()) return e;
        hlen = (size_t)head[8] | ((size_t)head[9] << 8) | ((size_t)head[10] << 16) | ((size_t)head[11] << 24);
        pos = 12;
    } else {
        return std::string(path) + ": unsupported .npy format version " + std::to_string(major);
    }
    if (hlen > (1u << 20)) return std::string(path) + ": implausible .npy header length";
    std::string dict(hlen, '\0');
    e = read_exact(fd, &dict[0], hlen, (off_t)pos, path);
    if (!e.empty()) return e;
    auto value_after = [&](const char *key) -> size_t {
        const size_t k = dict.find(key);
        if (k == std::string::npos) return std::string::npos;
        const size_t c = dict.find(':', k + strlen(key));
        return c == std::string::npos ? c : c + 1;
    };
    size_t d = value_after("'descr'");
    if (d == std::string::npos) return std::string(path) + ": .npy header has no descr";
    const size_t q0 = dict.find('\'', d);
    const size_t q1 = q0 == std::string::npos ? q0 : dict.find('\'', q0 + 1);
    if (q1 == std::string::npos) return std::string(path) + ": malformed descr";
    const std::string descr = dict.substr(q0 + 1, q1 - q0 - 1);
    if (descr != "<f4") return std::string(path) + ": dtype " + descr + " (only little-endian float32 '<f4' views are read)";
    size_t fo = value_after("'fortran_order'");
    if (fo == std::string::npos) return std::string(path) + ": .npy header has no fortran_order";
    while (fo < dict.size() && dict[fo] == ' ') fo++;
    if (dict.compare(fo, 5, "False") != 0) return std::string(path) + ": Fortran-ordered arrays are not read";
    size_t sh = value_after("'shape'");
    if (sh == std::string::npos) return std::string(path) + ": .npy header has no shape";
    const size_t open_p = dict.find('(', sh), close_p = dict.find(')', sh);
    if (open_p == std::string::npos || close_p == std::string::npos || close_p < open_p)
        return std::string(path) + ": malformed shape";
    info->ndim = 0;
    info->count = 1;
    size_t i = open_p + 1;
    while (i < close_p) {
        while (i < close_p && (dict[i] == ' ' || dict[i] == ',')) i++;
        if (i >= close_p) break;
        if (dict[i] < '0' || dict[i] > '9') return std::string(path) + ": malformed shape";
        long long v = 0;
        while (i < close_p && dict[i] >= '0' && dict[i] <= '9') v = v * 10 + (dict[i++] - '0');
        if (info->ndim >= MV3D_NPY_MAX_DIMS) return std::string(path) + ": more than 8 dimensions";
        info->shape[info->ndim++] = v;
        info->count *= (size_t)v;
    }
    info->data_offset = (long long)(pos + hlen);
    return "";
}

// ---- .npz (zip with one deflated member) -------------------------------------------------------------------
void put16(std::string &s, unsigned v) { s.push_back((char)(v & 0xff)); s.push_back((char)((v >> 8) & 0xff)); }
void put32(std::string &s, unsigned long v) { put16(s, (unsigned)(v & 0xffff)); put16(s, (unsigned)((v >> 16) & 0xffff)); }

std::string write_npz_file(const char *path, const char *member, const float *data, const long long *shape, int ndim,
                           size_t count, int level)
{
    const std::string prefix = npy_prefix(shape, ndim);
    const size_t data_bytes = count * sizeof(float);
    const size_t total = prefix.size() + data_bytes;
    if (total >= 0xffffffffull) return std::string(path) + ": arrays of 4 GiB and more need zip64, which is not written";
    const std::string name = std::string(member) + ".npy";

    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    // raw deflate, 32 KiB window, memLevel 8: what Python's zipfile (and so numpy) asks zlib for
    if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return "deflateInit2 failed";
    std::vector<unsigned char> out(deflateBound(&zs, (uLong)total) + 64);
    zs.next_out = out.data();
    zs.avail_out = (uInt)std::min<size_t>(out.size(), 0xffffffffu);
    zs.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(prefix.data()));
    zs.avail_in = (uInt)prefix.size();
    int rc = deflate(&zs, Z_NO_FLUSH);
    if (rc == Z_OK) {
        zs.next_in = reinterpret_cast<Bytef *>(const_cast<float *>(data));
        zs.avail_in = (uInt)data_bytes;
        rc = deflate(&zs, Z_FINISH);
    }
    const size_t csize = zs.total_out;
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) return std::string(path) + ": deflate failed";
    uLong crc = crc32(0L, Z_NULL, 0);
    crc = crc32(crc, reinterpret_cast<const Bytef *>(prefix.data()), (uInt)prefix.size());
    crc = crc32(crc, reinterpret_cast<const Bytef *>(data), (uInt)data_bytes);

    const unsigned dos_time = 0, dos_date = (0u << 9) | (1u << 5) | 1u;  // 1980-01-01 00:00: archives are reproducible
    std::string local;
    put32(local, 0x04034b50ul);
    put16(local, 20); put16(local, 0); put16(local, 8); put16(local, dos_time); put16(local, dos_date);
    put32(local, crc); put32(local, (unsigned long)csize); put32(local, (unsigned long)total);
    put16(local, (unsigned)name.size()); put16(local, 0);
    local += name;
    std::string central;
    put32(central, 0x02014b50ul);
    put16(central, (3u << 8) | 20u); put16(central, 20); put16(central, 0); put16(central, 8);
    put16(central, dos_time); put16(central, dos_date);
    put32(central, crc); put32(central, (unsigned long)csize); put32(central, (unsigned long)total);
    put16(central, (unsigned)name.size()); put16(central, 0); put16(central, 0); put16(central, 0); put16(central, 0);
    put32(central, 0600ul << 16); put32(central, 0);
    central += name;
    std::string end;
    put32(end, 0x06054b50ul);
    put16(end, 0); put16(end, 0); put16(end, 1); put16(end, 1);
    put32(end, (unsigned long)central.size()); put32(end, (unsigned long)(local.size() + csize)); put16(end, 0);

    Fd f;
    f.fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (f.fd < 0) return errno_text("open", path);
    std::string e = write_all(f.fd, local.data(), local.size(), path);
    if (e.empty()) e = write_all(f.fd, out.data(), csize, path);
    if (e.empty()) e = write_all(f.fd, central.data(), central.size(), path);
    if (e.empty()) e = write_all(f.fd, end.data(), end.size(), path);
    return e;
}

} // namespace

extern "C" int mv3d_io_stat_sweeps(const char *const *paths, int n, long long *points_out)
{
    if (n < 0 || (n > 0 && (!paths || !points_out))) return fail("mv3d_io_stat_sweeps: bad arguments");
    for (int i = 0; i < n; i++) {
        struct stat st;
        if (!paths[i] || stat(paths[i], &st) != 0)
            return fail("mv3d_io_stat_sweeps: %s", errno_text("stat", paths[i] ? paths[i] : "(null)").c_str());
        if (st.st_size % 16 != 0)
            return fail("mv3d_io_stat_sweeps: %s holds %lld bytes, not a whole number of (x, y, z, intensity) float32 "
                        "records (numpy's reshape((-1, 4)) raises here too)", paths[i], (long long)st.st_size);
        points_out[i] = (long long)(st.st_size / 16);
    }
    return 0;
}

extern "C" int mv3d_io_read_sweeps(const char *const *paths, int n, float *dst, const long long *offsets, int threads)
{
    if (n < 0 || (n > 0 && (!paths || !dst || !offsets))) return fail("mv3d_io_read_sweeps: bad arguments");
    for (int i = 0; i < n; i++)
        if (offsets[i + 1] < offsets[i]) return fail("mv3d_io_read_sweeps: offsets must not decrease");
    const std::string e = run_parallel(n, threads, [&](int i) -> std::string {
        Fd f;
        f.fd = open(paths[i], O_RDONLY);
        if (f.fd < 0) return errno_text("open", paths[i]);
        struct stat st;
        if (fstat(f.fd, &st) != 0) return errno_text("stat", paths[i]);
        const long long want = (offsets[i + 1] - offsets[i]) * 16;
        if ((long long)st.st_size != want)
            return std::string(paths[i]) + ": holds " + std::to_string((long long)st.st_size) + " bytes, the offsets say " +
                   std::to_string(want);
        return read_exact(f.fd, dst + 4 * offsets[i], (size_t)want, 0, paths[i]);
    });
    return e.empty() ? 0 : fail("mv3d_io_read_sweeps: %s", e.c_str());
}

extern "C" int mv3d_io_pack_sweeps(const float *const *sweeps, int n, float *dst, const long long *offsets, int threads)
{
    if (n < 0 || (n > 0 && (!sweeps || !dst || !offsets))) return fail("mv3d_io_pack_sweeps: bad arguments");
    for (int i = 0; i < n; i++) {
        if (offsets[i + 1] < offsets[i]) return fail("mv3d_io_pack_sweeps: offsets must not decrease");
        if (offsets[i + 1] > offsets[i] && !sweeps[i]) return fail("mv3d_io_pack_sweeps: sweep %d is NULL", i);
    }
    // The batch of a 64-sweep launch is 123 MB and the copy engine reads it next: streaming stores (no read-for-ownership
    // of the staging buffer) from a pool of threads that outlives the call -- starting sixteen threads per batch cost as
    // much as the copy. One pool per process and thread count, calls serialised on it.
    static std::mutex mu;
    static mv3d::HostPool *pool = nullptr;  // never destroyed: joining its threads at exit would hang a forked child
    static int pool_threads = 0;
    static pid_t pool_pid = 0;
    if (threads <= 0) {
        const char *env = getenv("MV3D_IO_THREADS");
        threads = env && atoi(env) > 0 ? atoi(env) : 16;
        const unsigned hw = std::thread::hardware_concurrency();
        if (hw > 0 && (unsigned)threads > hw) threads = (int)hw;
    }
    std::vector<const void *> srcs(n);
    std::vector<size_t> offs(n), bytes(n);
    for (int i = 0; i < n; i++) {
        srcs[i] = sweeps[i];
        offs[i] = (size_t)offsets[i] * 16;
        bytes[i] = (size_t)(offsets[i + 1] - offsets[i]) * 16;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (pool && pool_pid != getpid()) pool = nullptr;  // forked: the parent's worker threads do not exist here
    if (pool && pool_threads != threads) {
        delete pool;
        pool = nullptr;
    }
    if (!pool) {
        pool = new (std::nothrow) mv3d::HostPool(threads);
        pool_pid = getpid();
        if (!pool) return fail("mv3d_io_pack_sweeps: cannot create the copy pool");
        pool_threads = threads;
    }
    mv3d::copy_fill(*pool, dst, srcs.data(), offs.data(), bytes.data(), n);
    return 0;
}

extern "C" int mv3d_io_write_npy(const char *path, const float *data, const long long *shape, int ndim)
{
    size_t count;
    if (!path) return fail("mv3d_io_write_npy: path is NULL");
    if (check_shape("mv3d_io_write_npy", shape, ndim, &count)) return -1;
    if (count > 0 && !data) return fail("mv3d_io_write_npy: data is NULL");
    const std::string e = write_npy_file(path, data, shape, ndim, count);
    return e.empty() ? 0 : fail("mv3d_io_write_npy: %s", e.c_str());
}

extern "C" int mv3d_io_write_npy_batch(const char *const *paths, int n, const float *data, size_t stride_floats,
                                       const long long *shape, int ndim, int threads)
{
    size_t count;
    if (n < 0 || (n > 0 && !paths)) return fail("mv3d_io_write_npy_batch: bad arguments");
    if (check_shape("mv3d_io_write_npy_batch", shape, ndim, &count)) return -1;
    if (n > 0 && count > 0 && (!data || stride_floats < count)) return fail("mv3d_io_write_npy_batch: data/stride");
    const std::string e = run_parallel(n, threads, [&](int i) {
        return write_npy_file(paths[i], data + (size_t)i * stride_floats, shape, ndim, count);
    });
    return e.empty() ? 0 : fail("mv3d_io_write_npy_batch: %s", e.c_str());
}

extern "C" int mv3d_io_npy_info(const char *path, long long *shape_out, int *ndim_out, long long *data_offset_out)
{
    if (!path || !shape_out || !ndim_out) return fail("mv3d_io_npy_info: bad arguments");
    Fd f;
    f.fd = open(path, O_RDONLY);
    if (f.fd < 0) return fail("mv3d_io_npy_info: %s", errno_text("open", path).c_str());
    NpyInfo info;
    const std::string e = parse_npy(f.fd, path, &info);
    if (!e.empty()) return fail("mv3d_io_npy_info: %s", e.c_str());
    for (int i = 0; i < info.ndim; i++) shape_out[i] = info.shape[i];
    *ndim_out = info.ndim;
    if (data_offset_out) *data_offset_out = info.data_offset;
    return 0;
}

extern "C" int mv3d_io_read_npy(const char *path, float *dst, size_t capacity_floats)
{
    if (!path) return fail("mv3d_io_read_npy: path is NULL");
    Fd f;
    f.fd = open(path, O_RDONLY);
    if (f.fd < 0) return fail("mv3d_io_read_npy: %s", errno_text("open", path).c_str());
    NpyInfo info;
    std::string e = parse_npy(f.fd, path, &info);
    if (!e.empty()) return fail("mv3d_io_read_npy: %s", e.c_str());
    if (info.count > capacity_floats)
        return fail("mv3d_io_read_npy: %s holds %zu floats, the buffer %zu", path, info.count, capacity_floats);
    if (info.count > 0 && !dst) return fail("mv3d_io_read_npy: dst is NULL");
    e = read_exact(f.fd, dst, info.count * sizeof(float), (off_t)info.data_offset, path);
    return e.empty() ? 0 : fail("mv3d_io_read_npy: %s", e.c_str());
}

extern "C" int mv3d_io_write_npz_compressed(const char *path, const char *member, const float *data,
                                            const long long *shape, int ndim, int level)
{
    return mv3d_io_write_npz_compressed_batch(&path, 1, member, data, 0, shape, ndim, level, 1);
}

extern "C" int mv3d_io_write_npz_compressed_batch(const char *const *paths, int n, const char *member,
                                                  const float *data, size_t stride_floats, const long long *shape,
                                                  int ndim, int level, int threads)
{
    size_t count;
    if (n < 0 || (n > 0 && !paths) || !member || !*member) return fail("mv3d_io_write_npz_compressed: bad arguments");
    if (check_shape("mv3d_io_write_npz_compressed", shape, ndim, &count)) return -1;
    if (n > 1 && stride_floats < count) return fail("mv3d_io_write_npz_compressed: stride smaller than one array");
    if (n > 0 && count > 0 && !data) return fail("mv3d_io_write_npz_compressed: data is NULL");
    if (level <= 0 || level > 9) level = 6;
    const std::string e = run_parallel(n, threads, [&](int i) -> std::string {
        if (!paths[i]) return "NULL path";
        return write_npz_file(paths[i], member, data + (size_t)i * stride_floats, shape, ndim, count, level);
    });
    return e.empty() ? 0 : fail("mv3d_io_write_npz_compressed: %s", e.c_str());
}
