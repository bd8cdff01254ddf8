// raw volume files <-> device memory without a pageable host copy of the volume.
// The reference reads .mhd/.raw through SimpleITK into a NumPy array (read/mhd.py:157-196) and writes
// array.npy / MetaImage files from host arrays (structure/image.py:709-735, structure/rigid.py:697-716); with the
// volume resident in HBM the file payload moves through page-locked staging buffers: four worker threads each own two 8 MiB buffers and a stream,
// pread() of a worker's next chunk overlaps the cudaMemcpyAsync of its previous one (and the reverse for writes).

#include <fcntl.h>
#include <unistd.h>
#include <cerrno>
#include <thread>

namespace {

constexpr size_t kIoChunk = (size_t)8 << 20;
constexpr int kIoThreads = 4;          // one pread / pwrite stream of the page cache runs at ~5 GB/s; four keep up with PCIe better

struct IoLane {                        // one worker: two staging buffers, its own stream
    void *stage[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t s = nullptr;
};
struct IoCtx {
    IoLane lane[kIoThreads];
    bool init = false;
};
IoCtx g_io[kMaxDevices];
std::mutex g_io_mu;

int io_ctx(int device, IoCtx **out)
{
    IoCtx &c = g_io[device];
    if (!c.init) {
        for (int t = 0; t < kIoThreads; ++t) {
            for (int i = 0; i < 2; ++i) {
                CUDA_TRY(cudaHostAlloc(&c.lane[t].stage[i], kIoChunk, cudaHostAllocDefault));
                CUDA_TRY(cudaEventCreateWithFlags(&c.lane[t].ev[i], cudaEventDisableTiming));
            }
            CUDA_TRY(cudaStreamCreateWithFlags(&c.lane[t].s, cudaStreamNonBlocking));
        }
        c.init = true;
    }
    *out = &c;
    return MIA_OK;
}

// Worker t moves chunks t, t + T, t + 2T, ... of the payload.  Returns 0 or an error code (message in `err`).
int io_read_lane(int fd, IoLane &L, int t, int device, unsigned long long offset, unsigned long long bytes, char *d_dst,
                 std::string &err)
{
    if (cudaSetDevice(device) != cudaSuccess) { err = "cannot select device"; return MIA_ERR_CUDA; }
    const unsigned long long nchunks = (bytes + kIoChunk - 1) / kIoChunk;
    int n = 0, rc = MIA_OK;
    for (unsigned long long ch = (unsigned long long)t; ch < nchunks && rc == MIA_OK; ch += kIoThreads, ++n) {
        const int b = n & 1;
        const unsigned long long at = ch * kIoChunk;
        const size_t want = (size_t)((bytes - at) < kIoChunk ? (bytes - at) : kIoChunk);
        if (n >= 2 && cudaEventSynchronize(L.ev[b]) != cudaSuccess) { err = "event wait failed"; rc = MIA_ERR_CUDA; break; }
        size_t got = 0;
        while (got < want) {
            const ssize_t r = pread(fd, (char *)L.stage[b] + got, want - got, (off_t)(offset + at + got));
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) break;
            got += (size_t)r;
        }
        if (got != want) { err = "file is shorter than offset + bytes"; rc = MIA_ERR_ARG; break; }
        if (cudaMemcpyAsync(d_dst + at, L.stage[b], want, cudaMemcpyHostToDevice, L.s) != cudaSuccess ||
            cudaEventRecord(L.ev[b], L.s) != cudaSuccess) { err = "H2D copy failed"; rc = MIA_ERR_CUDA; }
    }
    if (cudaStreamSynchronize(L.s) != cudaSuccess && rc == MIA_OK) { err = "stream synchronize failed"; rc = MIA_ERR_CUDA; }
    return rc;
}

int io_write_lane(int fd, IoLane &L, int t, int device, unsigned long long offset, unsigned long long bytes,
                  const char *d_src, std::string &err)
{
    if (cudaSetDevice(device) != cudaSuccess) { err = "cannot select device"; return MIA_ERR_CUDA; }
    const unsigned long long nchunks = (bytes + kIoChunk - 1) / kIoChunk;
    auto flush = [&](int b, unsigned long long at, size_t len) {
        if (cudaEventSynchronize(L.ev[b]) != cudaSuccess) { err = "event wait failed"; return (int)MIA_ERR_CUDA; }
        size_t put = 0;
        while (put < len) {
            const ssize_t r = pwrite(fd, (const char *)L.stage[b] + put, len - put, (off_t)(offset + at + put));
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) { err = std::string("write failed: ") + strerror(errno); return (int)MIA_ERR_ARG; }
            put += (size_t)r;
        }
        return (int)MIA_OK;
    };
    int n = 0, rc = MIA_OK;
    unsigned long long prev_at = 0;
    size_t prev_len = 0;
    // chunk n of this lane is copied to stage[n & 1] while its chunk n - 1 is written out
    for (unsigned long long ch = (unsigned long long)t; ch < nchunks && rc == MIA_OK; ch += kIoThreads, ++n) {
        const int b = n & 1;
        const unsigned long long at = ch * kIoChunk;
        const size_t want = (size_t)((bytes - at) < kIoChunk ? (bytes - at) : kIoChunk);
        if (cudaMemcpyAsync(L.stage[b], d_src + at, want, cudaMemcpyDeviceToHost, L.s) != cudaSuccess ||
            cudaEventRecord(L.ev[b], L.s) != cudaSuccess) { err = "D2H copy failed"; rc = MIA_ERR_CUDA; break; }
        if (n >= 1) rc = flush(b ^ 1, prev_at, prev_len);
        prev_at = at;
        prev_len = want;
    }
    if (rc == MIA_OK && n >= 1) rc = flush((n - 1) & 1, prev_at, prev_len);
    cudaStreamSynchronize(L.s);
    return rc;
}

template <class Fn>
int io_run(const char *who, int device, Fn lane_fn)
{
    std::lock_guard<std::mutex> lock(g_io_mu);
    int prev = 0;
    if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess)
        return fail(MIA_ERR_CUDA, "%s: cannot select device %d", who, device);
    IoCtx *c = nullptr;
    int rc = io_ctx(device, &c);
    if (rc == MIA_OK) {
        int rcs[kIoThreads];
        std::string errs[kIoThreads];
        std::thread th[kIoThreads];
        for (int t = 1; t < kIoThreads; ++t) th[t] = std::thread([&, t] { rcs[t] = lane_fn(c->lane[t], t, errs[t]); });
        rcs[0] = lane_fn(c->lane[0], 0, errs[0]);
        for (int t = 1; t < kIoThreads; ++t) th[t].join();
        for (int t = 0; t < kIoThreads && rc == MIA_OK; ++t)
            if (rcs[t] != MIA_OK) rc = fail(rcs[t], "%s: %s", who, errs[t].c_str());
    }
    cudaSetDevice(prev);
    return rc;
}

}  // namespace

extern "C" {

int mia_read_file_to_device(const char *path, unsigned long long offset, unsigned long long bytes, void *d_dst, int device)
{
    if (!path || (!d_dst && bytes)) return fail(MIA_ERR_ARG, "mia_read_file_to_device: null pointer");
    if (device < 0 || device >= kMaxDevices) return fail(MIA_ERR_ARG, "mia_read_file_to_device: bad device");
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(MIA_ERR_ARG, "mia_read_file_to_device: cannot open %s: %s", path, strerror(errno));
    const int rc = io_run("mia_read_file_to_device", device, [&](IoLane &L, int t, std::string &err) {
        return io_read_lane(fd, L, t, device, offset, bytes, static_cast<char *>(d_dst), err);
    });
    close(fd);
    return rc;
}

int mia_write_file_from_device(const char *path, unsigned long long offset, const void *d_src, unsigned long long bytes,
                               int truncate, int device)
{
    if (!path || (!d_src && bytes)) return fail(MIA_ERR_ARG, "mia_write_file_from_device: null pointer");
    if (device < 0 || device >= kMaxDevices) return fail(MIA_ERR_ARG, "mia_write_file_from_device: bad device");
    const int fd = open(path, O_WRONLY | O_CREAT | (truncate ? O_TRUNC : 0), 0644);
    if (fd < 0) return fail(MIA_ERR_ARG, "mia_write_file_from_device: cannot open %s: %s", path, strerror(errno));
    const int rc = io_run("mia_write_file_from_device", device, [&](IoLane &L, int t, std::string &err) {
        return io_write_lane(fd, L, t, device, offset, bytes, static_cast<const char *>(d_src), err);
    });
    close(fd);
    return rc;
}

}  // extern "C"
