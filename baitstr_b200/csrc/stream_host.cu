// stream_host.cu -- bstr_run_files: the file loop of the select stage
// (/root/reference/src/select_STR_reads.pyx:359-370; reader = /root/reference/src/fastq.pyx:48-94).
//
// The reference parses a chunk, blocks in pool.map, prints, and only then parses the next
// chunk (pyx:362-367).  Here three kinds of host threads overlap instead:
//   reader   (1)        files in argv order, ".gz" by file-name suffix exactly as fastq.pyx:50;
//                       raw bytes go straight into page-locked batch buffers, cut at a record
//                       boundary (a whole number of 4-line groups; the host counts '\n', it
//                       never interprets FASTQ), the partial tail is carried to the next batch
//   device   (1 / GPU)  bstr lanes: submit (H2D + kernels) up to three batches per GPU, collect
//                       (D2H of the exact output) the oldest; batches are dealt to whichever
//                       GPU has a free lane -- reads are independent, there is no collective
//   writer   (caller)   writes the outputs in batch order to out_fd, "Done processing <path>"
//                       to err_fd after a file's last batch; the first failing batch in input
//                       order stops the run (everything before it has been written, as the
//                       reference prints the chunks before a bad record)
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#include <chrono>

#include "host_common.h"
#include "pinflate.h"
#include "ginflate.h"

namespace bstr {

static inline double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct PinnedBuf {
    uint8_t *p;
    size_t cap;
};

// GPU inflater of one device (csrc/ginflate.cu) + the device buffer its text lands in:
// [IN_SLACK][carry of the window before | text of this window][IN_SLACK], the batch k_fused_tile reads
struct GzDevice {
    ginflate::Device *dev = nullptr;
    size_t window = 0, range = 0;
    uint8_t *d_text = nullptr; size_t cap_text = 0;
};

constexpr int GZ_PAR = 2;          // .gz windows in flight per device: one decodes while the other uploads / chains / scans
struct StreamEngine {
    std::vector<PinnedBuf> pool;   // page-locked batch buffers, kept across calls
    std::vector<GzDevice> gz;      // GZ_PAR per entry of bstr_params.devices, created on first use
};

void stream_release(bstr_ctx *ctx)
{
    if (!ctx->engine) return;
    for (size_t d = 0; d < ctx->engine->gz.size(); d++) {
        GzDevice &g = ctx->engine->gz[d];
        if (!g.dev && !g.d_text) continue;
        cudaSetDevice(ctx->p.devices[d / GZ_PAR]);
        delete g.dev;
        cudaFree(g.d_text);
    }
    cudaSetDevice(ctx->dev);
    for (auto &b : ctx->engine->pool) if (b.p) cudaFreeHost(b.p);
    delete ctx->engine;
    ctx->engine = nullptr;
}

// Serial gzip reader (small files, BSTR_SERIAL_INFLATE): zlib's inflate() one member at a time,
// with the member chain of Python's gzip module, which the reference reads through
// (fastq.pyx:50-51): members may follow each other, zero padding after a member is skipped,
// anything else after a member -- and a file that does not start with a gzip header -- is an
// error ("Not a gzipped file").  zlib's own gzread() would accept both silently.
struct GzSerial {
    int fd = -1;
    z_stream zs;
    bool zs_live = false, in_member = false, eof = false, any_member = false;
    uint32_t crc = 0, isize = 0;
    std::vector<uint8_t> ibuf;
    const char *err = nullptr;
    bool open_fd(int f) { fd = f; ibuf.resize(1 << 20); memset(&zs, 0, sizeof(zs)); return true; }
    bool refill()
    {
        if (zs.avail_in) return true;
        for (;;) {
            ssize_t r = read(fd, ibuf.data(), ibuf.size());
            if (r < 0 && errno == EINTR) continue;
            if (r < 0) { err = strerror(errno); return false; }
            zs.next_in = ibuf.data(); zs.avail_in = (uInt)r;
            return true;                                          // avail_in == 0: end of file
        }
    }
    int get_byte()                                                // -1: end of file, -2: read error
    {
        if (!refill()) return -2;
        if (!zs.avail_in) return -1;
        zs.avail_in--;
        return *zs.next_in++;
    }
    // gzip member header as gzip.py reads it (_read_gzip_header): magic, method 8, flags, the
    // optional fields skipped; the header CRC is read and NOT checked, as there.
    bool read_header(int first)
    {
        uint8_t h[10];
        h[0] = (uint8_t)first;
        for (int i = 1; i < 10; i++) { const int c = get_byte(); if (c < 0) { if (c == -1) err = i < 2 ? "not a gzip file" : "truncated gzip header"; return false; } h[i] = (uint8_t)c; }
        if (h[0] != 0x1f || h[1] != 0x8b) { err = any_member ? "not a gzip file (bytes after the last member)" : "not a gzip file"; return false; }
        if (h[2] != 8) { err = "unknown compression method"; return false; }
        const int flg = h[3];
        auto skip = [&](int n) { while (n-- > 0) if (get_byte() < 0) { if (!err) err = "truncated gzip header"; return false; } return true; };
        auto skipz = [&]() { for (;;) { const int c = get_byte(); if (c < 0) { if (!err) err = "truncated gzip header"; return false; } if (c == 0) return true; } };
        if (flg & 4) { const int lo = get_byte(), hi = get_byte(); if (lo < 0 || hi < 0) { if (!err) err = "truncated gzip header"; return false; } if (!skip(lo | hi << 8)) return false; }
        if ((flg & 8) && !skipz()) return false;
        if ((flg & 16) && !skipz()) return false;
        if ((flg & 2) && !skip(2)) return false;
        return true;
    }
    long read_some(uint8_t *dst, size_t want)                     // bytes (0 = EOF) or -1
    {
        if (want > (1u << 30)) want = 1u << 30;
        size_t got = 0;
        while (got < want && !eof) {
            if (!in_member) {
                // between members: skip zero padding, then a member must start (or the file ends)
                int c;
                do c = get_byte(); while (c == 0 && any_member);
                if (c == -2) return -1;
                if (c == -1) { eof = true; break; }
                if (!read_header(c)) return -1;
                if (zs_live) inflateReset2(&zs, -15);
                else if (inflateInit2(&zs, -15) != Z_OK) { err = "zlib initialisation failed"; return -1; }
                zs_live = true; in_member = true; crc = 0; isize = 0;
            }
            if (!refill()) return -1;
            if (!zs.avail_in) { err = "compressed file ended before the end-of-stream marker was reached"; return -1; }
            zs.next_out = dst + got; zs.avail_out = (uInt)(want - got);
            const int r = inflate(&zs, Z_NO_FLUSH);
            const size_t now = want - zs.avail_out;
            crc = (uint32_t)crc32(crc, dst + got, (uInt)(now - got));
            isize += (uint32_t)(now - got);
            got = now;
            if (r == Z_STREAM_END) {
                uint8_t t[8];
                for (int i = 0; i < 8; i++) { const int c = get_byte(); if (c < 0) { if (!err) err = "compressed file ended before the end-of-stream marker was reached"; return -1; } t[i] = (uint8_t)c; }
                const uint32_t want_crc = t[0] | t[1] << 8 | t[2] << 16 | (uint32_t)t[3] << 24, want_len = t[4] | t[5] << 8 | t[6] << 16 | (uint32_t)t[7] << 24;
                if (want_crc != crc) { err = "CRC check failed"; return -1; }
                if (want_len != isize) { err = "incorrect length check"; return -1; }
                in_member = false; any_member = true;
                continue;
            }
            if (r == Z_OK || (r == Z_BUF_ERROR && zs.avail_in == 0)) continue;
            err = zs.msg ? zs.msg : "corrupt gzip data";
            return -1;
        }
        return (long)got;
    }
    void close_all() { if (zs_live) inflateEnd(&zs); zs_live = false; if (fd >= 0) close(fd); fd = -1; }
};

struct Source {
    int fd = -1;
    GzSerial gz;
    bool is_gz = false;
    int open_path(const char *path)
    {
        const size_t n = strlen(path);
        is_gz = n >= 3 && strcmp(path + n - 3, ".gz") == 0;          // fastq.pyx:50
        fd = open(path, O_RDONLY);
        if (fd < 0) return -1;
#ifdef POSIX_FADV_SEQUENTIAL
        posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
        if (is_gz) { gz.open_fd(fd); fd = -1; }
        return 0;
    }
    // returns bytes read (0 = EOF) or -1
    long read_some(uint8_t *dst, size_t want)
    {
        if (want > (1u << 30)) want = 1u << 30;
        if (is_gz) return gz.read_some(dst, want);
        for (;;) {
            ssize_t r = read(fd, dst, want);
            if (r < 0 && errno == EINTR) continue;
            return (long)r;
        }
    }
    const char *error() const { return is_gz ? gz.err : nullptr; }
    void close_all()
    {
        gz.close_all();
        if (fd >= 0) close(fd);
        fd = -1;
    }
};

static int write_all(int fd, const uint8_t *p, size_t n)
{
    while (n) {
        ssize_t w = write(fd, p, n);
        if (w < 0) { if (errno == EINTR) continue; return -1; }
        p += w; n -= (size_t)w;
    }
    return 0;
}

// number of '\n' in [p, p+n) (auto-vectorised byte compare)
static uint64_t count_newlines(const uint8_t *p, size_t n)
{
    uint64_t total = 0;
    while (n) {
        const size_t m = n < 4096 ? n : 4096;
        uint32_t c = 0;
        for (size_t i = 0; i < m; i++) c += p[i] == '\n';
        total += c; p += m; n -= m;
    }
    return total;
}

// Largest prefix of buf[0, have) that is a whole number of 4-line groups (0 if none).
static size_t record_cut(const uint8_t *buf, size_t have)
{
    const uint64_t lines = count_newlines(buf, have);
    uint64_t drop = lines & 3;                 // newlines after the last whole group
    if (lines < 4) return 0;
    size_t end = have;
    // walk back over the partial tail: `drop` newlines, then to just past the previous one
    const uint8_t *q = (const uint8_t *)memrchr(buf, '\n', end);
    while (drop--) { end = (size_t)(q - buf); q = (const uint8_t *)memrchr(buf, '\n', end); }
    return (size_t)(q - buf) + 1;
}

// One "*.gz" input that goes through the GPU inflater.  Its windows are separate jobs that any
// device may take; what links them lives here, in host memory: where the next run must start, the
// last 32 KiB of text and the CRC of the open member (ginflate::FileState), and the partial record
// at the end of a window, which is the head of the next window's batch.  The two ordered stages of
// a window (chain, scan) take turns by window number.
struct GzFile {
    const uint8_t *map = nullptr; size_t size = 0;
    ginflate::FileState fs;
    std::mutex mu; std::condition_variable cv;
    int chain_turn = 0, scan_turn = 0;
    std::vector<uint8_t> carry;
    bool failed = false;
    std::atomic<int> refs{0};
    ~GzFile() { if (map) munmap((void *)map, size); }
};

// One batch travelling reader -> device worker -> writer.
struct Job {
    uint64_t seq;
    int file;
    PinnedBuf buf;            // input (returned to the pool by the device worker)
    size_t off, nbytes;       // the batch is buf.p[off, off + nbytes)
    bool final, file_end, reader_error;
    // filled by the device worker
    int rc, lane;
    bstr_batch_out bo;
    char msg[512];
    // a window of a .gz file for the GPU inflater: file bytes [gz_off, gz_off + gz_nb), the first gz_own its own
    GzFile *gz; size_t gz_off, gz_nb, gz_own; bool gz_eof; int gz_w;
};

static void job_free(Job *j)
{
    if (j->gz && j->gz->refs.fetch_sub(1) == 1) delete j->gz;
    delete j;
}

struct Pipeline {
    bstr_ctx *ctx;
    const char *const *paths;
    int npaths;
    std::mutex mu;
    std::condition_variable cv_free, cv_work, cv_done, cv_lane;
    std::deque<PinnedBuf> free_bufs;
    size_t bufs_made = 0, bufs_max = 0, buf_cap = 0;
    std::deque<Job *> work;
    std::map<uint64_t, Job *> done;
    std::vector<int> lane_state;          // 0 free, 1 submitted, 2 collected (writer still reads it)
    bool reader_finished = false, abort = false;
    uint64_t n_jobs = 0;                  // valid once reader_finished
    int n_readers = 4;                    // host threads the caller allows (bstr_params.host_threads, at least 4)
    int plain_readers = 4;                // threads that read one plain file in parallel segments
    int gz_decoders = 4, gz_fillers = 2, gz_pending = 3;   // parallel inflate: decoder threads, segment producers, segments decoded ahead
    std::vector<int> gz_active;           // .gz windows running per device (mu): a device takes another one only if no device has fewer
    bool gpu_inflate = true;              // "*.gz" files of some size go through the GPU inflater (BSTR_GPU_INFLATE=0: host threads)
    uint64_t gz_gpu_min = 8u << 20;       // ... from this compressed size on
    size_t gz_window = 0, gz_range = 0;   // its geometry (BSTR_GZ_WINDOW / BSTR_GZ_RANGE, bytes)
    double t_read = 0, t_cut = 0, t_alloc = 0, t_bufwait = 0;   // reader threads, ms, summed (BSTR_TRACE)
    bool abort_seen() { std::lock_guard<std::mutex> g(mu); return abort; }
};

static PinnedBuf get_buf(Pipeline &P)
{
    std::unique_lock<std::mutex> lk(P.mu);
    for (;;) {
        if (P.abort) return PinnedBuf{nullptr, 0};
        if (!P.free_bufs.empty()) { PinnedBuf b = P.free_bufs.front(); P.free_bufs.pop_front(); return b; }
        if (P.bufs_made < P.bufs_max) {
            P.bufs_made++;
            lk.unlock();
            PinnedBuf b{nullptr, P.buf_cap};
            const double t0 = now_ms();
            if (cudaHostAlloc((void **)&b.p, b.cap, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); b.p = nullptr; b.cap = 0; }
            lk.lock();
            P.t_alloc += now_ms() - t0;
            return b;
        }
        const double t0 = now_ms();
        P.cv_free.wait(lk);
        P.t_bufwait += now_ms() - t0;
    }
}

static void put_buf(Pipeline &P, PinnedBuf b)
{
    if (!b.p) return;
    if (b.cap != P.buf_cap) { cudaFreeHost(b.p); return; }      // one-off oversize buffer
    { std::lock_guard<std::mutex> g(P.mu); P.free_bufs.push_back(b); }
    P.cv_free.notify_all();
}

static void emit(Pipeline &P, Job *j)
{
    { std::lock_guard<std::mutex> g(P.mu); j->seq = P.n_jobs++; P.work.push_back(j); }
    P.cv_work.notify_all();
}

static Job *new_job(int file, PinnedBuf b, size_t nbytes, bool final, bool file_end)
{
    Job *j = new Job();
    memset(j, 0, sizeof(*j));
    j->file = file; j->buf = b; j->nbytes = nbytes; j->final = final; j->file_end = file_end;
    j->lane = -1; j->bo.bad_record = -1;
    return j;
}

static void reader_fail(Pipeline &P, int file, PinnedBuf b, const char *what, const char *path, const char *detail = nullptr)
{
    Job *j = new_job(file, b, 0, true, true);
    j->reader_error = true; j->rc = BSTR_ERR_IO;
    snprintf(j->msg, sizeof(j->msg), "%s %s: %s", what, path, detail ? detail : errno ? strerror(errno) : "I/O error");
    emit(P, j);
}

// ---- large files: several threads produce fixed-size segments of the (inflated) file at once ----
// A segment lands HEAD bytes into its page-locked buffer, its newlines are counted by the
// thread that produced it, and the segments are stitched in file order: the running line count
// tells where the last whole record of a segment ends (ordinal of the newline a multiple of
// four), the bytes after it are carried in front of the next segment (that is what HEAD is
// for).  Batches therefore still start and end on record boundaries and the host still never
// looks at anything but '\n'.  Where the segments come from is a SegmentSource: pread() for a
// plain file, the parallel inflater (pinflate.h) for a .gz file.
struct SegmentSource {
    virtual ~SegmentSource() {}
    // next segment in file order: index, -1 after the last, -2 on error.  Thread-safe; may block.
    virtual long claim(uint64_t *off, size_t *len, bool *last) = 0;
    // produce it at dst; *have = bytes written (<= len), *newlines = '\n' among them
    virtual bool fill(long k, uint64_t off, size_t len, uint8_t *dst, size_t *have, uint64_t *newlines, pinflate::SegmentInfo *info) = 0;
    virtual int file_of(long) { return file0; }            // index (into the path list) of the file segment k belongs to
    int file0 = 0;
    virtual bool fold(const pinflate::SegmentInfo &, bool) { return true; }   // called in segment order
    virtual void cancel() {}
    virtual const char *error() { return nullptr; }
    size_t seg_bytes = 0;
};

// A run of consecutive plain files as ONE stream of segments: segment numbers continue across the
// file boundaries, so while the last segments of one shard file are still being read, other
// threads already read the first segments of the next files -- several files are in flight at
// once and the reader threads never drain at a file boundary.  (The reference opens one file
// after the other, src/select_STR_reads.pyx:359; output order only needs the stitcher and the
// writer to go in file order, not the readers.)
struct PlainSource : SegmentSource {
    struct File { int fd; uint64_t size, seg0, nseg; };
    std::vector<File> files;
    uint64_t nseg = 0;
    std::atomic<uint64_t> next{0};
    PlainSource(int first_file, size_t S) { file0 = first_file; seg_bytes = S; }
    void add(int fd, uint64_t size) { const uint64_t n = (size + seg_bytes - 1) / seg_bytes; files.push_back(File{fd, size, nseg, n}); nseg += n; }
    const File &of(long k, int *idx = nullptr) const
    {
        size_t lo = 0, hi = files.size() - 1;
        while (lo < hi) { const size_t mid = (lo + hi + 1) / 2; if (files[mid].seg0 <= (uint64_t)k) lo = mid; else hi = mid - 1; }
        if (idx) *idx = (int)lo;
        return files[lo];
    }
    int file_of(long k) override { int i; of(k, &i); return file0 + i; }
    long claim(uint64_t *off, size_t *len, bool *last) override
    {
        const uint64_t k = next.fetch_add(1);
        if (k >= nseg) return -1;
        const File &f = of((long)k);
        const uint64_t j = k - f.seg0;
        *off = j * seg_bytes;
        *len = (size_t)(f.size - *off < seg_bytes ? f.size - *off : seg_bytes);
        *last = j + 1 == f.nseg;                          // last segment of ITS file
        return (long)k;
    }
    bool fill(long k, uint64_t off, size_t len, uint8_t *dst, size_t *have, uint64_t *newlines, pinflate::SegmentInfo *) override
    {
        const int fd = of(k).fd;
        size_t got = 0;
        uint64_t nl = 0;
        while (got < len) {
            // a piece that stays in the core's cache: its newlines are counted before it is evicted,
            // which saves one pass over host memory per byte (the file path is host-memory-bound)
            const size_t want = len - got < (256u << 10) ? len - got : (256u << 10);
            ssize_t r = pread(fd, dst + got, want, (off_t)(off + got));
            if (r < 0 && errno == EINTR) continue;
            if (r < 0) return false;
            if (r == 0) break;                           // file shrank: what is there is the segment
            nl += count_newlines(dst + got, (size_t)r);
            got += (size_t)r;
        }
        *have = got;
        *newlines = nl;
        return true;
    }
    void close_all() { for (auto &f : files) if (f.fd >= 0) { close(f.fd); f.fd = -1; } }
};

struct GzipSource : SegmentSource {
    pinflate::Stream z;
    const uint8_t *map = nullptr; size_t map_size = 0;
    ~GzipSource() override { z.close(); if (map) munmap((void *)map, map_size); }
    bool open(int fd, size_t size, int threads, size_t S, int pending)
    {
        void *m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return false;
        map = (const uint8_t *)m; map_size = size; seg_bytes = S;
        madvise(m, size, MADV_SEQUENTIAL);
        return z.open(map, size, threads, (size_t)1 << 20, S, pending);
    }
    long claim(uint64_t *off, size_t *len, bool *last) override { return z.claim(off, len, last); }
    bool fill(long, uint64_t off, size_t len, uint8_t *dst, size_t *have, uint64_t *newlines, pinflate::SegmentInfo *info) override
    {
        if (!z.fill(off, len, dst, info)) return false;
        *have = len; *newlines = info->newlines;
        return true;
    }
    bool fold(const pinflate::SegmentInfo &info, bool last) override { return z.fold(info, last); }
    void cancel() override { z.cancel(); }
    const char *error() override { return z.error(); }
};

struct Segment {
    PinnedBuf buf;
    int file;
    size_t len;
    uint64_t lines;
    bool failed, last;
    pinflate::SegmentInfo info;
};

static bool read_segments_parallel(Pipeline &P, SegmentSource &src, int nthreads)
{
    const size_t HEAD = 1u << 20;
    std::map<long, Segment> segs;
    std::mutex mu;
    long frontier = 0;
    uint64_t lines = 0;
    std::vector<uint8_t> tail;
    bool stop = false;

    auto fail = [&](PinnedBuf b, const char *what, int fi) {      // mu held
        const char *detail = src.error();
        if (detail && *detail) errno = 0;
        reader_fail(P, fi, b, what, P.paths[fi], detail && *detail ? detail : nullptr);
        stop = true;
        src.cancel();
    };
    auto advance = [&]() {                               // mu held
        while (!stop) {
            auto it = segs.find(frontier);
            if (it == segs.end()) break;
            Segment g = std::move(it->second);
            segs.erase(it);
            const bool lastseg = g.last;
            const int fi = g.file;
            if (g.failed) { fail(g.buf, "read error on", fi); break; }
            if (!src.fold(g.info, lastseg)) { fail(g.buf, "read error on", fi); break; }
            uint8_t *seg = g.buf.p + HEAD;
            const uint64_t total = lines + g.lines;
            size_t cut = 0;
            if (lastseg) cut = g.len;
            else if (total - (total & 3) > lines) {      // a record ends inside this segment
                const uint8_t *q = (const uint8_t *)memrchr(seg, '\n', g.len);
                for (uint64_t d = total & 3; d; d--) q = (const uint8_t *)memrchr(seg, '\n', (size_t)(q - seg));
                cut = (size_t)(q - seg) + 1;
            }
            if (cut == 0 && !lastseg) {                  // no record boundary in here: all of it is carried
                tail.insert(tail.end(), seg, seg + g.len);
                put_buf(P, g.buf);
            } else {
                std::vector<uint8_t> carry(seg + cut, seg + g.len);
                Job *j;
                if (tail.size() <= HEAD) {
                    memcpy(seg - tail.size(), tail.data(), tail.size());
                    j = new_job(fi, g.buf, tail.size() + cut, lastseg, lastseg);
                    j->off = HEAD - tail.size();
                } else {                                 // a record longer than HEAD: one-off buffer
                    PinnedBuf big{nullptr, tail.size() + cut + 1};
                    if (big.cap > (2ull << 30) || cudaHostAlloc((void **)&big.p, big.cap, cudaHostAllocPortable) != cudaSuccess) {
                        cudaGetLastError(); errno = ENOMEM;
                        fail(g.buf, "record too large reading", fi); break;
                    }
                    memcpy(big.p, tail.data(), tail.size());
                    memcpy(big.p + tail.size(), seg, cut);
                    put_buf(P, g.buf);
                    j = new_job(fi, big, tail.size() + cut, lastseg, lastseg);
                }
                emit(P, j);
                tail.swap(carry);
            }
            lines = lastseg ? 0 : total;                 // the next segment starts the next file of the run
            frontier++;
        }
    };

    auto worker = [&]() {
        cudaSetDevice(P.ctx->dev);
        for (;;) {
            PinnedBuf b = get_buf(P);                    // buffer first: the lowest unproduced segment always has one
            {
                std::lock_guard<std::mutex> g(mu);
                if (!b.p) { if (!P.abort_seen() && !stop) { errno = ENOMEM; fail(b, "out of page-locked memory reading", src.file_of(frontier)); } stop = true; src.cancel(); return; }
                if (stop) { put_buf(P, b); return; }
            }
            uint64_t off = 0; size_t len = 0; bool last = false;
            const long k = src.claim(&off, &len, &last);
            if (k == -1) { put_buf(P, b); return; }
            if (k < 0) {
                std::lock_guard<std::mutex> g(mu);
                if (!stop && !P.abort_seen()) fail(b, "read error on", src.file_of(frontier)); else put_buf(P, b);
                return;
            }
            Segment sg;
            sg.buf = b; sg.len = 0; sg.lines = 0; sg.last = last; sg.file = src.file_of(k);
            const double t0 = now_ms();
            sg.failed = !src.fill(k, off, len, b.p + HEAD, &sg.len, &sg.lines, &sg.info);
            const double t1 = now_ms();
            { std::lock_guard<std::mutex> g(P.mu); P.t_read += t1 - t0; }
            std::lock_guard<std::mutex> g(mu);
            segs[k] = std::move(sg);
            advance();
        }
    };
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; i++) th.emplace_back(worker);
    for (auto &t : th) t.join();
    for (auto &kv : segs) if (kv.second.buf.p) put_buf(P, kv.second.buf);     // aborted run
    return !stop && !P.abort_seen();
}

static void reader_main(Pipeline *pp)
{
    Pipeline &P = *pp;
    cudaSetDevice(P.ctx->dev);
    for (int fi = 0; fi < P.npaths; fi++) {
        Source src;
        errno = 0;
        if (src.open_path(P.paths[fi]) != 0) { reader_fail(P, fi, PinnedBuf{nullptr, 0}, "cannot open", P.paths[fi]); break; }
        struct stat sb;
        auto big_plain = [&](int fd, struct stat *st) {
            return P.n_readers > 1 && P.buf_cap >= (4u << 20) && fstat(fd, st) == 0 && S_ISREG(st->st_mode) && (uint64_t)st->st_size > P.buf_cap;
        };
        if (!src.is_gz && big_plain(src.fd, &sb)) {
            // this file and the plain files that follow it in argv order (each larger than one batch)
            // go through the parallel segment reader as ONE run, so their reads overlap
            PlainSource ps(fi, P.buf_cap - (1u << 20));
            ps.add(src.fd, (uint64_t)sb.st_size);
            src.fd = -1;
            int fj = fi + 1;
            for (; fj < P.npaths && fj - fi < 64; fj++) {
                const size_t n = strlen(P.paths[fj]);
                if (n >= 3 && strcmp(P.paths[fj] + n - 3, ".gz") == 0) break;
                const int fd = open(P.paths[fj], O_RDONLY);
                struct stat sj;
                if (fd < 0 || !big_plain(fd, &sj)) { if (fd >= 0) close(fd); break; }     // handled in its turn (errors stay in file order)
#ifdef POSIX_FADV_SEQUENTIAL
                posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
                ps.add(fd, (uint64_t)sj.st_size);
            }
            const bool ok = read_segments_parallel(P, ps, P.plain_readers);
            ps.close_all();
            src.close_all();
            if (!ok) break;
            fi = fj - 1;
            continue;
        }
        if (src.is_gz && P.gpu_inflate) {
            // a regular .gz file of some size: the GPU inflater (ginflate.h).  The file is mapped and
            // cut into windows; every window is a job of its own, so the windows of ONE file are dealt
            // to all GPUs, and the inflated text never visits the host.
            const int fd = open(P.paths[fi], O_RDONLY);
            bool done = false;
            if (fd >= 0 && fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && (uint64_t)sb.st_size >= P.gz_gpu_min) {
                void *m = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
                if (m != MAP_FAILED) {
                    madvise(m, (size_t)sb.st_size, MADV_SEQUENTIAL);
                    GzFile *gz = new GzFile();
                    gz->map = (const uint8_t *)m; gz->size = (size_t)sb.st_size;
                    const size_t W = P.gz_window, LA = ginflate::LOOK_AHEAD;
                    const int nw = (int)((gz->size + W - 1) / W);
                    gz->refs = nw;
                    for (int w = 0; w < nw; w++) {
                        Job *j = new_job(fi, PinnedBuf{nullptr, 0}, 0, w == nw - 1, w == nw - 1);
                        j->gz = gz; j->gz_w = w; j->gz_off = (size_t)w * W;
                        j->gz_own = gz->size - j->gz_off < W ? gz->size - j->gz_off : W;
                        j->gz_nb = gz->size - j->gz_off < W + LA ? gz->size - j->gz_off : W + LA;
                        j->gz_eof = j->gz_off + j->gz_nb == gz->size;
                        emit(P, j);
                    }
                    done = true;
                }
            }
            if (fd >= 0) close(fd);
            if (done) { src.close_all(); continue; }
        }
        if (src.is_gz && P.n_readers > 1 && P.buf_cap >= (4u << 20) && !getenv("BSTR_SERIAL_INFLATE")) {
            // a regular .gz file of some size: inflate it on all reader threads (pinflate.h)
            const int fd = open(P.paths[fi], O_RDONLY);
            bool done = false, ok = true;
            if (fd >= 0 && fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size >= (1 << 20)) {
                GzipSource gs;
                const size_t S = P.buf_cap - (1u << 20) < ((size_t)32 << 20) ? P.buf_cap - (1u << 20) : (size_t)32 << 20;
                if (gs.open(fd, (size_t)sb.st_size, P.gz_decoders, S, P.gz_pending)) {
                    gs.file0 = fi;
                    ok = read_segments_parallel(P, gs, P.gz_fillers);
                    done = true;
                    if (getenv("BSTR_TRACE")) {
                        const pinflate::Stats zs = gs.z.stats();
                        fprintf(stderr, "[bstr] inflate %s: %.1f MB -> %.1f MB, %llu ranges (%llu decoded again), %llu members | threads: decode %.1f search %.1f redo %.1f idle %.1f ms\n",
                                P.paths[fi], zs.in_bytes / 1e6, zs.out_bytes / 1e6, (unsigned long long)zs.ranges, (unsigned long long)zs.redone,
                                (unsigned long long)zs.members, zs.decode_ms, zs.search_ms, zs.redo_ms, zs.wait_ms);
                    }
                }
            }
            if (fd >= 0) close(fd);
            if (done) {
                src.close_all();
                if (!ok) break;
                continue;
            }
        }
        PinnedBuf b = get_buf(P);
        size_t have = 0;
        bool eof = false, stop = false;
        while (!stop) {
            if (!b.p) {                                           // abort, or out of pinned memory
                bool ab;
                { std::lock_guard<std::mutex> g(P.mu); ab = P.abort; }
                if (!ab) { errno = ENOMEM; reader_fail(P, fi, b, "out of page-locked memory reading", P.paths[fi]); }
                stop = true; break;
            }
            const double tr0 = now_ms();
            while (!eof && have < b.cap) {
                errno = 0;
                long r = src.read_some(b.p + have, b.cap - have);
                if (r < 0) {
                    if (src.error()) errno = 0;
                    reader_fail(P, fi, b, "read error on", P.paths[fi], src.error());
                    b.p = nullptr; stop = true; break;
                }
                if (r == 0) eof = true;
                have += (size_t)r;
            }
            { std::lock_guard<std::mutex> g(P.mu); P.t_read += now_ms() - tr0; }
            if (stop) break;
            if (eof) { emit(P, new_job(fi, b, have, true, true)); b.p = nullptr; break; }
            const double tc0 = now_ms();
            const size_t cut = record_cut(b.p, have);
            { std::lock_guard<std::mutex> g(P.mu); P.t_cut += now_ms() - tc0; }
            if (cut == 0) {
                // not even one whole record in a full buffer: continue in a larger one-off buffer
                if (b.cap >= (2ull << 30)) {
                    Job *j = new_job(fi, b, 0, true, true);
                    j->reader_error = true; j->rc = BSTR_ERR_TOO_LARGE;
                    snprintf(j->msg, sizeof(j->msg), "%s: record larger than 2 GiB", P.paths[fi]);
                    emit(P, j); b.p = nullptr; stop = true; break;
                }
                PinnedBuf nb{nullptr, b.cap * 2 > (2ull << 30) ? (size_t)(2ull << 30) : b.cap * 2};
                if (cudaHostAlloc((void **)&nb.p, nb.cap, cudaHostAllocPortable) != cudaSuccess) {
                    cudaGetLastError(); errno = ENOMEM;
                    reader_fail(P, fi, b, "out of page-locked memory reading", P.paths[fi]); b.p = nullptr; stop = true; break;
                }
                memcpy(nb.p, b.p, have);
                put_buf(P, b);
                b = nb;
                continue;
            }
            PinnedBuf nb = get_buf(P);
            if (!nb.p) {
                bool ab;
                { std::lock_guard<std::mutex> g(P.mu); ab = P.abort; }
                if (!ab) { errno = ENOMEM; reader_fail(P, fi, b, "out of page-locked memory reading", P.paths[fi]); }
                else put_buf(P, b);
                b.p = nullptr; stop = true; break;
            }
            memcpy(nb.p, b.p + cut, have - cut);
            emit(P, new_job(fi, b, cut, false, false));
            b = nb; have -= cut;
        }
        src.close_all();
        if (b.p) put_buf(P, b);
        bool ab;
        { std::lock_guard<std::mutex> g(P.mu); ab = P.abort; }
        if (stop || ab) break;
    }
    { std::lock_guard<std::mutex> g(P.mu); P.reader_finished = true; }
    P.cv_work.notify_all(); P.cv_done.notify_all();
}

static void publish(Pipeline &P, Job *j)
{
    { std::lock_guard<std::mutex> g(P.mu); P.done[j->seq] = j; }
    P.cv_done.notify_all();
}

// One window of a .gz file on device d, lane `lane`: all stages of ginflate.h, then k_fused_tile over
// [carry of the window before | this window's text] straight from device memory.  The stages that
// must go in window order (chain, scan) wait for their turn on the file's GzFile; everything else
// -- the upload, the decoding of all ranges, the CRC kernel, the copy of the output -- overlaps with
// the other windows of the file on the other GPUs.  Fills j->rc / j->bo / j->msg.
static void gz_window_job(Pipeline &P, int d, int slot, int lane, Job *j)
{
    bstr_ctx *ctx = P.ctx;
    Lane &ln = ctx->lanes[lane];
    GzFile &gz = *j->gz;
    GzDevice &G = ctx->engine->gz[(size_t)d * GZ_PAR + (size_t)slot];
    bool turn_released = false;
    auto fail = [&](int rc, const char *msg) {
        j->rc = rc;
        snprintf(j->msg, sizeof(j->msg), "read error on %.300s: %.180s", P.paths[j->file], msg);
        { std::lock_guard<std::mutex> g(gz.mu); gz.failed = true; }
        gz.cv.notify_all();
    };
    auto wait_turn = [&](const int &turn) -> bool {
        std::unique_lock<std::mutex> lk(gz.mu);
        while (turn != j->gz_w && !gz.failed) {
            if (gz.cv.wait_for(lk, std::chrono::milliseconds(50)) == std::cv_status::timeout) {
                lk.unlock();
                const bool ab = P.abort_seen();
                lk.lock();
                if (ab) return false;
            }
        }
        return !gz.failed;
    };
    auto next_turn = [&](int &turn) { { std::lock_guard<std::mutex> g(gz.mu); turn++; } gz.cv.notify_all(); };
    if (cudaSetDevice(ln.dev) != cudaSuccess) { cudaGetLastError(); return fail(BSTR_ERR_CUDA, "cudaSetDevice failed"); }
    if (!G.dev || G.window != P.gz_window || G.range != P.gz_range) {
        delete G.dev;
        char err[256] = "";
        G.dev = ginflate::Device::create(P.gz_window, ginflate::LOOK_AHEAD, P.gz_range, err, sizeof(err));
        G.window = P.gz_window; G.range = P.gz_range;
        if (!G.dev) return fail(BSTR_ERR_CUDA, err);
    }
    const double t0 = now_ms();
    if (G.dev->decode(gz.map + j->gz_off, j->gz_nb, j->gz_own, j->gz_off, j->gz_eof, ln.st)) return fail(BSTR_ERR_IO, G.dev->error());
    const double t1 = now_ms();
    if (!wait_turn(gz.chain_turn)) return fail(BSTR_ERR_IO, "an earlier part of the file failed");
    const double t2 = now_ms();
    if (G.dev->chain(gz.fs, ln.st)) return fail(BSTR_ERR_IO, G.dev->error());
    next_turn(gz.chain_turn);
    const double t3 = now_ms();
    if (G.dev->crc_launch(ln.st)) return fail(BSTR_ERR_IO, G.dev->error());
    if (!wait_turn(gz.scan_turn)) return fail(BSTR_ERR_IO, "an earlier part of the file failed");
    const double t4 = now_ms();
    if (G.dev->crc_fold(gz.fs, ln.st)) return fail(BSTR_ERR_IO, G.dev->error());
    const uint64_t tn = G.dev->text_bytes();
    const size_t cn = gz.carry.size();
    const uint64_t total = cn + tn;
    if (total > (2ull << 30)) return fail(BSTR_ERR_TOO_LARGE, "a window of the file inflates to more than 2 GiB (set BSTR_GZ_WINDOW to a smaller value)");
    if (2 * IN_SLACK + total > G.cap_text) {
        cudaFree(G.d_text); G.d_text = nullptr; G.cap_text = 0;
        const size_t cap = (size_t)(2 * IN_SLACK + total + total / 4 + (1u << 20));
        if (cudaMalloc((void **)&G.d_text, cap) != cudaSuccess) { cudaGetLastError(); return fail(BSTR_ERR_CUDA, "out of device memory for the inflated window"); }
        G.cap_text = cap;
    }
    uint8_t *base = G.d_text + IN_SLACK;
    std::vector<uint8_t> head((size_t)(total < (128u << 10) ? total : (128u << 10)));
    bool ok = cudaMemsetAsync(G.d_text, 0, IN_SLACK, ln.st) == cudaSuccess;
    if (ok && cn) ok = cudaMemcpyAsync(base, gz.carry.data(), cn, cudaMemcpyHostToDevice, ln.st) == cudaSuccess;
    if (ok) ok = G.dev->resolve(base + cn, ln.st) == 0;
    if (ok && !head.empty()) ok = cudaMemcpyAsync(head.data(), base, head.size(), cudaMemcpyDeviceToHost, ln.st) == cudaSuccess;
    if (ok) ok = cudaStreamSynchronize(ln.st) == cudaSuccess;
    if (!ok) { cudaGetLastError(); return fail(BSTR_ERR_CUDA, G.dev->error()[0] ? G.dev->error() : "CUDA error while the window's text was put together"); }
    int rc = lane_submit_device(ctx, ln, base, total, j->final ? 1 : 0, head.data(), head.size());
    if (rc == BSTR_OK) {
        const std::function<int(uint64_t)> on_state = [&](uint64_t consumed) -> int {
            // what follows the last whole record is the head of the next window's batch
            const uint64_t rest = total - (consumed < total ? consumed : total);
            if (rest > (256u << 20)) { set_error(ctx, "a record of more than 256 MiB"); return BSTR_ERR_TOO_LARGE; }
            gz.carry.resize((size_t)rest);
            if (rest && cudaMemcpy(gz.carry.data(), base + consumed, (size_t)rest, cudaMemcpyDeviceToHost) != cudaSuccess)
                return cuda_fail(ctx, cudaGetLastError(), "carry read-back", __LINE__);
            next_turn(gz.scan_turn);
            turn_released = true;
            return BSTR_OK;
        };
        rc = lane_collect_device(ctx, ln, &j->bo, &on_state);
    }
    const double t5 = now_ms();
    if (rc != BSTR_OK) {
        j->rc = rc;
        { std::lock_guard<std::mutex> g(*(std::mutex *)ctx->lock); snprintf(j->msg, sizeof(j->msg), "%s", ctx->err); }
        if (!turn_released) { { std::lock_guard<std::mutex> g(gz.mu); gz.failed = true; } gz.cv.notify_all(); }
        return;
    }
    if (getenv("BSTR_TRACE"))
        fprintf(stderr, "[bstr] gz window %d of %s on device %d: %.1f MB -> %.1f MB | upload+decode %.1f ms, wait %.1f, chain %.1f, crc+wait %.1f, "
                "text+scan+output %.1f ms\n", j->gz_w, P.paths[j->file], ln.dev, j->gz_own / 1e6, tn / 1e6, t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4);
}

// One thread per GPU: keeps up to BSTR_LANES_PER_DEVICE batches in flight on its device.
static void device_main(Pipeline *pp, int d)
{
    Pipeline &P = *pp;
    bstr_ctx *ctx = P.ctx;
    const int ndev = ctx->p.n_devices;
    std::deque<std::pair<int, Job *>> inflight;
    // windows of .gz files run on threads of their own, at most GZ_PAR per device (gz_busy: P.mu)
    std::vector<std::thread> gz_threads;
    bool gz_busy[GZ_PAR] = {false};
    auto join_gz = [&]() { for (auto &t : gz_threads) t.join(); gz_threads.clear(); };
    for (;;) {
        Job *job = nullptr;
        int lane = -1, gz_slot = -1;
        bool quit = false, aborting = false;
        {
            std::unique_lock<std::mutex> lk(P.mu);
            for (;;) {
                if (P.abort) { aborting = true; break; }
                lane = -1;
                for (int l = d; l < ctx->n_lanes; l += ndev) if (P.lane_state[l] == 0) { lane = l; break; }
                if (lane >= 0 && !P.work.empty()) {
                    gz_slot = -1;
                    if (P.work.front()->gz) {
                        int least = P.gz_active[(size_t)d];
                        for (int o = 0; o < ndev; o++) if (P.gz_active[(size_t)o] < least) least = P.gz_active[(size_t)o];
                        if (P.gz_active[(size_t)d] == least)      // spread the windows over the GPUs before doubling up on one
                            for (int g = 0; g < GZ_PAR; g++) if (!gz_busy[g]) { gz_slot = g; break; }
                    }
                    if (!P.work.front()->gz || gz_slot >= 0) {
                        job = P.work.front(); P.work.pop_front(); P.lane_state[lane] = 1;
                        if (gz_slot >= 0) { gz_busy[gz_slot] = true; P.gz_active[(size_t)d]++; P.cv_work.notify_all(); }   // the others look at the new front
                        break;
                    }
                }
                if (!inflight.empty()) break;                       // nothing to submit: collect the oldest
                bool gz_running = false;
                for (int g = 0; g < GZ_PAR; g++) gz_running |= gz_busy[g];
                if (P.reader_finished && P.work.empty() && !gz_running) { quit = true; break; }
                P.cv_work.wait(lk);
            }
        }
        if (aborting) {
            for (auto &it : inflight) {                             // leave the device idle
                bstr_batch_out bo;
                lane_collect(ctx, ctx->lanes[it.first], &bo);
                put_buf(P, it.second->buf);
                job_free(it.second);
            }
            join_gz();
            return;
        }
        if (quit) { join_gz(); return; }
        if (job && job->gz) {                                       // a window of a .gz file: GPU inflater + scan, all on this device
            gz_threads.emplace_back([&P, ctx, d, gz_slot, lane, job, &gz_busy]() {
                gz_window_job(P, d, gz_slot, lane, job);
                const bool hold = job->rc == BSTR_OK && job->bo.out_bytes;
                job->lane = hold ? lane : -1;
                ctx->lanes[lane].busy = false;
                { std::lock_guard<std::mutex> g(P.mu); P.lane_state[lane] = hold ? 2 : 0; gz_busy[gz_slot] = false; P.gz_active[(size_t)d]--; }
                publish(P, job);
                P.cv_work.notify_all();
            });
            continue;
        }
        if (job) {
            if (job->reader_error || job->nbytes == 0) {            // nothing for the GPU
                put_buf(P, job->buf); job->buf.p = nullptr;
                { std::lock_guard<std::mutex> g(P.mu); P.lane_state[lane] = 0; }
                publish(P, job);
                continue;
            }
            const int rc = lane_submit(ctx, ctx->lanes[lane], job->buf.p + job->off, job->nbytes, job->final ? 1 : 0);
            if (rc) {
                job->rc = rc;
                { std::lock_guard<std::mutex> g(*(std::mutex *)ctx->lock); snprintf(job->msg, sizeof(job->msg), "%s", ctx->err); }
                put_buf(P, job->buf); job->buf.p = nullptr;
                { std::lock_guard<std::mutex> g(P.mu); P.lane_state[lane] = 0; }
                publish(P, job);
                continue;
            }
            inflight.push_back(std::make_pair(lane, job));
            continue;
        }
        const int cl = inflight.front().first;
        Job *j = inflight.front().second;
        inflight.pop_front();
        j->rc = lane_collect(ctx, ctx->lanes[cl], &j->bo);
        if (j->rc) { std::lock_guard<std::mutex> g(*(std::mutex *)ctx->lock); snprintf(j->msg, sizeof(j->msg), "%s", ctx->err); }
        put_buf(P, j->buf); j->buf.p = nullptr;                     // the H2D copy is complete
        const bool hold = j->rc == BSTR_OK && j->bo.out_bytes;      // the writer reads the lane's pinned output
        j->lane = hold ? cl : -1;
        { std::lock_guard<std::mutex> g(P.mu); P.lane_state[cl] = hold ? 2 : 0; }
        publish(P, j);
    }
}

}  // namespace bstr

using namespace bstr;

extern "C" int bstr_run_files(bstr_ctx *ctx, const char *const *paths, int npaths, int out_fd,
                              int err_fd)
{
    if (!ctx || (!paths && npaths)) return BSTR_ERR_PARAM;
    const double t_begin = now_ms();
    CUDA_TRY(cudaSetDevice(ctx->dev));
    int rc = lanes_create(ctx);
    if (rc) return rc;
    const double t_lanes = now_ms();
    double t_wait = 0, t_write = 0;
    if (!ctx->engine) ctx->engine = new StreamEngine();
    Pipeline P;
    P.ctx = ctx; P.paths = paths; P.npaths = npaths;
    P.buf_cap = ctx->p.batch_bytes;
    auto knob = [](const char *name, int dflt) { const char *v = getenv(name); return v && atoi(v) > 0 ? atoi(v) : dflt; };
    int host_threads = ctx->p.host_threads;
    if (host_threads <= 0) {                                   // automatic: the CPUs this process may run on
        cpu_set_t cs;
        host_threads = sched_getaffinity(0, sizeof(cs), &cs) == 0 ? CPU_COUNT(&cs) : 4;
    }
    P.n_readers = host_threads > 4 ? (host_threads > 64 ? 64 : host_threads) : 4;
    // pread + newline count is memory-bound: 8 threads feed one GPU's PCIe share; more GPUs get more
    const int want_plain = 8 * ctx->p.n_devices;
    P.plain_readers = knob("BSTR_PLAIN_READERS", P.n_readers >= want_plain ? want_plain : (P.n_readers >= 8 ? P.n_readers : 4));
    P.gz_fillers = knob("BSTR_GZ_FILLERS", P.n_readers >= 8 ? 3 : 2);        // resolve markers, CRC-32, newline count
    P.gz_decoders = knob("BSTR_GZ_DECODERS", P.n_readers - P.gz_fillers);    // together they use the host threads allowed
    P.gz_pending = knob("BSTR_GZ_PENDING", 6);
    {
        const char *gi = getenv("BSTR_GPU_INFLATE");
        P.gpu_inflate = !(gi && gi[0] == '0');
        const char *gm = getenv("BSTR_GZ_GPU_MIN"), *gw = getenv("BSTR_GZ_WINDOW"), *gr = getenv("BSTR_GZ_RANGE");
        if (gm && atoll(gm) >= 0) P.gz_gpu_min = (uint64_t)atoll(gm);
        P.gz_window = gw && atoll(gw) > 0 ? (size_t)atoll(gw) : ginflate::DEFAULT_WINDOW;
        P.gz_range = gr && atoll(gr) > 0 ? (size_t)atoll(gr) : ginflate::DEFAULT_RANGE;
        if (P.gz_window < P.gz_range) P.gz_window = P.gz_range;
        if (ctx->engine->gz.size() != (size_t)ctx->p.n_devices * GZ_PAR) ctx->engine->gz.resize((size_t)ctx->p.n_devices * GZ_PAR);
    }
    P.bufs_max = (size_t)ctx->n_lanes + 2 + (size_t)(P.plain_readers > P.gz_fillers ? P.plain_readers : P.gz_fillers);
    P.lane_state.assign((size_t)ctx->n_lanes, 0);
    P.gz_active.assign((size_t)ctx->p.n_devices, 0);
    for (auto &b : ctx->engine->pool) {
        if (b.cap == P.buf_cap && P.bufs_made < P.bufs_max) { P.free_bufs.push_back(b); P.bufs_made++; }
        else cudaFreeHost(b.p);
    }
    ctx->engine->pool.clear();

    std::thread reader(reader_main, &P);
    std::vector<std::thread> workers;
    for (int d = 0; d < ctx->p.n_devices; d++) workers.emplace_back(device_main, &P, d);

    // writer: batch order == input order (select_STR_reads.pyx:363-367)
    uint64_t next = 0, rec_base = 0;
    int result = BSTR_OK;
    char errmsg[sizeof(ctx->err)] = "";
    for (;;) {
        Job *j = nullptr;
        {
            std::unique_lock<std::mutex> lk(P.mu);
            const double tw0 = now_ms();
            for (;;) {
                auto it = P.done.find(next);
                if (it != P.done.end()) { j = it->second; P.done.erase(it); break; }
                if (P.reader_finished && next == P.n_jobs) break;
                P.cv_done.wait(lk);
            }
            t_wait += now_ms() - tw0;
        }
        if (!j) break;
        const double tx0 = now_ms();
        if (j->rc) {
            result = j->rc;
            if (j->bo.bad_record >= 0)
                snprintf(errmsg, sizeof(errmsg), "%.200s: record %llu: %.250s", paths[j->file],
                         (unsigned long long)(rec_base + (uint64_t)j->bo.bad_record), j->msg);
            else snprintf(errmsg, sizeof(errmsg), "%.500s", j->msg);
        } else if (j->bo.out_bytes && write_all(out_fd, j->bo.out, j->bo.out_bytes) != 0) {
            result = BSTR_ERR_IO;
            snprintf(errmsg, sizeof(errmsg), "write error: %s", strerror(errno));
        }
        if (j->lane >= 0) {
            { std::lock_guard<std::mutex> g(P.mu); P.lane_state[j->lane] = 0; }
            P.cv_work.notify_all();
        }
        t_write += now_ms() - tx0;
        if (result) { job_free(j); break; }
        rec_base += j->bo.n_records;
        if (j->file_end) {
            rec_base = 0;
            if (err_fd >= 0) {
                char line[4200];
                int n = snprintf(line, sizeof(line), "Done processing %s\n", paths[j->file]);   // pyx:370
                if (n > 0) write_all(err_fd, (const uint8_t *)line, (size_t)n);
            }
        }
        job_free(j);
        next++;
    }
    if (result) {
        { std::lock_guard<std::mutex> g(P.mu); P.abort = true; }
        P.cv_work.notify_all(); P.cv_free.notify_all();
    }
    reader.join();
    for (auto &w : workers) w.join();
    for (Job *j : P.work) { if (j->buf.p) put_buf(P, j->buf); job_free(j); }
    for (auto &kv : P.done) { if (kv.second->buf.p) put_buf(P, kv.second->buf); job_free(kv.second); }
    for (auto &b : P.free_bufs) ctx->engine->pool.push_back(b);
    for (int l = 0; l < ctx->n_lanes; l++) ctx->lanes[l].busy = false;
    CUDA_TRY(cudaSetDevice(ctx->dev));
    { std::lock_guard<std::mutex> g(*(std::mutex *)ctx->lock); ctx->stats.wall_ms += now_ms() - t_begin; }
    if (getenv("BSTR_TRACE"))
        fprintf(stderr, "[bstr] run_files %.1f ms: lanes %.1f | reader: read %.1f cut %.1f pinned-alloc %.1f buffer-wait %.1f | "
                "writer: wait %.1f write %.1f | device: h2d %.1f kernels %.1f d2h %.1f (summed over %llu batches)\n",
                now_ms() - t_begin, t_lanes - t_begin, P.t_read, P.t_cut, P.t_alloc, P.t_bufwait, t_wait, t_write,
                ctx->stats.h2d_ms, ctx->stats.kernel_ms, ctx->stats.d2h_ms, (unsigned long long)ctx->stats.batches);
    if (result) set_error(ctx, "%s", errmsg);
    return result;
}

// Host-only hook (no GPU needed): inflate one .gz file with the parallel inflater that
// bstr_run_files uses and write the bytes to out_fd.  Exists so that the ingest side can be
// tested on its own against gzip/zlib.
extern "C" int bstr_gunzip(const char *path, int out_fd, int threads, uint64_t range_bytes, uint64_t seg_bytes,
                           uint64_t *stats6, char *err, size_t errlen)
{
    auto say = [&](const char *m) { if (err && errlen) snprintf(err, errlen, "%s", m); };
    if (threads <= 0) {                                   // the serial reader of small files (GzSerial)
        Source src;
        if (src.open_path(path) != 0) { say(strerror(errno)); return BSTR_ERR_IO; }
        src.is_gz = true;
        if (src.fd >= 0) { src.gz.open_fd(src.fd); src.fd = -1; }
        std::vector<uint8_t> buf(1 << 20);
        uint64_t total = 0;
        int rc = 0;
        for (;;) {
            const long r = src.read_some(buf.data(), buf.size());
            if (r < 0) { say(src.error() ? src.error() : "read error"); rc = BSTR_ERR_IO; break; }
            if (r == 0) break;
            total += (uint64_t)r;
            if (out_fd >= 0 && write_all(out_fd, buf.data(), (size_t)r) != 0) { say(strerror(errno)); rc = BSTR_ERR_IO; break; }
        }
        src.close_all();
        struct stat sb0;
        if (stats6) { memset(stats6, 0, 6 * sizeof(uint64_t)); stats6[0] = stat(path, &sb0) == 0 ? (uint64_t)sb0.st_size : 0; stats6[1] = total; }
        return rc;
    }
    const int fd = open(path, O_RDONLY);
    struct stat sb;
    if (fd < 0 || fstat(fd, &sb) != 0) { say(strerror(errno)); if (fd >= 0) close(fd); return BSTR_ERR_IO; }
    if (sb.st_size == 0) { close(fd); say("not a gzip file"); return BSTR_ERR_IO; }
    void *m = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (m == MAP_FAILED) { say(strerror(errno)); return BSTR_ERR_IO; }
    int rc = 0;
    {
        pinflate::Stream z;
        if (!z.open((const uint8_t *)m, (size_t)sb.st_size, threads, (size_t)range_bytes, (size_t)seg_bytes, 3)) { say(z.error()); rc = BSTR_ERR_IO; }
        std::mutex mu;
        std::map<long, std::pair<std::vector<uint8_t>, std::pair<pinflate::SegmentInfo, bool>>> ready;
        long next = 0;
        auto consumer = [&]() {
            for (;;) {
                uint64_t off = 0; size_t len = 0; bool last = false;
                const long k = z.claim(&off, &len, &last);
                if (k == -1) return;
                std::vector<uint8_t> buf(k >= 0 ? len : 0);
                pinflate::SegmentInfo info;
                const bool ok = k >= 0 && z.fill(off, len, buf.data(), &info);
                std::lock_guard<std::mutex> g(mu);
                if (!ok) { if (!rc) { rc = BSTR_ERR_IO; say(z.error()); } z.cancel(); return; }
                ready[k] = std::make_pair(std::move(buf), std::make_pair(std::move(info), last));
                for (auto it = ready.find(next); it != ready.end() && !rc; it = ready.find(next)) {
                    if (!z.fold(it->second.second.first, it->second.second.second)) { rc = BSTR_ERR_IO; say(z.error()); z.cancel(); }
                    else if (out_fd >= 0 && write_all(out_fd, it->second.first.data(), it->second.first.size()) != 0) { rc = BSTR_ERR_IO; say(strerror(errno)); z.cancel(); }
                    ready.erase(it);
                    next++;
                }
                if (rc) return;
            }
        };
        if (!rc) {
            std::vector<std::thread> th;
            for (int i = 0; i < (threads > 3 ? 3 : 2); i++) th.emplace_back(consumer);
            for (auto &t : th) t.join();
        }
        if (stats6) {
            const pinflate::Stats zs = z.stats();
            stats6[0] = zs.in_bytes; stats6[1] = zs.out_bytes; stats6[2] = zs.ranges; stats6[3] = zs.accepted; stats6[4] = zs.redone; stats6[5] = zs.members;
        }
        z.close();
    }
    munmap(m, (size_t)sb.st_size);
    return rc;
}
