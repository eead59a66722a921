// ginflate.cu -- gzip inflate on the GPU, host side (see ginflate.h for the stages).
#include "ginflate.h"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>                      // crc32_combine only

#include <vector>

#include "ginflate_dev.cuh"

namespace ginflate {

struct DeviceImpl {
    size_t window_bytes = 0, look_ahead = 0, range_bytes = 0;
    uint32_t sym_cap = 0;                       // symbols per range slot
    int max_ranges = 0;
    // device
    uint8_t *d_comp = nullptr;   size_t cap_comp = 0;
    uint16_t *d_arena = nullptr; size_t cap_arena = 0;          // symbols
    RangeDesc *d_desc = nullptr; RangeMeta *d_meta = nullptr;
    RunRef *d_runs = nullptr;    uint8_t *d_hist = nullptr;     // (max_ranges + 2) x 32 KiB
    CrcChunk *d_chunks = nullptr; uint32_t *d_crc = nullptr; size_t cap_chunks = 0;
    // pinned host mirrors
    RangeDesc *h_desc = nullptr; RangeMeta *h_meta = nullptr; RunRef *h_runs = nullptr;
    CrcChunk *h_chunks = nullptr; uint32_t *h_crc = nullptr;
    // the window in flight
    size_t nbytes = 0, own_bytes = 0; uint64_t file_off = 0; bool at_eof = false;
    int nranges = 0, nruns = 0;
    uint64_t text = 0;
    struct End { uint64_t off; uint32_t crc, isize; };          // off: text offset in the window where a member ends
    std::vector<End> ends;
    struct Piece { uint32_t first_chunk, nchunks; uint64_t len; int end; };   // end: index into ends, -1 = the member goes on
    std::vector<Piece> pieces;
    size_t nchunks = 0;
    bool stream_end = false;
    char err[256] = {0};
    size_t dev_bytes = 0;

    int fail(const char *fmt, ...)
    {
        va_list ap; va_start(ap, fmt); vsnprintf(err, sizeof(err), fmt, ap); va_end(ap);
        return -1;
    }
    int cuda(cudaError_t e, const char *what) { return e == cudaSuccess ? 0 : fail("CUDA error in the GPU inflater (%s): %s", what, cudaGetErrorString(e)); }
};

#define GI_TRY(call) do { if (p->cuda((call), #call)) return -1; } while (0)

Device *Device::create(size_t window_bytes, size_t look_ahead, size_t range_bytes, char *err, size_t errlen)
{
    Device *d = new Device();
    DeviceImpl *p = d->p = new DeviceImpl();
    auto bail = [&](const char *what, cudaError_t e) -> Device * {
        if (err && errlen) snprintf(err, errlen, "GPU inflater: %s: %s", what, cudaGetErrorString(e));
        delete d;
        return nullptr;
    };
    if (range_bytes < 4096) range_bytes = 4096;
    if (window_bytes < range_bytes) window_bytes = range_bytes;
    p->window_bytes = window_bytes; p->look_ahead = look_ahead; p->range_bytes = range_bytes;
    p->max_ranges = (int)((window_bytes + range_bytes - 1) / range_bytes);
    p->sym_cap = (uint32_t)(range_bytes * 10 + (256u << 10));
    cudaError_t e;
    p->cap_comp = window_bytes + look_ahead + 64;
    if ((e = cudaMalloc((void **)&p->d_comp, p->cap_comp)) != cudaSuccess) return bail("window buffer", e);
    p->cap_arena = (size_t)p->max_ranges * p->sym_cap;
    if ((e = cudaMalloc((void **)&p->d_arena, p->cap_arena * sizeof(uint16_t))) != cudaSuccess) return bail("symbol arena", e);
    const size_t nr = (size_t)p->max_ranges + 2;
    if ((e = cudaMalloc((void **)&p->d_desc, nr * sizeof(RangeDesc))) != cudaSuccess) return bail("range table", e);
    if ((e = cudaMalloc((void **)&p->d_meta, nr * sizeof(RangeMeta))) != cudaSuccess) return bail("range table", e);
    if ((e = cudaMalloc((void **)&p->d_runs, nr * sizeof(RunRef))) != cudaSuccess) return bail("run table", e);
    if ((e = cudaMalloc((void **)&p->d_hist, nr * 32768)) != cudaSuccess) return bail("history", e);
    if ((e = cudaHostAlloc((void **)&p->h_desc, nr * sizeof(RangeDesc), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    if ((e = cudaHostAlloc((void **)&p->h_meta, nr * sizeof(RangeMeta), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    if ((e = cudaHostAlloc((void **)&p->h_runs, nr * sizeof(RunRef), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    p->dev_bytes = p->cap_comp + p->cap_arena * 2 + nr * (sizeof(RangeDesc) + sizeof(RangeMeta) + sizeof(RunRef) + 32768);
    e = cudaFuncSetAttribute(k_decode_ranges, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(WARPS_PER_CTA * sizeof(WarpTables)));
    if (e != cudaSuccess) return bail("shared memory of the decode kernel", e);
    return d;
}

Device::~Device()
{
    if (!p) return;
    cudaFree(p->d_comp); cudaFree(p->d_arena); cudaFree(p->d_desc); cudaFree(p->d_meta); cudaFree(p->d_runs); cudaFree(p->d_hist);
    cudaFree(p->d_chunks); cudaFree(p->d_crc);
    if (p->h_desc) cudaFreeHost(p->h_desc);
    if (p->h_meta) cudaFreeHost(p->h_meta);
    if (p->h_runs) cudaFreeHost(p->h_runs);
    if (p->h_chunks) cudaFreeHost(p->h_chunks);
    if (p->h_crc) cudaFreeHost(p->h_crc);
    delete p;
}

const char *Device::error() const { return p->err; }
uint64_t Device::text_bytes() const { return p->text; }
size_t Device::device_bytes() const { return p->dev_bytes; }

static int launch_decode(DeviceImpl *p, int n, cudaStream_t st)
{
    const int ctas = (n + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    k_decode_ranges<<<ctas, 32 * WARPS_PER_CTA, WARPS_PER_CTA * sizeof(WarpTables), st>>>(p->d_comp, p->nbytes, p->at_eof ? 1 : 0, p->d_desc, n,
                                                                                         p->d_arena, p->d_meta);
    return p->cuda(cudaGetLastError(), "k_decode_ranges");
}

int Device::decode(const uint8_t *comp, size_t nbytes, size_t own_bytes, uint64_t file_off, bool at_eof, cudaStream_t st)
{
    if (nbytes + 64 > p->cap_comp || own_bytes > p->window_bytes || own_bytes > nbytes) return p->fail("GPU inflater: window larger than the working set");
    p->nbytes = nbytes; p->own_bytes = own_bytes; p->file_off = file_off; p->at_eof = at_eof;
    p->err[0] = 0;
    GI_TRY(cudaMemcpyAsync(p->d_comp, comp, nbytes, cudaMemcpyHostToDevice, st));
    GI_TRY(cudaMemsetAsync(p->d_comp + nbytes, 0, 64, st));
    const size_t R = p->range_bytes;
    p->nranges = (int)((own_bytes + R - 1) / R);
    if (p->nranges == 0) p->nranges = 1;
    for (int r = 0; r < p->nranges; r++) {
        RangeDesc &d = p->h_desc[r];
        const uint64_t lo = (uint64_t)r * R, hi = (uint64_t)(r + 1) * R < own_bytes ? (uint64_t)(r + 1) * R : own_bytes;
        d.from_bit = lo * 8; d.stop_bit = hi * 8;
        d.sym_off = (uint64_t)r * p->sym_cap; d.sym_cap = p->sym_cap;
        d.mode = (r == 0 && file_off == 0) ? 2 : 0;
    }
    GI_TRY(cudaMemcpyAsync(p->d_desc, p->h_desc, (size_t)p->nranges * sizeof(RangeDesc), cudaMemcpyHostToDevice, st));
    if (launch_decode(p, p->nranges, st)) return -1;
    GI_TRY(cudaMemcpyAsync(p->h_meta, p->d_meta, (size_t)p->nranges * sizeof(RangeMeta), cudaMemcpyDeviceToHost, st));
    GI_TRY(cudaStreamSynchronize(st));
    return 0;
}

static const char *status_text(int s)
{
    switch (s) {
        case D_ERROR: return "invalid deflate data (or the compressed file is truncated)";
        case D_OVERFLOW: return "a range inflates to more than the GPU inflater's symbol buffer holds";
        case D_UNSUPPORTED: return "more gzip members per range than the GPU inflater tracks, or a block longer than its look-ahead";
        default: return "invalid deflate data";
    }
}

int Device::chain(FileState &fs, cudaStream_t st)
{
    const uint64_t base_bit = p->file_off * 8;
    if (fs.next_bit < base_bit && p->file_off != 0) return p->fail("GPU inflater: windows out of order");
    uint64_t expected = p->file_off == 0 ? p->h_meta[0].start_bit : fs.next_bit - base_bit;
    p->nruns = 0; p->text = 0; p->ends.clear(); p->stream_end = false;
    const uint64_t own_end = (uint64_t)p->own_bytes * 8;
    fs.ranges += (uint64_t)p->nranges;
    // a slot beyond the ranges for re-decoded runs' metadata
    RangeMeta &redo = p->h_meta[p->max_ranges + 1];
    for (int r = 0; r < p->nranges && !p->stream_end; r++) {
        const RangeDesc &d = p->h_desc[r];
        if (expected >= d.stop_bit) continue;              // the run before went straight through this range
        const RangeMeta *m = &p->h_meta[r];
        if (!(m->status == D_OK && m->start_bit == expected)) {
            if (p->file_off == 0 && r == 0) return p->fail("%s", m->status == D_ERROR ? "not a gzip file, or invalid deflate data" : status_text(m->status));
            // nothing usable starts at `expected`: decode from the known position (one warp)
            RangeDesc x = d;
            x.from_bit = expected; x.mode = 1;
            GI_TRY(cudaMemcpyAsync(p->d_desc + p->max_ranges + 1, &x, sizeof(x), cudaMemcpyHostToDevice, st));
            k_decode_ranges<<<1, 32 * WARPS_PER_CTA, WARPS_PER_CTA * sizeof(WarpTables), st>>>(p->d_comp, p->nbytes, p->at_eof ? 1 : 0,
                                                                                             p->d_desc + p->max_ranges + 1, 1, p->d_arena,
                                                                                             p->d_meta + p->max_ranges + 1);
            GI_TRY(cudaGetLastError());
            GI_TRY(cudaMemcpyAsync(&redo, p->d_meta + p->max_ranges + 1, sizeof(RangeMeta), cudaMemcpyDeviceToHost, st));
            GI_TRY(cudaStreamSynchronize(st));
            fs.redone++;
            if (redo.status != D_OK) return p->fail("%s", status_text(redo.status));
            p->h_meta[r] = redo;
            m = &p->h_meta[r];
        }
        RunRef &run = p->h_runs[p->nruns++];
        run.sym_off = d.sym_off; run.out_off = p->text; run.n = m->n_sym; run.pad = (uint32_t)r;
        for (int k = 0; k < m->n_ends; k++) p->ends.push_back(DeviceImpl::End{p->text + m->end_off[k], m->end_crc[k], m->end_isize[k]});
        p->text += m->n_sym;
        expected = m->end_bit;
        if (m->stream_end) p->stream_end = true;
    }
    const bool last_window = p->at_eof && p->own_bytes == p->nbytes;     // at_eof alone: only the look-ahead reaches the end of the file
    if (last_window && !p->stream_end) return p->fail("unexpected end of compressed data");
    if (!last_window && !p->stream_end && expected < own_end) return p->fail("GPU inflater: the chain of runs stops inside the window");
    fs.next_bit = base_bit + expected;
    fs.ended = p->stream_end;
    // history through the runs: hist[0] = before the window, hist[nruns] = after it
    GI_TRY(cudaMemcpyAsync(p->d_hist, fs.hist, 32768, cudaMemcpyHostToDevice, st));
    if (p->nruns) {
        GI_TRY(cudaMemcpyAsync(p->d_runs, p->h_runs, (size_t)p->nruns * sizeof(RunRef), cudaMemcpyHostToDevice, st));
        k_propagate<<<1, 1024, 0, st>>>(p->d_arena, p->d_runs, p->nruns, p->d_hist);
        GI_TRY(cudaGetLastError());
    }
    GI_TRY(cudaMemcpyAsync(fs.hist, p->d_hist + (size_t)p->nruns * 32768, 32768, cudaMemcpyDeviceToHost, st));
    GI_TRY(cudaStreamSynchronize(st));
    fs.out_bytes += p->text;
    return 0;
}

int Device::crc_launch(cudaStream_t st)
{
    // pieces = the window's text cut at the member ends; chunks = pieces cut at run boundaries and every 128 KiB
    const uint32_t CH = 128u << 10;
    std::vector<CrcChunk> chunks;
    p->pieces.clear();
    size_t e = 0;
    DeviceImpl::Piece cur{0, 0, 0, -1};
    for (int k = 0; k < p->nruns; k++) {
        const RunRef &run = p->h_runs[k];
        uint64_t at = 0;                                   // symbols of this run already assigned
        while (at < run.n || (e < p->ends.size() && p->ends[e].off == run.out_off + at)) {
            if (e < p->ends.size() && p->ends[e].off == run.out_off + at) {          // a member ends here
                cur.end = (int)e++;
                p->pieces.push_back(cur);
                cur = DeviceImpl::Piece{(uint32_t)chunks.size(), 0, 0, -1};
                continue;
            }
            uint64_t upto = run.n;
            if (e < p->ends.size() && p->ends[e].off < run.out_off + upto) upto = p->ends[e].off - run.out_off;
            if (upto - at > CH) upto = at + CH;
            chunks.push_back(CrcChunk{run.sym_off + at, (uint32_t)(upto - at), (uint32_t)k});
            cur.nchunks++; cur.len += upto - at;
            at = upto;
        }
    }
    if (cur.nchunks || cur.len) p->pieces.push_back(cur);
    p->nchunks = chunks.size();
    if (p->nchunks > p->cap_chunks) {
        cudaFree(p->d_chunks); cudaFree(p->d_crc);
        if (p->h_chunks) cudaFreeHost(p->h_chunks);
        if (p->h_crc) cudaFreeHost(p->h_crc);
        p->d_chunks = nullptr; p->d_crc = nullptr; p->h_chunks = nullptr; p->h_crc = nullptr;
        const size_t cap = p->nchunks + p->nchunks / 2 + 1024;
        GI_TRY(cudaMalloc((void **)&p->d_chunks, cap * sizeof(CrcChunk)));
        GI_TRY(cudaMalloc((void **)&p->d_crc, cap * sizeof(uint32_t)));
        GI_TRY(cudaHostAlloc((void **)&p->h_chunks, cap * sizeof(CrcChunk), cudaHostAllocPortable));
        GI_TRY(cudaHostAlloc((void **)&p->h_crc, cap * sizeof(uint32_t), cudaHostAllocPortable));
        p->cap_chunks = cap;
    }
    if (!p->nchunks) return 0;
    memcpy(p->h_chunks, chunks.data(), p->nchunks * sizeof(CrcChunk));
    GI_TRY(cudaMemcpyAsync(p->d_chunks, p->h_chunks, p->nchunks * sizeof(CrcChunk), cudaMemcpyHostToDevice, st));
    k_crc_chunks<<<(unsigned)((p->nchunks + 127) / 128), 128, 0, st>>>(p->d_arena, p->d_chunks, (int)p->nchunks, p->d_hist, p->d_crc);
    GI_TRY(cudaGetLastError());
    GI_TRY(cudaMemcpyAsync(p->h_crc, p->d_crc, p->nchunks * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    return 0;
}

int Device::crc_fold(FileState &fs, cudaStream_t st)
{
    GI_TRY(cudaStreamSynchronize(st));
    for (const DeviceImpl::Piece &pc : p->pieces) {
        for (uint32_t c = 0; c < pc.nchunks; c++) {
            const uint32_t n = p->h_chunks[pc.first_chunk + c].n;
            fs.crc = fs.len == 0 ? p->h_crc[pc.first_chunk + c] : (uint32_t)crc32_combine(fs.crc, p->h_crc[pc.first_chunk + c], (z_off_t)n);
            fs.len += n;
        }
        if (pc.end >= 0) {
            const DeviceImpl::End &e = p->ends[(size_t)pc.end];
            if (fs.len == 0) fs.crc = 0;
            if (e.crc != fs.crc) return p->fail("CRC check failed in a gzip member (stored %08x, computed %08x)", e.crc, fs.crc);
            if (e.isize != (uint32_t)fs.len) return p->fail("incorrect length check in a gzip member");
            fs.crc = 0; fs.len = 0; fs.members++;
        }
    }
    return 0;
}

int Device::resolve(uint8_t *d_text, cudaStream_t st)
{
    if (!p->nruns || !p->text) return 0;
    const unsigned gy = (unsigned)(p->nruns < 65535 ? p->nruns : 65535);
    k_resolve<<<dim3(64, gy), 256, 0, st>>>(p->d_arena, p->d_runs, p->nruns, p->d_hist, d_text);
    return p->cuda(cudaGetLastError(), "k_resolve");
}

}  // namespace ginflate

// ---- host-only-interface test hook: inflate one .gz file on one GPU, bytes to out_fd -----------------
#include <errno.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/baitstr_select.h"

extern "C" int bstr_gunzip_gpu(const char *path, int out_fd, int device, uint64_t window_bytes, uint64_t range_bytes,
                               uint64_t *stats6, char *err, size_t errlen)
{
    auto say = [&](const char *m) { if (err && errlen) snprintf(err, errlen, "%s", m); };
    if (!path) return BSTR_ERR_PARAM;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); say("no such CUDA device (the GPU inflater has no CPU fallback)"); return BSTR_ERR_NO_DEVICE; }
    const int fd = open(path, O_RDONLY);
    struct stat sb;
    if (fd < 0 || fstat(fd, &sb) != 0) { say(strerror(errno)); if (fd >= 0) close(fd); return BSTR_ERR_IO; }
    if (sb.st_size == 0) { close(fd); say("not a gzip file"); return BSTR_ERR_IO; }
    const size_t size = (size_t)sb.st_size;
    void *map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) { say(strerror(errno)); return BSTR_ERR_IO; }
    const size_t W = window_bytes ? (size_t)window_bytes : (size_t)64 << 20, LA = (size_t)4 << 20;
    ginflate::Device *dev = ginflate::Device::create(W, LA, range_bytes ? (size_t)range_bytes : (size_t)64 << 10, err, errlen);
    if (!dev) { munmap(map, size); return BSTR_ERR_CUDA; }
    cudaStream_t st = nullptr;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    ginflate::FileState *fs = new ginflate::FileState();
    uint8_t *d_text = nullptr; size_t cap_text = 0;
    std::vector<uint8_t> host;
    int rc = BSTR_OK;
    for (size_t off = 0; off < size && rc == BSTR_OK; off += W) {
        const size_t own = size - off < W ? size - off : W;
        const size_t nb = size - off < W + LA ? size - off : W + LA;
        const bool eof = off + nb == size;
        if (dev->decode((const uint8_t *)map + off, nb, own, off, eof, st) || dev->chain(*fs, st) || dev->crc_launch(st) || dev->crc_fold(*fs, st)) {
            say(dev->error()); rc = BSTR_ERR_IO; break;
        }
        const uint64_t n = dev->text_bytes();
        if (n > cap_text) {
            cudaFree(d_text); d_text = nullptr;
            cap_text = (size_t)n + (size_t)n / 4 + 4096;
            if (cudaMalloc((void **)&d_text, cap_text) != cudaSuccess) { cudaGetLastError(); say("out of device memory for the inflated window"); rc = BSTR_ERR_CUDA; break; }
        }
        if (n) {
            host.resize((size_t)n);
            if (dev->resolve(d_text, st) || cudaMemcpyAsync(host.data(), d_text, (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                cudaStreamSynchronize(st) != cudaSuccess) { say(dev->error()[0] ? dev->error() : "CUDA error reading the inflated window"); rc = BSTR_ERR_CUDA; break; }
            const uint8_t *q = host.data(); size_t left = (size_t)n;
            while (out_fd >= 0 && left) {
                const ssize_t w = write(out_fd, q, left);
                if (w < 0) { if (errno == EINTR) continue; say(strerror(errno)); rc = BSTR_ERR_IO; break; }
                q += w; left -= (size_t)w;
            }
        }
        if (fs->ended) break;
    }
    if (rc == BSTR_OK && !fs->ended) { say("unexpected end of compressed data"); rc = BSTR_ERR_IO; }
    if (stats6) { stats6[0] = size; stats6[1] = fs->out_bytes; stats6[2] = fs->ranges; stats6[3] = fs->ranges - fs->redone; stats6[4] = fs->redone; stats6[5] = fs->members; }
    cudaFree(d_text);
    delete fs;
    delete dev;
    if (st) cudaStreamDestroy(st);
    munmap(map, size);
    return rc;
}
