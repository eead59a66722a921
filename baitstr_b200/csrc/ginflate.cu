// ginflate.cu -- gzip inflate on the GPU, host side (see ginflate.h for the stages).
#include "ginflate.h"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>                      // crc32_combine only

#include <chrono>
#include <vector>

#include "ginflate_dev.cuh"

namespace ginflate {

// ---- CRC-32 of a concatenation from the CRCs of its parts (as zlib's crc32_combine): crc(A || B) =
// crc(A) * x^(8 |B|) mod P  xor  crc(B) over GF(2), bit-reflected polynomial 0xEDB88320.  The operator
// x^(8 n) mod P for the fixed chunk length is computed once.
constexpr uint32_t CRC_CHUNK = 8u << 10;
static size_t TARGET_RANGES = 2800;             // ~ the warps a B200 holds for k_decode_ranges (148 SMs x 20); BSTR_GZ_TARGET_RANGES
constexpr size_t MIN_RANGE = 16u << 10;
constexpr size_t SYM_EXTRA = 256u << 10;        // symbols a range may produce beyond 10x its size (the block that straddles its end)
static uint32_t gf2_multmodp(uint32_t a, uint32_t b)
{
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) { p ^= b; if ((a & (m - 1)) == 0) break; }
        m >>= 1;
        b = b & 1 ? (b >> 1) ^ 0xEDB88320u : b >> 1;
    }
    return p;
}
static uint32_t gf2_x2n(uint64_t nbytes)           // x^(8 nbytes) mod P
{
    uint32_t p = 1u << 31, sq = 1u << 23;          // x^0, x^8
    for (uint64_t n = nbytes; n; n >>= 1) { if (n & 1) p = gf2_multmodp(sq, p); sq = gf2_multmodp(sq, sq); }
    return p;
}

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
    WarpTables *d_tables = nullptr;                             // one slot per range (global memory, L1-resident while a warp decodes)
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
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double ms_upload = 0, ms_decode = 0, ms_propagate = 0;      // CUDA-event times, summed over the windows (BSTR_TRACE)

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
    if (const char *tr = getenv("BSTR_GZ_TARGET_RANGES")) { if (atoll(tr) >= 64 && atoll(tr) <= 60000) TARGET_RANGES = (size_t)atoll(tr); }
    const size_t full = (window_bytes + range_bytes - 1) / range_bytes;         // ranges of a full window
    p->max_ranges = (int)(full > TARGET_RANGES + 256 ? full : TARGET_RANGES + 256); // a short window is cut into smaller ranges (decode)
    p->sym_cap = (uint32_t)(range_bytes * 10 + SYM_EXTRA);
    cudaError_t e;
    p->cap_comp = window_bytes + look_ahead + 64;
    if ((e = cudaMalloc((void **)&p->d_comp, p->cap_comp)) != cudaSuccess) return bail("window buffer", e);
    p->cap_arena = full * p->sym_cap;
    if ((e = cudaMalloc((void **)&p->d_arena, p->cap_arena * sizeof(uint16_t))) != cudaSuccess) return bail("symbol arena", e);
    const size_t nr = (size_t)p->max_ranges + 2;
    if ((e = cudaMalloc((void **)&p->d_desc, nr * sizeof(RangeDesc))) != cudaSuccess) return bail("range table", e);
    if ((e = cudaMalloc((void **)&p->d_meta, nr * sizeof(RangeMeta))) != cudaSuccess) return bail("range table", e);
    if ((e = cudaMalloc((void **)&p->d_runs, nr * sizeof(RunRef))) != cudaSuccess) return bail("run table", e);
    if ((e = cudaMalloc((void **)&p->d_hist, nr * 32768)) != cudaSuccess) return bail("history", e);
    if ((e = cudaMalloc((void **)&p->d_tables, nr * sizeof(WarpTables))) != cudaSuccess) return bail("Huffman tables", e);
    if ((e = cudaHostAlloc((void **)&p->h_desc, nr * sizeof(RangeDesc), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    if ((e = cudaHostAlloc((void **)&p->h_meta, nr * sizeof(RangeMeta), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    if ((e = cudaHostAlloc((void **)&p->h_runs, nr * sizeof(RunRef), cudaHostAllocPortable)) != cudaSuccess) return bail("pinned tables", e);
    p->dev_bytes = p->cap_comp + p->cap_arena * 2 + nr * (sizeof(RangeDesc) + sizeof(RangeMeta) + sizeof(RunRef) + 32768 + sizeof(WarpTables));
    e = cudaFuncSetAttribute(k_propagate, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    if (e != cudaSuccess) return bail("shared memory of the propagate kernel", e);
    cudaFuncSetAttribute(k_decode_ranges, cudaFuncAttributePreferredSharedMemoryCarveout, 0);     // all of the SM's L1 for the tables
    for (int i = 0; i < 4; i++) cudaEventCreate(&p->ev[i]);
    {
        static uint16_t kraft4[4096];
        for (int i = 0; i < 4096; i++) {
            int k = 0;
            for (int j = 0; j < 4; j++) { const int l = (i >> (3 * j)) & 7; k += l ? 128 >> l : 0; }
            kraft4[i] = (uint16_t)k;
        }
        if ((e = cudaMemcpyToSymbol(g_kraft4, kraft4, sizeof(kraft4))) != cudaSuccess) return bail("search table", e);
    }
    return d;
}

Device::~Device()
{
    if (!p) return;
    cudaFree(p->d_comp); cudaFree(p->d_arena); cudaFree(p->d_desc); cudaFree(p->d_meta); cudaFree(p->d_runs); cudaFree(p->d_hist);
    cudaFree(p->d_chunks); cudaFree(p->d_crc); cudaFree(p->d_tables);
    if (p->h_desc) cudaFreeHost(p->h_desc);
    if (p->h_meta) cudaFreeHost(p->h_meta);
    if (p->h_runs) cudaFreeHost(p->h_runs);
    if (p->h_chunks) cudaFreeHost(p->h_chunks);
    if (p->h_crc) cudaFreeHost(p->h_crc);
    for (int i = 0; i < 4; i++) if (p->ev[i]) cudaEventDestroy(p->ev[i]);
    delete p;
}

const char *Device::error() const { return p->err; }
uint64_t Device::text_bytes() const { return p->text; }
size_t Device::device_bytes() const { return p->dev_bytes; }
void Device::kernel_ms(double *upload, double *decode, double *propagate) const { *upload = p->ms_upload; *decode = p->ms_decode; *propagate = p->ms_propagate; }

static int launch_decode(DeviceImpl *p, int n, cudaStream_t st)
{
    const int ctas = (n + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    k_decode_ranges<<<ctas, 32 * WARPS_PER_CTA, 0, st>>>(p->d_comp, p->nbytes, p->at_eof ? 1 : 0, p->d_desc, n, p->d_arena, p->d_meta, p->d_tables);
    return p->cuda(cudaGetLastError(), "k_decode_ranges");
}

int Device::decode(const uint8_t *comp, size_t nbytes, size_t own_bytes, uint64_t file_off, bool at_eof, cudaStream_t st)
{
    if (nbytes + 64 > p->cap_comp || own_bytes > p->window_bytes || own_bytes > nbytes) return p->fail("GPU inflater: window larger than the working set");
    p->nbytes = nbytes; p->own_bytes = own_bytes; p->file_off = file_off; p->at_eof = at_eof;
    p->err[0] = 0;
    cudaEventRecord(p->ev[0], st);
    GI_TRY(cudaMemcpyAsync(p->d_comp, comp, nbytes, cudaMemcpyHostToDevice, st));
    GI_TRY(cudaMemsetAsync(p->d_comp + nbytes, 0, 64, st));
    // a window shorter than a full one (small files, the last window) is cut into smaller ranges, so that
    // there still are ~TARGET_RANGES warps: decoding is a serial chain per warp, its time is the size of a range
    size_t R = (own_bytes + TARGET_RANGES - 1) / TARGET_RANGES;
    R = (R + 4095) & ~(size_t)4095;
    if (R < MIN_RANGE) R = MIN_RANGE;
    if (R > p->range_bytes) R = p->range_bytes;
    p->nranges = (int)((own_bytes + R - 1) / R);
    if (p->nranges == 0) p->nranges = 1;
    if (p->nranges > p->max_ranges) return p->fail("GPU inflater: too many ranges");
    size_t cap = R * 10 + SYM_EXTRA;
    if (cap * (size_t)p->nranges > p->cap_arena) cap = p->cap_arena / (size_t)p->nranges;
    cap &= ~(size_t)7;
    for (int r = 0; r < p->nranges; r++) {
        RangeDesc &d = p->h_desc[r];
        const uint64_t lo = (uint64_t)r * R, hi = (uint64_t)(r + 1) * R < own_bytes ? (uint64_t)(r + 1) * R : own_bytes;
        d.from_bit = lo * 8; d.stop_bit = hi * 8;
        d.sym_off = (uint64_t)r * cap; d.sym_cap = (uint32_t)cap;
        d.mode = (r == 0 && file_off == 0) ? 2 : 0;
    }
    GI_TRY(cudaMemcpyAsync(p->d_desc, p->h_desc, (size_t)p->nranges * sizeof(RangeDesc), cudaMemcpyHostToDevice, st));
    cudaEventRecord(p->ev[1], st);
    if (launch_decode(p, p->nranges, st)) return -1;
    cudaEventRecord(p->ev[2], st);
    GI_TRY(cudaMemcpyAsync(p->h_meta, p->d_meta, (size_t)p->nranges * sizeof(RangeMeta), cudaMemcpyDeviceToHost, st));
    GI_TRY(cudaStreamSynchronize(st));
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, p->ev[0], p->ev[1]); cudaEventElapsedTime(&b, p->ev[1], p->ev[2]);
    p->ms_upload += a; p->ms_decode += b;
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
            k_decode_ranges<<<1, 32 * WARPS_PER_CTA, 0, st>>>(p->d_comp, p->nbytes, p->at_eof ? 1 : 0, p->d_desc + p->max_ranges + 1, 1, p->d_arena,
                                                            p->d_meta + p->max_ranges + 1, p->d_tables + p->max_ranges + 1);
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
        cudaEventRecord(p->ev[0], st);
        k_propagate<<<1, 1024, 65536, st>>>(p->d_arena, p->d_runs, p->nruns, p->d_hist);
        GI_TRY(cudaGetLastError());
        cudaEventRecord(p->ev[3], st);
    }
    GI_TRY(cudaMemcpyAsync(fs.hist, p->d_hist + (size_t)p->nruns * 32768, 32768, cudaMemcpyDeviceToHost, st));
    GI_TRY(cudaStreamSynchronize(st));
    if (p->nruns) { float a = 0; cudaEventElapsedTime(&a, p->ev[0], p->ev[3]); p->ms_propagate += a; }
    fs.out_bytes += p->text;
    return 0;
}

int Device::crc_launch(cudaStream_t st)
{
    // pieces = the window's text cut at the member ends; chunks = pieces cut at run boundaries and every 128 KiB
    const uint32_t CH = CRC_CHUNK;
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
    static const uint32_t op_full = gf2_x2n(CRC_CHUNK);
    for (const DeviceImpl::Piece &pc : p->pieces) {
        for (uint32_t c = 0; c < pc.nchunks; c++) {
            const uint32_t n = p->h_chunks[pc.first_chunk + c].n;
            const uint32_t cc = p->h_crc[pc.first_chunk + c];
            fs.crc = fs.len == 0 ? cc : gf2_multmodp(n == CRC_CHUNK ? op_full : gf2_x2n(n), fs.crc) ^ cc;
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
#include <stdlib.h>
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
    const size_t W = window_bytes ? (size_t)window_bytes : ginflate::DEFAULT_WINDOW, LA = ginflate::LOOK_AHEAD;
    ginflate::Device *dev = ginflate::Device::create(W, LA, range_bytes ? (size_t)range_bytes : ginflate::DEFAULT_RANGE, err, errlen);
    if (!dev) { munmap(map, size); return BSTR_ERR_CUDA; }
    cudaStream_t st = nullptr;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    ginflate::FileState *fs = new ginflate::FileState();
    uint8_t *d_text = nullptr; size_t cap_text = 0;
    std::vector<uint8_t> host;
    int rc = BSTR_OK;
    double t_dec = 0, t_chain = 0, t_crc = 0, t_out = 0;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    for (size_t off = 0; off < size && rc == BSTR_OK; off += W) {
        const size_t own = size - off < W ? size - off : W;
        const size_t nb = size - off < W + LA ? size - off : W + LA;
        const bool eof = off + nb == size;
        const double t0 = now();
        int bad = dev->decode((const uint8_t *)map + off, nb, own, off, eof, st);
        const double t1 = now();
        if (!bad) bad = dev->chain(*fs, st);
        const double t2 = now();
        if (!bad) bad = dev->crc_launch(st) || dev->crc_fold(*fs, st);
        const double t3 = now();
        t_dec += t1 - t0; t_chain += t2 - t1; t_crc += t3 - t2;
        if (bad) { say(dev->error()); rc = BSTR_ERR_IO; break; }
        const double t4 = now();
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
        t_out += now() - t4;
        if (fs->ended) break;
    }
    if (rc == BSTR_OK && !fs->ended) { say("unexpected end of compressed data"); rc = BSTR_ERR_IO; }
    double ku = 0, kd = 0, kp = 0;
    dev->kernel_ms(&ku, &kd, &kp);
    if (getenv("BSTR_TRACE"))
        fprintf(stderr, "[bstr] gunzip_gpu kernels: H2D of the compressed windows %.1f ms, k_decode_ranges %.1f ms, k_propagate %.1f ms\n", ku, kd, kp);
    if (getenv("BSTR_TRACE"))
        fprintf(stderr, "[bstr] gunzip_gpu %s: %.1f MB -> %.1f MB | upload+decode %.1f ms, chain+propagate %.1f ms (%llu of %llu ranges decoded again), "
                "crc %.1f ms, resolve+read back+write %.1f ms\n", path, size / 1e6, fs->out_bytes / 1e6, t_dec, t_chain,
                (unsigned long long)fs->redone, (unsigned long long)fs->ranges, t_crc, t_out);
    if (stats6) { stats6[0] = size; stats6[1] = fs->out_bytes; stats6[2] = fs->ranges; stats6[3] = fs->ranges - fs->redone; stats6[4] = fs->redone; stats6[5] = fs->members; }
    cudaFree(d_text);
    delete fs;
    delete dev;
    if (st) cudaStreamDestroy(st);
    munmap(map, size);
    return rc;
}
