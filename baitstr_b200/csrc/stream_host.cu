// stream_host.cu -- bstr_run_files: the file loop of the select stage
// (/root/reference/src/select_STR_reads.pyx:359-370; reader = /root/reference/src/fastq.pyx:48-94).
//
// Files are consumed in argv order; ".gz" is decided by the file-name suffix exactly as
// fastq.pyx:50 does.  The host never parses FASTQ: it moves raw bytes into batches, the
// device reports how many bytes were whole records, and the unconsumed tail is carried to
// the front of the next batch.
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <zlib.h>

#include "host_common.h"

namespace bstr {

struct StreamEngine {
    uint8_t *buf;        // pinned staging buffer
    size_t cap;
};

void stream_release(bstr_ctx *ctx)
{
    if (!ctx->engine) return;
    if (ctx->engine->buf) cudaFreeHost(ctx->engine->buf);
    delete ctx->engine;
    ctx->engine = nullptr;
}

struct Source {
    int fd = -1;
    gzFile gz = nullptr;
    bool is_gz = false;
    int open_path(const char *path)
    {
        const size_t n = strlen(path);
        is_gz = n >= 3 && strcmp(path + n - 3, ".gz") == 0;          // fastq.pyx:50
        if (is_gz) {
            gz = gzopen(path, "rb");
            if (!gz) return -1;
            gzbuffer(gz, 1 << 20);
        } else {
            fd = open(path, O_RDONLY);
            if (fd < 0) return -1;
        }
        return 0;
    }
    // returns bytes read (0 = EOF) or -1
    long read_some(uint8_t *dst, size_t want)
    {
        if (want > (1u << 30)) want = 1u << 30;
        if (is_gz) {
            int r = gzread(gz, dst, (unsigned)want);
            if (r < 0) return -1;
            if (r == 0 && !gzeof(gz)) return -1;
            return r;
        }
        for (;;) {
            ssize_t r = read(fd, dst, want);
            if (r < 0 && errno == EINTR) continue;
            return (long)r;
        }
    }
    void close_all()
    {
        if (gz) gzclose(gz);
        if (fd >= 0) close(fd);
        gz = nullptr; fd = -1;
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

static int ensure_buf(bstr_ctx *ctx, size_t need)
{
    StreamEngine *e = ctx->engine;
    if (e->cap >= need) return BSTR_OK;
    uint8_t *nb = nullptr;
    CUDA_TRY(cudaHostAlloc((void **)&nb, need, cudaHostAllocDefault));
    if (e->buf) cudaFreeHost(e->buf);
    e->buf = nb; e->cap = need;
    return BSTR_OK;
}

}  // namespace bstr

using namespace bstr;

extern "C" int bstr_run_files(bstr_ctx *ctx, const char *const *paths, int npaths, int out_fd,
                              int err_fd)
{
    if (!ctx || (!paths && npaths)) return BSTR_ERR_PARAM;
    CUDA_TRY(cudaSetDevice(ctx->dev));
    if (!ctx->engine) { ctx->engine = new StreamEngine(); ctx->engine->buf = nullptr; ctx->engine->cap = 0; }
    int rc = ensure_buf(ctx, ctx->p.batch_bytes);
    if (rc) return rc;
    for (int fi = 0; fi < npaths; fi++) {
        Source src;
        if (src.open_path(paths[fi]) != 0) {
            set_error(ctx, "cannot open %s: %s", paths[fi], strerror(errno));
            return BSTR_ERR_IO;
        }
        size_t have = 0;
        uint64_t rec_base = 0;
        bool eof = false;
        while (!eof || have) {
            StreamEngine *e = ctx->engine;
            while (!eof && have < e->cap) {
                long r = src.read_some(e->buf + have, e->cap - have);
                if (r < 0) { src.close_all(); set_error(ctx, "read error on %s", paths[fi]); return BSTR_ERR_IO; }
                if (r == 0) eof = true;
                have += (size_t)r;
            }
            if (have == 0) break;
            bstr_batch_out bo;
            rc = bstr_scan_batch(ctx, e->buf, have, eof ? 1 : 0, &bo);
            if (rc) {
                if (bo.bad_record >= 0) {
                    char tmp[400];
                    snprintf(tmp, sizeof(tmp), "%s", ctx->err);
                    set_error(ctx, "%s: record %llu: %s", paths[fi],
                              (unsigned long long)(rec_base + (uint64_t)bo.bad_record), tmp);
                }
                src.close_all();
                return rc;
            }
            if (bo.out_bytes && write_all(out_fd, bo.out, bo.out_bytes) != 0) {
                src.close_all(); set_error(ctx, "write error: %s", strerror(errno)); return BSTR_ERR_IO;
            }
            rec_base += bo.n_records;
            if (eof) { have = 0; break; }
            if (bo.consumed == 0) {
                // one record larger than the whole batch buffer: enlarge and retry
                if (e->cap >= (2ull << 30)) { src.close_all(); set_error(ctx, "%s: record larger than 2 GiB", paths[fi]); return BSTR_ERR_TOO_LARGE; }
                size_t ncap = e->cap * 2;
                if (ncap > (2ull << 30)) ncap = 2ull << 30;
                uint8_t *nb = nullptr;
                CUDA_TRY(cudaHostAlloc((void **)&nb, ncap, cudaHostAllocDefault));
                memcpy(nb, e->buf, have);
                cudaFreeHost(e->buf);
                e->buf = nb; e->cap = ncap;
                continue;
            }
            memmove(e->buf, e->buf + bo.consumed, have - bo.consumed);
            have -= bo.consumed;
        }
        src.close_all();
        if (err_fd >= 0) {
            char line[4200];
            int n = snprintf(line, sizeof(line), "Done processing %s\n", paths[fi]);   // pyx:370
            if (n > 0) write_all(err_fd, (const uint8_t *)line, (size_t)n);
        }
    }
    return BSTR_OK;
}
