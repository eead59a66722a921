// select_host.cu -- host side of libbaitstr_select.so: C ABI (include/baitstr_select.h),
// device buffers, batch pipeline launches.  The streaming file driver is in stream_host.cu.
//
// Host role of the reference that this replaces: Pool/Chunker/pool.map/print loop of
// /root/reference/src/select_STR_reads.pyx:339-373.  No CPU implementation of the scan
// exists in this library; every entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <mutex>
#include <vector>

#include "host_common.h"
#include "kernels_v1.cuh"
#include "fused.cuh"

namespace bstr {

void set_error(bstr_ctx *ctx, const char *fmt, ...)
{
    std::lock_guard<std::mutex> g(*(std::mutex *)ctx->lock);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
}

int cuda_fail(bstr_ctx *ctx, cudaError_t e, const char *what, int line)
{
    set_error(ctx, "CUDA error at %s (select_host.cu:%d): %s", what, line, cudaGetErrorString(e));
    return BSTR_ERR_CUDA;
}

static int grow(bstr_ctx *ctx, void **p, size_t *cap, size_t need, bool pinned_host)
{
    if (*cap >= need) return BSTR_OK;
    size_t ncap = need + need / 4 + 4096;
    if (*p) { if (pinned_host) cudaFreeHost(*p); else cudaFree(*p); *p = nullptr; *cap = 0; }
    cudaError_t e = pinned_host ? cudaHostAlloc(p, ncap, cudaHostAllocPortable) : cudaMalloc(p, ncap);
    if (e != cudaSuccess) return cuda_fail(ctx, e, pinned_host ? "cudaHostAlloc" : "cudaMalloc", __LINE__);
    *cap = ncap;
    return BSTR_OK;
}

#define GROW(ptr, cap, need, pinned)                                                     \
    do { int rc_ = grow(ctx, (void **)&(ptr), &(cap), (need), (pinned)); if (rc_) return rc_; } while (0)

// size the device buffers of the fused path for a batch of nbytes raw bytes (worst cases,
// see DESIGN.md); the staged generic kernels get theirs in reserve_generic()
int DeviceSlot::reserve(bstr_ctx *ctx, uint64_t nbytes, uint64_t out_need)
{
    const uint64_t max_rec = nbytes / 4 + 2;
    GROW(d_in_raw, cap_in, nbytes + 2 * IN_SLACK, false);
    d_in = d_in_raw + IN_SLACK;
    if (ctx->p.want_results) GROW(d_res, cap_res, max_rec * sizeof(ReadResult), false);
    if (out_need < nbytes + nbytes / 2 + 4096) out_need = nbytes + nbytes / 2 + 4096;
    GROW(d_out, cap_out, out_need, false);
    const uint64_t n_tiles_f = nbytes / FT_MINTILE + 2;
    GROW(d_lb, cap_lb, (2 * n_tiles_f + 2) * 8, false);
    if (!d_state) {
        CUDA_TRY(cudaMalloc((void **)&d_state, sizeof(BatchState) + 64));
        CUDA_TRY(cudaHostAlloc((void **)&h_state, sizeof(BatchState), cudaHostAllocPortable));
    }
    if (ctx->force_generic) return reserve_generic(ctx, nbytes);
    return BSTR_OK;
}

int DeviceSlot::reserve_generic(bstr_ctx *ctx, uint64_t nbytes)
{
    const uint64_t max_lines = nbytes + 4;                 // every byte a newline
    const uint64_t max_rec = nbytes / 4 + 2;
    const uint64_t n_tiles = nbytes / IDX_TILE + 2;
    GROW(d_nlpos, cap_nlpos, max_lines * 4, false);
    GROW(d_tiles, cap_tiles, n_tiles * 4, false);
    GROW(d_res, cap_res, max_rec * sizeof(ReadResult), false);
    GROW(d_outlen, cap_outlen, max_rec * 4, false);
    return BSTR_OK;
}

void DeviceSlot::release()
{
    cudaFree(d_in_raw); cudaFree(d_nlpos); cudaFree(d_tiles); cudaFree(d_res);
    cudaFree(d_outlen); cudaFree(d_out); cudaFree(d_state); cudaFree(d_lb);
    if (h_state) cudaFreeHost(h_state);
    if (h_out) cudaFreeHost(h_out);
    if (h_res) cudaFreeHost(h_res);
    memset(this, 0, sizeof(*this));
}

ScanParams scan_params(const bstr_ctx *ctx)
{
    ScanParams s;
    s.n = ctx->p.num_copies < 1 ? 1 : ctx->p.num_copies;
    s.f = ctx->p.flank < 0 ? 0 : ctx->p.flank;     // a negative flank never rejects anything
    s.kmask = ctx->p.kmask;
    s.rmhm = ctx->p.remove_homopolymers != 0;
    s.altalgo = ctx->p.altalgo != 0;
    return s;
}

// The staged generic pipeline (kernels_v1.cuh): any input, four passes over the data.
static int launch_generic(bstr_ctx *ctx, DeviceSlot &s, uint64_t nbytes, int final, cudaStream_t st,
                          cudaEvent_t *ev /* 5 events or NULL */)
{
    const ScanParams prm = scan_params(ctx);
    { int rc_ = s.reserve_generic(ctx, nbytes); if (rc_) return rc_; }
    const unsigned n_tiles = (unsigned)((nbytes + IDX_TILE - 1) / IDX_TILE);
    unsigned long long *d_total = (unsigned long long *)((uint8_t *)s.d_state + sizeof(BatchState));
    const int rec_blocks = ctx->sm_count * 8;
    if (ev) cudaEventRecord(ev[0], st);
    if (n_tiles) k_count_newlines<<<n_tiles, IDX_THREADS, 0, st>>>(s.d_in, nbytes, s.d_tiles);
    k_exclusive_scan<<<1, 1024, 0, st>>>(s.d_tiles, n_tiles, d_total);
    if (n_tiles) k_fill_newlines<<<n_tiles, IDX_THREADS, 0, st>>>(s.d_in, nbytes, s.d_tiles, s.d_nlpos);
    k_finish_index<<<1, 256, 0, st>>>(s.d_nlpos, d_total, s.d_in, nbytes, final, s.d_state);
    if (ev) cudaEventRecord(ev[1], st);
    k_scan_records<<<rec_blocks, 128, 0, st>>>(s.d_in, s.d_nlpos, s.d_state, prm, s.d_res, s.d_outlen);
    if (ev) cudaEventRecord(ev[2], st);
    k_scan_outlen<<<1, 1024, 0, st>>>(s.d_outlen, s.d_state);
    if (ev) cudaEventRecord(ev[3], st);
    k_format_records<<<rec_blocks, 128, 0, st>>>(s.d_in, s.d_nlpos, s.d_state, s.d_res, s.d_outlen,
                                                 ctx->p.illumina_quals, s.d_out, s.cap_out);
    if (ev) cudaEventRecord(ev[4], st);
    CUDA_TRY(cudaGetLastError());
    return BSTR_OK;
}

cudaError_t set_tile_attrs()
{
    cudaError_t e = cudaFuncSetAttribute(k_fused_tile<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_MAXSMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fused_tile<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_MAXSMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fused_tile<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_MAXSMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_fused_tile<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_MAXSMEM);
    return e;
}

// Tile geometry for a batch whose records average `rec_bytes` bytes (fused.cuh: TileGeom).
// The tile is sized so that it owns about as many records as the CTA has threads; the
// CTA size follows from the tile size.  BSTR_GEOM=tile,threads[,stage] overrides (sweeps).
static TileGeom geometry_for_tile(const bstr_ctx *ctx, double rec_bytes, uint32_t max_threads, uint32_t tile0);

TileGeom pick_geometry(const bstr_ctx *ctx, double rec_bytes, uint32_t max_threads)
{
    if (rec_bytes < 16) rec_bytes = 16;
    if (ctx->geom_tile || ctx->auto_tile_off) return geometry_for_tile(ctx, rec_bytes, max_threads, ctx->geom_tile);
    // Automatic tile: the largest one (1 KiB steps) that still lets two CTAs share an SM (228 KB
    // per SM, 1 KB reserved per CTA) -- per-tile latencies (ticket, bulk copy, look-backs) are
    // amortised over more records.  If the last record warp of such a tile would be less than
    // half full, the tile is cut back to a whole number of warps instead: a nearly empty warp
    // costs as many issue slots as a full one (C5: 140 records per tile ran 8 % slower than 127).
    const uint32_t budget = 228u * 1024u / 2u - 1024u;
    uint32_t tile = FT_DEFTILE;
    for (uint32_t t = 49152; t > FT_DEFTILE; t -= 1024) {
        const TileGeom g = geometry_for_tile(ctx, rec_bytes, max_threads, t);      // may come back smaller than t
        if (g.smem_bytes <= budget) { if (g.tile > tile) tile = g.tile; break; }
    }
    const double recs = tile / rec_bytes;
    const double warps = floor(recs / 32.0);
    if (warps >= 1.0 && recs / 32.0 - warps < 0.5) {
        const uint32_t cut = (uint32_t)(warps * 32.0 * rec_bytes * 0.99) & ~15u;
        if (cut >= FT_DEFTILE * 3 / 4) tile = cut;
    }
    return geometry_for_tile(ctx, rec_bytes, max_threads, tile);
}

static TileGeom geometry_for_tile(const bstr_ctx *ctx, double rec_bytes, uint32_t max_threads, uint32_t tile0)
{
    uint32_t tile = tile0 ? tile0 : FT_DEFTILE;
    if (tile < FT_MINTILE) tile = FT_MINTILE;
    tile &= ~15u;
    if (tile + FT_OVL > FT_MAXREGION) tile = (FT_MAXREGION - FT_OVL) & ~15u;
    uint32_t threads = ctx->geom_threads;
    if (!threads) {
        const double recs = tile / rec_bytes * 1.03 + 2;       // records a tile owns, with margin
        threads = (((uint32_t)recs + 31) & ~31u) + 32;         // + the scan warp (warp 0 owns no records)
    }
    const uint32_t chunks = (tile + FT_OVL) / 16;
    const uint32_t min_thr = ((chunks + FT_MAXCPT - 1) / FT_MAXCPT + 31) & ~31u;
    if (threads < min_thr) threads = min_thr;
    if (threads < 64) threads = 64;
    if (threads > max_threads) threads = max_threads & ~31u;
    // the index phase gives every thread at most FT_MAXCPT 16-byte chunks of the window
    const uint32_t max_tile = (threads * FT_MAXCPT * 16u - FT_OVL) & ~15u;
    if (tile > max_tile) tile = max_tile;
    // a tile may own at most 2 x (threads - 32) records: shrink it for very short records
    const double fit = 1.8 * (threads - 32) * rec_bytes;
    if (tile > fit && !ctx->geom_tile) { tile = (uint32_t)fit & ~15u; if (tile < FT_MINTILE) tile = FT_MINTILE; }
    uint32_t stage = ctx->geom_stage;
    if (!stage) stage = (uint32_t)(tile * (1.0 + 48.0 / rec_bytes) + rec_bytes) + 2048;
    // phase D parks the matcher's per-warp work lists in the staging buffer (fused.cuh: WarpWork)
    const uint32_t lists = (threads / 32) * 4u * warp_work_words<8>();
    if (stage < lists && !ctx->geom_stage) stage = lists;
    stage = (stage + 15) & ~15u;
    TileGeom g = make_geometry(tile, threads, stage);
    if (g.smem_bytes > (uint32_t)FT_MAXSMEM)       // tiny records: a smaller stage, overflow -> generic kernels
        g = make_geometry(tile, threads, stage - (g.smem_bytes - FT_MAXSMEM));
    return g;
}

}  // namespace bstr

// Host-only test hook (no GPU): the tile geometry pick_geometry() chooses for batches whose
// records average `record_bytes` bytes, for the kernel instantiation with `window_words`
// (2, 4, 6, 8).  out = {tile, threads, region, chunks, chunks per thread, recmax, stage, smem}.
extern "C" int bstr_geometry_probe(double record_bytes, int window_words, uint32_t out[8])
{
    if (!out || !(record_bytes > 0)) return BSTR_ERR_PARAM;
    bstr_ctx probe;                                     // zeroed: no overrides, automatic tile
    const uint32_t max_threads = window_words >= bstr::FT_WIDEFROM ? bstr::FT_WIDETHREADS : bstr::FT_MAXTHREADS;
    const bstr::TileGeom g = bstr::pick_geometry(&probe, record_bytes, max_threads);
    out[0] = g.tile; out[1] = g.threads; out[2] = g.region; out[3] = g.chunks; out[4] = g.cpt;
    out[5] = g.recmax; out[6] = g.stage; out[7] = g.smem_bytes;
    return BSTR_OK;
}

namespace bstr {

// first bytes of a host batch -> average record size (4 lines per record); *hdr_bytes gets
// the length of the first line (header), which separates read length from record size
double estimate_record_bytes(const uint8_t *raw, uint64_t nbytes, double *hdr_bytes)
{
    const uint64_t span = nbytes < (128u << 10) ? nbytes : (128u << 10);
    uint64_t lines = 0;
    const uint8_t *first = nullptr;
    for (const uint8_t *p = raw, *e = raw + span; p < e && (p = (const uint8_t *)memchr(p, '\n', (size_t)(e - p))); p++) {
        if (!first) first = p;
        lines++;
    }
    if (hdr_bytes) *hdr_bytes = first ? (double)(first - raw) : 0.0;
    return lines >= 8 ? 4.0 * (double)span / (double)lines : 256.0;
}

// The fused single-pass tile kernel (fused.cuh).  Sets BatchState::fallback when the input
// does not fit the tile geometry; the caller then runs launch_generic on the same slot.
static int launch_fused(bstr_ctx *ctx, DeviceSlot &s, uint64_t nbytes, int final, cudaStream_t st,
                        cudaEvent_t *ev)
{
    const ScanParams prm = scan_params(ctx);
    // widest bit-parallel window the batch is expected to need (longer reads stay exact through
    // the byte path): L = (record - header - 6) / 2, window = L - 2 (f - 6) for the default
    // algorithm; 4 % margin for reads longer than the average
    const int fe = prm.altalgo && prm.f > 6 ? prm.f - 6 : 0;
    const double win = (s.rec_bytes - s.hdr_bytes - 6.0) * 0.5 * 1.04 - 2.0 * fe;
    const int wmax = ctx->force_wmax ? ctx->force_wmax : (win <= 64 ? 2 : (win <= 128 ? 4 : (win <= 192 ? 6 : 8)));
    const uint32_t max_threads = wmax >= FT_WIDEFROM ? FT_WIDETHREADS : FT_MAXTHREADS;   // launch bound of k_fused_tile<wmax>
    const TileGeom g = pick_geometry(ctx, s.rec_bytes, max_threads);
    const uint32_t n_tiles = (uint32_t)((nbytes + g.tile - 1) / g.tile);
    unsigned long long *lb_lines = s.d_lb, *lb_out = s.d_lb + n_tiles, *ticket = s.d_lb + 2 * (size_t)n_tiles;
    const uint32_t lb_words = 2 * n_tiles + 1;
    const uint32_t init_blocks = (lb_words + 255) / 256 < 64 ? (lb_words + 255) / 256 : 64;
    if (ev) cudaEventRecord(ev[0], st);
    k_fused_init<<<init_blocks, 256, 0, st>>>(s.d_state, s.d_lb, lb_words);
    if (ev) cudaEventRecord(ev[1], st);
#define BSTR_LAUNCH_TILE(W)                                                                          \
    k_fused_tile<W><<<n_tiles, g.threads, g.smem_bytes, st>>>(                                       \
        s.d_in, nbytes, final, prm, ctx->p.illumina_quals, s.d_state, lb_lines, lb_out, ticket,      \
        n_tiles, ctx->p.want_results ? s.d_res : nullptr, s.d_out, s.cap_out, ctx->exp_mask, g)
    if (n_tiles) {
        if (wmax <= 2) BSTR_LAUNCH_TILE(2);
        else if (wmax <= 4) BSTR_LAUNCH_TILE(4);
        else if (wmax <= 6) BSTR_LAUNCH_TILE(6);
        else BSTR_LAUNCH_TILE(8);
    }
#undef BSTR_LAUNCH_TILE
    if (ev) { cudaEventRecord(ev[2], st); cudaEventRecord(ev[4], st); }
    CUDA_TRY(cudaGetLastError());
    return BSTR_OK;
}

// Launch the device pipeline for the nbytes at slot.d_in.  Nothing here synchronises.
int launch_pipeline(bstr_ctx *ctx, DeviceSlot &s, uint64_t nbytes, int final, cudaStream_t st,
                    cudaEvent_t *ev, bool generic)
{
    if (generic || ctx->force_generic) return launch_generic(ctx, s, nbytes, final, st, ev);
    return launch_fused(ctx, s, nbytes, final, st, ev);
}

int pipeline_stage_count(const bstr_ctx *ctx) { return ctx->force_generic ? 4 : 2; }
const char *pipeline_stage_name(const bstr_ctx *ctx, int k)
{
    static const char *gen[4] = {"index(newlines)", "k_scan_records", "k_scan_outlen", "k_format_records"};
    static const char *fus[2] = {"k_fused_init", "k_fused_tile"};
    return ctx->force_generic ? gen[k] : fus[k];
}
int pipeline_launches_per_batch(const bstr_ctx *ctx) { return ctx->force_generic ? 7 : 2; }

int decode_bad(bstr_ctx *ctx, unsigned long long first_bad, int64_t *bad_record)
{
    if (first_bad == ~0ull) return BSTR_OK;
    const int code = (int)(first_bad & 0xFF);
    const long long rec = (long long)(first_bad >> 8);
    if (bad_record) *bad_record = rec;
    static const char *msg[] = {"", "header should begin with a @ in a fastq file",
                                "read should only have ACGTN", "separator line in fastq file should begin with a +",
                                "file ends inside a record"};
    set_error(ctx, "malformed FASTQ at record %lld: %s", rec, msg[code <= 4 ? code : 0]);
    switch (code) {
        case BAD_HEADER: return BSTR_ERR_FASTQ_HEADER;
        case BAD_BASES: return BSTR_ERR_FASTQ_BASES;
        case BAD_SEPARATOR: return BSTR_ERR_FASTQ_SEPARATOR;
        default: return BSTR_ERR_FASTQ_TRUNCATED;
    }
}

}  // namespace bstr

using namespace bstr;

extern "C" {

int bstr_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

bstr_ctx *bstr_create(const bstr_params *p, char *err, size_t errlen)
{
    auto fail = [&](const char *m) -> bstr_ctx * {
        if (err && errlen) snprintf(err, errlen, "%s", m);
        return nullptr;
    };
    if (!p || p->abi_version != BSTR_ABI_VERSION) return fail("bstr_create: ABI version mismatch");
    if (p->num_copies < 0) return fail("bstr_create: numcopies must be >= 0");
    int ndev = bstr_device_count();
    if (ndev <= 0)
        return fail("bstr_create: no CUDA device available (this library has no CPU fallback)");
    bstr_ctx *ctx = new bstr_ctx();
    ctx->lock = new std::mutex();
    ctx->p = *p;
    if (ctx->p.kmask == 0) ctx->p.kmask = 0x7C;            // 2..6 (select_STR_reads:134-136)
    ctx->p.kmask &= 0x7C;
    if (ctx->p.kmask == 0) { delete (std::mutex *)ctx->lock; delete ctx; return fail("bstr_create: no period in 2..6 selected"); }
    if (ctx->p.batch_bytes == 0) {
        const char *bb = getenv("BSTR_BATCH_BYTES");         // tuning knob; the bstr_params field wins
        ctx->p.batch_bytes = bb && atoll(bb) > 0 ? (uint64_t)atoll(bb) : 8ull << 20;
    }
    if (ctx->p.batch_bytes > (2ull << 30)) ctx->p.batch_bytes = 2ull << 30;
    if (ctx->p.n_devices <= 0) {
        ctx->p.n_devices = ndev > 16 ? 16 : ndev;
        for (int i = 0; i < ctx->p.n_devices; i++) ctx->p.devices[i] = i;
    }
    for (int i = 0; i < ctx->p.n_devices; i++)
        if (ctx->p.devices[i] < 0 || ctx->p.devices[i] >= ndev) {
            delete (std::mutex *)ctx->lock;
            delete ctx;
            return fail("bstr_create: device ordinal out of range");
        }
    ctx->dev = ctx->p.devices[0];
    cudaError_t e = cudaSetDevice(ctx->dev);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, ctx->dev);
    if (e != cudaSuccess) {
        if (err && errlen) snprintf(err, errlen, "bstr_create: %s", cudaGetErrorString(e));
        delete (std::mutex *)ctx->lock;
        delete ctx;
        return nullptr;
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->bench_slots = new std::vector<DeviceSlot>();
    {
        const char *fg = getenv("BSTR_FORCE_GENERIC");
        ctx->force_generic = fg && fg[0] == '1';
        const char *ex = getenv("BSTR_EXP");           // profiling experiments only: skips phases
        ctx->exp_mask = ex ? atoi(ex) : 0;
        const char *ge = getenv("BSTR_GEOM");          // tile,threads[,stage] (0 = automatic)
        if (ge) sscanf(ge, "%u,%u,%u", &ctx->geom_tile, &ctx->geom_threads, &ctx->geom_stage);
        const char *at = getenv("BSTR_AUTOTILE");      // 0: the fixed 40 KiB default tile (A/B measurements)
        ctx->auto_tile_off = at && at[0] == '0';
        const char *wm = getenv("BSTR_WMAX");          // window-width instantiation: 2, 4, 6 or 8 (0 = automatic)
        const int w = wm ? atoi(wm) : 0;
        ctx->force_wmax = (w == 2 || w == 4 || w == 6 || w == 8) ? w : 0;
    }
    e = set_tile_attrs();
    if (e != cudaSuccess) {
        if (err && errlen) snprintf(err, errlen, "bstr_create: %s", cudaGetErrorString(e));
        bstr_destroy(ctx);
        return nullptr;
    }
    for (int i = 0; i < 8; i++) cudaEventCreate(&ctx->ev[i]);
    return ctx;
}

void bstr_destroy(bstr_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->dev);
    stream_release(ctx);
    lanes_release(ctx);
    ctx->slot.release();
    if (ctx->bench_slots) { for (auto &s : *ctx->bench_slots) s.release(); delete ctx->bench_slots; }
    cudaFree(ctx->d_seqs); cudaFree(ctx->d_off); cudaFree(ctx->d_flush);
    for (int i = 0; i < 8; i++) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    delete (std::mutex *)ctx->lock;
    delete ctx;
}

const char *bstr_last_error(const bstr_ctx *ctx) { return ctx ? ctx->err : "null context"; }

void bstr_get_stats(const bstr_ctx *ctx, bstr_stats *st) { if (ctx && st) *st = ctx->stats; }
void bstr_reset_stats(bstr_ctx *ctx) { if (ctx) memset(&ctx->stats, 0, sizeof(ctx->stats)); }

void *bstr_host_alloc(size_t nbytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, nbytes ? nbytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void bstr_host_free(void *p) { if (p) cudaFreeHost(p); }

}  // extern "C"

#include "batch_api.inc"
#include "noalloc_api.inc"
#include "kmer_api.inc"
