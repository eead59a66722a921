// fused.cuh -- the single-pass tile kernel of the select stage (sm_100a).
//
// One CTA owns TileGeom::tile bytes of raw FASTQ (plus FT_OVL bytes of look-ahead for the records
// that start in its tile) staged in shared memory, and does everything the reference does
// per record in one visit of the data:
//   fastq.next()            /root/reference/src/fastq.pyx:72-94            (line split, checks)
//   ProcessSequence         /root/reference/src/select_STR_reads.pyx:294-319 (both strands)
//   PrintSequence           /root/reference/src/select_STR_reads.pyx:171-185 (annotated FASTQ)
//   ordered output          /root/reference/src/select_STR_reads.pyx:363-367 (input order)
// HBM traffic is the algorithmic minimum: every input byte is read once, every output
// byte written once.  Tiles are chained by two decoupled look-backs (line count, output
// bytes), tickets are handed out in launch order so a tile only ever waits for tiles that
// already run.  Inputs the tile geometry cannot hold (records longer than FT_OVL, more
// lines or records than the tile's tables hold, ...) raise BatchState::fallback and the host re-runs the
// batch through the staged generic kernels (kernels_v1.cuh) -- still on the GPU.
#pragma once
#include <stdint.h>
#include "kernels_v1.cuh"

namespace bstr {

#ifndef BSTR_MAXTHREADS
#define BSTR_MAXTHREADS 256
#endif
#ifndef BSTR_MINCTAS
#define BSTR_MINCTAS 2
#endif
constexpr int FT_MAXTHREADS = BSTR_MAXTHREADS;   // launch bound; the host picks 64..FT_MAXTHREADS per batch
constexpr int FT_MAXCPT = 16;                  // 16-byte chunks per thread in the index phase
#ifndef BSTR_OVL
#define BSTR_OVL 4096
#endif
constexpr int FT_OVL = BSTR_OVL;               // look-ahead past the tile end
constexpr int FT_MAXREGION = 65504;            // line positions are 16-bit
constexpr int FT_MINTILE = 8192;
constexpr int FT_DEFTILE = 40960;
constexpr int FT_MAXSMEM = 200 * 1024;         // opt-in dynamic shared memory limit of the kernel

constexpr unsigned long long LB_MASK = (1ull << 62) - 1;
constexpr unsigned long long LB_AGG = 1ull << 62;      // tile aggregate published
constexpr unsigned long long LB_INC = 2ull << 62;      // inclusive prefix published

// Tile geometry, chosen by the host per batch from the average record size (select_host.cu:
// pick_geometry) so that a tile holds about as many records as the CTA has threads: the
// per-record phases (validate, match, format) then run with full warps and no warp idles
// at the barriers between the phases.
struct TileGeom {
    uint32_t threads;        // CTA size, multiple of 32
    uint32_t tile;           // bytes owned by one tile (multiple of 16)
    uint32_t region;         // tile + FT_OVL
    uint32_t chunks;         // region / 16
    uint32_t cpt;            // chunks per thread (contiguous), <= FT_MAXCPT
    uint32_t recmax;         // 2 * threads: records a tile may own
    uint32_t nlmax;          // line-table entries
    uint32_t stage;          // output staging bytes
    uint32_t off_in, off_stage, off_rec, off_outoff, off_nl;   // byte offsets in dynamic smem
    uint32_t smem_bytes;
};

struct FusedVars {                                      // at offset 0 of the dynamic smem
    unsigned long long mbar;                            // mbarrier of the tile's bulk copy
    long long gbase;                                    // global index of local line 0
    unsigned long long out_base;
    uint32_t tile, pad0;
    uint32_t warp_tmp[FT_MAXTHREADS / 32];
};

inline TileGeom make_geometry(uint32_t tile, uint32_t threads, uint32_t stage)
{
    TileGeom g;
    g.threads = threads; g.tile = tile; g.region = tile + FT_OVL; g.chunks = g.region / 16;
    g.cpt = (g.chunks + threads - 1) / threads;
    // a thread indexes cpt CONTIGUOUS 16-byte chunks, so the lanes of a quarter warp load chunks
    // cpt apart: an odd stride spreads them over all eight 16-byte bank groups, an even one over
    // four or fewer (4-way conflict on every LDS.128 of the index phase)
    if (!(g.cpt & 1u) && g.cpt < (uint32_t)FT_MAXCPT) g.cpt++;
    g.recmax = 2 * threads; g.nlmax = 5 * g.recmax + 64; g.stage = stage;
    uint32_t o = (sizeof(FusedVars) + 15) & ~15u;
    g.off_in = o;       o += 16 + g.region + 32;
    g.off_stage = o;    o += g.stage + 48;
    g.off_rec = o;      o += g.recmax * 16;
    g.off_outoff = o;   o += g.recmax * 4;
    g.off_nl = o;       o += (g.nlmax + 8) * 2;
    g.smem_bytes = (o + 15) & ~15u;
    return g;
}

// ---- TMA bulk copy global -> shared (cp.async.bulk, SASS UBLKCP) with an mbarrier ----------
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gsrc, uint32_t bytes, unsigned long long *bar)
{   // one thread: arm the barrier with the byte count, then one copy instruction for the whole tile
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    } while (!done);
}

// Decoupled look-back in two steps so that work can be placed between them.
// lookback_publish (one thread): make this tile's aggregate visible at once -- successors then
// only wait for tiles whose aggregate is not known yet.  st[] was zeroed before the launch.
__device__ __forceinline__ void lookback_publish(unsigned long long *st, int tile, unsigned long long agg)
{
    atomicExch(&st[tile], (tile == 0 ? LB_INC : LB_AGG) | agg);
}

// lookback_resolve (one full warp, after lookback_publish of the same tile): returns the
// exclusive prefix of `agg` over the tiles before `tile` and publishes the inclusive prefix.
__device__ __forceinline__ unsigned long long lookback_resolve(unsigned long long *st, int tile,
                                                               unsigned long long agg)
{
    const int lane = threadIdx.x & 31;
    if (tile == 0) return 0;
    unsigned long long excl = 0;
    int hi = tile - 1;
    for (;;) {
        const int p = hi - lane;
        unsigned long long v = LB_INC;                   // before tile 0: inclusive 0
        if (p >= 0) {
            do { v = *(volatile unsigned long long *)&st[p]; } while ((v >> 62) == 0);
        }
        const unsigned inc = __ballot_sync(0xFFFFFFFFu, (v >> 62) == 2);
        const int first = inc ? __ffs(inc) - 1 : 32;     // nearest predecessor with a prefix
        unsigned long long c = (lane <= first) ? (v & LB_MASK) : 0ull;
#pragma unroll
        for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
        excl += c;
        if (inc) break;
        hi -= 32;
    }
    if (lane == 0) atomicExch(&st[tile], LB_INC | (excl + agg));
    return excl;
}

__device__ __forceinline__ unsigned long long warp_lookback(unsigned long long *st, int tile,
                                                            unsigned long long agg)
{
    if ((threadIdx.x & 31) == 0) lookback_publish(st, tile, agg);
    __syncwarp();
    return lookback_resolve(st, tile, agg);
}

// 16-bit mask of the '\n' bytes of one 16-byte chunk
__device__ __forceinline__ uint32_t nl_mask_of(const uint4 v)
{
    // nl_flags leaves 0x80 in every newline byte; dp4a with weights 1,2,4,8 / 16..128 packs
    // eight of them into one byte (scaled by 128)
    const uint32_t lo = __dp4a(nl_flags(v.y), 0x80402010u, __dp4a(nl_flags(v.x), 0x08040201u, 0u));
    const uint32_t hi = __dp4a(nl_flags(v.w), 0x80402010u, __dp4a(nl_flags(v.z), 0x08040201u, 0u));
    return (lo >> 7) | (hi << 1);
}

// exclusive scan of one value per thread over the CTA; *total gets the sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_tmp, uint32_t *total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    __syncthreads();                       // warp_tmp may still be read from a previous scan
    if (lane == 31) warp_tmp[wid] = inc;
    __syncthreads();
    uint32_t wbase = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < FT_MAXTHREADS / 32; w++) {
        const uint32_t x = w < nw ? warp_tmp[w] : 0u;
        if (w < wid) wbase += x;
        tot += x;
    }
    *total = tot;
    return wbase + inc - v;
}

__global__ void k_fused_init(BatchState *st, unsigned long long *lb, uint32_t n_lb_words)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_lb_words; i += gridDim.x * blockDim.x)
        lb[i] = 0ull;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->first_bad = ~0ull; st->n_lines = 0; st->n_records = 0; st->n_selected = 0;
        st->out_bytes = 0; st->consumed = 0; st->fallback = 0;
    }
}

__device__ __forceinline__ RecordLines tile_record_lines(const uint16_t *nl, int i)
{
    RecordLines ln;
    const uint32_t e0 = nl[i], e1 = nl[i + 1], e2 = nl[i + 2], e3 = nl[i + 3], e4 = nl[i + 4];
    ln.s0 = e0; ln.e0 = e1 - 1; ln.s1 = e1; ln.e1 = e2 - 1;
    ln.s2 = e2; ln.e2 = e3 - 1; ln.s3 = e3; ln.e3 = e4 - 1;
    return ln;
}

// ---- PrintSequence (select_STR_reads.pyx:171-185) into shared memory, word-granular -------
enum { XF_NONE = 0, XF_UPPER = 1, XF_MINUS31 = 2 };

template <int MODE> __device__ __forceinline__ uint32_t xform_word(uint32_t w)
{
    if (MODE == XF_UPPER) return w & ~((w & 0x40404040u) >> 1);     // letters: clear bit 5 (fastq.pyx:83)
    if (MODE == XF_MINUS31)                                        // per-byte x - 31 (pyx:181), no borrows
        return ((w | 0x80808080u) - 0x1F1F1F1Fu) ^ (~w & 0x80808080u);
    return w;
}
template <int MODE> __device__ __forceinline__ uint8_t xform_byte(uint8_t c)
{
    if (MODE == XF_UPPER) return (c & 0x40) ? (uint8_t)(c & 0xDF) : c;
    if (MODE == XF_MINUS31) return (uint8_t)(c - 31);
    return c;
}

// copy n bytes s -> d (both shared memory, any alignment) with aligned 32-bit stores.  No byte
// loops: the up to three bytes in front of the first aligned word, the up to three words after
// the last 16-byte trip and the up to three bytes at the end are predicated single operations,
// so the lanes of a warp (different alignments) do not wait for each other's loop trips.
template <int MODE>
__device__ __forceinline__ uint8_t *copy_xform(uint8_t *d, const uint8_t *s, int n)
{
    if (n <= 0) return d;
    int k0 = (int)((4u - ((uint32_t)(uintptr_t)d & 3u)) & 3u);
    if (k0 > n) k0 = n;
    if (k0 > 0) d[0] = xform_byte<MODE>(s[0]);
    if (k0 > 1) d[1] = xform_byte<MODE>(s[1]);
    if (k0 > 2) d[2] = xform_byte<MODE>(s[2]);
    d += k0; s += k0; n -= k0;
    const int nw = n >> 2;
    if (nw > 0) {
        const uint32_t *sp = (const uint32_t *)(s - ((uintptr_t)s & 3));   // keeps s's address space (LDS, not LD)
        const int sh = (int)((uintptr_t)s & 3) * 8;
        uint32_t *dp = (uint32_t *)d;
        uint32_t lo = sp[0];
        int i = 0;
        for (; i + 4 <= nw; i += 4) {                       // 16 bytes per trip: 4 LDS, 4 SHF, 4 STS
            const uint32_t w1 = sp[i + 1], w2 = sp[i + 2], w3 = sp[i + 3], w4 = sp[i + 4];
            dp[i] = xform_word<MODE>(__funnelshift_r(lo, w1, sh));
            dp[i + 1] = xform_word<MODE>(__funnelshift_r(w1, w2, sh));
            dp[i + 2] = xform_word<MODE>(__funnelshift_r(w2, w3, sh));
            dp[i + 3] = xform_word<MODE>(__funnelshift_r(w3, w4, sh));
            lo = w4;
        }
        const int rw = nw - i;                              // 0..3 words left
        if (rw > 0) {
            const uint32_t w1 = sp[i + 1];
            dp[i] = xform_word<MODE>(__funnelshift_r(lo, w1, sh));
            if (rw > 1) {
                const uint32_t w2 = sp[i + 2];
                dp[i + 1] = xform_word<MODE>(__funnelshift_r(w1, w2, sh));
                if (rw > 2) dp[i + 2] = xform_word<MODE>(__funnelshift_r(w2, sp[i + 3], sh));
            }
        }
        d += 4 * nw; s += 4 * nw; n -= 4 * nw;
    }
    if (n > 0) d[0] = xform_byte<MODE>(s[0]);
    if (n > 1) d[1] = xform_byte<MODE>(s[1]);
    if (n > 2) d[2] = xform_byte<MODE>(s[2]);
    return d + n;
}

// decimal of v < 1000 at p, no branches: three characters are stored, the first being the leading
// digit; the one or two characters behind a shorter number are overwritten by what follows
__device__ __forceinline__ uint8_t *put_dec3(uint8_t *p, uint32_t v)
{
    const uint32_t h = v / 100u, r = v - 100u * h, t = r / 10u, u = r - 10u * t;
    const bool g100 = v >= 100u, g10 = v >= 10u;
    p[0] = (uint8_t)('0' + (g100 ? h : (g10 ? t : u)));
    p[1] = (uint8_t)('0' + (g100 ? t : u));
    p[2] = (uint8_t)('0' + u);
    return p + 1 + (g10 ? 1 : 0) + (g100 ? 1 : 0);
}
__device__ __forceinline__ uint8_t upper_bf(uint8_t c) { return (uint8_t)(c & ~((c & 0x40) >> 1)); }      // letters only lose bit 5
__device__ __forceinline__ uint8_t comp_upper_bf(uint8_t c)
{   // A <-> T (xor 0x15), C <-> G (xor 0x04), N and '.' (bit 3) unchanged -- select_STR_reads.pyx:187-189 on upper case
    const uint8_t x = (c & 8) ? 0 : ((c & 2) ? 0x04 : 0x15);
    return (uint8_t)(c ^ x);
}

__device__ __forceinline__ void format_record_words(const uint8_t *in, const RecordLines &ln,
                                                    const ReadResult &rr, int illumina, uint8_t *p)
{
    const int L = (int)(ln.e1 - ln.s1);
    const int fk = rr.fck & 7, fc = rr.fck >> 3, rk = rr.rck & 7, rcn = rr.rck >> 3;
    p = copy_xform<XF_NONE>(p, in + ln.s0, (int)(ln.e0 - ln.s0));          // '@' + name (validated)
    if (L < 1000) {
        // Annotation without data-dependent branches: six motif characters and three digits are
        // always stored and the pointer advances by the true width; the surplus is overwritten by
        // the field that follows (the last one by '\n' and the bases, L >= 4 for a selected read).
        // The surplus reads stay inside the staged window (16 bytes of slack on both sides).
        *p = '\t';
        const uint8_t *fm = in + ln.s1 + rr.fstart;
#pragma unroll
        for (int j = 0; j < 6; j++) p[1 + j] = upper_bf(fm[j]);
        p += 1 + fk;
        *p++ = '\t'; p = put_dec3(p, fc);
        *p++ = '\t'; p = put_dec3(p, rr.fstart);
        *p++ = '\t'; p = put_dec3(p, rr.fstart + fk * fc);
        *p = '\t';
        const uint8_t *rm = in + ln.s1 + (L - 1 - (int)rr.rstart);
#pragma unroll
        for (int j = 0; j < 6; j++) p[1 + j] = comp_upper_bf(upper_bf(rm[-j]));
        p += 1 + rk;
        *p++ = '\t'; p = put_dec3(p, rcn);
        *p++ = '\t'; p = put_dec3(p, rr.rstart);
        *p++ = '\t'; p = put_dec3(p, rr.rstart + rk * rcn);
        *p++ = '\n';
    } else {
    *p++ = '\t';
    for (int j = 0; j < fk; j++) *p++ = upper(in[ln.s1 + rr.fstart + j]);
    *p++ = '\t'; p = put_dec(p, fc);
    *p++ = '\t'; p = put_dec(p, rr.fstart);
    *p++ = '\t'; p = put_dec(p, rr.fstart + fk * fc);
    *p++ = '\t';
    for (int j = 0; j < rk; j++) *p++ = comp_base(upper(in[ln.s1 + (L - 1 - (int)rr.rstart - j)]));
    *p++ = '\t'; p = put_dec(p, rcn);
    *p++ = '\t'; p = put_dec(p, rr.rstart);
    *p++ = '\t'; p = put_dec(p, rr.rstart + rk * rcn);
    *p++ = '\n';
    }
    p = copy_xform<XF_UPPER>(p, in + ln.s1, L);
    *p++ = '\n'; *p++ = '+'; *p++ = '\n';
    const int Q = (int)(ln.e3 - ln.s3);
    p = illumina ? copy_xform<XF_MINUS31>(p, in + ln.s3, Q) : copy_xform<XF_NONE>(p, in + ln.s3, Q);
    *p++ = '\n';
}

// Line index of this thread's chunks: packs the 16-bit newline masks of up to FT_MAXCPT
// chunks into mk[] and returns (newlines in the window) << 16 | (newlines in the owned tile).
// CAREFUL = the window reaches the end of the batch: bytes past n are masked and, on the
// final batch, an unterminated last line gets its virtual newline.
template <bool CAREFUL>
__device__ __forceinline__ uint32_t index_chunks(const uint8_t *sin, const TileGeom &g, uint32_t c0,
                                                 uint64_t ts, uint64_t n, int final, uint64_t (&mk)[FT_MAXCPT / 4])
{
    uint32_t cnt = 0;
    const uint32_t own = g.tile / 16;
#pragma unroll
    for (int i = 0; i < FT_MAXCPT; i++) {
        if (i < (int)g.cpt) {
            const uint32_t c = c0 + i;
            uint32_t m = 0;
            if (c < g.chunks) {
                if (!CAREFUL) {
                    m = nl_mask_of(*reinterpret_cast<const uint4 *>(sin + 16 * c));
                } else {
                    const uint64_t gpos = ts + 16ull * c;
                    if (gpos < n) {
                        m = nl_mask_of(*reinterpret_cast<const uint4 *>(sin + 16 * c));
                        const uint64_t rem = n - gpos;
                        if (rem < 16) m &= (1u << rem) - 1u;
                        // an unterminated last line: its last byte is dropped by line[:-1]
                        // (fastq.pyx:78,83,90), i.e. it plays the newline
                        if (final && rem <= 16) m |= 1u << (rem - 1);
                    }
                }
            }
            mk[i >> 2] |= (uint64_t)m << (16 * (i & 3));
            const uint32_t pc = __popc(m);
            cnt += (pc << 16) | (c < own ? pc : 0u);
        }
    }
    return cnt;
}

// ---- per-warp work list of the matcher (matcher.cuh: first_trip / extend_trips) -----------------
// The (read, period) pairs of a record warp whose copy-count scan continues are parked in shared
// memory and run lane-per-pair, 32 at a time after the last period, so the loop over copy numbers is executed by full
// warps once instead of once per period by the few lanes that hold a long repeat.  Results come
// back through a key (K*c << 16 | c << 3 | K: the tuple order of take_if_better) and a candidate
// mask per owner lane and strand.  The list lives in the staging buffer, which nothing else uses
// before phase F.
template <int W> struct WarpWork {
    static constexpr int ITEM_WORDS = W + 1, CAP = 160;   // 5 periods x 32 lanes: the list cannot overflow
    static constexpr int WORDS = CAP * ITEM_WORDS + 64 + 64 * W;   // items, 2 x 32 keys, 2 x 32 x W candidates
    uint32_t *items, *key, *cand;
    int count;                                              // warp-uniform
    __device__ __forceinline__ void init(uint32_t *base)
    {
        items = base; key = base + CAP * ITEM_WORDS; cand = key + 64; count = 0;
    }
    __device__ __forceinline__ void reset_keys()
    {
        const int lane = threadIdx.x & 31;
        key[lane] = 0; key[32 + lane] = 0;
        __syncwarp();
    }
    // run the top nitems (<= 32) pairs of the list, one per lane
    __device__ __forceinline__ void run(int nitems)
    {
        const int lane = threadIdx.x & 31;
        __syncwarp();
        const bool has = lane < nitems;
        Deferred<W> d;
        int cf = 0, cr = 0;
        Bits<W> df, dr;
#pragma unroll
        for (int i = 0; i < W; i++) { d.cur.w[i] = 0; df.w[i] = 0; dr.w[i] = 0; }
        d.meta = 0;
        if (has) {
            const uint32_t *it = items + (count - 1 - lane) * ITEM_WORDS;
#pragma unroll
            for (int i = 0; i < W; i++) d.cur.w[i] = it[i];
            d.meta = it[W];
            extend_trips<W>(d, cf, df, cr, dr);
        }
        const uint32_t owner = d.meta >> 27, K = d.meta & 7u;
        const uint32_t kf = (K * cf) << 16 | (uint32_t)cf << 3 | K, kr = (K * cr) << 16 | (uint32_t)cr << 3 | K;
        if (cf) atomicMax(&key[owner], kf);
        if (cr) atomicMax(&key[32 + owner], kr);
        __syncwarp();
        if (cf && key[owner] == kf) {
#pragma unroll
            for (int i = 0; i < W; i++) cand[owner * W + i] = df.w[i];
        }
        if (cr && key[32 + owner] == kr) {
#pragma unroll
            for (int i = 0; i < W; i++) cand[(32 + owner) * W + i] = dr.w[i];
        }
        count -= nitems;
        __syncwarp();
    }
    // called by every lane of the warp, once per period
    __device__ __forceinline__ void push(bool need, const Deferred<W> &d)
    {
        const uint32_t lane = threadIdx.x & 31;
        const unsigned m = __ballot_sync(0xFFFFFFFFu, need);
        if (!m) return;
        if (need) {
            uint32_t *it = items + (count + __popc(m & ((1u << lane) - 1u))) * ITEM_WORDS;
#pragma unroll
            for (int i = 0; i < W; i++) it[i] = d.cur.w[i];
            it[W] = d.meta | lane << 27;
        }
        count += __popc(m);
    }
    // after the last period: run what is left and fold the results into the owner's candidates
    __device__ __forceinline__ void finish(ScanState<W> &S)
    {
        const int lane = threadIdx.x & 31;
        while (count) run(count < 32 ? count : 32);
        const uint32_t kf = key[lane], kr = key[32 + lane];
        if (kf) {
            Bits<W> c;
#pragma unroll
            for (int i = 0; i < W; i++) c.w[i] = cand[lane * W + i];
            take_if_better(S.bf, c, (int)(kf & 7u), (int)(kf >> 3 & 0x1FFFu));
        }
        if (kr) {
            Bits<W> c;
#pragma unroll
            for (int i = 0; i < W; i++) c.w[i] = cand[(32 + lane) * W + i];
            take_if_better(S.br, c, (int)(kr & 7u), (int)(kr >> 3 & 0x1FFFu));
        }
        __syncwarp();
    }
};

// Both strands of the warp's 32 reads with a W-word window; every lane of the warp calls it
// (L = 0 for a lane without a read).
template <int W>
__device__ __forceinline__ void match_read_warp(const uint8_t *seq, int L, const ScanParams &prm, uint32_t *work,
                                                Hit &fwd, Hit &rc)
{
    WarpWork<W> wl;
    wl.init(work);
    wl.reset_keys();
    ScanState<W> S;
    scan_first<W>(seq, L, prm, S, [&](bool need, const Deferred<W> &d) { wl.push(need, d); });
    wl.finish(S);
    scan_finish<W>(S, L, prm, fwd, rc);
}

// words of shared memory one record warp needs for its work list (widest window compiled in)
template <int WMAX> __host__ __device__ constexpr uint32_t warp_work_words() { return (uint32_t)WarpWork<(WMAX >= 4 ? 4 : 2)>::WORDS; }

// Dispatch on the window width as match_read (matcher.cuh), with the warp work list.  `wmax` is
// warp-uniform; reads wider than the compiled windows take the exact byte path inline.
#ifndef BSTR_WORKLIST_MAXW
#define BSTR_WORKLIST_MAXW 4       // widest window whose continuing pairs go through the warp work list (A/B builds: 0, 2)
#endif
template <int WMAX>
__device__ __forceinline__ void match_read_tile(const uint8_t *seq, int L, const ScanParams &prm, int wmax,
                                                uint32_t *work, Hit &fwd, Hit &rc)
{
    if (wmax <= 2) {
        if (BSTR_WORKLIST_MAXW >= 2) match_read_warp<2>(seq, L, prm, work, fwd, rc);
        else match_read_bits<2>(seq, L, prm, fwd, rc);
    } else if (WMAX >= 4 && wmax <= 4) {
        if (BSTR_WORKLIST_MAXW >= 4) match_read_warp<(WMAX >= 4 ? 4 : 2)>(seq, L, prm, work, fwd, rc);
        else match_read_bits<(WMAX >= 4 ? 4 : 2)>(seq, L, prm, fwd, rc);
    }
    // 6- and 8-word windows keep the continuing pairs inline: with the scan state of such a window
    // live across the list run the kernel spills (180 / 504 bytes at the 168-register cap)
    else if (WMAX == 6 && wmax <= 6) match_read_bits<(WMAX == 6 ? 6 : 2)>(seq, L, prm, fwd, rc);
    else if (WMAX >= 8 && wmax <= 8) match_read_bits<(WMAX >= 8 ? 8 : 2)>(seq, L, prm, fwd, rc);
    else {
        fwd.k = rc.k = 0; fwd.copies = rc.copies = fwd.start = rc.start = 0;
        if (L > 0) match_read_bytes(seq, L, prm, fwd, rc);
    }
}

// CTA size limit per instantiation: batches of long reads have few records per tile, so the wide
// windows run with at most FT_WIDETHREADS threads and may keep 168 registers (the 8-word matcher
// spilled 344 bytes under the 128-register cap of 256-thread CTAs and ran 1.5x slower for it)
constexpr int FT_WIDETHREADS = 192;
#ifndef BSTR_WIDE_FROM
#define BSTR_WIDE_FROM 6
#endif
constexpr int FT_WIDEFROM = BSTR_WIDE_FROM;    // narrowest window that gets the smaller CTA limit
template <int WMAX> constexpr int ft_max_threads() { return WMAX >= FT_WIDEFROM ? FT_WIDETHREADS : FT_MAXTHREADS; }

template <int WMAX>
__global__ void __launch_bounds__(ft_max_threads<WMAX>(), BSTR_MINCTAS)
k_fused_tile(const uint8_t *__restrict__ in, uint64_t n, int final, ScanParams prm, int illumina,
             BatchState *st, unsigned long long *lb_lines, unsigned long long *lb_out,
             unsigned long long *ticket, uint32_t ntiles, ReadResult *res, uint8_t *out,
             uint64_t outcap, int exp_mask, const TileGeom g)
{
    extern __shared__ __align__(16) uint8_t fused_smem[];
    FusedVars &S = *reinterpret_cast<FusedVars *>(fused_smem);
    uint8_t *const in_raw = fused_smem + g.off_in;
    uint8_t *const sin = in_raw + 16;                        // tile bytes; 16 B of zero slack before
    uint8_t *const stage = fused_smem + g.off_stage;
    ReadResult *const s_rec = reinterpret_cast<ReadResult *>(fused_smem + g.off_rec);
    uint32_t *const s_outoff = reinterpret_cast<uint32_t *>(fused_smem + g.off_outoff);
    uint16_t *const s_nl = reinterpret_cast<uint16_t *>(fused_smem + g.off_nl);   // newline position + 1
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nthr = blockDim.x;
    if (tid == 0) {
        S.tile = (uint32_t)atomicAdd(ticket, 1ull);
        mbar_init(&S.mbar, 1);
    }
    if (tid < 4) {
        reinterpret_cast<uint32_t *>(in_raw)[tid] = 0u;
        reinterpret_cast<uint32_t *>(sin + g.region)[tid] = 0u;
        reinterpret_cast<uint32_t *>(sin + g.region + 16)[tid] = 0u;
    }
    __syncthreads();
    const uint32_t tile = S.tile;
    const uint64_t ts = (uint64_t)tile * g.tile;
    const bool reach = ts + g.region >= n;           // the staged window covers the batch end
    const bool last = tile == ntiles - 1;

    // ---- A: stage the tile (+ look-ahead) in shared memory: ONE bulk copy (TMA, cp.async.bulk)
    //         issued by one thread, completion on an mbarrier; every input byte is read from HBM
    //         once.  Past the batch end the device buffer holds >= 16 zero bytes (IN_SLACK), the
    //         rest of the window is zero-filled here.
    const uint32_t nb = reach ? (uint32_t)((n - ts + 15) & ~15ull) : g.region;   // bytes to fetch
    if (tid == 0) bulk_load(sin, in + ts, nb, &S.mbar);
    for (uint32_t c = nb / 16 + tid; c < g.chunks; c += nthr)
        *reinterpret_cast<uint4 *>(sin + 16 * c) = make_uint4(0, 0, 0, 0);
    mbar_wait(&S.mbar, 0);
    __syncthreads();

    // ---- B: line index of the window -----------------------------------------------------
    uint64_t mk[FT_MAXCPT / 4] = {0, 0, 0, 0};
    const uint32_t c0 = tid * g.cpt;
    const uint32_t cnt = reach ? index_chunks<true>(sin, g, c0, ts, n, final, mk)
                               : index_chunks<false>(sin, g, c0, ts, n, final, mk);   // all << 16 | own
    uint32_t tot;
    const uint32_t ex = block_excl_scan(cnt, S.warp_tmp, &tot);
    const uint32_t shift = tile == 0 ? 1u : 0u;      // tile 0: virtual newline before byte 0
    const uint32_t real_own = tot & 0xFFFFu;
    uint32_t cnt_own = real_own + shift, cnt_all = (tot >> 16) + shift;
    bool bail = cnt_all + 2 > g.nlmax;
    if ((nthr / 32 - 1) * warp_work_words<WMAX>() * 4u > g.stage) bail = true;   // no room for the matcher's work lists
    {
        uint32_t idx = (ex >> 16) + shift;
#pragma unroll
        for (int w = 0; w < FT_MAXCPT / 4; w++) {
            uint64_t m = mk[w];
            const uint32_t base = 16u * (c0 + 4 * w) + 1u;
            while (m) {
                const int b = __ffsll((long long)m) - 1;
                m &= m - 1;
                if (idx < g.nlmax) s_nl[idx] = (uint16_t)(base + b);
                idx++;
            }
        }
        if (tid == 0 && shift) s_nl[0] = 0;
    }

    // ---- C: which local lines start records?  That needs the global line number of the tile's
    //         first line (look-back #1) only modulo 4, and a valid FASTQ tile gives the phase
    //         away: a header line starts with '@' and the line two below with '+', while a
    //         non-empty sequence line can start with neither.  If exactly one phase fits the
    //         first two records of the window, the records are processed with it at once and
    //         warp 0 -- the scan warp, it owns no records -- resolves the look-back meanwhile;
    //         the guess is verified afterwards (mismatch: the batch goes to the generic
    //         kernels).  Tiles that touch the batch end, and tile 0, take no guess.
    if (tid == 0) lookback_publish(lb_lines, (int)tile, real_own);
    __syncthreads();                                  // s_nl complete
    uint32_t phi = 0, nfit = 0;
    if (!reach && tile != 0 && !bail && !(exp_mask & 64)) {
#pragma unroll
        for (uint32_t p = 0; p < 4; p++) {
            bool fit = p + 7 < cnt_all;               // two whole records visible
            if (fit)
                fit = sin[s_nl[p]] == '@' && sin[s_nl[p + 2]] == '+' && sin[s_nl[p + 4]] == '@' && sin[s_nl[p + 6]] == '+';
            if (fit) { phi = p; nfit++; }
        }
    }
    const bool spec = nfit == 1;                      // CTA-uniform: every thread reads the same bytes
    if (!spec) {
        if (wid == 0) {
            const unsigned long long base = lookback_resolve(lb_lines, (int)tile, real_own);
            if (lane == 0) S.gbase = tile == 0 ? -1ll : (long long)base;
        }
        __syncthreads();
    }
    long long gbase = spec ? 0ll : S.gbase;           // spec: known after phase D
    const long long G = gbase + (long long)cnt_all;   // lines of the whole batch, if `reach` (never with spec)
    if (final && reach && (G & 3) == 3 && !bail) {
        // EOF after the separator line: accepted with an empty quality string
        // (readline() returns '' at fastq.pyx:89)
        if (tid == 0) s_nl[cnt_all] = (uint16_t)(n - ts + 1);
        cnt_all++;
    }
    if (last && tid == 0) {
        const unsigned long long lines = (unsigned long long)G + ((final && (G & 3) == 3) ? 1ull : 0ull);
        st->n_lines = lines;
        st->n_records = lines >> 2;
        if (final) {
            if ((G & 3) == 1 || (G & 3) == 2)        // file ends inside a record (fastq.pyx:86)
                atomicMin(&st->first_bad, ((unsigned long long)(G >> 2) << 8) | BAD_TRUNCATED);
            atomicMax(&st->consumed, (unsigned long long)n);
        }
    }
    const uint32_t i0 = spec ? phi : (uint32_t)((3 - gbase) & 3);
    const uint32_t nrec_own = cnt_own > i0 ? (cnt_own - i0 + 3) / 4 : 0u;
    uint32_t nproc = cnt_all >= i0 + 5 ? (cnt_all - 1 - i0) / 4 : 0u;
    if (nproc > nrec_own) nproc = nrec_own;
    if (nproc < nrec_own && !reach) bail = true;      // a record runs past the look-ahead
    if (nproc > g.recmax) bail = true;
    if (!bail && final && reach && nproc < nrec_own && tid == 0) {
        // a record cut short by EOF: the reference still tests its header (fastq.pyx:77) and a
        // complete second line (fastq.pyx:81) before the missing line raises (fastq.pyx:86)
        const uint32_t i = i0 + 4 * nproc, after = cnt_all - 1 - i;
        if (after >= 1) {
            const unsigned long long r = (unsigned long long)(G >> 2);
            if (sin[s_nl[i]] != '@') atomicMin(&st->first_bad, (r << 8) | BAD_HEADER);
            else if (after >= 2 && bases_invalid(sin + s_nl[i + 1], (int)s_nl[i + 2] - 1 - (int)s_nl[i + 1]))
                atomicMin(&st->first_bad, (r << 8) | BAD_BASES);
        }
    }
    if (bail) nproc = 0;
    __syncthreads();                                  // s_nl complete (virtual entry included)

    // ---- D: per record: fastq.next() checks + both-strand match (thread per record; warp 0
    //         resolves look-back #1 instead).  A failed check is kept in the record's slot
    //         (fck = 0, rstart = code) and reported in phase F, when the record index is known.
    const uint32_t rthr = nthr - 32, rtid = tid - 32;
    uint32_t nsel = 0;
    if (wid == 0) {
        if (spec) {
            const unsigned long long base = lookback_resolve(lb_lines, (int)tile, real_own);
            if (lane == 0) S.gbase = (long long)base;
        }
        for (uint32_t j = nproc + lane; j < g.recmax; j += 32) s_outoff[j] = 0;
    } else {
        uint32_t *const work = reinterpret_cast<uint32_t *>(stage) + (wid - 1) * warp_work_words<WMAX>();
        for (uint32_t jb = 0; jb < nproc; jb += rthr) {
            const uint32_t j = jb + rtid;
            const bool live = j < nproc;
            RecordLines ln = {};
            int L = 0, bad = 0;
            if (live) {
                ln = tile_record_lines(s_nl, (int)(i0 + 4 * j)); L = (int)(ln.e1 - ln.s1);
                if (sin[ln.s0] != '@') bad = BAD_HEADER;                       // fastq.pyx:77
                else if (!(exp_mask & 2) && bases_invalid(sin + ln.s1, L)) bad = BAD_BASES;        // fastq.pyx:81
                else if (sin[ln.s2] != '+') bad = BAD_SEPARATOR;                // fastq.pyx:86
            }
            const int wmax = __reduce_max_sync(0xFFFFFFFFu, (unsigned)window_words(L, prm));
            Hit f, rc;
            // every lane of the warp takes part (the work list votes); a lane without a good record scans nothing
            const int Lm = (live && !bad && !(exp_mask & 1)) ? L : 0;
            match_read_tile<WMAX>(sin + ln.s1, Lm, prm, wmax, work, f, rc);
            const bool sel = f.k && rc.k;
            if (live) {
                ReadResult o;
                o.fstart = sel ? f.start : 0; o.rstart = sel ? rc.start : (uint32_t)bad;
                o.fck = sel ? ((uint32_t)f.copies << 3 | f.k) : 0;
                o.rck = sel ? ((uint32_t)rc.copies << 3 | rc.k) : 0;
                s_rec[j] = o;
                s_outoff[j] = sel ? record_out_len(ln.e0 - ln.s0 - 1, L, ln.e3 - ln.s3, f, rc) : 0u;
            }
            nsel += __popc(__ballot_sync(0xFFFFFFFFu, sel));
        }
    }
    __syncthreads();
    if (spec) {
        gbase = S.gbase;
        if ((uint32_t)((3 - gbase) & 3) != i0) bail = true;     // wrong guess: generic kernels redo the batch
    }
    if (!bail && nrec_own && tid == 0) {
        const uint32_t iL = i0 + 4 * (nrec_own - 1);
        if (!(shift && iL == 0)) atomicMax(&st->consumed, ts + s_nl[iL]);
    }

    // ---- E: output offsets of the records inside the tile; the tile's byte count is published
    //         at once, the look-back over the earlier tiles is resolved AFTER formatting: a tile
    //         needs the totals of every tile before it, i.e. it waits for the slowest matcher
    //         among several hundred tiles in flight (11 % of the kernel when waited for here)
    const uint32_t o0 = s_outoff[2 * tid], o1 = s_outoff[2 * tid + 1];
    uint32_t total;
    const uint32_t oex = block_excl_scan(o0 + o1, S.warp_tmp, &total);
    s_outoff[2 * tid] = oex;
    s_outoff[2 * tid + 1] = oex + o0;
    if (bail) { total = 0; nproc = 0; }
    if (total > g.stage - 32 && !(exp_mask & 16)) { bail = true; total = 0; nproc = 0; }   // bit 16: occupancy experiments with a dummy stage (needs bits 4|8)
    if (tid == 0) {
        if (bail) atomicExch(&st->fallback, 1ull);
        lookback_publish(lb_out, (int)tile, total);
    }
    if (lane == 0 && nsel && !bail && wid != 0) atomicAdd(&st->n_selected, (unsigned long long)nsel);
    __syncthreads();

    // ---- F: PrintSequence into the staging buffer (dense, from offset 0) ---------------------
    if (wid != 0)
        for (uint32_t j = rtid; j < nproc; j += rthr) {
            const ReadResult rr = s_rec[j];
            const unsigned long long r = (unsigned long long)((gbase + (long long)(i0 + 4 * j) + 1) >> 2);
            if (res) {
                ReadResult o = rr;
                if (o.fck == 0) o.rstart = 0;
                res[r] = o;
            }
            if (rr.fck == 0) {
                if (rr.rstart) atomicMin(&st->first_bad, (r << 8) | rr.rstart);
                continue;
            }
            if (exp_mask & 4) continue;
            const RecordLines ln = tile_record_lines(s_nl, (int)(i0 + 4 * j));
            format_record_words(sin, ln, rr, illumina, stage + s_outoff[j]);
        }
    if (wid == 0) {
        // exp bit 32 (timing experiments only): no look-back wait, tiles write at disjoint fake offsets
        const unsigned long long ob = (exp_mask & 32) ? (unsigned long long)tile * g.stage
                                                      : lookback_resolve(lb_out, (int)tile, total);
        if (lane == 0) {
            S.out_base = ob;
            if (last) st->out_bytes = ob + total;
        }
    }
    __syncthreads();
    const unsigned long long out_base = S.out_base;
    if (total == 0 || out_base + total > outcap) return;     // host grows the buffer and re-runs

    // ---- G: flush the staging buffer: aligned 16-byte stores; the source is shifted by the
    //         destination's phase (uniform over the CTA, so the word select does not diverge) ----
    if (exp_mask & 8) return;
    uint8_t *dst = out + out_base;
    const uint32_t head = min(total, (16u - (uint32_t)(out_base & 15)) & 15u);
    if (tid < head) dst[tid] = stage[tid];
    const uint32_t nchunk = (total - head) >> 4;
    const uint32_t q = head >> 2, r8 = 8 * (head & 3);
#define BSTR_FLUSH(Q0, W0, W1, W2, W3, W4)                                                          \
    if (q == Q0)                                                                                    \
        for (uint32_t c = tid; c < nchunk; c += nthr) {                                             \
            const uint4 A = *reinterpret_cast<const uint4 *>(stage + 16 * c);                       \
            const uint4 B = *reinterpret_cast<const uint4 *>(stage + 16 * c + 16);                  \
            uint4 o;                                                                                \
            o.x = __funnelshift_r(W0, W1, r8); o.y = __funnelshift_r(W1, W2, r8);                   \
            o.z = __funnelshift_r(W2, W3, r8); o.w = __funnelshift_r(W3, W4, r8);                   \
            *reinterpret_cast<uint4 *>(dst + head + 16 * c) = o;                                    \
        }
    BSTR_FLUSH(0, A.x, A.y, A.z, A.w, B.x)
    BSTR_FLUSH(1, A.y, A.z, A.w, B.x, B.y)
    BSTR_FLUSH(2, A.z, A.w, B.x, B.y, B.z)
    BSTR_FLUSH(3, A.w, B.x, B.y, B.z, B.w)
#undef BSTR_FLUSH
    const uint32_t done = head + (nchunk << 4);
    if (tid < total - done) dst[done + tid] = stage[done + tid];
}

}  // namespace bstr
