// Parallel gzip inflate (see pinflate.h).  Deflate decoding follows RFC 1951, the gzip container
// RFC 1952; nothing here is taken from the reference (it calls Python's gzip, src/fastq.pyx:50-51).
#include "pinflate.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>                      // crc32, crc32_combine only

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <thread>

#if defined(__x86_64__)
#include <immintrin.h>
#elif defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace pinflate {
namespace {

static double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------
// Huffman tables.  One 32-bit entry per table slot:
//   bits 0-3   code bits to drop (for a link: the primary index width)
//   bits 4-7   extra bits to read (for a link: index width of the secondary table)
//   bits 8-10  kind
//   bits 16-31 literal byte / length base / distance base / offset of the secondary table;
//              for K_LIT2 two literal bytes (first in the low byte) whose codes together fit the
//              primary index - sequence and quality lines are runs of 2..6-bit literals
// ---------------------------------------------------------------------------------------------
enum : uint32_t { K_LIT = 0, K_BASE = 1, K_EOB = 2, K_LINK = 3, K_LIT2 = 4, K_BAD = 5 };   // (kind & 3) == 0: literals
static inline uint32_t mk(uint32_t kind, uint32_t nbits, uint32_t extra, uint32_t val) { return nbits | (extra << 4) | (kind << 8) | (val << 16); }
static inline uint32_t e_nbits(uint32_t e) { return e & 15; }
static inline uint32_t e_extra(uint32_t e) { return (e >> 4) & 15; }
static inline uint32_t e_kind(uint32_t e) { return (e >> 8) & 7; }
static inline uint32_t e_val(uint32_t e) { return e >> 16; }

static const int LB = 11, DB = 8, CB = 7;         // primary index widths: literal/length, distance, code-length code
static const int LIT_CAP = (1 << LB) + 2400, DIST_CAP = (1 << DB) + 2048;

static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

enum TableKind { T_CODELEN, T_LITLEN, T_DIST };

static inline uint32_t symbol_entry(TableKind t, int s)
{
    if (t == T_CODELEN) return mk(K_LIT, 0, 0, (uint32_t)s);
    if (t == T_DIST) return s < 30 ? mk(K_BASE, 0, dist_extra[s], dist_base[s]) : mk(K_BAD, 0, 0, 0);
    if (s < 256) return mk(K_LIT, 0, 0, (uint32_t)s);
    if (s == 256) return mk(K_EOB, 0, 0, 0);
    if (s < 286) return mk(K_BASE, 0, len_extra[s - 257], len_base[s - 257]);
    return mk(K_BAD, 0, 0, 0);
}

static inline uint32_t bit_reverse(uint32_t c, int n)
{
    uint32_t r = 0;
    for (int i = 0; i < n; i++) { r = (r << 1) | (c & 1); c >>= 1; }
    return r;
}

// Canonical code from code lengths.  Returns 0 (complete), 1 (no code at all, or the one-code
// case zlib also accepts), -1 (over-subscribed / incomplete / too large).
static int build_table(uint32_t *tab, int cap, int pb, const uint8_t *lens, int n, TableKind t)
{
    int count[16] = {0};
    for (int i = 0; i < n; i++) count[lens[i]]++;
    int maxlen = 15;
    while (maxlen > 0 && !count[maxlen]) maxlen--;
    const uint32_t bad = mk(K_BAD, 0, 0, 0);
    if (maxlen == 0) {
        for (int i = 0; i < (1 << pb); i++) tab[i] = bad;
        return t == T_DIST ? 1 : -1;
    }
    int left = 1;
    for (int l = 1; l <= 15; l++) { left = (left << 1) - count[l]; if (left < 0) return -1; }
    if (left > 0 && (t == T_CODELEN || maxlen != 1)) return -1;
    if (left > 0 || maxlen > pb) for (int i = 0; i < (1 << pb); i++) tab[i] = bad;      // else every slot is written below
    uint32_t next[16], code = 0;
    count[0] = 0;
    for (int l = 1; l <= 15; l++) { code = (code + (uint32_t)count[l - 1]) << 1; next[l] = code; }
    uint16_t rev[288];
    for (int s = 0; s < n; s++) if (lens[s]) rev[s] = (uint16_t)bit_reverse(next[lens[s]]++, lens[s]);
    const uint32_t pmask = (1u << pb) - 1;
    int used = 1 << pb;
    if (maxlen > pb) {
        uint8_t deepest[1 << LB];
        memset(deepest, 0, (size_t)1 << pb);
        for (int s = 0; s < n; s++) if (lens[s] > pb) { uint8_t &d = deepest[rev[s] & pmask]; if (lens[s] > d) d = lens[s]; }
        for (int s = 0; s < n; s++) {
            if (lens[s] <= pb) continue;
            const uint32_t p = rev[s] & pmask;
            if (e_kind(tab[p]) == K_LINK) continue;
            const int sb = deepest[p] - pb, sz = 1 << sb;
            if (used + sz > cap) return -1;
            tab[p] = mk(K_LINK, (uint32_t)pb, (uint32_t)sb, (uint32_t)used);
            for (int k = 0; k < sz; k++) tab[used + k] = bad;
            used += sz;
        }
    }
    for (int s = 0; s < n; s++) {
        const int l = lens[s];
        if (!l) continue;
        uint32_t e = symbol_entry(t, s);
        if (l <= pb) {
            e |= (uint32_t)l;
            for (uint32_t i = rev[s]; i < (1u << pb); i += 1u << l) tab[i] = e;
        } else {
            const uint32_t link = tab[rev[s] & pmask];
            const uint32_t off = e_val(link), sb = e_extra(link);
            e |= (uint32_t)(l - pb);
            for (uint32_t i = (uint32_t)rev[s] >> pb; i < (1u << sb); i += 1u << (l - pb)) tab[off + i] = e;
        }
    }
    if (t == T_LITLEN) {                                         // pair up short literals
        uint32_t one[1 << LB];
        memcpy(one, tab, sizeof(one));
        for (uint32_t i = 0; i < (1u << LB); i++) {
            const uint32_t a = one[i];
            if (e_kind(a) != K_LIT || e_nbits(a) >= (uint32_t)LB) continue;
            const uint32_t b = one[i >> e_nbits(a)];
            if (e_kind(b) != K_LIT || e_nbits(a) + e_nbits(b) > (uint32_t)LB) continue;
            tab[i] = mk(K_LIT2, e_nbits(a) + e_nbits(b), 0, e_val(a) | e_val(b) << 8);
        }
    }
    return left > 0 ? 1 : 0;
}

struct Huff {
    uint32_t lit[LIT_CAP];
    uint32_t dist[DIST_CAP];
};

static Huff *fixed_tables()
{
    static Huff *h = [] {
        Huff *f = new Huff;
        uint8_t l[288];
        for (int i = 0; i < 288; i++) l[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
        build_table(f->lit, LIT_CAP, LB, l, 288, T_LITLEN);
        for (int i = 0; i < 32; i++) l[i] = 5;
        build_table(f->dist, DIST_CAP, DB, l, 32, T_DIST);
        return f;
    }();
    return h;
}

// ---------------------------------------------------------------------------------------------
// Bit reader: LSB-first, 64-bit window.  Past the end of the input it feeds zero bytes (a few)
// and then reports failure, so a decoder can never run away.
// ---------------------------------------------------------------------------------------------
struct BitReader {
    const uint8_t *in;
    size_t size, pos;          // pos: next byte to load (may run past size: padding)
    uint64_t buf;
    int cnt;
    void init(const uint8_t *data, size_t n, uint64_t bit)
    {
        in = data; size = n; pos = (size_t)(bit >> 3); buf = 0; cnt = 0;
        refill();
        drop((int)(bit & 7));
    }
    inline bool refill()
    {
        if (pos + 8 <= size) {
            uint64_t w;
            memcpy(&w, in + pos, 8);
            buf |= w << cnt;
            pos += (size_t)((63 - cnt) >> 3);
            cnt |= 56;
            return true;
        }
        while (cnt <= 56) {
            if (pos < size) buf |= (uint64_t)in[pos] << cnt;
            else if (pos > size + 16) return false;
            pos++; cnt += 8;
        }
        return true;
    }
    inline void drop(int n) { buf >>= n; cnt -= n; }
    inline uint32_t take(int n) { const uint32_t v = (uint32_t)(buf & ((1ull << n) - 1)); drop(n); return v; }
    inline uint64_t bitpos() const { return (uint64_t)pos * 8 - (uint64_t)cnt; }
    inline bool overrun() const { return bitpos() > (uint64_t)size * 8; }
    void align_byte() { drop(cnt & 7); }
};

static const uint8_t codelen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

// Dynamic block header (after the 3 block bits).  strict: only what a sane compressor writes
// (used when the position is a guess).  false: invalid.
static bool read_dynamic_header(BitReader &br, Huff &h, bool strict)
{
    if (!br.refill()) return false;
    const int hlit = (int)br.take(5) + 257, hdist = (int)br.take(5) + 1, hclen = (int)br.take(4) + 4;
    if (hlit > 286 || hdist > 30) return false;
    uint8_t cl[19] = {0};
    for (int i = 0; i < hclen; i++) {
        if (br.cnt < 3 && !br.refill()) return false;
        cl[codelen_order[i]] = (uint8_t)br.take(3);
    }
    uint32_t ctab[1 << CB];
    if (build_table(ctab, 1 << CB, CB, cl, 19, T_CODELEN) != 0) return false;
    uint8_t lens[286 + 30];
    const int total = hlit + hdist;
    int i = 0;
    while (i < total) {
        if (br.cnt < 14 && !br.refill()) return false;
        const uint32_t e = ctab[br.buf & ((1u << CB) - 1)];
        if (e_kind(e) != K_LIT) return false;
        br.drop((int)e_nbits(e));
        const uint32_t s = e_val(e);
        if (s < 16) { lens[i++] = (uint8_t)s; continue; }
        int rep; uint8_t v = 0;
        if (s == 16) { if (i == 0) return false; v = lens[i - 1]; rep = 3 + (int)br.take(2); }
        else if (s == 17) rep = 3 + (int)br.take(3);
        else rep = 11 + (int)br.take(7);
        if (i + rep > total) return false;
        while (rep--) lens[i++] = v;
    }
    if (lens[256] == 0) return false;
    const int rl = build_table(h.lit, LIT_CAP, LB, lens, hlit, T_LITLEN);
    if (rl < 0 || (strict && rl != 0)) return false;
    const int rd = build_table(h.dist, DIST_CAP, DB, lens + hlit, hdist, T_DIST);
    if (rd < 0) return false;
    return !br.overrun();
}

// ---------------------------------------------------------------------------------------------
// Symbols of one decoded range: 0..255 a byte, 0x8000|i the byte i of the (unknown) 32 KiB before
// the range (i = 32767 is the byte right before it).
// ---------------------------------------------------------------------------------------------
struct MemberEnd { uint64_t off; uint32_t crc, isize; };      // off: symbols of this range before the member ended

// Symbol buffers go round: a run takes one that a finished run gave back, so the pages stay
// mapped and warm instead of being faulted in again for every range.
struct SymPool {
    std::mutex mu;
    std::vector<std::pair<uint16_t *, size_t>> idle;
    size_t keep = 8;
    ~SymPool() { for (auto &b : idle) free(b.first); }
    bool take(uint16_t **p, size_t *cap)
    {
        std::lock_guard<std::mutex> g(mu);
        if (idle.empty()) return false;
        size_t best = 0;
        for (size_t i = 1; i < idle.size(); i++) if (idle[i].second > idle[best].second) best = i;
        *p = idle[best].first; *cap = idle[best].second;
        idle[best] = idle.back(); idle.pop_back();
        return true;
    }
    void give(uint16_t *p, size_t cap)
    {
        if (!p) return;
        {
            std::lock_guard<std::mutex> g(mu);
            if (idle.size() < keep && cap <= ((size_t)64 << 20)) { idle.emplace_back(p, cap); return; }
        }
        free(p);
    }
};

struct Run {
    uint16_t *sym = nullptr;
    size_t n = 0, cap = 0;
    uint64_t start_bit = 0, end_bit = 0;
    bool stream_end = false;
    std::vector<MemberEnd> ends;
    SymPool *pool = nullptr;
    explicit Run(SymPool *p = nullptr) : pool(p) {}
    ~Run() { if (pool) pool->give(sym, cap); else free(sym); }
    bool reserve(size_t extra)
    {
        if (n + extra <= cap) return true;
        if (!sym && pool && pool->take(&sym, &cap) && n + extra <= cap) return true;
        size_t nc = cap ? cap + cap / 2 : (size_t)1 << 20;
        while (nc < n + extra) nc += nc / 2;
        uint16_t *p = (uint16_t *)realloc(sym, nc * sizeof(uint16_t));
        if (!p) return false;
        sym = p; cap = nc;
        return true;
    }
};

static uint8_t texty[256];             // bytes a FASTQ file is made of (guessed block starts must decode to these)
static const bool texty_init = [] {
    for (int c = 0; c < 256; c++) texty[c] = (c >= 32 && c < 127) || c == '\n' || c == '\t' || c == '\r';
    return true;
}();

enum { D_OK = 0, D_ERROR = -1, D_REJECT = -2 };

// Copy a match of `len` symbols from `dist` back; positions before the start of the run become markers.
static inline void copy_match(uint16_t *base, uint16_t *out, uint32_t len, uint32_t dist)
{
    if (dist > (size_t)(out - base)) {                               // reaches into the unknown history
        uint32_t idx = 32768u - (dist - (uint32_t)(out - base));
        while (len && idx < 32768u) { *out++ = (uint16_t)(0x8000u | idx++); len--; }
    }
    const uint16_t *src = out - dist;
    if (dist >= 4) {
        for (uint32_t j = 0; j < len; j += 4) memcpy(out + j, src + j, 8);      // may write 3 symbols too many: the slack covers it
    } else {
        for (uint32_t j = 0; j < len; j++) out[j] = src[j];
    }
}

template <bool STRICT>
static int inflate_block(BitReader &br, const Huff &h, Run &o)
{
    const uint32_t LMASK = (1u << LB) - 1, DMASK = (1u << DB) - 1;
    for (;;) {
        if (o.n + 320 > o.cap && !o.reserve(4096)) return D_ERROR;
        // fast loop: whole words of input left, room for output; state in locals
        if (br.pos + 8 <= br.size) {
            const uint8_t *in = br.in;
            const size_t fast_end = br.size - 8;
            uint64_t buf = br.buf;
            int cnt = br.cnt;
            size_t pos = br.pos;
            uint16_t *const base = o.sym, *out = o.sym + o.n, *const out_lim = o.sym + o.cap - 320;
            int rc = 1;                                               // 1: left the fast loop, keep going below
            while (pos <= fast_end && out < out_lim) {
                uint64_t w;
                memcpy(&w, in + pos, 8);
                buf |= w << cnt;
                pos += (size_t)((63 - cnt) >> 3);
                cnt |= 56;
                uint32_t e = h.lit[buf & LMASK];
                if ((e & 0x300) == 0) {                               // one or two literals with short codes; up to three such entries per refill
#define PINFLATE_LITERALS()                                                                                      \
    {                                                                                                            \
        const uint32_t v = e >> 16, two = (v & 0xff) | (v >> 8) << 16;                                           \
        if (STRICT && !(texty[v & 0xff] && ((e & 0x400) == 0 || texty[v >> 8]))) { rc = D_REJECT; break; }       \
        memcpy(out, &two, 4);                                                                                    \
        out += 1 + ((e >> 10) & 1);                                                                              \
        buf >>= (e & 15); cnt -= (int)(e & 15);                                                                  \
    }
                    PINFLATE_LITERALS();
                    e = h.lit[buf & LMASK];
                    if ((e & 0x300) != 0) continue;
                    PINFLATE_LITERALS();
                    e = h.lit[buf & LMASK];
                    if ((e & 0x300) != 0) continue;
                    PINFLATE_LITERALS();
                    continue;
                }
                if (e_kind(e) == K_LINK) { buf >>= LB; cnt -= LB; e = h.lit[e_val(e) + (buf & ((1u << e_extra(e)) - 1))]; }
                buf >>= e_nbits(e); cnt -= (int)e_nbits(e);
                const uint32_t kind = e_kind(e);
                if (kind == K_LIT) {
                    if (STRICT && !texty[e_val(e)]) { rc = D_REJECT; break; }
                    *out++ = (uint16_t)e_val(e);
                    continue;
                }
                if (kind == K_BASE) {
                    const uint32_t xl = e_extra(e);
                    const uint32_t len = e_val(e) + (uint32_t)(buf & ((1u << xl) - 1));
                    buf >>= xl; cnt -= (int)xl;
                    uint32_t d = h.dist[buf & DMASK];
                    if (e_kind(d) == K_LINK) { buf >>= DB; cnt -= DB; d = h.dist[e_val(d) + (buf & ((1u << e_extra(d)) - 1))]; }
                    if (e_kind(d) != K_BASE) { rc = STRICT ? D_REJECT : D_ERROR; break; }
                    buf >>= e_nbits(d); cnt -= (int)e_nbits(d);
                    const uint32_t xd = e_extra(d);
                    const uint32_t dist = e_val(d) + (uint32_t)(buf & ((1u << xd) - 1));
                    buf >>= xd; cnt -= (int)xd;
                    copy_match(base, out, len, dist);
                    out += len;
                    continue;
                }
                rc = kind == K_EOB ? D_OK : (STRICT ? D_REJECT : D_ERROR);
                break;
            }
            br.buf = buf; br.cnt = cnt; br.pos = pos;
            o.n = (size_t)(out - base);
            if (rc == D_OK) return br.overrun() ? (STRICT ? D_REJECT : D_ERROR) : D_OK;
            if (rc != 1) return rc;
            if (o.n + 320 > o.cap) continue;                           // out of room: grow, then back to the fast loop
        }
        // careful path: the last few bytes of the input
        if (!br.refill()) return D_ERROR;
        uint32_t e = h.lit[br.buf & LMASK];
        if (e_kind(e) == K_LINK) { br.drop(LB); e = h.lit[e_val(e) + (br.buf & ((1u << e_extra(e)) - 1))]; }
        br.drop((int)e_nbits(e));
        const uint32_t kind = e_kind(e);
        if (kind == K_LIT) {
            if (STRICT && !texty[e_val(e)]) return D_REJECT;
            o.sym[o.n++] = (uint16_t)e_val(e);
            continue;
        }
        if (kind == K_LIT2) {
            if (STRICT && !(texty[e_val(e) & 0xff] && texty[e_val(e) >> 8])) return D_REJECT;
            o.sym[o.n++] = (uint16_t)(e_val(e) & 0xff);
            o.sym[o.n++] = (uint16_t)(e_val(e) >> 8);
            continue;
        }
        if (kind == K_BASE) {
            const uint32_t len = e_val(e) + br.take((int)e_extra(e));
            uint32_t d = h.dist[br.buf & DMASK];
            if (e_kind(d) == K_LINK) { br.drop(DB); d = h.dist[e_val(d) + (br.buf & ((1u << e_extra(d)) - 1))]; }
            if (e_kind(d) != K_BASE) return STRICT ? D_REJECT : D_ERROR;
            br.drop((int)e_nbits(d));
            const uint32_t dist = e_val(d) + br.take((int)e_extra(d));
            copy_match(o.sym, o.sym + o.n, len, dist);
            o.n += len;
            continue;
        }
        if (kind == K_EOB) return br.overrun() ? (STRICT ? D_REJECT : D_ERROR) : D_OK;
        return STRICT ? D_REJECT : D_ERROR;
    }
}

static inline uint32_t le32(const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }

// gzip member header at data[at...].  Returns the offset of the deflate stream, 0 if this is
// not a gzip header (or it is cut short).
static size_t gzip_header(const uint8_t *d, size_t size, size_t at)
{
    if (at + 18 > size || d[at] != 0x1f || d[at + 1] != 0x8b || d[at + 2] != 8) return 0;
    const uint8_t flg = d[at + 3];
    if (flg & 0xe0) return 0;
    size_t p = at + 10;
    if (flg & 4) { if (p + 2 > size) return 0; p += 2 + ((size_t)d[p] | (size_t)d[p + 1] << 8); }
    if (flg & 8) { while (p < size && d[p]) p++; p++; }
    if (flg & 16) { while (p < size && d[p]) p++; p++; }
    if (flg & 2) p += 2;
    return p < size ? p : 0;
}

// Decode deflate blocks from `bit` (a block start) up to the first block boundary at or after
// stop_bit, crossing gzip member boundaries on the way.  guessed: the first block must pass the
// strict checks (D_REJECT if not).  On D_OK o.end_bit is where the next block starts, or
// o.stream_end is set.
static int decode_run(const uint8_t *in, size_t size, uint64_t bit, uint64_t stop_bit, bool guessed, size_t soft_cap, Run &o,
                      const std::atomic<bool> &quit)
{
    std::unique_ptr<Huff> dyn(new Huff);
    BitReader br;
    br.init(in, size, bit);
    o.start_bit = bit;
    bool first = true;
    for (;;) {
        const uint64_t here = br.bitpos();
        if (!first && (here >= stop_bit || o.n >= soft_cap)) { o.end_bit = here; return D_OK; }
        if (quit.load(std::memory_order_relaxed)) return D_ERROR;
        if (!br.refill()) return D_ERROR;
        const uint32_t bfinal = br.take(1), btype = br.take(2);
        const bool strict = guessed && first;
        int r;
        if (btype == 0) {
            if (strict) return D_REJECT;
            br.align_byte();
            if (br.cnt < 32 && !br.refill()) return D_ERROR;
            const uint32_t len = br.take(16), nlen = br.take(16);
            if ((len ^ nlen) != 0xffffu) return D_ERROR;
            size_t at = (size_t)(br.bitpos() >> 3);
            if (at + len > size) return D_ERROR;
            if (!o.reserve(len + 320)) return D_ERROR;
            for (uint32_t j = 0; j < len; j++) o.sym[o.n + j] = in[at + j];
            o.n += len;
            br.init(in, size, (uint64_t)(at + len) * 8);
            r = D_OK;
        } else if (btype == 1) {
            if (strict) return D_REJECT;
            r = inflate_block<false>(br, *fixed_tables(), o);
        } else if (btype == 2) {
            if (!read_dynamic_header(br, *dyn, strict)) return strict ? D_REJECT : D_ERROR;
            r = strict ? inflate_block<true>(br, *dyn, o) : inflate_block<false>(br, *dyn, o);
        } else
            return strict ? D_REJECT : D_ERROR;
        if (r != D_OK) return r;
        first = false;
        if (!bfinal) continue;
        // end of a gzip member: trailer, then maybe another member
        br.align_byte();
        size_t at = (size_t)(br.bitpos() >> 3);
        if (at + 8 > size) return D_ERROR;
        o.ends.push_back(MemberEnd{(uint64_t)o.n, le32(in + at), le32(in + at + 4)});
        at += 8;
        while (at < size && in[at] == 0) at++;                 // zero padding between members (tape blocks); gzip(1) accepts it
        const size_t body = at < size ? gzip_header(in, size, at) : 0;
        if (!body) {
            // end of file -- or bytes that are no gzip member: Python's gzip module, which the
            // reference reads through (fastq.pyx:50-51), raises "Not a gzipped file" for them
            if (at < size) return strict ? D_REJECT : D_ERROR;
            o.stream_end = true;
            o.end_bit = (uint64_t)size * 8;
            return D_OK;
        }
        br.init(in, size, (uint64_t)body * 8);
    }
}

static uint16_t kraft4[4096];          // sum of 2^(7-len) over four 3-bit code lengths (len 0 counts nothing)
static uint8_t head13[8192];           // first 13 bits of a dynamic block header that can be one
static const bool kraft4_init = [] {
    for (int i = 0; i < 8192; i++) head13[i] = (i & 6) == 4 && ((i >> 3) & 31) <= 29 && ((i >> 8) & 31) <= 29;
    for (int i = 0; i < 4096; i++) {
        int k = 0;
        for (int j = 0; j < 4; j++) { const int l = (i >> (3 * j)) & 7; k += l ? 128 >> l : 0; }
        kraft4[i] = (uint16_t)k;
    }
    return true;
}();

// First position in [from, to) that looks like the start of a dynamic block.
static uint64_t find_dynamic_block(const uint8_t *in, size_t size, uint64_t from, uint64_t to, Huff &scratch)
{
    for (uint64_t pos = from; pos < to;) {
        const size_t byte = (size_t)(pos >> 3);
        if (byte + 16 > size) return UINT64_MAX;
        uint64_t w;
        memcpy(&w, in + byte, 8);                                // 8 candidate positions share this word (13 + 7 bits used)
        unsigned cand = 0;                                       // BTYPE 10 (either BFINAL), HLIT <= 29, HDIST <= 29: no branches
        for (unsigned k = 0; k < 8; k++) cand |= (unsigned)head13[(w >> k) & 0x1fff] << k;
        cand &= 0xffu << (pos & 7);
        const uint64_t base = pos & ~7ull;
        pos = base + 8;
        while (cand) {
            const unsigned k = (unsigned)__builtin_ctz(cand);
            cand &= cand - 1;
            const uint64_t at = base + k;
            if (at >= to) return UINT64_MAX;
            const int hclen = (int)((w >> (k + 13)) & 15) + 4;
            // the code-length code must be complete: sum 2^(7-len) == 128 (four 3-bit lengths per table look-up)
            uint64_t v;
            const uint64_t p2 = at + 17;
            memcpy(&v, in + (size_t)(p2 >> 3), 8);
            v >>= (p2 & 7);                                      // >= 57 bits = 19 codes of 3
            v &= (1ull << (3 * hclen)) - 1;
            const int kraft = kraft4[v & 4095] + kraft4[(v >> 12) & 4095] + kraft4[(v >> 24) & 4095] + kraft4[(v >> 36) & 4095] + kraft4[v >> 48];
            if (kraft != 128) continue;
            BitReader br;
            br.init(in, size, at + 3);
            if (read_dynamic_header(br, scratch, true)) return at;
        }
    }
    return UINT64_MAX;
}

// BGZF (bgzip) members start byte-aligned with a fixed 16-byte header; the deflate data follows 2 bytes later.
static uint64_t find_bgzf_member(const uint8_t *in, size_t size, uint64_t from, uint64_t to)
{
    size_t h = (size_t)((from + 7) >> 3);
    h = h >= 18 ? h - 18 : 0;
    while (h + 26 <= size) {
        const uint8_t *q = (const uint8_t *)memchr(in + h, 0x1f, size - 25 - h);
        if (!q) break;
        h = (size_t)(q - in);
        const uint64_t b = (uint64_t)(h + 18) * 8;               // where the member's deflate data starts
        if (b >= to) break;
        if (b >= from && q[1] == 0x8b && q[2] == 8 && q[3] == 4 && q[10] == 6 && q[11] == 0 && q[12] == 'B' && q[13] == 'C' &&
            q[14] == 2 && q[15] == 0)
            return b;
        h++;
    }
    return UINT64_MAX;
}

// Symbols to bytes through `lut` (identity for 0..255, the 32 KiB before the run at 0x8000...: in
// DNA most bases are copies of copies that lead back into that history, so about half of all
// symbols are markers and a branch per symbol would mispredict).  Returns the number of '\n' written.
static uint64_t resolve_markers(const uint16_t *s, size_t n, const uint8_t *lut, uint8_t *d)
{
    size_t i = 0;
    uint64_t newlines = 0;
#if defined(__SSE2__)
    const __m128i hi = _mm_set1_epi16((short)0xff00), nl = _mm_set1_epi8('\n');
    for (; i + 16 <= n; i += 16) {
        const __m128i a = _mm_loadu_si128((const __m128i *)(s + i)), b = _mm_loadu_si128((const __m128i *)(s + i + 8));
        if (_mm_movemask_epi8(_mm_cmpeq_epi16(_mm_and_si128(_mm_or_si128(a, b), hi), _mm_setzero_si128())) == 0xffff)
            _mm_storeu_si128((__m128i *)(d + i), _mm_packus_epi16(a, b));
        else
            for (size_t j = i; j < i + 16; j++) d[j] = lut[s[j]];
        newlines += (unsigned)__builtin_popcount((unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(d + i)), nl)));
    }
#endif
    for (; i < n; i++) { d[i] = lut[s[i]]; newlines += d[i] == '\n'; }
    return newlines;
}

// CRC-32 (gzip polynomial) by carry-less multiplication: fold 64 bytes per step, then reduce
// (Gopal et al., "Fast CRC computation for generic polynomials using PCLMULQDQ", Intel 2009;
// the folding constants are x^(512+-32), x^(128+-32), x^64 mod P and the Barrett pair for the
// bit-reflected P = 0x1DB710641).  len is a multiple of 16 and >= 64; crc is the raw register.
#if defined(__x86_64__)
__attribute__((target("pclmul,sse4.1")))
static uint32_t crc32_clmul(const uint8_t *buf, size_t len, uint32_t crc)
{
    static const uint64_t __attribute__((aligned(16))) k1k2[] = {0x0154442bd4ull, 0x01c6e41596ull};
    static const uint64_t __attribute__((aligned(16))) k3k4[] = {0x01751997d0ull, 0x00ccaa009eull};
    static const uint64_t __attribute__((aligned(16))) k5k0[] = {0x0163cd6124ull, 0x0000000000ull};
    static const uint64_t __attribute__((aligned(16))) poly[] = {0x01db710641ull, 0x01f7011641ull};
    __m128i x0, x1, x2, x3, x4, x5, x6, x7, x8, y5, y6, y7, y8;
    x1 = _mm_loadu_si128((const __m128i *)(buf + 0x00));
    x2 = _mm_loadu_si128((const __m128i *)(buf + 0x10));
    x3 = _mm_loadu_si128((const __m128i *)(buf + 0x20));
    x4 = _mm_loadu_si128((const __m128i *)(buf + 0x30));
    x1 = _mm_xor_si128(x1, _mm_cvtsi32_si128((int)crc));
    x0 = _mm_load_si128((const __m128i *)k1k2);
    buf += 64; len -= 64;
    while (len >= 64) {
        x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x6 = _mm_clmulepi64_si128(x2, x0, 0x00);
        x7 = _mm_clmulepi64_si128(x3, x0, 0x00); x8 = _mm_clmulepi64_si128(x4, x0, 0x00);
        x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x2 = _mm_clmulepi64_si128(x2, x0, 0x11);
        x3 = _mm_clmulepi64_si128(x3, x0, 0x11); x4 = _mm_clmulepi64_si128(x4, x0, 0x11);
        y5 = _mm_loadu_si128((const __m128i *)(buf + 0x00)); y6 = _mm_loadu_si128((const __m128i *)(buf + 0x10));
        y7 = _mm_loadu_si128((const __m128i *)(buf + 0x20)); y8 = _mm_loadu_si128((const __m128i *)(buf + 0x30));
        x1 = _mm_xor_si128(_mm_xor_si128(x1, x5), y5); x2 = _mm_xor_si128(_mm_xor_si128(x2, x6), y6);
        x3 = _mm_xor_si128(_mm_xor_si128(x3, x7), y7); x4 = _mm_xor_si128(_mm_xor_si128(x4, x8), y8);
        buf += 64; len -= 64;
    }
    x0 = _mm_load_si128((const __m128i *)k3k4);
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x1 = _mm_xor_si128(_mm_xor_si128(x1, x2), x5);
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x1 = _mm_xor_si128(_mm_xor_si128(x1, x3), x5);
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x1 = _mm_xor_si128(_mm_xor_si128(x1, x4), x5);
    while (len >= 16) {
        x2 = _mm_loadu_si128((const __m128i *)buf);
        x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x1 = _mm_xor_si128(_mm_xor_si128(x1, x2), x5);
        buf += 16; len -= 16;
    }
    x2 = _mm_clmulepi64_si128(x1, x0, 0x10);
    x3 = _mm_setr_epi32(~0, 0, ~0, 0);
    x1 = _mm_srli_si128(x1, 8);
    x1 = _mm_xor_si128(x1, x2);
    x0 = _mm_loadl_epi64((const __m128i *)k5k0);
    x2 = _mm_srli_si128(x1, 4);
    x1 = _mm_and_si128(x1, x3);
    x1 = _mm_clmulepi64_si128(x1, x0, 0x00);
    x1 = _mm_xor_si128(x1, x2);
    x0 = _mm_load_si128((const __m128i *)poly);
    x2 = _mm_and_si128(x1, x3);
    x2 = _mm_clmulepi64_si128(x2, x0, 0x10);
    x2 = _mm_and_si128(x2, x3);
    x2 = _mm_clmulepi64_si128(x2, x0, 0x00);
    x1 = _mm_xor_si128(x1, x2);
    return (uint32_t)_mm_extract_epi32(x1, 1);
}
#endif

static uint32_t fast_crc32(uint32_t crc, const uint8_t *p, size_t n)           // same contract as zlib's crc32_z
{
#if defined(__x86_64__)
    static const bool have = __builtin_cpu_supports("pclmul") && __builtin_cpu_supports("sse4.1");
    if (have && n >= 80) {
        const size_t body = n & ~(size_t)15;
        crc = ~crc32_clmul(p, body, ~crc);
        p += body; n -= body;
    }
#endif
    return n ? (uint32_t)crc32_z(crc, p, n) : crc;
}

} // namespace

// ---------------------------------------------------------------------------------------------
// Engine: decoder threads, the in-order chain that accepts their runs, segment hand-out.
// ---------------------------------------------------------------------------------------------
namespace {

struct Piece {                          // an accepted run: inflated bytes [out_off, out_off + n)
    uint16_t *sym = nullptr;
    size_t n = 0;
    uint64_t out_off = 0;
    uint8_t *window = nullptr;          // symbol -> byte: 64 KiB table, identity at 0..255, the 32 KiB before the run at 0x8000...
    size_t resolved = 0, cap = 0;
    SymPool *pool = nullptr;
    ~Piece() { if (pool) pool->give(sym, cap); else free(sym); free(window); }
};

struct Range {
    std::unique_ptr<Run> run;
    int state = 0;                      // 0 to do, 1 being decoded, 2 decoded from a start it found, 3 no start found
};

} // namespace

class Engine {
public:
    const uint8_t *in = nullptr;
    size_t size = 0, C = 0, S = 0;
    int nthreads = 0;
    size_t look = 0, limit = 0, soft_cap = 0;

    std::mutex mu;
    std::condition_variable cv_dec, cv_seg;
    std::vector<Range> ranges;
    size_t next_range = 0, cur = 0;
    bool cur_consumed = false, advancing = false;
    uint64_t pos = 0;
    uint8_t window[32768];
    std::vector<std::unique_ptr<Piece>> pieces;
    size_t first_live = 0;
    std::vector<MemberEnd> ends;        // off = offset in the inflated stream
    uint64_t accepted_out = 0, filled_out = 0;
    bool ended = false, failed = false;
    std::atomic<bool> quit{false};
    char err[256] = {0};
    long next_seg = 0;
    uint32_t run_crc = 0;               // fold()
    uint64_t run_len = 0, fold_off = 0;
    size_t next_end = 0;
    Stats st;
    SymPool pool;
    std::vector<std::thread> threads;
    ~Engine() { pieces.clear(); ranges.clear(); }        // hand every buffer back while the pool is still alive

    bool stopping() const { return ended || failed || quit.load(std::memory_order_relaxed); }
    void fail(const char *msg)
    {
        if (!failed) { failed = true; snprintf(err, sizeof(err), "%s", msg); }
        cv_dec.notify_all(); cv_seg.notify_all();
    }

    void accept(std::unique_ptr<Run> run)                    // mu held
    {
        for (const MemberEnd &m : run->ends) { ends.push_back(MemberEnd{accepted_out + m.off, m.crc, m.isize}); st.members++; }
        if (run->n) {
            std::unique_ptr<Piece> p(new Piece);
            p->window = (uint8_t *)malloc(65536);
            if (!p->window) { fail("out of memory"); return; }
            for (int v = 0; v < 256; v++) p->window[v] = (uint8_t)v;
            memcpy(p->window + 0x8000, window, 32768);
            p->sym = run->sym; p->n = run->n; p->cap = run->cap; p->pool = run->pool; p->out_off = accepted_out;
            run->sym = nullptr;
            if (p->n >= 32768) resolve_markers(p->sym + p->n - 32768, 32768, p->window, window);
            else {
                memmove(window, window + p->n, 32768 - p->n);
                resolve_markers(p->sym, p->n, p->window, window + 32768 - p->n);
            }
            accepted_out += p->n;
            pieces.push_back(std::move(p));
        }
        st.accepted++;
        pos = run->end_bit;
        if (run->stream_end) { ended = true; cv_dec.notify_all(); cv_seg.notify_all(); return; }
        if (pos >= (uint64_t)size * 8) { fail("unexpected end of compressed data"); return; }
        const size_t nc = (size_t)(pos / ((uint64_t)C * 8));
        if (nc != cur) {
            for (size_t k = cur + 1; k < nc; k++) ranges[k].run.reset();      // ranges the run went straight through
            cur = nc; cur_consumed = false;
        }
        cv_seg.notify_all();
    }

    // Extend the accepted stream as far as the decoded ranges allow.  mu held on entry and exit.
    void advance(std::unique_lock<std::mutex> &lk)
    {
        while (!stopping()) {
            Range &r = ranges[cur];
            if (!cur_consumed) {
                if (r.state < 2) return;
                cur_consumed = true;
                if (r.state == 2 && r.run && r.run->start_bit == pos) { accept(std::move(r.run)); continue; }
                r.run.reset();
            }
            // nothing usable starts at pos: decode from the known position
            const uint64_t from = pos, stop = (uint64_t)(cur + 1) * C * 8;
            lk.unlock();
            std::unique_ptr<Run> run(new Run(&pool));
            const double t0 = now_ms();
            const int rc = decode_run(in, size, from, stop, false, soft_cap, *run, quit);
            const double t1 = now_ms();
            lk.lock();
            st.redo_ms += t1 - t0; st.redone++;
            if (rc != D_OK) { if (!quit.load()) fail("invalid deflate data"); return; }
            accept(std::move(run));
        }
    }

    void decode_range(size_t j, std::unique_ptr<Run> &out, int &state, double &t_search, double &t_decode)
    {
        const uint64_t lo = (uint64_t)j * C * 8, hi = (uint64_t)(j + 1) * C * 8 < (uint64_t)size * 8 ? (uint64_t)(j + 1) * C * 8 : (uint64_t)size * 8;
        state = 3;
        if (j == 0) {                                           // the first range starts where the gzip header ends
            out.reset(new Run(&pool));
            const double t0 = now_ms();
            const int rc = decode_run(in, size, pos0, hi, false, soft_cap, *out, quit);
            t_decode += now_ms() - t0;
            if (rc == D_OK) state = 2; else out.reset();
            return;
        }
        std::unique_ptr<Huff> scratch(new Huff);
        uint64_t from = lo;
        while (!quit.load(std::memory_order_relaxed)) {
            const double t0 = now_ms();
            const uint64_t d = find_dynamic_block(in, size, from, hi, *scratch);
            const uint64_t b = find_bgzf_member(in, size, from, d < hi ? d : hi);      // a bgzip member that starts before it
            const uint64_t c = d < b ? d : b;
            const double t1 = now_ms();
            t_search += t1 - t0;
            if (c == UINT64_MAX) return;
            out.reset(new Run(&pool));
            const int rc = decode_run(in, size, c, hi, true, soft_cap, *out, quit);
            t_decode += now_ms() - t1;
            if (rc == D_OK) { state = 2; return; }
            out.reset();
            from = c + 1;
        }
    }

    uint64_t pos0 = 0;

    void worker()
    {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            if (stopping()) break;
            if (!advancing && (cur_consumed || ranges[cur].state >= 2)) {
                advancing = true;
                advance(lk);
                advancing = false;
                cv_dec.notify_all();
                continue;
            }
            if (next_range < cur) next_range = cur;
            if (next_range < ranges.size() && next_range < cur + look && accepted_out - filled_out < limit) {
                const size_t j = next_range++;
                ranges[j].state = 1;
                lk.unlock();
                std::unique_ptr<Run> run;
                int state = 3;
                double ts = 0, td = 0;
                decode_range(j, run, state, ts, td);
                lk.lock();
                st.search_ms += ts; st.decode_ms += td; st.ranges++;
                if (j < cur || (j == cur && cur_consumed)) run.reset();     // overtaken meanwhile
                else ranges[j].run = std::move(run);
                ranges[j].state = state;
                cv_dec.notify_all();
                continue;
            }
            const double t0 = now_ms();
            cv_dec.wait(lk);
            st.wait_ms += now_ms() - t0;
        }
        cv_dec.notify_all(); cv_seg.notify_all();
    }
};

Stream::Stream() : e(nullptr) {}
Stream::~Stream() { close(); }

bool Stream::open(const uint8_t *data, size_t size, int threads, size_t range_bytes, size_t seg_bytes, int pending_segments)
{
    close();
    e = new Engine;
    e->in = data; e->size = size;
    const size_t body = gzip_header(data, size, 0);
    if (!body) { snprintf(e->err, sizeof(e->err), "not a gzip file"); e->failed = true; return false; }
    e->pos0 = e->pos = (uint64_t)body * 8;
    e->C = range_bytes < 4096 ? 4096 : range_bytes;
    e->S = seg_bytes < 4096 ? 4096 : seg_bytes;
    e->nthreads = threads < 1 ? 1 : threads;
    e->look = (size_t)e->nthreads * 2 + 2;
    e->pool.keep = e->look + 8;
    e->limit = (size_t)(pending_segments < 2 ? 2 : pending_segments) * e->S;
    e->soft_cap = e->S > ((size_t)64 << 20) ? e->S : (size_t)64 << 20;
    e->ranges.resize((size + e->C - 1) / e->C);
    e->cur = (size_t)(e->pos / ((uint64_t)e->C * 8));
    memset(e->window, 0, sizeof(e->window));
    e->st.in_bytes = size;
    if (e->cur > 0) e->cur_consumed = true;                    // a header longer than a range: decode from pos directly
    for (int i = 0; i < e->nthreads; i++) e->threads.emplace_back([this] { e->worker(); });
    return true;
}

long Stream::claim(uint64_t *off, size_t *len, bool *last)
{
    std::unique_lock<std::mutex> lk(e->mu);
    for (;;) {
        if (e->failed || e->quit.load()) return -2;
        const uint64_t s = (uint64_t)e->next_seg, S = e->S;
        if (e->ended) {
            const uint64_t nseg = e->accepted_out ? (e->accepted_out + S - 1) / S : 1;
            if (s >= nseg) return -1;
            *off = s * S;
            *len = (size_t)(e->accepted_out - s * S < S ? e->accepted_out - s * S : S);
            *last = s + 1 == nseg;
            return e->next_seg++;
        }
        if (e->accepted_out > (s + 1) * S) { *off = s * S; *len = (size_t)S; *last = false; return e->next_seg++; }
        e->cv_seg.wait(lk);
    }
}

bool Stream::fill(uint64_t off, size_t len, uint8_t *dst, SegmentInfo *info)
{
    struct Src { Piece *p; size_t from, count; };
    std::vector<Src> srcs;
    std::vector<uint64_t> cuts;
    {
        std::lock_guard<std::mutex> g(e->mu);
        if (e->failed) return false;
        size_t lo = e->first_live, hi = e->pieces.size();
        while (lo < hi) {                                       // first piece that ends after off
            const size_t mid = (lo + hi) / 2;
            if (e->pieces[mid]->out_off + e->pieces[mid]->n > off) hi = mid; else lo = mid + 1;
        }
        uint64_t at = off;
        for (size_t k = lo; at < off + len && k < e->pieces.size(); k++) {
            Piece *p = e->pieces[k].get();
            const size_t from = (size_t)(at - p->out_off);
            const size_t count = (size_t)(p->n - from < off + len - at ? p->n - from : off + len - at);
            srcs.push_back(Src{p, from, count});
            at += count;
        }
        if (at != off + len) return false;
        for (const MemberEnd &m : e->ends) if (m.off > off && m.off < off + len) cuts.push_back(m.off);
    }
    // resolve, count newlines and checksum block by block while the bytes are in cache
    cuts.push_back(off + len);
    size_t w = 0, ci = 0;
    uint64_t newlines = 0, part_len = 0;
    uint32_t crc = 0;
    if (info) info->parts.clear();
    for (const Src &s : srcs) {
        for (size_t done = 0; done < s.count;) {
            const size_t blk = s.count - done < 65536 ? s.count - done : 65536;
            newlines += resolve_markers(s.p->sym + s.from + done, blk, s.p->window, dst + w);
            if (info) {
                for (size_t a = w; a < w + blk;) {
                    const size_t c = (size_t)(cuts[ci] - off) < w + blk ? (size_t)(cuts[ci] - off) : w + blk;
                    crc = fast_crc32(crc, dst + a, c - a);
                    part_len += c - a;
                    a = c;
                    if (off + c == cuts[ci]) { info->parts.push_back(SegmentInfo::Part{part_len, crc}); crc = 0; part_len = 0; ci++; }
                }
            }
            w += blk; done += blk;
        }
    }
    if (info) info->newlines = newlines;
    {
        std::lock_guard<std::mutex> g(e->mu);
        for (const Src &s : srcs) {
            s.p->resolved += s.count;
            if (s.p->resolved == s.p->n) { e->pool.give(s.p->sym, s.p->cap); free(s.p->window); s.p->sym = nullptr; s.p->window = nullptr; }
        }
        while (e->first_live < e->pieces.size() && !e->pieces[e->first_live]->sym) e->first_live++;
        e->filled_out += len;
        e->st.out_bytes += len;
    }
    e->cv_dec.notify_all();
    return true;
}

bool Stream::fold(const SegmentInfo &info, bool last)
{
    std::lock_guard<std::mutex> g(e->mu);
    auto members_ending_here = [&]() -> bool {
        while (e->next_end < e->ends.size() && e->ends[e->next_end].off == e->fold_off) {
            const MemberEnd &m = e->ends[e->next_end++];
            if (m.crc != e->run_crc || m.isize != (uint32_t)e->run_len) {
                e->fail("gzip CRC or length check failed");
                return false;
            }
            e->run_crc = 0; e->run_len = 0;
        }
        return true;
    };
    for (const SegmentInfo::Part &p : info.parts) {
        if (!members_ending_here()) return false;
        e->run_crc = (uint32_t)crc32_combine(e->run_crc, p.crc, (z_off_t)p.len);
        e->run_len += p.len;
        e->fold_off += p.len;
    }
    if (last) {
        if (!members_ending_here()) return false;
        if (e->next_end != e->ends.size() || e->run_len != 0) { e->fail("gzip member bookkeeping is inconsistent"); return false; }
    }
    return true;
}

void Stream::cancel()
{
    if (!e) return;
    e->quit.store(true);
    std::lock_guard<std::mutex> g(e->mu);
    e->cv_dec.notify_all(); e->cv_seg.notify_all();
}

void Stream::close()
{
    if (!e) return;
    cancel();
    for (auto &t : e->threads) t.join();
    delete e;
    e = nullptr;
}

const char *Stream::error() const { return e ? e->err : "closed"; }
Stats Stream::stats() const { std::lock_guard<std::mutex> g(e->mu); return e->st; }

} // namespace pinflate
