// ginflate_dev.cuh -- DEFLATE range decoder on the GPU (sm_100a), device side.
//
// Replaces, for "*.gz" inputs, the single-threaded gzip.open of the reference
// (/root/reference/src/fastq.pyx:50-51).  Same two-pass scheme as the host inflater (pinflate.cpp,
// after pugz/rapidgzip): the compressed stream is cut into ranges; one WARP per range finds the
// first block start in its range and decodes from there WITHOUT the 32 KiB of history -- a
// back-reference that reaches before the range start is written as a 16-bit marker
// 0x8000 | window index, everything else as the byte itself.  A chain check on the host accepts
// the runs whose start is the end of the run before; k_propagate then resolves only the last
// 32 KiB of every run (which is the next run's history), k_resolve turns all symbols into bytes
// in parallel and k_crc_chunks checks every gzip member's CRC-32.  RFC 1951 / RFC 1952 decoding;
// the table layout follows pinflate.cpp so that both decoders accept the same streams.
//
// Decoding is serial per range, and ALL 32 lanes of the warp execute it redundantly: identical
// instructions on identical data (same addresses: broadcast loads, merged stores), so the extra
// lanes cost no issue slots -- and a match copy, the one place with per-symbol memory latency, is
// naturally lane-parallel (lane j copies symbol j): one L2 round trip per match, not one per symbol.
#pragma once
#include <stdint.h>

namespace ginflate {

// ---- table entries (as pinflate.cpp) -----------------------------------------------------------
//   bits 0-3 code bits to drop | 4-7 extra bits (link: width of the secondary index) | 8-10 kind |
//   16-31 literal / length base / distance base / offset of the secondary table; K_LIT2: two literals
enum : uint32_t { K_LIT = 0, K_BASE = 1, K_EOB = 2, K_LINK = 3, K_LIT2 = 4, K_BAD = 5 };
__host__ __device__ inline uint32_t mk(uint32_t kind, uint32_t nbits, uint32_t extra, uint32_t val) { return nbits | (extra << 4) | (kind << 8) | (val << 16); }
__host__ __device__ inline uint32_t e_nbits(uint32_t e) { return e & 15; }
__host__ __device__ inline uint32_t e_extra(uint32_t e) { return (e >> 4) & 15; }
__host__ __device__ inline uint32_t e_kind(uint32_t e) { return (e >> 8) & 7; }
__host__ __device__ inline uint32_t e_val(uint32_t e) { return e >> 16; }

constexpr int LB = 11, DB = 8, CB = 7;                 // primary index widths: literal/length, distance, code-length code
constexpr int LIT_CAP = (1 << LB) + 2400, DIST_CAP = (1 << DB) + 2048;

// per-range table slot in GLOBAL memory: the two tables of the current block + scratch of the table
// builder.  Decoding is a serial, latency-bound chain per warp (one active lane), so throughput is
// resident warps x chain speed: with the tables in shared memory (30 KB per warp) an SM held 7
// warps; in global memory they stay L1-resident (the kernel carves out no shared memory) and an SM
// holds 20 (measured: decode stage 3.0 -> see profiles/r02_ginflate_stages.txt)
struct WarpTables {
    uint32_t lit[LIT_CAP];
    uint32_t dist[DIST_CAP];
    uint32_t ctab[1 << CB];
    uint16_t rev[288];
    uint8_t lens[320];
    uint8_t deepest[1 << LB];
};
constexpr int WARPS_PER_CTA = 4;

__constant__ uint16_t c_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ uint8_t c_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_codelen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

enum TableKind { T_CODELEN, T_LITLEN, T_DIST };

__device__ inline uint32_t symbol_entry(TableKind t, int s)
{
    if (t == T_CODELEN) return mk(K_LIT, 0, 0, (uint32_t)s);
    if (t == T_DIST) return s < 30 ? mk(K_BASE, 0, c_dist_extra[s], c_dist_base[s]) : mk(K_BAD, 0, 0, 0);
    if (s < 256) return mk(K_LIT, 0, 0, (uint32_t)s);
    if (s == 256) return mk(K_EOB, 0, 0, 0);
    if (s < 286) return mk(K_BASE, 0, c_len_extra[s - 257], c_len_base[s - 257]);
    return mk(K_BAD, 0, 0, 0);
}

// Canonical code from code lengths into tab (primary 2^pb entries + secondary tables).  Returns 0
// (complete), 1 (no code at all / the one-code case zlib also accepts), -1 (over-subscribed,
// incomplete, or too large).  One thread; rev / deepest are the warp's scratch.
__device__ inline int build_table(uint32_t *tab, int cap, int pb, const uint8_t *lens, int n, TableKind t,
                                  uint16_t *rev, uint8_t *deepest)
{
    int count[16];
    for (int i = 0; i < 16; i++) count[i] = 0;
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
    for (int s = 0; s < n; s++)
        if (lens[s]) rev[s] = (uint16_t)(__brev(next[lens[s]]++) >> (32 - lens[s]));
    const uint32_t pmask = (1u << pb) - 1;
    int used = 1 << pb;
    if (maxlen > pb) {
        for (int i = 0; i < (1 << pb); i++) deepest[i] = 0;
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
    if (t == T_LITLEN) {
        // pair up short literals (sequence and quality lines are runs of 2..6-bit literals).  In
        // place and from the top down: the partner entry lives at a smaller index, still unpaired.
        for (int i = (1 << LB) - 1; i >= 0; i--) {
            const uint32_t a = tab[i];
            if (e_kind(a) != K_LIT || e_nbits(a) >= (uint32_t)LB || e_nbits(a) == 0) continue;
            const uint32_t b = tab[(uint32_t)i >> e_nbits(a)];
            if (e_kind(b) != K_LIT || e_nbits(b) == 0 || e_nbits(a) + e_nbits(b) > (uint32_t)LB) continue;
            tab[i] = mk(K_LIT2, e_nbits(a) + e_nbits(b), 0, e_val(a) | e_val(b) << 8);
        }
    }
    return left > 0 ? 1 : 0;
}

// ---- bit reader over the compressed window in global memory: LSB first, 64-bit buffer ------------
// The buffer behind `in` is readable (zero-filled) for 32 bytes past `size`; reading further than
// 16 bytes past the end fails, so a decoder can never run away.
struct BitReader {
    const uint8_t *in;
    uint64_t size, pos;        // pos: next byte to load
    uint64_t buf;
    int cnt;
    __device__ inline uint64_t load8(uint64_t p) const
    {   // eight bytes at any alignment from two aligned words
        const uint64_t *w = reinterpret_cast<const uint64_t *>(in) + (p >> 3);
        const unsigned sh = (unsigned)(p & 7) * 8;
        const uint64_t lo = __ldg(w);
        if (sh == 0) return lo;
        return (lo >> sh) | (__ldg(w + 1) << (64 - sh));
    }
    __device__ inline bool refill()
    {
        if (pos + 8 <= size) {
            buf |= load8(pos) << cnt;
            pos += (uint64_t)((63 - cnt) >> 3);
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
    __device__ inline void init(const uint8_t *data, uint64_t n, uint64_t bit)
    {
        in = data; size = n; pos = bit >> 3; buf = 0; cnt = 0;
        refill();
        drop((int)(bit & 7));
    }
    __device__ inline void drop(int n) { buf >>= n; cnt -= n; }
    __device__ inline uint32_t take(int n) { const uint32_t v = (uint32_t)(buf & ((1ull << n) - 1)); drop(n); return v; }
    __device__ inline uint64_t bitpos() const { return pos * 8 - (uint64_t)cnt; }
    __device__ inline bool overrun() const { return bitpos() > size * 8; }
    __device__ inline void align_byte() { drop(cnt & 7); }
};

// Dynamic block header (after the 3 block bits).  strict: only what a sane compressor writes (the
// position is a guess).
__device__ inline bool read_dynamic_header(BitReader &br, WarpTables &T, bool strict)
{
    if (!br.refill()) return false;
    const int hlit = (int)br.take(5) + 257, hdist = (int)br.take(5) + 1, hclen = (int)br.take(4) + 4;
    if (hlit > 286 || hdist > 30) return false;
    uint8_t cl[19];
    for (int i = 0; i < 19; i++) cl[i] = 0;
    for (int i = 0; i < hclen; i++) {
        if (br.cnt < 3 && !br.refill()) return false;
        cl[c_codelen_order[i]] = (uint8_t)br.take(3);
    }
    if (build_table(T.ctab, 1 << CB, CB, cl, 19, T_CODELEN, T.rev, T.deepest) != 0) return false;
    uint8_t *lens = T.lens;
    const int total = hlit + hdist;
    int i = 0;
    while (i < total) {
        if (br.cnt < 14 && !br.refill()) return false;
        const uint32_t e = T.ctab[br.buf & ((1u << CB) - 1)];
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
    const int rl = build_table(T.lit, LIT_CAP, LB, lens, hlit, T_LITLEN, T.rev, T.deepest);
    if (rl < 0 || (strict && rl != 0)) return false;
    const int rd = build_table(T.dist, DIST_CAP, DB, lens + hlit, hdist, T_DIST, T.rev, T.deepest);
    if (rd < 0) return false;
    return !br.overrun();
}

__device__ inline void fixed_tables(WarpTables &T)
{
    uint8_t *l = T.lens;
    for (int i = 0; i < 288; i++) l[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
    build_table(T.lit, LIT_CAP, LB, l, 288, T_LITLEN, T.rev, T.deepest);
    for (int i = 0; i < 32; i++) l[i] = 5;
    build_table(T.dist, DIST_CAP, DB, l, 32, T_DIST, T.rev, T.deepest);
}


// ---- one run: symbols of a range ------------------------------------------------------------------
enum { D_OK = 0, D_ERROR = -1, D_REJECT = -2, D_OVERFLOW = -3, D_UNSUPPORTED = -4 };
constexpr int MAX_ENDS = 48;                           // gzip members that may end inside one run (bgzip: ~8 per 128 KiB)
constexpr uint32_t SYM_SLACK = 320;                    // a match is at most 258 symbols

struct RangeDesc {                                     // host -> device
    uint64_t from_bit;      // mode 0: search from here; mode 1: a block starts exactly here; mode 2: a gzip header starts at this BYTE * 8
    uint64_t stop_bit;      // decode up to the first block boundary at or after this bit (mode 0: also the end of the search)
    uint64_t sym_off;       // where the run's symbols go in the arena
    uint32_t sym_cap;
    int32_t mode;
};
struct RangeMeta {                                     // device -> host
    uint64_t start_bit, end_bit;
    uint32_t n_sym;
    int32_t status;         // D_OK: a run; 1: no block start in the range; < 0: D_*
    int32_t stream_end;     // the run ended with the last member of the file
    int32_t n_ends;
    uint64_t end_off[MAX_ENDS];     // symbols of this run before the member ended
    uint32_t end_crc[MAX_ENDS], end_isize[MAX_ENDS];
};

__device__ inline bool texty(uint32_t c) { return (c >= 32 && c < 127) || c == '\n' || c == '\t' || c == '\r'; }

// Copy a match of `len` symbols from `dist` back, the whole warp: symbol j of the match is symbol
// (j mod dist) of the `dist` symbols before it -- all of which exist before the copy starts, so the
// lanes are independent.  Positions before the start of the run become markers.
__device__ inline void copy_match(uint16_t *base, uint16_t *out, uint32_t len, uint32_t dist)
{
    const int have = (int)(out - base);
    for (uint32_t j = threadIdx.x & 31; j < len; j += 32) {
        const int t = have - (int)dist + (int)(j < dist ? j : j % dist);
        out[j] = t < 0 ? (uint16_t)(0x8000u | (uint32_t)(32768 + t)) : base[t];
    }
    __syncwarp();                                                    // later matches read these symbols from other lanes
}

// The symbols of one block.  The hot loop does not shift a 64-bit buffer (two SHF per shift here):
// it keeps three consecutive 32-bit words of the input (the third is the prefetch) and a bit offset,
// and takes a 32-bit window with ONE funnel shift -- enough for a literal/length code with its extra
// bits (15 + 5) and, after re-normalising, for a distance code with its extra bits (15 + 13).  The
// last bytes of the input (and everything the fast loop does not like) go through the careful
// BitReader loop below, which is the same code as before.
template <bool STRICT>
__device__ inline int inflate_block(BitReader &br, const WarpTables &T, uint16_t *base, uint32_t &n, uint32_t cap)
{
    const uint32_t LMASK = (1u << LB) - 1, DMASK = (1u << DB) - 1;
    uint16_t *out = base + n;
    uint16_t *const lim = base + (cap > SYM_SLACK ? cap - SYM_SLACK : 0);
    // ---- fast loop ------------------------------------------------------------------------------
    {
        const uint64_t bit0 = br.bitpos();
        const uint32_t *wp = reinterpret_cast<const uint32_t *>(br.in);
        uint64_t wi = bit0 >> 5;                                       // word that holds the next bit
        const uint64_t wi_end = br.size >= 64 ? (br.size - 32) >> 2 : 0;   // stay clear of the end: the careful loop finishes
        if (wi + 4 < wi_end) {
            uint32_t off = (uint32_t)(bit0 & 31);
            uint32_t lo = __ldg(wp + wi), hi = __ldg(wp + wi + 1), nx = __ldg(wp + wi + 2);
            int rc = 1;                                                // 1: left the loop near the end of the input / output
#define GI_NORMALISE() if (off >= 32) { lo = hi; hi = nx; wi++; nx = __ldg(wp + wi + 2); off -= 32; }
            while (wi + 4 < wi_end && out < lim) {
                uint32_t win = __funnelshift_r(lo, hi, off);
                uint32_t e = T.lit[win & LMASK];
                if ((e & 0x300) == 0) {                               // K_LIT or K_LIT2: one or two literals
                    const uint32_t v = e >> 16;
                    if (STRICT && !(texty(v & 0xff) && ((e & 0x400) == 0 || texty(v >> 8)))) { rc = D_REJECT; break; }
                    out[0] = (uint16_t)(v & 0xff);
                    out[1] = (uint16_t)(v >> 8);                      // overwritten by what follows unless this is a pair
                    out += 1 + ((e >> 10) & 1);
                    off += e & 15;
                    GI_NORMALISE();
                    continue;
                }
                uint32_t used = e_nbits(e);
                if (e_kind(e) == K_LINK) {
                    e = T.lit[e_val(e) + ((win >> LB) & ((1u << e_extra(e)) - 1))];
                    used = LB + e_nbits(e);
                }
                const uint32_t kind = e_kind(e);
                if (kind == K_LIT) {
                    if (STRICT && !texty(e_val(e))) { rc = D_REJECT; break; }
                    *out++ = (uint16_t)e_val(e);
                    off += used;
                    GI_NORMALISE();
                    continue;
                }
                if (kind == K_BASE) {
                    const uint32_t xl = e_extra(e);
                    const uint32_t len = e_val(e) + ((win >> used) & ((1u << xl) - 1));
                    off += used + xl;
                    GI_NORMALISE();
                    win = __funnelshift_r(lo, hi, off);
                    uint32_t d = T.dist[win & DMASK];
                    uint32_t du = e_nbits(d);
                    if (e_kind(d) == K_LINK) {
                        d = T.dist[e_val(d) + ((win >> DB) & ((1u << e_extra(d)) - 1))];
                        du = DB + e_nbits(d);
                    }
                    if (e_kind(d) != K_BASE) { rc = STRICT ? D_REJECT : D_ERROR; break; }
                    const uint32_t xd = e_extra(d);
                    const uint32_t dist = e_val(d) + ((win >> du) & ((1u << xd) - 1));
                    off += du + xd;
                    GI_NORMALISE();
                    copy_match(base, out, len, dist);
                    out += len;
                    continue;
                }
                if (kind == K_EOB) { off += used; rc = D_OK; break; }
                rc = STRICT ? D_REJECT : D_ERROR;
                break;
            }
#undef GI_NORMALISE
            n = (uint32_t)(out - base);
            if (rc < 0) return rc;
            br.init(br.in, br.size, wi * 32 + off);                    // hand the position back to the careful reader
            if (rc == D_OK) return br.overrun() ? (STRICT ? D_REJECT : D_ERROR) : D_OK;
        }
    }
    // ---- careful loop: the end of the input, the end of the output buffer --------------------------
    for (;;) {
        if (out >= lim) { n = (uint32_t)(out - base); return D_OVERFLOW; }
        // a literal/length code needs at most 15 bits; the rest of a length/distance pair is topped up below
        if (br.cnt < 15 && !br.refill()) { n = (uint32_t)(out - base); return D_ERROR; }
        uint32_t e = T.lit[br.buf & LMASK];
        if ((e & 0x300) == 0) {                                       // K_LIT or K_LIT2: one or two literals
            const uint32_t v = e >> 16;
            if (STRICT && !(texty(v & 0xff) && ((e & 0x400) == 0 || texty(v >> 8)))) { n = (uint32_t)(out - base); return D_REJECT; }
            out[0] = (uint16_t)(v & 0xff);
            out[1] = (uint16_t)(v >> 8);                              // overwritten by what follows unless this is a pair
            out += 1 + ((e >> 10) & 1);
            br.drop((int)(e & 15));
            continue;
        }
        if (e_kind(e) == K_LINK) { br.drop(LB); e = T.lit[e_val(e) + (uint32_t)(br.buf & ((1u << e_extra(e)) - 1))]; }
        br.drop((int)e_nbits(e));
        const uint32_t kind = e_kind(e);
        if (kind == K_LIT) {
            if (STRICT && !texty(e_val(e))) { n = (uint32_t)(out - base); return D_REJECT; }
            *out++ = (uint16_t)e_val(e);
            continue;
        }
        if (kind == K_BASE) {
            if (br.cnt < 33 && !br.refill()) { n = (uint32_t)(out - base); return D_ERROR; }     // 5 + 15 + 13 bits to come
            const uint32_t len = e_val(e) + br.take((int)e_extra(e));
            uint32_t d = T.dist[br.buf & DMASK];
            if (e_kind(d) == K_LINK) { br.drop(DB); d = T.dist[e_val(d) + (uint32_t)(br.buf & ((1u << e_extra(d)) - 1))]; }
            if (e_kind(d) != K_BASE) { n = (uint32_t)(out - base); return STRICT ? D_REJECT : D_ERROR; }
            br.drop((int)e_nbits(d));
            const uint32_t dist = e_val(d) + br.take((int)e_extra(d));
            copy_match(base, out, len, dist);
            out += len;
            continue;
        }
        n = (uint32_t)(out - base);
        if (kind == K_EOB) return br.overrun() ? (STRICT ? D_REJECT : D_ERROR) : D_OK;
        return STRICT ? D_REJECT : D_ERROR;
    }
}

__device__ inline uint32_t le32(const uint8_t *p) { return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }

// gzip member header at d[at...]: offset of the deflate stream, 0 if this is not a gzip header (or it is cut short)
__device__ inline uint64_t gzip_header(const uint8_t *d, uint64_t size, uint64_t at)
{
    if (at + 18 > size || d[at] != 0x1f || d[at + 1] != 0x8b || d[at + 2] != 8) return 0;
    const uint8_t flg = d[at + 3];
    if (flg & 0xe0) return 0;
    uint64_t p = at + 10;
    if (flg & 4) { if (p + 2 > size) return 0; p += 2 + ((uint64_t)d[p] | (uint64_t)d[p + 1] << 8); }
    if (flg & 8) { while (p < size && d[p]) p++; p++; }
    if (flg & 16) { while (p < size && d[p]) p++; p++; }
    if (flg & 2) p += 2;
    return p < size ? p : 0;
}

// Decode deflate blocks from `bit` (a block start) up to the first block boundary at or after
// stop_bit, crossing gzip member boundaries on the way (decode_run of pinflate.cpp).  guessed: the
// first block must pass the strict checks.  at_eof: the buffer ends where the file ends.
__device__ inline int decode_run(const uint8_t *in, uint64_t size, bool at_eof, uint64_t bit, uint64_t stop_bit, bool guessed,
                                 WarpTables &T, uint16_t *sym, uint32_t cap, RangeMeta &m)
{
    BitReader br;
    br.init(in, size, bit);
    m.start_bit = bit; m.n_sym = 0; m.stream_end = 0; m.n_ends = 0;
    uint32_t n = 0;
    bool first = true;
    for (;;) {
        const uint64_t here = br.bitpos();
        if (!first && here >= stop_bit) { m.end_bit = here; m.n_sym = n; return D_OK; }
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
            const uint64_t at = br.bitpos() >> 3;
            if (at + len > size) return D_ERROR;
            if ((uint64_t)n + len + SYM_SLACK > cap) return D_OVERFLOW;
            for (uint32_t j = threadIdx.x & 31; j < len; j += 32) sym[n + j] = in[at + j];
            __syncwarp();
            n += len;
            br.init(in, size, (at + len) * 8);
            r = D_OK;
        } else if (btype == 1) {
            if (strict) return D_REJECT;
            fixed_tables(T);
            r = inflate_block<false>(br, T, sym, n, cap);
        } else if (btype == 2) {
            if (!read_dynamic_header(br, T, strict)) return strict ? D_REJECT : D_ERROR;
            r = strict ? inflate_block<true>(br, T, sym, n, cap) : inflate_block<false>(br, T, sym, n, cap);
        } else
            return strict ? D_REJECT : D_ERROR;
        if (r != D_OK) return r;
        first = false;
        if (!bfinal) continue;
        // end of a gzip member: trailer, then maybe another member
        br.align_byte();
        uint64_t at = br.bitpos() >> 3;
        if (at + 8 > size) return D_ERROR;
        if (m.n_ends >= MAX_ENDS) return D_UNSUPPORTED;
        m.end_off[m.n_ends] = n; m.end_crc[m.n_ends] = le32(in + at); m.end_isize[m.n_ends] = le32(in + at + 4);
        m.n_ends++;
        at += 8;
        while (at < size && in[at] == 0) at++;                 // zero padding between members; gzip(1) accepts it
        const uint64_t body = at < size ? gzip_header(in, size, at) : 0;
        if (!body) {
            // end of file -- or bytes that are no gzip member ("Not a gzipped file" in Python's gzip
            // module, which the reference reads through, fastq.pyx:50-51)
            if (at < size) return strict ? D_REJECT : D_ERROR;
            if (!at_eof) return D_UNSUPPORTED;                 // the run left the window's look-ahead
            m.stream_end = 1; m.end_bit = size * 8; m.n_sym = n;
            return D_OK;
        }
        br.init(in, size, body * 8);
    }
}

// sum of 2^(7-len) over four 3-bit code lengths (len 0 counts nothing): the Kraft sum of the
// code-length code in five look-ups instead of a 19-trip loop (filled by the host, ginflate.cu)
__device__ uint16_t g_kraft4[4096];

// ---- block-start search, the whole warp: lane l looks at bit position pos + l --------------------
// A dynamic block header: BTYPE = 10, HLIT <= 29, HDIST <= 29, and a complete code-length code
// (Kraft sum of the HCLEN + 4 three-bit lengths = 1).  Returns the first such position in
// [from, to) or ~0; the full header parse and the strict decode of the block decide.
__device__ inline uint64_t find_dynamic_block(const uint8_t *in, uint64_t size, uint64_t from, uint64_t to)
{
    const unsigned lane = threadIdx.x & 31;
    for (uint64_t base = from; base < to; base += 32) {
        const uint64_t at = base + lane;
        bool ok = false;
        const uint64_t byte = at >> 3;
        if (at < to && byte + 16 <= size) {
            BitReader br; br.in = in; br.size = size;
            uint64_t w = br.load8(byte) >> (at & 7);                  // >= 57 bits
            ok = ((w >> 1) & 3) == 2 && ((w >> 3) & 31) <= 29 && ((w >> 8) & 31) <= 29;
            if (ok) {
                const int hclen = (int)((w >> 13) & 15) + 4;
                const uint64_t p2 = at + 17;
                uint64_t v = br.load8(p2 >> 3) >> (p2 & 7);           // >= 57 bits = 19 codes of 3
                v &= (1ull << (3 * hclen)) - 1;
                const int kraft = g_kraft4[v & 4095] + g_kraft4[(v >> 12) & 4095] + g_kraft4[(v >> 24) & 4095] + g_kraft4[(v >> 36) & 4095] +
                                  g_kraft4[v >> 48];
                ok = kraft == 128;
            }
        }
        const unsigned hit = __ballot_sync(0xFFFFFFFFu, ok);
        if (hit) return base + (unsigned)(__ffs((int)hit) - 1);
    }
    return ~0ull;
}

// BGZF (bgzip) members start byte-aligned with a fixed 16-byte header; the deflate data follows 2 bytes later.
__device__ inline uint64_t find_bgzf_member(const uint8_t *in, uint64_t size, uint64_t from, uint64_t to)
{
    const unsigned lane = threadIdx.x & 31;
    uint64_t h0 = (from + 7) >> 3;
    h0 = h0 >= 18 ? h0 - 18 : 0;
    for (; h0 + 26 <= size && (h0 + 18) * 8 < to; h0 += 32) {
        const uint64_t h = h0 + lane;
        bool ok = false;
        if (h + 26 <= size) {
            const uint8_t *q = in + h;
            const uint64_t b = (h + 18) * 8;
            ok = b >= from && b < to && q[0] == 0x1f && q[1] == 0x8b && q[2] == 8 && q[3] == 4 && q[10] == 6 && q[11] == 0 &&
                 q[12] == 'B' && q[13] == 'C' && q[14] == 2 && q[15] == 0;
        }
        const unsigned hit = __ballot_sync(0xFFFFFFFFu, ok);
        if (hit) return (h0 + (unsigned)(__ffs((int)hit) - 1) + 18) * 8;
    }
    return ~0ull;
}

// ---- kernels ------------------------------------------------------------------------------------------
// One warp per range.  mode 0: search for a block start (dynamic block or bgzip member) in
// [from_bit, stop_bit) and decode from the first one that passes the strict checks; mode 1: decode
// from from_bit (a known block start); mode 2: from_bit / 8 is the byte where a gzip header starts.
__global__ void __launch_bounds__(32 * WARPS_PER_CTA, 5)
k_decode_ranges(const uint8_t *__restrict__ in, uint64_t size, int at_eof, const RangeDesc *__restrict__ ranges, int nranges,
                uint16_t *__restrict__ arena, RangeMeta *__restrict__ meta, WarpTables *__restrict__ tables)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = blockIdx.x * WARPS_PER_CTA + warp;
    if (r >= nranges) return;
    WarpTables &T = tables[r];
    const RangeDesc d = ranges[r];
    RangeMeta &m = meta[r];
    uint16_t *sym = arena + d.sym_off;
    if (d.mode != 0) {
        uint64_t bit = d.from_bit;
        int rc = D_OK;
        if (d.mode == 2) {
            const uint64_t body = gzip_header(in, size, d.from_bit >> 3);
            if (!body) rc = D_ERROR;
            bit = body * 8;
        }
        RangeMeta mm;                                          // per lane, identical; lane 0 publishes
        if (rc == D_OK) rc = decode_run(in, size, at_eof != 0, bit, d.stop_bit, false, T, sym, d.sym_cap, mm);
        if (lane == 0) { mm.status = rc; m = mm; }
        return;
    }
    uint64_t from = d.from_bit;
    for (;;) {
        const uint64_t dyn = find_dynamic_block(in, size, from, d.stop_bit);
        const uint64_t bg = find_bgzf_member(in, size, from, dyn < d.stop_bit ? dyn : d.stop_bit);     // a bgzip member that starts before it
        const uint64_t c = dyn < bg ? dyn : bg;
        if (c == ~0ull) { if (lane == 0) { m.status = 1; m.n_sym = 0; m.start_bit = m.end_bit = 0; m.n_ends = 0; m.stream_end = 0; } return; }
        RangeMeta mm;
        const int rc = decode_run(in, size, at_eof != 0, c, d.stop_bit, true, T, sym, d.sym_cap, mm);     // all lanes, same result
        if (rc == D_OK || rc == D_OVERFLOW || rc == D_UNSUPPORTED) { if (lane == 0) { mm.status = rc; m = mm; } return; }
        from = c + 1;                                          // not a block start after all (or broken data: the chain decides)
    }
}

// History of every accepted run: hist[k] = the 32 KiB of text before run k (hist[0] is given).  One
// CTA walks the runs in order; a step only resolves the last 32 KiB of a run (a gather through
// the history before it), which is the history of the next run.
struct RunRef { uint64_t sym_off; uint64_t out_off; uint32_t n; uint32_t pad; };
__global__ void __launch_bounds__(1024, 1)
k_propagate(const uint16_t *__restrict__ arena, const RunRef *__restrict__ runs, int nruns, uint8_t *hist /* (nruns + 1) x 32768 */)
{
    extern __shared__ __align__(16) uint8_t pr_smem[];         // two histories: the gathers of a step stay on chip
    uint8_t *cur = pr_smem, *nxt = pr_smem + 32768;
    for (uint32_t i = threadIdx.x * 16; i < 32768u; i += blockDim.x * 16)
        *reinterpret_cast<uint4 *>(cur + i) = *reinterpret_cast<const uint4 *>(hist + i);
    __syncthreads();
    // thread t owns bytes t, t + 1024, ... of every history: consecutive lanes read consecutive symbols
    // (coalesced), 32 independent loads per thread go out together, then the gathers through the old
    // history stay in shared memory
    for (int k = 0; k < nruns; k++) {
        const uint32_t n = runs[k].n;
        const uint16_t *s = arena + runs[k].sym_off;
        uint8_t *o = hist + (size_t)(k + 1) * 32768;
        uint16_t sy[32];
#pragma unroll
        for (int j = 0; j < 32; j++) {
            const uint32_t i = threadIdx.x + 1024u * j;
            // byte i of the new history = byte (n + i - 32768) of (old history ++ run), counted from the run's start
            sy[j] = (n >= 32768u || i >= 32768u - n) ? s[n - (32768u - i)] : (uint16_t)0;
        }
#pragma unroll
        for (int j = 0; j < 32; j++) {
            const uint32_t i = threadIdx.x + 1024u * j;
            const bool from_run = n >= 32768u || i >= 32768u - n;
            const uint32_t v = sy[j];
            const uint8_t b = (uint8_t)(!from_run ? cur[(i + n) & 0x7fffu] : (v < 0x8000u ? v : cur[v & 0x7fffu]));
            nxt[i] = b;
            o[i] = b;
        }
        __syncthreads();
        uint8_t *t = cur; cur = nxt; nxt = t;
    }
}

// All symbols to bytes: text[out_off + j] = symbol, or the byte of the run's history a marker names.
__global__ void __launch_bounds__(256)
k_resolve(const uint16_t *__restrict__ arena, const RunRef *__restrict__ runs, int nruns, const uint8_t *__restrict__ hist,
          uint8_t *__restrict__ text)
{
    for (int k = blockIdx.y; k < nruns; k += gridDim.y) {
        const uint8_t *h = hist + (size_t)k * 32768;
        const uint16_t *s = arena + runs[k].sym_off;
        uint8_t *o = text + runs[k].out_off;
        const uint32_t n = runs[k].n;
        for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
            const uint16_t v = s[j];
            o[j] = v < 0x8000u ? (uint8_t)v : h[v & 0x7fffu];
        }
    }
}

// CRC-32 (gzip polynomial, reflected) of chunks of the inflated stream, straight from the symbols:
// thread per chunk, byte at a time through a 256-entry table in shared memory.  crc[i] is the
// standard CRC of the chunk alone (as zlib's crc32(0, ...)), so the host can chain the chunks of a
// member with crc32_combine.
struct CrcChunk { uint64_t sym_off; uint32_t n; uint32_t run; };
__global__ void __launch_bounds__(128)
k_crc_chunks(const uint16_t *__restrict__ arena, const CrcChunk *__restrict__ chunks, int nchunks, const uint8_t *__restrict__ hist,
             uint32_t *__restrict__ crc)
{
    __shared__ uint32_t tab[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        uint32_t c = (uint32_t)i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1u)));
        tab[i] = c;
    }
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nchunks) return;
    const CrcChunk ch = chunks[i];
    const uint16_t *s = arena + ch.sym_off;
    const uint8_t *h = hist + (size_t)ch.run * 32768;
    uint32_t c = 0xFFFFFFFFu;
    for (uint32_t j = 0; j < ch.n; j++) {
        const uint16_t v = s[j];
        const uint8_t b = v < 0x8000u ? (uint8_t)v : h[v & 0x7fffu];
        c = tab[(c ^ b) & 0xffu] ^ (c >> 8);
    }
    crc[i] = ~c;
}

}  // namespace ginflate
