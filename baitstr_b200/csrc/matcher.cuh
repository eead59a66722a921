// matcher.cuh -- per-read STR detection on the GPU (sm_100a), thread-per-read, bit-parallel.
//
// Replaces the regex scan loop of the reference:
//   AltMatchPatternsAndReportBest   /root/reference/src/select_STR_reads.pyx:58-104
//   MatchPatternsAndReportBest (-x) /root/reference/src/select_STR_reads.pyx:106-169
//   pattern (?=((?=(.{k}))\2{n,}))  /root/reference/src/select_STR_reads.pyx:350
//   ProcessSequence (both strands)  /root/reference/src/select_STR_reads.pyx:294-319
//
// Formulation (DESIGN.md "Scan"): a read is turned into three bit-planes over the only
// window that can hold an admissible repeat, [f-6, L-f+6).  For every period k the
// equality mask E_k[u] = (t[u] == t[u+k]) is three funnel-shift/XOR groups per 32 bases.
// Starts with >= n copies are the positions where E_k has k(n-1) consecutive ones
// (shift/AND doubling).  The reference's max over the tuple (len, copies, STR, motif,
// start) is found without enumerating matches: "cur &= cur >> k" keeps the starts that
// still have one more copy, so the last non-empty mask holds exactly the starts with the
// largest copy count; only that handful of bits gets its motif compared.  The 3' flank
// rule (a greedy match that runs past L-f is DROPPED, not shortened) makes every run that
// ends early enough fully admissible and at most one run per period partially admissible;
// that one run is handled exactly by a short scalar path.
#pragma once
#include <stdint.h>

namespace bstr {

struct ScanParams {
    int n;        // minimum copies (-n), already clamped to >= 1
    int f;        // flank (-f)
    int kmask;    // bit k set: period k is searched
    int rmhm;     // -r
    int altalgo;  // 1 = default algorithm, 0 = -x
};

// One strand's winner.  k == 0: no admissible match (None).
struct Hit {
    int k, copies, start;
};

// 16-byte per-read result record (include/baitstr_select.h: bstr_read_result)
struct ReadResult {
    uint32_t fstart, rstart;
    uint32_t fck, rck;       // copies << 3 | k ; fck == 0 <=> read not selected
};

__device__ __forceinline__ uint32_t lowmask32(int nbits)
{   // bits [0, nbits) set, nbits clamped to [0, 32]
    return nbits <= 0 ? 0u : (nbits >= 32 ? 0xFFFFFFFFu : ((1u << nbits) - 1u));
}

template <int W> struct Bits {
    uint32_t w[W];
    __device__ __forceinline__ bool any() const
    {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < W; i++) o |= w[i];
        return o != 0;
    }
};

// dst = src >> s (towards bit 0), s in [1,31] compile-time or uniform; `fill` enters at the top
template <int W>
__device__ __forceinline__ void shr(Bits<W> &d, const Bits<W> &a, int s, uint32_t fill = 0)
{
#pragma unroll
    for (int i = 0; i < W; i++)
        d.w[i] = __funnelshift_r(a.w[i], (i + 1 < W) ? a.w[i + 1] : fill, s);
}

template <int W>
__device__ __forceinline__ void and_shr(Bits<W> &d, const Bits<W> &a, const Bits<W> &b, int s)
{   // d = a & (b >> s)
#pragma unroll
    for (int i = 0; i < W; i++)
        d.w[i] = a.w[i] & __funnelshift_r(b.w[i], (i + 1 < W) ? b.w[i + 1] : 0u, s);
}

template <int W>
__device__ __forceinline__ void lowmask(Bits<W> &m, int nbits)
{
#pragma unroll
    for (int i = 0; i < W; i++) m.w[i] = lowmask32(nbits - 32 * i);
}

template <int W>
__device__ __forceinline__ bool testbit(const Bits<W> &a, int pos)
{
    uint32_t v = 0;
#pragma unroll
    for (int i = 0; i < W; i++)
        if ((pos >> 5) == i) v = a.w[i];
    return (v >> (pos & 31)) & 1u;
}

// 32 bits starting at bit position pos (pos >= 0); zero beyond the array
template <int W>
__device__ __forceinline__ uint32_t extract32(const Bits<W> &a, int pos)
{
    uint32_t lo = 0, hi = 0;
    const int wi = pos >> 5;
#pragma unroll
    for (int i = 0; i < W; i++)
        if (wi == i) { lo = a.w[i]; hi = (i + 1 < W) ? a.w[i + 1] : 0u; }
    return __funnelshift_r(lo, hi, pos & 31);
}

// highest set bit position of (a & lowmask(limit+1)), or -1
template <int W>
__device__ __forceinline__ int highest_set_le(const Bits<W> &a, int limit)
{
    int r = -1;
#pragma unroll
    for (int i = 0; i < W; i++) {
        uint32_t v = a.w[i] & lowmask32(limit + 1 - 32 * i);
        if (v) r = 32 * i + 31 - __clz(v);
    }
    return r;
}

// lowest set bit position of a at position > after, or 32*W if none
template <int W>
__device__ __forceinline__ int lowest_set_gt(const Bits<W> &a, int after)
{
    int r = 32 * W;
#pragma unroll
    for (int i = W - 1; i >= 0; i--) {
        uint32_t v = a.w[i] & ~lowmask32(after + 1 - 32 * i);
        if (v) r = 32 * i + __ffs(v) - 1;
    }
    return r;
}

// ---------------------------------------------------------------------------------------
// Base encoding.  ASCII facts used (upper and lower case differ only in bit 5):
//   A 0x41  C 0x43  G 0x47  T 0x54  N 0x4E  . 0x2E
//   bit3 = 1 only for N and '.'          -> "special" plane
//   bit6 = 0 only for '.'                -> tells '.' from N
//   (bit1^bit2, bit2) = A 00, C 10, G 01, T 11  -> 2-bit code A0 C1 G2 T3 (p0 = low bit)
// Planes: p0, p1 = 2-bit code (N: 00, '.': p0=1), p2 = special.  Complement flips p0 and p1
// of the non-special bases.  Order-preserving 3-bit symbol for motif comparison:
//   (~p2, p1, p0): N 0 < '.' 1 < A 4 < C 5 < G 6 < T 7   (bytes: '.' 46 < A < C < G < T)
// Positions past the encoded window get the code p1 = p2 = 1, which equals no real symbol.
// ---------------------------------------------------------------------------------------
template <int W> struct Planes {
    Bits<W> p0, p1, p2;
};

// Gather bit b of each of the 4 bytes of `w` into the top nibble (bits 28..31) of the
// product; lower bits are don't-care.  No two partial products collide, so no carries.
__device__ __forceinline__ uint32_t gather_top(uint32_t w, int b)
{
    return (w & (0x01010101u << b)) * (0x10204080u >> b);
}

// Encode Gn bases starting at src (any alignment, generic address space) into raw planes.
// Reads whole aligned 32-bit words, up to 3 bytes before src and up to 7 bytes past the
// last base: callers keep 16 bytes of slack on both sides of their buffers.
template <int W>
__device__ __forceinline__ void encode_window(const uint8_t *src, int Gn, Planes<W> &P,
                                              Bits<W> &dotless_n /* N positions */)
{
    const uintptr_t addr = (uintptr_t)src;
    const uint32_t *wp = (const uint32_t *)(addr & ~(uintptr_t)3);
    const int mis8 = (int)(addr & 3) * 8;
#pragma unroll
    for (int i = 0; i < W; i++) {
        uint32_t a1 = 0, a2 = 0, a3 = 0, a6 = 0;
        if (32 * i < Gn) {
            // groups in descending order so that (acc << 4 | nibble) ends with group 0 lowest
#pragma unroll
            for (int g = 7; g >= 0; g--) {
                uint32_t v = 0;
                if (32 * i + 4 * g < Gn) {
                    const uint32_t lo = wp[8 * i + g], hi = wp[8 * i + g + 1];
                    v = __funnelshift_r(lo, hi, mis8);
                }
                a1 = __funnelshift_l(gather_top(v, 1), a1, 4);
                a2 = __funnelshift_l(gather_top(v, 2), a2, 4);
                a3 = __funnelshift_l(gather_top(v, 3), a3, 4);
                a6 = __funnelshift_l(gather_top(v, 6), a6, 4);
            }
        }
        const uint32_t valid = lowmask32(Gn - 32 * i);
        const uint32_t s = a3 & valid;
        P.p2.w[i] = s;
        P.p1.w[i] = a2 & ~s & valid;
        P.p0.w[i] = (((a1 ^ a2) & ~s) | (s & ~a6)) & valid;
        dotless_n.w[i] = s & a6;
    }
}

// Mark positions >= Gn with the "no symbol" code.
template <int W>
__device__ __forceinline__ void mark_invalid(Planes<W> &P, int Gn)
{
#pragma unroll
    for (int i = 0; i < W; i++) {
        const uint32_t inv = ~lowmask32(Gn - 32 * i);
        P.p1.w[i] |= inv;
        P.p2.w[i] |= inv;
    }
}

// a = reverse of the low Gn bits of a (bit v -> bit Gn-1-v), zeros above
template <int W>
__device__ __forceinline__ void reverse_bits(Bits<W> &a, int Gn)
{
    Bits<W> r;
#pragma unroll
    for (int i = 0; i < W; i++) r.w[i] = __brev(a.w[W - 1 - i]);
    const int rs = 32 * W - Gn;        // shift towards bit 0 by rs
    const int sb = rs & 31;
    Bits<W> t;
#pragma unroll
    for (int i = 0; i < W; i++)
        t.w[i] = __funnelshift_r(r.w[i], (i + 1 < W) ? r.w[i + 1] : 0u, sb);
    const int q = rs >> 5;
#pragma unroll
    for (int i = 0; i < W; i++) {
        uint32_t v = 0;
#pragma unroll
        for (int j = i; j < W; j++)
            if (j - i == q) v = t.w[j];
        a.w[i] = v;
    }
}

// Reverse complement of the raw planes (ReverseComplement, select_STR_reads.pyx:187-189:
// A<->T, C<->G, N and '.' unchanged).
template <int W>
__device__ __forceinline__ void revcomp_planes(Planes<W> &P, Bits<W> &nmask, int Gn)
{
    reverse_bits(P.p0, Gn);
    reverse_bits(P.p1, Gn);
    reverse_bits(P.p2, Gn);
    reverse_bits(nmask, Gn);
#pragma unroll
    for (int i = 0; i < W; i++) {
        const uint32_t reg = ~P.p2.w[i] & lowmask32(Gn - 32 * i);
        P.p0.w[i] ^= reg;
        P.p1.w[i] ^= reg;
    }
}

// Packed candidate key, compared as one unsigned 64-bit integer; mirrors the reference's
// tuple order (len(STR), copies, STR/motif, start)  (select_STR_reads.pyx:93,101-102):
//   bits 44..   len      bits 32..43 copies      bits 12..29 motif (6 x 3-bit symbols,
//   first base most significant, left-aligned to 6 symbols is NOT needed: equal len and
//   copies imply equal k)      bits 0..11 local start
__device__ __forceinline__ uint64_t pack_key(int k, int c, uint32_t motif, int start)
{
    return ((uint64_t)(uint32_t)(k * c) << 44) | ((uint64_t)(uint32_t)c << 32) |
           ((uint64_t)motif << 12) | (uint64_t)(uint32_t)start;
}

template <int W>
__device__ __forceinline__ uint32_t motif_key(const Planes<W> &P, int pos, int k)
{
    const uint32_t b0 = extract32(P.p0, pos), b1 = extract32(P.p1, pos),
                   b2 = ~extract32(P.p2, pos);
    uint32_t key = 0;
#pragma unroll
    for (int j = 0; j < 6; j++)
        if (j < k)
            key = (key << 3) | (((b2 >> j) & 1u) << 2) | (((b1 >> j) & 1u) << 1) | ((b0 >> j) & 1u);
    return key;
}

// State carried across the periods of one strand.
template <int W> struct StrandState {
    Bits<W> bcur;       // starts of the best (len, copies) class found so far
    int bk, bc;         // its period and copy count (bk == 0: none)
    uint64_t pbest;     // best fully evaluated candidate from partially admissible runs
};

template <int K, int W>
__device__ __forceinline__ void scan_period(const Planes<W> &P, const Bits<W> &nm,
                                            const Bits<W> &hm, int sh, int end_max, int n,
                                            bool use_hm, StrandState<W> &S)
{
    // local candidate start u must be >= sh, its greedy end must be <= end_max
    if (end_max - sh < K * n) return;
    Bits<W> E;
#pragma unroll
    for (int i = 0; i < W; i++) {
        const uint32_t x0 = __funnelshift_r(P.p0.w[i], (i + 1 < W) ? P.p0.w[i + 1] : 0u, K);
        const uint32_t x1 = __funnelshift_r(P.p1.w[i], (i + 1 < W) ? P.p1.w[i + 1] : ~0u, K);
        const uint32_t x2 = __funnelshift_r(P.p2.w[i], (i + 1 < W) ? P.p2.w[i + 1] : ~0u, K);
        E.w[i] = ~((P.p0.w[i] ^ x0) | (P.p1.w[i] ^ x1) | (P.p2.w[i] ^ x2));
    }
    E.w[0] &= ~lowmask32(sh);                 // 5' flank: no start below sh
    // A = starts with at least n copies = K(n-1) consecutive ones of E
    Bits<W> A;
    if (n <= 1) {
        // one copy always matches as long as the motif itself fits; E cannot tell (it is
        // zero where t[u+K] is past the window), so admit every start whose motif fits.
        lowmask(A, end_max - K + 1);
        A.w[0] &= ~lowmask32(sh);
    } else {
        Bits<W> t;
        and_shr(t, E, E, 1);                                   // run >= 2
        if (K == 2) A = t;
        else if (K == 3) and_shr(A, t, E, 2);
        else if (K == 4) and_shr(A, t, t, 2);
        else if (K == 5) { Bits<W> u; and_shr(u, t, t, 2); and_shr(A, u, E, 4); }
        else             { Bits<W> u; and_shr(u, t, t, 2); and_shr(A, u, t, 4); }
        for (int c = 2; c < n; c++) { Bits<W> u = A; and_shr(A, u, u, K); }
    }
#pragma unroll
    for (int i = 0; i < W; i++) {
        A.w[i] &= ~nm.w[i];                                    // 'N' in motif (pyx:76)
        if (use_hm) A.w[i] &= ~hm.w[i];                        // homopolymer motif (pyx:89)
    }
    if (!A.any()) return;

    // Runs that end at b <= p0 = end_max-K have every start admissible (greedy end b+K at
    // most end_max).  a0 = start of the run covering p0 (or p0+1 if E[p0] == 0).
    const int p0 = end_max - K;
    Bits<W> Z;
#pragma unroll
    for (int i = 0; i < W; i++) Z.w[i] = ~E.w[i];
    const int a0 = highest_set_le(Z, p0) + 1;
    Bits<W> As;
    {
        Bits<W> m; lowmask(m, a0);
#pragma unroll
        for (int i = 0; i < W; i++) As.w[i] = A.w[i] & m.w[i];
    }
    if (a0 <= p0) {
        // The run [a0, b0) covers p0.  A start i in it ends at the largest i + K*c <= b0+K;
        // the reference drops it unless that end is <= end_max (pyx:84) -- it does not
        // fall back to fewer copies.
        Bits<W> m1; lowmask(m1, p0 + 1);
        bool some = false;
#pragma unroll
        for (int i = 0; i < W; i++) some |= (A.w[i] & ~As.w[i] & m1.w[i]) != 0;
        if (some) {
            const int b0 = lowest_set_gt(Z, p0);
            if (b0 < end_max) {
                const int x = b0 + K - end_max;                // 1..K-1
#pragma unroll
                for (int d = 0; d < K; d++) {
                    const int i = a0 + d;
                    if (i <= p0 && testbit(A, i)) {
                        const int rr = b0 - i;
                        if (rr % K >= x) {
                            const uint64_t key = pack_key(K, 1 + rr / K, motif_key(P, i, K), i);
                            if (key > S.pbest) S.pbest = key;
                        }
                    }
                }
            }
        }
    }
    if (!As.any()) return;
    // largest copy count among the fully admissible starts
    int c = n;
    Bits<W> cur = As;
    bool more = true;
    if (n <= 1) {
        // As holds every start whose motif fits (one copy).  "One more copy" is K ones of E.
        Bits<W> t, q;
        and_shr(t, E, E, 1);
        if (K == 2) q = t;
        else if (K == 3) and_shr(q, t, E, 2);
        else if (K == 4) and_shr(q, t, t, 2);
        else if (K == 5) { Bits<W> u; and_shr(u, t, t, 2); and_shr(q, u, E, 4); }
        else             { Bits<W> u; and_shr(u, t, t, 2); and_shr(q, u, t, 4); }
        Bits<W> nx;
#pragma unroll
        for (int i = 0; i < W; i++) nx.w[i] = cur.w[i] & q.w[i];
        if (nx.any()) { cur = nx; c = 2; } else more = false;
    }
    while (more) {
        Bits<W> nx;
        and_shr(nx, cur, cur, K);
        if (!nx.any()) break;
        cur = nx;
        c++;
    }
    if (K * c > S.bk * S.bc || (K * c == S.bk * S.bc && c > S.bc)) {
        S.bk = K; S.bc = c; S.bcur = cur;
    }
}

// Best match of one strand given its planes.  Coordinates are window-local.
template <int W>
__device__ __forceinline__ uint64_t strand_best(const Planes<W> &P, const Bits<W> &nmask,
                                                int sh, int end_max, int n, int kmask,
                                                bool use_hm)
{
    StrandState<W> S;
    S.bk = 0; S.bc = 0; S.pbest = 0;
#pragma unroll
    for (int i = 0; i < W; i++) S.bcur.w[i] = 0;

    // nm_k = starts whose k-window holds an N; hm_k = starts whose k-window is one symbol
    Bits<W> nm = nmask, hm, E1;
    if (use_hm) {
#pragma unroll
        for (int i = 0; i < W; i++) {
            const uint32_t x0 = __funnelshift_r(P.p0.w[i], (i + 1 < W) ? P.p0.w[i + 1] : 0u, 1);
            const uint32_t x1 = __funnelshift_r(P.p1.w[i], (i + 1 < W) ? P.p1.w[i + 1] : ~0u, 1);
            const uint32_t x2 = __funnelshift_r(P.p2.w[i], (i + 1 < W) ? P.p2.w[i + 1] : ~0u, 1);
            E1.w[i] = ~((P.p0.w[i] ^ x0) | (P.p1.w[i] ^ x1) | (P.p2.w[i] ^ x2));
        }
        hm = E1;
    } else {
#pragma unroll
        for (int i = 0; i < W; i++) { hm.w[i] = 0; E1.w[i] = 0; }
    }
#define BSTR_PERIOD(K)                                                              \
    {                                                                               \
        Bits<W> t_;                                                                 \
        shr(t_, nmask, (K) - 1);                                                    \
        _Pragma("unroll") for (int i = 0; i < W; i++) nm.w[i] |= t_.w[i];           \
        if (use_hm && (K) > 2) { Bits<W> h_ = hm; and_shr(hm, h_, E1, (K) - 2); }   \
        if (kmask & (1 << (K))) scan_period<K, W>(P, nm, hm, sh, end_max, n, use_hm, S); \
    }
    BSTR_PERIOD(2) BSTR_PERIOD(3) BSTR_PERIOD(4) BSTR_PERIOD(5) BSTR_PERIOD(6)
#undef BSTR_PERIOD

    uint64_t best = S.pbest;
    if (S.bk) {
        // every set bit of bcur is a start with exactly bc copies of period bk; the tuple
        // order then prefers the greatest motif bytes, then the greatest start.
#pragma unroll
        for (int i = W - 1; i >= 0; i--) {
            uint32_t v = S.bcur.w[i];
            while (v) {
                const int pos = 32 * i + 31 - __clz(v);
                v &= ~(1u << (pos & 31));
                const uint64_t key = pack_key(S.bk, S.bc, motif_key(P, pos, S.bk), pos);
                if (key > best) best = key;
            }
        }
    }
    return best;
}

__device__ __forceinline__ Hit unpack_hit(uint64_t key, int base)
{
    Hit h;
    if (key == 0) { h.k = 0; h.copies = 0; h.start = 0; return h; }
    const int len = (int)(key >> 44), c = (int)((key >> 32) & 0xFFF);
    h.copies = c; h.k = len / c; h.start = base + (int)(key & 0xFFF);
    return h;
}

// -x: flank and homopolymer tests apply to the winner only (select_STR_reads.pyx:153-167)
template <int W>
__device__ __forceinline__ uint64_t xalgo_filter(uint64_t key, const ScanParams &prm, int L)
{
    if (key == 0) return 0;
    const int len = (int)(key >> 44), c = (int)((key >> 32) & 0xFFF), start = (int)(key & 0xFFF);
    const int k = len / c;
    if (start < prm.f || L - (start + len) < prm.f) return 0;
    if (prm.rmhm) {
        const uint32_t m = (uint32_t)(key >> 12) & 0x3FFFF;
        bool same = true;
        for (int j = 1; j < k; j++) same &= ((m >> (3 * j)) & 7u) == (m & 7u);
        if (same) return 0;
    }
    return key;
}

// Both strands of one read with a W-word window.  seq = first base, L = read length.
template <int W>
__device__ __forceinline__ void match_read_bits(const uint8_t *seq, int L, const ScanParams &prm,
                                                Hit &fwd, Hit &rc)
{
    fwd.k = 0; rc.k = 0; fwd.copies = rc.copies = fwd.start = rc.start = 0;
    const int fe = prm.altalgo ? prm.f : 0;          // -x: candidates ignore the flanks
    const int g0 = fe > 6 ? fe - 6 : 0;              // window = [g0, L - g0)
    const int Gn = L - 2 * g0;
    const int sh = fe - g0;
    const int end_max = sh + (L - 2 * fe);           // local bound on a candidate's end
    const int n = prm.n < 1 ? 1 : prm.n;
    if (Gn <= 0 || end_max - sh < 2 * n) return;
    const bool use_hm = prm.altalgo && prm.rmhm;

    Planes<W> raw;
    Bits<W> rawn;
    encode_window<W>(seq + g0, Gn, raw, rawn);
    uint64_t key0 = 0, key1 = 0;
    // one copy of the strand code, run twice (keeps the kernel inside the instruction cache)
#pragma unroll 1
    for (int strand = 0; strand < 2; strand++) {
        Planes<W> Q = raw;
        Bits<W> qn = rawn;
        if (strand) revcomp_planes(Q, qn, Gn);       // ReverseComplement, pyx:187-189
        mark_invalid(Q, Gn);
        uint64_t k = strand_best<W>(Q, qn, sh, end_max, n, prm.kmask, use_hm);
        if (!prm.altalgo) k = xalgo_filter<W>(k, prm, L);
        if (strand) key1 = k; else key0 = k;
        if (k == 0) break;                           // ProcessSequence: a strand with None drops the read
    }
    if (key0 == 0) return;
    fwd = unpack_hit(key0, g0);
    rc = unpack_hit(key1, g0);
}

// ---------------------------------------------------------------------------------------
// Long-read path (window wider than 256 bases): exact scalar scan straight off the bytes,
// O(L) per period and strand, O(1) state.  Same tuple order as above.  Still GPU code: the
// product has no CPU path.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint8_t strand_char(const uint8_t *seq, int L, bool rc, int i)
{
    uint8_t c = rc ? seq[L - 1 - i] : seq[i];
    if (c >= 'a' && c <= 'z') c -= 32;                       // fastq.pyx:83
    if (rc) {                                                // select_STR_reads.pyx:187-189
        if (c == 'A') c = 'T'; else if (c == 'T') c = 'A';
        else if (c == 'C') c = 'G'; else if (c == 'G') c = 'C';
    }
    return c;
}

__device__ inline Hit match_strand_bytes(const uint8_t *seq, int L, bool rc, const ScanParams &prm)
{
    const int n = prm.n < 1 ? 1 : prm.n;
    int b_len = 0, b_c = 0, b_s = 0, b_k = 0;
    for (int k = 2; k <= 6; k++) {
        if (!(prm.kmask & (1 << k)) || L < k * n) continue;
        int run = 0;
        for (int i = L - k - 1; i >= -1; i--) {
            // here run = run[i+1]; candidate start is i+1 .. handled one step late so that
            // start L-k (run 0) is visited too
            const int s = i + 1;
            {
                const int c = 1 + run / k;
                if (c >= n) {
                    const int len = k * c;
                    bool ok = true, homo = true;
                    const uint8_t m0 = strand_char(seq, L, rc, s);
                    for (int j = 0; j < k; j++) {
                        const uint8_t ch = strand_char(seq, L, rc, s + j);
                        ok &= ch != 'N';
                        homo &= ch == m0;
                    }
                    if (prm.altalgo) {
                        ok &= s >= prm.f && L - (s + len) >= prm.f;
                        if (prm.rmhm && homo) ok = false;
                    }
                    if (ok) {
                        bool better = false;
                        if (len != b_len) better = len > b_len;
                        else if (c != b_c) better = c > b_c;
                        else {
                            int cmp = 0;
                            for (int j = 0; j < k && cmp == 0; j++)
                                cmp = (int)strand_char(seq, L, rc, s + j) - (int)strand_char(seq, L, rc, b_s + j);
                            better = cmp > 0 || (cmp == 0 && s > b_s);
                        }
                        if (better) { b_len = len; b_c = c; b_s = s; b_k = k; }
                    }
                }
            }
            if (i >= 0)
                run = (strand_char(seq, L, rc, i) == strand_char(seq, L, rc, i + k)) ? run + 1 : 0;
        }
    }
    Hit h; h.k = 0; h.copies = 0; h.start = 0;
    if (b_k == 0) return h;
    if (!prm.altalgo) {
        if (b_s < prm.f || L - (b_s + b_len) < prm.f) return h;
        if (prm.rmhm) {
            bool homo = true;
            for (int j = 1; j < b_k; j++)
                homo &= strand_char(seq, L, rc, b_s + j) == strand_char(seq, L, rc, b_s);
            if (homo) return h;
        }
    }
    h.k = b_k; h.copies = b_c; h.start = b_s;
    return h;
}

__device__ inline void match_read_bytes(const uint8_t *seq, int L, const ScanParams &prm,
                                        Hit &fwd, Hit &rc)
{
    rc.k = 0; rc.copies = 0; rc.start = 0;
    fwd = match_strand_bytes(seq, L, false, prm);
    if (fwd.k == 0) return;
    rc = match_strand_bytes(seq, L, true, prm);
}

// number of 32-bit words the bit-parallel window of a read of length L needs (0: no window)
__device__ __forceinline__ int window_words(int L, const ScanParams &prm)
{
    const int fe = prm.altalgo ? prm.f : 0;
    const int g0 = fe > 6 ? fe - 6 : 0;
    const int Gn = L - 2 * g0;
    return Gn <= 0 ? 0 : (Gn + 31) >> 5;
}

// Dispatch on the window width.  `wmax` is warp-uniform (max over the warp's reads) so the
// branch does not diverge.
__device__ __forceinline__ void match_read(const uint8_t *seq, int L, const ScanParams &prm,
                                           int wmax, Hit &fwd, Hit &rc)
{
    if (wmax <= 2) match_read_bits<2>(seq, L, prm, fwd, rc);
    else if (wmax <= 4) match_read_bits<4>(seq, L, prm, fwd, rc);
    else if (wmax <= 8) match_read_bits<8>(seq, L, prm, fwd, rc);
    else match_read_bytes(seq, L, prm, fwd, rc);
}

// Bit 0 of every byte lane of the result is set iff that byte is NOT one of acgtnACGTN.
// (fastq.pyx:62-67); the other bits are don't-care.  Derivation in DESIGN.md: with x_i =
// bit i of the byte, T = x2&~x1&~x3 (only 'T' has it), a valid byte has x7 = 0,
// x4 = T, x0 = ~x3&~T, x3 -> x2&x1, and x6 = 1 unless it is '.' (x3 & x5).
__device__ __forceinline__ uint32_t swar_invalid(uint32_t w)
{
    const uint32_t s1 = w >> 1, s2 = w >> 2, s3 = w >> 3, s4 = w >> 4, s5 = w >> 5, s6 = w >> 6,
                   s7 = w >> 7;
    const uint32_t t = s2 & ~s1 & ~s3;
    const uint32_t b = s3 & ~(s2 & s1);
    const uint32_t c = (s4 ^ t) | b;
    const uint32_t d = w ^ (~s3 & ~t);
    const uint32_t e = ~s6 & ~(s3 & s5);
    return c | d | e | s7;
}

// Any byte of [p, p+L) outside acgtnACGTN. ?  Reads whole aligned words around the range.
__device__ __forceinline__ bool bases_invalid(const uint8_t *p, int L)
{
    const uintptr_t a = (uintptr_t)p;
    const uint32_t *wp = (const uint32_t *)(a & ~(uintptr_t)3);
    if (L <= 0) return false;
    const int head = (int)(a & 3), total = head + L, nw = (total + 3) >> 2;
    // bytes outside the range are replaced by 'A'; only the first and last word need it
    const uint32_t mfirst = 0xFFFFFFFFu << (8 * head);
    const int r = total - 4 * (nw - 1);                        // 1..4 bytes in the last word
    const uint32_t mlast = r < 4 ? (1u << (8 * r)) - 1u : 0xFFFFFFFFu;
    uint32_t acc;
    if (nw == 1) {
        const uint32_t m = mfirst & mlast;
        acc = swar_invalid((wp[0] & m) | (0x41414141u & ~m));
    } else {
        acc = swar_invalid((wp[0] & mfirst) | (0x41414141u & ~mfirst));
        for (int i = 1; i < nw - 1; i++) acc |= swar_invalid(wp[i]);
        acc |= swar_invalid((wp[nw - 1] & mlast) | (0x41414141u & ~mlast));
    }
    return (acc & 0x01010101u) != 0;
}

}  // namespace bstr
