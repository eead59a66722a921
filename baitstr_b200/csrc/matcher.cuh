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
    // aligned base by pointer arithmetic (not an integer round trip), so that the compiler keeps
    // the address space of `src` and emits LDS for shared-memory callers instead of generic LD
    const uintptr_t addr = (uintptr_t)src;
    const uint32_t *wp = (const uint32_t *)(src - (addr & 3));
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

// ---------------------------------------------------------------------------------------
// Joint two-strand scan.  For a period K let M(c) be the set of window positions s with at
// least c copies of t[s:s+K] starting at s (M(c+1) = M(c) & M(c) >> K for c >= 2, M(2) = K
// consecutive ones of E_K).  The regex is greedy, so a position counts for the copy number
// c only if the repeat cannot be extended by one more copy:
//   forward strand   start s:  c copies and s   not in M(c+1)          (extends rightwards)
//   reverse strand   the same repeat read from its right end e = s + K*c; there the greedy
//                    extension runs leftwards in forward coordinates:   s-K not in M(c+1)
// and it is admissible iff s >= sh and s + K*c <= end_max on either strand (a greedy match
// that violates a flank is DROPPED by the reference, never shortened: pyx:81-86).  Both
// strands therefore come out of ONE pass over c with the same masks; the reverse
// complement is never materialised.  The winner of the tuple order (len, copies, motif,
// start) is the largest admissible c per period, the largest K*c over the periods, and
// among the surviving positions a bit-parallel lexicographic filter on the motif symbols.
// No data-dependent scalar candidate loops: every lane of a warp runs the same code.
// ---------------------------------------------------------------------------------------
template <int W> struct Best {
    Bits<W> cand;      // positions (leftmost base of the repeat, window-local)
    int k, c;          // k == 0: nothing admissible yet
};

// d = a << s (away from bit 0), s in [1,31]
template <int W>
__device__ __forceinline__ void shl(Bits<W> &d, const Bits<W> &a, int s)
{
#pragma unroll
    for (int i = W - 1; i >= 0; i--)
        d.w[i] = __funnelshift_l((i > 0) ? a.w[i - 1] : 0u, a.w[i], s);
}

template <int W>
__device__ __forceinline__ void take_if_better(Best<W> &b, const Bits<W> &cand, int K, int c)
{
    if (c && (K * c > b.k * b.c || (K * c == b.k * b.c && c > b.c))) { b.k = K; b.c = c; b.cand = cand; }
}

// The copy-count scan of one period is split in two so that a warp does not run as many trips
// as its longest repeat for every period (measured round 1: a read needs 1.1 trips for K = 2, its
// warp ran 5.9 -- scan_period executed with 10 of 32 lanes):
//   first_trip   every lane, no loop: E_K, M(n), the filters, and the exactly-n-copies candidates
//                of both strands; reports whether any start can still be admissible with more
//                copies (M(n+1) & lim != 0) -- true for ~0.5 (read, period) pairs per read
//   extend_trips the loop over c = n+1, n+2, ... for such a pair, period as a run-time value; the
//                fused kernel collects the pairs of a warp in a shared-memory list and runs them
//                lane-per-pair (fused.cuh), other callers run it inline
// Deferred: one (read, period) pair whose scan continues.  meta = K | c << 3 | sh << 12 | end_max << 16
// (11 bits); the fused kernel's work list keeps the owner lane in bits 27..31.
template <int W> struct Deferred {
    Bits<W> cur;           // M(c), filtered
    uint32_t meta;
};
__device__ __forceinline__ uint32_t deferred_meta(int K, int c, int sh, int end_max)
{
    return (uint32_t)K | (uint32_t)c << 3 | (uint32_t)sh << 12 | (uint32_t)end_max << 16;
}

template <int K, int W>
__device__ __forceinline__ bool first_trip(const Planes<W> &P, const Bits<W> &nm, const Bits<W> &hm,
                                           bool use_hm, int sh, int end_max, int n,
                                           Best<W> &bf, Best<W> &br, Deferred<W> &more)
{
    Bits<W> E;
#pragma unroll
    for (int i = 0; i < W; i++) {
        const uint32_t x0 = __funnelshift_r(P.p0.w[i], (i + 1 < W) ? P.p0.w[i + 1] : 0u, K);
        const uint32_t x1 = __funnelshift_r(P.p1.w[i], (i + 1 < W) ? P.p1.w[i + 1] : ~0u, K);
        const uint32_t x2 = __funnelshift_r(P.p2.w[i], (i + 1 < W) ? P.p2.w[i + 1] : ~0u, K);
        E.w[i] = ~((P.p0.w[i] ^ x0) | (P.p1.w[i] ^ x1) | (P.p2.w[i] ^ x2));
    }
    // M2 = K consecutive ones of E = at least two copies
    Bits<W> M2;
    {
        Bits<W> t;
        and_shr(t, E, E, 1);
        if (K == 2) M2 = t;
        else if (K == 3) and_shr(M2, t, E, 2);
        else if (K == 4) and_shr(M2, t, t, 2);
        else if (K == 5) { Bits<W> u; and_shr(u, t, t, 2); and_shr(M2, u, E, 4); }
        else             { Bits<W> u; and_shr(u, t, t, 2); and_shr(M2, u, t, 4); }
    }
    // positions whose motif may be reported: no 'N' (pyx:76), not a homopolymer with -r (pyx:89).
    // A repeat's copies are identical strings, so filtering M(c) up front equals filtering
    // the candidates afterwards.
    Bits<W> cur;
#pragma unroll
    for (int i = 0; i < W; i++) {
        uint32_t ok = ~nm.w[i];
        if (use_hm) ok &= ~hm.w[i];
        cur.w[i] = ok;
        M2.w[i] &= ok;
    }
    const int c = n <= 1 ? 1 : n;
    Bits<W> lim;                                   // s >= sh and s + K*c <= end_max
    lowmask(lim, end_max - K * c + 1);
    const uint32_t lo_keep = ~lowmask32(sh);       // sh <= 6: word 0 only
    lim.w[0] &= lo_keep;
    Bits<W> nxt;
    if (n <= 1) nxt = M2;                          // c = 1: every (filtered) position is a one-copy match
    else {
        cur = M2;
        for (int cc = 2; cc < n; cc++) { Bits<W> u = cur; and_shr(cur, u, u, K); }
        and_shr(nxt, cur, cur, K);
    }
    Bits<W> nl_;
    shl(nl_, nxt, K);
    Bits<W> xf, xr;
    uint32_t af = 0, ar = 0, live = 0;
    {
        Bits<W> lim2;                              // the bound for c + 1 copies
        shr(lim2, lim, K);
        lim2.w[0] &= lo_keep;
#pragma unroll
        for (int i = 0; i < W; i++) {
            const uint32_t l = lim.w[i];
            xf.w[i] = cur.w[i] & ~nxt.w[i] & l;
            xr.w[i] = cur.w[i] & ~nl_.w[i] & l;
            af |= xf.w[i]; ar |= xr.w[i];
            live |= nxt.w[i] & lim2.w[i];
        }
    }
    if (af) take_if_better(bf, xf, K, c);
    if (ar) take_if_better(br, xr, K, c);
    more.cur = nxt;
    more.meta = deferred_meta(K, c + 1, sh, end_max);
    return live != 0;
}

// Trips c, c+1, ... of one period from M(c) (meta: deferred_meta).  cf / cr = the largest copy
// number with an admissible start on the forward / reverse strand (0: none), df / dr those starts.
template <int W>
__device__ __forceinline__ void extend_trips(const Deferred<W> &d, int &cf, Bits<W> &df, int &cr, Bits<W> &dr)
{
    const int K = (int)(d.meta & 7u), sh = (int)(d.meta >> 12 & 15u), end_max = (int)(d.meta >> 16 & 0x7FFu);   // bits 27..31: the owner lane of a listed pair
    int c = (int)(d.meta >> 3 & 511u);
    Bits<W> cur = d.cur, lim;
    lowmask(lim, end_max - K * c + 1);
    const uint32_t lo_keep = ~lowmask32(sh);
    lim.w[0] &= lo_keep;
    cf = 0; cr = 0;
#pragma unroll
    for (int i = 0; i < W; i++) { df.w[i] = 0; dr.w[i] = 0; }
    for (;;) {
        // starts that can still be admissible with c or more copies (lim only shrinks with c and is
        // empty once K * c > end_max - sh).  Testing cur alone would keep a lane in the loop on
        // repeats outside the admissible range -- above all the run of "no symbol" codes past the
        // window end, which equal each other
        uint32_t live = 0;
#pragma unroll
        for (int i = 0; i < W; i++) live |= cur.w[i] & lim.w[i];
        if (!live) break;
        Bits<W> nxt, nl_;
        and_shr(nxt, cur, cur, K);
        shl(nl_, nxt, K);
        Bits<W> xf, xr;
        uint32_t af = 0, ar = 0;
#pragma unroll
        for (int i = 0; i < W; i++) {
            const uint32_t l = lim.w[i];
            xf.w[i] = cur.w[i] & ~nxt.w[i] & l;
            xr.w[i] = cur.w[i] & ~nl_.w[i] & l;
            af |= xf.w[i]; ar |= xr.w[i];
        }
        if (af) { cf = c; df = xf; }
        if (ar) { cr = c; dr = xr; }
        cur = nxt;
        c++;
        {   // lim for the next copy count: shift the upper bound down by K, keep s >= sh
            Bits<W> u = lim;
            shr(lim, u, K);
            lim.w[0] &= lo_keep;
        }
    }
}

// Forward winner among `cand` (all same k and copies): greatest motif bytes, then greatest
// start (tuple order of pyx:93).  Symbol order '.' < A < C < G < T = (~p2, p1, p0).
template <int W>
__device__ __forceinline__ int pick_forward(const Planes<W> &P, Bits<W> cand, int k)
{
    for (int j = 0; j < k; j++) {
        Bits<W> x2, x1, x0;
        if (j) { shr(x2, P.p2, j, ~0u); shr(x1, P.p1, j, ~0u); shr(x0, P.p0, j); }
        else { x2 = P.p2; x1 = P.p1; x0 = P.p0; }
        Bits<W> t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & ~x2.w[i];
        if (t.any()) cand = t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & x1.w[i];
        if (t.any()) cand = t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & x0.w[i];
        if (t.any()) cand = t;
    }
    return highest_set_le(cand, 32 * W - 1);
}

// Reverse-strand winner: the reported motif is revcomp(t[s:s+k]), compared from its first
// base = complement of t[s+k-1]; complement reverses A<C<G<T and leaves '.' lowest.  Then
// the greatest reverse-strand start = the smallest forward position.
template <int W>
__device__ __forceinline__ int pick_reverse(const Planes<W> &P, Bits<W> cand, int k)
{
    for (int j = k - 1; j >= 0; j--) {
        Bits<W> x2, x1, x0;
        if (j) { shr(x2, P.p2, j, ~0u); shr(x1, P.p1, j, ~0u); shr(x0, P.p0, j); }
        else { x2 = P.p2; x1 = P.p1; x0 = P.p0; }
        Bits<W> t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & ~x2.w[i];
        if (t.any()) cand = t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & ~x1.w[i];
        if (t.any()) cand = t;
#pragma unroll
        for (int i = 0; i < W; i++) t.w[i] = cand.w[i] & ~x0.w[i];
        if (t.any()) cand = t;
    }
    return lowest_set_gt(cand, -1);
}

// all k symbols at pos equal?  (IsHomopolymer, pyx:51-56)
template <int W>
__device__ __forceinline__ bool motif_is_homopolymer(const Planes<W> &P, int pos, int k)
{
    const uint32_t m = (1u << k) - 1u;
    const uint32_t a = extract32(P.p0, pos) & m, b = extract32(P.p1, pos) & m, c = extract32(P.p2, pos) & m;
    return (a == 0 || a == m) && (b == 0 || b == m) && (c == 0 || c == m);
}

// What the owner of a read keeps between the first trips and the winner pick.
template <int W> struct ScanState {
    Planes<W> P;
    Best<W> bf, br;
    int g0;
};

// First trips of every period for one read with a W-word window (seq = first base, L = read
// length; L <= 0 or a window too small for any match: nothing happens, but emit is still called
// for every period so that a warp may vote inside it).  emit(need, deferred) is called once per
// period K = 2..6, in that order, by every caller thread; need = the pair continues in
// extend_trips.
template <int W, class Emit>
__device__ __forceinline__ void scan_first(const uint8_t *seq, int L, const ScanParams &prm, ScanState<W> &S, Emit &&emit)
{
    const int fe = prm.altalgo ? prm.f : 0;          // -x: candidates ignore the flanks
    const int g0 = fe > 6 ? fe - 6 : 0;              // window = [g0, L - g0)
    const int Gn = L - 2 * g0;
    const int sh = fe - g0;
    const int end_max = sh + (L - 2 * fe);           // local bound on a candidate's end
    const int n = prm.n < 1 ? 1 : prm.n;
    const bool alive = Gn > 0 && end_max - sh >= 2 * n;
    const bool use_hm = prm.altalgo && prm.rmhm;
    S.g0 = g0;
    S.bf.k = S.bf.c = S.br.k = S.br.c = 0;
#pragma unroll
    for (int i = 0; i < W; i++) { S.bf.cand.w[i] = 0; S.br.cand.w[i] = 0; }

    Bits<W> nmask;
    encode_window<W>(seq + g0, alive ? Gn : 0, S.P, nmask);
    mark_invalid(S.P, alive ? Gn : 0);
    const Planes<W> &P = S.P;

    // nm_k = positions whose k-window holds an N; hm_k = positions whose k-window is one symbol
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
        Deferred<W> d_;                                                             \
        bool need_ = false;                                                         \
        if (alive && (prm.kmask & (1 << (K))) && end_max - sh >= (K) * n)           \
            need_ = first_trip<K, W>(P, nm, hm, use_hm, sh, end_max, n, S.bf, S.br, d_); \
        emit(need_, d_);                                                            \
    }
    BSTR_PERIOD(2) BSTR_PERIOD(3) BSTR_PERIOD(4) BSTR_PERIOD(5) BSTR_PERIOD(6)
#undef BSTR_PERIOD
}

// The winners of both strands from the candidates collected in S (after every deferred pair of
// the read has been folded into S.bf / S.br with take_if_better).
template <int W>
__device__ __forceinline__ void scan_finish(const ScanState<W> &S, int L, const ScanParams &prm, Hit &fwd, Hit &rc)
{
    fwd.k = 0; rc.k = 0; fwd.copies = rc.copies = fwd.start = rc.start = 0;
    const Planes<W> &P = S.P;
    const int g0 = S.g0;
    if (S.bf.k == 0) return;                         // ProcessSequence: forward None -> drop
    {
        const int pos = pick_forward(P, S.bf.cand, S.bf.k);
        const int start = g0 + pos, len = S.bf.k * S.bf.c;
        if (!prm.altalgo) {                          // -x: tests on the winner only (pyx:153-167)
            if (start < prm.f || L - (start + len) < prm.f) return;
            if (prm.rmhm && motif_is_homopolymer(P, pos, S.bf.k)) return;
        }
        fwd.k = S.bf.k; fwd.copies = S.bf.c; fwd.start = start;
    }
    if (S.br.k == 0) return;
    {
        const int pos = pick_reverse(P, S.br.cand, S.br.k);
        const int len = S.br.k * S.br.c, rstart = L - (g0 + pos + len);
        if (!prm.altalgo) {
            if (rstart < prm.f || g0 + pos < prm.f) return;
            if (prm.rmhm && motif_is_homopolymer(P, pos, S.br.k)) return;
        }
        rc.k = S.br.k; rc.copies = S.br.c; rc.start = rstart;
    }
}

// Both strands of one read with a W-word window, the continuing pairs run inline (callers
// without a warp work list: k_match_reads, the staged kernels, the host emulation of the tests).
template <int W>
__device__ __forceinline__ void match_read_bits(const uint8_t *seq, int L, const ScanParams &prm,
                                                Hit &fwd, Hit &rc)
{
    ScanState<W> S;
    scan_first<W>(seq, L, prm, S, [&](bool need, const Deferred<W> &d) {
        if (!need) return;
        int cf, cr;
        Bits<W> df, dr;
        extend_trips<W>(d, cf, df, cr, dr);
        take_if_better(S.bf, df, (int)(d.meta & 7u), cf);
        take_if_better(S.br, dr, (int)(d.meta & 7u), cr);
    });
    scan_finish<W>(S, L, prm, fwd, rc);
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

__device__ inline Hit match_strand_bytes(const uint8_t *seq, int L, bool rc, const ScanParams &prm,
                                         Hit *pre = nullptr /* -x: the winner before its flank / -r tests */)
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
    if (pre) { pre->k = b_k; pre->copies = b_c; pre->start = b_s; }
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

// ---------------------------------------------------------------------------------------
// -d/--debug (select_STR_reads.pyx:76-87, 138-140): one code per regex hit the reference
// rejects with a message, in the reference's order (patterns by ascending period, starts
// ascending): 'N' = "N in motif.", '5' = "Insufficient 5' flank.", '3' = "Insufficient 3'
// flank." (the last two only in the default algorithm; homopolymer hits under -r are dropped
// silently, pyx:89).  Returns the number of codes; out may be null (count only).
// ---------------------------------------------------------------------------------------
__device__ inline uint32_t debug_strand_codes(const uint8_t *seq, int L, bool rc, const ScanParams &prm,
                                              uint8_t *out)
{
    const int n = prm.n < 1 ? 1 : prm.n;
    uint32_t cnt = 0;
    for (int k = 2; k <= 6; k++) {
        if (!(prm.kmask & (1 << k)) || L < k * n) continue;
        int reach = -1;                      // first p >= i with t[p] != t[p+k] (or p = L-k): run(i) = reach - i
        for (int i = 0; i + k * n <= L; i++) {
            if (reach < i) {
                reach = i;
                while (reach + k < L && strand_char(seq, L, rc, reach) == strand_char(seq, L, rc, reach + k)) reach++;
            }
            const int c = 1 + (reach - i) / k;
            if (c < n) continue;
            bool has_n = false;
            for (int j = 0; j < k; j++) has_n |= strand_char(seq, L, rc, i + j) == 'N';
            uint8_t code = 0;
            if (has_n) code = 'N';
            else if (prm.altalgo) {
                if (i < prm.f) code = '5';
                else if (L - (i + k * c) < prm.f) code = '3';
            }
            if (code) {
                if (out) out[cnt] = code;
                cnt++;
            }
        }
    }
    return cnt;
}

// The -d report of one read (ProcessSequence, pyx:294-311): o[0..3] = forward strand {number of
// codes, 1 if the strand has no best match, tail code, tail value}, o[4..7] the same for the
// reverse complement, which the reference only scans when the forward strand has a match.
// tail = the -x algorithm's message about its winner (pyx:157-163).
__device__ inline void debug_read(const uint8_t *seq, int L, const ScanParams &prm, uint8_t *codes_fwd,
                                  uint8_t *codes_rc, uint32_t (&o)[8])
{
    o[0] = 0; o[1] = 1; o[2] = 0; o[3] = 0; o[4] = 0; o[5] = 1; o[6] = 0; o[7] = 0;
    for (int strand = 0; strand < 2; strand++) {
        const bool rc = strand == 1;
        o[4 * strand] = debug_strand_codes(seq, L, rc, prm, rc ? codes_rc : codes_fwd);
        Hit pre;
        const Hit h = match_strand_bytes(seq, L, rc, prm, &pre);
        o[4 * strand + 1] = h.k == 0;
        if (!prm.altalgo && pre.k) {
            const int end = pre.start + pre.k * pre.copies;
            if (pre.start < prm.f) { o[4 * strand + 2] = '5'; o[4 * strand + 3] = (uint32_t)pre.start; }
            else if (L - end < prm.f) { o[4 * strand + 2] = '3'; o[4 * strand + 3] = (uint32_t)(L - end); }
        }
        if (h.k == 0) break;                               // pyx:303-305: no second strand
    }
}

// upper bound of debug_strand_codes for a read of length L (every start of every period)
__host__ __device__ inline uint32_t debug_codes_cap(int L, int n, int kmask)
{
    if (n < 1) n = 1;
    uint32_t cap = 0;
    for (int k = 2; k <= 6; k++)
        if ((kmask & (1 << k)) && L >= k * n) cap += (uint32_t)(L - k * n + 1);
    return cap;
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
// branch does not diverge.  WMAX (2, 4, 6 or 8) is the widest bit-parallel window compiled into
// the calling kernel: a kernel built for short reads carries neither the code nor the
// registers of the wide windows; reads that need more words take the exact byte path.
template <int WMAX = 8>
__device__ __forceinline__ void match_read(const uint8_t *seq, int L, const ScanParams &prm,
                                           int wmax, Hit &fwd, Hit &rc)
{
    if (wmax <= 2) match_read_bits<2>(seq, L, prm, fwd, rc);
    else if (WMAX >= 4 && wmax <= 4) match_read_bits<(WMAX >= 4 ? 4 : 2)>(seq, L, prm, fwd, rc);
    else if (WMAX == 6 && wmax <= 6) match_read_bits<(WMAX == 6 ? 6 : 2)>(seq, L, prm, fwd, rc);
    else if (WMAX >= 8 && wmax <= 8) match_read_bits<(WMAX >= 8 ? 8 : 2)>(seq, L, prm, fwd, rc);
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

// Bit 0 of every byte lane is set iff that byte is NOT one of acgtACGT -- the cheap test (9 ops per
// 4 bases instead of 22): bits 7,6,3 must read 0,1,0 (bit 5 = case is ignored) and, of the
// remaining bits, b4 = 0 needs b0 & (b1 | ~b2) (A 001, C 011, G 111) and b4 = 1 needs b2 & ~b1 & ~b0
// (T 100).  Exactly eight byte values pass, all of them valid.  N, n and '.' fail it and go through
// swar_invalid.
__device__ __forceinline__ uint32_t swar_not_acgt(uint32_t w)
{
    const uint32_t s1 = w >> 1, s2 = w >> 2, s4 = w >> 4;
    const uint32_t t = w & (s1 | ~s2);
    const uint32_t g = s2 & ~s1 & ~w;
    const uint32_t okbit = (s4 & g) | (~s4 & t);
    const uint32_t hi = (w ^ 0x40404040u) & 0xC8C8C8C8u;           // bits 7, 6, 3 differ from 0, 1, 0
    return (~okbit & 0x01010101u) | hi;                            // any bit set: some byte is not ACGT/acgt
}

// Any byte of [p, p+L) outside acgtnACGTN. ?  Reads whole aligned words around the range.
__device__ __forceinline__ bool bases_invalid(const uint8_t *p, int L)
{
    const uintptr_t a = (uintptr_t)p;
    const uint32_t *wp = (const uint32_t *)(p - (a & 3));          // keeps p's address space (LDS, not LD)
    if (L <= 0) return false;
    const int head = (int)(a & 3), total = head + L, nw = (total + 3) >> 2;
    // bytes outside the range are replaced by 'A'; only the first and last word need it
    const uint32_t mfirst = 0xFFFFFFFFu << (8 * head);
    const int r = total - 4 * (nw - 1);                        // 1..4 bytes in the last word
    const uint32_t mlast = r < 4 ? (1u << (8 * r)) - 1u : 0xFFFFFFFFu;
    uint32_t acc = 0;
    // Four words per trip: the loads go out together and the cheap ACGT test decides for the group;
    // only a group with an N, a '.' or an invalid byte (one trip in ten of a warp on the bench
    // workloads) gets the exact test.
    for (int i = 0; i < nw; i += 4) {
        uint32_t w0 = wp[i], w1 = i + 1 < nw ? wp[i + 1] : 0x41414141u, w2 = i + 2 < nw ? wp[i + 2] : 0x41414141u,
                 w3 = i + 3 < nw ? wp[i + 3] : 0x41414141u;
        if (i == 0) w0 = (w0 & mfirst) | (0x41414141u & ~mfirst);
        if (i + 4 >= nw) {                                       // the group holds the last word
            const int k = nw - 1 - i;
            if (k == 0) w0 = (w0 & mlast) | (0x41414141u & ~mlast);
            else if (k == 1) w1 = (w1 & mlast) | (0x41414141u & ~mlast);
            else if (k == 2) w2 = (w2 & mlast) | (0x41414141u & ~mlast);
            else w3 = (w3 & mlast) | (0x41414141u & ~mlast);
        }
        if (swar_not_acgt(w0) | swar_not_acgt(w1) | swar_not_acgt(w2) | swar_not_acgt(w3))
            acc |= swar_invalid(w0) | swar_invalid(w1) | swar_invalid(w2) | swar_invalid(w3);
    }
    return (acc & 0x01010101u) != 0;
}

}  // namespace bstr
