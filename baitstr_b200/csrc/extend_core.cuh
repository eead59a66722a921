// extend_core.cuh -- what extend_STR_reads does with ONE merged block after its k-mer counting pass:
// the walk through the k-mer graph on both sides of the STR.  Written against an executor X so that
// the same text runs as a warp on the device (kmer_api.inc: eight lanes probe the eight neighbours
// of a k-mer at once, all lanes scan the k-mers already used) and as plain code on the host
// (tests/emul/emul_extend.cpp).  Every thread of an executor runs this text with the same values.
//
// What it restates, from /root/reference/src/extend_STR_reads.c:
//   RvKmers / FwKmers            :195-216   the four predecessors / successors of a k-mer
//   CheckForSNPBackwards/Forwards :218-438  a fork is passed if both arms meet again after klen + 1 steps
//   ExtendBackward / ExtendForward :471-734 walk while exactly one neighbour is in the table (or the
//                                           fork is such a bubble), stop at a k-mer already used or
//                                           after `flanks` k-mers; the flank as a string
//   PercentIdentity              :740-772   the flank against the block it grew from (>= 95 % or dropped)
//   FindFirstGoodKmer            :809-850   where the walks start
//   the per-block driver         :915-993
// The table is the reference's map after its pruning: canonical k-mers whose count, saturated at 255,
// lies within the thresholds.  Look-ups are made with the k-mer AS IT IS (not canonicalised), as the
// reference does: of a k-mer and its reverse complement only the smaller one is ever a key.
//   X.lookup8(cand)          bit p set iff cand[p] is in the table (ABSENT = a value that never is)
//   X.seen(list, n, a, b)    whether a or b is among list[0 .. n)
//   X.leader() / X.sync()    one thread writes the lists and strings; sync makes that visible
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define BE_X __device__ __forceinline__
#define BE_POPC(x) __popc(x)
#else
#define BE_X inline
#define BE_POPC(x) __builtin_popcount(x)
#endif

namespace bextend {

constexpr uint64_t ABSENT = ~0ull;

enum { ST_PRINTED = 1, ST_NO_LEFT_ROOM = 2, ST_LEFT_PID = 3, ST_RIGHT_PID = 4, ST_RIGHT_SHORT = 5 };

struct Block {                     // one record of the merge stage's output (extend_STR_reads.c:921-927)
    uint32_t off;                  // offset of its bases in the blob
    uint32_t slen;
    int32_t zstart, end;
};

struct Result {
    int32_t status;                // ST_*
    uint32_t indx1, indx2;         // the block's own bases printed are [indx1, indx2)
    uint32_t llen, rlen;           // lengths of the flanks
};

BE_X long long code_of(unsigned char c)
{   // fasta_encoding as it stands (utilities.h:61-75): ACGT in either case 0..3, N and n 0, everything else -1
    const unsigned char u = c & 0xDF;
    return (u == 'A' || u == 'N') ? 0 : u == 'C' ? 1 : u == 'G' ? 2 : u == 'T' ? 3 : -1;
}

BE_X uint64_t revcomp(uint64_t word, int klen)
{   // ReverseComplementKmer (kmer.c:36-47): the low klen digits reversed, each d -> 3 - d
    uint64_t rc = 0;
    for (int i = 0; i < klen; i++) { rc = (rc << 2) + (3 - (word & 3)); word >>= 2; }
    return rc;
}

BE_X uint64_t next_kmer(uint64_t word, long long code, int klen)
{   // GetNextKmer (kmer.c:24-33)
    return ((word << (64 - 2 * klen + 2)) >> (64 - 2 * klen)) + (uint64_t)code;
}

BE_X uint64_t kmer_at(const char *s, int klen)
{   // ConvertStringToKmer (kmer.c:64-69): BuildIndex over klen - 1 characters, then one GetNextKmer
    uint64_t word = 0;
    for (int i = 0; i < klen - 1; i++) word = (word << 2) + (uint64_t)code_of((unsigned char)s[i]);
    return next_kmer(word, code_of((unsigned char)s[klen - 1]), klen);
}

// the eight neighbours of curr on one side: own[i] continue curr's strand, other[i] are the same steps
// seen from the reverse complement; walk[p] = the k-mer the walk would stand on after taking cand[p]
template <bool BACKWARD, bool OWN_FIRST>
BE_X void neighbours(uint64_t curr, int klen, uint64_t cand[8], uint64_t walk[8])
{
    const uint64_t rev = revcomp(curr, klen);
    for (int i = 0; i < 4; i++) {
        uint64_t own, other;
        if (BACKWARD) { own = (curr >> 2) + ((uint64_t)i << (2 * klen - 2)); other = next_kmer(rev, i, klen); }     // RvKmers(curr), FwKmers(rev)
        else          { own = next_kmer(curr, i, klen); other = (rev >> 2) + ((uint64_t)i << (2 * klen - 2)); }     // FwKmers(curr), RvKmers(rev)
        const int po = OWN_FIRST ? 2 * i : 2 * i + 1, pt = OWN_FIRST ? 2 * i + 1 : 2 * i;
        cand[po] = own; walk[po] = own;
        cand[pt] = other; walk[pt] = revcomp(other, klen);
    }
}

// CheckForSNPBackwards / CheckForSNPForwards: called at a fork with exactly two arms
template <bool BACKWARD, class X>
BE_X bool fork_is_a_bubble(X &x, uint64_t curr, int klen)
{
    uint64_t cand[8], walk[8];
    neighbours<BACKWARD, true>(curr, klen, cand, walk);
    unsigned hits = x.lookup8(cand);
    uint64_t arm[2] = {0, 0}, reached[2] = {0, 0};
    int n = 0;
    for (int p = 0; p < 8 && n < 2; p++) if (hits >> p & 1) arm[n++] = walk[p];       // the first two, in the reference's order
    for (int a = 0; a < 2; a++) {
        uint64_t at = arm[a];
        int steps = 0;
        while (steps <= klen) {
            neighbours<BACKWARD, true>(at, klen, cand, walk);
            hits = x.lookup8(cand);
            if (hits == 0 || (hits & (hits - 1))) break;                              // not exactly one way on
            int p = 0;
            while (!(hits >> p & 1)) p++;
            at = walk[p];
            steps++;
        }
        if (steps != klen + 1) return false;
        reached[a] = at;
    }
    return reached[0] == reached[1];
}

// ExtendBackward / ExtendForward up to the list of k-mers used; returns how many (>= 1).  `start` is the
// k-mer the walk starts on, `used` has room for `flanks` k-mers
template <bool BACKWARD, class X>
BE_X uint32_t walk_side(X &x, uint64_t start, int klen, uint32_t flanks, uint64_t *used)
{
    uint32_t n = 0;
    if (x.leader()) used[0] = start;
    n = 1;
    x.sync();
    uint64_t curr = start;
    for (;;) {
        uint64_t cand[8], walk[8];
        // the main loops test rvs[i] before fws[i] on both sides (:497-508, :649-660): own first when
        // walking backwards, the other strand first when walking forwards
        neighbours<BACKWARD, BACKWARD>(curr, klen, cand, walk);
        const unsigned hits = x.lookup8(cand);
        const int count = BE_POPC(hits);
        if (count == 2) { if (!fork_is_a_bubble<BACKWARD>(x, curr, klen)) break; }   // rescued if it is only a substitution (:511-517)
        else if (count != 1) break;
        int last = 7;
        while (!(hits >> last & 1)) last--;                                           // `extension` is the last one found
        const uint64_t extension = cand[last], rev_extension = revcomp(extension, klen);
        if (x.seen(used, n, extension, rev_extension)) break;                         // already in this contig (:528-536)
        curr = walk[last];
        if (x.leader()) used[n] = curr;
        n++;
        x.sync();
        if (n == flanks) break;                                                       // :563-565 (not INFEXPAND)
    }
    return n;
}

BE_X char base_of(uint64_t digit) { return "ACGT"[digit & 3]; }

// the left flank (:568-577): the last k-mer spelled out, then the last base of every earlier one, newest first
BE_X uint32_t spell_left(const uint64_t *used, uint32_t n, int klen, char *out)
{
    const uint64_t last = used[n - 1];
    for (int i = 0; i < klen; i++) out[i] = base_of(last >> (2 * (klen - 1 - i)));
    for (uint32_t i = 0; i + 1 < n; i++) out[n + klen - i - 2] = base_of(used[i]);
    out[klen + n - 1] = 0;
    return (uint32_t)klen + n - 1;
}

// the right flank (:713-722): the first base of every k-mer, then the rest of the last one
BE_X uint32_t spell_right(const uint64_t *used, uint32_t n, int klen, char *out)
{
    for (uint32_t i = 0; i < n; i++) out[i] = base_of(used[i] >> (2 * (klen - 1)));
    const uint64_t last = used[n - 1];
    for (int i = 1; i < klen; i++) out[n + i - 1] = base_of(last >> (2 * (klen - 1 - i)));
    out[n + klen - 1] = 0;
    return n + (uint32_t)klen - 1;
}

// PercentIdentity(...) < 95.00 (:740-772).  0 / 0 is NaN there and NaN < 95 is false
BE_X bool pid_below_95(const char *flank, uint32_t flen, const char *bases, uint32_t blen, uint32_t index, bool reverse)
{
    uint32_t matches = 0, total = 0;
    if (!reverse) {
        const uint32_t n = blen - index < flen ? blen - index : flen;
        for (uint32_t i = 0; i < n; i++) { matches += flank[i] == bases[index + i]; total++; }
    } else {
        for (long long a = (long long)flen - 1, b = (long long)index - 1; a >= 0 && b >= 0; a--, b--) { matches += flank[a] == bases[b]; total++; }
    }
    return total > 0 && 100ull * matches < 95ull * total;
}

// FindFirstGoodKmer (:809-850) over bases[0 .. num_kmers + klen - 1)
template <class X>
BE_X uint32_t first_good_kmer(X &x, const char *bases, uint32_t num_kmers, bool return_on_first, int klen)
{
    uint32_t i = 0, j = 0;
    for (i = 0; i < num_kmers; i += 8) {                    // eight positions per look-up round
        uint64_t cand[8];
        for (uint32_t p = 0; p < 8; p++) {
            if (i + p < num_kmers) {
                const uint64_t word = kmer_at(bases + i + p, klen), anti = revcomp(word, klen);
                cand[p] = word < anti ? word : anti;
            } else {
                cand[p] = ABSENT;
            }
        }
        const unsigned hits = x.lookup8(cand);
        if (!hits) continue;
        if (return_on_first) {
            uint32_t p = 0;
            while (!(hits >> p & 1)) p++;
            return i + p + (uint32_t)klen;
        }
        uint32_t p = 7;
        while (!(hits >> p & 1)) p--;
        j = i + p;
    }
    if (return_on_first) return num_kmers + (uint32_t)klen;  // none: i == num_kmers when the reference's loop ends
    return j;
}

// the driver's loop body (:915-993).  lflank / rflank have room for klen + flanks bytes, used for flanks k-mers
template <class X>
BE_X void extend_block(X &x, const Block &b, const char *blob, int klen, uint32_t flanks, uint64_t *used, char *lflank, char *rflank, Result *out)
{
    const char *bases = blob + b.off;
    Result r = {0, 0, 0, 0, 0};
    if (klen > b.zstart) {                                  // no room for a k-mer left of the STR: never printed (:931-939, :783)
        r.status = ST_NO_LEFT_ROOM;
    } else if (b.slen - (uint32_t)b.end < (uint32_t)klen) { // the reference walks on from bytes behind the block here (:958-969): refused
        r.status = ST_RIGHT_SHORT;
    } else {
        r.indx1 = first_good_kmer(x, bases, (uint32_t)(b.zstart - klen + 1), true, klen);
        uint32_t n = walk_side<true>(x, kmer_at(bases + r.indx1 - klen, klen), klen, flanks, used);
        if (x.leader()) r.llen = spell_left(used, n, klen, lflank);
        r.llen = (uint32_t)klen + n - 1;
        x.sync();
        if (pid_below_95(lflank, r.llen, bases, b.slen, r.indx1, true)) {
            r.status = ST_LEFT_PID;
        } else {
            r.indx2 = (uint32_t)b.end + first_good_kmer(x, bases + b.end, b.slen - (uint32_t)b.end - (uint32_t)klen + 1, false, klen);
            n = walk_side<false>(x, kmer_at(bases + r.indx2, klen), klen, flanks, used);
            if (x.leader()) spell_right(used, n, klen, rflank);
            r.rlen = n + (uint32_t)klen - 1;
            x.sync();
            r.status = pid_below_95(rflank, r.rlen, bases, b.slen, r.indx2, false) ? ST_RIGHT_PID : ST_PRINTED;
        }
    }
    if (x.leader()) *out = r;
}

}  // namespace bextend
