// merge_core.cuh -- the per-read step of merge_STR_reads (first slice), written
// once for the device (a warp per chain of reads, merge.cu) and for the host build of the same code
// that the CPU suite drives (tests/emul/emul_merge.cpp).
//
// What it restates, from /root/reference/src/merge_STR_reads.c:
//   Align        :43-213   local alignment (match 1, mismatch -1, gap -3) of two flanks, then ONE
//                          traceback from the last best cell that builds the consensus flank,
//                          quality by quality, and counts gaps / matches / mismatches
//   AlignFlanks  :215-306  left flanks, right flanks, the 90 % identity and <= 2 gaps tests, and
//                          the new block = left consensus + one motif + right consensus
// The quirks are kept, each marked "as the reference": they decide which reads merge.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define BM_HD __host__ __device__ __forceinline__
#else
#define BM_HD inline
#endif

namespace bmerge {

// Consensus flanks can outgrow the reads (the right flank's overhang is taken from column j of the
// LONGER sequence, see align_trace), so the bounds leave room beyond a 250-base read's ~215.
constexpr int MAXF = 383;          // longest flank that can be aligned (DP side = flank + 1 <= 384)
constexpr int PITCH = 401;         // row pitch of the traceback matrix, in cells: the cells of one anti-diagonal lie PITCH - 1 = 400
                                   // cells = 25 words apart, an odd number of shared-memory banks, so 32 lanes hit 32 banks
constexpr int MAXB = 800;          // longest block / consensus sequence held
constexpr int MAXSUP = 24;         // distinct copy numbers remembered per block
constexpr int T_BYTES = ((MAXF + 1) * PITCH / 4 + 15) / 16 * 16;   // traceback matrix: two bits per cell

enum { ERR_NONE = 0, ERR_FLANK_TOO_LONG = 1, ERR_BLOCK_TOO_LONG = 2, ERR_TOO_MANY_ALLELES = 3 };

// what the reference keeps per block (sparse_word_hash.h:31-41) -- without the sequence
struct BlockHead {
    int32_t created_by;            // index of the read that created it, -1 = the key is not in the map
    uint16_t zstart, end, slen, support;             // uint16_t as in the reference (support wraps there too)
    uint16_t nsup;
    uint16_t sup_copies[MAXSUP], sup_n[MAXSUP];      // in order of insertion; the reference's list is head-inserted, i.e. read backwards
};

struct Read {                      // one record of the select stage's output, parsed (merge_STR_reads.c:350-355)
    uint32_t off;                  // offset of its bases in the forward blob and in the reverse-complement blob, and of its qualities
    uint16_t slen;
    uint16_t fcopies, fzstart, fend, rcopies, rzstart, rend;
    uint8_t fmlen, rmlen;
    char fmotif[7], rmotif[7];
    uint32_t fkey, rkey;           // ids of "motif lflank rflank" for the read as it is / reverse-complemented (:361-365, :385-390)
};

// the traceback matrix holds B[i][j] only (0 diagonal, 1 up, 2 left): the scores the reference reads
// back from A during the traceback follow from the moves (a diagonal move came from A - (+1 | -1), a
// gap move from A + 3), so A is never stored
BM_HD void t_put(uint8_t *T, int pitch, int i, int j, unsigned v)
{
    const int at = i * pitch + j, sh = (at & 3) * 2;
    uint8_t &b = T[at >> 2];
    b = (uint8_t)((b & ~(3u << sh)) | (v << sh));
}
BM_HD unsigned t_get(const uint8_t *T, int pitch, int i, int j)
{
    const int at = i * pitch + j;
    return (T[at >> 2] >> ((at & 3) * 2)) & 3u;
}

// one cell of Align's recurrence (:86-110): returns A[i][j], *code = B[i][j]
BM_HD int dp_cell(int left, int up, int diag, bool same, unsigned *code)
{
    const int score1 = left - 3, score2 = up - 3, score3 = diag + (same ? 1 : -1);
    int a = score1 > score2 ? score1 : score2;
    const int b = score3 > 0 ? score3 : 0;
    a = a > b ? a : b;
    const unsigned bt = a == score3 ? 0u : a == score2 ? 1u : a == score1 ? 2u : 0u;   // B stays 0 when no score is the maximum (calloc)
    *code = bt;
    return a;
}

// "if (A[i][j] >= best)" over a row-major scan keeps the LAST best cell: the maximum of this key
BM_HD uint64_t best_key(int score, int i, int j) { return ((uint64_t)(uint32_t)score << 32) | ((uint64_t)(uint32_t)i << 16) | (uint64_t)(uint32_t)j; }

struct AlignOut {
    int n;                         // consensus length
    unsigned gaps;
    int matches, mismatches;
    int err;
};

// A geometry smaller than the largest (MAXF, PITCH, MAXB) lets more warps share an SM's shared memory; a
// chain that outgrows it is walked again with the largest one (merge.cu).  The row pitch is chosen so
// that the cells of one anti-diagonal, pitch - 1 apart, fall into different shared-memory banks:
// (pitch - 1) / 16 words, an odd number
BM_HD int pitch_for(int maxf) { int p = maxf + 1; while ((p - 1) % 32 != 16) p++; return p; }
BM_HD int t_bytes_for(int maxf) { return ((maxf + 1) * pitch_for(maxf) / 4 + 15) / 16 * 16; }

// the traceback of Align (:112-205) from the best cell; consensus into seq/qual (capacity cap <= MAXB)
BM_HD void align_trace(const char *t1, const char *q1, int len1, const char *t2, const char *q2, int len2,
                       const uint8_t *T, int pitch, int cap, uint64_t best, bool right_gapped, char *seq, char *qual, AlignOut *o)
{
    int sindx = 0, err = ERR_NONE;
    unsigned i = (unsigned)(best >> 16) & 0xFFFFu, j = (unsigned)best & 0xFFFFu;
    int max_score = (int)(best >> 32);
    int matches = 0, mismatches = 0;
    unsigned num_gaps = 0;
#define BM_EMIT(c, q) do { if (sindx < cap - 1) { seq[sindx] = (c); qual[sindx] = (q); sindx++; } else err = ERR_BLOCK_TOO_LONG; } while (0)
    if (right_gapped) {
        // as the reference: the overhang of the LONGER flank beyond the best cell's column j -- compared
        // with j also when it walks the first sequence (:128-133)
        if (len1 > len2) { for (int a = len1 - 1; a > (int)j; a--) BM_EMIT(t1[a - 1], q1[a - 1]); }
        else             { for (int a = len2 - 1; a > (int)j; a--) BM_EMIT(t2[a - 1], q2[a - 1]); }
    } else {
        num_gaps = (unsigned)(len2 - (int)j - 1);
    }
    while (max_score > 0 && i >= 1 && j >= 1) {
        const unsigned b = t_get(T, pitch, (int)i, (int)j);
        if (b == 0) {
            if (q1[i - 1] > q2[j - 1]) BM_EMIT(t1[i - 1], q1[i - 1]); else BM_EMIT(t2[j - 1], q2[j - 1]);
            if (t1[i - 1] != t2[j - 1]) { mismatches++; max_score += 1; } else { matches++; max_score -= 1; }
            i--; j--;
        } else if (b == 1) {
            if (q1[i - 1] > '5') BM_EMIT(t1[i - 1], q1[i - 1]);
            i--; num_gaps++; max_score += 3;
        } else {
            if (q2[j - 1] > '5') BM_EMIT(t2[j - 1], q2[j - 1]);
            j--; num_gaps++; max_score += 3;
        }
        // max_score is now A[i][j] of the cell moved to ("max_score = A[i][j]", :170)
    }
    if (right_gapped) {
        num_gaps = i > j ? i - j : j - i;                                   // as the reference (:174-175): replaces the count
    } else {
        if (len1 > len2) { for (int a = (int)i; a > 0; a--) BM_EMIT(t1[a - 1], q1[a - 1]); }
        else             { for (int a = (int)j; a > 0; a--) BM_EMIT(t2[a - 1], q2[a - 1]); }
    }
#undef BM_EMIT
    for (int a = 0, b = sindx - 1; a < b; a++, b--) {
        char c = seq[a]; seq[a] = seq[b]; seq[b] = c;
        c = qual[a]; qual[a] = qual[b]; qual[b] = c;
    }
    o->n = sindx; o->gaps = num_gaps; o->matches = matches; o->mismatches = mismatches; o->err = err;
}

// "(pid < pid_threshold) || (gaps > 2)" (:234, :251): pid = matches * 100.0 / (matches + mismatches)
// is NaN for 0 / 0 and NaN < 90 is false -- an empty alignment passes, as in the reference
BM_HD bool align_rejected(const AlignOut &o)
{
    const int t = o.matches + o.mismatches;
    return (t > 0 && 100 * o.matches < 90 * t) || o.gaps > 2;
}

// the tail of AlignFlanks (:262-303): the block becomes left consensus + motif + right consensus
BM_HD int apply_merge(BlockHead *h, char *bseq, char *bqual, const char *lseq, const char *lqual, int ln,
                      const char *rseq, const char *rqual, int rn, const char *motif, int mlen, unsigned copies)
{
    if (ln + mlen + rn > MAXB - 1) return ERR_BLOCK_TOO_LONG;
    h->zstart = (uint16_t)ln;
    h->support++;
    bool found = false;
    for (int s = 0; s < h->nsup; s++)
        if (h->sup_copies[s] == (uint16_t)copies) { h->sup_n[s]++; found = true; break; }
    if (!found) {
        if (h->nsup >= MAXSUP) return ERR_TOO_MANY_ALLELES;
        h->sup_copies[h->nsup] = (uint16_t)copies; h->sup_n[h->nsup] = 1; h->nsup++;
    }
    h->end = (uint16_t)(ln + mlen);
    h->slen = (uint16_t)(ln + mlen + rn);
    for (int a = 0; a < ln; a++) { bseq[a] = lseq[a]; bqual[a] = lqual[a]; }
    for (int a = 0; a < mlen; a++) { bseq[ln + a] = motif[a]; bqual[ln + a] = '!'; }
    for (int a = 0; a < rn; a++) { bseq[ln + mlen + a] = rseq[a]; bqual[ln + mlen + a] = rqual[a]; }
    bseq[h->slen] = 0; bqual[h->slen] = 0;
    return ERR_NONE;
}

// a new block from a read that merged nowhere (:405-416): the read as it stands (reverse-complemented)
BM_HD int create_block(BlockHead *h, char *bseq, char *bqual, const Read &r, const char *bases, const char *quals, int index)
{
    if (r.slen > MAXB - 1) return ERR_BLOCK_TOO_LONG;
    h->created_by = index;
    h->zstart = r.rzstart; h->end = r.rend; h->slen = r.slen; h->support = 1;
    h->nsup = 1; h->sup_copies[0] = r.rcopies; h->sup_n[0] = 1;
    for (int a = 0; a < r.slen; a++) { bseq[a] = bases[a]; bqual[a] = quals[a]; }
    bseq[r.slen] = 0; bqual[r.slen] = 0;
    return ERR_NONE;
}

// what the reference would print of a block: the first (up to three) copy numbers with >= 2
// supporting reads, newest entry of the list first (:445-455), and whether the block is printed (:457-460)
struct Verdict { unsigned copies[3], maxcopies; bool print; };
BM_HD Verdict block_verdict(const BlockHead &h, unsigned min_threshold, unsigned max_threshold, bool include_all)
{
    Verdict v = {{0, 0, 0}, 0, false};
    int i = 0;
    for (int s = (int)h.nsup - 1; s >= 0 && i < 3; s--)      // the reference's list has the newest entry first
        if (h.sup_n[s] >= 2) { v.copies[i] = h.sup_copies[s]; if (v.copies[i] > v.maxcopies) v.maxcopies = v.copies[i]; i++; }
    v.print = h.support >= min_threshold && h.support <= max_threshold &&
              ((!include_all && v.copies[2] == 0 && v.copies[1] != 0) || (include_all && v.copies[2] == 0));
    return v;
}

}  // namespace bmerge
