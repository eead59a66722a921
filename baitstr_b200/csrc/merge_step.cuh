// merge_step.cuh -- what merge_STR_reads does with ONE read (merge_STR_reads.c:357-422), written
// against an executor X so that the same text runs as a warp on the device (merge.cu: the 32 lanes
// fill the alignment matrix together, lane 0 walks the traceback) and as plain code on the host
// (tests/emul/emul_merge.cpp):
//   X.dp_fill(t1, len1, t2, len2, T)   Align's matrix (:86-110) into T, returns the best cell's key
//   X.leader()                         the one thread that runs the sequential parts
//   X.bcast(v)                         the leader's value to everybody
//   X.sync()                           the leader's writes are visible to everybody afterwards
#pragma once
#include "merge_core.cuh"

#ifdef __CUDACC__
#define BM_X __device__ __forceinline__
#else
#define BM_X inline
#endif

namespace bmerge {

struct Scratch {                   // per executor
    uint8_t *T;                    // t_bytes_for(maxf)
    char *lseq, *lqual, *rseq, *rqual;   // cap each
    int maxf, pitch, cap;          // the geometry: longest flank, row pitch of T (pitch_for(maxf)), room for a consensus flank
};

// AlignFlanks (:215-306).  Returns 1 merged, 0 not, < 0 = -(ERR_*)
template <class X>
BM_X int try_merge(X &x, BlockHead *h, char *bseq, char *bqual, const char *bases, const char *quals, int slen,
                   const char *motif, int mlen, unsigned copies, int zstart, int end, const Scratch &s)
{
    const int zb = h->zstart, eb = h->end, sb = h->slen;
    // the sequences to the left of the STR: block [0, zstart], read [0, zstart]
    int len1 = zb + 1, len2 = zstart + 1;
    if (len1 - 1 > s.maxf || len2 - 1 > s.maxf || sb - eb > s.maxf || slen - end > s.maxf) return -ERR_FLANK_TOO_LONG;
    uint64_t best = x.dp_fill(bseq, len1, bases, len2, s.T);
    int verdict = 0, ln = 0;
    if (x.leader()) {
        AlignOut o;
        align_trace(bseq, bqual, len1, bases, quals, len2, s.T, s.pitch, s.cap, best, false, s.lseq, s.lqual, &o);
        ln = o.n;
        verdict = o.err ? -o.err : (align_rejected(o) ? 0 : 1);
    }
    verdict = x.bcast(verdict);
    if (verdict <= 0) return verdict;
    // to the right of it: block [end, slen], read [end, slen]
    len1 = sb - eb + 1; len2 = slen - end + 1;
    best = x.dp_fill(bseq + eb, len1, bases + end, len2, s.T);
    if (x.leader()) {
        AlignOut o;
        align_trace(bseq + eb, bqual + eb, len1, bases + end, quals + end, len2, s.T, s.pitch, s.cap, best, true, s.rseq, s.rqual, &o);
        verdict = o.err ? -o.err : (align_rejected(o) ? 0 : 1);
        if (verdict == 1) {
            const int e = apply_merge(h, bseq, bqual, s.lseq, s.lqual, ln, s.rseq, s.rqual, o.n, motif, mlen, copies);
            if (e) verdict = -e;
        }
    }
    verdict = x.bcast(verdict);
    x.sync();
    return verdict;
}

// one trip of the reference's read loop.  0, or an ERR_* code
template <class X>
BM_X int process_read(X &x, const Read &r, int index, BlockHead *heads, char *bseqs, char *bquals,
                      const char *fwd, const char *qfwd, const char *rev, const char *qrev, const Scratch &s)
{
    int merged = 0;
    BlockHead *hf = heads + r.fkey;
    if (hf->created_by >= 0)                                                         // :367-378
        merged = try_merge(x, hf, bseqs + (size_t)r.fkey * MAXB, bquals + (size_t)r.fkey * MAXB, fwd + r.off, qfwd + r.off, r.slen,
                           r.fmotif, r.fmlen, r.fcopies, r.fzstart, r.fend, s);
    if (merged < 0) return -merged;
    if (merged) return ERR_NONE;
    BlockHead *hr = heads + r.rkey;                                                  // :380-403, the read reverse-complemented
    if (hr->created_by >= 0) {
        merged = try_merge(x, hr, bseqs + (size_t)r.rkey * MAXB, bquals + (size_t)r.rkey * MAXB, rev + r.off, qrev + r.off, r.slen,
                           r.rmotif, r.rmlen, r.rcopies, r.rzstart, r.rend, s);
        // merged nowhere although its key is in the map: as the reference, the new block is hung on a
        // local copy of the list head (:417-419) and never seen again -- the read is dropped
        return merged < 0 ? -merged : ERR_NONE;
    }
    int e = 0;                                                                       // :405-421
    if (x.leader()) e = create_block(hr, bseqs + (size_t)r.rkey * MAXB, bquals + (size_t)r.rkey * MAXB, r, rev + r.off, qrev + r.off, index);
    e = x.bcast(e);
    x.sync();
    return e;
}

}  // namespace bmerge
