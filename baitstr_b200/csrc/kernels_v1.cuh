// kernels_v1.cuh -- the staged device pipeline (index -> scan -> offsets -> format).
//
// One batch = a byte range of raw FASTQ that starts on a record boundary.
//   k_count_newlines / k_fill_newlines   line index (replaces readline(), fastq.pyx:72-90)
//   k_scan_records                       validation + both-strand match per record
//                                        (fastq.pyx:77-87, select_STR_reads.pyx:294-319)
//   k_exclusive_scan                     ordered output offsets (input order is preserved,
//                                        select_STR_reads.pyx:363-367)
//   k_format_records                     PrintSequence (select_STR_reads.pyx:171-185)
#pragma once
#include <stdint.h>
#include "matcher.cuh"

namespace bstr {

constexpr int IDX_THREADS = 256;
constexpr int IDX_TILE = IDX_THREADS * 16;     // bytes per index block

struct BatchState {                 // lives in device memory, one per in-flight batch
    unsigned long long first_bad;   // min over bad records of (record << 8 | code); ~0 = none
    unsigned long long n_lines;     // terminated lines (incl. the virtual one at EOF)
    unsigned long long n_records;
    unsigned long long n_selected;
    unsigned long long out_bytes;
    unsigned long long consumed;
    unsigned long long fallback;    // fused path only: input needs the generic kernels
};

enum { BAD_HEADER = 1, BAD_BASES = 2, BAD_SEPARATOR = 3, BAD_TRUNCATED = 4 };

// 0x80 in every byte of w that equals '\n'
__device__ __forceinline__ uint32_t nl_flags(uint32_t w)
{
    const uint32_t x = w ^ 0x0A0A0A0Au;
    const uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | x | 0x7F7F7F7Fu);
}

// 16-bit mask of newline positions in the 16 bytes at `base + 16*chunk` (bytes >= n ignored)
__device__ __forceinline__ uint32_t nl_mask16(const uint8_t *base, uint64_t chunk, uint64_t n)
{
    const uint64_t off = chunk * 16;
    if (off >= n) return 0;
    const uint4 v = *reinterpret_cast<const uint4 *>(base + off);
    uint32_t m;
    const uint32_t f0 = nl_flags(v.x), f1 = nl_flags(v.y), f2 = nl_flags(v.z), f3 = nl_flags(v.w);
    // compress 0x80 flags (bits 7,15,23,31) to 4 bits each
    m = (((f0 >> 7) & 0x01010101u) * 0x01020408u >> 24) |
        ((((f1 >> 7) & 0x01010101u) * 0x01020408u >> 24) << 4) |
        ((((f2 >> 7) & 0x01010101u) * 0x01020408u >> 24) << 8) |
        ((((f3 >> 7) & 0x01010101u) * 0x01020408u >> 24) << 12);
    const uint64_t rem = n - off;
    if (rem < 16) m &= (1u << rem) - 1u;
    return m;
}

__global__ void __launch_bounds__(IDX_THREADS)
k_count_newlines(const uint8_t *in, uint64_t n, uint32_t *tile_counts)
{
    __shared__ uint32_t s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const uint64_t chunk = (uint64_t)blockIdx.x * IDX_THREADS + threadIdx.x;
    uint32_t c = __popc(nl_mask16(in, chunk, n));
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = s_cnt;
}

// Single-block exclusive scan, in place; total stored to *total.  (V1 plumbing; the fused
// path uses a decoupled look-back instead.)
__device__ __forceinline__ void block_exclusive_scan(uint32_t *data, uint64_t n,
                                                     unsigned long long *total)
{
    __shared__ uint32_t s_warp[32];
    __shared__ unsigned long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (uint64_t base = 0; base < n; base += 1024 * 4) {
        const uint64_t i0 = base + (uint64_t)threadIdx.x * 4;
        uint32_t v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = (i0 + j < n) ? data[i0 + j] : 0u;
        const uint32_t mine = v[0] + v[1] + v[2] + v[3];
        uint32_t inc = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            uint32_t w = s_warp[lane], winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, winc, d);
                if (lane >= d) winc += t;
            }
            s_warp[lane] = winc - w;
        }
        __syncthreads();
        const unsigned long long carry = s_carry;
        uint32_t ex = (uint32_t)carry + s_warp[wid] + (inc - mine);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (i0 + j < n) data[i0 + j] = ex;
            ex += v[j];
        }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[31] + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = s_carry;
}

__global__ void __launch_bounds__(1024)
k_exclusive_scan(uint32_t *data, uint64_t n, unsigned long long *total)
{
    block_exclusive_scan(data, n, total);
}

// nlpos[j] = byte offset of the j-th '\n'.  tile_prefix = exclusive scan of tile_counts.
__global__ void __launch_bounds__(IDX_THREADS)
k_fill_newlines(const uint8_t *in, uint64_t n, const uint32_t *tile_prefix, uint32_t *nlpos)
{
    __shared__ uint32_t s_warp[IDX_THREADS / 32];
    const uint64_t chunk = (uint64_t)blockIdx.x * IDX_THREADS + threadIdx.x;
    uint32_t m = nl_mask16(in, chunk, n);
    const uint32_t c = __popc(m);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < IDX_THREADS / 32; w++)
        if (w < wid) wbase += s_warp[w];
    uint32_t pos = tile_prefix[blockIdx.x] + wbase + inc - c;
    while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        nlpos[pos++] = (uint32_t)(chunk * 16 + b);
    }
}

// After the line index: number of lines / records the batch holds.
//   final batch, last byte not '\n': the last char of the unterminated line is dropped by
//   the reference's line[:-1] (fastq.pyx:78,83,90), i.e. it acts as that line's newline.
__device__ __forceinline__ bool base_ok(uint8_t b)
{   // fastq.pyx:65
    const uint8_t u = b & 0xDF;     // clear bit 5: upper-case letters; '.' (0x2E) -> 0x0E
    return u == 'A' || u == 'C' || u == 'G' || u == 'T' || u == 'N' || b == '.';
}

// One block.  A record cut short by EOF still gets the reference's earlier checks first:
// header '@' (fastq.pyx:77) and the bases of a complete second line (fastq.pyx:81) are
// tested before the missing separator raises (fastq.pyx:86), so those errors win.
__global__ void k_finish_index(uint32_t *nlpos, const unsigned long long *n_newlines,
                               const uint8_t *in, uint64_t n, int final, BatchState *st)
{
    __shared__ uint32_t s_b0, s_b1;
    __shared__ unsigned long long s_rec;
    if (threadIdx.x == 0) {
        s_b0 = s_b1 = 0;
        unsigned long long lines = *n_newlines;
        if (final && n > 0 && in[n - 1] != '\n') {
            nlpos[lines] = (uint32_t)(n - 1);
            lines++;
        }
        unsigned long long nrec = lines / 4;
        st->first_bad = ~0ull;
        if (final && (lines & 3)) {
            // EOF inside a record.  1 line (header only) or 2 lines (no separator) is the
            // reference's IndexError at fastq.pyx:86; 3 lines (separator present, quality
            // line missing) is ACCEPTED with an empty quality string (readline() returns '').
            if ((lines & 3) == 3) {
                nlpos[lines] = (uint32_t)n;          // empty 4th line: start == end == n
                lines++;
                nrec++;
            } else {
                unsigned long long code = BAD_TRUNCATED;
                const uint32_t s0 = nrec ? nlpos[4 * nrec - 1] + 1 : 0;
                if (in[s0] != '@') code = BAD_HEADER;
                else if ((lines & 3) == 2) { s_b0 = nlpos[4 * nrec] + 1; s_b1 = nlpos[4 * nrec + 1]; }
                st->first_bad = (nrec << 8) | code;
                s_rec = nrec;
            }
        }
        st->n_lines = lines;
        st->n_records = nrec;
        st->consumed = final ? n : (nrec ? (unsigned long long)nlpos[4 * nrec - 1] + 1 : 0);
        st->n_selected = 0;
        st->out_bytes = 0;
        st->fallback = 0;
    }
    __syncthreads();
    bool bad = false;
    for (uint32_t i = s_b0 + threadIdx.x; i < s_b1; i += blockDim.x) bad |= !base_ok(in[i]);
    if (bad) atomicMin(&st->first_bad, (s_rec << 8) | BAD_BASES);
}

__device__ __forceinline__ int dec_digits(uint32_t v)
{   // branch-free: every lane of a warp runs the same nine compares
    return 1 + (v >= 10u) + (v >= 100u) + (v >= 1000u) + (v >= 10000u) + (v >= 100000u) +
           (v >= 1000000u) + (v >= 10000000u) + (v >= 100000000u) + (v >= 1000000000u);
}

// bytes PrintSequence emits for one selected record
__device__ __forceinline__ uint32_t record_out_len(uint32_t name_len, uint32_t L, uint32_t Q,
                                                   const Hit &f, const Hit &r)
{
    uint32_t a = 1 + name_len;                                  // '@' name
    if (L < 1000u) {
        // copies, starts and ends are all bounded by the read length: three digits at most
        const uint32_t fe = f.start + f.k * f.copies, re = r.start + r.k * r.copies;
        a += 9 + f.k + r.k + 6 +                                // 8 tabs, '\n', two motifs, one digit per number
             (f.copies >= 10u) + (f.copies >= 100u) + (f.start >= 10u) + (f.start >= 100u) + (fe >= 10u) + (fe >= 100u) +
             (r.copies >= 10u) + (r.copies >= 100u) + (r.start >= 10u) + (r.start >= 100u) + (re >= 10u) + (re >= 100u);
    } else {
    a += 1 + f.k + 1 + dec_digits(f.copies) + 1 + dec_digits(f.start) + 1 +
         dec_digits(f.start + f.k * f.copies);
    a += 1 + r.k + 1 + dec_digits(r.copies) + 1 + dec_digits(r.start) + 1 +
         dec_digits(r.start + r.k * r.copies) + 1;              // ... '\n'
    }
    a += L + 1 + 2 + Q + 1;                                     // seq \n + \n qual \n
    return a;
}

struct RecordLines {
    uint32_t s0, e0, s1, e1, s2, e2, s3, e3;   // [start, end) of each line's content
};

__device__ __forceinline__ RecordLines record_lines(const uint32_t *nlpos, uint64_t r)
{
    RecordLines x;
    x.s0 = r ? nlpos[4 * r - 1] + 1 : 0;
    x.e0 = nlpos[4 * r];
    x.s1 = x.e0 + 1; x.e1 = nlpos[4 * r + 1];
    x.s2 = x.e1 + 1; x.e2 = nlpos[4 * r + 2];
    x.s3 = x.e2 + 1; x.e3 = nlpos[4 * r + 3];
    if (x.e3 < x.s3) x.e3 = x.s3;               // synthetic empty 4th line at EOF
    return x;
}

__global__ void __launch_bounds__(128)
k_scan_records(const uint8_t *in, const uint32_t *nlpos, BatchState *st, ScanParams prm,
               ReadResult *res, uint32_t *outlen)
{
    const uint64_t nrec = st->n_records;
    for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x; base < nrec;
         base += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = base + threadIdx.x;
    const bool live = r < nrec;
    RecordLines ln = {};
    int L = 0;
    if (live) { ln = record_lines(nlpos, r); L = (int)(ln.e1 - ln.s1); }
    const int wmax = __reduce_max_sync(0xFFFFFFFFu, (unsigned)window_words(L, prm));
    int bad = 0;
    bool sel = false;
    Hit f, rc;
    f.k = rc.k = 0; f.copies = rc.copies = f.start = rc.start = 0;
    if (live) {
    // line[0] tests look at the raw first byte of the line, which at EOF may be the byte
    // that line[:-1] drops ("+" as an unterminated last line is still a separator).
    if (in[ln.s0] != '@') bad = BAD_HEADER;
    if (!bad)
        for (int i = 0; i < L; i++)
            if (!base_ok(in[ln.s1 + i])) { bad = BAD_BASES; break; }
    if (!bad && in[ln.s2] != '+') bad = BAD_SEPARATOR;
    if (bad) atomicMin(&st->first_bad, (unsigned long long)((r << 8) | (unsigned)bad));
    else match_read(in + ln.s1, L, prm, wmax, f, rc);
    sel = f.k && rc.k;
    }
    const unsigned selmask = __ballot_sync(0xFFFFFFFFu, sel);
    if ((threadIdx.x & 31) == 0 && selmask)
        atomicAdd(&st->n_selected, (unsigned long long)__popc(selmask));
    if (!live) continue;
    ReadResult out;
    out.fstart = sel ? f.start : 0; out.rstart = sel ? rc.start : 0;
    out.fck = sel ? ((uint32_t)f.copies << 3 | f.k) : 0;
    out.rck = sel ? ((uint32_t)rc.copies << 3 | rc.k) : 0;
    res[r] = out;
    outlen[r] = sel ? record_out_len(ln.e0 - ln.s0 - 1, L, ln.e3 - ln.s3, f, rc) : 0;
    }
}

__device__ __forceinline__ uint8_t *put_dec(uint8_t *p, uint32_t v)
{
    if (v < 1000u) {                           // copies, starts and ends of short reads: no loop
        const uint32_t h = v / 100u, r = v - 100u * h, t = r / 10u, u = r - 10u * t;
        if (v >= 100u) *p++ = (uint8_t)('0' + h);
        if (v >= 10u) *p++ = (uint8_t)('0' + t);
        *p++ = (uint8_t)('0' + u);
        return p;
    }
    const int d = dec_digits(v);
    for (int i = d - 1; i >= 0; i--) { p[i] = (uint8_t)('0' + v % 10); v /= 10; }
    return p + d;
}

__device__ __forceinline__ uint8_t upper(uint8_t c) { return (c >= 'a' && c <= 'z') ? c - 32 : c; }

__device__ __forceinline__ uint8_t comp_base(uint8_t c)
{
    switch (c) {
        case 'A': return 'T'; case 'T': return 'A';
        case 'C': return 'G'; case 'G': return 'C';
        default: return c;
    }
}

// PrintSequence for one record, byte at a time (V1).
__device__ inline void format_record(const uint8_t *in, const RecordLines &ln, const ReadResult &rr,
                                     int illumina, uint8_t *p)
{
    const int L = (int)(ln.e1 - ln.s1);
    const int fk = rr.fck & 7, fc = rr.fck >> 3, rk = rr.rck & 7, rcn = rr.rck >> 3;
    *p++ = '@';
    for (uint32_t i = ln.s0 + 1; i < ln.e0; i++) *p++ = in[i];
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
    for (int i = 0; i < L; i++) *p++ = upper(in[ln.s1 + i]);
    *p++ = '\n'; *p++ = '+'; *p++ = '\n';
    const uint8_t sub = illumina ? 31 : 0;                 // chr(ord(x)-64+33), pyx:181
    for (uint32_t i = ln.s3; i < ln.e3; i++) *p++ = (uint8_t)(in[i] - sub);
    *p++ = '\n';
}

__global__ void __launch_bounds__(128)
k_format_records(const uint8_t *in, const uint32_t *nlpos, BatchState *st, const ReadResult *res,
                 const uint32_t *outoff, int illumina, uint8_t *out, uint64_t outcap)
{
    const uint64_t nrec = st->n_records;
    if (st->out_bytes > outcap) return;          // host grows the buffer and re-runs
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < nrec;
         r += (uint64_t)gridDim.x * blockDim.x) {
        const ReadResult rr = res[r];
        if (rr.fck == 0) continue;
        const RecordLines ln = record_lines(nlpos, r);
        format_record(in, ln, rr, illumina, out + outoff[r]);
    }
}

// exclusive scan of outlen over the (device-side) record count
__global__ void __launch_bounds__(1024)
k_scan_outlen(uint32_t *outlen, BatchState *st)
{
    block_exclusive_scan(outlen, st->n_records, &st->out_bytes);
}

__global__ void __launch_bounds__(256)
k_match_reads(const uint8_t *seqs, const uint64_t *off, uint64_t n, ScanParams prm, ReadResult *res)
{
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = r < n;
    const uint64_t o = live ? off[r] : 0;
    const int L = live ? (int)(off[r + 1] - o) : 0;
    const int wmax = __reduce_max_sync(0xFFFFFFFFu, (unsigned)window_words(L, prm));
    if (!live) return;
    Hit f, rc;
    match_read(seqs + o, L, prm, wmax, f, rc);
    ReadResult out;
    const bool sel = f.k && rc.k;
    // the hook reports the forward hit even when the reverse strand has none (fck != 0,
    // rck == 0), so tests can check each strand against MatchPatternsAndReportBest
    out.fstart = f.k ? f.start : 0; out.rstart = sel ? rc.start : 0;
    out.fck = f.k ? ((uint32_t)f.copies << 3 | f.k) : 0;
    out.rck = sel ? ((uint32_t)rc.copies << 3 | rc.k) : 0;
    res[r] = out;
}

// -d/--debug report of ProcessSequence (select_STR_reads.pyx:294-311) for one read per thread:
// the rejection codes of the forward strand, and -- only if the forward strand has a best
// match, as in the reference -- those of the reverse complement.  codes of read r live at
// [coff[r], coff[r+1]): forward strand first, reverse strand from the middle of the slot.
// info[8 r ..] = {fwd count, fwd is None, fwd tail code, fwd tail value, same four for rc};
// tail = the -x algorithm's message about its winner ("Insufficient 5'/3' flank.: <value>").
__global__ void __launch_bounds__(128)
k_debug_reads(const uint8_t *seqs, const uint64_t *off, uint64_t n, ScanParams prm, const uint64_t *coff,
              uint8_t *codes, uint32_t *info)
{
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint8_t *seq = seqs + off[r];
    const int L = (int)(off[r + 1] - off[r]);
    const uint64_t half = (coff[r + 1] - coff[r]) / 2;
    uint32_t o[8];
    debug_read(seq, L, prm, codes + coff[r], codes + coff[r] + half, o);
    for (int i = 0; i < 8; i++) info[8 * r + i] = o[i];
}

__global__ void k_flush_l2(uint32_t *buf, uint64_t nwords, uint32_t v)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords;
         i += (uint64_t)gridDim.x * blockDim.x)
        buf[i] = v;
}

}  // namespace bstr
