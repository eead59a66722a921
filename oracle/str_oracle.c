/*
 * oracle/str_oracle.c -- CPU restatement of BaitSTR's select stage.  TEST INFRASTRUCTURE.
 *
 * This file is the checker the CUDA path is compared with.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it; the product
 * library (baitstr_b200/csrc) never links or calls it and has no CPU fallback.
 *
 * Parity pin: validated in tests/test_oracle.py against the UNMODIFIED reference
 * (oracle/_ref, built from /root/reference/src (fastq.pyx, select_STR_reads.pyx) by oracle/build_ref.py) on the
 * shipped test_data (golden md5 79a4e5db... / 51248b2b...), on the
 * known-answer table of SURVEY.md 8(a) and on random fuzz strings.
 *
 * What is restated (reference file:line):
 *   strand matcher, default algo      src/select_STR_reads.pyx:58-104
 *   strand matcher, -x algo           src/select_STR_reads.pyx:106-169
 *   the regex (?=((?=(.{k}))\2{n,}))  src/select_STR_reads.pyx:350  (CPython _sre:
 *       zero-width lookahead tried at every start i; group 2 = t[i:i+k]; \2{n,} is
 *       greedy, whole copies only, at least n copies INCLUDING the first)
 *   IsHomopolymer                     src/select_STR_reads.pyx:51-56
 *   ReverseComplement                 src/select_STR_reads.pyx:187-189
 *   ProcessSequence                   src/select_STR_reads.pyx:294-319
 *   PrintSequence                     src/select_STR_reads.pyx:171-185
 *   fastq.next / checkbases           src/fastq.pyx:62-94
 *
 * Plain C, no dependencies.  Build: gcc -O2 -shared -fPIC (see oracle/Makefile).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

typedef struct {
    int found;    /* 0 = None */
    int k;        /* motif length */
    int copies;
    int start;    /* 0-based, in the coordinates of the strand string that was scanned */
    int end;      /* start + k*copies */
} oracle_match_t;

typedef struct {
    int n;        /* -n  minimum copies             (select_STR_reads:80)  */
    int f;        /* -f  flanking distance          (select_STR_reads:81)  */
    int kmask;    /* bit k set => period k searched (select_STR_reads:134-136) */
    int rmhm;     /* -r                              */
    int altalgo;  /* 1 = default ("alt") algorithm; 0 = -x */
    int illumina; /* -i: quality bytes - 31 on output */
} oracle_params_t;

/* tuple order of select_STR_reads.pyx:93 = (len(STR), copies, STR, motif, start, end);
 * with len and copies equal k is equal, STR = motif*copies, so (STR, motif) compare
 * like the STR bytes themselves; end is implied by start. */
static int cand_greater(const char *t, int len_a, int c_a, int s_a,
                        int len_b, int c_b, int s_b)
{
    if (len_a != len_b) return len_a > len_b;
    if (c_a != c_b) return c_a > c_b;
    int cmp = memcmp(t + s_a, t + s_b, (size_t)len_a);
    if (cmp != 0) return cmp > 0;
    return s_a > s_b;
}

/* IsHomopolymer: len(set(motif)) == 1  (select_STR_reads.pyx:51-56) */
static int is_homopolymer(const char *m, int k)
{
    for (int j = 1; j < k; j++) if (m[j] != m[0]) return 0;
    return 1;
}

/* One strand.  t is already upper-cased (fastq.pyx:83) so re.IGNORECASE is moot. */
void oracle_match_strand(const char *t, int L, const oracle_params_t *p, oracle_match_t *out)
{
    int have = 0, b_len = 0, b_c = 0, b_s = 0, b_k = 0;
    int *run = (int *)malloc(sizeof(int) * (size_t)(L + 8));
    for (int k = 2; k <= 6; k++) {
        if (!(p->kmask & (1 << k))) continue;
        if (L < k) continue;
        /* run[i] = number of consecutive j >= i with t[j] == t[j+k] */
        for (int i = L; i >= L - k && i >= 0; i--) run[i] = 0;
        for (int i = L - k - 1; i >= 0; i--) run[i] = (t[i] == t[i + k]) ? run[i + 1] + 1 : 0;
        for (int i = 0; i + k <= L; i++) {            /* finditer: every start position */
            int c = 1 + run[i] / k;                   /* greedy whole copies, \2{n,} */
            if (c < p->n) continue;                   /* regex fails at this start */
            int len = k * c, end = i + len;
            if (memchr(t + i, 'N', (size_t)k)) continue;              /* pyx:76 / :138 */
            if (p->altalgo) {
                if (i < p->f) continue;                               /* pyx:81 */
                if (L - end < p->f) continue;                         /* pyx:84 */
                if (p->rmhm && is_homopolymer(t + i, k)) continue;    /* pyx:89 */
            }
            if (!have || cand_greater(t, len, c, i, b_len, b_c, b_s)) {
                have = 1; b_len = len; b_c = c; b_s = i; b_k = k;
            }
        }
    }
    free(run);
    out->found = 0;
    if (!have) return;
    if (!p->altalgo) {                                /* pyx:153-167: tests on the winner only */
        if (b_s < p->f) return;
        if (L - (b_s + b_len) < p->f) return;
        if (p->rmhm && is_homopolymer(t + b_s, b_k)) return;
    }
    out->found = 1; out->k = b_k; out->copies = b_c; out->start = b_s; out->end = b_s + b_len;
}

/* ReverseComplement: translate 'atcgnATCGN' -> 'tagcnTAGCN', everything else unchanged,
 * then reverse (select_STR_reads.pyx:187-189). */
void oracle_revcomp(const char *s, int L, char *out)
{
    for (int i = 0; i < L; i++) {
        char c = s[L - 1 - i], r = c;
        switch (c) {
            case 'a': r = 't'; break; case 't': r = 'a'; break;
            case 'c': r = 'g'; break; case 'g': r = 'c'; break;
            case 'A': r = 'T'; break; case 'T': r = 'A'; break;
            case 'C': r = 'G'; break; case 'G': r = 'C'; break;
            default: break;   /* n, N, '.', anything else */
        }
        out[i] = r;
    }
}

/* ProcessSequence (select_STR_reads.pyx:294-319): both strands must yield a match.
 * seq must already be upper-cased.  Returns 1 if the read is selected. */
int oracle_process_read(const char *seq, int L, const oracle_params_t *p,
                        oracle_match_t *fwd, oracle_match_t *rc, char *rcbuf)
{
    oracle_match_strand(seq, L, p, fwd);
    if (!fwd->found) { rc->found = 0; return 0; }
    oracle_revcomp(seq, L, rcbuf);
    oracle_match_strand(rcbuf, L, p, rc);
    return rc->found;
}

/* ---- whole-buffer select: fastq.next + ProcessSequence + PrintSequence ------------ */

enum {
    ORACLE_OK = 0,
    ORACLE_ERR_HEADER = -1,     /* fastq.pyx:77  line[0] != '@'            */
    ORACLE_ERR_BASES = -2,      /* fastq.pyx:81  checkbases failed         */
    ORACLE_ERR_SEPARATOR = -3,  /* fastq.pyx:86  line[0] != '+'            */
    ORACLE_ERR_TRUNCATED = -4,  /* fastq.pyx:86  ""[0] IndexError          */
    ORACLE_ERR_OUTCAP = -5
};

/* readline(): returns the line INCLUDING its '\n' if present; length 0 at EOF. */
static size_t next_line(const uint8_t *buf, size_t n, size_t pos)
{
    if (pos >= n) return 0;
    const uint8_t *nl = (const uint8_t *)memchr(buf + pos, '\n', n - pos);
    return nl ? (size_t)(nl - (buf + pos)) + 1 : n - pos;
}

static int base_ok(uint8_t b)
{
    switch (b) {
        case 'a': case 'c': case 'g': case 't': case 'n':
        case 'A': case 'C': case 'G': case 'T': case 'N': case '.': return 1;
        default: return 0;
    }
}

/*
 * Process a whole FASTQ file image.  Writes the annotated FASTQ of the selected reads
 * to out (capacity outcap) and, if results != NULL, 8 ints per input record:
 *   {selected, fk, fcopies, fstart, rk, rcopies, rstart, seqlen}.
 * Returns bytes written (>= 0) or a negative ORACLE_ERR_*; *nrec / *nsel get counts and
 * *bad_record the 0-based index of the offending record on error.
 *
 * line[:-1] semantics of fastq.pyx:78,83,90 are kept literally: the last character of
 * every line is dropped whether or not it is '\n' (so a final line without '\n' loses
 * its last quality character) and an empty line at a header/separator position is the
 * IndexError / assertion failure the reference raises.
 */
long oracle_select_buffer(const uint8_t *raw, size_t n, const oracle_params_t *p,
                          uint8_t *out, size_t outcap, int32_t *results, size_t results_cap,
                          long *nrec, long *nsel, long *bad_record)
{
    size_t pos = 0, o = 0;
    long rec = 0, sel = 0;
    size_t cap = 1024;
    char *seq = (char *)malloc(cap), *rcb = (char *)malloc(cap);
    long rv = 0;
    for (;;) {
        size_t l1 = next_line(raw, n, pos);
        if (l1 == 0) break;                                         /* fastq.pyx:73-75 */
        if (raw[pos] != '@') { rv = ORACLE_ERR_HEADER; break; }     /* fastq.pyx:77 */
        const uint8_t *name = raw + pos + 1; size_t name_len = l1 - 1 - (l1 >= 1 ? 1 : 0);
        if (l1 < 2) name_len = 0;                                   /* "@"[1:-1] == "" */
        pos += l1;
        size_t l2 = next_line(raw, n, pos);
        size_t L = l2 ? l2 - 1 : 0;                                 /* line[:-1] */
        for (size_t i = 0; i < L; i++)
            if (!base_ok(raw[pos + i])) { rv = ORACLE_ERR_BASES; goto done; }  /* fastq.pyx:62-67,81 */
        if (L + 1 > cap) { cap = 2 * (L + 1); seq = (char *)realloc(seq, cap); rcb = (char *)realloc(rcb, cap); }
        for (size_t i = 0; i < L; i++) {                            /* .upper(), fastq.pyx:83 */
            uint8_t b = raw[pos + i];
            seq[i] = (char)((b >= 'a' && b <= 'z') ? b - 32 : b);
        }
        pos += l2;
        size_t l3 = next_line(raw, n, pos);
        if (l3 == 0) { rv = ORACLE_ERR_TRUNCATED; break; }          /* ""[0] */
        if (raw[pos] != '+') { rv = ORACLE_ERR_SEPARATOR; break; }  /* fastq.pyx:86 */
        pos += l3;
        size_t l4 = next_line(raw, n, pos);
        const uint8_t *qual = raw + pos; size_t Q = l4 ? l4 - 1 : 0; /* line[:-1] */
        pos += l4;

        oracle_match_t fm, rm;
        int selected = oracle_process_read(seq, (int)L, p, &fm, &rm, rcb);
        if (results && (size_t)rec < results_cap) {
            int32_t *r = results + 8 * rec;
            r[0] = selected; r[1] = fm.found ? fm.k : 0; r[2] = fm.found ? fm.copies : 0;
            r[3] = fm.found ? fm.start : 0; r[4] = selected ? rm.k : 0;
            r[5] = selected ? rm.copies : 0; r[6] = selected ? rm.start : 0; r[7] = (int32_t)L;
        }
        rec++;
        if (!selected) continue;
        sel++;
        /* PrintSequence, select_STR_reads.pyx:174-185 */
        size_t need = name_len + 2 * L + Q + 128;
        if (o + need > outcap) { rv = ORACLE_ERR_OUTCAP; break; }
        out[o++] = '@'; memcpy(out + o, name, name_len); o += name_len;
        out[o++] = '\t'; memcpy(out + o, seq + fm.start, (size_t)fm.k); o += (size_t)fm.k;
        o += (size_t)sprintf((char *)out + o, "\t%d\t%d\t%d\t", fm.copies, fm.start, fm.end);
        memcpy(out + o, rcb + rm.start, (size_t)rm.k); o += (size_t)rm.k;
        o += (size_t)sprintf((char *)out + o, "\t%d\t%d\t%d\n", rm.copies, rm.start, rm.end);
        memcpy(out + o, seq, L); o += L; out[o++] = '\n';
        out[o++] = '+'; out[o++] = '\n';
        if (p->illumina) for (size_t i = 0; i < Q; i++) out[o++] = (uint8_t)(qual[i] - 64 + 33);
        else { memcpy(out + o, qual, Q); o += Q; }
        out[o++] = '\n';
    }
done:
    free(seq); free(rcb);
    if (nrec) *nrec = rec;
    if (nsel) *nsel = sel;
    if (bad_record) *bad_record = rv < 0 ? rec : -1;
    return rv < 0 ? rv : (long)o;
}
