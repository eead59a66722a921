/*
 * synth_gen.c -- deterministic synthetic FASTQ generator (bench / test tooling, plain C).
 *
 * Workload definition of SURVEY.md 8(d): header "@syn<cfg>.<idx %010llu>/<mate>", bases
 * i.i.d. uniform ACGT; with p = 0.30 one planted STR (k ~ U{2..6}, motif ~ U(ACGT)^k,
 * copies ~ U{3..12} truncated at the read end, position ~ U[0, L - k*copies]); with
 * p = 0.02 one base overwritten by N; with p = 0.01 the sequence line lower-cased;
 * qualities Sanger, i.i.d. U{Q2..Q40} ('#'..'I').  Every read is seeded from
 * (seed, index) alone (splitmix64), so any prefix or shard is reproducible and shards can
 * be generated independently.
 */
#include <stdint.h>
#include <stdio.h>
#include <string.h>

static inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline uint32_t below(uint64_t *s, uint32_t n)   /* uniform in [0, n) */
{
    return (uint32_t)(((splitmix64(s) >> 32) * (uint64_t)n) >> 32);
}

/* bytes one record of read length L and index digits takes (fixed-width header) */
uint64_t synth_record_bytes(int cfg, int L)
{
    char tmp[64];
    int h = snprintf(tmp, sizeof tmp, "@syn%d.%010llu/%d\n", cfg, 0ull, 1);
    return (uint64_t)h + (uint64_t)L + 1 + 2 + (uint64_t)L + 1;
}

/*
 * Write reads [first, first + count) into out (capacity cap).  mate = 1 or 2 (header
 * suffix; it also perturbs the seed so mates differ).  Returns bytes written, or 0 if the
 * buffer is too small.
 */
uint64_t synth_fastq(uint8_t *out, uint64_t cap, int cfg, int L, uint64_t first, uint64_t count,
                     int mate, uint64_t seed)
{
    static const char B[4] = {'A', 'C', 'G', 'T'};
    const uint64_t rec = synth_record_bytes(cfg, L);
    if (count * rec > cap) return 0;
    uint8_t *p = out;
    for (uint64_t idx = first; idx < first + count; idx++) {
        uint64_t s = seed ^ (idx * 0xD1342543DE82EF95ull) ^ ((uint64_t)mate << 56);
        splitmix64(&s);
        p += sprintf((char *)p, "@syn%d.%010llu/%d\n", cfg, (unsigned long long)idx, mate);
        uint8_t *seq = p;
        for (int i = 0; i < L; i += 32) {
            uint64_t r = splitmix64(&s);
            for (int j = 0; j < 32 && i + j < L; j++) { seq[i + j] = (uint8_t)B[r & 3]; r >>= 2; }
        }
        uint64_t r = splitmix64(&s);
        if (L > 0 && (r & 0xFFFF) < (uint64_t)(0.30 * 65536)) {
            const int k = 2 + (int)below(&s, 5);
            char motif[6];
            uint64_t m = splitmix64(&s);
            for (int j = 0; j < k; j++) { motif[j] = B[m & 3]; m >>= 2; }
            int copies = 3 + (int)below(&s, 10);
            int span = k * copies;
            int pos = span >= L ? 0 : (int)below(&s, (uint32_t)(L - span + 1));
            for (int j = 0; j < span && pos + j < L; j++) seq[pos + j] = (uint8_t)motif[j % k];
        }
        if (L > 0 && ((r >> 16) & 0xFFFF) < (uint64_t)(0.02 * 65536)) seq[below(&s, (uint32_t)L)] = 'N';
        if (((r >> 32) & 0xFFFF) < (uint64_t)(0.01 * 65536))
            for (int i = 0; i < L; i++) seq[i] |= 0x20;
        p += L;
        *p++ = '\n'; *p++ = '+'; *p++ = '\n';
        for (int i = 0; i < L; i += 10) {
            uint64_t q = splitmix64(&s);
            for (int j = 0; j < 10 && i + j < L; j++) { *p++ = (uint8_t)('#' + (q & 63) * 39 / 64); q >>= 6; }
        }
        *p++ = '\n';
    }
    return (uint64_t)(p - out);
}
