"""TEST INFRASTRUCTURE -- CPU restatement of the k-mer counting pass of extend_STR_reads
(ReadAndCountNonSingletonKmers, /root/reference/src/extend_STR_reads.c:31-194) for the parity tests
of bstr_kmers_* (include/baitstr_select.h).  Never imported by the product path.

Pinned against the reference itself: tests/test_kmers.py compares it with the output of
oracle/_ref/downstream/kmer_count_ref (the reference's own function, compiled from its sources in
place by oracle/build_downstream.sh) and with the golden that binary wrote for the shipped
test_data (tests/golden/kmers_testdata_k27.txt.gz, made by tests/golden/make_kmer_golden.sh).

What the reference computes, restated:
  * reads come from fastq_seq.c:48-117; a read takes part iff slen >= klen (extend_STR_reads.c:58),
    and yields slen - klen + 1 k-mers of its base line;
  * the rolling 2-bit word is kmer.c:4-47 on uint64_t: BuildIndex over the first klen-1 bases,
    GetNextKmer = ((word << (64-2k+2)) >> (64-2k)) + code, the stored key is min(word,
    ReverseComplementKmer(word)); code = fasta_encoding[c] (utilities.h:61-75) = 0..3 for ACGTacgt,
    0 for N and n as well, and -1 (added as a 64-bit two's complement number) for anything else;
  * a k-mer seen once goes to the Bloom filter only, from the second sighting on it is counted in
    the map, the count saturating at 255 (:160-165); at the end entries outside
    [min_threshold, max_threshold] are erased (:187-193).  With min_threshold >= 2 the result is
    exactly { key : min(occurrences, 255) } filtered by the thresholds.
"""
M64 = (1 << 64) - 1


def code(c):
    # the table as it stands in utilities.h:61-75: N and n sit on a 0 like A (not on a -1)
    return {65: 0, 67: 1, 71: 2, 84: 3, 78: 0, 97: 0, 99: 1, 103: 2, 116: 3, 110: 0}.get(c, -1)


def revcomp_word(word, klen):
    rc = 0
    for _ in range(klen):
        rc = ((rc << 2) + (3 - (word & 3))) & M64
        word >>= 2
    return rc


def read_kmers(bases, klen):
    """canonical keys of one read's k-mers, in read order (kmer.c:4-47)"""
    word = 0
    for c in bases[:klen - 1]:
        word = ((word << 2) + code(c)) & M64
    sh1, sh2 = 64 - 2 * klen + 2, 64 - 2 * klen
    out = []
    for i in range(len(bases) - klen + 1):
        word = ((((word << sh1) & M64) >> sh2) + code(bases[i + klen - 1])) & M64
        out.append(min(word, revcomp_word(word, klen)))
    return out


def count_kmers(fastq_bytes_list, klen, min_threshold=2, max_threshold=255):
    """[(key, count)] sorted by key, for raw FASTQ byte strings (whole 4-line records each)"""
    assert min_threshold >= 2, "below 2 the reference's result depends on Bloom-filter false positives"
    counts = {}
    for raw in fastq_bytes_list:
        lines = raw.split(b"\n")
        if lines and lines[-1] == b"":
            lines.pop()
        assert len(lines) % 4 == 0
        for r in range(len(lines) // 4):
            bases, quals = lines[4 * r + 1], lines[4 * r + 3]
            bases = bases[:len(quals)]          # fastq_seq.c:100-102 cuts the bases at the quality line's length
            if len(quals) < klen or len(bases) < klen:
                continue
            for key in read_kmers(bases, klen):
                counts[key] = counts.get(key, 0) + 1
    out = []
    for key in sorted(counts):
        c = min(counts[key], 255)
        if min_threshold <= c <= max_threshold:
            out.append((key, c))
    return out
