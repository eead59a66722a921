// TEST INFRASTRUCTURE.  Calls the reference's own k-mer counting pass --
// ReadAndCountNonSingletonKmers, /root/reference/src/extend_STR_reads.c:31-194 -- on FASTQ files and
// dumps the surviving (k-mer, count) pairs, sorted by k-mer, one "kmer count" pair per line.
// The reference's source file is compiled where it lies (its main() renamed by the build recipe,
// oracle/build_downstream.sh); nothing of it is copied into this repository.  The Bloom filter of
// the reference is seeded from time() (bloom_filter.c:112-115): its false positives enter the map
// with count 1 and are pruned again as long as min_threshold >= 2, which the harness insists on, so
// the dump is deterministic.
//   usage: kmer_count_ref KLEN MIN_THRESHOLD MAX_THRESHOLD EXPECTED_KMERS reads1.fq [reads2.fq ...]
#include <algorithm>
#include <vector>

#include REFERENCE_EXTEND_C        // -DREFERENCE_EXTEND_C='"/root/reference/src/extend_STR_reads.c"' -Dmain=reference_extend_main

#undef main
int main(int argc, char **argv)
{
    if (argc < 6) { fprintf(stderr, "usage: kmer_count_ref KLEN MIN MAX EXPECTED reads.fq ...\n"); return 2; }
    const uint klen = (uint)atoi(argv[1]), lo = (uint)atoi(argv[2]), hi = (uint)atoi(argv[3]);
    const uint64_t expected = strtoull(argv[4], NULL, 10);
    if (lo < 2) { fprintf(stderr, "min_threshold must be >= 2 (Bloom-filter false positives would survive)\n"); return 2; }
    t0 = time(0);
    debug_flag = FALSE;
    kmer_buffer = (char *)calloc(klen + 1, 1);
    // the function reads argv[4 .. nameidx): shift so that the files sit there
    std::vector<char *> av(4, (char *)"");
    for (int i = 5; i < argc; i++) av.push_back(argv[i]);
    SparseHashMap kmers;
    kmers.rehash(expected);
    kmers.set_deleted_key(0xFFFFFFFFFFFFFFFF);
    ReadAndCountNonSingletonKmers(kmers, expected, klen, av.data(), (uint)av.size(), 1u << 30, lo, hi);
    std::vector<std::pair<unsigned long long, unsigned> > out;
    for (SparseHashMap::iterator it = kmers.begin(); it != kmers.end(); ++it)
        out.push_back(std::make_pair((unsigned long long)(*it).first, (unsigned)(*it).second.count));
    std::sort(out.begin(), out.end());
    for (size_t i = 0; i < out.size(); i++) printf("%llu %u\n", out[i].first, out[i].second);
    return 0;
}
