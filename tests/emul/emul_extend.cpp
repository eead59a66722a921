// TEST INFRASTRUCTURE.  Host build of the product's extension walk (baitstr_b200/csrc/extend_core.cuh,
// the text the device runs as a warp) with a sequential executor over a k-mer table read from a dump
// ("kmer count" lines, as oracle/_ref/downstream/kmer_count_ref prints them for the same files and
// thresholds), so that the GPU-less container can compare it with the reference's own
// extend_STR_reads.  Nothing in the product uses this file.
//   usage: emul_extend KLEN FLANKS kmers.txt merged.reads.str.fq     (contigs to stdout)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <unordered_set>
#include <vector>

#include "../../baitstr_b200/csrc/extend_core.cuh"
#include "../../baitstr_b200/csrc/extend_plan.h"

using namespace bextend;

struct HostExec {
    const std::unordered_set<uint64_t> *table;
    unsigned lookup8(const uint64_t cand[8]) const
    {
        unsigned m = 0;
        for (int p = 0; p < 8; p++) if (cand[p] != ABSENT && table->count(cand[p])) m |= 1u << p;
        return m;
    }
    bool seen(const uint64_t *list, uint32_t n, uint64_t a, uint64_t b) const
    {
        for (uint32_t i = 0; i < n; i++) if (list[i] == a || list[i] == b) return true;
        return false;
    }
    bool leader() const { return true; }
    void sync() const {}
};

int main(int argc, char **argv)
{
    if (argc != 5) { fprintf(stderr, "usage: emul_extend KLEN FLANKS kmers.txt merged.reads.str.fq\n"); return 2; }
    int klen = atoi(argv[1]);
    if (klen % 2 == 0) klen--;
    const uint32_t flanks = (uint32_t)atoi(argv[2]);
    std::unordered_set<uint64_t> table;
    {
        FILE *f = fopen(argv[3], "r");
        if (!f) { perror(argv[3]); return 1; }
        unsigned long long k; unsigned c;
        while (fscanf(f, "%llu %u", &k, &c) == 2) table.insert(k);
        fclose(f);
    }
    std::string raw;
    {
        FILE *f = fopen(argv[4], "rb");
        if (!f) { perror(argv[4]); return 1; }
        char buf[1 << 16]; size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) raw.append(buf, n);
        fclose(f);
    }
    Plan P;
    if (!make_plan((const uint8_t *)raw.data(), raw.size(), P)) { fprintf(stderr, "%s\n", P.error.c_str()); return 1; }
    HostExec x{&table};
    std::vector<uint64_t> used(flanks + 1);
    std::vector<char> lf((size_t)klen + flanks + 1), rf((size_t)klen + flanks + 1);
    std::string out;
    uint64_t not_extended = 0;
    for (size_t i = 0; i < P.blocks.size(); i++) {
        Result r;
        extend_block(x, P.blocks[i], P.blob.data(), klen, flanks, used.data(), lf.data(), rf.data(), &r);
        not_extended += ends_not_extended(P.blocks[i], r, klen);
        print_contig(out, P, i, r, lf.data(), rf.data());
    }
    fwrite(out.data(), 1, out.size(), stdout);
    fprintf(stderr, "%zu blocks, %llu ends not extended\n", P.blocks.size(), (unsigned long long)not_extended);
    return 0;
}
