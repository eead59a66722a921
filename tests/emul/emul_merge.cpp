// TEST INFRASTRUCTURE.  Host build of the merge stage's product sources -- merge_plan.h (parsing,
// keys, chains, the reference's table order, printing) and merge_core.cuh / merge_step.cuh (the
// per-read step that merge.cu runs as a warp) -- with a sequential executor, so that the GPU-less
// container can compare them with the reference's own merge_STR_reads (oracle/_ref/downstream).
// Nothing in the product uses this file.
//   usage: emul_merge KLENGTH MIN_THRESHOLD MAX_THRESHOLD ALL(0|1) reads.str.fq   (blocks to stdout)
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../baitstr_b200/csrc/merge_plan.h"
#include "../../baitstr_b200/csrc/merge_step.cuh"

using namespace bmerge;

struct HostExec {
    // Align's matrix, row by row as the reference fills it (:86-110)
    uint64_t dp_fill(const char *t1, int len1, const char *t2, int len2, uint8_t *T)
    {
        std::vector<int> prev((size_t)len2, 0), cur((size_t)len2, 0);
        uint64_t best = 0;
        for (int i = 1; i < len1; i++) {
            cur[0] = 0;
            for (int j = 1; j < len2; j++) {
                unsigned code;
                const int a = dp_cell(cur[j - 1], prev[j], prev[j - 1], t1[i - 1] == t2[j - 1], &code);
                cur[j] = a;
                t_put(T, PITCH, i, j, code);
                const uint64_t key = best_key(a, i, j);
                if (key > best) best = key;
            }
            prev.swap(cur);
        }
        return best;
    }
    bool leader() const { return true; }
    int bcast(int v) const { return v; }
    void sync() const {}
};

int main(int argc, char **argv)
{
    if (argc != 6) { fprintf(stderr, "usage: emul_merge KLENGTH MIN MAX ALL reads.str.fq\n"); return 2; }
    int klength = atoi(argv[1]);
    if (klength % 2 == 0) klength--;                      // merge_STR_reads.c:557-560
    FILE *f = fopen(argv[5], "rb");
    if (!f) { perror(argv[5]); return 1; }
    std::string raw;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) raw.append(buf, n);
    fclose(f);
    Plan P;
    if (!make_plan((const uint8_t *)raw.data(), raw.size(), klength, P)) { fprintf(stderr, "%s\n", P.error.c_str()); return 1; }
    std::vector<BlockHead> heads(P.keys.size());
    for (auto &h : heads) { memset(&h, 0, sizeof h); h.created_by = -1; }
    std::vector<char> bseqs(P.keys.size() * (size_t)MAXB), bquals(P.keys.size() * (size_t)MAXB);
    std::vector<uint8_t> T(T_BYTES);
    std::vector<char> sc(4 * MAXB);
    Scratch s{T.data(), sc.data(), sc.data() + MAXB, sc.data() + 2 * MAXB, sc.data() + 3 * MAXB, MAXF, PITCH, MAXB};
    HostExec x;
    // chain by chain, as the device does (the order of the chains does not matter)
    for (size_t c = 0; c + 1 < P.chain_start.size(); c++)
        for (uint32_t at = P.chain_start[c]; at < P.chain_start[c + 1]; at++) {
            const uint32_t i = P.chain_reads[at];
            const int e = process_read(x, P.reads[i], (int)i, heads.data(), bseqs.data(), bquals.data(), P.fwd.data(), P.qfwd.data(),
                                       P.rev.data(), P.qrev.data(), s);
            if (e) { fprintf(stderr, "read %u: merge error %d\n", i, e); return 1; }
        }
    // keys in the order their blocks were created = the order of the reads that created them
    std::vector<std::pair<int32_t, uint32_t>> made;
    for (uint32_t k = 0; k < heads.size(); k++) if (heads[k].created_by >= 0) made.push_back({heads[k].created_by, k});
    std::sort(made.begin(), made.end());
    std::vector<uint32_t> inserted;
    for (auto &m : made) inserted.push_back(m.second);
    std::string out;
    unsigned bindex = 1;
    for (uint32_t k : table_order(P.keys, inserted))
        print_block(out, bindex, P.keys.key(k), heads[k], &bseqs[(size_t)k * MAXB], &bquals[(size_t)k * MAXB], (unsigned)atoi(argv[2]),
                    (unsigned)atoi(argv[3]), atoi(argv[4]) != 0);
    fwrite(out.data(), 1, out.size(), stdout);
    fprintf(stderr, "%zu reads, %zu keys, %zu chains, %zu blocks\n", P.reads.size(), P.keys.size(), P.chain_start.size() - 1, inserted.size());
    return 0;
}
