// extend_plan.h -- host side of the extension: the merged blocks parsed (extend_STR_reads.c:915-927,
// fastq_seq.c:48-117) and the contigs printed (PrintContig, :776-807).  Plain C++, shared by
// kmer_api.inc and the host build the CPU suite drives (tests/emul/emul_extend.cpp).
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "extend_core.cuh"

namespace bextend {

struct Plan {
    std::vector<Block> blocks;
    std::vector<std::string> name, motif, copies;   // "%s\t%s\t%s\t%d\t%d" of the block's first line
    std::string blob;                                // bases, NUL-terminated, back to back
    std::string error;
};

inline bool make_plan(const uint8_t *raw, uint64_t nbytes, Plan &P)
{
    char msg[256];
    const char *p = (const char *)raw, *end = p + nbytes;
    uint64_t recno = 0;
    auto line = [&](const char *&s, const char *&e) {
        if (p >= end) return false;
        s = p;
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        e = nl ? nl : end;
        p = nl ? nl + 1 : end;
        return true;
    };
    if (nbytes && raw[nbytes - 1] != '\n') {
        // GetZippedLine returns -1 for a line that ends at the end of the file (utilities.c:205-207): the reference dies
        P.error = "extend: the last line is not terminated by a newline (the reference's reader fails on it, utilities.c:205)";
        return false;
    }
    for (;;) {
        const char *ns, *ne, *bs, *be, *ps, *pe, *qs, *qe;
        if (!line(ns, ne)) break;                            // only the end of the text ends the input (fastq_seq.c:44-52)
        recno++;
        if (ns == ne || !line(bs, be) || !line(ps, pe) || !line(qs, qe) || *ns != '@' || ps == pe || *ps != '+') {
            snprintf(msg, sizeof msg, "extend: block %llu is not a 4-line FASTQ record", (unsigned long long)recno);
            P.error = msg; return false;
        }
        // sscanf(name, "%s\t%s\t%s\t%d\t%d"): five blank-separated fields
        std::vector<std::string> f;
        for (const char *q = ns; q < ne && f.size() < 5;) {
            while (q < ne && (*q == ' ' || *q == '\t')) q++;
            const char *w = q;
            while (q < ne && *q != ' ' && *q != '\t') q++;
            if (q > w) f.emplace_back(w, (size_t)(q - w));
        }
        char *e1 = nullptr, *e2 = nullptr;
        const long zs = f.size() == 5 ? strtol(f[3].c_str(), &e1, 10) : -1, en = f.size() == 5 ? strtol(f[4].c_str(), &e2, 10) : -1;
        if (f.size() != 5 || e1 == f[3].c_str() || e2 == f[4].c_str() || f[1].size() > 6) {
            snprintf(msg, sizeof msg, "extend: error in parsing the name of block %llu (extend_STR_reads.c:921-927)", (unsigned long long)recno);
            P.error = msg; return false;
        }
        const size_t L = (size_t)(be - bs) < (size_t)(qe - qs) ? (size_t)(be - bs) : (size_t)(qe - qs);   // fastq_seq.c:100-102
        if (zs < 0 || en < zs || (size_t)en > L || P.blob.size() + L + 1 > 0xFFFFFFF0ull) {
            snprintf(msg, sizeof msg, "extend: block %llu: the STR [%ld, %ld) does not lie in its %zu bases", (unsigned long long)recno, zs, en, L);
            P.error = msg; return false;
        }
        Block b;
        b.off = (uint32_t)P.blob.size(); b.slen = (uint32_t)L; b.zstart = (int32_t)zs; b.end = (int32_t)en;
        P.blob.append(bs, L); P.blob.push_back('\0');
        P.blocks.push_back(b);
        P.name.push_back(f[0]); P.motif.push_back(f[1]); P.copies.push_back(f[2]);
    }
    return true;
}

// how many ends of this block the reference counts as "not extended" (:933, :957)
inline unsigned ends_not_extended(const Block &b, const Result &r, int klen)
{
    unsigned n = klen > b.zstart ? 1u : 0u;
    if (r.status != ST_LEFT_PID && b.slen - (uint32_t)b.end < (uint32_t)klen) n++;
    return n;
}

// PrintContig (:776-807); nothing unless the block was extended on both sides
inline void print_contig(std::string &out, const Plan &P, size_t i, const Result &r, const char *lflank, const char *rflank)
{
    if (r.status != ST_PRINTED) return;
    const Block &b = P.blocks[i];
    char buf[64];
    out.append(">").append(P.name[i], 1, std::string::npos).append("\t").append(P.motif[i]).append(":").append(P.copies[i]);
    snprintf(buf, sizeof buf, ":%d:%d\n", (int)(r.llen + (uint32_t)b.zstart - r.indx1), (int)(r.llen + (uint32_t)b.end - r.indx1));
    out.append(buf);
    out.append(lflank, r.llen);
    if (r.indx2 > r.indx1) out.append(P.blob, b.off + r.indx1, r.indx2 - r.indx1);
    out.append(rflank, r.rlen);
    out.append("\n");
}

}  // namespace bextend
