// merge_plan.h -- host side of the merge stage's first slice (merge.cu):
// everything of merge_STR_reads that is not the alignment work.  Plain C++, shared with the host
// build the CPU suite drives (tests/emul/emul_merge.cpp).
//
//   parse      the select stage's records and their annotated names  (merge_STR_reads.c:350-355,
//              fastq_seq.c:48-117), both strands of every read       (fastq_seq.c:153-176)
//   keys       "motif lflank rflank" of the read as it is and reverse-complemented (:361-365, :385-390)
//   chains     reads that can ever meet share a key: connected components over (fkey, rkey); the
//              reference's loop is sequential only WITHIN a component, so components run in parallel
//   order      the reference prints its sparse_hash_map in table order (:433-436): the table is
//              replayed here -- MurmurHash3 x64 128 of the key (murmur_hash.c:68-178, low word),
//              divided by 8 (a pointer key: hashtable-common.h:351-361), 32 buckets to start with, doubling at 80 % occupancy with a rehash in table order,
//              triangular probing (sparsehashtable.h:610-658, 695-729, JUMP_ = num_probes)
//   print      the polymorphic blocks (:438-505)
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "merge_core.cuh"

namespace bmerge {


inline const unsigned char *complement_table()
{   // the IUPAC complement in either case, a blank for everything else (the table of fastq_seq.c:146-150)
    static unsigned char table[256];
    static const bool made = [] {
        static const char pairs[] = "ATCGBVDHKMRYNNSSWWXX";
        for (int c = 0; c < 256; c++) table[c] = ' ';
        for (int i = 0; pairs[i]; i += 2) {
            const unsigned char a = (unsigned char)pairs[i], b = (unsigned char)pairs[i + 1];
            table[a] = b; table[b] = a; table[a | 0x20] = (unsigned char)(b | 0x20); table[b | 0x20] = (unsigned char)(a | 0x20);
        }
        return true;
    }();
    (void)made;
    return table;
}
inline unsigned char complement_of(unsigned char c) { return complement_table()[c]; }

// the distinct keys of a run, numbered in order of first sight: open addressing over one arena of
// key bytes (a std::unordered_map<std::string, ...> costs an allocation per key)
struct KeyTable {
    std::string arena;                          // key bytes back to back
    std::vector<uint32_t> start, len;           // per key id
    std::vector<uint64_t> hash;                 // per key id: the hash the table order needs anyway
    std::vector<uint32_t> slot;                 // id + 1, 0 = empty
    uint64_t mask = 0;
    static uint64_t hash_of(const char *p, size_t n);
    uint32_t intern(const char *p, size_t n)
    {
        if ((start.size() + 1) * 2 > slot.size()) grow();
        const uint64_t h = hash_of(p, n);
        for (uint64_t at = h & mask;; at = (at + 1) & mask) {
            const uint32_t id1 = slot[at];
            if (!id1) {
                const uint32_t id = (uint32_t)start.size();
                start.push_back((uint32_t)arena.size()); len.push_back((uint32_t)n); hash.push_back(h);
                arena.append(p, n);
                slot[at] = id + 1;
                return id;
            }
            const uint32_t id = id1 - 1;
            if (hash[id] == h && len[id] == n && memcmp(arena.data() + start[id], p, n) == 0) return id;
        }
    }
    void grow()
    {
        const size_t n = slot.empty() ? 1024 : slot.size() * 2;
        slot.assign(n, 0); mask = n - 1;
        for (uint32_t id = 0; id < start.size(); id++) {
            uint64_t at = hash[id] & mask;
            while (slot[at]) at = (at + 1) & mask;
            slot[at] = id + 1;
        }
    }
    std::string key(uint32_t id) const { return arena.substr(start[id], len[id]); }
    size_t size() const { return start.size(); }
};

struct Plan {
    std::vector<Read> reads;
    std::string fwd, rev, qfwd, qrev;           // blobs: bases / qualities, as read and reversed(-complemented)
    KeyTable keys;                              // key id -> "motif lflank rflank"
    std::vector<uint32_t> chain_start;          // chains (components): reads of chain c = chain_reads[chain_start[c] .. chain_start[c+1])
    std::vector<uint32_t> chain_reads;          // in input order within a chain
    std::string error;
};

inline bool parse_name(const char *s, const char *e, Read &r)
{   // "%[^\t]\t%s\t%d\t%d\t%d\t%s\t%d\t%d\t%d" (:350-355)
    const char *p = s;
    while (p < e && *p != '\t') p++;
    if (p == s || p == e) return false;
    auto skip = [&]() { while (p < e && (*p == ' ' || *p == '\t')) p++; };
    auto word = [&](char *out, uint8_t &len) {
        skip();
        int n = 0;
        while (p < e && *p != ' ' && *p != '\t') { if (n < 6) out[n] = *p; n++; p++; }
        if (n == 0 || n > 6) return false;      // fmotif[7] in the reference
        out[n] = 0; len = (uint8_t)n;
        return true;
    };
    auto number = [&](uint16_t &out) {
        skip();
        long v = 0; int n = 0;
        while (p < e && *p >= '0' && *p <= '9') { v = v * 10 + (*p - '0'); if (v > 65535) return false; n++; p++; }
        out = (uint16_t)v;
        return n > 0;
    };
    return word(r.fmotif, r.fmlen) && number(r.fcopies) && number(r.fzstart) && number(r.fend) &&
           word(r.rmotif, r.rmlen) && number(r.rcopies) && number(r.rzstart) && number(r.rend);
}

// reads + blobs + key ids + chains; false (plan.error set) on input the reference would die on or
// read out of bounds for
inline bool make_plan(const uint8_t *raw, uint64_t nbytes, int klength, Plan &P)
{
    char msg[256];
    const char *p = (const char *)raw, *end = p + nbytes;
    uint64_t recno = 0;
    P.fwd.reserve(nbytes / 2); P.qfwd.reserve(nbytes / 2); P.rev.reserve(nbytes / 2); P.qrev.reserve(nbytes / 2);
    P.reads.reserve(nbytes / 256);
    auto line = [&](const char *&s, const char *&e) {
        if (p >= end) return false;
        s = p;
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        e = nl ? nl : end;
        p = nl ? nl + 1 : end;
        return true;
    };
    for (;;) {
        const char *ns, *ne, *bs, *be, *ps, *pe, *qs, *qe;
        if (!line(ns, ne) || ns == ne) break;                                // end of input (fastq_seq.c:44-51)
        recno++;
        if (!line(bs, be) || !line(ps, pe) || !line(qs, qe) || *ns != '@' || ps == pe || *ps != '+') {
            snprintf(msg, sizeof msg, "merge: record %llu is not a 4-line FASTQ record", (unsigned long long)recno);
            P.error = msg; return false;
        }
        Read r;
        memset(&r, 0, sizeof r);
        if (!parse_name(ns, ne, r)) {
            snprintf(msg, sizeof msg, "merge: error in parsing read name of record %llu (merge_STR_reads.c:350-355)", (unsigned long long)recno);
            P.error = msg; return false;
        }
        const size_t L = (size_t)(be - bs);
        if ((size_t)(qe - qs) != L || L == 0 || L > 65535) {
            snprintf(msg, sizeof msg, "merge: record %llu: bases and qualities differ in length (or are empty)", (unsigned long long)recno);
            P.error = msg; return false;
        }
        r.slen = (uint16_t)L;
        r.off = (uint32_t)P.fwd.size();
        if (P.fwd.size() + L + 1 > 0xFFFFFFF0ull) { P.error = "merge: more than 4 GiB of reads in one call"; return false; }
        P.fwd.append(bs, L); P.fwd.push_back('\0');
        P.qfwd.append(qs, L); P.qfwd.push_back('\0');
        P.rev.resize(P.fwd.size()); P.qrev.resize(P.fwd.size());
        {
            const unsigned char *comp = complement_table();
            char *rv = &P.rev[r.off], *qv = &P.qrev[r.off];
            for (size_t i = 0; i < L; i++) { rv[i] = (char)comp[(unsigned char)bs[L - 1 - i]]; qv[i] = qs[L - 1 - i]; }
            rv[L] = 0; qv[L] = 0;
        }
        const int k = klength;
        if (r.fzstart < k || r.fend + k > (int)L || r.rzstart < k || r.rend + k > (int)L || r.fend < r.fzstart || r.rend < r.rzstart) {
            snprintf(msg, sizeof msg, "merge: record %llu: a flank is shorter than klength %d (the reference reads outside the read there)", (unsigned long long)recno, k);
            P.error = msg; return false;
        }
        char key[8 + 2 * 1000 + 4];
        auto make_key = [&](const char *motif, int mlen, const char *bases, int zstart, int end_) {
            char *q = key;
            memcpy(q, motif, (size_t)mlen); q += mlen; *q++ = ' ';
            memcpy(q, bases + zstart - k, (size_t)k); q += k; *q++ = ' ';
            memcpy(q, bases + end_, (size_t)k); q += k;
            return (size_t)(q - key);
        };
        r.fkey = P.keys.intern(key, make_key(r.fmotif, r.fmlen, P.fwd.data() + r.off, r.fzstart, r.fend));
        r.rkey = P.keys.intern(key, make_key(r.rmotif, r.rmlen, P.rev.data() + r.off, r.rzstart, r.rend));
        P.reads.push_back(r);
    }
    // chains: union-find over key ids, a read joins its two keys
    std::vector<uint32_t> parent(P.keys.size());
    for (uint32_t i = 0; i < parent.size(); i++) parent[i] = i;
    auto find = [&](uint32_t x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
    for (const Read &r : P.reads) {
        const uint32_t a = find(r.fkey), b = find(r.rkey);
        if (a != b) parent[a < b ? b : a] = a < b ? a : b;
    }
    std::vector<uint32_t> chain_of(P.keys.size(), 0xFFFFFFFFu), count;
    std::vector<uint32_t> read_chain(P.reads.size());
    for (size_t i = 0; i < P.reads.size(); i++) {
        const uint32_t root = find(P.reads[i].fkey);
        if (chain_of[root] == 0xFFFFFFFFu) { chain_of[root] = (uint32_t)count.size(); count.push_back(0); }
        read_chain[i] = chain_of[root]; count[chain_of[root]]++;
    }
    P.chain_start.assign(count.size() + 1, 0);
    for (size_t c = 0; c < count.size(); c++) P.chain_start[c + 1] = P.chain_start[c] + count[c];
    P.chain_reads.resize(P.reads.size());
    std::vector<uint32_t> fill(P.chain_start.begin(), P.chain_start.end() - 1);
    for (size_t i = 0; i < P.reads.size(); i++) P.chain_reads[fill[read_chain[i]]++] = (uint32_t)i;
    return true;
}

// ---------------------------------------------------------------- the order of the reference's map

inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
inline uint64_t fmix64(uint64_t k)
{
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

// low 64 bits of MurmurHash3 x64 128, seed 0 -- what MurmurHasher returns as size_t (sparse_word_hash.h:51-58)
inline uint64_t murmur3_x64_128_low(const void *data, size_t len)
{
    const uint8_t *d = (const uint8_t *)data;
    const size_t nblocks = len / 16;
    uint64_t h1 = 0, h2 = 0;
    const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
    for (size_t i = 0; i < nblocks; i++) {
        uint64_t k1, k2;
        memcpy(&k1, d + 16 * i, 8); memcpy(&k2, d + 16 * i + 8, 8);
        k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
        h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
        k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
        h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
    }
    const uint8_t *tail = d + nblocks * 16;
    uint64_t k1 = 0, k2 = 0;
    const size_t rem = len & 15;
    for (size_t i = rem; i > 8; i--) k2 ^= (uint64_t)tail[i - 1] << (8 * (i - 9));
    if (rem > 8) { k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2; }
    for (size_t i = rem < 8 ? rem : 8; i > 0; i--) k1 ^= (uint64_t)tail[i - 1] << (8 * (i - 1));
    if (rem > 0) { k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1; }
    h1 ^= (uint64_t)len; h2 ^= (uint64_t)len;
    h1 += h2; h2 += h1;
    h1 = fmix64(h1); h2 = fmix64(h2);
    h1 += h2;
    return h1;
}

// key ids in the order the reference's table would list them, given the order they were inserted in
inline uint64_t KeyTable::hash_of(const char *p, size_t n) { return murmur3_x64_128_low(p, n); }

inline std::vector<uint32_t> table_order(const KeyTable &keys, const std::vector<uint32_t> &inserted)
{
    std::vector<uint64_t> hash(inserted.size());
    // the map's key type is a pointer (const char *): sparsehash drops the "known-0" low bits of the
    // hash of a pointer key, hash / sizeof(void *) (hashtable-common.h:351-361)
    for (size_t i = 0; i < inserted.size(); i++) hash[i] = keys.hash[inserted[i]] / 8;
    std::vector<int64_t> bucket(32, -1);        // index into `inserted`
    size_t n = 0;
    auto threshold = [](size_t buckets) { return (size_t)(buckets * 0.8f); };
    auto place = [&](std::vector<int64_t> &b, int64_t who) {
        const size_t mask = b.size() - 1;
        size_t at = (size_t)hash[(size_t)who] & mask, probes = 0;
        while (b[at] >= 0) { probes++; at = (at + probes) & mask; }
        b[at] = who;
    };
    for (size_t i = 0; i < inserted.size(); i++) {
        if (n + 1 > threshold(bucket.size())) {
            size_t sz = 4;
            while (n + 1 >= threshold(sz)) sz *= 2;
            if (sz > bucket.size()) {
                std::vector<int64_t> bigger(sz, -1);
                for (int64_t who : bucket) if (who >= 0) place(bigger, who);
                bucket.swap(bigger);
            }
        }
        place(bucket, (int64_t)i);
        n++;
    }
    std::vector<uint32_t> order;
    order.reserve(n);
    for (int64_t who : bucket) if (who >= 0) order.push_back(inserted[(size_t)who]);
    return order;
}

// ---------------------------------------------------------------- printing (:438-505)

// what the reference would print of a block: the first (up to three) copy numbers with >= 2
// supporting reads, newest entry of the list first (:445-455), and whether the block is printed (:457-460)
struct Verdict { unsigned copies[3], maxcopies; bool print; };
inline Verdict block_verdict(const BlockHead &h, unsigned min_threshold, unsigned max_threshold, bool include_all)
{
    Verdict v = {{0, 0, 0}, 0, false};
    int i = 0;
    for (int s = (int)h.nsup - 1; s >= 0 && i < 3; s--)      // the reference's list has the newest entry first
        if (h.sup_n[s] >= 2) { v.copies[i] = h.sup_copies[s]; if (v.copies[i] > v.maxcopies) v.maxcopies = v.copies[i]; i++; }
    v.print = h.support >= min_threshold && h.support <= max_threshold &&
              ((!include_all && v.copies[2] == 0 && v.copies[1] != 0) || (include_all && v.copies[2] == 0));
    return v;
}

// appends the block's four lines if it passes the thresholds; returns whether it did
inline bool print_block(std::string &out, unsigned &bindex, const std::string &key, const BlockHead &h, const char *seq, const char *qual,
                        unsigned min_threshold, unsigned max_threshold, bool include_all)
{
    const std::string motif = key.substr(0, key.find(' '));
    const Verdict v = block_verdict(h, min_threshold, max_threshold, include_all);
    if (!v.print) return false;
    const unsigned *copies = v.copies, maxcopies = v.maxcopies;
    char buf[96];
    snprintf(buf, sizeof buf, "@Block%u\t", bindex++);
    out.append(buf).append(motif).append("\t");
    if (include_all && copies[1] == 0) snprintf(buf, sizeof buf, "%u", copies[0]);
    else snprintf(buf, sizeof buf, "%u,%u", copies[0], copies[1]);
    out.append(buf);
    snprintf(buf, sizeof buf, "\t%d\t%d\n", (int)h.zstart, (int)(h.zstart + maxcopies * motif.size()));
    out.append(buf);
    out.append(seq, h.zstart);
    for (unsigned c = 0; c < maxcopies; c++) out.append(motif);
    if (h.slen > h.end) out.append(seq + h.end, (size_t)(h.slen - h.end));
    out.append("\n+\n");
    out.append(qual, h.zstart);
    out.append((size_t)maxcopies * motif.size(), '!');
    if (h.slen > h.end) out.append(qual + h.end, (size_t)(h.slen - h.end));
    out.append("\n");
    return true;
}

}  // namespace bmerge
