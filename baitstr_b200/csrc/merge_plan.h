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
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <functional>
#include <memory>
#include <string>
#include <thread>
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
    uint32_t intern(const char *p, size_t n) { return intern(p, n, hash_of(p, n)); }
    uint32_t intern(const char *p, size_t n, uint64_t h)
    {
        if ((start.size() + 1) * 2 > slot.size()) grow();
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

struct Blob {                                    // bytes that are written before they are read: not cleared on allocation
    std::unique_ptr<char[]> p;
    size_t n = 0;
    void resize(size_t bytes) { p.reset(new char[bytes ? bytes : 1]); n = bytes; }
    char *data() { return p.get(); }
    const char *data() const { return p.get(); }
    size_t size() const { return n; }
    char &operator[](size_t i) { return p[i]; }
};

// the distinct keys of the input: ids, hashes (for the table order) and texts (for the motif when printing)
struct Keys {
    std::vector<uint64_t> hash;                 // per key id
    std::vector<const char *> text;             // into the shards' arenas
    std::vector<uint32_t> len;
    std::vector<KeyTable> shard;                // interned per hash range, by one thread each
    size_t size() const { return hash.size(); }
    std::string key(uint32_t id) const { return std::string(text[id], len[id]); }
};

struct Plan {
    std::vector<Read> reads;
    Blob fwd, rev, qfwd, qrev;                  // bases / qualities, as read and reversed(-complemented)
    Keys keys;                                  // key id -> "motif lflank rflank"
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
// read out of bounds for.  Parsing, both strands and the key hashes run on `threads` host threads over
// ranges of records; the keys are interned and the chains built by one
inline bool make_plan(const uint8_t *raw, uint64_t nbytes, int klength, Plan &P, unsigned threads = 0, bool host_strands = true)
{
    const char *text = (const char *)raw;
    const bool trace = getenv("BSTR_TRACE_PLAN") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto tp = now();
    auto lap = [&](const char *what) { if (trace) { auto t = now(); fprintf(stderr, "[merge plan] %-28s %.1f ms\n", what, std::chrono::duration<double, std::milli>(t - tp).count()); tp = t; } };
    if (threads == 0) { threads = std::thread::hardware_concurrency(); if (threads == 0) threads = 1; if (threads > 16) threads = 16; }
    if (nbytes < (4u << 20)) threads = 1;
    auto parallel = [&](size_t n, const std::function<void(unsigned, size_t, size_t)> &fn) {
        if (threads <= 1 || n < threads) { fn(0, 0, n); return; }
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < threads; t++) pool.emplace_back(fn, t, n * t / threads, n * (t + 1) / threads);
        for (auto &th : pool) th.join();
    };
    // 1. line ends
    std::vector<std::vector<uint64_t>> part(threads);
    parallel((size_t)nbytes, [&](unsigned t, size_t lo, size_t hi) {
        std::vector<uint64_t> &v = part[t];
        v.reserve((hi - lo) / 64 + 16);
        for (const char *q = text + lo, *e = text + hi; q < e;) {
            const char *nl = (const char *)memchr(q, '\n', (size_t)(e - q));
            if (!nl) break;
            v.push_back((uint64_t)(nl - text));
            q = nl + 1;
        }
    });
    std::vector<uint64_t> nl;
    { size_t total = 1; for (auto &v : part) total += v.size(); nl.reserve(total); }
    for (auto &v : part) nl.insert(nl.end(), v.begin(), v.end());
    if (nbytes && text[nbytes - 1] != '\n') {
        // GetZippedLine returns -1 for a line that ends at the end of the file (utilities.c:205-207): the reference dies
        P.error = "merge: the last line is not terminated by a newline (the reference's reader fails on it, utilities.c:205)";
        return false;
    }
    lap("line ends");
    auto line_start = [&](size_t i) { return i ? nl[i - 1] + 1 : 0; };
    size_t nrec = nl.size() / 4;
    const bool ragged = nl.size() % 4 != 0;
    // 2. names, lengths, checks.  Only the end of the text ends the input: a blank line where a record should start
    // fails the reference's assertion (fastq_seq.c:52)
    P.reads.assign(nrec, Read());
    std::vector<uint64_t> bad(threads, UINT64_MAX);
    std::vector<int> why(threads, 0);
    const int k = klength;
    parallel(nrec, [&](unsigned t, size_t lo, size_t hi) {
        for (size_t r = lo; r < hi; r++) {
            const char *ns = text + line_start(4 * r), *ne = text + nl[4 * r], *bs = text + line_start(4 * r + 1), *be = text + nl[4 * r + 1],
                       *ps = text + line_start(4 * r + 2), *pe = text + nl[4 * r + 2], *qs = text + line_start(4 * r + 3), *qe = text + nl[4 * r + 3];
            Read &rd = P.reads[r];
            memset(&rd, 0, sizeof rd);
            int e = 0;
            const size_t L = (size_t)(be - bs);
            if (ns == ne || *ns != '@' || ps == pe || *ps != '+') e = 1;
            else if (!parse_name(ns, ne, rd)) e = 2;
            else if ((size_t)(qe - qs) != L || L == 0 || L > 65535) e = 3;
            else if (rd.fzstart < k || rd.fend + k > (int)L || rd.rzstart < k || rd.rend + k > (int)L || rd.fend < rd.fzstart || rd.rend < rd.rzstart) e = 4;
            if (e) { bad[t] = r; why[t] = e; return; }
            rd.slen = (uint16_t)L;
        }
    });
    lap("names, lengths, checks");
    uint64_t first_bad = UINT64_MAX;
    int reason = 0;
    for (unsigned t = 0; t < threads; t++)
        if (bad[t] < first_bad) { first_bad = bad[t]; reason = why[t]; }
    if (first_bad >= nrec && ragged) { first_bad = nrec; reason = 1; }          // lines behind the last whole record: half a record
    if (first_bad <= nrec && first_bad != UINT64_MAX) {
        char msg[256];
        const unsigned long long recno = (unsigned long long)first_bad + 1;
        if (reason == 1) snprintf(msg, sizeof msg, "merge: record %llu is not a 4-line FASTQ record", recno);
        else if (reason == 2) snprintf(msg, sizeof msg, "merge: error in parsing read name of record %llu (merge_STR_reads.c:350-355)", recno);
        else if (reason == 3) snprintf(msg, sizeof msg, "merge: record %llu: bases and qualities differ in length (or are empty)", recno);
        else snprintf(msg, sizeof msg, "merge: record %llu: a flank is shorter than klength %d (the reference reads outside the read there)", recno, k);
        P.error = msg;
        return false;
    }
    P.reads.resize(nrec);
    // 3. offsets of the reads in the blobs
    uint64_t total = 0;
    for (size_t r = 0; r < nrec; r++) { P.reads[r].off = (uint32_t)total; total += (uint64_t)P.reads[r].slen + 1; }
    if (total > 0xFFFFFFF0ull) { P.error = "merge: more than 4 GiB of reads in one call"; return false; }
    // host_strands = false: the caller reverses the reads itself (merge.cu does it on the device); only the
    // two flanks of the reverse strand's key are complemented here
    P.fwd.resize(total); P.qfwd.resize(total);
    if (host_strands) { P.rev.resize(total); P.qrev.resize(total); }
    lap("offsets, blob allocation");
    // 4. both strands, and the two keys of every read with their hashes
    struct KeyRef { uint64_t hash; uint32_t at, len; };
    std::vector<KeyRef> kref(2 * nrec);
    std::vector<std::string> arena(threads);
    std::vector<unsigned> owner(nrec ? nrec : 1);
    parallel(nrec, [&](unsigned t, size_t lo, size_t hi) {
        const unsigned char *comp = complement_table();
        std::string &ar = arena[t];
        ar.reserve((hi - lo) * (size_t)(2 * (2 * k + 8)));
        for (size_t r = lo; r < hi; r++) {
            Read &rd = P.reads[r];
            const size_t L = rd.slen;
            const char *bs = text + line_start(4 * r + 1), *qs = text + line_start(4 * r + 3);
            char *fw = &P.fwd[rd.off], *qf = &P.qfwd[rd.off];
            memcpy(fw, bs, L); fw[L] = 0;
            memcpy(qf, qs, L); qf[L] = 0;
            if (host_strands) {
                char *rv = &P.rev[rd.off], *qv = &P.qrev[rd.off];
                for (size_t i = 0; i < L; i++) { rv[i] = (char)comp[(unsigned char)bs[L - 1 - i]]; qv[i] = qs[L - 1 - i]; }
                rv[L] = 0; qv[L] = 0;
            }
            auto make_key = [&](const char *motif, int mlen, const char *bases, int zstart, int end_, KeyRef &out) {
                out.at = (uint32_t)ar.size();
                ar.append(motif, (size_t)mlen).push_back(' ');
                ar.append(bases + zstart - k, (size_t)k).push_back(' ');
                ar.append(bases + end_, (size_t)k);
                out.len = (uint32_t)(ar.size() - out.at);
                out.hash = KeyTable::hash_of(ar.data() + out.at, out.len);
            };
            make_key(rd.fmotif, rd.fmlen, fw, rd.fzstart, rd.fend, kref[2 * r]);
            {   // the reverse strand's key: rev[i] = complement(fwd[L - 1 - i])
                KeyRef &out = kref[2 * r + 1];
                out.at = (uint32_t)ar.size();
                ar.append(rd.rmotif, (size_t)rd.rmlen).push_back(' ');
                for (int i = rd.rzstart - k; i < rd.rzstart; i++) ar.push_back((char)comp[(unsigned char)bs[L - 1 - i]]);
                ar.push_back(' ');
                for (int i = rd.rend; i < rd.rend + k; i++) ar.push_back((char)comp[(unsigned char)bs[L - 1 - i]]);
                out.len = (uint32_t)(ar.size() - out.at);
                out.hash = KeyTable::hash_of(ar.data() + out.at, out.len);
            }
            owner[r] = t;
        }
    });
    lap("strands, keys, hashes");
    // 5. key ids: the keys are split by hash over `threads` tables, every thread interns its share of all
    // keys (no locks), then the tables' local ids are made global by their tables' offsets
    P.keys.shard.assign(threads, KeyTable());
    parallel(threads, [&](unsigned, size_t lo, size_t hi) {
        for (size_t sh = lo; sh < hi; sh++) {
            KeyTable &T = P.keys.shard[sh];
            for (size_t i = 0; i < 2 * nrec; i++) {
                const KeyRef &kr = kref[i];
                if ((kr.hash >> 40) % threads != sh) continue;
                const uint32_t id = T.intern(arena[owner[i >> 1]].data() + kr.at, kr.len, kr.hash);
                if (i & 1) P.reads[i >> 1].rkey = id; else P.reads[i >> 1].fkey = id;
            }
        }
    });
    std::vector<uint32_t> base(threads + 1, 0);
    for (unsigned sh = 0; sh < threads; sh++) base[sh + 1] = base[sh] + (uint32_t)P.keys.shard[sh].size();
    P.keys.hash.resize(base[threads]); P.keys.text.resize(base[threads]); P.keys.len.resize(base[threads]);
    parallel(threads, [&](unsigned, size_t lo, size_t hi) {
        for (size_t sh = lo; sh < hi; sh++) {
            const KeyTable &T = P.keys.shard[sh];
            for (uint32_t id = 0; id < T.size(); id++) {
                P.keys.hash[base[sh] + id] = T.hash[id]; P.keys.text[base[sh] + id] = T.arena.data() + T.start[id]; P.keys.len[base[sh] + id] = T.len[id];
            }
        }
    });
    parallel(nrec, [&](unsigned, size_t lo, size_t hi) {
        for (size_t r = lo; r < hi; r++) {
            P.reads[r].fkey += base[(kref[2 * r].hash >> 40) % threads];
            P.reads[r].rkey += base[(kref[2 * r + 1].hash >> 40) % threads];
        }
    });
    lap("key ids");
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
    lap("chains");
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

inline std::vector<uint32_t> table_order(const Keys &keys, const std::vector<uint32_t> &inserted)
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
