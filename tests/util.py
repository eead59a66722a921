"""Shared helpers for the test-suite (test infrastructure)."""
import gzip
import json
import os
import random

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# SURVEY.md 8(a) known-answer table, straight from the reference matcher
# (n=2, K=2..6, f=0, default algorithm unless stated).  fwd/rc = (motif, copies, start, end).
KNOWN = [
    ("GGCATATATATCG", {}, ("AT", 4, 3, 11), ("AT", 4, 2, 10)),
    ("CGTAAAAAAATGC", {}, ("AA", 3, 4, 10), ("TT", 3, 4, 10)),
    ("CGTAAAAAAATGC", {"rmhm": True}, None, None),
    ("CGTANANANATGC", {}, None, None),
    ("CGTACACNCACACTG", {}, ("CA", 2, 8, 12), ("TG", 2, 3, 7)),
    ("TTACACGGTGTGAA", {}, ("TG", 2, 8, 12), ("GT", 2, 8, 12)),
    ("TTACACGGACACAA", {}, ("CA", 2, 9, 13), ("TG", 2, 1, 5)),
    ("AGAGAGAGAGAGCTTCATGCA", {"f": 5}, ("GA", 3, 5, 11), None),
    ("CTTCATGCATAGAGAGAGAGAG", {"f": 5}, None, ("TC", 3, 5, 11)),
    ("TTAGAGAGAGAGAGAGCTTCATGCATTT", {"f": 8}, ("AG", 4, 8, 16), None),
    ("TTAGAGAGAGAGAGAGCTTCATGCATTT", {"f": 8, "altalgo": False}, None, None),
    ("ACACACACACACACTTGTGTGGGCAATCAGC", {"f": 9}, ("TG", 3, 15, 21), ("CA", 3, 10, 16)),
    ("ACACACACACACACTTGTGTGGGCAATCAGC", {"f": 9, "altalgo": False}, None, None),
    ("CGT....TGC", {}, ("..", 2, 3, 7), ("..", 2, 3, 7)),
    ("TTACGTAGACGTAGACGTAGACGTT", {}, ("TAGACG", 3, 5, 23), ("TCTACG", 3, 4, 22)),
    ("TTACGACGACGTT", {"n": 3}, ("ACG", 3, 2, 11), ("CGT", 3, 2, 11)),
    ("TTACGACGTT", {"n": 3}, None, None),
    ("TTACACACACACTT", {"ks": [3]}, None, None),
    ("TTACACACACACTT", {"ks": [4]}, ("CACA", 2, 3, 11), ("TGTG", 2, 3, 11)),
]


def gz_read(name):
    with gzip.open(os.path.join(GOLDEN, name), "rb") as fh:
        return fh.read()


def manifest():
    with open(os.path.join(GOLDEN, "MANIFEST.json")) as fh:
        return json.load(fh)


def fuzz_cases():
    return json.loads(gz_read("fuzz_cases.json.gz"))


def edge_cases():
    with open(os.path.join(GOLDEN, "edge_cases.json")) as fh:
        return json.load(fh)


def flags_to_kwargs(flags):
    """reference CLI flags -> keyword arguments shared by the oracle and the GPU API."""
    kw = {"n": 2, "f": 31, "ks": None, "rmhm": False, "altalgo": True, "illumina": False}
    ks, i = [], 0
    while i < len(flags):
        a = flags[i]
        if a == "-n":
            kw["n"] = int(flags[i + 1]); i += 1
        elif a == "-f":
            kw["f"] = int(flags[i + 1]); i += 1
        elif a == "-i":
            kw["illumina"] = True
        elif a == "-r":
            kw["rmhm"] = True
        elif a == "-x":
            kw["altalgo"] = False
        elif a[:1] == "-" and a[1:] in "23456":
            ks.append(int(a[1]))
        elif a[:2] == "--" and a[3:] == "mer":
            ks.append(int(a[2]))
        i += 1
    kw["ks"] = sorted(set(ks)) or None
    return kw


def synth_fastq(nreads, L, seed, p_str=0.30, p_n=0.02, p_lower=0.01, tag="t", vary_len=False,
                p_dot=0.0):
    """Small pure-Python synthetic FASTQ (tests only; the bench uses the C generator)."""
    rng = random.Random(seed)
    out = []
    for idx in range(nreads):
        ll = L if not vary_len else rng.randint(0, L)
        s = [rng.choice("ACGT") for _ in range(ll)]
        if ll and rng.random() < p_str:
            k = rng.randint(2, 6)
            motif = [rng.choice("ACGT") for _ in range(k)]
            c = rng.randint(3, 12)
            pos = rng.randint(0, max(0, ll - 1))
            for j in range(k * c):
                if pos + j < ll:
                    s[pos + j] = motif[j % k]
        if ll and rng.random() < p_n:
            s[rng.randrange(ll)] = "N"
        if ll and rng.random() < p_dot:
            s[rng.randrange(ll)] = "."
        s = "".join(s)
        if rng.random() < p_lower:
            s = s.lower()
        q = "".join(chr(rng.randint(35, 73)) for _ in range(ll))
        out.append("@%s.%010d/1\n%s\n+\n%s\n" % (tag, idx, s, q))
    return "".join(out).encode("latin-1")


def str_locus_reads(nloci, coverage, L, seed, err=0.01, indel=0.002, ctx=260, shared_flank=0.0):
    """Reads (both strands, substitution and indel errors, mixed qualities) from `nloci` STR loci
    with two alleles each: the input the merge stage sees after the select stage.  With
    shared_flank > 0 some loci reuse another locus's left context (reads of different loci then
    share keys: chains with several blocks)."""
    rng = random.Random(seed)
    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}
    out = []
    contexts = []
    idx = 0
    for locus in range(nloci):
        k = rng.randint(2, 6)
        while True:
            motif = "".join(rng.choice("ACGT") for _ in range(k))
            if len(set(motif)) > 1:
                break
        left = "".join(rng.choice("ACGT") for _ in range(ctx))
        right = "".join(rng.choice("ACGT") for _ in range(ctx))
        if contexts and rng.random() < shared_flank:
            left = rng.choice(contexts)
        contexts.append(left)
        c1 = rng.randint(4, 12)
        alleles = [c1, c1 if rng.random() < 0.3 else max(3, c1 + rng.choice([-3, -2, -1, 1, 2, 3]))]
        for _ in range(coverage):
            hap = left + motif * rng.choice(alleles) + right
            start = rng.randint(0, len(hap) - L)
            s = list(hap[start:start + L + 4])
            i = 0
            while i < len(s):
                u = rng.random()
                if u < err:
                    s[i] = rng.choice([b for b in "ACGT" if b != s[i]])
                elif u < err + indel:
                    if rng.random() < 0.5:
                        del s[i]
                        continue
                    s.insert(i, rng.choice("ACGT"))
                    i += 1
                i += 1
            s = "".join(s[:L])
            if rng.random() < 0.5:
                s = "".join(comp[b] for b in reversed(s))
            q = "".join(chr(rng.choice([35, 40, 50, 53, 54, 60, 70, 72, 73])) if rng.random() < 0.2 else "I" for _ in range(len(s)))
            out.append("@m.%08d/1\n%s\n+\n%s\n" % (idx, s, q))
            idx += 1
    rng.shuffle(out)
    return "".join(out).encode("latin-1")


def genome_reads(nreads, L, genome_len, seed, err=0.01):
    """reads of both strands from one random genome with substitution errors: real coverage, so
    most k-mers repeat (what extend_STR_reads is run on), the erroneous ones stay singletons"""
    import numpy as np
    rng = np.random.default_rng(seed)
    genome = rng.integers(0, 4, genome_len, dtype=np.uint8)
    starts = rng.integers(0, genome_len - L, nreads)
    codes = genome[starts[:, None] + np.arange(L)[None, :]]
    flip = rng.random(nreads) < 0.5
    codes[flip] = 3 - codes[flip][:, ::-1]
    hit = rng.random(codes.shape) < err
    codes[hit] = (codes[hit] + rng.integers(1, 4, int(hit.sum()), dtype=np.uint8)) & 3
    bases = np.frombuffer(b"ACGT", np.uint8)[codes]
    rec = np.empty((nreads, 12 + L + 3 + L + 1), np.uint8)
    names = np.array([list(b"@g%09d\n" % i) for i in range(nreads)], np.uint8)
    rec[:, :12] = names
    rec[:, 12:12 + L] = bases
    rec[:, 12 + L:12 + L + 3] = np.frombuffer(b"\n+\n", np.uint8)
    rec[:, 12 + L + 3:12 + 2 * L + 3] = ord("I")
    rec[:, -1] = 10
    return rec.tobytes()


def diploid_str_reads(genome_len, nloci, coverage, L, seed, snp_every=250, err=0.004, repeat_len=0):
    """Reads of both strands from a diploid genome with `nloci` STR loci (different copy numbers on the
    two haplotypes) and a SNP about every `snp_every` bases: what the extend stage walks through --
    long unique stretches, heterozygous forks that close again (bubbles), forks that do not (indels of
    the STR alleles), and with repeat_len > 0 a repeated segment (walks that meet themselves)."""
    rng = random.Random(seed)
    step = genome_len // (nloci + 1)
    hap_a, hap_b = [], []
    repeat = "".join(rng.choice("ACGT") for _ in range(repeat_len))
    for locus in range(nloci + 1):
        unique = "".join(rng.choice("ACGT") for _ in range(step))
        if repeat_len and locus % 3 == 1:
            unique = unique[:step // 2] + repeat + unique[step // 2:]
        other = list(unique)
        for i in range(len(other)):
            if rng.random() < 1.0 / snp_every:
                other[i] = rng.choice([b for b in "ACGT" if b != other[i]])
        hap_a.append(unique)
        hap_b.append("".join(other))
        if locus < nloci:
            k = rng.randint(2, 6)
            while True:
                motif = "".join(rng.choice("ACGT") for _ in range(k))
                if len(set(motif)) > 1:
                    break
            c = rng.randint(5, 12)
            hap_a.append(motif * c)
            hap_b.append(motif * (c if rng.random() < 0.3 else max(3, c + rng.choice([-2, -1, 1, 2, 3]))))
    haps = ["".join(hap_a), "".join(hap_b)]
    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}
    nreads = coverage * len(haps[0]) // L
    out = []
    for idx in range(nreads):
        hap = haps[idx & 1]
        start = rng.randint(0, len(hap) - L)
        s = list(hap[start:start + L])
        for i in range(L):
            if rng.random() < err:
                s[i] = rng.choice([b for b in "ACGT" if b != s[i]])
        s = "".join(s)
        if rng.random() < 0.5:
            s = "".join(comp[b] for b in reversed(s))
        q = "".join(chr(rng.choice([40, 50, 53, 54, 60, 70, 73])) if rng.random() < 0.1 else "I" for _ in range(L))
        out.append("@d.%08d/1\n%s\n+\n%s\n" % (idx, s, q))
    return "".join(out).encode("latin-1")


def diploid_str_reads_np(genome_len, nloci, coverage, L, seed, snp_every=250, err=0.004):
    """diploid_str_reads for large inputs (numpy): the same kind of genome -- STR loci with different copy numbers
    on the two haplotypes, a SNP about every `snp_every` bases -- and reads of both strands with substitution errors"""
    import numpy as np
    rng = np.random.default_rng(seed)
    step = genome_len // (nloci + 1)
    parts_a, parts_b = [], []
    for locus in range(nloci + 1):
        unique = rng.integers(0, 4, step, dtype=np.uint8)
        other = unique.copy()
        snp = rng.random(step) < 1.0 / snp_every
        other[snp] = (other[snp] + rng.integers(1, 4, int(snp.sum()), dtype=np.uint8)) & 3
        parts_a.append(unique)
        parts_b.append(other)
        if locus < nloci:
            k = int(rng.integers(2, 7))
            while True:
                motif = rng.integers(0, 4, k, dtype=np.uint8)
                if len(set(motif.tolist())) > 1:
                    break
            c = int(rng.integers(5, 13))
            c2 = c if rng.random() < 0.3 else max(3, c + int(rng.choice([-2, -1, 1, 2, 3])))
            parts_a.append(np.tile(motif, c))
            parts_b.append(np.tile(motif, c2))
    haps = [np.concatenate(parts_a), np.concatenate(parts_b)]
    nreads = coverage * len(haps[0]) // L
    half = nreads // 2
    codes = np.empty((nreads, L), np.uint8)
    for h, (lo, hi) in enumerate(((0, half), (half, nreads))):
        starts = rng.integers(0, len(haps[h]) - L, hi - lo)
        codes[lo:hi] = haps[h][starts[:, None] + np.arange(L)[None, :]]
    rng.shuffle(codes, axis=0)
    hit = rng.random(codes.shape) < err
    codes[hit] = (codes[hit] + rng.integers(1, 4, int(hit.sum()), dtype=np.uint8)) & 3
    flip = rng.random(nreads) < 0.5
    codes[flip] = 3 - codes[flip][:, ::-1]
    bases = np.frombuffer(b"ACGT", np.uint8)[codes]
    quals = np.full((nreads, L), ord("I"), np.uint8)
    low = rng.random(quals.shape) < 0.1
    quals[low] = rng.choice(np.frombuffer(b"(25<FI", np.uint8), int(low.sum()))
    rec = np.empty((nreads, 14 + L + 3 + L + 1), np.uint8)
    rec[:, :14] = np.array([list(b"@d.%08d/1\n" % i) for i in range(nreads)], np.uint8)
    rec[:, 14:14 + L] = bases
    rec[:, 14 + L:14 + L + 3] = np.frombuffer(b"\n+\n", np.uint8)
    rec[:, 17 + L:17 + 2 * L] = quals
    rec[:, -1] = 10
    return rec.tobytes()
