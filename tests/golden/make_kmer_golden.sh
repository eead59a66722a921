#!/bin/sh
# Writes tests/golden/kmers_testdata_k27.txt.gz: the (k-mer, count) pairs the REFERENCE's own k-mer
# counting pass (extend_STR_reads.c:31-194, behind oracle/kmer_ref_harness.cc) keeps for the shipped
# test_data at the defaults of extend_STR_reads (k = 27 as in the README's walk-through, thresholds
# 2..255).  Needs oracle/_ref/downstream/kmer_count_ref (oracle/build_downstream.sh, this container).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
BIN="$HERE/../../oracle/_ref/downstream/kmer_count_ref"
T=$(mktemp -d)
gzip -dc "$HERE/testdata_1.fq.gz" > "$T/1.fq"
gzip -dc "$HERE/testdata_2.fq.gz" > "$T/2.fq"
"$BIN" 27 2 255 100000 "$T/1.fq" "$T/2.fq" 2>/dev/null | gzip -9n > "$HERE/kmers_testdata_k27.txt.gz"
rm -rf "$T"
gzip -dc "$HERE/kmers_testdata_k27.txt.gz" | md5sum
