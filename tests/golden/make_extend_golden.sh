#!/bin/sh
# Writes tests/golden/extend_readme.txt: what the REFERENCE's own extend_STR_reads (compiled in place by
# oracle/build_downstream.sh) prints for the README's walk-through (test_data/Makefile:10-12):
# extend_STR_reads 4000 20 27 merged.reads.str.fq reads1 reads2, the merged blocks being the reference's
# own merge output (tests/golden/merge_readme_k27.txt).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
BIN="$HERE/../../oracle/_ref/downstream/extend_STR_reads"
T=$(mktemp -d)
gzip -dc "$HERE/testdata_1.fq.gz" > "$T/1.fq"
gzip -dc "$HERE/testdata_2.fq.gz" > "$T/2.fq"
"$BIN" 4000 20 27 "$HERE/merge_readme_k27.txt" "$T/1.fq" "$T/2.fq" 2>/dev/null > "$HERE/extend_readme.txt"
rm -rf "$T"
md5sum "$HERE/extend_readme.txt"; head -c 200 "$HERE/extend_readme.txt"
