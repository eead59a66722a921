#!/bin/sh
# Writes tests/golden/merge_*.txt: what the REFERENCE's own merge_STR_reads (compiled in place by
# oracle/build_downstream.sh) prints for the reference-generated select goldens of the shipped
# test_data.  readme = the README's walk-through (select -i -n 3 -f 29, merge 27); the others
# exercise --all and the thresholds on the default select run.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
BIN="$HERE/../../oracle/_ref/downstream/merge_STR_reads"
T=$(mktemp -d)
gzip -dc "$HERE/testdata_readme.out.gz" > "$T/readme.fq"
gzip -dc "$HERE/testdata_default.out.gz" > "$T/default.fq"
"$BIN" 27 "$T/readme.fq" 2>/dev/null > "$HERE/merge_readme_k27.txt"
"$BIN" --all --min_threshold=2 27 "$T/readme.fq" 2>/dev/null > "$HERE/merge_readme_k27_all_min2.txt"
"$BIN" --all --min_threshold=2 27 "$T/default.fq" 2>/dev/null > "$HERE/merge_default_k27_all_min2.txt"
"$BIN" --min_threshold=2 --max_threshold=6 23 "$T/default.fq" 2>/dev/null > "$HERE/merge_default_k23_min2_max6.txt"
rm -rf "$T"
md5sum "$HERE"/merge_*.txt; wc -l "$HERE"/merge_*.txt
