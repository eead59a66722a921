#!/bin/bash
# TEST INFRASTRUCTURE.  Builds the reference's own downstream stages -- merge_STR_reads and
# extend_STR_reads, the consumers that pin the select stage's output format -- from the sources
# where they lie under /root/reference, into oracle/_ref/downstream (git-ignored; travels to the
# GPU box with gpurun).  Recipe = /root/reference/src/Makefile:52-73 without its in-tree writes:
# sparsehash is configured out of tree (only the generated sparseconfig.h is needed).
# Used by tests/test_downstream.py (the acceptance test of SURVEY.md section 4 / 8c).
set -e
REF=${REF:-/root/reference/src}
HERE=$(cd "$(dirname "$0")" && pwd)
B=$HERE/_ref/downstream
[ -d "$REF" ] || { echo "no reference sources at $REF"; exit 2; }
mkdir -p "$B/sparse"
if [ ! -f "$B/sparse/src/sparsehash/internal/sparseconfig.h" ]; then
  (cd "$B/sparse" && "$REF/sparsehash/configure" > configure.log 2>&1 && make src/sparsehash/internal/sparseconfig.h >> configure.log 2>&1)
fi
OBJS=""
for f in utilities sllist clparsing kmer murmur_hash bloom_filter fastq_seq; do
  gcc -O3 -w -c "$REF/$f.c" -o "$B/$f.o"
  OBJS="$OBJS $B/$f.o"
done
V=$(cat "$REF/VERSION")
for prog in merge_STR_reads extend_STR_reads; do
  g++ -O3 -w -D"VERSION=\"$V\"" -I"$B/sparse/src" -I"$REF/sparsehash/src" -I"$REF" -o "$B/$prog" $OBJS "$REF/$prog.c" -lz -lm
done
# the reference's k-mer counting pass alone (extend_STR_reads.c:31-194) behind a dump harness: the oracle
# of the GPU k-mer counter (bstr_kmers_*)
g++ -O3 -w -D"VERSION=\"$V\"" -Dmain=reference_extend_main -DREFERENCE_EXTEND_C="\"$REF/extend_STR_reads.c\"" \
    -I"$B/sparse/src" -I"$REF/sparsehash/src" -I"$REF" -o "$B/kmer_count_ref" $OBJS "$HERE/kmer_ref_harness.cc" -lz -lm
rm -f $OBJS
echo "oracle/_ref/downstream: ok"
