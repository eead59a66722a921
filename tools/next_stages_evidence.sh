# Evidence for the two new stages on ONE GPU: parity tests, timing beside the reference's own code, then one
# `ncu --set full` capture of each kernel (only after the plain commands have exited 0 without ncu).
set -x
python -m pytest tests/test_kmers.py tests/test_merge.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r02_pytest_next_stages.log || exit 1
cat gpurun_out/r02_pytest_next_stages.log
BSTR_TRACE=1 python tools/kmer_bench.py 2> gpurun_out/kmer_trace.txt | tail -1 > gpurun_out/r02_kmer_bench.json || exit 1
cat gpurun_out/r02_kmer_bench.json; grep bstr_kmers gpurun_out/kmer_trace.txt | tail -1
ncu --set full --clock-control none --import-source on -k regex:k_kmer_count -s 1 -c 1 -f -o gpurun_out/r02_kmer_count python tools/kmer_bench.py 300000 150 1000000 27 > gpurun_out/ncu_k.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_merge_chains -s 1 -c 1 -f -o gpurun_out/r02_merge_chains python tools/merge_bench.py 4000 40 150 > gpurun_out/ncu_m.log 2>&1
ls -la gpurun_out/*.ncu-rep
