set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_b.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu_b.log
for WL in C2 C3 C4 C5; do
for rep in 1 2; do for L in tools/ab/lib_base.so tools/ab/lib_m2.so baitstr_b200/libbaitstr_select.so; do
  BSTR_LIB=$PWD/$L python bench.py --workload $WL --reads 4000000 --steps 5 --warmup 3 --quick 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$WL', '$L', 'reads/s %.4g' % d['value'], 'kernel_us', d['roofline']['kernel_us_per_launch'], d['out_checksum_fnv1a'])"
done; done; done 2>&1 | grep -v "^+" | tee gpurun_out/r02_ab_m3.txt
ncu --set full --clock-control none --import-source on -k regex:k_fused_tile -s 3 -c 1 -f -o gpurun_out/r02_fused_m3 python bench.py --reads 2000000 --steps 2 --warmup 1 --quick > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -5
