# Round-2 evidence on ONE GPU (each ncu pass only after the plain command has exited 0 without ncu):
#   full bench line, reference arm, ncu launch list of the bench command, one --set full capture of the hot kernel
set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_ref.err
python bench.py --steps 2 --warmup 1 --quick > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_fused_m2.csv python bench.py --steps 2 --warmup 1 --quick > gpurun_out/ncu_l.log 2>&1
for WL in C2 C3; do
  ncu --set full --clock-control none --import-source on -k regex:k_fused_tile -s 3 -c 1 -f -o gpurun_out/r02_fused_m2_$WL python bench.py --workload $WL --reads 2000000 --steps 2 --warmup 1 --quick > gpurun_out/ncu_f_$WL.log 2>&1
done
ls -la gpurun_out | tail -8
