# round-end evidence: ncu launch list of the bench command and one full-set capture of the hot kernel
set -x
V=${V:-v12}
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$V.csv python bench.py --steps 2 --warmup 1 --quick > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fused_tile -s 3 -c 2 -f -o gpurun_out/fused_$V python bench.py --reads 2000000 --steps 2 --warmup 1 --quick > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out/ | tail -5
