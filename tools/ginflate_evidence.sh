# GPU inflate evidence: stage timings (CUDA events), ingest.gz both ways, ncu launch list of a .gz run through bstr_run_files
set -x
python tools/ginflate_bench.py 2000000 > gpurun_out/gi_stages.txt 2>&1
for g in 1 0; do BSTR_GPU_INFLATE=$g python -c "
import bench, json
for wl, n in ((\"C2\", 8000000), (\"C4\", 4000000), (\"C5\", 6000000)):
    r = bench.measure_ingest(wl, n, 0)
    print(\"BSTR_GPU_INFLATE=$g\", wl, n, \"reads: gz\", json.dumps(r.get(\"gz\")), \"plain\", json.dumps(r.get(\"plain\")), flush=True)
"; done > gpurun_out/gi_ingest.txt 2>&1
cat > /tmp/gz_run.py <<'PY'
import os, subprocess, sys, tempfile
sys.path.insert(0, os.getcwd())
from baitstr_b200 import capi, synth
tmp = tempfile.mkdtemp(prefix="gzr_", dir="/dev/shm")
p = os.path.join(tmp, "r.fq")
synth.generate(2, 100, 0, 1000000, 1).tofile(p)
subprocess.check_call("gzip -1 -k %s" % p, shell=True)
with capi.Context(devices=[0]) as ctx, open(os.devnull, "wb") as dn:
    ctx.run_files([p + ".gz"], out_fd=dn.fileno(), err_fd=dn.fileno())
    print(ctx.stats())
subprocess.call("rm -rf %s" % tmp, shell=True)
PY
python /tmp/gz_run.py && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_ncu_launches_gz_run_files.csv python /tmp/gz_run.py > gpurun_out/ncu_gz.log 2>&1
tail -3 gpurun_out/ncu_gz.log
