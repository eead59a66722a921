import os, subprocess, sys, tempfile
sys.path.insert(0, "/root/repo")
from baitstr_b200 import capi, synth
tmp = tempfile.mkdtemp(prefix="gib_", dir="/dev/shm")
p = os.path.join(tmp, "r.fq")
synth.generate(2, 100, 0, 2000000, 1).tofile(p)
subprocess.check_call("gzip -1 -k %s" % p, shell=True)
print(capi.gunzip_gpu(p + ".gz", -1, 0, 64 << 20, 32 << 10))
subprocess.call("rm -rf %s" % tmp, shell=True)
