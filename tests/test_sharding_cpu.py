"""N>1 host logic on CPU: world_size-2 gloo run of the shard arithmetic and the ordered merge
(the per-shard scan is done by the oracle here -- there is no GPU in this tier)."""
import os
import socket
import subprocess
import sys

from baitstr_b200 import shard
from tests import util

WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from baitstr_b200 import shard
from oracle import oracle
from tests import util
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
raw = util.synth_fastq(1001, 100, seed=7, p_str=0.5)
recs = raw.split(b"\n")[:-1]
first, count = shard.read_range(1001, rank, world)
mine = b"\n".join(recs[4 * first:4 * (first + count)]) + b"\n"
out = oracle.select_buffer(mine)
dist.barrier()
whole = shard.gather_in_rank_order(dist, out)
t = shard.reduce_scalar(dist, float(rank + 1), "max")
s = shard.reduce_scalar(dist, float(count), "sum")
assert t == world and s == 1001, (t, s)
if rank == 0:
    assert whole == oracle.select_buffer(raw)
    print("SHARD_OK", len(whole))
dist.destroy_process_group()
'''


def test_read_range_partitions():
    for total in (0, 1, 7, 1001, 10_000_000):
        for world in (1, 2, 3, 8):
            blocks = [shard.read_range(total, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and sum(c for _, c in blocks) == total
            for (f0, c0), (f1, _) in zip(blocks, blocks[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in blocks) - min(c for _, c in blocks) <= 1
    assert shard.file_block(list("abcdefgh"), 3, 8) == ["d"]
    assert shard.file_block(list("abc"), 1, 2) == ["c"] and shard.file_block(list("abc"), 0, 2) == ["a", "b"]


def test_world2_gloo_ordered_merge(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), util.ROOT],
                       capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "SHARD_OK" in r.stdout
