"""CPU tier: the N>1 path is pure host-side sharding (images are independent; no collective on
the data path).  World-size-2 gloo run of the shard arithmetic bench.py uses."""
import os
import subprocess
import sys

from tests.conftest import ROOT

WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["B2_ROOT"])
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
from jpeg_decoder_b200 import sharding
n = 256
lo, hi = sharding.weak_shard(n, rank)                             # what bench.py decodes on this rank
seeds = list(range(lo, hi))
t = torch.tensor([sharding.max_over_ranks(float(rank + 1))], dtype=torch.float64)   # time = max over ranks
mine = torch.tensor([seeds[0], seeds[-1]], dtype=torch.int64)
allr = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
dist.all_gather(allr, mine)
if rank == 0:
    ranges = sorted((int(a[0]), int(a[1])) for a in allr)
    assert all(ranges[i][1] + 1 == ranges[i + 1][0] for i in range(world - 1)), ranges   # disjoint, contiguous
    assert float(t) == world
    print("SHARD_OK", ranges)
dist.destroy_process_group()
'''


def test_weak_scaling_shards_are_disjoint(tmp_path):
    p = tmp_path / "w.py"
    p.write_text(WORKER)
    env = dict(os.environ, B2_ROOT=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29531", str(p)],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "SHARD_OK" in out.stdout


def test_shard_range_partitions_any_batch():
    from jpeg_decoder_b200.sharding import shard_range
    for total in (0, 1, 7, 256, 16384, 16385):
        for world in (1, 2, 3, 4, 8):
            parts = [shard_range(total, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
