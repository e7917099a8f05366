"""World-size-2 gloo test of the N>1 host logic: row-band partition + the single reduce
that assembles the five accumulation images (and their halo rows) on rank 0."""
import os
import subprocess
import sys

from conftest import ROOT

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["GDPT_ROOT"])
import numpy as np, torch, torch.distributed as dist
from gradient_based_path_tracing_b200.multi import partition_rows, reduce_bands
from oracle import cport
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % os.environ["GDPT_PORT"], rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
h, w = 37, 20
rng = np.random.default_rng(11)
full = [rng.standard_normal((h, w, 3)) for _ in range(5)]          # what a single GPU would accumulate
y0, y1 = partition_rows(h, world)[rank]
mine = [torch.zeros(h * w * 3, dtype=torch.float64) for _ in range(5)]
for m, f in zip(mine, full):
    m.view(h, w, 3)[y0:y1] = torch.from_numpy(f[y0:y1])             # each rank only fills its band
reduce_bands(mine, dist, dst=0)
if rank == 0:
    got = [m.view(h, w, 3).numpy() for m in mine]
    for a, b in zip(got, full):
        assert np.array_equal(a, b)
    # the merge across the band border needs the neighbour's last row: check against the oracle's merge
    cx, cy = cport.merge(got[1], got[3], got[2], got[4])
    wx, wy = cport.merge(full[1], full[3], full[2], full[4])
    assert np.array_equal(cx, wx) and np.array_equal(cy, wy)
    print("OK")
dist.destroy_process_group()
'''


def test_two_rank_band_reduce_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", GDPT_PORT=port, GDPT_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0]
