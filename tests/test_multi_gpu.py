"""NCCL tests of the N>1 path on a box with at least two GPUs (skipped otherwise): one process per GPU, the image
split into row bands, the distributed screened-Poisson solve (all-to-all transposes over NVLink) against the
gather-to-rank-0 solve on the same accumulated images."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["GDPT_ROOT"])
import numpy as np, torch, torch.distributed as dist
import gradient_based_path_tracing_b200 as g
from gradient_based_path_tracing_b200.multi import DistributedRenderer
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%s" % os.environ["GDPT_PORT"], rank=rank, world_size=world,
                        device_id=torch.device("cuda", rank))
sc = g.Scene.named(os.environ.get("GDPT_SCENE", "cbox_gdpt"))
res = {}
for solver in ("gather", "distributed"):
    r = DistributedRenderer(sc, rank, dist, torch, solver=solver)
    assert r.solver == solver
    res[solver] = r.render(spp=2, precision="f64", to_host=True)
if rank == 0:
    a, b = res["gather"], res["distributed"]
    for k in ("primal", "dx", "dy"):
        assert np.array_equal(a[k], b[k]), k                       # same samples, same merge
    rel = np.linalg.norm(a["poisson"] - b["poisson"]) / np.linalg.norm(a["poisson"])
    assert rel < 1e-12, rel                                        # gate 1e-5; only the DC summation order differs
    single = g.gradient_path_render(sc, spp=2, precision="f64")
    assert np.abs(single.primal - b["primal"]).max() < 1e-12
    rel = np.linalg.norm(single.poisson - b["poisson"]) / np.linalg.norm(single.poisson)
    assert rel < 1e-10, rel
    print("OK")
# the plain path integrator over the same bands
r = DistributedRenderer(sc, rank, dist, torch, solver="gather")
img = r.render_path(spp=2, precision="f64", to_host=True)
if rank == 0:
    want = g.path_render(sc, spp=2, precision="f64")
    assert np.abs(img - want).max() < 1e-12
    print("PATH OK")
# the BVH built on rank 0 only and broadcast (gdpt_create_ex + gdpt_bvh_adopt)
sc2 = g.Scene(sc.a)
r = DistributedRenderer(sc2, rank, dist, torch, solver="gather", share_bvh=True)
assert sc2.bvh_info(rank)[:3] == sc.bvh_info(rank)[:3]
shared = r.render(spp=2, precision="f64", to_host=True)
if rank == 0:
    assert np.array_equal(shared["primal"], res["gather"]["primal"]) and np.array_equal(shared["poisson"], res["gather"]["poisson"])
    print("SHARED BVH OK")
dist.barrier()
dist.destroy_process_group()
'''


def _gpus():
    import gradient_based_path_tracing_b200 as g
    return g._capi.lib().gdpt_device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_distributed_solve_equals_gather_solve_nccl(tmp_path, world):
    if _gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = str(33500 + (os.getpid() * 3 + world) % 2000)
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), GDPT_PORT=port, GDPT_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0] and "PATH OK" in outs[0]
