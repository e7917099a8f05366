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


# ---------------------------------------------------------------- distributed Poisson solve (plumbing)
# The halo exchange, the two all-to-all transposes (uneven splits), the DC all-reduce and the band gather of
# multi.DistributedPoisson run over gloo with the local stages restated in numpy (scipy's DCT-I), and the
# assembled result must equal the oracle's fourierSolve on the whole image.  The CUDA stages themselves are
# checked on the GPU (tests/test_parity_gpu.py::test_distributed_poisson_*).
NUMPY_STAGES = r'''
import numpy as np, scipy.fft


class NumpyStages:
    """Same contracts as multi.CudaStages (include/gdpt.h, gdpt_poisson_dist_*), on CPU torch tensors."""
    @staticmethod
    def cols(g):
        n, bw, w = g["n_ranks"], g["band_w"], g["w"]
        return [(s * bw, w if s == n - 1 else (s + 1) * bw) for s in range(n)]
    def sync(self):
        pass
    def merge_rows(self, g, ya, yb, cx0, cx1, cy0, cy1, cx, cy):
        w, h = g["w"], g["h"]
        v = lambda t: t.numpy().reshape(h, w, 3)
        for y in range(ya, yb):
            v(cx)[y] = v(cx0)[y]
            v(cx)[y, 1:] += v(cx1)[y, :-1]
            v(cy)[y] = v(cy0)[y] + (v(cy1)[y - 1] if y > 0 else 0.0)
    def rows_forward(self, g, u, gx, gy, alpha, send, dc_partial):
        w, h, y0, y1 = g["w"], g["h"], g["y0"], g["y1"]
        U, GX, GY = (t.numpy().reshape(h, w, 3) for t in (u, gx, gy))
        rows = y1 - y0
        rhs = np.zeros((rows, w, 3)); dc = np.zeros(3)
        for y in range(y0, y1):
            for x in range(w):
                ix, iy = 0 < x < w - 1, 0 < y < h - 1
                m = (2.0 if ix else 1.0) * (2.0 if iy else 1.0)
                dc += m * U[y, x]
                v = alpha * U[y, x]
                v = v - ((GX[y, x + 1] - GX[y, x]) if ix else (-2.0 * GX[y, x]))
                v = v - ((GY[y + 1, x] - GY[y, x]) if iy else (-2.0 * GY[y, x]))
                rhs[y - y0, x] = v
        dc_partial.numpy()[:] = dc
        t = scipy.fft.dct(rhs, type=1, axis=1)                    # rows transformed: [rows][w][3]
        chunks = [np.ascontiguousarray(t[:, c0:c1, :].transpose(0, 2, 1)).ravel() for c0, c1 in self.cols(g)]   # [rows][3][w_s]
        send.numpy()[:] = np.concatenate(chunks)
    def columns(self, g, alpha, dc, recv, send_back):
        w, h = g["w"], g["h"]
        x0, x1 = self.cols(g)[g["rank"]]
        ws = x1 - x0
        m = recv.numpy().reshape(h, 3, ws)
        f = scipy.fft.dct(m, type=1, axis=0)
        lapx = 2.0 * np.cos(np.pi * np.arange(x0, x1) / (w - 1))
        lapy = -4.0 + 2.0 * np.cos(np.pi * np.arange(h) / (h - 1))
        resp = (lapy[:, None] + lapx[None, :]).astype(np.float32).astype(np.float64)
        f = f / (alpha - resp)[:, None, :]
        if x0 == 0:
            f[0, :, 0] = dc.numpy()
        send_back.numpy()[:] = scipy.fft.dct(f, type=1, axis=0).ravel()
    def rows_inverse(self, g, recv_back, out):
        w, h, y0, y1 = g["w"], g["h"], g["y0"], g["y1"]
        rows = y1 - y0
        t = np.zeros((rows, w, 3)); off = 0
        r = recv_back.numpy()
        for c0, c1 in self.cols(g):
            n = rows * 3 * (c1 - c0)
            t[:, c0:c1, :] = r[off:off + n].reshape(rows, 3, c1 - c0).transpose(0, 2, 1); off += n
        o = scipy.fft.dct(t, type=1, axis=1) / (4.0 * (w - 1) * (h - 1))
        out.numpy().reshape(h, w, 3)[y0:y1] = o
'''

DIST_WORKER = NUMPY_STAGES + r'''
import os, sys
sys.path.insert(0, os.environ["GDPT_ROOT"])
import torch, torch.distributed as dist
from gradient_based_path_tracing_b200.multi import partition_rows, DistributedPoisson
from oracle import cport
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % os.environ["GDPT_PORT"], rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
for (h, w) in ((37, 22), (24, 31)):
    rng = np.random.default_rng(5)
    full = [rng.standard_normal((h, w, 3)) for _ in range(5)]          # img, cx0, cy0, cx1, cy1 of a single GPU
    bands = partition_rows(h, world)
    y0, y1 = bands[rank]
    mine = [torch.zeros(h * w * 3, dtype=torch.float64) for _ in range(5)]
    for m, f in zip(mine, full):
        m.view(h, w, 3)[y0:y1] = torch.from_numpy(f[y0:y1])             # each rank only holds its band
    cx, cy, out = (torch.zeros(h * w * 3, dtype=torch.float64) for _ in range(3))
    dp = DistributedPoisson(w, h, bands, dist, torch, NumpyStages())
    dp.solve(*mine, cx, cy, out, alpha=0.04)
    dp.gather_rows([mine[0], cx, cy, out], dst=0)
    if rank == 0:
        wx, wy = cport.merge(full[1], full[3], full[2], full[4])
        want = cport.poisson(full[0], wx, wy, 0.04)
        assert np.array_equal(cx.view(h, w, 3).numpy(), wx) and np.array_equal(cy.view(h, w, 3).numpy(), wy)
        assert np.array_equal(mine[0].view(h, w, 3).numpy(), full[0])
        rel = np.linalg.norm(out.view(h, w, 3).numpy() - want) / np.linalg.norm(want)
        assert rel < 1e-10, rel
if rank == 0:
    print("OK")
dist.destroy_process_group()
'''


def _run_ranks(tmp_path, text, world):
    script = tmp_path / "worker.py"
    script.write_text(text)
    port = str(31500 + (os.getpid() * 7 + world) % 2000)
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), GDPT_PORT=port, GDPT_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0]


def test_distributed_poisson_two_ranks_gloo(tmp_path):
    _run_ranks(tmp_path, DIST_WORKER, 2)


def test_distributed_poisson_three_ranks_gloo(tmp_path):
    _run_ranks(tmp_path, DIST_WORKER, 3)


def test_column_bands_cover_the_image():
    from gradient_based_path_tracing_b200.multi import column_bands
    for w in (9, 10, 22, 31, 1920, 3840):
        for n in (1, 2, 3, 4, 8):
            b = column_bands(w, n)
            assert b[0][0] == 0 and b[-1][1] == w and all(b[i][1] == b[i + 1][0] for i in range(n - 1))
            assert all(x1 > x0 for x0, x1 in b)
            assert len({x1 - x0 for x0, x1 in b[:-1]}) <= 1
