"""TEST INFRASTRUCTURE — not part of the product.

Runs the compiled reference program oracle/_ref/smallgdpt_ref (built by oracle/smallgdpt/build_smallgdpt.sh from
/root/reference/small_gdpt.cpp, unmodified) and reads back the four PFM files it writes (small_gdpt.cpp:546-561).
"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
EXE = os.path.join(HERE, "..", "_ref", "smallgdpt_ref")
NAMES = ("image", "image_dx", "image_dy", "image_poisson")


def available():
    return os.path.exists(EXE)


def read_pfm(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"PF":
            raise ValueError("not a colour PFM: " + path)
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        a = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4")
    return a.reshape(h, w, 3).astype(np.float32)


def run(width, height, spp, threads=None, timeout=600):
    """The reference program at a film of width x height: {image, image_dx, image_dy, image_poisson} as float32
    [height][width][3], rows in file order, plus 'seconds' (wall clock of the whole program)."""
    import time
    env = dict(os.environ)
    if threads:
        env["OMP_NUM_THREADS"] = str(int(threads))
    with tempfile.TemporaryDirectory() as d:
        t0 = time.perf_counter()
        subprocess.run([EXE, str(int(width)), str(int(height)), str(int(spp)), d], check=True, env=env, timeout=timeout,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        dt = time.perf_counter() - t0
        out = {n: read_pfm(os.path.join(d, n + ".pfm")) for n in NAMES}
    out["seconds"] = dt
    return out


def stale_read_mask(width, height):
    """Pixels of the gradient files that can depend on memory the program never initialised.

    shiftPath (small_gdpt.cpp:200-285) fills offsetPath.verts / x / y but not offsetPath.rnds, and pathContrib / pathProb
    (:291-292, :337-338) form the offset path's first incoming direction from rnds[0..1]: for a first vertex on the glass
    sphere (the only material whose value depends on that direction, :308-322 / :134-147) the result depends on what the
    thread's stack held.  Observed here: zeros (exactly the pixel corner) in most runs, something else on one thread of a
    3-thread run — so a LIVE run of the program is compared outside this mask only; the committed fixtures are a run in which
    the words were zero everywhere (the host compilation of the restatement, which uses 0, reproduces them bit for bit).
    The mask is geometric and generous: every pixel whose footprint, grown by one pixel, has a camera ray passing within the
    glass sphere's radius of its centre, plus the right / lower neighbour that receives the pixel's cx1 / cy1 in the merge
    (:530-539).  [height][width] bool, rows in file order (the program flips them, :522)."""
    cam_o = np.array([50.0, 50.0, 295.6])
    cam_d = np.array([0.0, -0.042612, -1.0]); cam_d /= np.linalg.norm(cam_d)
    cx = np.array([width * .5135 / height, 0.0, 0.0])
    cy = np.cross(cx, cam_d); cy = cy / np.linalg.norm(cy) * .5135
    centre, rad = np.array([73.0, 16.5, 78.0]), 16.5
    hit = np.zeros((height, width), bool)
    ys, xs = np.mgrid[0:height, 0:width]
    for jx in (-1.0, 0.5, 2.0):
        for jy in (-1.0, 0.5, 2.0):
            d = (cx[None, None, :] * ((xs[..., None] + jx) / width - .5) + cy[None, None, :] * ((ys[..., None] + jy) / height - .5) + cam_d)
            d /= np.linalg.norm(d, axis=-1, keepdims=True)
            oc = centre - cam_o
            t = d @ oc
            dist2 = (oc @ oc) - t * t
            hit |= dist2 <= (rad * 1.02) ** 2
    hit = hit[::-1]                                   # file row = height - y - 1
    m = hit.copy()
    m[:, 1:] |= hit[:, :-1]
    m[1:] |= hit[:-1]
    return m
