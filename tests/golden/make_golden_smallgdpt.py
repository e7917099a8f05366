"""Generates tests/golden/smallgdpt_*.npz: the four PFM images of the reference program small_gdpt.cpp (compiled, unmodified,
by oracle/smallgdpt/build_smallgdpt.sh) at small films.  Run here, where /root/reference exists:
    python tests/golden/make_golden_smallgdpt.py
"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
from oracle import smallgdpt as ref  # noqa: E402

CASES = [(64, 48, 16), (128, 96, 4), (100, 60, 6)]      # width, height, spp (the last one is not 4:3)

if __name__ == "__main__":
    if not ref.available():
        sys.exit("oracle/_ref/smallgdpt_ref is not built (oracle/smallgdpt/build_smallgdpt.sh)")
    for w, h, spp in CASES:
        r = ref.run(w, h, spp)
        path = os.path.join(ROOT, "tests", "golden", "smallgdpt_%dx%d_%dspp.npz" % (w, h, spp))
        np.savez_compressed(path, width=w, height=h, spp=spp, **{n: r[n] for n in ref.NAMES})
        print(path, os.path.getsize(path), "bytes; mean", float(r["image"].mean()))
