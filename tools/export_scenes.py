#!/usr/bin/env python
"""Flatten the reference's scenes into GDPT1 containers under scenes/ (run where
/root/reference and oracle/_ref exist).  The reference's OWN parser and loaders read the
XML / OBJ / .serialized / JPG files (oracle/ref.py -> parse_scene), so the product is fed
exactly the Scene the reference would render; variants of the XML (integrator type, film
size, sample count, an added area emitter) are generated under oracle/_ref/scenes/, never
in the reference tree.

Configs (BASELINE.json / SURVEY §8): C1 cbox_gdpt, C2 veach_mi gradpath, C3 disney
(matpreview geometry + disneybsdf + added area emitter: gradpath ignores envmaps),
C4 sponza 1920x1080, C5 sponza 3840x2160 (camera-only overlay on C4's container).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from gradient_based_path_tracing_b200.scene import read_container, write_container  # noqa: E402

OUT = os.path.join(ROOT, "scenes")
FULL = ["cbox_gdpt", "cbox_gdpt_64", "cbox_gdpt_d2", "veach_mi_gradpath", "veach_mi_small", "disney_gradpath", "matpreview_small", "sponza_1080p"]
from gradient_based_path_tracing_b200.scene import OVERLAY_BASE as CAMERA_ONLY  # noqa: E402


def main():
    os.makedirs(OUT, exist_ok=True)
    ref.stage_scenes()
    only = sys.argv[1:]
    for name in FULL + list(CAMERA_ONLY):
        if only and name not in only:
            continue
        s = ref.RefScene(ref.scene_path(name))
        raw = os.path.join(OUT, "_tmp.gdpt")
        s.export(raw)
        s.close()
        arrays = read_container(raw)
        os.remove(raw)
        if name in CAMERA_ONLY:
            arrays = {k: v for k, v in arrays.items() if k.startswith("camera.") or k == "options"}
            path = os.path.join(OUT, name + ".camera.gdpt")
        else:
            path = os.path.join(OUT, name + ".gdpt.z")
        write_container(path, arrays)
        print(name, "->", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
