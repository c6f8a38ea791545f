"""Writes tests/golden/dg_golden.json: DGDetectorMetrics of the reference's own dg_detector_hme_level0 (compiled by
oracle/Makefile from /root/reference, C kernels) for seeded synthetic picture pairs.  The pictures are regenerated
from the seeds by the tests (svtme/synth.py), only the metrics are stored.  Run here (needs /root/reference):
    python tests/golden/make_dg_golden.py"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as H  # noqa: E402
import refbind  # noqa: E402

# (width, height, input_resolution, clip kind, seed, source picture, reference picture)
CASES = [(320, 184, 0, "pan", 1, 2, 0), (640, 360, 1, "pan", 2, 1, 0), (832, 480, 2, "pan", 3, 2, 0),
         (832, 488, 2, "noise", 4, 1, 0), (1280, 720, 3, "pan", 5, 2, 0), (448, 264, 3, "pan", 6, 2, 1),
         (448, 264, 3, "zero", 7, 1, 0), (448, 264, 3, "static", 8, 1, 0), (1920, 1080, 4, "pan", 9, 2, 0)]

if __name__ == "__main__":
    ref = refbind.Reference()
    ref.set_isa(0)
    out = []
    for w, h, res, kind, seed, cur, rf in CASES:
        _, pics = H.make_pictures(ref, w, h, 3, seed=seed, kind=kind)
        m, _, _ = ref.dg_hme_level0(pics[cur], pics[rf], res)
        out.append({"case": [w, h, res, kind, seed, cur, rf], "metrics": m})
        print(out[-1])
    with open(os.path.join(HERE, "dg_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
