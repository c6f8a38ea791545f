#!/usr/bin/env python
"""Generates tests/golden/me_golden.npz from the COMPILED REFERENCE (oracle/_ref/libsvtref.so = the reference's own
motion_estimation.c / enc_mode_config.c / compute_sad_c.c ... built by oracle/Makefile from /root/reference).

The reference ships no golden vectors for the open-loop ME path, so the fixtures are outputs of
the unmodified reference code on a small deterministic clip: input luma, the SvtMeKnobs the reference's own
svt_aom_sig_deriv_me produced, and every output array of the path.  Run here (needs /root/reference for the build):

    make -C oracle ref && python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import refbind  # noqa: E402
import helpers as H  # noqa: E402

W, HH, N = 192, 136, 7
# (enc_mode, temporal layer, picture, extra preset fields)
CASES = [(6, 0, 4, {}), (6, 2, 3, {}), (10, 1, 2, {}), (2, 2, 3, {}), (13, 0, 4, {}), (8, 1, 2, {"slice_type_b": 0})]


def main():
    ref = refbind.Reference()
    ref.set_isa(0)  # plain C kernels of the reference
    out = {}
    frames, pics = H.make_pictures(ref, W, HH, N, seed=77, kind="pan")
    out["frames"] = np.stack(frames)
    for lvl in (1, 2):  # decimated planes incl. padding, as svt_aom_downsample_filtering_input_picture leaves them
        out["plane%d_pic3" % lvl] = pics[3].planes[lvl].copy()
    for n, (em, tl, cur, extra) in enumerate(CASES):
        cfg = dict(extra)
        k = H.preset_knobs(ref, em, W, HH, cur, tl, hl=3, **cfg)
        l0, l1 = H.ref_lists(cur, N, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
        res = ref.me_picture(k, pics[cur], H.refs_dict(pics, l0, l1), n_threads=1, extended=True)
        out["case%d_knobs" % n] = np.frombuffer(bytes(k), dtype=np.uint8).copy()
        out["case%d_meta" % n] = np.array([em, tl, cur, len(l0), len(l1)] + l0 + l1, dtype=np.int32)
        for name, arr in res.arrays.items():
            out["case%d_%s" % (n, name)] = arr.copy()
    path = os.path.join(HERE, "me_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote %s: %d arrays, %.1f KB" % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
