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
# temporal-filter mode: (tf hme_me_level, me_exit_th, centre picture, neighbouring picture)
TF_CASES = [(0, 0, 3, 2), (1, 256, 3, 5), (2, 256, 3, 4), (4, 3000, 3, 1)]
TF_ARRAYS = ("sb_best_sad", "sb_best_mv", "hme_sc", "do_ref", "tf_exit_sb", "tf_blk_counts")
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
    for n, (lvl, th, cur, nb) in enumerate(TF_CASES):
        k = ref.derive_tf_knobs(hme_me_level=lvl, qp_opt=1, me_exit_th=th, qp=35, width=W, height=HH,
                                picture_number=cur, temporal_layer_index=0, is_ref=1)
        res = ref.me_picture(k, pics[cur], {(0, 0): pics[nb]}, n_threads=1, extended=True)
        out["tf%d_knobs" % n] = np.frombuffer(bytes(k), dtype=np.uint8).copy()
        out["tf%d_meta" % n] = np.array([lvl, th, cur, nb], dtype=np.int32)
        for name in TF_ARRAYS:
            out["tf%d_%s" % (n, name)] = res.arrays[name].copy()
    path = os.path.join(HERE, "me_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote %s: %d arrays, %.1f KB" % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
