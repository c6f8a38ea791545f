"""Shared helpers of the parity tests (TEST INFRASTRUCTURE)."""
import numpy as np

import refbind
from svtme import synth

Q16 = 1 << 16


def make_pictures(binding, w, h, n, seed=0, kind="pan", hme_level1=1):
    frames = synth.make_clip(w, h, n, seed=seed, kind=kind)
    pics = [refbind.PaddedPicture(f, i) for i, f in enumerate(frames)]
    for p in pics:
        binding.build_planes(p, hme_level1)
    return frames, pics


def ref_lists(cur, n_frames, n0, n1):
    """Deterministic pseudo prediction structure: nearest past pictures in list 0, nearest future in list 1
    (falls back to past pictures when the future runs out, as base-layer pictures do)."""
    past = [p for p in range(cur - 1, -1, -1)]
    fut = [p for p in range(cur + 1, n_frames)]
    l0 = (past + fut)[:n0]
    l1 = (fut + past)[:n1]
    return l0, l1


def refs_dict(pics, l0, l1):
    d = {(0, i): pics[p] for i, p in enumerate(l0)}
    d.update({(1, i): pics[p] for i, p in enumerate(l1)})
    return d


def preset_knobs(binding, enc_mode, w, h, poc, tl, hl=3, **kw):
    args = dict(enc_mode=enc_mode, qp=35, frame_rate_q16=30 * Q16, width=w, height=h, hierarchical_levels=hl,
                picture_number=poc, temporal_layer_index=tl, is_ref=1 if tl < hl else 0, slice_type_b=1)
    args.update(kw)
    return binding.derive_knobs(**args)


def assert_same(a, b, what=""):
    bad = a.diff(b)
    if bad:
        n = bad[0]
        idx = np.argwhere(a[n] != b[n])[:5]
        raise AssertionError("%s: arrays differ: %s; first diffs of %s at %s: %s vs %s" % (
            what, bad, n, idx.tolist(), [a[n][tuple(i)] for i in idx], [b[n][tuple(i)] for i in idx]))
