"""Pins the oracle restatement (oracle/svtme_oracle.c) against the compiled reference (oracle/_ref): every output
array of the picture-level ME -- MVs, candidates, counts, the eight per-SB statistics, and the per-reference SAD/MV
tables, do_ref flags and HME centres -- must be byte-identical on the same inputs.  CPU only."""
import itertools

import numpy as np
import pytest

import helpers as H

W, HH = 448, 264  # 7 x 5 SB64 with a partial (8-row) last SB row; 480p-class... (small: whole file runs in ~1 min)


@pytest.fixture(scope="module")
def clips(reference, oracle):
    out = {}
    for kind in ("pan", "noise", "zero", "static"):
        _, pr = H.make_pictures(reference, W, HH, 9, seed=3, kind=kind)
        _, po = H.make_pictures(oracle, W, HH, 9, seed=3, kind=kind)
        for a, b in zip(pr, po):
            for lvl in range(3):
                assert np.array_equal(a.planes[lvl], b.planes[lvl]), "decimation/padding differs at level %d" % lvl
        out[kind] = (pr, po)
    return out


@pytest.mark.parametrize("enc_mode,tl", list(itertools.product([10, 8, 6, 4, 2], [0, 1, 3])))
def test_presets_pan(reference, oracle, clips, enc_mode, tl):
    pr, po = clips["pan"]
    k = H.preset_knobs(reference, enc_mode, W, HH, 4, tl)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    a = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=4)
    b = oracle.me_picture(k, po[4], H.refs_dict(po, l0, l1), n_threads=4)
    H.assert_same(a, b, "M%d tl%d" % (enc_mode, tl))


@pytest.mark.parametrize("kind", ["noise", "zero", "static"])
@pytest.mark.parametrize("enc_mode", [10, 6, 4])
def test_adversarial_clips(reference, oracle, clips, kind, enc_mode):
    pr, po = clips[kind]
    k = H.preset_knobs(reference, enc_mode, W, HH, 4, 2)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    a = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=4)
    b = oracle.me_picture(k, po[4], H.refs_dict(po, l0, l1), n_threads=4)
    H.assert_same(a, b, "%s M%d" % (kind, enc_mode))


@pytest.mark.parametrize("cfg", [
    dict(enc_mode=9, rtc_tune=1), dict(enc_mode=11, rtc_tune=1), dict(enc_mode=7, rtc_tune=1, sc_class1=1),
    dict(enc_mode=6, sc_class1=1), dict(enc_mode=10, sc_class1=1), dict(enc_mode=6, slice_type_b=0),
    dict(enc_mode=-1), dict(enc_mode=13), dict(enc_mode=5, qp=20), dict(enc_mode=6, qp=63),
    dict(enc_mode=7, similar_brightness_refs=1, hl=3, tl=3), dict(enc_mode=4, ref_list0_count_try=1, ref_list1_count_try=0),
    dict(enc_mode=12, rate_control_vbr=1),
])
def test_modes(reference, oracle, clips, cfg):
    """Low-delay (rtc), screen-content, P-slice, extreme presets/QPs: the branches the default presets do not take."""
    pr, po = clips["pan"]
    cfg = dict(cfg)
    tl, hl = cfg.pop("tl", 1), cfg.pop("hl", 3)
    em = cfg.pop("enc_mode")
    k = H.preset_knobs(reference, em, W, HH, 5, tl, hl=hl, **cfg)
    l0, l1 = H.ref_lists(5, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    a = reference.me_picture(k, pr[5], H.refs_dict(pr, l0, l1), n_threads=4)
    b = oracle.me_picture(k, po[5], H.refs_dict(po, l0, l1), n_threads=4)
    H.assert_same(a, b, str(cfg))


def test_single_thread_equals_multi_thread(reference, clips):
    """The reference's per-thread MeContext carries no state between SBs that changes results."""
    pr, _ = clips["pan"]
    k = H.preset_knobs(reference, 6, W, HH, 4, 1)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    a = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=1)
    b = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=5)
    H.assert_same(a, b)


def test_avx2_twin_equals_c(reference, clips):
    """The reference's AVX2 kernels give the same picture-level results as its C kernels (C/AVX2 bit-exactness)."""
    pr, _ = clips["pan"]
    k = H.preset_knobs(reference, 4, W, HH, 4, 1)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    reference.set_isa(0)
    a = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=4)
    if reference.set_isa(1) != 1:
        reference.set_isa(0)
        pytest.skip("no AVX2 on this host")
    try:
        b = reference.me_picture(k, pr[4], H.refs_dict(pr, l0, l1), n_threads=4)
    finally:
        reference.set_isa(0)
    H.assert_same(a, b)
