"""Dynamic-GOP detector HME level 0: dg_detector_hme_level0 (pd_process.c:493-590), the
TASK_DG_DETECTOR_HME branch of the ME kernel thread (me_process.c:327-329).

CPU: the oracle restatement against the compiled reference (C and AVX2 kernels) and against the committed golden
metrics (tests/golden/dg_golden.json, written by the compiled reference).
GPU: svtme_dg_detector_hme_level0 through the C ABI against the oracle (metrics, per-block SAD / vector, algorithmic
work) and against the golden metrics."""
import json
import os

import numpy as np
import pytest

import helpers as H
from svtme import api

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "dg_golden.json")))


def _ids(cases):
    return ["%dx%d-res%d-%s" % tuple(c["case"][:4]) for c in cases]


@pytest.mark.parametrize("g", GOLD, ids=_ids(GOLD))
def test_oracle_reproduces_golden_metrics(oracle, g):
    w, h, res, kind, seed, cur, rf = g["case"]
    _, po = H.make_pictures(oracle, w, h, 3, seed=seed, kind=kind)
    m, sad, mv = oracle.dg_hme_level0(po[cur], po[rf], res)
    assert m == g["metrics"]
    # the metrics are the sums of the per-block outputs
    assert int(sad.sum()) == m["tot_dist"] and int((sad > 16 * 16 * 30).sum()) == m["tot_cplx"]
    assert int((mv != 0).any(axis=1).sum()) == m["tot_active"]
    assert (mv % 4 == 0).all()


@pytest.mark.parametrize("isa", [0, 1], ids=["c", "avx2"])
@pytest.mark.parametrize("case", [(320, 184, 0, "pan"), (320, 184, 1, "noise"), (640, 360, 2, "pan"), (448, 264, 3, "pan"),
                                  (448, 264, 5, "noise"), (456, 272, 3, "pan"), (448, 264, 3, "zero"), (64, 64, 3, "pan"),
                                  (1280, 720, 3, "pan")],
                         ids=lambda c: "%dx%d-res%d-%s" % c)
def test_oracle_matches_reference(reference, oracle, case, isa):
    w, h, res, kind = case
    if reference.set_isa(isa) != isa:
        pytest.skip("AVX2 kernels not available")
    try:
        _, pr = H.make_pictures(reference, w, h, 3, seed=w + res, kind=kind)
        _, po = H.make_pictures(oracle, w, h, 3, seed=w + res, kind=kind)
        for cur, rf in ((2, 0), (0, 1)):
            a, _, _ = reference.dg_hme_level0(pr[cur], pr[rf], res)
            b, _, _ = oracle.dg_hme_level0(po[cur], po[rf], res)
            assert a == b, (case, cur, rf)
    finally:
        reference.set_isa(0)


# ---- GPU ----
@pytest.fixture(scope="module")
def dev():
    d = api.MeDevice(device=0)
    yield d
    d.close()


def _gpu_check(dev, oracle, w, h, res, kind, seed, cur, rf):
    frames, po = H.make_pictures(oracle, w, h, 3, seed=seed, kind=kind)
    want_m, want_sad, want_mv = oracle.dg_hme_level0(po[cur], po[rf], res)
    want_px = oracle.last_px_compares()
    a, b = dev.register(frames[cur], cur), dev.register(frames[rf], rf)
    try:
        n0 = dev.launch_count()
        m, sad, mv, ms, px = dev.dg_hme_level0(a, b, res)
        assert dev.launch_count() == n0 + 1 and ms > 0
    finally:
        a.release()
        b.release()
    assert np.array_equal(sad, want_sad), np.argwhere(sad != want_sad)[:5].tolist()
    assert np.array_equal(mv, want_mv), np.argwhere(mv != want_mv)[:5].tolist()
    assert m == want_m
    assert px == want_px
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("g", GOLD, ids=_ids(GOLD))
def test_gpu_matches_oracle_and_golden(dev, oracle, g):
    m = _gpu_check(dev, oracle, *g["case"])
    assert m == g["metrics"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", [(64, 64, 3, "pan"), (72, 40, 3, "noise"), (128, 64, 0, "pan"), (456, 272, 1, "pan"),
                                  (456, 272, 2, "noise"), (1000, 560, 3, "pan"), (448, 264, 6, "noise"), (3840, 2160, 5, "pan")],
                         ids=lambda c: "%dx%d-res%d-%s" % c)
def test_gpu_matches_oracle_more_shapes(dev, oracle, case):
    """Tiny pictures (windows clipped on every side, widths below 8), sizes that are not multiples of 64, every window
    class, and the benchmark's 4K size."""
    w, h, res, kind = case
    _gpu_check(dev, oracle, w, h, res, kind, seed=w + h, cur=1, rf=0)


@pytest.mark.gpu
def test_gpu_per_block_outputs_are_optional(dev, oracle):
    frames, po = H.make_pictures(oracle, 448, 264, 2, seed=5, kind="pan")
    want, _, _ = oracle.dg_hme_level0(po[1], po[0], 3)
    a, b = dev.register(frames[1], 1), dev.register(frames[0], 0)
    m, sad, mv, _, _ = dev.dg_hme_level0(a, b, 3, per_block=False)
    assert m == want and sad is None and mv is None


# ---- picture-level sums of the ME results ----
def _sums_of(res):
    return [int(res["me_8x8_distortion"].astype(np.uint64).sum()), int(res["rc_me_distortion"].astype(np.uint64).sum()),
            int(res["stationary_block_present_sb"].sum()), int(res["rc_me_allow_gm"].sum())]


@pytest.mark.parametrize("enc_mode", [4, 10])
def test_picture_sums_reference_and_oracle(reference, oracle, enc_mode):
    """picture_sums = what Initial RC (norm_me_dist, initial_rc_process.c:683-691) and svt_aom_global_motion_estimation
    (total_me_sad / total_stationary_sb / total_gm_sbs, global_me.c:162-168) sum over the per-block arrays."""
    w, h, n = 448, 264, 5
    _, pr = H.make_pictures(reference, w, h, n, seed=21, kind="pan")
    _, po = H.make_pictures(oracle, w, h, n, seed=21, kind="pan")
    k = H.preset_knobs(reference, enc_mode, w, h, 2, 1)
    l0, l1 = H.ref_lists(2, n, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    a = reference.me_picture(k, pr[2], H.refs_dict(pr, l0, l1), n_threads=2)
    b = oracle.me_picture(k, po[2], H.refs_dict(po, l0, l1), n_threads=2)
    assert a["picture_sums"].tolist() == _sums_of(a) == b["picture_sums"].tolist()
    assert a["picture_sums"][0] > 0


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0], ids=["staged"])
def test_gpu_picture_sums(oracle, pipeline):
    w, h, n = 832, 488, 5
    frames, po = H.make_pictures(oracle, w, h, n, seed=22, kind="pan")
    k = api.derive_knobs(enc_mode=6, qp=35, frame_rate_q16=30 << 16, width=w, height=h, hierarchical_levels=3,
                         picture_number=2, temporal_layer_index=1, is_ref=1, slice_type_b=1)
    l0, l1 = H.ref_lists(2, n, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    want = oracle.me_picture(k, po[2], H.refs_dict(po, l0, l1), n_threads=8)
    d = api.MeDevice(device=0, pipeline=pipeline)
    try:
        pd = [d.register(f, i) for i, f in enumerate(frames)]
        for rep in range(2):  # the sums are re-zeroed per search
            got = d.me_picture(k, pd[2], H.refs_dict(pd, l0, l1), extended=False)
            assert got["picture_sums"].tolist() == want["picture_sums"].tolist() == _sums_of(got)
    finally:
        d.close()


# ---- picture-analysis statistics on the sixteenth plane (second half) ----
STAT_CASES = [(448, 264, "pan"), (1280, 720, "pan"), (320, 184, "zero"), (1920, 1080, "noise"), (72, 136, "pan")]


@pytest.mark.parametrize("case", STAT_CASES, ids=lambda c: "%dx%d-%s" % c)
def test_picture_statistics_oracle_matches_reference(reference, oracle, case):
    """svt_aom_gathering_picture_statistics (pic_analysis_process.c:1559) of the compiled reference vs the restatement."""
    w, h, kind = case
    _, pr = H.make_pictures(reference, w, h, 1, seed=w, kind=kind)
    _, po = H.make_pictures(oracle, w, h, 1, seed=w, kind=kind)
    for step in (1, 4):
        for rw, rh in ((4, 4), (1, 1), (2, 3)):
            a, b = reference.picture_statistics(pr[0], rw, rh, step), oracle.picture_statistics(po[0], rw, rh, step)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2], (case, step, rw, rh)


def test_picture_statistics_properties(oracle):
    """Bins start at 1 and are scaled by 16 step^2: with step 1 the bins of a region sum to 16 (256 + its pixels)."""
    w, h = 448, 264
    _, po = H.make_pictures(oracle, w, h, 1, seed=3, kind="pan")
    hist, avg, luma = oracle.picture_statistics(po[0], 4, 4, 1)
    rw, rh = (w // 4) // 4, (h // 4) // 4
    assert int(hist[0][0].sum()) == 16 * (256 + rw * rh)
    assert int(hist.sum()) == 16 * (16 * 256 + (w // 4) * (h // 4))
    assert abs(luma - float(po[0].interior(2).mean())) <= 1


@pytest.mark.gpu
@pytest.mark.parametrize("case", STAT_CASES + [(3840, 2160, "pan"), (64, 64, "noise")], ids=lambda c: "%dx%d-%s" % c)
def test_gpu_picture_statistics(dev, oracle, case):
    w, h, kind = case
    frames, po = H.make_pictures(oracle, w, h, 1, seed=w, kind=kind)
    pic = dev.register(frames[0], 0)
    try:
        for step in (1, 4):
            for rw, rh in ((4, 4), (1, 1), (2, 3)):
                a, b = oracle.picture_statistics(po[0], rw, rh, step), dev.picture_statistics(pic, rw, rh, step)
                assert np.array_equal(a[0], b[0]), (case, step, rw, rh, np.argwhere(a[0] != b[0])[:4].tolist())
                assert np.array_equal(a[1], b[1]) and a[2] == b[2], (case, step, rw, rh)
    finally:
        pic.release()
